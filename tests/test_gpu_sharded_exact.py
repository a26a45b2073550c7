"""Exact serving from document-range shards (query partitioning): dcd_search_batch_sharded on R shards,
concatenated in rank order, must equal Index.Search on the whole corpus -- EVERY page, position and rank, including the
documents next to a shard cut.  operator* ignores page and document boundaries (IndexSequence.cs:218-284: a group ends only on
a gap > absR), so gap-runs straddle the cuts; the corpora below are built so that they do: frequent words at the last and the
first positions of adjacent shards, dist 255 / null options, chains that span more than one document, a chain that runs
through a whole (tiny) shard, an empty shard between two bridged ones."""
import os

import numpy as np
import pytest

import docodo_b200 as dcd
from test_gpu_parity import RANK_RTOL, configure
from test_gpu_sharded import SEED, run_members, sharded_synthetic

pytestmark = pytest.mark.gpu


def sharded_search(members, comms, reqs, opts):
    """Collective exact search: every member on its own thread; per request the pages of all members in rank order."""
    res = [None] * len(members)

    def member(r):
        def go():
            res[r] = members[r].SearchBatch(reqs, opts, comm=comms[r])
        return go

    run_members([member(r) for r in range(len(members))])
    out = []
    for qi in range(len(reqs)):
        pages, base = [], 0
        for r, m in enumerate(members):
            assert res[r][qi].Success
            pages += [(p.ordinal + base, p.id, tuple(p.pos), p.rank) for p in res[r][qi].foundPages]
            base += int(m.counters.pages)
        out.append(pages)
    return out, res


def assert_exact(got_pages, want, req, max_pages=30000):
    exp = [(p["page_index"], p["id"], tuple(p["pos"]), p["rank"]) for p in want.foundPages]
    got = got_pages[:len(exp)] if len(got_pages) > max_pages else got_pages   # the cap is applied after the merge
    assert [g[:3] for g in got] == [e[:3] for e in exp], req
    for g, e in zip(got, exp):
        assert abs(g[3] - e[3]) <= RANK_RTOL * abs(e[3]), (req, g[3], e[3])


@pytest.mark.parametrize("R", [2, 3])
def test_exact_sharded_search_synthetic(R, data_dir, golden, stop_words, en_voc):
    """Zipf corpus: the frequent words occur every few dozen characters, so with dist 255 / null options the chains of the
    union of a request's lists cross every cut."""
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:20000]
    members, syns, comms, ora = sharded_synthetic(R, 240, words, stop_words, en_voc, golden)
    w = [x for x in words if x not in set(stop_words)]
    reqs = [f"{w[20]} {w[31]}", f"\"{w[2]} {w[3]}\"", f"{w[25]} ({w[36]} | {w[47]})", w[4], f"{w[0]} {w[1]}", f"{w[0]} {w[1]} {w[2]}",
            f"{w[5]} {w[300]}", f"\"{w[0]} {w[1]}\" {w[7]}", f"{w[0]} | {w[9]}", f"({w[0]} | {w[3]}) ({w[1]} | {w[2]}) {w[6]}",
            f"{w[1500]} {w[2]}", "zzzzqq", f"{w[0]} zzzzqq", f"{w[11]} {w[12]} {w[13]} {w[14]}", f"{w[3]}*", f"?{w[8][1:]}"]
    dists = [40, None, 100, None, 12, 255, None, 40, 40, 255, None, 40, 40, 255, 40, None]
    opts = [None if d is None else dcd.SearchOptions(dist=d) for d in dists]
    got, res = sharded_search(members, comms, reqs, opts)
    for req, d, g in zip(reqs, dists, got):
        assert_exact(g, ora.search(req, d), req)
    # pages-on-device form (the payload of the NCCL gather): same pages, same position counts
    import torch
    pres = [None] * R

    def member(r):
        def go():
            pres[r] = members[r].SearchBatchPages(reqs[:6], opts[:6], page_base=sum(int(m.counters.pages) for m in members[:r]),
                                                  device_out=lambda n: torch.empty(n, dtype=torch.int32, device="cuda"), comm=comms[r])
        return go

    run_members([member(r) for r in range(R)])
    for qi in range(6):
        pairs = []
        for r in range(R):
            cnts, pr = pres[r]
            at = int(cnts[:qi].sum())
            pairs += pr[at:at + int(cnts[qi])].cpu().tolist()
        assert pairs == [[g[0], len(g[2])] for g in got[qi]], reqs[qi]
    for m, s, cm in zip(members, syns, comms):
        if s is not None:
            m.synth_free(s)
        m.Dispose()
        cm.destroy()


def _build_custom(parts, stop_words, en_voc, golden):
    """parts: per shard a list of documents, each a list of page strings.  Returns (members, comms, oracle)."""
    ix0, ora = configure("envoc", stop_words, en_voc, golden)
    ix0.Dispose()
    all_pages, dfp = [], [0]
    for docs in parts:
        for pages in docs:
            all_pages += pages
            dfp.append(len(all_pages))
    text = "".join(all_pages).encode("latin-1")
    offs = np.concatenate([[0], np.cumsum([len(p) for p in all_pages])]).astype(np.uint64)
    ora.build_bulk(np.frombuffer(text, dtype=np.uint8), 1, offs, np.array(dfp, dtype=np.uint64), nthreads=1)
    R = len(parts)
    members = [configure("envoc", stop_words, en_voc, golden)[0] for _ in range(R)]
    comms = dcd.ShardComm.local(R)
    bases, doc_bases, acc, dacc = [], [], 0, 0
    for docs in parts:
        bases.append(acc)
        doc_bases.append(dacc)
        acc += sum(len(p) for pages in docs for p in pages)
        dacc += len(docs)

    def member(r):
        def go():
            b = members[r].GetBuilder()
            b.SetOrigin(bases[r], doc_bases[r])
            pages = [p for d in parts[r] for p in d]
            if pages:
                po = np.concatenate([[0], np.cumsum([len(p) for p in pages])]).astype(np.uint64)
                df = np.concatenate([[0], np.cumsum([len(d) for d in parts[r]])]).astype(np.uint64)
                b.AddPagesBulk(np.frombuffer("".join(pages).encode("latin-1"), dtype=np.uint8), dcd.ENC_LATIN1, po, df, source="doc")
            b.BuildSharded(comms[r])
        return go

    run_members([member(r) for r in range(R)])
    return members, comms, ora


def test_exact_sharded_search_adversarial(golden, stop_words, en_voc):
    """Hand-made cuts: a phrase split by the cut ("quick" is the last word of shard 0, "brown" the first of shard 1); a dense
    chain of "zebra" that starts two documents before a cut, runs through the whole tiny shard 2 and ends inside shard 3
    with the only "yak" beyond the far end; an EMPTY shard between two bridged ones; tags that meet only across a cut."""
    filler = " ".join("pad%d" % (i % 7) for i in range(60))            # ~ 300 characters without query words
    z = lambda n: " ".join(["zebra"] * n)                              # noqa: E731  dense chain, gap 6
    parts = [
        [[filler + " lion", "tiger " + filler], [filler, filler + " the quick"]],                 # shard 0 ends with "quick"
        [["brown fox " + filler, filler], [filler + " " + z(40)], [z(50)]],                        # shard 1: chain starts 2 docs before the cut
        [[z(30)]],                                                                                 # shard 2: nothing but the chain
        [],                                                                                        # shard 3: empty
        [[z(20) + " yak " + filler, filler + " lion"], [filler + " gnu"]],                         # shard 4: chain ends, then "yak"
        [["emu " + filler, filler]],                                                               # shard 5: "gnu emu" across the cut
    ]
    members, comms, ora = _build_custom(parts, stop_words, en_voc, golden)
    reqs = ["\"quick brown\"", "quick brown", "zebra yak", "\"zebra yak\"", "yak zebra", "gnu emu", "\"gnu emu\"", "\"emu gnu\"",
            "lion tiger", "zebra (yak | lion)", "zebra", "lion", "zebra yak lion", "fox quick", "brown (quick | fox)"]
    for d in (0, 5, 40, 255, None):
        opts = None if d is None else dcd.SearchOptions(dist=d)
        got, _ = sharded_search(members, comms, reqs, opts)
        for req, g in zip(reqs, got):
            assert_exact(g, ora.search(req, d), (req, d))
    # the shard-local answers (no exchange) do differ here: the cuts are straddled
    differs = 0
    for req in ("\"quick brown\"", "zebra yak", "gnu emu"):
        local, base = [], 0
        for m in members:
            r = m.Search(req, dcd.SearchOptions(dist=5))
            local += [(p.ordinal + base, tuple(p.pos)) for p in r.foundPages]
            base += int(m.counters.pages)
        differs += local != [(p["page_index"], tuple(p["pos"])) for p in ora.search(req, 5).foundPages]
    assert differs == 3
    want = ora.search("zebra yak", 5)
    assert len(want.foundPages) >= 4 and sum(len(p["pos"]) for p in want.foundPages) == 141   # the whole chain + "yak", 4 shards
    assert len(ora.search("\"quick brown\"", 0).foundPages) == 2
    for m, cm in zip(members, comms):
        m.Dispose()
        cm.destroy()
