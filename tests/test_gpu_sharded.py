"""GPU parity of the document-range sharded build: R shards with global coordinates, dictionary unification,
first-delta stitch and segment exchange must give, concatenated in rank order, exactly the single-builder `.index` /
`.index.list` bytes of the oracle (canonical semantics; replaces MergeAll / MergeFiles(shift_coords), Index.cs:537-752).
The in-process loopback transport runs R contexts on ONE device (each member on its own thread); the NCCL transport is
covered by tests/test_gpu_nccl.py when the box has >= 2 GPUs."""
import os
import threading

import numpy as np
import pytest

import docodo_b200 as dcd
from docsets import feed, indextest_source
from test_gpu_parity import configure

pytestmark = pytest.mark.gpu

SEED = 0xD0C0D00003


def run_members(fns):
    """Run one callable per shard member concurrently (the collectives rendezvous inside the library)."""
    errs = [None] * len(fns)

    def wrap(i):
        try:
            fns[i]()
        except BaseException as e:   # noqa: BLE001
            errs[i] = e

    th = [threading.Thread(target=wrap, args=(i,)) for i in range(len(fns))]
    for t in th:
        t.start()
    for t in th:
        t.join(300)
    for t in th:
        assert not t.is_alive(), "shard member hung"
    for e in errs:
        if e is not None:
            raise e


def sharded_synthetic(R, ndocs, words, stop_words, en_voc, golden, ranges=None):
    """Whole corpus on the oracle; R members each generate + index their own document range on the device."""
    ix0, ora = configure("envoc", stop_words, en_voc, golden)
    syn = ix0.synth_generate(words, 1.1, SEED, ndocs=ndocs, pages_per_doc=20, min_page_chars=200)
    text = ix0.d2h(syn.d_text, syn.text_bytes)
    offs = ix0.d2h(syn.d_page_offs, 8 * (syn.npages + 1), np.uint64)
    dfp = ix0.d2h(syn.d_doc_first_page, 8 * (syn.ndocs + 1), np.uint64)
    ora.build_bulk(text, 1, offs, dfp, nthreads=1)
    ix0.synth_free(syn)
    ix0.Dispose()

    ranges = ranges or dcd.shard_ranges(ndocs, R)
    members = [configure("envoc", stop_words, en_voc, golden)[0] for _ in range(R)]
    syns = [m.synth_generate(words, 1.1, SEED, ndocs=max(cnt, 0), pages_per_doc=20, min_page_chars=200, first_doc=first)
            if cnt else None for m, (first, cnt) in zip(members, ranges)]
    bases = dcd.coord_bases([s.text_bytes if s else 0 for s in syns])
    comms = dcd.ShardComm.local(R)

    def member(r):
        def go():
            b = members[r].GetBuilder()
            b.SetOrigin(bases[r], ranges[r][0])
            s = syns[r]
            if s is not None:
                b.AddPagesBulk((s.d_text, s.d_page_offs, s.d_doc_first_page, s.npages, s.ndocs), dcd.ENC_LATIN1, None, None,
                               source="doc", on_device=True)
            b.BuildSharded(comms[r])
        return go

    run_members([member(r) for r in range(R)])
    return members, syns, comms, ora


def check_merged(members, ora):
    parts = [m.serialize_merged() for m in members]
    lists = [m.serialize()[1] for m in members]
    merged = b"".join(parts)
    want = ora.index_bytes()
    assert len(merged) == len(want), (len(merged), len(want), [len(p) for p in parts])
    assert merged == want
    assert b"".join(lists) == ora.list_bytes()
    ks = [m.counters for m in members]
    assert sum(k.postings for k in ks) == ora.total_postings
    assert len({int(k.codes) for k in ks}) == 1                      # one global dictionary
    assert [int(k.merged_code_begin) for k in ks][1:] == [int(k.merged_code_end) for k in ks][:-1]
    assert int(ks[0].merged_code_begin) == 0 and int(ks[-1].merged_code_end) == int(ks[0].codes)
    return ks


@pytest.mark.parametrize("R", [2, 3])
def test_sharded_build_equals_single_builder(R, data_dir, golden, stop_words, en_voc):
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:20000]
    members, syns, comms, ora = sharded_synthetic(R, 240, words, stop_words, en_voc, golden)
    ks = check_merged(members, ora)
    assert all(int(k.shard_p2p) == 1 for k in ks)   # fused peer-memory scatter
    # every shard answers queries for its document range with global coordinates (the decoder starts each list of a shard
    # from the code's first-delta base); the collective exact search (halo exchange, tests/test_gpu_sharded_exact.py) makes
    # the concatenation in rank order equal to the whole-corpus answer on EVERY page, the documents next to a cut included.
    w = [x for x in words if x not in set(stop_words)]
    shard_reqs = ((f"{w[20]} {w[31]}", 40), (f"\"{w[2]} {w[3]}\"", None), (f"{w[25]} ({w[36]} | {w[47]})", 100), (w[4], None),
                  (f"{w[0]} {w[1]}", 12))
    res = [None] * R

    def searcher(r):
        def go():
            res[r] = members[r].SearchBatch([q for q, _ in shard_reqs], [None if d is None else dcd.SearchOptions(dist=d) for _, d in shard_reqs],
                                            comm=comms[r])
        return go

    run_members([searcher(r) for r in range(R)])
    total_pages = 0
    for qi, (req, dist) in enumerate(shard_reqs):
        want = ora.search(req, dist)
        got, base = [], 0
        for r, m in enumerate(members):
            assert res[r][qi].Success
            got += [(p.ordinal + base, p.id, tuple(p.pos)) for p in res[r][qi].foundPages]
            base += int(m.counters.pages)
        exp = [(p["page_index"], p["id"], tuple(p["pos"])) for p in want.foundPages]
        assert got == exp, req
        assert any(o >= int(members[0].counters.pages) for o, _, _ in got), req   # the later shards contribute
        total_pages += len(got)
    assert total_pages > 1000
    # flat batch export of every shard (the payload of the sharded gather) == the per-request results
    reqs = [f"{w[20]} {w[31]}", f"\"{w[2]} {w[3]}\"", w[4]]
    base = 0
    for m in members:
        cnts, pairs = m.SearchBatchPages(reqs, dcd.SearchOptions(dist=40), page_base=base)
        rs = m.SearchBatch(reqs, dcd.SearchOptions(dist=40))
        assert cnts.tolist() == [len(r.foundPages) for r in rs]
        assert pairs.tolist() == [[p.ordinal + base, len(p.pos)] for r in rs for p in r.foundPages]
        # pages kept on the device (no host shaping): same counts, same pairs
        import torch
        dc, dp = m.SearchBatchPages(reqs, dcd.SearchOptions(dist=40), page_base=base,
                                    device_out=lambda n: torch.empty(n, dtype=torch.int32, device="cuda"))
        assert dc.tolist() == cnts.tolist() and dp.is_cuda and dp.cpu().tolist() == pairs.tolist()
        base += int(m.counters.pages)
    # the merged files load like any Docodo index (Index.Load, Index.cs:312-380) and answer the same requests exactly as the
    # single-builder oracle does -- pages, positions, ranks, documents -- including runs that straddle a shard cut
    from test_gpu_parity import assert_same_result
    merged_index = b"".join(m.serialize_merged() for m in members)
    merged_list = b"".join(m.serialize()[1] for m in members)
    whole = configure("envoc", stop_words, en_voc, golden)[0]
    whole.LoadBytes(merged_index, merged_list)
    for req, dist in shard_reqs + (("time people", 40), ("\"new york\"", None), ("government (state | world)", 255), ("years", None),
                                   ("world war", 255), ("?ork", None)):
        assert_same_result(whole.Search(req, None if dist is None else dcd.SearchOptions(dist=dist)), ora.search(req, dist), req)
    whole.Dispose()
    for m, s, cm in zip(members, syns, comms):
        if s is not None:
            m.synth_free(s)
        m.Dispose()
        cm.destroy()


def test_sharded_build_all_to_all_fallback(monkeypatch, data_dir, golden, stop_words, en_voc):
    """DCD_SHARD_P2P=0 selects the all-to-all + owner-concat path (used when peer mappings are unavailable)."""
    monkeypatch.setenv("DCD_SHARD_P2P", "0")
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:5000]
    members, syns, comms, ora = sharded_synthetic(3, 120, words, stop_words, en_voc, golden)
    check_merged(members, ora)
    assert all(int(m.counters.shard_p2p) == 0 for m in members)
    for m, s, cm in zip(members, syns, comms):
        m.synth_free(s)
        m.Dispose()
        cm.destroy()


@pytest.mark.parametrize("pipeline,ce_min", [("1", "64"), ("0", "64"), ("0", "0"), ("1", "0")])
def test_sharded_build_copy_engine_and_pipeline_variants(monkeypatch, pipeline, ce_min, data_dir, golden, stop_words, en_voc):
    """The exchange has two switches: DCD_SHARD_PIPELINE (lengths pass + encode in parts, each part scattered while the next
    is encoded) and DCD_SHARD_CE_MIN (segments of at least that many ushorts travel by cudaMemcpyAsync instead of the scatter
    kernel; 64 makes the hot lists of this small corpus take that road).  Every combination gives the same bytes."""
    monkeypatch.setenv("DCD_SHARD_CE_MIN", ce_min)
    monkeypatch.setenv("DCD_SHARD_PIPELINE", pipeline)
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:5000]
    members, syns, comms, ora = sharded_synthetic(3, 150, words, stop_words, en_voc, golden, ranges=[(0, 60), (60, 20), (80, 70)])
    check_merged(members, ora)
    for m, s, cm in zip(members, syns, comms):
        m.synth_free(s)
        m.Dispose()
        cm.destroy()


def test_sharded_build_uneven_and_empty_shard(data_dir, golden, stop_words, en_voc):
    """Ragged partition: one rank holds most documents, one holds none (its segments and dictionary are empty)."""
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:5000]
    ranges = [(0, 7), (7, 0), (7, 90), (97, 3)]
    members, syns, comms, ora = sharded_synthetic(4, 100, words, stop_words, en_voc, golden, ranges=ranges)
    check_merged(members, ora)
    assert members[1].counters.postings == 0
    for m, s, cm in zip(members, syns, comms):
        if s is not None:
            m.synth_free(s)
        m.Dispose()
        cm.destroy()


def test_sharded_build_single_rank_is_plain_build(data_dir, golden, stop_words, en_voc):
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:5000]
    members, syns, comms, ora = sharded_synthetic(1, 40, words, stop_words, en_voc, golden)
    check_merged(members, ora)
    ib, lb = members[0].serialize()
    assert ib == ora.index_bytes() and lb == ora.list_bytes()
    members[0].synth_free(syns[0])
    members[0].Dispose()
    comms[0].destroy()


def test_sharded_build_explicit_documents(golden, stop_words, en_voc):
    """The IndexTest documents (header pages, host-side postings) split over two shards."""
    docs = indextest_source(golden, 6)
    ix_all, ora = configure("nostem_stop", stop_words, en_voc, golden)
    feed(ora, docs)
    ora.finish()
    half = len(docs) // 2
    # coordinate base of shard 1 = end coordinate of shard 0's last page (exact text length of shard 0)
    ix_all.Create(docs[:half])
    npg = int(ix_all.counters.pages)
    base1 = int(ix_all.page_info(npg - 1)[3]) if npg else 0
    ix_all.Dispose()
    members = [configure("nostem_stop", stop_words, en_voc, golden)[0] for _ in range(2)]
    comms = dcd.ShardComm.local(2)
    parts = [docs[:half], docs[half:]]
    bases = [0, base1]

    def member(r):
        def go():
            b = members[r].GetBuilder()
            b.SetOrigin(bases[r], 0)
            for source, name, pages in parts[r]:
                b.AddDoc(source, name)
                for pid, text in pages:
                    b.AddPage(pid, text)
            b.BuildSharded(comms[r])
        return go

    run_members([member(0), member(1)])
    check_merged(members, ora)
    for m, cm in zip(members, comms):
        m.Dispose()
        cm.destroy()


def test_pages_on_device_edge_cases(golden, stop_words, en_voc):
    """dcd_search_batch_pages / dcd_search_pages_fetch: unbuilt index, syntax errors, empty results, host destination,
    too small a buffer, fetch without a pending batch."""
    import ctypes as C

    import torch
    from docodo_b200.capi import check, lib
    L = lib()
    ix = configure("envoc", stop_words, en_voc, golden)[0]
    cnt, pairs = ix.SearchBatchPages(["old lady"], None, device_out=lambda n: torch.empty(n, dtype=torch.int32, device="cuda"))
    assert cnt.tolist() == [0] and pairs.shape[0] == 0           # "Index is not built": no pages
    assert L.dcd_search_pages_fetch(ix._ctx, None, 0, 1) != 0    # nothing pending any more
    ix.Create(indextest_source(golden, 6))
    reqs = ["old lady", "zzzzqqq", "(((", "lady", "old (lady | man)"]
    rs = ix.SearchBatch(reqs, dcd.SearchOptions(dist=40))
    cnt, pairs = ix.SearchBatchPages(reqs, dcd.SearchOptions(dist=40), page_base=7,
                                     device_out=lambda n: torch.empty(n, dtype=torch.int32, device="cuda"))
    assert cnt.tolist() == [len(r.foundPages) if r.Success else 0 for r in rs] and cnt[0] > 0 and cnt[1] == 0
    assert pairs.cpu().tolist() == [[p.ordinal + 7, len(p.pos)] for r in rs if r.Success for p in r.foundPages]
    # host destination and capacity check
    chars, offs = dcd.index._pack(reqs)
    has = np.ones(len(reqs), dtype=np.int32)
    dd = np.full(len(reqs), 40, dtype=np.int32)
    counts = np.zeros(len(reqs), dtype=np.int32)
    total = C.c_uint64()
    check(ix._ctx, L.dcd_search_batch_pages(ix._ctx, dcd.index._ptr(chars), dcd.index._ptr(offs), len(reqs), dcd.index._ptr(has),
                                            dcd.index._ptr(dd), 7, dcd.index._ptr(counts), C.byref(total)))
    host = np.zeros((total.value, 2), dtype=np.int32)
    assert L.dcd_search_pages_fetch(ix._ctx, dcd.index._ptr(host), total.value - 1, 0) != 0   # too small: refused, batch still pending
    check(ix._ctx, L.dcd_search_pages_fetch(ix._ctx, dcd.index._ptr(host), total.value, 0))
    assert host.tolist() == pairs.cpu().tolist()
    ix.Dispose()
