#!/usr/bin/env python
"""Parity of the NCCL sharded build on N GPUs of one box (launched by torch.distributed.run, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/nccl_shard_check.py [--docs 400]

Every rank generates and indexes its own document range of the synthetic corpus on its GPU; the merged `.index` slices and the
`.index.list` parts are gathered on rank 0 and byte-compared with the CPU oracle's single-builder index of the whole corpus.
Also checks a few queries answered shard by shard (hit pages concatenated in rank order).  Prints NCCL_SHARD_CHECK_OK."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--docs", type=int, default=400)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist

    import docodo_b200 as dcd
    from conftest import read_stop_words
    from docodo_b200.sharding import broadcast_unique_id, exchange_origins, gather_hit_pages
    from oracle_py import STEM_EN, OracleIndex

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    data = os.path.join(ROOT, "tests", "data")
    stop = read_stop_words(os.path.join(data, "stop.txt"))
    voc = open(os.path.join(data, "en.voc"), "rb").read()
    words = open(os.path.join(data, "en_words.txt")).read().split()[:20000]
    seed = 0xD0C0D00004

    ix = dcd.Index(device=local)
    ix.AddStopWords(stop)
    ix.AddVoc(dcd.Vocab(data=voc, stemmer=dcd.STEM_EN))
    first, cnt = dcd.shard_ranges(args.docs, world)[rank]
    syn = ix.synth_generate(words, 1.1, seed, ndocs=cnt, pages_per_doc=20, min_page_chars=200, first_doc=first)
    coord_base, doc_base, total_units, total_docs = exchange_origins(dist, syn.text_bytes, cnt)
    assert doc_base == first and total_docs == args.docs
    comm = dcd.ShardComm.nccl(ix, broadcast_unique_id(dist, dcd.ShardComm.unique_id), rank, world)
    for _ in range(2):   # twice: buffers of the first build are recycled by the second
        b = ix.GetBuilder()
        b.SetOrigin(coord_base, doc_base)
        b.AddPagesBulk((syn.d_text, syn.d_page_offs, syn.d_doc_first_page, syn.npages, syn.ndocs), dcd.ENC_LATIN1, None, None,
                       source="doc", on_device=True)
        b.BuildSharded(comm)
    k = ix.counters
    part = ix.serialize_merged()
    lst = ix.serialize()[1]
    parts = [None] * world
    dist.all_gather_object(parts, (part, lst, int(k.postings), float(k.shard_ms[0]), float(k.shard_ms[1]), float(k.shard_ms[2]),
                                   int(k.exchange_bytes)))
    # frequent words: every shard contributes hundreds of pages (the shard decoder starts each list from its first-delta base)
    w = [x for x in words if x not in set(stop)]
    reqs = [(f"{w[20]} {w[31]}", 40), (f"\"{w[2]} {w[3]}\"", None), (f"{w[25]} ({w[36]} | {w[47]})", 100), (w[4], None)]
    # EXACT sharded serving (dcd_search_batch_sharded: halo exchange over the same NCCL communicator): every page is compared,
    # the documents next to a shard cut included -- with dist 100 / 255 the chains of these frequent words cross every cut
    reqs += [(f"{w[0]} {w[1]} {w[2]}", 255), (f"{w[0]} {w[1]}", None)]
    opts = [None if d is None else dcd.SearchOptions(dist=d) for _, d in reqs]
    per_q = []
    for r in ix.SearchBatch([r for r, _ in reqs], opts, comm=comm):
        rows = [[p.ordinal + 0, int(p.id), len(p.pos), int(p.pos[0]), int(p.pos[-1])] for p in r.foundPages]
        per_q.append(np.array(rows, dtype=np.int64).reshape(len(rows), 5))
    # page ordinals are local to the shard: make them global by the page count of the ranks before
    npg = torch.tensor([int(k.pages)], dtype=torch.int64, device="cuda")
    allp = [torch.zeros_like(npg) for _ in range(world)]
    dist.all_gather(allp, npg)
    page_base = sum(int(t.item()) for t in allp[:rank])
    for a in per_q:
        if len(a):
            a[:, 0] += page_base
    merged_hits = gather_hit_pages(dist, per_q)
    # batch form: flat int32 pairs exported into page-locked staging, merged on the device after two NCCL all-gathers
    from docodo_b200.sharding import gather_pages_flat, pinned_buffer
    cnts, pairs = ix.SearchBatchPages([r for r, _ in reqs], opts, page_base=page_base, out=lambda n: pinned_buffer("pairs", n, torch.int32),
                                      comm=comm)
    merged_flat = gather_pages_flat(dist, cnts, pairs)
    # the same with the pages kept on the device from the query kernel to the NCCL gather
    dcnts, dpairs = ix.SearchBatchPages([r for r, _ in reqs], opts, page_base=page_base,
                                        device_out=lambda n: torch.empty(n, dtype=torch.int32, device="cuda"), comm=comm)
    merged_dev = gather_pages_flat(dist, dcnts, dpairs)
    if rank == 0:
        assert merged_dev[0].tolist() == merged_flat[0].tolist() and merged_dev[1].tolist() == merged_flat[1].tolist()
    if rank == 0:
        ora = OracleIndex()
        ora.set_stopwords(stop)
        ora.add_voc_file(voc, STEM_EN)
        from oracle_py import synth_corpus
        text, po, dfp = synth_corpus(words, 1.1, seed, args.docs, 20, 200, 0, nthreads=4)
        ora.build_bulk(text, 1, po, dfp, nthreads=1)
        merged = b"".join(p[0] for p in parts)
        want = ora.index_bytes()
        assert len(merged) == len(want), (len(merged), len(want))
        assert merged == want, "merged .index differs from the single-builder oracle"
        assert b"".join(p[1] for p in parts) == ora.list_bytes(), ".index.list differs"
        assert sum(p[2] for p in parts) == ora.total_postings
        for (req, d), got in zip(reqs, merged_hits):
            w = ora.search(req, d)
            wrows = [[p["page_index"], int(p["id"]), len(p["pos"]), int(p["pos"][0]), int(p["pos"][-1])] for p in w.foundPages]
            assert got.tolist() == wrows, req
            assert len(wrows) > 50 and wrows[-1][0] // 20 >= dcd.shard_ranges(args.docs, world)[-1][0], req
        mc, mp = merged_flat
        at = 0
        for qi, (req, d) in enumerate(reqs):
            w = ora.search(req, d)
            wrows = [[p["page_index"], len(p["pos"])] for p in w.foundPages]
            assert mp[at:at + mc[qi]].tolist() == wrows, req
            at += mc[qi]
        assert at == len(mp)
        print("shard stage ms per rank (unify, bases, exchange+concat):", [[round(x, 3) for x in p[3:6]] for p in parts],
              "exchange bytes:", [p[6] for p in parts])
        print("NCCL_SHARD_CHECK_OK world=%d docs=%d index_bytes=%d" % (world, args.docs, len(want)), flush=True)
    dist.barrier()
    comm.destroy()
    ix.synth_free(syn)
    ix.Dispose()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
