#!/usr/bin/env python
"""bench.py -- Docodo index-build throughput (GB/s of text) + proximity queries/s on B200.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run, one rank per GPU)
    python bench.py --impl reference ...                     (the CPU path -- the oracle port of Docodo's C# code -- on host cores)

A step = one full index build (tokenize -> word codes -> sort -> delta/varint encode -> page table) of the BASELINE.json
config-3 corpus (synthetic Zipf(1.1) English-like text, 1 M documents x 20 pages, ~4 GiB) that is already resident in HBM;
at N GPUs every rank builds its own document-range shard of the same size with global coordinates (weak scaling, config 4).
`e2e` runs the same build through the C ABI with HOST buffers (pinned text in, encoded index out) every step.
One JSON line is printed by rank 0.  See DESIGN.md section "Measurement" for the byte model behind `roofline`.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "index_build_text_GBps"
UNIT = "GB/s"
ZIPF_S = 1.1
PAGES_PER_DOC = 20
MIN_PAGE_CHARS = 200


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--docs", type=int, default=1_000_000, help="documents per GPU (config 3: 1 M x 20 pages ~ 4 GiB)")
    ap.add_argument("--queries", type=int, default=8192, help="size of the mixed query batch (secondary metric)")
    ap.add_argument("--query-docs", type=int, default=0, help="documents of the query index (0: same index as the build)")
    ap.add_argument("--config5-queries", type=int, default=1_000_000,
                    help="BASELINE config 5: requests evaluated against a 10 M-page index (500 K documents x 20 pages) split over the "
                         "GPUs of the run (strong scaling); 0 skips it")
    ap.add_argument("--config5-docs", type=int, default=500_000, help="documents of the config-5 index (total over all GPUs)")
    ap.add_argument("--config5-chunk", type=int, default=65536, help="requests per dcd_search_batch call of the config-5 stream")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-docs", type=int, default=0, help="documents of the CPU-baseline sample (0: auto, ~10-30 s)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-queries", action="store_true")
    ap.add_argument("--no-verify", action="store_true", help="skip the parity check of the timed build against the CPU oracle")
    ap.add_argument("--verify-docs", type=int, default=50_000,
                    help="N > 1: documents per GPU of the sharded verification build (real NCCL + IPC transport) checked against the oracle")
    ap.add_argument("--verify-queries", type=int, default=128, help="requests of the query batch re-evaluated by the oracle")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with nvidia-smi while the timed region runs."""

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device = device
        self.stop_flag = False
        self.sm = []
        self.reasons = set()
        self.sm_max = None

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_throttle_reasons.hw_slowdown,clocks_throttle_reasons.hw_thermal_slowdown," \
            "clocks_throttle_reasons.sw_thermal_slowdown,clocks_throttle_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                out = [x.strip() for x in out]
                self.sm.append(float(out[0]))
                self.sm_max = float(out[1])
                for nm, v in zip(names, out[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.05)

    def result(self):
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def measured_peaks():
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(pk["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6532.2, "fallback of /opt/skills/guides/B200_PROFILING.md (MEASURED_PEAKS.json absent)"


def bind_to_gpu_numa(device):
    """Pin this process to the host cores next to its GPU (NVML CPU affinity) BEFORE the page-locked buffers are allocated, so
    that they land in the GPU's NUMA node: at N = 8 every rank moves 6.3 GB per build over its own PCIe link and the host
    memory of the wrong socket is the shared bottleneck.  Returns the cores, or None when NVML does not say."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device)
        ncpu = os.cpu_count() or 64
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return sorted(cpus)
    except Exception:
        pass
    return None


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's C# build path (oracle/, test infrastructure) on the host cores
# ---------------------------------------------------------------------------------------------------------------
def make_oracle():
    from docodo_b200.corpus import read_stop_words
    from oracle_py import STEM_EN, OracleIndex
    ora = OracleIndex()
    ora.set_stopwords(read_stop_words())
    ora.add_voc_file(open(os.path.join(ROOT, "tests", "data", "en.voc"), "rb").read(), STEM_EN)
    return ora


def cpu_build_once(words, seed, ndocs, cores, first_doc=0):
    """Index `ndocs` documents of the synthetic corpus with the oracle on `cores` threads (MaxDegreeOfParallelism = cores:
    contiguous document ranges + MergeIndexes-style stitch).  Returns (seconds, text_bytes)."""
    from oracle_py import synth_corpus
    text, po, dfp = synth_corpus(words, ZIPF_S, seed, ndocs, PAGES_PER_DOC, MIN_PAGE_CHARS, first_doc, nthreads=cores)
    ora = make_oracle()
    sec = ora.build_bulk(text, 1, po, dfp, nthreads=cores)
    nbytes = int(len(text))
    del ora
    return sec, nbytes


def cpu_sample_docs(words, seed, cores, target_s=12.0):
    sec, nb = cpu_build_once(words, seed, 2000 * cores, cores)
    rate = 2000 * cores / max(sec, 1e-3)
    return int(max(2000 * cores, min(rate * target_s, 4_000_000)))


def run_reference(args, words, seed):
    """--impl reference: Docodo's own CPU algorithm for the path.  The C# sources cannot be compiled on this image (no .NET),
    so this is the oracle port (kind 'port'), all host threads, each step a bounded sample of the config-3 workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    ndocs = args.cpu_docs or cpu_sample_docs(words, seed, cores, target_s=8.0)
    times, nbytes = [], 0
    for i in range(args.warmup + args.steps):
        sec, nbytes = cpu_build_once(words, seed, ndocs, cores)
        if i >= args.warmup:
            times.append(sec)
    t = float(np.mean(times))
    val = nbytes / t / 1e9
    sample = "first %d documents (%d pages, %.1f MB) of the config-3 corpus per step" % (ndocs, ndocs * PAGES_PER_DOC, nbytes / 1e6)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u16/u64", "data": "synthetic",
            "config": {"workload": "config3-sample: Zipf(1.1) English-like corpus, %d docs x %d pages, en.voc + stop list"
                                   % (ndocs, PAGES_PER_DOC), "l2": "n/a (CPU)"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------
# --verify: the build that is timed (and the batch that is timed) against the CPU oracle -- checker only, outside the
# timed regions.  N = 1: the whole config-3 corpus is indexed by the oracle on all host threads and the sha256 of its
# `.index` / `.index.list` bytes is compared with dcd_index_serialize of the GPU build.  N > 1: a sharded build of
# --verify-docs documents per GPU through the SAME communicator (real NCCL + CUDA-IPC peer memory) is compared slice by slice
# (dcd_merged_serialize of every rank, in rank order) with the oracle's single-builder index of the same corpus.
# ---------------------------------------------------------------------------------------------------------------
def sha256_np(a):
    import hashlib
    return hashlib.sha256(memoryview(np.ascontiguousarray(a)).cast("B")).hexdigest() if a.size else hashlib.sha256(b"").hexdigest()


def gpu_serialized(ix, L, check, merged=False):
    """(index bytes ndarray, list bytes ndarray) of the resident index (merged=True: this rank's slice of the merged .index)."""
    ni, nl = C.c_uint64(), C.c_uint64()
    check(ix._ctx, L.dcd_index_serialized_size(ix._ctx, C.byref(ni), C.byref(nl)))
    lb = np.empty(max(nl.value, 1), dtype=np.uint8)
    if merged:
        nm = C.c_uint64()
        check(ix._ctx, L.dcd_merged_serialized_size(ix._ctx, C.byref(nm)))
        ib = np.empty(max(nm.value, 1), dtype=np.uint8)
        check(ix._ctx, L.dcd_merged_serialize(ix._ctx, ib.ctypes.data_as(C.c_void_p), nm.value))
        tmp = np.empty(max(ni.value, 1), dtype=np.uint8)
        check(ix._ctx, L.dcd_index_serialize(ix._ctx, tmp.ctypes.data_as(C.c_void_p), ni.value, lb.ctypes.data_as(C.c_void_p), nl.value))
        return ib[:nm.value], lb[:nl.value]
    ib = np.empty(max(ni.value, 1), dtype=np.uint8)
    check(ix._ctx, L.dcd_index_serialize(ix._ctx, ib.ctypes.data_as(C.c_void_p), ni.value, lb.ctypes.data_as(C.c_void_p), nl.value))
    return ib[:ni.value], lb[:nl.value]


def same_result(got, want, rtol=1e-6):
    """Index.SearchBatchFlat entry vs the oracle's SearchResult: found pages (ordinal, positions), page ranks, found documents
    (ordinal order as ranked) and document ranks (1e-6 relative, the north star's floating-point tolerance)."""
    if got["status"] != 0 or not want.success:
        return got["status"] != 0 and not want.success
    wp = want.foundPages
    if len(wp) != len(got["ordinals"]):
        return False
    if len(wp):
        if not np.array_equal(got["ordinals"], np.array([p["page_index"] for p in wp], dtype=np.int64)):
            return False
        if not np.array_equal(got["npos"], np.array([len(p["pos"]) for p in wp], dtype=np.int32)):
            return False
        if not np.array_equal(got["pos"], np.concatenate([np.asarray(p["pos"], dtype=np.int32) for p in wp])):
            return False
        wr = np.array([p["rank"] for p in wp], dtype=np.float64)
        if np.any(np.abs(got["ranks"].astype(np.float64) - wr) > rtol * np.abs(wr)):
            return False
    if [d[0] for d in got["docs"]] != [d["doc"] for d in want.foundDocs]:
        return False
    return all(abs(a[1] - b["rank"]) <= rtol * abs(b["rank"]) for a, b in zip(got["docs"], want.foundDocs))


# ---------------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------------
def stage_bytes(k, text_bytes):
    """Algorithmic bytes per stage (DESIGN.md 'Measurement'): T text, N tokens x 4 B records, P postings x 8 B
    keys, d digit passes, O encoded ushort bytes (+ per-code offsets), G page table."""
    T, N, P, d = float(text_bytes), float(k.tokens), float(k.postings), int(k.sort_passes)
    O = 2.0 * k.encoded_u16 + 8.0 * (k.codes + 1)
    G = 8.0 * k.pages
    R = 4.0   # bytes per token record of the bulk-page tokenizer (wordSlot:22 | offset:10)
    st = {"tokenize": T + R * N, "expand": R * N + 8 * P, "onesweep_sort": 2.0 * d * 8 * P, "encode": 8 * P + O, "pages": 2 * G}
    return st, sum(st.values())


STAGE_NAMES = ["tokenize", "word_codes_rank", "expand", "onesweep_sort", "encode", "pages", "host_key_table"]
STAGE_KERNEL = {"tokenize": "k_tokenize2<1>", "expand": "k_expand2", "onesweep_sort": "k_onesweep_wide", "encode": "k_encode"}


def ncu_traffic(kernel, text_bytes):
    """dram__bytes_read.sum + dram__bytes_write.sum of `kernel` per launch from the committed `ncu --set full` capture of this
    workload (profiles/ncu_traffic.json: kernel -> {text_bytes, dram_bytes}); null when no capture matches the workload size."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[kernel]
        if abs(t["text_bytes"] - text_bytes) <= 0.01 * text_bytes:
            return float(t["dram_bytes"])
    except Exception:
        pass
    return None


def run_b200(args, words, seed):
    import torch
    import torch.distributed as dist

    import docodo_b200 as dcd
    from docodo_b200.capi import check, lib
    from docodo_b200.corpus import make_queries, read_stop_words
    from docodo_b200.sharding import gather_pages_flat  # noqa: F401

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- docodo_b200 has no CPU fallback (use --impl reference for the CPU path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_cores = bind_to_gpu_numa(local) if world > 1 and not os.environ.get("DCD_NO_NUMA_BIND") else None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allsum(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    L = lib()
    st = torch.cuda.Stream(device=dev)
    ix = dcd.Index(device=local)
    ix.set_stream(st.cuda_stream)
    ix.AddStopWords(read_stop_words())
    ix.AddVoc(dcd.Vocab(data=open(os.path.join(ROOT, "tests", "data", "en.voc"), "rb").read(), stemmer=dcd.STEM_EN))

    ndocs = args.docs
    syn = ix.synth_generate(words, ZIPF_S, seed, ndocs, PAGES_PER_DOC, MIN_PAGE_CHARS, first_doc=rank * ndocs)
    T = int(syn.text_bytes)
    # global coordinates: this shard starts where the previous ranks' text ends (exact text lengths)
    lens = torch.zeros(world, dtype=torch.int64, device=dev)
    lens[rank] = T
    if world > 1:
        dist.all_reduce(lens)
    coord_base = int(lens[:rank].sum().item())

    # N > 1: the collective sharded build (dictionary unification, first-delta stitch, NCCL all-to-all of the encoded posting
    # segments to the rank owning each code range, owner concat) -- BASELINE config 4 at config-3 shard size per GPU
    comm = None
    if world > 1:
        from docodo_b200.sharding import broadcast_unique_id
        comm = dcd.ShardComm.nccl(ix, broadcast_unique_id(dist, dcd.ShardComm.unique_id), rank, world)

    def finish(b):
        if comm is not None:
            b.BuildSharded(comm)
        else:
            b.Build()

    def build_device():
        b = ix.GetBuilder()
        b.SetOrigin(coord_base, rank * ndocs)
        b.AddPagesBulk((syn.d_text, syn.d_page_offs, syn.d_doc_first_page, syn.npages, syn.ndocs), dcd.ENC_LATIN1, None, None,
                       source="doc", on_device=True)
        finish(b)
        return ix.counters

    # ---- --verify: the exact build that is timed below, against the CPU oracle (outside the timed region) --------------------
    verify, ora_v = None, None
    if not args.no_verify:
        tv0 = time.perf_counter()
        cores = host_cores()
        if world == 1:
            build_device()
            gi, gl = gpu_serialized(ix, L, check)
            g = (int(gi.size), sha256_np(gi), int(gl.size), sha256_np(gl))
            del gi, gl
            text = ix.d2h(syn.d_text, T)
            po = ix.d2h(syn.d_page_offs, 8 * (int(syn.npages) + 1), np.uint64)
            dfp = ix.d2h(syn.d_doc_first_page, 8 * (int(syn.ndocs) + 1), np.uint64)
            ora_v = make_oracle()
            osec = ora_v.build_bulk(text, 1, po, dfp, nthreads=cores)
            del text
            ni, di, nl, dl = ora_v.digests()
            verify = {"match": g == (ni, di[0], nl, dl[0]) and int(ix.counters.postings) == int(ora_v.total_postings),
                      "what": "sha256 of dcd_index_serialize (.index, .index.list) of the timed config-3 build == the CPU oracle's "
                              "single-builder files of the same %d documents" % ndocs,
                      "index_bytes": g[0], "index_sha": g[1][:16], "list_bytes": g[2], "list_sha": g[3][:16],
                      "oracle_index_bytes": ni, "oracle_index_sha": di[0][:16], "oracle_list_sha": dl[0][:16],
                      "postings": int(ix.counters.postings), "oracle_postings": int(ora_v.total_postings),
                      "oracle_build_s": round(osec, 1), "oracle_threads": cores}
        else:
            # a sharded build of --verify-docs documents per GPU through the SAME communicator as the timed build (real NCCL +
            # CUDA-IPC peer-memory scatter); rank r's merged slice must be the r-th slice of the oracle's single-builder `.index`
            vd = args.verify_docs
            syn_v = ix.synth_generate(words, ZIPF_S, seed, vd, PAGES_PER_DOC, MIN_PAGE_CHARS, first_doc=rank * vd)
            lens_v = torch.zeros(world, dtype=torch.int64, device=dev)
            lens_v[rank] = int(syn_v.text_bytes)
            dist.all_reduce(lens_v)
            b = ix.GetBuilder()
            b.SetOrigin(int(lens_v[:rank].sum().item()), rank * vd)
            b.AddPagesBulk((syn_v.d_text, syn_v.d_page_offs, syn_v.d_doc_first_page, syn_v.npages, syn_v.ndocs), dcd.ENC_LATIN1, None,
                           None, source="doc", on_device=True)
            b.BuildSharded(comm)
            kv = ix.counters
            gi, gl = gpu_serialized(ix, L, check, merged=True)
            mine = (int(gi.size), sha256_np(gi), int(gl.size), sha256_np(gl), int(kv.postings), int(kv.shard_p2p))
            del gi, gl
            ix.synth_free(syn_v)
            # queries on the verification shards: exact sharded search + NCCL gather, compared with the oracle below
            vq = None
            if args.verify_queries > 0:
                from docodo_b200.sharding import gather_pages_flat
                from oracle_py import synth_corpus
                stext = synth_corpus(words, ZIPF_S, seed, 2000, PAGES_PER_DOC, MIN_PAGE_CHARS, 0, nthreads=1)[0]
                vreqs, vdists = make_queries(words, args.verify_queries, sample_text=stext[:4 << 20])
                pgs = torch.zeros(world, dtype=torch.int64, device=dev)
                pgs[rank] = int(kv.pages)
                dist.all_reduce(pgs)
                vcnt, vpairs = ix.SearchBatchPages(vreqs, [None if d is None else dcd.SearchOptions(dist=d) for d in vdists],
                                                   page_base=int(pgs[:rank].sum().item()),
                                                   device_out=lambda n: torch.empty(n, dtype=torch.int32, device=dev), comm=comm)
                vq = gather_pages_flat(dist, vcnt, vpairs)
                if vq is not None:
                    vq = (np.array(vq[0]).copy(), np.array(vq[1]).copy())
            parts = [None] * world
            dist.all_gather_object(parts, mine)
            if rank == 0:
                from oracle_py import synth_corpus
                text, po, dfp = synth_corpus(words, ZIPF_S, seed, vd * world, PAGES_PER_DOC, MIN_PAGE_CHARS, 0, nthreads=cores)
                ora = make_oracle()
                osec = ora.build_bulk(text, 1, po, dfp, nthreads=cores)
                del text
                ci = np.concatenate([[0], np.cumsum([p[0] for p in parts])]).tolist()
                cl = np.concatenate([[0], np.cumsum([p[2] for p in parts])]).tolist()
                ni, di, nl, dl = ora.digests(ci, cl)
                ok = (ni == ci[-1] and nl == cl[-1] and di == [p[1] for p in parts] and dl == [p[3] for p in parts]
                      and sum(p[4] for p in parts) == int(ora.total_postings))
                verify = {"match": bool(ok),
                          "what": "per-rank sha256 of dcd_merged_serialize / .index.list of a %d-rank sharded build (%d documents per "
                                  "GPU, same NCCL communicator and peer-memory transport as the timed build) == the same slices of the "
                                  "CPU oracle's single-builder files" % (world, vd),
                          "transport_p2p": [p[5] for p in parts], "index_bytes": ci[-1], "oracle_index_bytes": ni,
                          "slice_sha": [p[1][:12] for p in parts], "oracle_slice_sha": [d[:12] for d in di],
                          "list_bytes": cl[-1], "oracle_list_bytes": nl, "postings": sum(p[4] for p in parts),
                          "oracle_postings": int(ora.total_postings), "oracle_build_s": round(osec, 1), "oracle_threads": cores}
                if vq is not None:
                    mc, mp = vq
                    bad, at, pages_cmp = [], 0, 0
                    for qi, (rq, d_) in enumerate(zip(vreqs, vdists)):
                        w_ = ora.search(rq, d_)
                        exp = [[p["page_index"], len(p["pos"])] for p in w_.foundPages]
                        pages_cmp += len(exp)
                        if mp[at:at + int(mc[qi])].tolist() != exp:
                            bad.append(rq)
                        at += int(mc[qi])
                    verify.update({"queries_checked": len(vreqs), "queries_match": not bad, "query_pages_compared": pages_cmp,
                                   "queries_what": "dcd_search_batch_pages_sharded (halo exchange) + NCCL gather on the verification "
                                                   "shards == the oracle's Index.Search on the whole corpus: every page ordinal and "
                                                   "position count, documents next to the shard cuts included"})
                    if bad:
                        verify["match"] = False
                        verify["queries_bad"] = bad[:5]
                del ora
            dist.barrier()
        if verify is not None:
            verify["seconds"] = round(time.perf_counter() - tv0, 1)
            if not verify["match"]:
                print("bench.py: VERIFY FAILED: " + json.dumps(verify), file=sys.stderr, flush=True)

    stage_ms = np.zeros(8)
    shard_ms = np.zeros(4)
    launches = 0
    for _ in range(args.warmup):
        build_device()
    torch.cuda.synchronize()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(st)
    dev_ms = 0.0
    for _ in range(args.steps):
        k = build_device()
        stage_ms += np.array(list(k.stage_ms))
        shard_ms += np.array(list(k.shard_ms))
        dev_ms += k.build_ms_device
        launches += int(k.kernel_launches)
    ev1.record(st)
    torch.cuda.synchronize()
    barrier()
    sampler.stop_flag = True
    ms_total = allmax(ev0.elapsed_time(ev1))
    ms_step = ms_total / args.steps
    total_T = allsum(T)
    value = total_T / (ms_step * 1e-3) / 1e9
    k = ix.counters
    stage_ms /= args.steps
    sbytes, b_build = stage_bytes(k, T)
    peak, peak_src = measured_peaks()
    timed = {n: stage_ms[i] for i, n in enumerate(STAGE_NAMES) if n in sbytes and n in STAGE_KERNEL}
    dom = max(timed, key=timed.get)
    nl = k.sort_passes if dom == "onesweep_sort" else 1
    ach = sbytes[dom] / (timed[dom] * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": STAGE_KERNEL[dom], "launches_per_step": int(nl),
                "ms_per_launch": timed[dom] / nl, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": ncu_traffic(STAGE_KERNEL[dom], T), "algorithmic_bytes_per_launch": sbytes[dom] / nl, "peak_source": peak_src,
                "whole_build": {"algorithmic_bytes": b_build, "achieved": b_build / (ms_step * 1e-3) / 1e9,
                                "frac": b_build / (ms_step * 1e-3) / 1e9 / peak},
                "stages_ms": {n: round(float(stage_ms[i]), 3) for i, n in enumerate(STAGE_NAMES)},
                "stages_GBps": {n: round(sbytes[n] / (stage_ms[STAGE_NAMES.index(n)] * 1e-3) / 1e9, 1) for n in sbytes
                                if stage_ms[STAGE_NAMES.index(n)] > 0}}
    counters = {"text_bytes": T, "tokens": int(k.tokens), "postings": int(k.postings), "codes": int(k.codes),
                "distinct_words": int(k.distinct_words), "encoded_u16": int(k.encoded_u16), "pages": int(k.pages),
                "sort_passes": int(k.sort_passes), "device_ms_per_step": dev_ms / args.steps,
                "host_finish_wall_ms": float(k.build_ms_wall), "host_add_pages_ms": float(k.stage_ms[7])}

    # ---- e2e: the same build through the C ABI with HOST buffers (pinned text in, encoded index out) ------------------
    e2e = None
    if not args.no_e2e:
        h_text = torch.empty(T, dtype=torch.uint8, pin_memory=True)
        h_po = torch.empty(int(syn.npages) + 1, dtype=torch.int64, pin_memory=True)
        h_dfp = torch.empty(int(syn.ndocs) + 1, dtype=torch.int64, pin_memory=True)
        check(ix._ctx, L.dcd_memcpy_d2h(ix._ctx, h_text.data_ptr(), syn.d_text, T))
        check(ix._ctx, L.dcd_memcpy_d2h(ix._ctx, h_po.data_ptr(), syn.d_page_offs, 8 * (int(syn.npages) + 1)))
        check(ix._ctx, L.dcd_memcpy_d2h(ix._ctx, h_dfp.data_ptr(), syn.d_doc_first_page, 8 * (int(syn.ndocs) + 1)))
        o_arena = torch.empty(int(max(k.encoded_u16, k.merged_u16) * (1.25 if world > 1 else 1)) + 16, dtype=torch.int16, pin_memory=True)
        o_offs = torch.empty(int(k.codes) + 1, dtype=torch.int64, pin_memory=True)
        o_pages = torch.empty(int(k.pages) + 1, dtype=torch.int64, pin_memory=True)
        src = dcd.index._u16("doc")

        # Two builds in flight (world == 1): a second context with its own stream and output buffers runs build k + 1 while
        # build k's encoded index is still on its way back, so the two directions of the bus work at the same time -- what a
        # host that indexes batch after batch does.  N > 1 keeps one collective build at a time (one communicator).
        pipe = 2 if world == 1 else 1
        ctxs = [(ix, o_arena, o_offs, o_pages)]
        if pipe == 2:
            ixb = dcd.Index(device=local)
            ixb.AddStopWords(read_stop_words())
            ixb.AddVoc(dcd.Vocab(data=open(os.path.join(ROOT, "tests", "data", "en.voc"), "rb").read(), stemmer=dcd.STEM_EN))
            ctxs.append((ixb, torch.empty_like(o_arena).pin_memory(), torch.empty_like(o_offs).pin_memory(),
                         torch.empty_like(o_pages).pin_memory()))

        e2e_trace = [] if os.environ.get("DCD_E2E_TRACE") else None
        # one build at a time uploads and computes; the other context is then downloading its index (device -> host).  (Measured:
        # letting the next upload start as soon as the text has arrived -- before the build's last 20 ms of compute -- is slower
        # on this box, 114 vs 101 ms per build: the two directions of the bus do not run at full rate together.)
        h2d_turn = threading.Lock()

        def build_host(slot=0):
            cx, oa, oo, op_ = ctxs[slot]
            h2d_turn.acquire()
            tt = [time.perf_counter()]
            b = cx.GetBuilder()
            b.SetOrigin(coord_base, rank * ndocs)
            check(cx._ctx, L.dcd_build_add_pages_bulk(b._b, h_text.data_ptr(), dcd.ENC_LATIN1, h_po.data_ptr(), int(syn.npages),
                                                      h_dfp.data_ptr(), int(syn.ndocs), dcd.index._ptr(src), len(src), 0))
            tt.append(time.perf_counter())
            if comm is not None:   # the product of the sharded build: this rank's code range of the merged index + its page list
                b.BuildSharded(comm)
                h2d_turn.release()
                check(cx._ctx, L.dcd_merged_download_raw(cx._ctx, oa.data_ptr(), oa.numel(), oo.data_ptr(), oo.numel()))
                check(cx._ctx, L.dcd_index_download_raw(cx._ctx, None, 0, None, op_.data_ptr()))
            else:
                b.Build()
                h2d_turn.release()
                tt.append(time.perf_counter())
                check(cx._ctx, L.dcd_index_download_raw(cx._ctx, oa.data_ptr(), oa.numel(), oo.data_ptr(), op_.data_ptr()))
                tt.append(time.perf_counter())
                if e2e_trace is not None:
                    e2e_trace.append((slot, [round((x - tt[0]) * 1e3, 1) for x in tt], round(tt[0] * 1e3 % 100000, 1)))

        for sl in range(pipe):
            build_host(sl)   # warm-up (allocator, page-locked mappings)
        torch.cuda.synchronize()
        barrier()
        nsteps = args.e2e_steps * pipe
        t0 = time.perf_counter()
        if pipe == 1:
            for _ in range(nsteps):
                build_host(0)
        else:
            def worker(sl):
                for _ in range(args.e2e_steps):
                    build_host(sl)
            th = [threading.Thread(target=worker, args=(sl,)) for sl in range(pipe)]
            for t_ in th:
                t_.start()
            for t_ in th:
                t_.join()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        barrier()
        e_ms = allmax(wall) / nsteps
        if e2e_trace:
            print("e2e trace (slot, [add, build, download] ms from start, start):", e2e_trace, file=sys.stderr)
        if pipe == 2:
            assert torch.equal(ctxs[0][1][:int(k.encoded_u16)], ctxs[1][1][:int(k.encoded_u16)])   # both contexts built the same index
            ctxs[1][0].Dispose()
        k2 = ix.counters
        e2e = {"value": total_T / (e_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": e_ms, "steps": nsteps, "builds_in_flight": pipe,
               "host_cores_bound": len(numa_cores) if numa_cores else None,
               "h2d_bytes_per_step": T + 8 * (int(syn.npages) + 1) + 8 * (int(syn.ndocs) + 1),
               "d2h_bytes_per_step": (2 * int(k2.merged_u16) + 8 * (int(k2.merged_code_end) - int(k2.merged_code_begin) + 1)
                                      if comm is not None else 2 * int(k2.encoded_u16) + 8 * (int(k2.codes) + 1)) + 8 * int(k2.pages),
               "api": "dcd_build_begin/add_pages_bulk(host)/finish + dcd_index_download_raw"}
        del h_text, o_arena

    # ---- secondary metric: batch of mixed AND / phrase / near-k queries against the index just built ----------------------
    queries = None
    if not args.no_queries and args.queries > 0:
        nsample = min(T, 4 << 20)
        sample = np.zeros(nsample, dtype=np.uint8)
        check(ix._ctx, L.dcd_memcpy_d2h(ix._ctx, sample.ctypes.data_as(C.c_void_p), syn.d_text, nsample))
        if world > 1:   # one batch for every rank (collective search): the phrases are sampled from rank 0's text
            ns = torch.tensor([nsample], dtype=torch.int64, device=dev)
            dist.broadcast(ns, 0)
            tsample = torch.zeros(int(ns.item()), dtype=torch.uint8, device=dev)
            if rank == 0:
                tsample.copy_(torch.from_numpy(sample))
            dist.broadcast(tsample, 0)
            sample = tsample.cpu().numpy()
        reqs, dists = make_queries(words, args.queries, sample_text=sample)
        chars, offs = dcd.index._pack(reqs)
        has = np.array([0 if d is None else 1 for d in dists], dtype=np.int32)
        dd = np.array([0 if d is None else d for d in dists], dtype=np.int32)

        # sharded serving (N > 1): the per-shard page lists of every request are gathered on rank 0 over NCCL and concatenated
        # in rank order (disjoint ascending document ranges), MAX_FOUND_PAGES applied after the merge -- inside the timed region
        page_base = 0
        if world > 1:
            from docodo_b200.sharding import gather_pages_flat, pinned_buffer
            pg = torch.zeros(world, dtype=torch.int64, device=dev)
            pg[rank] = int(k.pages)
            dist.all_reduce(pg)
            page_base = int(pg[:rank].sum().item())
        gather_info = {}

        def run_queries():
            h = C.c_void_p()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            if world > 1:
                # sharded serving: the found pages stay on the device from the query kernel to the NCCL gather (no host shaping)
                total = C.c_uint64()
                counts = np.zeros(len(reqs), dtype=np.int32)
                # EXACT sharded serving: halo exchange between the shards over the build's communicator (dcd_search_batch_pages_sharded)
                check(ix._ctx, L.dcd_search_batch_pages_sharded(ix._ctx, comm._h, dcd.index._ptr(chars), dcd.index._ptr(offs), len(reqs),
                                                                dcd.index._ptr(has), dcd.index._ptr(dd), page_base,
                                                                dcd.index._ptr(counts), C.byref(total)))
                t1 = time.perf_counter()
                pairs = torch.empty(2 * max(total.value, 1), dtype=torch.int32, device=dev)
                torch.cuda.current_stream().synchronize()
                check(ix._ctx, L.dcd_search_pages_fetch(ix._ctx, pairs.data_ptr(), total.value, 1))
                t2 = time.perf_counter()
                trace = {} if os.environ.get("DCD_GATHER_TRACE") else None
                res = gather_pages_flat(dist, counts, pairs[:2 * total.value].view(-1, 2), trace=trace)
                if rank == 0:
                    gather_info["merged_pages"] = int(res[0].sum())
                    gather_info["gather_ms"] = (time.perf_counter() - t1) * 1e3
                    gather_info["export_ms"] = (t2 - t1) * 1e3
                    if trace:
                        gather_info["trace"] = {k_: round(v_, 2) for k_, v_ in trace.items()}
            else:
                check(ix._ctx, L.dcd_search_batch(ix._ctx, dcd.index._ptr(chars), dcd.index._ptr(offs), len(reqs), dcd.index._ptr(has),
                                                  dcd.index._ptr(dd), C.byref(h)))
                t1 = time.perf_counter()
            dt = time.perf_counter() - t0
            if world > 1 and os.environ.get("DCD_GATHER_TRACE"):
                walls = [None] * world
                dist.all_gather_object(walls, round((t1 - t0) * 1e3, 1))
                gather_info["search_wall_ms_per_rank"] = walls
            npages = 0
            if h.value:
                npages = sum(L.dcd_hits_num_pages(h, q) for q in range(0, len(reqs), max(1, len(reqs) // 256)))
                L.dcd_hits_free(h)
            return dt, npages

        cold_dt, _ = run_queries()   # first batch on this index: also fills the decoded-list cache (hot lists decoded once)
        qt = []
        for _ in range(3):
            dt, _np = run_queries()
            qt.append(dt)
        kq = ix.counters
        wall = allmax(float(np.median(qt)))
        qdev = allmax(kq.search_ms_device * 1e-3)
        queries = {"metric": "proximity_queries_per_s", "batch": len(reqs), "index_pages_per_gpu": int(k.pages),
                   "value": len(reqs) / wall, "unit": "queries/s", "cold_value": len(reqs) / allmax(cold_dt),
                   "cold_note": "first batch after the build: includes decoding the batch's long lists into the per-index decoded-list cache",
                   "note": ("wall time of dcd_search_batch incl. host planning, device waves, hit download and result shaping"
                            if world == 1 else
                            "every rank evaluates the whole batch on its shard (SURVEY.md 8e) with dcd_search_batch_pages_sharded (exact: "
                            "halo exchange at the shard cuts; found pages stay on the device as int32 pairs), then the NCCL gather of "
                            "the per-shard page lists on rank 0, merged in rank order on the device and copied to the host; wall time "
                            "of all of it"),
                   "index_pages_total": int(k.pages) * world, "gather": gather_info or None,
                   "device_only_qps": len(reqs) / qdev if qdev > 0 else None, "device_ms": kq.search_ms_device,
                   "hits": int(kq.search_hits), "algorithmic_GBps": kq.search_bytes / max(qdev, 1e-9) / 1e9,
                   "frac_hbm": kq.search_bytes / max(qdev, 1e-9) / 1e9 / peak,
                   "mix": "40% 2-term AND d255, 20% 3-term AND d40, 25% quoted phrase, 10% near-k, 5% a (b|c)"}

    # ---- --verify, queries: a fixed sample of the timed batch re-evaluated by the oracle on its index of the same corpus ---------
    if ora_v is not None and queries is not None and args.verify_queries > 0:
        tq0 = time.perf_counter()
        step = max(1, len(reqs) // args.verify_queries)
        idx = list(range(0, len(reqs), step))[:args.verify_queries]
        got = ix.SearchBatchFlat([reqs[i] for i in idx], [None if dists[i] is None else dcd.SearchOptions(dist=dists[i]) for i in idx])
        bad, pages_cmp = [], 0
        for i, g_ in zip(idx, got):
            w_ = ora_v.search(reqs[i], dists[i])
            pages_cmp += len(w_.foundPages)
            if not same_result(g_, w_):
                bad.append(reqs[i])
        verify["queries_checked"] = len(idx)
        verify["queries_match"] = not bad
        verify["query_pages_compared"] = pages_cmp
        verify["query_seconds"] = round(time.perf_counter() - tq0, 1)
        if bad:
            verify["match"] = False
            verify["queries_bad"] = bad[:5]
            print("bench.py: QUERY VERIFY FAILED: " + json.dumps(bad[:5]), file=sys.stderr, flush=True)
    ora_v = None

    # ---- BASELINE config 5: 1 M mixed requests against a 10 M-page index split over the GPUs of the run (strong scaling) -------
    config5 = None
    if not args.no_queries and args.config5_queries > 0:
        from docodo_b200.corpus import SEED_C5
        docs5 = args.config5_docs // world
        syn5 = ix.synth_generate(words, ZIPF_S, SEED_C5, docs5, PAGES_PER_DOC, MIN_PAGE_CHARS, first_doc=rank * docs5)
        lens5 = torch.zeros(world, dtype=torch.int64, device=dev)
        lens5[rank] = int(syn5.text_bytes)
        if world > 1:
            dist.all_reduce(lens5)
        b5 = ix.GetBuilder()
        b5.SetOrigin(int(lens5[:rank].sum().item()), rank * docs5)
        b5.AddPagesBulk((syn5.d_text, syn5.d_page_offs, syn5.d_doc_first_page, syn5.npages, syn5.ndocs), dcd.ENC_LATIN1, None, None,
                        source="doc", on_device=True)
        finish(b5)
        k5 = ix.counters
        ns5 = min(int(syn5.text_bytes), 4 << 20)
        sample5 = np.zeros(ns5, dtype=np.uint8)
        check(ix._ctx, L.dcd_memcpy_d2h(ix._ctx, sample5.ctypes.data_as(C.c_void_p), syn5.d_text, ns5))
        if world > 1:
            t5 = torch.zeros(4 << 20, dtype=torch.uint8, device=dev)
            if rank == 0:
                t5[:ns5].copy_(torch.from_numpy(sample5))
            dist.broadcast(t5, 0)
            sample5 = t5.cpu().numpy()
        reqs5, dists5 = make_queries(words, args.config5_queries, sample_text=sample5)
        pbase5 = 0
        if world > 1:
            pg5 = torch.zeros(world, dtype=torch.int64, device=dev)
            pg5[rank] = int(k5.pages)
            dist.all_reduce(pg5)
            pbase5 = int(pg5[:rank].sum().item())

        def run5(limit=None):
            """The request stream in chunks of --config5-chunk requests per call; returns (seconds, pages found, device ms)."""
            n5 = len(reqs5) if limit is None else min(limit, len(reqs5))
            pages, dev_ms = 0, 0.0
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            for a in range(0, n5, args.config5_chunk):
                rq = reqs5[a:a + args.config5_chunk]
                ds = dists5[a:a + args.config5_chunk]
                chars5, offs5 = dcd.index._pack(rq)
                has5 = np.array([0 if d is None else 1 for d in ds], dtype=np.int32)
                dd5 = np.array([0 if d is None else d for d in ds], dtype=np.int32)
                if world > 1:
                    total = C.c_uint64()
                    counts = np.zeros(len(rq), dtype=np.int32)
                    check(ix._ctx, L.dcd_search_batch_pages_sharded(ix._ctx, comm._h, dcd.index._ptr(chars5), dcd.index._ptr(offs5), len(rq),
                                                                    dcd.index._ptr(has5), dcd.index._ptr(dd5), pbase5,
                                                                    dcd.index._ptr(counts), C.byref(total)))
                    pairs = torch.empty(2 * max(total.value, 1), dtype=torch.int32, device=dev)
                    torch.cuda.current_stream().synchronize()
                    check(ix._ctx, L.dcd_search_pages_fetch(ix._ctx, pairs.data_ptr(), total.value, 1))
                    res = gather_pages_flat(dist, counts, pairs[:2 * total.value].view(-1, 2))
                    if rank == 0:
                        pages += int(res[0].sum())
                else:
                    h = C.c_void_p()
                    check(ix._ctx, L.dcd_search_batch(ix._ctx, dcd.index._ptr(chars5), dcd.index._ptr(offs5), len(rq), dcd.index._ptr(has5),
                                                      dcd.index._ptr(dd5), C.byref(h)))
                    tot = C.c_uint64()
                    check(ix._ctx, L.dcd_hits_export_pages(h, None, None, None, None, 0, C.byref(tot)))
                    pages += int(tot.value)
                    L.dcd_hits_free(h)
                dev_ms += float(ix.counters.search_ms_device)
            return time.perf_counter() - t0, pages, dev_ms

        run5(limit=args.config5_chunk)                      # warm-up: allocator, decoded-list cache of the hottest lists
        sec5, pages5, dev5 = run5()
        sec5 = allmax(sec5)
        config5 = {"metric": "proximity_queries_per_s", "value": len(reqs5) / sec5, "unit": "queries/s", "queries": len(reqs5),
                   "index_pages_total": int(k5.pages) * world if world > 1 else int(k5.pages), "index_docs_total": docs5 * world,
                   "n_gpus": world, "scaling": "strong", "seconds": sec5, "device_only_qps": len(reqs5) / max(allmax(dev5) * 1e-3, 1e-9),
                   "found_pages": pages5, "chunk": args.config5_chunk,
                   "note": ("BASELINE config 5: the whole request stream through dcd_search_batch in chunks (host planning, device "
                            "waves, hit download, result shaping), single index" if world == 1 else
                            "BASELINE config 5: the index is split by document range over the GPUs; every rank evaluates every chunk "
                            "on its shard with dcd_search_batch_pages_sharded (exact: halo exchange at the cuts) and the per-shard "
                            "page lists are gathered on rank 0 over NCCL, MAX_FOUND_PAGES applied after the merge")}
        ix.synth_free(syn5)


    # ---- CPU baseline beside it: the oracle port on the host cores, bounded sample (rank 0, N = 1 only) -------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = host_cores()
        cdocs = args.cpu_docs or cpu_sample_docs(words, seed, cores, target_s=12.0)
        sec, nb = cpu_build_once(words, seed, cdocs, cores)
        cpu = {"value": nb / sec / 1e9, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "first %d documents (%.1f MB) of the same corpus, oracle build_bulk with %d threads, %.1f s"
                         % (cdocs, nb / 1e6, cores, sec)}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/u64",
                "data": "synthetic",
                "config": {"workload": "config3: synthetic Zipf(1.1) English-like corpus, %d docs x %d pages per GPU (%.2f GB text "
                                       "per GPU), en.voc (Porter2 on device) + stop list, single-GPU index build per rank"
                                       % (ndocs, PAGES_PER_DOC, T / 1e9),
                           "parallelism": "doc-range shard per GPU, global coordinates, NCCL posting-segment exchange + owner merge (config 4 "
                                          "at config-3 shard size)" if world > 1 else "1 GPU",
                           "l2": "inputs (%.2f GB text, %.1f GB keys) exceed the 126 MB L2; no flush needed" % (T / 1e9, 8 * k.postings / 1e9)},
                "verify": verify, "clocks": sampler.result(), "e2e": e2e, "gpu_launches": launches, "queries": queries,
                "config5": config5,
                "shard": None if comm is None else {
                    "transport": "nccl + peer-memory scatter over NVLink (CUDA IPC)" if int(k.shard_p2p) else "nccl all-to-all + owner concat",
                    "ranks": world,
                    "stages_ms": {"unify_dictionaries": round(float(shard_ms[0]) / args.steps, 3),
                                  "first_delta_bases": round(float(shard_ms[1]) / args.steps, 3),
                                  "owner_split_plan": round(float(shard_ms[2]) / args.steps, 3),
                                  "segment_exchange_merge": round(float(shard_ms[3]) / args.steps, 3)},
                    "exchange_bytes_rank0": int(k.exchange_bytes), "merged_u16_rank0": int(k.merged_u16),
                    # bytes that arrive over NVLink per build ~ (R - 1) / R of the rank's merged slice, over the exchange stage
                    "nvlink_in_GBps_rank0": round(2.0 * int(k.merged_u16) * (world - 1) / world / 1e9
                                                  / max(float(shard_ms[3]) / args.steps * 1e-3, 1e-9), 1),
                    "nvlink_fabric_GBps": 620.0,   # all-to-all by copy engines on this box (profiles/r02_exchange_variants.md)
                    "note": "rank 0's figures; the step time is the max over ranks of the whole collective build"},
                "roofline": roofline, "cpu_baseline": cpu, "counters": counters}
        print(json.dumps(line), flush=True)
    ix.synth_free(syn)
    if comm is not None:
        comm.destroy()
    ix.Dispose()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    from docodo_b200.corpus import SEED_C3, ranked_vocabulary
    words = ranked_vocabulary()
    if args.impl == "reference":
        run_reference(args, words, SEED_C3)
    else:
        run_b200(args, words, SEED_C3)


if __name__ == "__main__":
    main()
