"""Host plumbing of the document-range sharded build and of sharded query serving over ``torch.distributed``
(one process per GPU).  Only rendezvous data travels here -- text lengths, the 128-byte NCCL id, per-query
hit pages; the posting segments themselves move inside ``dcd_build_finish_sharded`` over NCCL / NVLink.

Works with any backend: ``nccl`` on the GPU box (tensors on the rank's device), ``gloo`` in the CPU tests.
"""
import numpy as np


def _tensor_device(dist):
    import torch
    if dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def exchange_origins(dist, text_units, ndocs):
    """All-gather (text length, document count) of every shard -> (coord_base, doc_number_base, total_units, total_docs) of
    this rank: the shard's first character sits at the sum of the EXACT text lengths of the ranks before it (the shift
    MergeFiles(shift_coords) applies, Index.cs:584-593, without the maxCoord quirk A.8)."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = _tensor_device(dist)
    mine = torch.tensor([int(text_units), int(ndocs)], dtype=torch.int64, device=dev)
    out = [torch.zeros(2, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(out, mine)
    vals = [[int(v) for v in t.cpu().tolist()] for t in out]
    coord_base = sum(v[0] for v in vals[:rank])
    doc_base = sum(v[1] for v in vals[:rank])
    return coord_base, doc_base, sum(v[0] for v in vals), sum(v[1] for v in vals)


def broadcast_unique_id(dist, make_id):
    """Rank 0 creates the NCCL unique id (``make_id() -> 128 bytes``); every rank returns the same bytes."""
    import torch
    dev = _tensor_device(dist)
    if dist.get_rank() == 0:
        raw = bytes(make_id())
        assert len(raw) == 128
        t = torch.tensor(list(raw), dtype=torch.uint8, device=dev)
    else:
        t = torch.zeros(128, dtype=torch.uint8, device=dev)
    dist.broadcast(t, src=0)
    return bytes(t.cpu().tolist())


def gather_hit_pages(dist, per_query_pages, max_found_pages=30000, dst=0):
    """Merge per-shard hit lists of a query batch on rank `dst`: shards are disjoint ascending document ranges with global
    coordinates, so the merged list of a query is the concatenation in rank order; MAX_FOUND_PAGES is applied after the
    merge (PrepareSearchResult stops once Count > 30000, Search.cs:409).  per_query_pages: list (one entry per query)
    of int64 arrays [n, k] (page ordinal, ...payload).  Returns the merged list on `dst`, None elsewhere."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = _tensor_device(dist)
    nq = len(per_query_pages)
    width = 1
    for a in per_query_pages:
        if len(a):
            width = int(np.asarray(a).reshape(len(a), -1).shape[1])
            break
    counts = torch.tensor([len(a) for a in per_query_pages], dtype=torch.int64, device=dev)
    all_counts = [torch.zeros(nq, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(all_counts, counts)
    all_counts = [t.cpu().numpy() for t in all_counts]
    flat = np.concatenate([np.asarray(a, dtype=np.int64).reshape(len(a), width) for a in per_query_pages] +
                          [np.zeros((0, width), np.int64)], axis=0)
    maxrows = max(int(c.sum()) for c in all_counts)
    pad = np.zeros((maxrows, width), dtype=np.int64)
    pad[:len(flat)] = flat
    mine = torch.from_numpy(pad).to(dev)
    bufs = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(bufs, mine)
    if rank != dst:
        return None
    bufs = [b.cpu().numpy() for b in bufs]
    starts = [np.concatenate([[0], np.cumsum(c)]) for c in all_counts]
    merged = []
    for q in range(nq):
        rows = [bufs[r][starts[r][q]:starts[r][q + 1]] for r in range(world)]
        m = np.concatenate(rows, axis=0)
        if len(m) > max_found_pages + 1:
            m = m[:max_found_pages + 1]
        merged.append(m)
    return merged


_PINNED = {}


def pinned_buffer(name, numel, dtype):
    """Grow-only page-locked staging tensors, kept between batches (allocating page-locked memory costs more than the
    copy it speeds up).  Returns a 1-d tensor of `numel` elements; plain CPU memory when there is no CUDA device."""
    import torch
    t = _PINNED.get((name, dtype))
    if t is None or t.numel() < numel:
        t = torch.empty(max(int(numel) * 5 // 4, 1024), dtype=dtype, pin_memory=torch.cuda.is_available())
        _PINNED[(name, dtype)] = t
    return t[:numel]


def gather_pages_flat(dist, counts, payload, max_found_pages=30000, dst=0, trace=None):
    """Vectorised form of gather_hit_pages for whole batches.  counts: [nq] pages per request on this shard; payload:
    integer array or CPU tensor [sum(counts), k] with the pages of request 0, 1, ... back to back (e.g. global page ordinal,
    positions in page; int32 halves the traffic; a page-locked tensor goes to the device without staging).  Two collectives
    on the rank's backend (NCCL on GPUs): all-gather of the counts, all-gather of the padded payload; the merge itself is
    index arithmetic on the backend's device (no sort): the rows of rank r for request q go to
    start[q] + sum_{r' < r} counts[r'][q].  On `dst` returns (merged_counts [nq] int64 numpy, merged_payload numpy) with each
    request's pages in rank order (= coordinate order), cut after page max_found_pages + 1 (Search.cs:409); None elsewhere.
    On NCCL the merged payload is a view of a page-locked buffer that the next call reuses.  trace: optional dict that
    receives the wall ms of every phase (adds a device synchronize per phase)."""
    import time

    import torch
    t_last = [time.perf_counter()]

    def mark(name):
        if trace is None:
            return
        if dist.get_backend() == "nccl":
            torch.cuda.synchronize()
        now = time.perf_counter()
        trace[name] = trace.get(name, 0.0) + (now - t_last[0]) * 1e3
        t_last[0] = now

    world, rank = dist.get_world_size(), dist.get_rank()
    dev = _tensor_device(dist)
    nccl = dist.get_backend() == "nccl"
    counts = np.ascontiguousarray(counts, dtype=np.int64)
    nq = len(counts)
    n_mine = int(counts.sum())
    if torch.is_tensor(payload):
        tp = payload
    else:
        payload = np.ascontiguousarray(payload)
        if payload.dtype not in (np.int32, np.int64):
            payload = payload.astype(np.int64)
        tp = torch.from_numpy(payload)
    tp = tp.reshape(n_mine, -1) if tp.numel() else tp.reshape(0, 1)
    width = tp.shape[1]
    c_all = torch.zeros(world, nq, dtype=torch.int64, device=dev)
    if nccl:
        dist.all_gather_into_tensor(c_all.view(-1), torch.from_numpy(counts).to(dev))
    else:
        dist.all_gather(list(c_all.unbind(0)), torch.from_numpy(counts).to(dev))
    hc = c_all.cpu().numpy()                                     # [world, nq]: the only host round trip before the merge
    mark("counts_allgather_ms")
    per_rank = hc.sum(axis=1)
    maxrows = max(int(per_rank.max()) if nq else 0, 1)
    bufs = torch.empty(world, maxrows, width, dtype=tp.dtype, device=dev)
    mine = bufs[rank]                                            # in-place all-gather: this rank's slice is its send buffer
    if n_mine:
        mine[:n_mine].copy_(tp, non_blocking=True)
    mark("h2d_ms")
    if nccl:
        dist.all_gather_into_tensor(bufs.view(-1), mine.reshape(-1))
    else:
        dist.all_gather(list(bufs.unbind(0)), mine.clone())
    mark("payload_allgather_ms")
    if rank != dst:
        return None
    mcounts_h = hc.sum(axis=0)                                   # [nq]
    total = int(mcounts_h.sum()) if nq else 0
    mcounts = c_all.sum(dim=0)
    mstart = torch.cumsum(mcounts, 0) - mcounts                  # start of every request in the merged list
    before = torch.cumsum(c_all, 0) - c_all                      # [world, nq]: pages of the request on earlier ranks
    cap = max_found_pages + 1
    over = bool(nq and (mcounts_h > cap).any())
    merged = torch.empty(total, width, dtype=tp.dtype, device=dev)
    within_all = torch.empty(total, dtype=torch.int64, device=dev) if over else None
    for r in range(world):
        n_r = int(per_rank[r])
        if not n_r:
            continue
        src_start = torch.cumsum(c_all[r], 0) - c_all[r]
        shift = torch.repeat_interleave(mstart + before[r] - src_start, c_all[r], output_size=n_r)
        dest = shift + torch.arange(n_r, device=dev)
        merged[dest] = bufs[r, :n_r]
        if over:
            within_all[dest] = dest - torch.repeat_interleave(mstart, c_all[r], output_size=n_r)
    if over:
        if width >= 2:   # page max_found_pages + 1 of the MERGED list keeps one position (PrepareSearchResult breaks right after it, Search.cs:409)
            merged[within_all == cap - 1, 1] = 1
        merged = merged[within_all < cap]
        mcounts_h = np.minimum(mcounts_h, cap)
    mark("merge_ms")
    if nccl:
        out = pinned_buffer("merged", merged.numel(), merged.dtype).view(merged.shape)
        out.copy_(merged, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        mark("d2h_ms")
        return mcounts_h.astype(np.int64), out.numpy()
    return mcounts_h.astype(np.int64), merged.cpu().numpy()
