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


def gather_pages_flat(dist, counts, payload, max_found_pages=30000, dst=0):
    """Vectorised form of gather_hit_pages for whole batches.  counts: [nq] pages per request on this shard; payload:
    integer array [sum(counts), k] with the pages of request 0, 1, ... back to back (e.g. global page ordinal, positions in
    page; int32 halves the traffic).  Two collectives on the rank's backend (NCCL on GPUs): all-gather of the counts,
    all-gather of the padded payload; the merge itself is index arithmetic on the backend's device (no sort): the rows of
    rank r for request q go to  start[q] + sum_{r' < r} counts[r'][q].  On `dst` returns (merged_counts [nq] int64 numpy,
    merged_payload numpy) with each request's pages in rank order (= coordinate order), cut after page
    max_found_pages + 1 (Search.cs:409); None elsewhere."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = _tensor_device(dist)
    counts = np.ascontiguousarray(counts, dtype=np.int64)
    nq = len(counts)
    payload = np.ascontiguousarray(payload)
    if payload.dtype not in (np.int32, np.int64):
        payload = payload.astype(np.int64)
    payload = payload.reshape(int(counts.sum()), -1) if payload.size else payload.reshape(0, 1)
    width = payload.shape[1]
    c_all = torch.zeros(world, nq, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(c_all.view(-1), torch.from_numpy(counts).to(dev)) if dist.get_backend() == "nccl" else \
        dist.all_gather(list(c_all.unbind(0)), torch.from_numpy(counts).to(dev))
    per_rank = c_all.sum(dim=1)
    maxrows = max(int(per_rank.max().item()) if nq else 0, 1)
    mine = torch.zeros(maxrows, width, dtype=torch.from_numpy(payload).dtype, device=dev)
    if len(payload):
        mine[:len(payload)] = torch.from_numpy(payload).to(dev)
    bufs = torch.zeros(world, maxrows, width, dtype=mine.dtype, device=dev)
    if dist.get_backend() == "nccl":
        dist.all_gather_into_tensor(bufs.view(-1), mine.view(-1))
    else:
        dist.all_gather(list(bufs.unbind(0)), mine)
    if rank != dst:
        return None
    mcounts = c_all.sum(dim=0)                                   # [nq]
    mstart = torch.cumsum(mcounts, 0) - mcounts                  # start of every request in the merged list
    before = torch.cumsum(c_all, 0) - c_all                      # [world, nq]: pages of the request on earlier ranks
    total = int(mcounts.sum().item()) if nq else 0
    merged = torch.zeros(total, width, dtype=mine.dtype, device=dev)
    within_all = torch.zeros(total, dtype=torch.int64, device=dev)
    for r in range(world):
        n_r = int(per_rank[r].item())
        if not n_r:
            continue
        src_start = torch.cumsum(c_all[r], 0) - c_all[r]
        shift = torch.repeat_interleave(mstart + before[r] - src_start, c_all[r])      # [n_r]
        dest = shift + torch.arange(n_r, device=dev)
        merged[dest] = bufs[r, :n_r]
        within_all[dest] = dest - torch.repeat_interleave(mstart, c_all[r])
    cap = max_found_pages + 1
    if nq and bool((mcounts > cap).any().item()):
        merged = merged[within_all < cap]
        mcounts = torch.clamp(mcounts, max=cap)
    return mcounts.cpu().numpy(), merged.cpu().numpy()
