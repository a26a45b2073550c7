"""Host plumbing of the sharded build / sharded query serving (docodo_b200/sharding.py) over torch.distributed with the
gloo backend, world_size 2 and 3, on CPU (the device side of the same path is tests/test_gpu_sharded.py and
tools/nccl_shard_check.py)."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from docodo_b200 import coord_bases, shard_ranges
    from docodo_b200.sharding import broadcast_unique_id, exchange_origins, gather_hit_pages
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ndocs_total = 1001
        ranges = shard_ranges(ndocs_total, world)
        first, cnt = ranges[rank]
        lens = [1000 * c + 7 * r for r, (_, c) in enumerate(ranges)]          # pretend text lengths
        cb, db, tot_units, tot_docs = exchange_origins(dist, lens[rank], cnt)
        assert cb == coord_bases(lens)[rank] and db == first
        assert tot_units == sum(lens) and tot_docs == ndocs_total
        uid = broadcast_unique_id(dist, lambda: bytes((i * 7 + 3) % 256 for i in range(128)))
        assert uid == bytes((i * 7 + 3) % 256 for i in range(128))
        # hit pages: query q has (rank + q) % 3 pages on this rank; payload = (page ordinal, position)
        nq = 5
        pages = []
        for qi in range(nq):
            n = (rank + qi) % 3
            pages.append(np.array([[1000 * rank + 10 * qi + k, k] for k in range(n)], dtype=np.int64).reshape(n, 2))
        merged = gather_hit_pages(dist, pages, max_found_pages=3)
        if rank == 0:
            for qi in range(nq):
                want = [[1000 * r + 10 * qi + k, k] for r in range(world) for k in range((r + qi) % 3)][:4]
                assert merged[qi].tolist() == want, (qi, merged[qi].tolist(), want)
        else:
            assert merged is None
        from docodo_b200.sharding import gather_pages_flat
        counts = np.array([len(a) for a in pages])
        flat = np.concatenate(pages + [np.zeros((0, 2), np.int64)], axis=0)
        res = gather_pages_flat(dist, counts, flat, max_found_pages=3)
        if rank == 0:
            mc, mp = res
            at = 0
            for qi in range(nq):
                full = [[1000 * r + 10 * qi + k, k] for r in range(world) for k in range((r + qi) % 3)]
                want = full[:4]
                if len(full) > 3:
                    want[3][1] = 1   # page max_found_pages + 1 of the merged list keeps one position (Search.cs:409)
                assert mc[qi] == len(want) and mp[at:at + mc[qi]].tolist() == want, (qi, mc[qi], want)
                at += mc[qi]
            assert at == len(mp)
        else:
            assert res is None
        # the cut after page max_found_pages + 1 (here 2) applies to the MERGED list; payload given as a torch tensor
        import torch
        res = gather_pages_flat(dist, counts, torch.from_numpy(flat.astype(np.int32)), max_found_pages=1)
        if rank == 0:
            mc, mp = res
            at = 0
            for qi in range(nq):
                full = [[1000 * r + 10 * qi + k, k] for r in range(world) for k in range((r + qi) % 3)]
                want = full[:2]
                if len(full) > 1:
                    want[1][1] = 1
                assert mc[qi] == len(want) and mp[at:at + mc[qi]].tolist() == want, (qi, mc[qi], want)
                at += mc[qi]
            assert at == len(mp) and mp.dtype == np.int32
        q.put((rank, "ok"))
    except BaseException as e:   # noqa: BLE001
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharding_plumbing_gloo(world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(60)
    assert sorted(res) == [(r, "ok") for r in range(world)], res


def test_shard_ranges_and_bases():
    sys.path.insert(0, ROOT)
    from docodo_b200 import coord_bases, shard_ranges
    assert shard_ranges(10, 3) == [(0, 4), (4, 3), (7, 3)]
    assert shard_ranges(2, 4) == [(0, 1), (1, 1), (2, 0), (2, 0)]
    assert shard_ranges(0, 2) == [(0, 0), (0, 0)]
    assert coord_bases([5, 0, 7]) == [0, 5, 5]
