"""What the NVLink fabric of the box gives an all-to-all of the size the sharded build moves (1.69 GB out of and into every
GPU at 8 GPUs): copy-engine peer copies from one process.  Prints GB/s per GPU and direction.  Usage: python tools/fabric_probe.py [MB per pair]"""
import sys
import time

import torch

n = torch.cuda.device_count()
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 240
src = [torch.empty(mb << 20, dtype=torch.uint8, device=f"cuda:{i}").fill_(i) for i in range(n)]
dst = [[torch.empty(mb << 20, dtype=torch.uint8, device=f"cuda:{j}") for i in range(n)] for j in range(n)]
streams = [[torch.cuda.Stream(device=i) for j in range(n)] for i in range(n)]


def sync():
    for i in range(n):
        torch.cuda.synchronize(i)


def a2a(order):
    sync()
    t0 = time.perf_counter()
    for k in range(1, n):
        for i in range(n):
            j = (i + k) % n if order == "rot" else k if k != i else 0
            if j == i:
                continue
            with torch.cuda.device(i), torch.cuda.stream(streams[i][j]):
                dst[j][i].copy_(src[i], non_blocking=True)
    sync()
    return time.perf_counter() - t0


for order in ("rot", "rot", "rot"):
    t = a2a(order)
    print(f"all-to-all {n} GPUs, {mb} MB per pair, {order}: {t * 1e3:.2f} ms = {mb * (n - 1) / 1024 / t:.0f} GB/s out of (and into) every GPU", flush=True)
# one pair, one direction, and both directions
for both in (False, True):
    sync()
    t0 = time.perf_counter()
    for rep in range(4):
        with torch.cuda.device(0), torch.cuda.stream(streams[0][1]):
            dst[1][0].copy_(src[0], non_blocking=True)
        if both:
            with torch.cuda.device(1), torch.cuda.stream(streams[1][0]):
                dst[0][1].copy_(src[1], non_blocking=True)
    sync()
    t = time.perf_counter() - t0
    print(f"pair 0->1{' and 1->0' if both else ''}: {4 * mb / 1024 / t:.0f} GB/s per direction", flush=True)
