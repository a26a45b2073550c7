"""Diagnostic: (1) queries on an index whose coordinates exceed 2^32 (shard origin), (2) decoded list values of a shard member."""
import os
import sys
import threading

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import docodo_b200 as dcd  # noqa: E402

data = os.path.join(ROOT, "tests", "data")
words = open(os.path.join(data, "en_words.txt")).read().split()[:5000]


def build(origin, ndocs=120, first_doc=0):
    ix = dcd.Index(InMemory=True)
    syn = ix.synth_generate(words, 1.1, 0xD0C0D00003, ndocs=ndocs, pages_per_doc=20, min_page_chars=200, first_doc=first_doc)
    b = ix.GetBuilder()
    b.SetOrigin(origin, first_doc)
    b.AddPagesBulk((syn.d_text, syn.d_page_offs, syn.d_doc_first_page, syn.npages, syn.ndocs), dcd.ENC_LATIN1, None, None,
                   source="doc", on_device=True)
    return ix, b, syn


w = words
reqs = [f"{w[0]} {w[1]}", f"\"{w[2]} {w[3]}\"", w[4], f"{w[5]} ({w[6]} | {w[7]})", f"{w[0]} {w[0]}"]
res = {}
for origin in (0, 5_000_000_000, (1 << 39)):
    ix, b, syn = build(origin)
    b.Build()
    rs = ix.SearchBatch(reqs, dcd.SearchOptions(dist=40))
    res[origin] = [[(p.id, p.ordinal, tuple(p.pos)) for p in r.foundPages] for r in rs]
    print("origin", origin, "pages per request", [len(x) for x in res[origin]], "first values of 'time':", ix[w[0]].values()[:3],
          "max_coord", ix.MaxCoord, flush=True)
    ix.synth_free(syn)
    ix.Dispose()
print("origin-independent results:", res[0] == res[5_000_000_000] == res[1 << 39])

# sharded loopback, R = 2
R = 2
ranges = dcd.shard_ranges(120, R)
mem = [build(0, ndocs=c, first_doc=f) for f, c in ranges]
bases = dcd.coord_bases([m[2].text_bytes for m in mem])
comms = dcd.ShardComm.local(R)


def go(r):
    ix, b, syn = mem[r]
    b.SetOrigin(bases[r], ranges[r][0])


# builders were created with origin 0 above; rebuild them with the right origin
for ix, b, syn in mem:
    del b
mem2 = []
for r, (f, c) in enumerate(ranges):
    ix, _, syn = mem[r]
    b = ix.GetBuilder()
    b.SetOrigin(bases[r], f)
    b.AddPagesBulk((syn.d_text, syn.d_page_offs, syn.d_doc_first_page, syn.npages, syn.ndocs), dcd.ENC_LATIN1, None, None,
                   source="doc", on_device=True)
    mem2.append((ix, b, syn))
th = [threading.Thread(target=lambda r=r: mem2[r][1].BuildSharded(comms[r])) for r in range(R)]
[t.start() for t in th]
[t.join() for t in th]
for r in range(R):
    ix = mem2[r][0]
    v = ix[w[0]].values()
    print("member", r, "coord base", bases[r], "'time' values", v[:3], "...", v[-1:], "page 0 span", ix.page_info(0)[3:], flush=True)
