"""R-shard build through the in-process loopback group on ONE GPU (every kernel of the sharded path, no NCCL), sized for
profiling: `ncu --metrics gpu__time_duration.sum` over this script gives the per-kernel times of a sharded build, which ncu
cannot take from a multi-rank run.  Usage: python tools/loopback_profile.py [R] [docs_per_shard] [repeats]"""
import os
import sys
import threading

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import docodo_b200 as dcd   # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 4
DOCS = int(sys.argv[2]) if len(sys.argv) > 2 else 250000
REPS = int(sys.argv[3]) if len(sys.argv) > 3 else 2
from docodo_b200.corpus import SEED_C3, ranked_vocabulary, read_stop_words   # noqa: E402

ROOT = os.path.join(os.path.dirname(__file__), "..")
words = ranked_vocabulary()
members = []
for _ in range(R):
    ix = dcd.Index()
    ix.AddStopWords(read_stop_words())
    ix.AddVoc(dcd.Vocab(data=open(os.path.join(ROOT, "tests", "data", "en.voc"), "rb").read(), stemmer=dcd.STEM_EN))
    members.append(ix)
ranges = dcd.shard_ranges(DOCS * R, R)
syns = [m.synth_generate(words, 1.1, SEED_C3, ndocs=cnt, pages_per_doc=20, min_page_chars=200, first_doc=first)
        for m, (first, cnt) in zip(members, ranges)]
bases = dcd.coord_bases([s.text_bytes for s in syns])
comms = dcd.ShardComm.local(R)
print("text bytes per shard:", [s.text_bytes for s in syns], flush=True)


def member(r):
    b = members[r].GetBuilder()
    b.SetOrigin(bases[r], ranges[r][0])
    s = syns[r]
    b.AddPagesBulk((s.d_text, s.d_page_offs, s.d_doc_first_page, s.npages, s.ndocs), dcd.ENC_LATIN1, None, None, source="doc",
                   on_device=True)
    b.BuildSharded(comms[r])


for rep in range(REPS):
    th = [threading.Thread(target=member, args=(r,)) for r in range(R)]
    [t.start() for t in th]
    [t.join() for t in th]
    k = members[0].counters
    print(rep, "rank0 build ms", round(k.build_ms_device, 2), "stages", [round(x, 2) for x in k.stage_ms][:7], "shard",
          [round(x, 2) for x in k.shard_ms], "postings", int(k.postings), "codes", int(k.codes), flush=True)
