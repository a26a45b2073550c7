import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import json
import numpy as np
import docodo_b200 as dcd
from docodo_b200.corpus import SEED_C3, ranked_vocabulary, make_queries, read_stop_words
from test_gpu_sharded import run_members
from oracle_py import OracleIndex, STEM_EN, synth_corpus

R = int(sys.argv[1]) if len(sys.argv) > 1 else 2
vd = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
nq = int(sys.argv[3]) if len(sys.argv) > 3 else 256
words = ranked_vocabulary()
stop = read_stop_words()
voc = open(os.path.join(ROOT, "tests", "data", "en.voc"), "rb").read()
def mk():
    ix = dcd.Index(); ix.AddStopWords(stop); ix.AddVoc(dcd.Vocab(data=voc, stemmer=dcd.STEM_EN)); return ix
members = [mk() for _ in range(R)]
syns = [m.synth_generate(words, 1.1, SEED_C3, vd, 20, 200, first_doc=r * vd) for r, m in enumerate(members)]
bases = dcd.coord_bases([s.text_bytes for s in syns])
comms = dcd.ShardComm.local(R)
def member(r):
    def go():
        b = members[r].GetBuilder(); b.SetOrigin(bases[r], r * vd); s = syns[r]
        b.AddPagesBulk((s.d_text, s.d_page_offs, s.d_doc_first_page, s.npages, s.ndocs), dcd.ENC_LATIN1, None, None, source="doc", on_device=True)
        b.BuildSharded(comms[r])
    return go
run_members([member(r) for r in range(R)])
text, po, dfp = synth_corpus(words, 1.1, SEED_C3, vd * R, 20, 200, 0, nthreads=8)
ora = OracleIndex(); ora.set_stopwords(stop); ora.add_voc_file(voc, STEM_EN)
ora.build_bulk(text, 1, po, dfp, nthreads=8)
stext = synth_corpus(words, 1.1, SEED_C3, 2000, 20, 200, 0, nthreads=1)[0]
reqs, dists = make_queries(words, nq, sample_text=stext[:4 << 20])
opts = [None if d is None else dcd.SearchOptions(dist=d) for d in dists]
res = [None] * R
def searcher(r):
    def go():
        res[r] = members[r].SearchBatch(reqs, opts, comm=comms[r])
    return go
run_members([searcher(r) for r in range(R)])
bad = 0
for qi, (rq, d) in enumerate(zip(reqs, dists)):
    got, base = [], 0
    for r, m in enumerate(members):
        got += [(p.ordinal + base, tuple(p.pos)) for p in res[r][qi].foundPages]
        base += int(m.counters.pages)
    w = ora.search(rq, d)
    exp = [(p["page_index"], tuple(p["pos"])) for p in w.foundPages]
    if got != exp:
        bad += 1
        loc, base = [], 0
        for r, m in enumerate(members):
            rr = m.Search(rq, opts[qi])
            loc += [(p.ordinal + base, tuple(p.pos)) for p in rr.foundPages]
            base += int(m.counters.pages)
        print("BAD", repr(rq), d, "got", len(got), "exp", len(exp), "local", len(loc), "local==exp", loc == exp)
        print("   got-exp", sorted(set(got) - set(exp))[:5], "exp-got", sorted(set(exp) - set(got))[:5])
        print("   codes", [(t, members[0].GetWordCodes(t)) for t in rq.replace('"', ' ').replace('(', ' ').replace(')', ' ').replace('|', ' ').split()])
print("checked", len(reqs), "bad", bad)
