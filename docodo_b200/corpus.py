"""Synthetic workload definitions of BASELINE.json configs 3-5: the ranked vocabulary fed to the
corpus generator (device: csrc/synth.cu; CPU twin for the baseline: oracle/oracle_synth.cpp) and the mixed query batch.
Benchmark/test utility: nothing here is part of the Docodo API.

Vocabulary ranking (rank 1 = most frequent under Zipf(s)):
  1. the English stop words of Dict/stop.txt (they are the most frequent words of real English text; the index build
     drops them through the stop list, Build.cs:185),
  2. the distinct surface words of Dict/en/*.txt (tests/data/en_words.txt) in a seeded shuffle,
  3. an out-of-vocabulary tail of pseudo-words (3-12 letters, consonant/vowel alternation) that exercises the "$stem"
     branch of GetWordCodes (Build.cs:206-238).
"""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.path.join(ROOT, "tests", "data")

SEED_C3 = 0xD0C0D00003
SEED_C4 = 0xD0C0D00004
SEED_C5 = 0xD0C0D00005


def read_stop_words(path=None):
    """Index.LoadStopWords (Index.cs:227-230): non-blank lines without ';'."""
    path = path or os.path.join(DATA, "stop.txt")
    out = []
    for s in open(path, encoding="utf-8-sig").read().splitlines():
        if len(s.strip(" ")) > 0 and ";" not in s:
            out.append(s)
    return out


def ranked_vocabulary(n_oov=100_000, seed=SEED_C3):
    stops = [w for w in read_stop_words() if w.isascii() and w.isalpha()]
    base = open(os.path.join(DATA, "en_words.txt")).read().split()
    sset = set(stops)
    base = [w for w in base if w not in sset]
    rng = np.random.RandomState(seed & 0x7FFFFFFF)
    order = rng.permutation(len(base))
    words = stops + [base[i] for i in order]
    if n_oov:
        cons, vow = "bcdfghjklmnprstvwz", "aeiou"
        lens = rng.randint(3, 13, size=n_oov)
        ci = rng.randint(0, len(cons), size=(n_oov, 12))
        vi = rng.randint(0, len(vow), size=(n_oov, 12))
        start = rng.randint(0, 2, size=n_oov)
        seen = set(words)
        for k in range(n_oov):
            w = "".join(cons[ci[k, j]] if (j + start[k]) % 2 == 0 else vow[vi[k, j]] for j in range(lens[k]))
            if w not in seen:
                seen.add(w)
                words.append(w)
    return words


def make_queries(words, nq, seed=SEED_C5, zipf_s=1.1, sample_text=None, min_rank=50):
    """Config-5 mix: 40 % two-term AND (dist 255 == opt null), 20 % three-term AND (dist 40), 25 % quoted phrase of 2-3
    adjacent corpus words (from sample_text, so that hits exist), 10 % near-k two-term (k in 10/40/100), 5 % a (b|c).
    Ranks ~ Zipf(s); at least one term per query has rank >= min_rank.  Returns (requests, dists) with dist None == null opt."""
    rng = np.random.RandomState(seed & 0x7FFFFFFF)
    n = len(words)
    p = np.arange(1, n + 1, dtype=np.float64) ** -zipf_s
    cdf = np.cumsum(p / p.sum())
    stops = set(read_stop_words())

    def draw(rare=False):
        while True:
            r = int(np.searchsorted(cdf, rng.random_sample()))
            r = min(r, n - 1)
            if rare and r < min_rank:
                continue
            w = words[r]
            if len(w) >= 3 and w not in stops:
                return w

    toks = None
    if sample_text is not None:
        import re
        toks = [t for t in re.findall(r"[a-z]+|[0-9]+", bytes(sample_text).decode("latin-1").lower())]
    reqs, dists = [], []
    for _ in range(nq):
        u = rng.random_sample()
        if u < 0.40:
            reqs.append("%s %s" % (draw(True), draw()))
            dists.append(None)
        elif u < 0.60:
            reqs.append("%s %s %s" % (draw(True), draw(), draw()))
            dists.append(40)
        elif u < 0.85 and toks:
            k = 2 + int(rng.randint(0, 2))
            for _try in range(50):
                i = int(rng.randint(0, len(toks) - k))
                ph = toks[i:i + k]
                if all(len(t) >= 3 and t not in stops and t.isalpha() for t in ph):
                    break
            reqs.append('"%s"' % " ".join(ph))
            dists.append(40)
        elif u < 0.95:
            reqs.append("%s %s" % (draw(True), draw()))
            dists.append(int(rng.choice([10, 40, 100])))
        else:
            reqs.append("%s (%s|%s)" % (draw(True), draw(), draw()))
            dists.append(None)
    return reqs, dists
