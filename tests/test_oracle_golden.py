"""Pin the CPU oracle against the reference's own known answers (XUnitDocodoTest/*.cs).
CPU only; no product code is involved."""
import hashlib
import os
import struct

import numpy as np
import pytest

from docsets import feed, indextest_source, text_file_source
from oracle_py import STEM_EN, STEM_NONE, OracleIndex


def build(golden, npages, stop_words=None, stemmers=True, voc_pairs=None, voc_file=None):
    ix = OracleIndex()
    if not stemmers:
        ix.clear_stemmers()
    if voc_pairs:
        ix.add_voc_pairs(voc_pairs)
    if voc_file:
        ix.add_voc_file(voc_file, STEM_EN)
    if stop_words:
        ix.set_stopwords(stop_words)
    feed(ix, indextest_source(golden, npages))
    ix.finish()
    return ix


# ---- IndexSequenceTest.cs -------------------------------------------------------------------
def test_convert_roundtrip():   # IndexSequenceTest.cs:17-26
    o = OracleIndex()
    rng = np.random.default_rng(1)
    vals = np.cumsum(rng.integers(0, 0x10000, 100)).astype(np.uint64)
    enc = o.encode(vals)
    assert np.array_equal(o.decode(enc), vals)
    # codec definition, IndexSequence.cs:63-84
    assert list(o.encode([62])) == [62]
    assert list(o.encode([0])) == [0]
    assert list(o.encode([0x8000])) == [0x8000, 1]
    assert list(o.encode([5, 5])) == [5, 0]
    assert list(o.encode([(1 << 45) + 3])) == [0x8003, 0x8000, 0x8000, 1]
    assert list(o.encode([0xFFFFFFFFFFFFFFFF])) == [0xFFFF, 0xFFFF, 0xFFFF, 0xFFFF, 0xF]


def test_shift():   # IndexSequenceTest.cs:28-79
    o = OracleIndex()
    assert list(o.shift(o.encode([62]), 0)) == [62]
    rng = np.random.default_rng(2)
    vals = np.concatenate([[0], np.cumsum(rng.integers(0, 0x10000, 100))]).astype(np.uint64)
    enc = o.encode(vals)
    for s in (100, 0xFFFFF, 0xFFFFFFFF):
        sh = o.shift(enc, s)
        dec = o.decode(sh)
        assert len(dec) == len(vals)
        assert np.array_equal(dec - vals, np.full(len(vals), s, dtype=np.uint64))


# ---- IndexTest.cs ---------------------------------------------------------------------------
def test_coord_test(golden):   # IndexTest.cs:126-162 (N reduced from 1000 to 50 pages for CPU time)
    N = 50
    ix = build(golden, N, stemmers=False)
    for word, pos in golden["CoordTest"].items():
        res = ix.search(word)
        assert len(res.foundPages) == 2 * N
        for p in res.foundPages:
            assert p["pos"] == pos


def test_request_syntax(golden):   # IndexTest.cs:164-226
    g = golden["RequestSyntaxTest"]
    N = g["npages"]
    ix = build(golden, N)
    res = ix.search("and (tupman|old)")
    assert len(res.foundDocs) == 2
    assert len(res.foundDocs[0]["pages"]) == N and len(res.foundDocs[-1]["pages"]) == N
    res = ix.search("and (tupman|old) {Name=Dump}")
    assert len(res.foundDocs) == 1
    assert len(res.foundDocs[0]["pages"]) == N
    assert res.foundDocs[0]["name"] == "Test:Dump"
    for q in range(2 * N):
        assert len(res.foundPages[q]["pos"]) == 42
    assert res.foundPages[0]["pos"] == res.foundPages[1]["pos"]
    for req, want in g["dist40"].items():
        res = ix.search(req, dist=40)
        if want is None:
            assert len(res.foundPages) == 0, req
        else:
            assert len(res.foundPages[0]["pos"]) == want, req


def test_vocab(golden):   # IndexTest.cs:228-270
    N = 100
    plain = build(golden, N)
    voc = [tuple(x) for x in golden["VocabTest"]["voc"]]
    vix = build(golden, N, voc_pairs=voc)
    # vocindex["#1"].Count() == #and * 2 * Npages
    ib = vix.index_bytes()
    seqs = parse_index(ib)[1]
    o = OracleIndex()
    assert len(o.decode(seqs["#1"])) == golden["VocabTest"]["hash1_count_per_page"] * 2 * N
    for req in ("Tupman", "and", "Tupman and"):
        a, b = vix.search(req), plain.search(req)
        assert [(p["id"], p["pos"]) for p in a.foundPages] == [(p["id"], p["pos"]) for p in b.foundPages]
        assert len(a.foundPages) > 0


def test_builder(golden):   # IndexTest.cs:272-319
    ix = OracleIndex()
    ix.add_voc_pairs([tuple(x) for x in golden["VocabTest"]["voc"]])
    ix.add_doc("A", "")
    words = ["and", "tupman", "everybody", "old"]
    pos = {w: [] for w in words}
    for w, i in golden["BuilderTest_words"]:
        if w in pos:
            pos[w].append(i)
        ix.builder_add_word(w, i)
    ix.builder_end_page("1")
    ix.finish()
    for w in words:
        res = ix.search(w)
        assert len(res.foundPages) == 1
        assert res.foundPages[0]["pos"] == pos[w]


# ---- helpers + survey anchors -----------------------------------------------------------------
def parse_index(b):
    """.index layout, Index.cs:321-354."""
    max_coord = struct.unpack_from("<Q", b, 0)[0]
    i = 8
    out = {}
    while i < len(b):
        n = 0
        sh = 0
        while True:
            c = b[i]
            i += 1
            n |= (c & 0x7F) << sh
            sh += 7
            if not c & 0x80:
                break
        key = b[i:i + n].decode("utf-8")
        i += n
        cnt = struct.unpack_from("<i", b, i)[0]
        i += 4
        out[key] = np.frombuffer(b, dtype=np.uint16, count=cnt, offset=i).copy()
        i += 2 * cnt
    return max_coord, out


@pytest.mark.parametrize("fname,end,maxc,codes,postings,nshort,size,sha", [
    ("pickwick.cp1251.txt", 1731138, 1731132, 14886, 195374, 251856, 688004, "4bb968a3274a274b"),
    ("voyna.cp1251.txt", 733165, 733156, 20863, 81544, 116363, 656812, "20ab1460e23eda77"),
])
def test_survey_anchors(data_dir, stop_words, fname, end, maxc, codes, postings, nshort, size, sha):
    """stemmer-free, voc-free anchor numbers measured at survey time with an independent
    scratch restatement; file names differ here (the header page carries the file name), so only the
    name-independent quantities are compared when the name length differs."""
    orig = {"pickwick.cp1251.txt": "Dickens Charles. The Pickwick Papers - royallib.ru.txt",
            "voyna.cp1251.txt": "voyna-i-mir-tom-1.txt"}[fname]
    ix = OracleIndex()
    ix.clear_stemmers()
    ix.set_stopwords(stop_words)
    feed(ix, text_file_source(os.path.join(data_dir, fname), name=orig))
    ix.finish()
    ib = ix.index_bytes()
    mc, seqs = parse_index(ib)
    assert mc == maxc
    assert len(seqs) == codes
    assert ix.total_postings == postings
    assert sum(len(v) for v in seqs.values()) == nshort
    assert len(ib) == size
    assert hashlib.sha256(ib).hexdigest()[:16] == sha
    lb = ix.list_bytes()
    last_end = struct.unpack_from("<Q", lb, len(lb) - 8 - 1 - len(":%d" % ((os.path.getsize(os.path.join(data_dir, fname)) + 2999) // 3000)))[0]
    assert last_end == end
