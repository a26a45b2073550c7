"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same inputs.
Bit-exact for `.index` / `.index.list` bytes, hit lists and positions; ranks within 1e-6 relative (north star)."""
import hashlib
import os

import numpy as np
import pytest

import docodo_b200 as dcd
from docsets import feed, indextest_source, text_file_source
from oracle_py import STEM_EN, OracleIndex

pytestmark = pytest.mark.gpu

RANK_RTOL = 1e-6


def configure(cfg, stop_words, en_voc, golden):
    ix, ora = dcd.Index(), OracleIndex()
    if cfg in ("nostem", "nostem_stop"):
        ix.Stemmers.Clear()
        ora.clear_stemmers()
    if cfg in ("nostem_stop", "envoc"):
        ix.AddStopWords(stop_words)
        ora.set_stopwords(stop_words)
    if cfg == "vocpairs":
        pairs = [tuple(x) for x in golden["VocabTest"]["voc"]]
        v = dcd.Vocab()
        v.update(dict(pairs))
        v.Range = ("a", "z")
        ix.AddVoc(v)
        ora.add_voc_pairs(pairs)
    if cfg == "envoc":
        ix.AddVoc(dcd.Vocab(data=en_voc, stemmer=dcd.STEM_EN))
        ora.add_voc_file(en_voc, STEM_EN)
    return ix, ora


def build_both(cfg, docs, stop_words, en_voc, golden):
    ix, ora = configure(cfg, stop_words, en_voc, golden)
    ix.Create(docs)
    feed(ora, docs)
    ora.finish()
    return ix, ora


def assert_same_index(ix, ora):
    ib, lb = ix.serialize()
    ob, ol = ora.index_bytes(), ora.list_bytes()
    assert len(ib) == len(ob), (len(ib), len(ob), ix.counters.postings, ora.total_postings)
    assert ib == ob
    assert lb == ol
    assert ix.counters.postings == ora.total_postings


def assert_same_result(got, want, req):
    assert got.Success == want.success, (req, got.Error, want.error)
    if not got.Success:
        assert got.Error == want.error, req
        return
    assert [(p.id, p.pos) for p in got.foundPages] == [(p["id"], p["pos"]) for p in want.foundPages], req
    assert [p.ordinal for p in got.foundPages] == [p["page_index"] for p in want.foundPages], req
    for a, b in zip(got.foundPages, want.foundPages):
        assert abs(a.rank - b["rank"]) <= RANK_RTOL * abs(b["rank"]), (req, a.rank, b["rank"])
    assert [d.Name for d in got.foundDocs] == [d["name"] for d in want.foundDocs], req
    for a, b in zip(got.foundDocs, want.foundDocs):
        assert abs(a.rank - b["rank"]) <= RANK_RTOL * abs(b["rank"]), (req, a.rank, b["rank"])
        assert [p.ordinal for p in a.pages] == [want.foundPages[k]["page_index"] for k in b["pages"]], req
    assert [(w["Word"], w["nFound"]) for w in got.words] == [(w["word"], w["nFound"]) for w in want.words], req


INDEXTEST_QUERIES = [
    "and (tupman|old)", "and (tupman|old) {Name=Dump}", "lady old", "\"lady\" old", "\"old lady\"", "\"lady old\"",
    "lady (old | young)", "\"old ladies were\"", "\"old lady were\"", "\"old (lady|ladies) (who|were|looked)\"", "?an?",
    "tupman", "Tupman and", "everybody", "pickwick", "the and", "", "old +", "(old", "mr. pickwick, snuff!", "{Name=Sample}",
    "{Title=pickwick}", "{Size=190}", "old {Source=test}", "twenty seven minutes", "\"twenty seven\"", "190", "zzzzqq", "old zzzzqq",
    "old | zzzzqq", "gentleman (old | fat) wife", "a b", "laugh*", "merr? laugh?", "OLD", "old -filter:Dump lady",
]


@pytest.mark.parametrize("cfg", ["plain", "nostem", "nostem_stop", "vocpairs", "envoc"])
def test_indextest_build_bytes(cfg, golden, stop_words, en_voc):
    ix, ora = build_both(cfg, indextest_source(golden, 12), stop_words, en_voc, golden)
    assert_same_index(ix, ora)
    ix.Dispose()


@pytest.mark.parametrize("cfg", ["plain", "nostem", "vocpairs", "envoc"])
def test_indextest_search_parity(cfg, golden, stop_words, en_voc):
    ix, ora = build_both(cfg, indextest_source(golden, 30), stop_words, en_voc, golden)
    for dist in (None, 0, 40):
        opts = None if dist is None else dcd.SearchOptions(dist=dist)
        got = ix.SearchBatch(INDEXTEST_QUERIES, opts)
        for req, g in zip(INDEXTEST_QUERIES, got):
            assert_same_result(g, ora.search(req, dist), (cfg, dist, req))
    ix.Dispose()


def test_reference_golden_vectors_on_gpu(golden):
    """XUnitDocodoTest/IndexTest.cs:164-226 straight through the CUDA path."""
    g = golden["RequestSyntaxTest"]
    N = g["npages"]
    with dcd.Index() as ix:
        ix.Create(indextest_source(golden, N))
        res = ix.Search("and (tupman|old)")
        assert len(res.foundDocs) == 2
        assert len(res.foundDocs[0].pages) == N and len(res.foundDocs[-1].pages) == N
        res = ix.Search("and (tupman|old) {Name=Dump}")
        assert len(res.foundDocs) == 1 and len(res.foundDocs[0].pages) == N
        assert all(len(res.foundPages[q].pos) == 42 for q in range(2 * N))
        for req, want in g["dist40"].items():
            res = ix.Search(req, dcd.SearchOptions(dist=40))
            if want is None:
                assert len(res.foundPages) == 0, req
            else:
                assert len(res.foundPages[0].pos) == want, req


def test_coordtest_on_gpu(golden):
    """XUnitDocodoTest/IndexTest.cs:126-162 (N = 200 pages per document)."""
    N = 200
    with dcd.Index() as ix:
        ix.Stemmers.Clear()
        ix.Create(indextest_source(golden, N))
        for word, pos in golden["CoordTest"].items():
            res = ix.Search(word)
            assert len(res.foundPages) == 2 * N
            assert all(p.pos == pos for p in res.foundPages)


def test_builder_api_on_gpu(golden):
    """XUnitDocodoTest/IndexTest.cs:272-319."""
    pairs = [tuple(x) for x in golden["VocabTest"]["voc"]]
    ora = OracleIndex()
    ora.add_voc_pairs(pairs)
    ora.add_doc("A", "")
    with dcd.Index() as ix:
        v = dcd.Vocab()
        v.update(dict(pairs))
        v.Range = ("a", "z")
        ix.AddVoc(v)
        b = ix.GetBuilder()
        b.AddDoc("A", "")
        words = ["and", "tupman", "everybody", "old"]
        pos = {w: [] for w in words}
        for w, i in golden["BuilderTest_words"]:
            if w in pos:
                pos[w].append(i)
            b.AddWord(w, i)
            ora.builder_add_word(w, i)
        b.EndPage("1")
        ora.builder_end_page("1")
        b.Build()
        ora.finish()
        assert_same_index(ix, ora)
        for w in words:
            res = ix.Search(w)
            assert len(res.foundPages) == 1 and res.foundPages[0].pos == pos[w]


PICKWICK_QUERIES = ["pickwick", "mr pickwick said", "\"sam weller\"", "winkle (gun|rook)", "1827", "and (tupman|old)", "lady old",
                    "\"old lady\"", "lady (old | young)", "?an?", "{Name=pickwick}", "{Source=doc} weller", "fat boy asleep",
                    "\"fat boy\"", "pickwick -filter:pickwick club"]


@pytest.mark.parametrize("cfg", ["envoc", "nostem_stop"])
def test_config1_pickwick(cfg, data_dir, golden, stop_words, en_voc):
    """BASELINE config 1: Pickwick Papers with Dict/en + stop list (and the stemmer-free anchor)."""
    docs = text_file_source(os.path.join(data_dir, "pickwick.cp1251.txt"), name="Dickens Charles. The Pickwick Papers - royallib.ru.txt")
    ix, ora = build_both(cfg, docs, stop_words, en_voc, golden)
    assert_same_index(ix, ora)
    if cfg == "nostem_stop":
        assert hashlib.sha256(ix.serialize()[0]).hexdigest()[:16] == "4bb968a3274a274b"   # BASELINE.md section 5 anchor
    for dist in (None, 0, 40):
        opts = None if dist is None else dcd.SearchOptions(dist=dist)
        for req, g in zip(PICKWICK_QUERIES, ix.SearchBatch(PICKWICK_QUERIES, opts)):
            assert_same_result(g, ora.search(req, dist), (cfg, dist, req))
    ix.Dispose()


VOYNA_QUERIES = ["князь андрей", "\"князь андрей\"", "сказал (он|она)", "наташа ростова", "prince", "пьер", "\"анна павловна\""]


@pytest.mark.parametrize("cfg", ["plain", "envoc"])
def test_config2_voyna(cfg, data_dir, golden, stop_words, en_voc):
    """BASELINE config 2: Voyna i mir with (i) the Russian Snowball stemmer-only branch, (ii) en.voc + stop list."""
    docs = text_file_source(os.path.join(data_dir, "voyna.cp1251.txt"), name="voyna-i-mir-tom-1.txt")
    ix, ora = build_both(cfg, docs, stop_words, en_voc, golden)
    assert_same_index(ix, ora)
    for dist in (40, None):
        opts = None if dist is None else dcd.SearchOptions(dist=dist)
        for req, g in zip(VOYNA_QUERIES, ix.SearchBatch(VOYNA_QUERIES, opts)):
            assert_same_result(g, ora.search(req, dist), (cfg, dist, req))
    ix.Dispose()


def test_load_roundtrip(golden, stop_words, en_voc):
    """`.index` / `.index.list` written by one instance are served by another (Index.Load, Index.cs:312-380)."""
    ix, ora = build_both("envoc", indextest_source(golden, 10), stop_words, en_voc, golden)
    ib, lb = ix.serialize()
    ix2, _ = configure("envoc", stop_words, en_voc, golden)
    ix2.LoadBytes(ib, lb)
    assert ix2.serialize() == (ib, lb)
    for req in INDEXTEST_QUERIES[:12]:
        assert_same_result(ix2.Search(req, dcd.SearchOptions(dist=40)), ora.search(req, 40), req)
    seq = ix2["old"]
    assert list(seq.values()) == list(ora.decode(seq.self))
    ix.Dispose()
    ix2.Dispose()


def test_sequence_algebra_fuzz():
    """operator* / operator+ (IndexSequence.cs:218-322) on random lists: device state-machine-free form vs the statement-by-
    statement oracle, all four sign combinations of R, dense and sparse gaps, equal elements, empty operands."""
    rng = np.random.default_rng(7)
    o = OracleIndex()
    with dcd.Index() as ix:
        for trial in range(60):
            na, nb = int(rng.integers(0, 400)), int(rng.integers(0, 400))
            gap = int(rng.choice([3, 30, 300, 40000]))
            a = np.unique(np.cumsum(rng.integers(1, gap, na))).astype(np.uint64)
            b = np.unique(np.cumsum(rng.integers(1, gap, nb))).astype(np.uint64)
            if trial % 3 == 0 and len(a) and len(b):   # force shared coordinates
                b = np.unique(np.concatenate([b, rng.choice(a, min(len(a), 20))])).astype(np.uint64)
            ea, eb = o.encode(a), o.encode(b)
            assert np.array_equal(dcd.IndexSequence.from_values(ix, a).self, ea)
            assert np.array_equal(dcd.IndexSequence(ix, ea).values(), a)
            for Ra in (int(rng.integers(1, 60)), -int(rng.integers(1, 60)), 0):
                for Rb in (int(rng.integers(1, 60)), -int(rng.integers(1, 60))):
                    for op in ("*", "+"):
                        want, wantR = o.op(op, ea, Ra, eb, Rb)
                        A, B = dcd.IndexSequence(ix, ea, Ra), dcd.IndexSequence(ix, eb, Rb)
                        got = A * B if op == "*" else A + B
                        assert got.R == wantR
                        assert np.array_equal(got.self, want), (trial, op, Ra, Rb)


@pytest.mark.parametrize("na,nb", [(30000, 52000), (9000, 70001), (2_600_000, 1_900_000)])
def test_sequence_algebra_long_lists(na, nb):
    """The same operators on lists spanning many 4096-element tiles of the fused tile pipeline, and (last case) several
    512-tile chunks of merge-path cuts: tile halos, carried scan state, staged next tiles and chunk boundaries."""
    rng = np.random.default_rng(na + nb)
    o = OracleIndex()
    with dcd.Index() as ix:
        for gap_a, gap_b in ((40, 25), (3, 900), (300, 300)):
            a = np.unique(np.cumsum(rng.integers(1, gap_a, na))).astype(np.uint64)
            b = np.unique(np.cumsum(rng.integers(1, gap_b, nb))).astype(np.uint64)
            b = np.unique(np.concatenate([b, rng.choice(a, min(len(a), 5000))])).astype(np.uint64)   # shared coordinates
            ea, eb = o.encode(a), o.encode(b)
            assert np.array_equal(dcd.IndexSequence.from_values(ix, a).self, ea)
            assert np.array_equal(dcd.IndexSequence(ix, eb).values(), b)
            for Ra, Rb in ((12, 30), (-9, -14), (0, 7), (-200, 45)):
                for op in ("*", "+"):
                    want, wantR = o.op(op, ea, Ra, eb, Rb)
                    A, B = dcd.IndexSequence(ix, ea, Ra), dcd.IndexSequence(ix, eb, Rb)
                    got = A * B if op == "*" else A + B
                    assert got.R == wantR
                    assert len(got.self) == len(want), (op, Ra, Rb, len(got.self), len(want))
                    assert np.array_equal(got.self, want), (op, Ra, Rb)
            if na > 1_000_000:
                break   # one gap profile is enough at this size


def test_sequence_algebra_skewed_sizes(monkeypatch):
    """operator* with one list far shorter than the other takes the galloping path (work only around the short list's
    elements); it must give the statement-by-statement oracle's result for both operand orders, ordered and unordered R,
    shared coordinates, sparse and dense long lists (dense: the neighbourhood walk gives up and the merge path takes over),
    and the same result as with galloping switched off."""
    rng = np.random.default_rng(21)
    o = OracleIndex()
    with dcd.Index() as ix:
        for ns, ng, gap_s, gap_g in ((1, 50_000, 5, 400), (40, 30_000, 20_000, 700), (900, 120_000, 90_000, 1100), (300, 60_000, 9000, 9),
                                     (2000, 400_000, 220_000, 1100), (64, 2000, 30_000, 1000), (7, 90_000, 3, 3)):
            sm_ = np.unique(np.cumsum(rng.integers(1, gap_s, ns))).astype(np.uint64)
            g = np.unique(np.cumsum(rng.integers(1, gap_g, ng))).astype(np.uint64)
            sm_ = np.unique(np.concatenate([sm_, rng.choice(g, min(len(g), 1 + ns // 4))])).astype(np.uint64)   # shared coordinates
            es, eg = o.encode(sm_), o.encode(g)
            for Rs, Rg in ((259, 258), (-9, -14), (45, -200), (-7, 300), (1, 1)):
                for first_is_small in (True, False):
                    ea, Ra, eb, Rb = (es, Rs, eg, Rg) if first_is_small else (eg, Rg, es, Rs)
                    want, wantR = o.op("*", ea, Ra, eb, Rb)
                    monkeypatch.setenv("DCD_SEARCH_GALLOP", "16")
                    got = dcd.IndexSequence(ix, ea, Ra) * dcd.IndexSequence(ix, eb, Rb)
                    assert got.R == wantR and np.array_equal(got.self, want), (ns, ng, gap_g, Rs, Rg, first_is_small, len(got.self), len(want))
                    monkeypatch.setenv("DCD_SEARCH_GALLOP", "0")
                    plain = dcd.IndexSequence(ix, ea, Ra) * dcd.IndexSequence(ix, eb, Rb)
                    assert np.array_equal(plain.self, want)


def test_codec_edge_cases():
    """IndexSequence.Builder.Add (IndexSequence.cs:63-84): zero delta, group boundaries, first value 0."""
    o = OracleIndex()
    with dcd.Index() as ix:
        for vals in ([0], [62], [0x7FFF], [0x8000], [5, 5], [0, 0x7FFF, 0x8000 + 0x7FFF, 1 << 30, (1 << 30) + (1 << 15)],
                     [(1 << 40) - 1], list(range(0, 5000, 7))):
            v = np.array(vals, dtype=np.uint64)
            enc = dcd.IndexSequence.from_values(ix, v).self
            assert np.array_equal(enc, o.encode(v)), vals[:4]
            assert np.array_equal(dcd.IndexSequence(ix, enc).values(), v)
        assert dcd.IndexSequence.from_values(ix, []).Count == 0


def test_bulk_pages_with_empty_and_ragged(golden, stop_words, en_voc):
    """Bulk entry (host buffers): ragged page lengths, empty pages (skipped, Build.cs:483), a page that ends inside a word,
    a run longer than MAX_WORD_LENGTH, digits glued to letters, Latin-1 and UTF-16 encodings."""
    rng = np.random.default_rng(3)
    base = golden["TestText1"].replace("�", "'")
    pieces, docs_first = [], [0]
    for d in range(40):
        for p in range(int(rng.integers(0, 6))):
            k = int(rng.integers(0, len(base) - 400))
            n = int(rng.choice([0, 1, 2, 3, 7, 150, 333, 400]))
            pieces.append(base[k:k + n])
        pieces.append("x" * 40 + " abc123def " + "Z" * 32 + " 1234567 " + "y" * 33)
        docs_first.append(len(pieces))
    offs = np.concatenate([[0], np.cumsum([len(p) for p in pieces])]).astype(np.uint64)
    text = "".join(pieces)
    for enc in (dcd.ENC_LATIN1, dcd.ENC_UTF16):
        buf = np.frombuffer(text.encode("latin-1" if enc == dcd.ENC_LATIN1 else "utf-16-le"),
                            dtype=np.uint8 if enc == dcd.ENC_LATIN1 else np.uint16)
        ix, ora = configure("envoc", stop_words, en_voc, golden)
        b = ix.GetBuilder()
        b.AddPagesBulk(buf, enc, offs, docs_first, source="doc")
        b.Build()
        ora.build_bulk(buf, 1 if enc == dcd.ENC_LATIN1 else 2, offs, docs_first, nthreads=1)
        assert_same_index(ix, ora)
        for req in ("old lady", "abc", "123", "\"old lady\"", "def"):
            assert_same_result(ix.Search(req, dcd.SearchOptions(dist=40)), ora.search(req, 40), req)
        ix.Dispose()


def test_bulk_host_text_arrives_in_chunks(monkeypatch, data_dir, golden, stop_words, en_voc):
    """Host-buffer bulk entry: the text is copied in chunks on the copy stream and tokenized chunk by chunk (one launch per
    chunk over the tiles whose window is resident).  4 KB and 7 KB chunks (not multiples of the 960-unit tile) over a
    1.6 MB corpus, Latin-1 and UTF-16, must give the oracle's bytes."""
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:5000]
    ix0, ora = configure("envoc", stop_words, en_voc, golden)
    syn = ix0.synth_generate(words, 1.1, 0xD0C0D00003, ndocs=60, pages_per_doc=20, min_page_chars=200)
    text = ix0.d2h(syn.d_text, syn.text_bytes)
    offs = ix0.d2h(syn.d_page_offs, 8 * (syn.npages + 1), np.uint64)
    dfp = ix0.d2h(syn.d_doc_first_page, 8 * (syn.ndocs + 1), np.uint64)
    ix0.synth_free(syn)
    ix0.Dispose()
    ora.build_bulk(text, 1, offs, dfp, nthreads=1)
    for kb, enc in ((4, dcd.ENC_LATIN1), (7, dcd.ENC_LATIN1), (1, dcd.ENC_UTF16), (1 << 20, dcd.ENC_LATIN1)):
        monkeypatch.setenv("DCD_H2D_CHUNK_KB", str(kb))
        ix = configure("envoc", stop_words, en_voc, golden)[0]
        buf = text if enc == dcd.ENC_LATIN1 else text.astype(np.uint16)
        b = ix.GetBuilder()
        b.AddPagesBulk(buf, enc, offs, dfp, source="doc")
        b.Build()
        assert_same_index(ix, ora)
        ix.Dispose()
    # a builder dropped with copies in flight drains them before its text buffer returns to the pool
    monkeypatch.setenv("DCD_H2D_CHUNK_KB", "4")
    ix = configure("envoc", stop_words, en_voc, golden)[0]
    b = ix.GetBuilder()
    b.AddPagesBulk(text, dcd.ENC_LATIN1, offs, dfp, source="doc")
    del b
    b = ix.GetBuilder()
    b.AddPagesBulk(text, dcd.ENC_LATIN1, offs, dfp, source="doc")
    b.Build()
    assert_same_index(ix, ora)
    ix.Dispose()


def test_query_waves_overlap(monkeypatch, data_dir, golden, stop_words, en_voc):
    """A batch cut into many small waves (DCD_SEARCH_WAVE_KB) that overlap on the device (DCD_SEARCH_INFLIGHT streams), with
    the shaping threads active (bulk index, >= 64 requests): every depth gives the oracle's results."""
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:3000]
    ix, ora = configure("envoc", stop_words, en_voc, golden)
    syn = ix.synth_generate(words, 1.1, 0xD0C0D00003, ndocs=120, pages_per_doc=20, min_page_chars=200)
    text = ix.d2h(syn.d_text, syn.text_bytes)
    offs = ix.d2h(syn.d_page_offs, 8 * (syn.npages + 1), np.uint64)
    dfp = ix.d2h(syn.d_doc_first_page, 8 * (syn.ndocs + 1), np.uint64)
    b = ix.GetBuilder()
    b.AddPagesBulk((syn.d_text, syn.d_page_offs, syn.d_doc_first_page, syn.npages, syn.ndocs), dcd.ENC_LATIN1, None, None,
                   source="doc", on_device=True)
    b.Build()
    ora.build_bulk(text, 1, offs, dfp, nthreads=1)
    rng = np.random.default_rng(11)
    pick = lambda: words[int(rng.integers(0, 400))]   # noqa: E731
    reqs = []
    for i in range(96):
        k = i % 4
        reqs.append(f"{pick()} {pick()}" if k == 0 else f"\"{pick()} {pick()}\"" if k == 1 else f"{pick()} {pick()} {pick()}" if k == 2
                    else f"{pick()} ({pick()} | {pick()})")
    want = [ora.search(r, 40) for r in reqs]
    for kb, depth in ((None, None), (256, 1), (256, 3), (64, 8), (1, 2)):
        if kb is not None:
            monkeypatch.setenv("DCD_SEARCH_WAVE_KB", str(kb))
            monkeypatch.setenv("DCD_SEARCH_INFLIGHT", str(depth))
        got = ix.SearchBatch(reqs, dcd.SearchOptions(dist=40))
        for r, g, w in zip(reqs, got, want):
            assert_same_result(g, w, (kb, depth, r))
    assert sum(len(w.foundPages) for w in want) > 1000
    ix.synth_free(syn)
    ix.Dispose()


def test_queries_with_large_coordinate_origin(data_dir, golden, stop_words, en_voc):
    """A document-range shard far into the corpus: coordinates beyond 2^32 and close to the 2^40 limit (dcd_build_set_origin).
    Index bytes differ only by the first delta of every list; found pages, positions and ranks do not depend on the origin."""
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:5000]
    w = [x for x in words if x not in set(stop_words)]
    reqs = [f"{w[0]} {w[1]}", f"\"{w[2]} {w[3]}\"", w[4], f"{w[5]} ({w[6]} | {w[7]})", f"{w[0]} {w[0]}"]
    res, first = {}, {}
    for origin in (0, 5_000_000_000, (1 << 40) - (1 << 20)):
        ix = configure("envoc", stop_words, en_voc, golden)[0]
        syn = ix.synth_generate(words, 1.1, 0xD0C0D00003, ndocs=60, pages_per_doc=20, min_page_chars=200)
        b = ix.GetBuilder()
        b.SetOrigin(origin, 0)
        b.AddPagesBulk((syn.d_text, syn.d_page_offs, syn.d_doc_first_page, syn.npages, syn.ndocs), dcd.ENC_LATIN1, None, None,
                       source="doc", on_device=True)
        b.Build()
        rs = ix.SearchBatch(reqs, dcd.SearchOptions(dist=40))
        res[origin] = [[(p.id, p.ordinal, tuple(p.pos), p.rank) for p in r.foundPages] for r in rs]
        first[origin] = ix[w[0]].values()[:4].tolist()
        assert ix.MaxCoord >= origin and ix.page_info(0)[3] > origin
        ix.synth_free(syn)
        ix.Dispose()
    assert all(len(x) > 20 for x in res[0])
    assert res[0] == res[5_000_000_000] == res[(1 << 40) - (1 << 20)]
    assert [v - 5_000_000_000 for v in first[5_000_000_000]] == first[0]
    # one unit past the limit is refused (coordinates are 40-bit)
    ix = configure("envoc", stop_words, en_voc, golden)[0]
    syn = ix.synth_generate(words, 1.1, 0xD0C0D00003, ndocs=60, pages_per_doc=20, min_page_chars=200)
    b = ix.GetBuilder()
    b.SetOrigin((1 << 40) - 1000, 0)
    with pytest.raises(Exception):
        b.AddPagesBulk((syn.d_text, syn.d_page_offs, syn.d_doc_first_page, syn.npages, syn.ndocs), dcd.ENC_LATIN1, None, None,
                       source="doc", on_device=True)
        b.Build()
    ix.synth_free(syn)
    ix.Dispose()


def test_synthetic_corpus_device_resident(data_dir, golden, stop_words, en_voc):
    """BASELINE config 3 at test size: Zipf corpus generated on the device, indexed from device pointers; the same bytes are
    copied to the host and indexed by the oracle (single- and multi-threaded stitch)."""
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:20000]
    ix, ora = configure("envoc", stop_words, en_voc, golden)
    syn = ix.synth_generate(words, 1.1, 0xD0C0D00003, ndocs=300, pages_per_doc=20, min_page_chars=200)
    assert syn.npages == 6000 and syn.text_bytes > 6000 * 200
    text = ix.d2h(syn.d_text, syn.text_bytes)
    offs = ix.d2h(syn.d_page_offs, 8 * (syn.npages + 1), np.uint64)
    dfp = ix.d2h(syn.d_doc_first_page, 8 * (syn.ndocs + 1), np.uint64)
    assert offs[0] == 0 and offs[-1] == syn.text_bytes and np.all(np.diff(offs.astype(np.int64)) >= 200)
    b = ix.GetBuilder()
    b.AddPagesBulk((syn.d_text, syn.d_page_offs, syn.d_doc_first_page, syn.npages, syn.ndocs), dcd.ENC_LATIN1, None, None,
                   source="doc", on_device=True)
    b.Build()
    ora.build_bulk(text, 1, offs, dfp, nthreads=1)
    assert_same_index(ix, ora)
    ora4 = configure("envoc", stop_words, en_voc, golden)[1]
    ora4.build_bulk(text, 1, offs, dfp, nthreads=4)
    assert ora4.index_bytes() == ora.index_bytes() and ora4.list_bytes() == ora.list_bytes()
    k = ix.counters
    assert k.text_units == syn.text_bytes and k.pages == syn.npages
    ix.synth_free(syn)
    ix.Dispose()


def test_word_in_last_table_slot(golden, stop_words, en_voc):
    """The bulk tokenizer's 4-byte records use slot 0x3FFFFF as "no word" while the default word table has exactly 2^22
    slots: words whose hash (or whose linear probe) lands on the last slot must not be dropped.  The words below hash to
    slots 0x3FFFFF, 0x3FFFFE and 0x3FFFFD of a 2^22-slot table (found by brute force over wordkey_slot_hash); two words per
    slot make the second one probe forward, across the reserved last slot and around to slot 0."""
    last = ["plyqklm", "nbnmlxs", "okhwxgh", "hqqwjac", "hhfvohp", "upxxnrt", "nodyupx", "jopspvr", "ckizebi", "wqtggrh"]
    filler = golden["TestText1"].replace("�", "'")[:600]
    pieces, docs_first = [], [0]
    for d in range(6):
        for p in range(3):
            pieces.append(" ".join(last[(d + p + k) % len(last)] for k in range(7)) + " " + filler)
        docs_first.append(len(pieces))
    offs = np.concatenate([[0], np.cumsum([len(p) for p in pieces])]).astype(np.uint64)
    buf = np.frombuffer("".join(pieces).encode("latin-1"), dtype=np.uint8)
    ix, ora = configure("envoc", stop_words, en_voc, golden)
    b = ix.GetBuilder()
    b.AddPagesBulk(buf, dcd.ENC_LATIN1, offs, docs_first, source="doc")
    b.Build()
    ora.build_bulk(buf, 1, offs, docs_first, nthreads=1)
    assert_same_index(ix, ora)
    for w in last:
        assert ix[w].Count > 0, w
        assert_same_result(ix.Search(w, dcd.SearchOptions(dist=40)), ora.search(w, 40), w)
    ix.Dispose()


def _utf8_pages(pieces):
    """pieces (str pages) -> (utf-8 bytes, byte offsets, utf-16 units, unit offsets)."""
    b8 = [p.encode("utf-8") for p in pieces]
    u16 = [np.frombuffer(p.encode("utf-16-le"), dtype=np.uint16) for p in pieces]
    offs8 = np.concatenate([[0], np.cumsum([len(x) for x in b8])]).astype(np.uint64)
    offs16 = np.concatenate([[0], np.cumsum([len(x) for x in u16])]).astype(np.uint64)
    return (np.frombuffer(b"".join(b8), dtype=np.uint8), offs8,
            np.concatenate(u16) if offs16[-1] else np.zeros(0, np.uint16), offs16)


@pytest.mark.parametrize("kind", ["voyna", "mixed", "latin1", "ascii"])
def test_utf8_page_input(kind, data_dir, golden, stop_words, en_voc):
    """DCD_ENC_UTF8 (north star: "UTF-8/UTF-16 tokenizer"): UTF-8 bytes with BYTE page offsets must give the index the oracle
    builds from the decoded UTF-16 string -- coordinates are UTF-16 units (Build.cs:529-531), a 4-byte sequence counts two.
    voyna: config 2's text re-encoded (2-byte Cyrillic, Russian Snowball); mixed: 1/2/3/4-byte sequences, pages that end
    inside words, empty pages, sequences that straddle the 32-byte lane and 1024-byte block boundaries of the transcoder;
    latin1: code points < U+0100 only (Latin-1 target); ascii: indexed in place."""
    rng = np.random.default_rng(5)
    if kind == "voyna":
        text = open(os.path.join(data_dir, "voyna.cp1251.txt"), "rb").read().decode("cp1251")
        pieces = [text[k:k + 3000] for k in range(0, len(text), 3000)]
        docs_first = [0, len(pieces) // 2, len(pieces)]
    else:
        base = golden["TestText1"].replace("�", "'")
        ru = open(os.path.join(data_dir, "voyna.cp1251.txt"), "rb").read().decode("cp1251")[5000:9000]
        pieces, docs_first = [], [0]
        for d in range(30):
            for p in range(int(rng.integers(1, 5))):
                k = int(rng.integers(0, len(base) - 500))
                n = int(rng.choice([0, 1, 5, 31, 33, 200, 480]))
                s = base[k:k + n]
                if kind == "mixed":
                    j = int(rng.integers(0, len(ru) - 300))
                    s = s[:n // 2] + " " + ru[j:j + int(rng.integers(0, 300))] + " \U0001D400\U0001F600x日本語 naïve ÉCOLE " + s[n // 2:] + \
                        "€" * int(rng.integers(0, 40)) + "ж" * int(rng.integers(0, 36)) + " " + "ÿ" * 5
                elif kind == "latin1":
                    s = s[:n // 2] + " naïve ÉCOLE façade Ærø " + s[n // 2:] + "é" * int(rng.integers(0, 35)) + " ÿ×÷ "
                pieces.append(s)
            docs_first.append(len(pieces))
    buf8, offs8, buf16, offs16 = _utf8_pages(pieces)
    if kind == "voyna":
        ix, ora = dcd.Index(), OracleIndex()   # vocs == []: Russian Snowball through the stemmer-only branch (Build.cs:211-237)
    else:
        ix, ora = configure("envoc", stop_words, en_voc, golden)
    b = ix.GetBuilder()
    b.AddPagesBulk(buf8, dcd.ENC_UTF8, offs8, docs_first, source="doc")
    b.Build()
    ora.build_bulk(buf16, 2, offs16, docs_first, nthreads=1)
    assert_same_index(ix, ora)
    assert ix.counters.text_units == len(buf16)
    reqs = ("князь андрей", "\"князь андрей\"", "сказал (он|она)") if kind == "voyna" else ("old lady", "naïve", "école", "\"old lady\"")
    for req in reqs:
        assert_same_result(ix.Search(req, dcd.SearchOptions(dist=40)), ora.search(req, 40), req)
    ix.Dispose()
    if kind in ("mixed", "ascii"):   # device-resident UTF-8 text (on_device): same bytes from a device buffer
        import torch
        ix2 = configure("envoc", stop_words, en_voc, golden)[0]
        t = torch.from_numpy(buf8.copy()).cuda()
        po = torch.from_numpy(offs8.astype(np.int64)).cuda()
        dfp = torch.tensor(docs_first, dtype=torch.int64).cuda()
        torch.cuda.synchronize()
        b = ix2.GetBuilder()
        b.AddPagesBulk((t.data_ptr(), po.data_ptr(), dfp.data_ptr(), len(offs8) - 1, len(docs_first) - 1), dcd.ENC_UTF8, None, None,
                       source="doc", on_device=True)
        b.Build()
        assert_same_index(ix2, ora)
        ix2.Dispose()


def test_sort_lookback_segments(monkeypatch, data_dir, golden, stop_words, en_voc):
    """Builds beyond 2^30 postings: the look-back of the posting sort is cut into segments with 64-bit bases (sort.cu).  With
    DCD_SORT_SEG_TILES=2 a 60 K-page corpus crosses a hundred segment boundaries in both passes; the index must not change."""
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()[:20000]
    ix, ora = configure("envoc", stop_words, en_voc, golden)
    syn = ix.synth_generate(words, 1.1, 0xD0C0D00003, ndocs=3000, pages_per_doc=20, min_page_chars=200)
    text = ix.d2h(syn.d_text, syn.text_bytes)
    offs = ix.d2h(syn.d_page_offs, 8 * (syn.npages + 1), np.uint64)
    dfp = ix.d2h(syn.d_doc_first_page, 8 * (syn.ndocs + 1), np.uint64)
    ora.build_bulk(text, 1, offs, dfp, nthreads=8)
    for seg in ("2", "3", "65536"):
        monkeypatch.setenv("DCD_SORT_SEG_TILES", seg)
        b = ix.GetBuilder()
        b.AddPagesBulk((syn.d_text, syn.d_page_offs, syn.d_doc_first_page, syn.npages, syn.ndocs), dcd.ENC_LATIN1, None, None,
                       source="doc", on_device=True)
        b.Build()
        assert ix.counters.postings > 8192 * 200
        assert_same_index(ix, ora)
    ix.synth_free(syn)
    ix.Dispose()
