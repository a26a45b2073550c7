"""Dictionary-side queries on the device against the oracle's restatement of Search.cs:160-188 and
Index.cs:46-89,284-307: wildcard words, Levenshtein corrections, suggestions, the most-popular-words histogram."""
import os

import pytest

import docodo_b200 as dcd
from docsets import feed, text_file_source
from test_gpu_parity import configure

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pickwick(data_dir, golden, stop_words, en_voc):
    docs = text_file_source(os.path.join(data_dir, "pickwick.cp1251.txt"), name="Dickens Pickwick Papers.txt")
    ix, ora = configure("envoc", stop_words, en_voc, golden)
    ix.Create(docs)
    feed(ora, docs)
    ora.finish()
    ora.load(ora.index_bytes(), ora.list_bytes())
    yield ix, ora
    ix.Dispose()


def test_like_words(pickwick):
    ix, ora = pickwick
    for w in ("_an_", "pick_", "_wick", "w_ller", "_", "a_", "_zzzz_", "tup_an", "plain", "s_m_w_", "_1_"):
        assert ix.GetLikeWords(w) == ora.like_words(w), w
    assert len(ix.GetLikeWords("_an_")) == 100 and "pickwick" in ix.GetLikeWords("pick_")


def test_close_words(pickwick):
    ix, ora = pickwick
    for w in ("pikwick", "wellar", "tupmann", "gentelman", "x", "", "zzzzzzzzzzzzzzzzzzzzzzzzzzzzzzzz", "#1", "$pick", "1827"):
        assert ix.GetCloseWords(w) == ora.close_words(w), w
    assert ix.GetCloseWords("pikwick")[0] == "pickwick"
    with pytest.raises(dcd.DocodoError):
        ix.GetCloseWords("a" * 65)


def test_suggestions(pickwick):
    ix, ora = pickwick
    for req, n in (("mr pick", 10), ("pi", 10), ("the wel", 5), ("p", 10), ("sam ", 10), ("gentle", 3), ("Mr. PICK", 10), ("zzzq", 10),
                   ("hello, ", 10), ("snod", 50)):
        assert ix.GetSuggessions(req, n) == ora.suggestions(req, n), req
    assert ix.GetSuggessions("mr pick", 10)[0] == "wick"


def test_histogram(pickwick):
    ix, ora = pickwick
    got, want = ix.CalcHistogram(1000), ora.histogram(1000)
    assert got == want and list(got.items()) == list(want.items())
    assert len(got) > 100 and any(k.startswith("(") for k in got)
    assert ix.CalcHistogram(7) == ora.histogram(7)


def test_on_loaded_index(pickwick):
    """Index.Load path: the key table comes from the file, the device copy is made on first use."""
    ix, ora = pickwick
    ib, lb = ix.serialize()
    ix2 = dcd.Index()
    ix2.LoadBytes(ib, lb)
    assert ix2.GetCloseWords("pikwick") == ora.close_words("pikwick")
    assert ix2.GetSuggessions("wel", 4) == ora.suggestions("wel", 4)
    ix2.Dispose()


def test_search_json_from_device_results(pickwick, data_dir):
    """/search body (server.cs:85-96) from the device's hit lists: snippets cut from the page texts the data source returns."""
    import json

    from docodo_b200.results import BEGIN_MATCHED_SYMBOL, search_json, shape_documents
    ix, ora = pickwick
    docs = text_file_source(os.path.join(data_dir, "pickwick.cp1251.txt"), name="Dickens Pickwick Papers.txt")
    texts = {(f"{src}:{name}", pid): text for src, name, pages in docs for pid, text in pages}
    r = ix.Search("\"old lady\"", dcd.SearchOptions(dist=0))
    want = ora.search("\"old lady\"", 0)
    shape_documents(r, lambda name, pid: texts.get((name, pid)))
    js = json.loads(search_json(r))
    assert js["found"] == len(want.foundDocs) == 1
    d = js["result"][0]
    assert d["Name"] == "doc:Dickens Pickwick Papers.txt" and d["headers"]["Source"] == "doc"
    assert [p["id"] for p in d["pages"]] == [p["id"] for p in want.foundPages] and len(d["pages"]) >= 5
    for p, w in zip(d["pages"], want.foundPages):
        assert p["pos"] == w["pos"] and abs(p["rank"] - w["rank"]) <= 1e-6 * w["rank"]
        assert (BEGIN_MATCHED_SYMBOL + "old") in p["text"].lower() and len(p["text"]) <= 400
    assert sorted(w.lower() for w in d["foundWords"]) == ["lady", "old"] and " ... " in d["summary"]
