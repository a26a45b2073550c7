"""Result shaping for REST: snippets (Search.cs:619-751), headers / summary (Search.cs:558-596) and the
/search JSON (server.cs:85-96) -- pure host string work, checked against values derived by hand from the reference code.
The reference cannot run here (no .NET), so the JSON member order / number formatting of Json.NET is restated, not pinned."""
import json

from docodo_b200.index import ResultDocPage, ResultDocument, SearchResult
from docodo_b200.results import (BEGIN_MATCHED_SYMBOL as B, END_MATCHED_SYMBOL as E, MakeHeaders, PreparePageText, SpannableString,
                                 search_json, shape_documents)


def test_spannable_builder_and_string():
    text = "The old lady looked at the old man."
    sp = SpannableString.build(text, [4, 8])
    assert [list(x) for x in sp] == [[0, "The "], [1, "old"], [0, " "], [1, "lady"], [0, " looked at the old man."]]
    assert str(sp) == f"The {B}old{E} {B}lady{E} looked at the old man."
    # Substring keeps found words whole and cuts plain spans (Search.cs:632-669)
    assert str(sp.substring(2, 8)) == f"e {B}old{E} {B}lady{E}"
    assert str(sp.substring(0, 100)) == str(sp)


def test_prepare_page_text_window_and_punctuation():
    text = ("x" * 200) + " Mr. Pickwick ,,, said :  the old lady !!! Yes ... " + ("y" * 400)
    pos = [text.index("old"), text.index("lady")]
    page = ResultDocPage("3", pos, 1.0, 0)
    out = str(PreparePageText(page, text, 320))
    # window: [min(pos) - 80, min(max(pos) + 80, start + 320)); runs of punctuation between words collapse to one mark + space
    # (RegexReplace works span by span, Search.cs:671-678: the "!!!" that directly follows a found word starts its span, so no
    # \b precedes it and it stays)
    assert f"the {B}old{E} {B}lady{E} !!! Yes. " in out and "Pickwick, said: the" in out
    assert out.startswith("x" * 50 + " Mr. Pickwick") and not out.startswith("x" * 51) and len(out) < 330


def test_make_headers():
    assert MakeHeaders("Name=Sample\nSource=Test\n") == {"Name": "Sample", "Source": "Test"}


def test_shape_and_json():
    texts = {("Test:Sample", "0"): "Name=Sample\nTitle=The Pickwick Papers\n",
             ("Test:Sample", "1"): "It was an old lady. The old lady looked out.",
             ("Test:Sample", "2"): "No match here but one lady, old."}
    r = SearchResult()
    p0 = ResultDocPage("0", [22], 1.0, 0)          # "Pickwick" in the header page: 12 + len("Title=The ")
    p1 = ResultDocPage("1", [10, 14, 24, 28], 2.5, 1)
    p2 = ResultDocPage("2", [22, 28], 1.5, 2)
    d = ResultDocument("Test:Sample", 3.25, [p0, p1, p2])
    r.foundDocs, r.foundPages = [d], [p0, p1, p2]
    shape_documents(r, lambda name, pid: texts.get((name, pid)))
    assert d.headers == {"Name": "Sample", "Title": f"The {B}Pickwick{E} Papers"}
    assert [p.id for p in d.pages] == ["1", "2"]                                   # page "0" is removed (Search.cs:575)
    assert d.pages[0].text == f"It was an {B}old{E} {B}lady{E}. The {B}old{E} {B}lady{E} looked out."
    assert d.pages[1].text == f"No match here but one {B}lady{E}, {B}old{E}."
    assert d.foundWords == ["old", "lady"]
    assert d.summary == d.pages[0].text + " ... " + d.pages[1].text   # lowest ranks first, then id order
    js = json.loads(search_json(r))
    assert js["found"] == 1 and list(js["result"][0].keys()) == ["pages", "headers", "rank", "summary", "foundWords", "Name", "nPages"]
    assert list(js["result"][0]["pages"][0].keys()) == ["id", "pos", "text", "rank"]
    assert js["result"][0]["rank"] == 3.25 and js["result"][0]["pages"][0]["pos"] == [10, 14, 24, 28]
    assert search_json(r).startswith('{\n  "found": 1,\n  "result": [\n    {\n      "pages": [')


def test_no_text_source_leaves_pages_untouched():
    r = SearchResult()
    p = ResultDocPage("1", [3], 1.0, 0)
    r.foundDocs, r.foundPages = [ResultDocument("a:b", 1.0, [p])], [p]
    shape_documents(r, None)
    assert r.foundDocs[0].pages == [p] and json.loads(search_json(r))["result"][0]["summary"] is None
