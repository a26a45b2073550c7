"""Result shaping above the C ABI: what Index.Search does with the found pages once the posting-list
work is done -- document headers, found-text snippets, summaries (Search.cs:552-596, 619-751) -- and the `/search` JSON of
DocodoServer (server.cs:85-96).  Host-side string work over the page texts a data source hands back; the page lists,
positions and ranks come from the device through docodo_b200.Index.Search.

    text_source(doc_name, page_id) -> str or None      stands in for IIndexDirectDataSource[doc][page].text
"""
import json

import regex

BEGIN_MATCHED_SYMBOL = "ˋ"   # SearchResult.BEGIN_MATCHED_SYMBOL, Search.cs:25
END_MATCHED_SYMBOL = "ˊ"     # Search.cs:26
MAX_FOUND_PAGE_TEXT = 320         # Index.MAX_FOUND_PAGE_TEXT, Index.cs:103 (Search.cs:579 uses the Index field)

_WORD_END = regex.compile(r"(?<=\w)\b")
_PUNCT = [(regex.compile(r"\b\W*\.+\W*\b"), ". "), (regex.compile(r"\b\W*\?+\W*\b"), "? "), (regex.compile(r"\b\W*!+\W*\b"), "! "),
          (regex.compile(r"\b\W*:+\W*\b"), ": "), (regex.compile(r"\b\W*,+\W*\b"), ", ")]   # Search.cs:742-746


class SpannableString(list):
    """Search.cs:630-726: a list of [format, text] spans; format != 0 marks a found word."""

    @staticmethod
    def build(text, startwords):   # Builder.Add(string, int[]), Search.cs:701-713
        sp = SpannableString()
        lastpos = 0
        for pos in startwords:
            sp.append([0, text[lastpos:pos]])
            m = _WORD_END.search(text[pos:])
            wordend = m.start() if m else 0   # Match.Index of a failed match is 0
            sp.append([1, text[pos:pos + wordend]])
            lastpos = pos + wordend
        sp.append([0, text[lastpos:]])
        return sp

    def substring(self, start, length):   # Search.cs:632-669
        res = SpannableString()
        l = 0
        for fmt, text in self:
            l += len(text)
            if not res and l > start:
                if fmt != 0:
                    res.append([fmt, text])
                else:
                    res.append([0, text[start - l + len(text):]])
            elif res:
                if l >= start + length:
                    if fmt != 0:
                        res.append([fmt, text])
                    else:
                        res.append([0, text[:start + length - l + len(text)]])
                    break
                res.append([fmt, text])
        return res

    def regex_replace(self, pattern, replace):   # Search.cs:671-678
        for sp in self:
            sp[1] = pattern.sub(replace, sp[1])
        return self

    def __str__(self):   # implicit operator string, Search.cs:680-691
        return "".join(BEGIN_MATCHED_SYMBOL + t + END_MATCHED_SYMBOL if f != 0 else t for f, t in self)


def PreparePageText(page, text, Len=MAX_FOUND_PAGE_TEXT):   # Search.cs:729-750
    spans = SpannableString.build(text, page.pos)
    r0 = min(max(0, min(page.pos) - Len // 4), len(text))
    r1 = min(min(max(page.pos) + Len // 4, len(text)), r0 + Len)
    res = spans.substring(r0, r1 - r0)
    for pat, rep in _PUNCT:
        res.regex_replace(pat, rep)
    return res


def MakeHeaders(s):   # ResultDocument.MakeHeaders, Search.cs:80-88
    headers = {}
    splits = regex.split(r"[=\n]", s)
    for q in range(0, len(splits) - 1, 2):
        if splits[q] in headers:
            raise KeyError(splits[q])   # Dictionary.Add throws on a duplicate header name
        headers[splits[q]] = splits[q + 1]
    return headers


def shape_documents(result, text_source):
    """Search.cs:558-596 for every found document: headers from page "0", found-text snippets per page, summary, found words.
    Mutates and returns `result` (docodo_b200.SearchResult)."""
    for doc in result.foundDocs:
        doc.foundWords = []
        doc.headers = None
        doc.summary = None
        headers = text_source(doc.Name, "0") if text_source else None
        if headers is None:
            continue
        pages = list(doc.pages)
        if pages and pages[0].id == "0":
            headers = str(SpannableString.build(headers, pages[0].pos))
        doc.headers = MakeHeaders(headers)
        pages = [p for p in pages if p.id != "0"]
        for page in pages:
            sp = PreparePageText(page, text_source(doc.Name, page.id) or "", MAX_FOUND_PAGE_TEXT)
            seen = []
            for f, t in sp:
                if f != 0 and t not in seen:
                    seen.append(t)
            doc.foundWords += seen
            page.text = str(sp)
        doc.pages = pages
        if pages:   # the three lowest-ranked pages, then in id order (string compare), joined by " ... "
            top = sorted(sorted(pages, key=lambda p: p.rank)[:3], key=lambda p: p.id)
            doc.summary = " ... ".join(p.text for p in top)
        uniq = []
        for w in doc.foundWords:
            if w not in uniq:
                uniq.append(w)
        doc.foundWords = uniq
    return result


def _page_obj(p):   # Json.NET: public fields in declaration order, then properties (ResultDocPage, Search.cs:92-123)
    return {"id": p.id, "pos": list(p.pos), "text": p.text, "rank": _f32(p.rank)}


def _f32(x):
    """float -> the shortest round-trip text of the float32 value (what Json.NET writes for a C# float)."""
    import numpy as np
    return float(np.format_float_positional(np.float32(x), unique=True, trim="0")) if x == x else x


def search_json(result):
    """The body DocodoServer answers `GET /search?req=` with (server.cs:85-96): {"found": N, "result": [ResultDocument...]},
    Json.NET Formatting.Indented (two spaces).  ResultDocument fields: pages, headers, rank, summary, foundWords, then the
    inherited Document.Name / nPages (Search.cs:59-90, Build.cs:23-29)."""
    docs = []
    for d in result.foundDocs:
        docs.append({"pages": [_page_obj(p) for p in d.pages], "headers": getattr(d, "headers", None), "rank": _f32(d.rank),
                     "summary": getattr(d, "summary", None), "foundWords": getattr(d, "foundWords", None), "Name": d.Name,
                     "nPages": getattr(d, "nPages", 0)})
    return json.dumps({"found": len(result.foundDocs), "result": docs}, indent=2, ensure_ascii=False)
