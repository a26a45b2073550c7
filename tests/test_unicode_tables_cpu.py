"""The character tables behind the tokenizer (\\p{L}, \\p{N}, \\w, ToLower / ToUpper per UTF-16 unit; Build.cs:482,526) are
generated from Python's `unicodedata` for both the product (docodo_b200/csrc/unicode_tables.h) and the oracle -- one source,
so a wrong rule would be invisible to the parity tests.  This test checks the committed product header against an
INDEPENDENT implementation of the same classes, the third-party `regex` module (its own Unicode database and its own
\\p{..} / \\w code), over every BMP code unit, and the case tables against closed-form rules for the scripts the reference's
tests pin (ASCII, Latin-1, Greek, Cyrillic, Armenian)."""
import os
import re

import regex

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLS_L, CLS_N, CLS_W = 1, 2, 4


def parse_header(path, prefix):
    h = open(path).read()
    body = lambda name: h[h.index("%s_%s[]" % (prefix, name)):].split("};")[0]   # noqa: E731
    cls = [0] * 0x10000
    for a, b, v in re.findall(r"\{0x([0-9A-F]{4}),0x([0-9A-F]{4}),(\d+)\}", body("CLS_RUNS")):
        for c in range(int(a, 16), int(b, 16) + 1):
            cls[c] = int(v)
    maps = {}
    for name in ("LOWER", "UPPER"):
        m = list(range(0x10000))
        for a, b in re.findall(r"\{0x([0-9A-F]{4}),0x([0-9A-F]{4})\}", body(name + "_PAIRS")):
            m[int(a, 16)] = int(b, 16)
        maps[name] = m
    return cls, maps["LOWER"], maps["UPPER"]


def test_product_and_oracle_tables_are_the_same_table():
    p = parse_header(os.path.join(ROOT, "docodo_b200", "csrc", "unicode_tables.h"), "DCD")
    o = parse_header(os.path.join(ROOT, "oracle", "unicode_tables.h"), "ORC")
    assert p == o


def test_classes_against_the_regex_module():
    cls, _, _ = parse_header(os.path.join(ROOT, "docodo_b200", "csrc", "unicode_tables.h"), "DCD")
    pL, pN, pW = regex.compile(r"\p{L}"), regex.compile(r"\p{N}"), regex.compile(r"[\p{L}\p{Mn}\p{Nd}\p{Pc}]")
    unassigned = regex.compile(r"\p{Cn}")
    diffs = []
    for c in range(0x10000):
        if 0xD800 <= c <= 0xDFFF:
            assert cls[c] == 0          # a lone surrogate unit is never a letter / digit / word character
            continue
        ch = chr(c)
        want = (CLS_L if pL.fullmatch(ch) else 0) | (CLS_N if pN.fullmatch(ch) else 0) | (CLS_W if pW.fullmatch(ch) else 0)
        if cls[c] != want:
            diffs.append((c, cls[c], want, bool(unassigned.fullmatch(ch))))
    # the two Unicode databases may be of different versions: only code points one of them does not know may differ
    real = [d for d in diffs if not d[3] and d[1] != 0]
    assert not real, [(hex(c), a, b) for c, a, b, _ in real[:20]]
    assert len(diffs) < 600, len(diffs)
    # everything the reference's own tests reach (ASCII, Latin-1, Greek, Cyrillic) agrees exactly
    assert not [d for d in diffs if d[0] < 0x0530]


def test_case_tables_closed_forms():
    _, lower, upper = parse_header(os.path.join(ROOT, "docodo_b200", "csrc", "unicode_tables.h"), "DCD")
    for c in range(ord("A"), ord("Z") + 1):
        assert lower[c] == c + 32 and upper[c + 32] == c
    for c in range(0xC0, 0xDF):           # Latin-1 capitals except the multiplication sign
        if c != 0xD7:
            assert lower[c] == c + 32 and upper[c + 32] == c, hex(c)
    assert lower[0xD7] == 0xD7 and upper[0xF7] == 0xF7 and upper[0xDF] == 0xDF and upper[0xFF] == 0x178 and lower[0x178] == 0xFF
    for c in range(0x0410, 0x0430):       # Cyrillic А..Я
        assert lower[c] == c + 0x20 and upper[c + 0x20] == c
    for c in range(0x0400, 0x0410):       # Ѐ..Џ
        assert lower[c] == c + 0x50 and upper[c + 0x50] == c
    for c in range(0x0391, 0x03AA):       # Greek capitals
        if c != 0x03A2:
            assert lower[c] == c + 0x20, hex(c)
    for c in range(0x0531, 0x0557):       # Armenian
        assert lower[c] == c + 0x30 and upper[c + 0x30] == c
    assert lower[0x0130] == 0x69          # İ: the simple mapping .NET's invariant ToLower uses
    for c in list(range(0x30, 0x3A)) + [0x20, 0x5F, 0x2E]:
        assert lower[c] == c and upper[c] == c
