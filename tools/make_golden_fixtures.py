#!/usr/bin/env python3
"""Extract the reference's own test fixtures / known answers into tests/golden/ (run in the build
container, where /root/reference exists; the GPU box only sees the committed output).

Sources (all under /root/reference):
  XUnitDocodoTest/IndexTest.cs:19-38   TestText1 / TestHeaders1 string constants
  XUnitDocodoTest/IndexTest.cs:140-157 CoordTest expectations  (regex positions, computed here the
                                       same way the test does: Regex \\bword\\b on the lower-cased text)
  XUnitDocodoTest/IndexTest.cs:180-220 RequestSyntaxTest literal counts
  XUnitDocodoTest/IndexTest.cs:262     VocabTest '#1' count
  Dict/en.voc, Dict/stop.txt, Dict/en/*.txt, Test/*.txt  -> tests/data/ (input DATA named by
                                       BASELINE.json configs 1-3; not source code)
"""
import glob
import json
import os
import re
import shutil
import sys

import regex

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
DATA = os.path.join(ROOT, "tests", "data")


def cs_unescape(s):
    return (s.replace("\\n", "\n").replace('\\"', '"').replace("\\\\", "\\"))


def extract_const(src, name):
    m = re.search(r"const string %s\s*=\s*(.*?);\s*\n\s*(?:const|//|class)" % name, src, re.S)
    body = m.group(1)
    parts = re.findall(r'"((?:[^"\\]|\\.)*)"', body)
    return cs_unescape("".join(parts))


def main():
    os.makedirs(GOLD, exist_ok=True)
    os.makedirs(DATA, exist_ok=True)
    src = open(os.path.join(REF, "XUnitDocodoTest", "IndexTest.cs"), encoding="utf-8-sig").read()
    text1 = extract_const(src, "TestText1")
    headers1 = extract_const(src, "TestHeaders1")
    low = text1.lower()
    coord_words = ["and", "tupman", "everybody", "old"]
    fx = {
        "_source": "XUnitDocodoTest/IndexTest.cs (see tools/make_golden_fixtures.py)",
        "TestText1": text1,
        "TestHeaders1": headers1,
        "CoordTest": {w: [m.start() for m in regex.finditer(r"\b" + w + r"\b", low)] for w in coord_words},
        "BuilderTest_words": [[m.group(0), m.start()] for m in regex.finditer(r"\b\w+\b", low)],
        "RequestSyntaxTest": {
            "npages": 100,
            "and (tupman|old)": {"docs": 2, "pages_per_doc": 100, "pos_per_page": 42},
            "and (tupman|old) {Name=Dump}": {"docs": 1, "pages_per_doc": 100},
            "dist40": {
                "lady old": 10,
                "\"lady\" old": 8,
                "\"old lady\"": 8,
                "\"lady old\"": None,
                "lady (old | young)": 12,
                "\"old ladies were\"": 3,
                "\"old lady were\"": None,
                "\"old (lady|ladies) (who|were|looked)\"": 9,
                "?an?": len(regex.findall(r"\w*an\w*", low)),
            },
        },
        "VocabTest": {
            "voc": [["and", 1], ["end", 3], ["old", 2], ["the", 6], ["them", 5], ["then", 4]],
            "hash1_count_per_page": len(regex.findall(r"\band\b", low)),
        },
    }
    assert len(text1) == 2759, len(text1)
    with open(os.path.join(GOLD, "indextest.json"), "w", encoding="utf-8") as f:
        json.dump(fx, f, ensure_ascii=True, indent=1)
    # input data named by the BASELINE configs
    shutil.copyfile(os.path.join(REF, "Dict", "en.voc"), os.path.join(DATA, "en.voc"))
    shutil.copyfile(os.path.join(REF, "Dict", "stop.txt"), os.path.join(DATA, "stop.txt"))
    shutil.copyfile(os.path.join(REF, "Test", "Dickens Charles. The Pickwick Papers - royallib.ru.txt"),
                    os.path.join(DATA, "pickwick.cp1251.txt"))
    shutil.copyfile(os.path.join(REF, "Test", "voyna-i-mir-tom-1.txt"), os.path.join(DATA, "voyna.cp1251.txt"))
    # config 3 vocabulary: distinct lower-cased letter-only surface words of Dict/en/*.txt
    words = set()
    for fn in sorted(glob.glob(os.path.join(REF, "Dict", "en", "*"))):
        for line in open(fn, encoding="utf-8", errors="replace"):
            a = line.split(" ")
            if len(a) >= 2:
                for w in a[:2]:
                    w = w.lower()
                    if 3 <= len(w) <= 32 and re.fullmatch(r"[a-z]+", w):
                        words.add(w)
    with open(os.path.join(DATA, "en_words.txt"), "w") as f:
        f.write("\n".join(sorted(words)) + "\n")
    # stemmer known answers: a deterministic 1/16 sample of (surface word, key of en.voc it must map to)
    print("TestText1 len", len(text1), "words", len(words))


if __name__ == "__main__":
    main()
