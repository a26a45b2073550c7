"""CPU tier: the C-ABI library loads and exports every symbol include/docodo_b200.h declares; the host build of the
device stemmer sources agrees with the oracle; creating a context without a GPU fails loudly (no fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import docodo_b200 as dcd
from docodo_b200 import capi
from oracle_py import STEM_EN, STEM_RU, OracleIndex

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    h = open(os.path.join(ROOT, "include", "docodo_b200.h")).read()
    h = re.sub(r"/\*.*?\*/", "", h, flags=re.S)
    return sorted(set(re.findall(r"\b(dcd_[a-z0-9_]+)\s*\(", h)))


def test_every_header_symbol_is_exported_and_bound():
    L = dcd.lib()
    syms = header_symbols()
    assert len(syms) >= 40
    for s in syms:
        assert hasattr(L, s), "missing export " + s
        assert s in capi.SIGNATURES, "python binding lacks " + s
    for s in capi.SIGNATURES:
        assert s in syms, "bound symbol not declared in the header: " + s


def test_library_is_sm100a_only():
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", dcd.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(dcd.DocodoError) as e:
        dcd.Index()
    assert e.value.code == -2   # DCD_E_NO_DEVICE


def test_english_stemmer_host_build_matches_oracle(data_dir):
    o = OracleIndex()
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()
    extra = ["skis", "skies", "dying", "only", "generously", "communion", "arsenal", "'twas", "yes", "say", "saying", "happy",
             "geology", "geologist", "inning", "proceeding", "agreed", "feed", "hopping", "hoping", "filing", "fizzed", "abc's'"]
    bad = []
    for w in words[::3] + extra:
        if len(w) > 47:
            continue
        a, b = dcd.stem_host(STEM_EN, w), o.stem(STEM_EN, w)
        if a != b:
            bad.append((w, a, b))
    assert not bad, bad[:20]


def test_english_stemmer_pinned_by_en_voc(data_dir, en_voc):
    """Known-answer test for the third-party stemmer (StemmersNet 1.1.1): en.voc keys are exactly the stems that stemmer
    produced for the FreeLing word list (Dict.cs:260-296), so every surface word must stem onto a key of en.voc."""
    import struct
    keys = set()
    i = 0
    b = en_voc
    while i < len(b):
        n = b[i]
        i += 1
        keys.add(b[i:i + n].decode("utf-8"))
        i += n + 4
    o = OracleIndex()
    words = open(os.path.join(data_dir, "en_words.txt")).read().split()
    miss = [w for w in words[::2] if o.stem(STEM_EN, w) not in keys]
    assert len(miss) == 0, miss[:20]
    miss = [w for w in words[1::2] if len(w) <= 47 and dcd.stem_host(STEM_EN, w) not in keys]
    assert len(miss) == 0, miss[:20]


def test_russian_stemmer_host_build_matches_oracle(data_dir):
    """Parity of the two independent restatements (the reference DLL itself is unpinned for Russian)."""
    import regex
    o = OracleIndex()
    text = open(os.path.join(data_dir, "voyna.cp1251.txt"), "rb").read().decode("cp1251").lower()
    words = sorted(set(w for w in regex.findall(r"\p{L}+", text) if 3 <= len(w) <= 32 and regex.fullmatch(r"[а-яё]+", w)))
    assert len(words) > 15000
    bad = []
    for w in words:
        a, b = dcd.stem_host(STEM_RU, w), o.stem(STEM_RU, w)
        if a != b:
            bad.append((w, a, b))
    assert not bad, bad[:20]
    # spot checks from the published Snowball sample vocabulary
    for w, s in (("важная", "важн"), ("бегали", "бега"), ("человеческого", "человеческ"), ("князь", "княз"), ("сказал", "сказа")):
        assert o.stem(STEM_RU, w) == s, (w, o.stem(STEM_RU, w))


def test_csharp_shim_declares_every_entry_point():
    """csharp/DocodoB200.cs is the reference-side binding: one DllImport per function of the header."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "docodo_b200.h")).read()
    shim = open(os.path.join(root, "csharp", "DocodoB200.cs")).read()
    declared = set(re.findall(r"\b(dcd_[a-z0-9_]+)\s*\(", header))
    imported = set(re.findall(r"extern\s+\w+\s+(dcd_[a-z0-9_]+)\s*\(", shim))
    assert declared and declared == imported, (sorted(declared - imported), sorted(imported - declared))
