"""ctypes binding of the CPU oracle (oracle/_build/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  The product package (docodo_b200) never does.
"""
import ctypes as C
import os
import struct
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_DIR = os.path.join(os.path.dirname(_HERE), "oracle")
_LIB = None


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
        if not os.path.exists(path):
            build_oracle()
        L = C.CDLL(path)
        L.orc_new.restype = C.c_void_p
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_add_voc_file.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_int]
        L.orc_add_voc_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint16, C.c_uint16, C.c_int]
        L.orc_set_stopwords.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_clear_stemmers.argtypes = [C.c_void_p]
        L.orc_set_caps.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_add_doc.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.orc_index_page.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        L.orc_builder_add.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint64]
        L.orc_builder_add_word.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint64]
        L.orc_builder_end_page.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64]
        L.orc_finish.argtypes = [C.c_void_p]
        L.orc_total_postings.argtypes = [C.c_void_p]
        L.orc_total_postings.restype = C.c_uint64
        L.orc_num_codes.argtypes = [C.c_void_p]
        L.orc_num_codes.restype = C.c_uint64
        for f in (L.orc_index_bytes, L.orc_list_bytes):
            f.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
            f.restype = C.c_size_t
        L.orc_load.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
        L.orc_build_bulk.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_int]
        L.orc_build_bulk.restype = C.c_double
        L.orc_search.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.orc_search.restype = C.c_size_t
        L.orc_encode.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.orc_encode.restype = C.c_size_t
        L.orc_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.orc_decode.restype = C.c_size_t
        L.orc_shift.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint64, C.POINTER(C.c_void_p)]
        L.orc_shift.restype = C.c_size_t
        L.orc_op.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t, C.c_int,
                             C.POINTER(C.c_int), C.POINTER(C.c_void_p)]
        L.orc_op.restype = C.c_size_t
        L.orc_stem.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.orc_stem.restype = C.c_size_t
        L.orc_word_codes.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        L.orc_word_codes.restype = C.c_size_t
        L.orc_lower.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_void_p)]
        L.orc_lower.restype = C.c_size_t
        L.orc_char_class.argtypes = [C.c_uint16]
        for f in (L.orc_like_words, L.orc_close_words):
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_int), C.POINTER(C.c_void_p)]
            f.restype = C.c_size_t
        L.orc_suggestions.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_void_p)]
        L.orc_suggestions.restype = C.c_size_t
        L.orc_histogram.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.orc_histogram.restype = C.c_size_t
        L.orc_vb_new.argtypes = [C.c_int]
        L.orc_vb_new.restype = C.c_void_p
        L.orc_vb_free.argtypes = [C.c_void_p]
        L.orc_vb_add_group.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_vb_build.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_vb_build.restype = C.c_size_t
        L.orc_synth_generate.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_uint64, C.c_int64, C.c_int, C.c_int,
                                         C.c_int64, C.c_int]
        L.orc_synth_generate.restype = C.c_void_p
        L.orc_synth_text.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.orc_synth_text.restype = C.c_uint64
        L.orc_synth_pages.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.orc_synth_pages.restype = C.c_uint64
        L.orc_synth_free.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


def synth_corpus(words, zipf_s, seed, ndocs, pages_per_doc=20, min_page_chars=200, first_doc=0, nthreads=1):
    """CPU twin of Index.synth_generate (oracle/oracle_synth.cpp): returns (text u8, page_offs u64, doc_first_page u64)."""
    L = lib()
    enc = [w.encode("ascii") for w in words]
    offs = np.zeros(len(enc) + 1, dtype=np.uint32)
    offs[1:] = np.cumsum([len(w) for w in enc])
    blob = np.frombuffer(b"".join(enc), dtype=np.uint8).copy()
    h = L.orc_synth_generate(blob.ctypes.data_as(C.c_void_p), offs.ctypes.data_as(C.c_void_p), len(enc), zipf_s, seed, ndocs,
                             pages_per_doc, min_page_chars, first_doc, nthreads)
    tp, pp, dp = C.c_void_p(), C.c_void_p(), C.c_void_p()
    nt = L.orc_synth_text(h, C.byref(tp))
    npg = L.orc_synth_pages(h, C.byref(pp), C.byref(dp))
    text = np.frombuffer(C.string_at(tp, nt), dtype=np.uint8).copy() if nt else np.zeros(0, np.uint8)
    page_offs = np.frombuffer(C.string_at(pp, 8 * (npg + 1)), dtype=np.uint64).copy()
    dfp = np.frombuffer(C.string_at(dp, 8 * (ndocs + 1)), dtype=np.uint64).copy()
    L.orc_synth_free(h)
    return text, page_offs, dfp


def u16(s):
    """str -> numpy uint16 array of UTF-16 code units."""
    return np.frombuffer(s.encode("utf-16-le", "surrogatepass"), dtype=np.uint16).copy()


def from_u16(a):
    return np.asarray(a, dtype=np.uint16).tobytes().decode("utf-16-le", "surrogatepass")


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and len(a) else None


def pack_strings(strs):
    arrs = [u16(s) for s in strs]
    offs = np.zeros(len(arrs) + 1, dtype=np.uint32)
    if arrs:
        offs[1:] = np.cumsum([len(a) for a in arrs])
    chars = np.concatenate(arrs) if arrs and offs[-1] else np.zeros(1, dtype=np.uint16)
    return chars, offs


STEM_NONE, STEM_EN, STEM_RU = 0, 1, 2


class SearchResult:
    def __init__(self, buf):
        o = 0

        def i32():
            nonlocal o
            v = struct.unpack_from("<i", buf, o)[0]
            o += 4
            return v

        def f32():
            nonlocal o
            v = struct.unpack_from("<f", buf, o)[0]
            o += 4
            return v

        def s16():
            nonlocal o
            n = i32()
            s = buf[o:o + 2 * n].decode("utf-16-le", "surrogatepass")
            o += 2 * n
            return s

        self.success = bool(i32())
        n = i32()
        self.error = buf[o:o + n].decode()
        o += n
        self.foundPages = []
        for _ in range(i32()):
            pi = i32()
            pid = s16()
            rank = f32()
            npos = i32()
            pos = list(struct.unpack_from("<%di" % npos, buf, o))
            o += 4 * npos
            self.foundPages.append({"page_index": pi, "id": pid, "rank": rank, "pos": pos})
        self.foundDocs = []
        for _ in range(i32()):
            d = i32()
            name = s16()
            rank = f32()
            npg = i32()
            pages = list(struct.unpack_from("<%di" % npg, buf, o))
            o += 4 * npg
            self.foundDocs.append({"doc": d, "name": name, "rank": rank, "pages": pages})
        self.words = []
        for _ in range(i32()):
            w = s16()
            self.words.append({"word": w, "nFound": i32()})


class OracleIndex:
    """Mirror of the slice of Docodo.Index the path needs, backed by the C++ oracle."""

    def __init__(self):
        self.L = lib()
        self.h = C.c_void_p(self.L.orc_new())

    def __del__(self):
        try:
            self.L.orc_free(self.h)
        except Exception:
            pass

    # configuration
    def add_voc_file(self, data, stemmer=STEM_EN):
        assert self.L.orc_add_voc_file(self.h, data, len(data), stemmer) == 0

    def add_voc_pairs(self, pairs, begin="a", end="z", stemmer=STEM_NONE):
        chars, offs = pack_strings([k for k, _ in pairs])
        vals = np.array([v for _, v in pairs], dtype=np.int32)
        self.L.orc_add_voc_pairs(self.h, _p(chars), _p(offs), _p(vals), len(pairs), ord(begin), ord(end), stemmer)

    def set_stopwords(self, words):
        chars, offs = pack_strings(words)
        self.L.orc_set_stopwords(self.h, _p(chars), _p(offs), len(words))

    def clear_stemmers(self):
        self.L.orc_clear_stemmers(self.h)

    def set_caps(self, max_pages, max_docs):
        self.L.orc_set_caps(self.h, max_pages, max_docs)

    # build
    def add_doc(self, source, name):
        a, b = u16(source), u16(name)
        self.L.orc_add_doc(self.h, _p(a), len(a), _p(b), len(b))

    def index_page(self, pid, text):
        a, b = u16(pid), u16(text)
        self.L.orc_index_page(self.h, _p(a), len(a), _p(b), len(b))

    def builder_add(self, code, coord):
        a = u16(code)
        self.L.orc_builder_add(self.h, _p(a), len(a), coord)

    def builder_add_word(self, word, coord):
        a = u16(word)
        self.L.orc_builder_add_word(self.h, _p(a), len(a), coord)

    def builder_end_page(self, pid, coord=None):
        a = u16(pid)
        self.L.orc_builder_end_page(self.h, _p(a), len(a), 0 if coord is None else 1, coord or 0)

    def finish(self):
        self.L.orc_finish(self.h)

    def build_bulk(self, text, enc, page_offs, doc_first_page, nthreads=1):
        """text: bytes/ndarray; enc 1 = single-byte, 2 = UTF-16.  Returns seconds."""
        page_offs = np.ascontiguousarray(page_offs, dtype=np.uint64)
        doc_first_page = np.ascontiguousarray(doc_first_page, dtype=np.uint64)
        t = np.ascontiguousarray(text)
        return self.L.orc_build_bulk(self.h, t.ctypes.data_as(C.c_void_p), enc, _p(page_offs), len(page_offs) - 1,
                                     _p(doc_first_page), len(doc_first_page) - 1, nthreads)

    @property
    def total_postings(self):
        return self.L.orc_total_postings(self.h)

    @property
    def num_codes(self):
        return self.L.orc_num_codes(self.h)

    def index_bytes(self):
        p = C.c_void_p()
        n = self.L.orc_index_bytes(self.h, C.byref(p))
        return C.string_at(p, n) if n else b""

    def list_bytes(self):
        p = C.c_void_p()
        n = self.L.orc_list_bytes(self.h, C.byref(p))
        return C.string_at(p, n) if n else b""

    def digests(self, cuts_index=None, cuts_list=None):
        """sha256 of the `.index` / `.index.list` bytes without copying them into Python (bench.py --verify at config-3 size).
        cuts_*: optional slice boundaries [0, ..., len] -> one digest per slice (the rank slices of a sharded build).
        Returns (index_len, [index digests], list_len, [list digests])."""
        import hashlib
        out = []
        for fn, cuts in ((self.L.orc_index_bytes, cuts_index), (self.L.orc_list_bytes, cuts_list)):
            p = C.c_void_p()
            n = fn(self.h, C.byref(p))
            cuts = list(cuts) if cuts is not None else [0, n]
            ds = []
            for a, b in zip(cuts[:-1], cuts[1:]):
                a, b = min(a, n), min(b, n)
                ds.append(hashlib.sha256((C.c_ubyte * (b - a)).from_address(p.value + a) if b > a else b"").hexdigest())
            out += [n, ds]
        return tuple(out)

    def load(self, index_bytes, list_bytes):
        assert self.L.orc_load(self.h, index_bytes, len(index_bytes), list_bytes, len(list_bytes)) == 0

    # search
    def search(self, req, dist=None):
        a = u16(req)
        p = C.c_void_p()
        n = self.L.orc_search(self.h, _p(a), len(a), 0 if dist is None else 1, dist or 0, C.byref(p))
        return SearchResult(C.string_at(p, n))

    # dictionary-side queries (oracle_aux.cpp)
    def _strs(self, fn, text, *extra):
        a = u16(text)
        p, cnt = C.c_void_p(), C.c_int()
        n = fn(self.h, _p(a), len(a), *extra, C.byref(cnt), C.byref(p))
        s = C.string_at(p, 2 * n).decode("utf-16-le", "surrogatepass")
        return s.split("\0") if cnt.value else []

    def like_words(self, word):
        return self._strs(self.L.orc_like_words, word)

    def close_words(self, word):
        return self._strs(self.L.orc_close_words, word)

    def suggestions(self, req, n=10):
        return self._strs(self.L.orc_suggestions, req, n)

    def histogram(self, top=1000):
        p, pc, cnt = C.c_void_p(), C.c_void_p(), C.c_int()
        n = self.L.orc_histogram(self.h, top, C.byref(cnt), C.byref(p), C.byref(pc))
        keys = C.string_at(p, 2 * n).decode("utf-16-le", "surrogatepass").split("\0") if cnt.value else []
        vals = np.frombuffer(C.string_at(pc, 8 * cnt.value), dtype=np.uint64) if cnt.value else []
        return dict(zip(keys, [int(x) for x in vals]))

    # helpers
    def encode(self, values):
        v = np.ascontiguousarray(values, dtype=np.uint64)
        p = C.c_void_p()
        n = self.L.orc_encode(self.h, _p(v), len(v), C.byref(p))
        return np.frombuffer(C.string_at(p, 2 * n), dtype=np.uint16).copy()

    def decode(self, words):
        w = np.ascontiguousarray(words, dtype=np.uint16)
        p = C.c_void_p()
        n = self.L.orc_decode(self.h, _p(w), len(w), C.byref(p))
        return np.frombuffer(C.string_at(p, 8 * n), dtype=np.uint64).copy()

    def shift(self, words, s):
        w = np.ascontiguousarray(words, dtype=np.uint16)
        p = C.c_void_p()
        n = self.L.orc_shift(self.h, _p(w), len(w), s, C.byref(p))
        return np.frombuffer(C.string_at(p, 2 * n), dtype=np.uint16).copy()

    def op(self, op, a, Ra, b, Rb):
        a = np.ascontiguousarray(a, dtype=np.uint16)
        b = np.ascontiguousarray(b, dtype=np.uint16)
        p = C.c_void_p()
        R = C.c_int()
        n = self.L.orc_op(self.h, 0 if op == "*" else 1, _p(a), len(a), Ra, _p(b), len(b), Rb, C.byref(R), C.byref(p))
        return np.frombuffer(C.string_at(p, 2 * n), dtype=np.uint16).copy(), R.value

    def stem(self, kind, word):
        a = u16(word)
        p = C.c_void_p()
        n = self.L.orc_stem(self.h, kind, _p(a), len(a), C.byref(p))
        return C.string_at(p, 2 * n).decode("utf-16-le", "surrogatepass")

    def word_codes(self, word):
        a = u16(word)
        p = C.c_void_p()
        n = self.L.orc_word_codes(self.h, _p(a), len(a), C.byref(p))
        s = C.string_at(p, 2 * n).decode("utf-16-le", "surrogatepass")
        return s.split("\0") if n else []

    def lower(self, s, upper=False):
        a = u16(s)
        p = C.c_void_p()
        n = self.L.orc_lower(self.h, _p(a), len(a), 1 if upper else 0, C.byref(p))
        return C.string_at(p, 2 * n).decode("utf-16-le", "surrogatepass")


class OracleVocBuilder:
    """VocBuilder (Dict.cs:109-296) restated in the oracle."""

    def __init__(self, stemmer=STEM_EN):
        self.L = lib()
        self.h = C.c_void_p(self.L.orc_vb_new(stemmer))

    def __del__(self):
        try:
            self.L.orc_vb_free(self.h)
        except Exception:
            pass

    def AddWordsGroup(self, words):
        chars, offs = pack_strings(words)
        self.L.orc_vb_add_group(self.h, _p(chars), _p(offs), len(words))

    def build(self):
        n = self.L.orc_vb_build(self.h, None, 0)
        buf = (C.c_uint8 * max(n, 1))()
        self.L.orc_vb_build(self.h, buf, n)
        return bytes(buf[:n])
