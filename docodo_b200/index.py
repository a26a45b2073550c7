"""Host-side mirror of the slice of Docodo.NET's public API that the B200 path sits behind.

Same member names, argument meaning and error behaviour as the C# classes (file:line into
/root/reference/Docodo.NET): ``Index`` (Index.cs:94-140, Search.cs:430-617), ``Index.Builder`` (Build.cs:258-437),
``IndexSequence`` (IndexSequence.cs), ``Vocab`` (Dict.cs:13-104), ``SearchOptions`` (Search.cs:262-276), the result types
(Search.cs:20-123).  All work is done by libdocodo_b200.so through the C ABI; nothing here computes postings.
"""
import ctypes as C
import os

import numpy as np

from . import capi
from .capi import DocodoError, check, dcd_counters, dcd_options, dcd_synth, lib

STEM_NONE, STEM_EN, STEM_RU = 0, 1, 2
ENC_LATIN1, ENC_UTF16, ENC_UTF8 = 1, 2, 3
_KNOWN_STEMMERS = {"digit": 0, "ru": 1, "en": 2, "de": 3, "fr": 4}   # Index.KnownStemmers order, Index.cs:175-181


def _u16(s):
    return np.frombuffer(s.encode("utf-16-le", "surrogatepass"), dtype=np.uint16).copy()


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def _pack(strs):
    arrs = [_u16(s) for s in strs]
    offs = np.zeros(len(arrs) + 1, dtype=np.uint32)
    if arrs:
        offs[1:] = np.cumsum([len(a) for a in arrs])
    chars = np.concatenate(arrs) if arrs and offs[-1] else np.zeros(1, dtype=np.uint16)
    return chars, offs


def _str_at(p, n):
    return C.string_at(p, 2 * n).decode("utf-16-le", "surrogatepass") if n else ""


def stem_host(stemmer, word):
    """IStemmer.Stem evaluated by the host build of the device stemmer source (no GPU needed)."""
    a = _u16(word)
    out = np.zeros(max(len(a), 1) + 2, dtype=np.uint16)
    n = C.c_int32()
    st = lib().dcd_stem_host(stemmer, _ptr(a), len(a), _ptr(out), len(out), C.byref(n))
    if st != 0:
        raise DocodoError(st, "dcd_stem_host")
    return out[:n.value].tobytes().decode("utf-16-le", "surrogatepass")


class Vocab(dict):
    """Docodo.Vocab (Dict.cs:13-104): stem -> group id, a first-letter Range and a stemmer."""

    def __init__(self, filename=None, data=None, stemmer=STEM_NONE, name=None):
        super().__init__()
        self.Range = ("\0", "\0")
        self.Name = name
        self.stemmer = stemmer
        self._bytes = None
        if filename is not None:
            lang = os.path.basename(filename).split(".")[0]   # Dict.cs:35-37
            self.Name = lang
            self.stemmer = {"en": STEM_EN, "ru": STEM_RU}.get(lang, STEM_NONE)
            self._bytes = open(filename, "rb").read()
        elif data is not None:
            self._bytes = bytes(data)


class SearchOptions:
    """Index.SearchOptions (Search.cs:262-276)."""

    def __init__(self, dist=0, DoCorrection=False, RemoveWordBreaks=True):
        self.dist = dist
        self.DoCorrection = DoCorrection
        self.RemoveWordBreaks = RemoveWordBreaks


class ResultDocPage:
    """Index.ResultDocPage (Search.cs:92-123)."""

    def __init__(self, id, pos, rank, ordinal):
        self.id = id
        self.pos = pos
        self.rank = rank
        self.ordinal = ordinal
        self.text = None

    def __eq__(self, other):
        return isinstance(other, ResultDocPage) and self.id == other.id and self.pos == other.pos


class ResultDocument:
    """Index.ResultDocument (Search.cs:59-90)."""

    def __init__(self, name, rank, pages):
        self.Name = name
        self.rank = rank
        self.pages = pages

    def __eq__(self, other):
        return isinstance(other, ResultDocument) and self.Name == other.Name

    def __hash__(self):
        return hash(self.Name)


class SearchResult:
    """Index.SearchResult / ErrorSearchResult (Search.cs:20-56)."""

    def __init__(self):
        self.foundDocs = []
        self.foundPages = []
        self.Success = True
        self.Error = ""
        self.words = []

    def __eq__(self, other):   # Search.cs:28-37
        return isinstance(other, SearchResult) and self.foundPages == other.foundPages


class IndexSequence:
    """Docodo.IndexSequence (IndexSequence.cs): ascending ulongs stored as delta + 15-bit-group ushorts.
    Encode / decode / '*' / '+' run on the device through dcd_seq_*; R carries the distance / order convention."""

    def __init__(self, index, words=None, R=0):
        self._ix = index
        self.self = np.asarray(words if words is not None else [], dtype=np.uint16)
        self.R = R

    @property
    def Count(self):   # number of ushorts, IndexSequence.cs:45
        return int(self.self.size)

    @classmethod
    def from_values(cls, index, values, R=0):   # new Builder().AddRange(values)
        v = np.ascontiguousarray(values, dtype=np.uint64)
        out = np.zeros(max(5 * len(v), 1), dtype=np.uint16)
        n = C.c_uint64()
        check(index._ctx, lib().dcd_seq_encode(index._ctx, _ptr(v), len(v), _ptr(out), len(out), C.byref(n)))
        return cls(index, out[:n.value].copy(), R)

    def values(self):   # enumerator, IndexSequence.cs:104-154
        w = np.ascontiguousarray(self.self)
        out = np.zeros(max(len(w), 1), dtype=np.uint64)
        n = C.c_uint64()
        check(self._ix._ctx, lib().dcd_seq_decode(self._ix._ctx, _ptr(w), len(w), _ptr(out), len(out), C.byref(n)))
        return out[:n.value].copy()

    def _op(self, other, op):
        a, b = np.ascontiguousarray(self.self), np.ascontiguousarray(other.self)
        cap = 5 * (len(a) + len(b)) + 8
        out = np.zeros(cap, dtype=np.uint16)
        n = C.c_uint64()
        R = C.c_int32()
        check(self._ix._ctx, lib().dcd_seq_op(self._ix._ctx, op, _ptr(a), len(a), self.R, _ptr(b), len(b), other.R, _ptr(out), cap,
                                              C.byref(n), C.byref(R)))
        return IndexSequence(self._ix, out[:n.value].copy(), R.value)

    def __mul__(self, other):   # operator *, IndexSequence.cs:218-284
        return self._op(other, 0)

    __and__ = __mul__           # operator &, IndexSequence.cs:205-215

    def __add__(self, other):   # operator +, IndexSequence.cs:286-322
        return self._op(other, 1)


class Index:
    """Docodo.Index restricted to the build + query path, executed on one B200.

    >>> ix = Index()
    >>> ix.AddVoc(Vocab("Dict/en.voc")); ix.LoadStopWords("Dict/stop.txt")
    >>> b = ix.GetBuilder(); b.AddDoc("src", "doc"); b.AddPage("1", text); b.Build()
    >>> ix.Search("old lady", SearchOptions(dist=40)).foundPages[0].pos
    """

    def __init__(self, path=None, InMemory=False, vocs=None, device=0):
        L = lib()
        ctx = C.c_void_p()
        st = L.dcd_ctx_create(device, C.byref(ctx))
        if st != 0:
            msg = L.dcd_last_error(None)
            raise DocodoError(st, msg.decode() if msg else "dcd_ctx_create failed")
        self._ctx = ctx
        self._opt = dcd_options()
        L.dcd_options_default(C.byref(self._opt))
        self.WorkPath = path
        self.InMemory = InMemory
        self.vocs = []
        self.stopWords = set()
        self.CanSearch = False
        for v in vocs or []:
            self.AddVoc(v)
        if path is not None and os.path.exists(os.path.join(path, ".index")):
            self.Load()

    def Dispose(self):
        if getattr(self, "_ctx", None):
            lib().dcd_ctx_destroy(self._ctx)
            self._ctx = None

    close = Dispose

    def __del__(self):
        try:
            self.Dispose()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.Dispose()

    # ---- tunables (public fields of Docodo.Index, Index.cs:96-115,154,183) ----
    def _push_options(self):
        check(self._ctx, lib().dcd_options_set(self._ctx, C.byref(self._opt)))

    def _opt_prop(name):
        def g(self):
            return getattr(self._opt, name)

        def s(self, v):
            setattr(self._opt, name, int(v))
            self._push_options()
        return property(g, s)

    MIN_WORD_LENGTH = _opt_prop("min_word_length")
    MAX_WORD_LENGTH = _opt_prop("max_word_length")
    MAX_FOUND_PAGES = _opt_prop("max_found_pages")
    MAX_FOUND_DOCS = _opt_prop("max_found_docs")
    bKeepForms = _opt_prop("keep_forms")
    dict_capacity_log2 = _opt_prop("dict_capacity_log2")

    class _Stemmers:
        def __init__(self, ix):
            self._ix = ix

        def Clear(self):   # index.Stemmers.Clear(), IndexTest.cs:135
            self._ix._opt.stemmers_mask = 0
            self._ix._push_options()

        def Set(self, langs):
            m = 0
            for l in langs:
                m |= 1 << _KNOWN_STEMMERS[l]
            self._ix._opt.stemmers_mask = m
            self._ix._push_options()

    @property
    def Stemmers(self):
        return Index._Stemmers(self)

    # ---- configuration ----
    def AddVoc(self, voc):   # Index.cs:221
        L = lib()
        if voc._bytes is not None:
            check(self._ctx, L.dcd_vocab_load(self._ctx, voc._bytes, len(voc._bytes), voc.stemmer))
        else:
            keys = list(voc.keys())
            chars, offs = _pack(keys)
            vals = np.array([voc[k] for k in keys], dtype=np.int32)
            check(self._ctx, L.dcd_vocab_add_pairs(self._ctx, _ptr(chars), _ptr(offs), _ptr(vals), len(keys), ord(voc.Range[0]),
                                                   ord(voc.Range[1]), voc.stemmer))
        self.vocs.append(voc)

    def LoadStopWords(self, file):   # Index.cs:227-230
        lines = open(file, encoding="utf-8-sig").read().splitlines()
        self.stopWords = set(s for s in lines if len(s.strip(" ")) > 0 and ";" not in s)
        self._push_stop()

    def AddStopWords(self, words):   # Index.cs:231-235
        self.stopWords |= set(words)
        self._push_stop()

    def _push_stop(self):
        words = sorted(self.stopWords)
        chars, offs = _pack(words)
        check(self._ctx, lib().dcd_stopwords_set(self._ctx, _ptr(chars), _ptr(offs), len(words)))

    # ---- build ----
    def GetBuilder(self):   # Build.cs:443
        return Index.Builder(self)

    class Builder:
        """Index.Builder (Build.cs:258-437) + the page loop of IndexTask (Build.cs:477-536) as AddPage."""

        def __init__(self, parent):
            self.Parent = parent
            b = C.c_void_p()
            check(parent._ctx, lib().dcd_build_begin(parent._ctx, C.byref(b)))
            self._b = b

        def _ck(self, st):
            check(self.Parent._ctx, st)

        def SetOrigin(self, coord_base, doc_number_base=0):
            """Shard origin (dcd_build_set_origin): the coordinate shift of MergeFiles(shift_coords), Index.cs:584-593."""
            self._ck(lib().dcd_build_set_origin(self._b, coord_base, doc_number_base))

        def AddDoc(self, sourceid, name):   # Build.cs:348-352
            if len(sourceid) == 0:
                raise Exception("sourceid must not be empty")
            a, b = _u16(sourceid), _u16(name)
            self._ck(lib().dcd_build_add_doc(self._b, _ptr(a), len(a), _ptr(b), len(b)))

        def AddPage(self, id, text):   # one IndexPage through Build.cs:477-536
            a, t = _u16(id), _u16(text)
            self._ck(lib().dcd_build_add_page(self._b, _ptr(a), len(a), _ptr(t), len(t)))

        def AddPagesBulk(self, text, encoding, page_offs, doc_first_page, source="doc", on_device=False):
            """Many documents at once (ids "1".."k", names "<source>:d<n>").  text/page_offs/doc_first_page are numpy arrays,
            or raw device pointers (ints) when on_device."""
            s = _u16(source)
            if on_device:
                tp, po, dp, npages, ndocs = text
                self._ck(lib().dcd_build_add_pages_bulk(self._b, tp, encoding, po, npages, dp, ndocs, _ptr(s), len(s), 1))
                return
            t = np.ascontiguousarray(text)
            po = np.ascontiguousarray(page_offs, dtype=np.uint64)
            dp = np.ascontiguousarray(doc_first_page, dtype=np.uint64)
            self._ck(lib().dcd_build_add_pages_bulk(self._b, t.ctypes.data_as(C.c_void_p), encoding, _ptr(po), len(po) - 1, _ptr(dp),
                                                    len(dp) - 1, _ptr(s), len(s), 0))

        def AddWord(self, word, coord):   # Build.cs:302-316
            a = _u16(word)
            self._ck(lib().dcd_build_add_word(self._b, _ptr(a), len(a), coord))

        def Add(self, code, coord):   # Build.cs:324-341
            a = _u16(code)
            self._ck(lib().dcd_build_add_posting(self._b, _ptr(a), len(a), coord))

        def EndPage(self, id, maxcoord=None):   # Build.cs:359-367
            a = _u16(id)
            self._ck(lib().dcd_build_end_page(self._b, _ptr(a), len(a), 0 if maxcoord is None else 1, maxcoord or 0))

        def Build(self):   # Build.cs:407-434
            b, self._b = self._b, None
            self._ck(lib().dcd_build_finish(b))
            self.Parent.CanSearch = True
            if self.Parent.WorkPath:
                os.makedirs(self.Parent.WorkPath, exist_ok=True)
                self.Parent.Save(self.Parent.WorkPath)
            return self.Parent

        def BuildSharded(self, comm):
            """Collective finish of a document-range sharded build (dcd_build_finish_sharded): every rank of `comm` calls it.
            Stands in for MergeAll / MergeFiles(shift_coords) / MergeIndexes (Index.cs:537-752)."""
            b, self._b = self._b, None
            self._ck(lib().dcd_build_finish_sharded(b, comm._h))
            self.Parent.CanSearch = True
            return self.Parent

        def __del__(self):
            if getattr(self, "_b", None):
                lib().dcd_build_abort(self._b)

    def Create(self, docs):
        """CreateAsync (Index.cs:424-534) over an in-memory document set: [(source, name, [(page_id, text), ...]), ...]
        in indexing order (MaxDegreeOfParallelism = 1 semantics)."""
        b = self.GetBuilder()
        for source, name, pages in docs:
            b.AddDoc(source, name)
            for pid, text in pages:
                b.AddPage(pid, text)
        return b.Build()

    # ---- persistence ----
    def serialize(self):
        """(bytes of `.index`, bytes of `.index.list`) for the resident index."""
        L = lib()
        ni, nl = C.c_uint64(), C.c_uint64()
        check(self._ctx, L.dcd_index_serialized_size(self._ctx, C.byref(ni), C.byref(nl)))
        ib = np.zeros(max(ni.value, 1), dtype=np.uint8)
        lb = np.zeros(max(nl.value, 1), dtype=np.uint8)
        check(self._ctx, L.dcd_index_serialize(self._ctx, _ptr(ib), ni.value, _ptr(lb), nl.value))
        return ib[:ni.value].tobytes(), lb[:nl.value].tobytes()

    def serialize_merged(self):
        """This rank's slice of the merged `.index` after BuildSharded (concatenate the slices in rank order)."""
        L = lib()
        ni = C.c_uint64()
        check(self._ctx, L.dcd_merged_serialized_size(self._ctx, C.byref(ni)))
        ib = np.zeros(max(ni.value, 1), dtype=np.uint8)
        check(self._ctx, L.dcd_merged_serialize(self._ctx, _ptr(ib), ni.value))
        return ib[:ni.value].tobytes()

    def Save(self, path):
        check(self._ctx, lib().dcd_index_save(self._ctx, path.encode()))

    def Load(self, path=None):   # Index.cs:312-380
        p = path or self.WorkPath
        st = lib().dcd_index_load(self._ctx, p.encode())
        self.CanSearch = st == 0
        return st == 0

    def LoadBytes(self, index_bytes, list_bytes):
        check(self._ctx, lib().dcd_index_load_bytes(self._ctx, index_bytes, len(index_bytes), list_bytes, len(list_bytes)))
        self.CanSearch = True

    def __getitem__(self, key):   # Index.this[string], Index.cs:255-263
        a = _u16(key)
        p = C.c_void_p()
        n = C.c_int32()
        st = lib().dcd_index_sequence(self._ctx, _ptr(a), len(a), C.byref(p), C.byref(n))
        if st != 0:
            raise KeyError(key)
        words = np.frombuffer(C.string_at(p, 2 * n.value), dtype=np.uint16).copy() if n.value else np.zeros(0, np.uint16)
        return IndexSequence(self, words)

    def set_stream(self, cuda_stream):
        """Run the context on the caller's CUDA stream (int handle, e.g. torch.cuda.current_stream().cuda_stream)."""
        check(self._ctx, lib().dcd_ctx_set_stream(self._ctx, cuda_stream))

    @property
    def counters(self):
        k = dcd_counters()
        check(self._ctx, lib().dcd_counters_get(self._ctx, C.byref(k)))
        return k

    @property
    def Count(self):   # Index.cs:253
        return int(self.counters.codes)

    @property
    def MaxCoord(self):   # Index.cs:310
        return int(self.counters.max_coord)

    def GetWordCodes(self, word):   # Build.cs:175-247
        a = _u16(word)
        out = np.zeros(512, dtype=np.uint16)
        nu, nc = C.c_int32(), C.c_int32()
        check(self._ctx, lib().dcd_word_codes(self._ctx, _ptr(a), len(a), _ptr(out), len(out), C.byref(nu), C.byref(nc)))
        s = out[:nu.value].tobytes().decode("utf-16-le", "surrogatepass")
        return s.split("\0") if nc.value else []

    # ---- dictionary-side queries (device scans over the key table) ----
    def _strings(self, fn, *args):
        nu, nc = C.c_int32(), C.c_int32()
        check(self._ctx, fn(self._ctx, *args, None, 0, C.byref(nu), C.byref(nc)))
        out = np.zeros(max(nu.value, 1), dtype=np.uint16)
        check(self._ctx, fn(self._ctx, *args, _ptr(out), len(out), C.byref(nu), C.byref(nc)))
        s = out[:nu.value].tobytes().decode("utf-16-le", "surrogatepass")
        return s.split("\0") if nc.value else []

    def GetLikeWords(self, word):   # Search.cs:160-167
        a = _u16(word)
        return self._strings(lib().dcd_like_words, _ptr(a), len(a))

    def GetCloseWords(self, word):   # Search.cs:169-174
        a = _u16(word)
        return self._strings(lib().dcd_close_words, _ptr(a), len(a))

    def GetSuggessions(self, req, n=10):   # Search.cs:176-188
        a = _u16(req)
        return self._strings(lib().dcd_suggestions, _ptr(a), len(a), n)

    def CalcHistogram(self, top=1000):   # Index.CalcHistogram(index), Index.cs:284-307 -> {display key: count}
        L = lib()
        nu, nc = C.c_int32(), C.c_int32()
        check(self._ctx, L.dcd_histogram(self._ctx, top, None, 0, C.byref(nu), None, C.byref(nc)))
        out = np.zeros(max(nu.value, 1), dtype=np.uint16)
        cnt = np.zeros(max(nc.value, 1), dtype=np.uint64)
        check(self._ctx, L.dcd_histogram(self._ctx, top, _ptr(out), len(out), C.byref(nu), _ptr(cnt), C.byref(nc)))
        keys = out[:nu.value].tobytes().decode("utf-16-le", "surrogatepass").split("\0") if nc.value else []
        return dict(zip(keys, [int(x) for x in cnt[:nc.value]]))

    # ---- search ----
    def page_info(self, ordinal):
        L = lib()
        pid, nid, doc, st, en = C.c_void_p(), C.c_int32(), C.c_int64(), C.c_uint64(), C.c_uint64()
        check(self._ctx, L.dcd_index_page_info(self._ctx, ordinal, C.byref(pid), C.byref(nid), C.byref(doc), C.byref(st), C.byref(en)))
        return _str_at(pid, nid.value), doc.value, st.value, en.value

    def doc_name(self, ordinal):
        p, n = C.c_void_p(), C.c_int32()
        check(self._ctx, lib().dcd_index_doc_name(self._ctx, ordinal, C.byref(p), C.byref(n)))
        return _str_at(p, n.value)

    def SearchBatch(self, reqs, opts=None, comm=None):
        """Evaluate many requests in one device batch; opts is None, one SearchOptions or a list (None entries allowed).
        comm (ShardComm): collective, EXACT serving from the document-range shards of BuildSharded -- every rank calls it with
        the same batch; the per-rank results concatenate in rank order to Index.Search on the merged index."""
        L = lib()
        if opts is None or isinstance(opts, SearchOptions):
            opts = [opts] * len(reqs)
        chars, offs = _pack(reqs)
        has = np.array([0 if o is None else 1 for o in opts], dtype=np.int32)
        dist = np.array([0 if o is None else o.dist for o in opts], dtype=np.int32)
        h = C.c_void_p()
        if comm is not None:
            check(self._ctx, L.dcd_search_batch_sharded(self._ctx, comm._h, _ptr(chars), _ptr(offs), len(reqs), _ptr(has), _ptr(dist),
                                                        C.byref(h)))
        else:
            check(self._ctx, L.dcd_search_batch(self._ctx, _ptr(chars), _ptr(offs), len(reqs), _ptr(has), _ptr(dist), C.byref(h)))
        try:
            return [self._result(h, q) for q in range(len(reqs))]
        finally:
            L.dcd_hits_free(h)

    def SearchBatchFlat(self, reqs, opts=None):
        """Same evaluation as SearchBatch, results as flat numpy arrays per request (no per-page Python objects):
        {"status", "ordinals" int64, "ranks" float32, "npos" int32, "pos" int32 (all pages back to back),
        "docs" [(doc ordinal, rank, first page, pages)]} -- for callers that compare or ship whole results."""
        L = lib()
        if opts is None or isinstance(opts, SearchOptions):
            opts = [opts] * len(reqs)
        chars, offs = _pack(reqs)
        has = np.array([0 if o is None else 1 for o in opts], dtype=np.int32)
        dist = np.array([0 if o is None else o.dist for o in opts], dtype=np.int32)
        h = C.c_void_p()
        check(self._ctx, L.dcd_search_batch(self._ctx, _ptr(chars), _ptr(offs), len(reqs), _ptr(has), _ptr(dist), C.byref(h)))
        try:
            total = C.c_uint64()
            counts = np.zeros(max(len(reqs), 1), dtype=np.int32)
            check(self._ctx, L.dcd_hits_export_pages(h, _ptr(counts), None, None, None, 0, C.byref(total)))
            n = max(total.value, 1)
            ordinals, ranks, npos = np.zeros(n, np.int64), np.zeros(n, np.float32), np.zeros(n, np.int32)
            check(self._ctx, L.dcd_hits_export_pages(h, None, _ptr(ordinals), _ptr(ranks), _ptr(npos), n, None))
            out, at = [], 0
            for q in range(len(reqs)):
                k = int(counts[q])
                tp = C.c_uint64()
                check(self._ctx, L.dcd_hits_export_positions(h, q, None, 0, C.byref(tp)))
                pos = np.zeros(max(tp.value, 1), dtype=np.int32)
                check(self._ctx, L.dcd_hits_export_positions(h, q, _ptr(pos), tp.value, None))
                docs = []
                for j in range(L.dcd_hits_num_docs(h, q)):
                    d, rank, fp, npg = C.c_int64(), C.c_float(), C.c_int32(), C.c_int32()
                    L.dcd_hits_doc(h, q, j, C.byref(d), C.byref(rank), C.byref(fp), C.byref(npg))
                    docs.append((d.value, rank.value, fp.value, npg.value))
                out.append({"status": L.dcd_hits_status(h, q), "ordinals": ordinals[at:at + k], "ranks": ranks[at:at + k],
                            "npos": npos[at:at + k], "pos": pos[:tp.value], "docs": docs})
                at += k
            return out
        finally:
            L.dcd_hits_free(h)

    def SearchBatchPages(self, reqs, opts=None, page_base=0, out=None, device_out=None, comm=None):
        """Sharded serving: evaluate the batch on this shard and return (counts [nreq] int32, pairs [total, 2] int32) -- per
        request the found pages as (page ordinal + page_base, positions in page), the payload sharding.gather_pages_flat
        moves.  out: optional callable (numel) -> 1-d int32 CPU tensor to export into (page-locked staging).
        device_out: optional callable (numel) -> 1-d int32 CUDA tensor on this context's device: the pages then never
        leave the device and no per-request results are shaped on the host (dcd_search_batch_pages / dcd_search_pages_fetch)."""
        L = lib()
        if opts is None or isinstance(opts, SearchOptions):
            opts = [opts] * len(reqs)
        chars, offs = _pack(reqs)
        has = np.array([0 if o is None else 1 for o in opts], dtype=np.int32)
        dist = np.array([0 if o is None else o.dist for o in opts], dtype=np.int32)
        if device_out is not None:
            total = C.c_uint64()
            counts = np.zeros(len(reqs), dtype=np.int32)
            if comm is not None:   # exact (collective): halo exchange between the shards
                check(self._ctx, L.dcd_search_batch_pages_sharded(self._ctx, comm._h, _ptr(chars), _ptr(offs), len(reqs), _ptr(has),
                                                                  _ptr(dist), page_base, _ptr(counts), C.byref(total)))
            else:
                check(self._ctx, L.dcd_search_batch_pages(self._ctx, _ptr(chars), _ptr(offs), len(reqs), _ptr(has), _ptr(dist), page_base,
                                                          _ptr(counts), C.byref(total)))
            buf = device_out(2 * max(total.value, 1))
            check(self._ctx, L.dcd_search_pages_fetch(self._ctx, buf.data_ptr(), total.value, 1))
            return counts, buf[:2 * total.value].view(-1, 2)
        h = C.c_void_p()
        if comm is not None:
            check(self._ctx, L.dcd_search_batch_sharded(self._ctx, comm._h, _ptr(chars), _ptr(offs), len(reqs), _ptr(has), _ptr(dist),
                                                        C.byref(h)))
        else:
            check(self._ctx, L.dcd_search_batch(self._ctx, _ptr(chars), _ptr(offs), len(reqs), _ptr(has), _ptr(dist), C.byref(h)))
        try:
            total = C.c_uint64()
            counts = np.zeros(len(reqs), dtype=np.int32)
            check(self._ctx, L.dcd_hits_export_pairs32(h, page_base, _ptr(counts), None, 0, C.byref(total)))
            if out is not None:
                buf = out(2 * max(total.value, 1))
                check(self._ctx, L.dcd_hits_export_pairs32(h, page_base, None, buf.data_ptr(), total.value, None))
                return counts, buf[:2 * total.value].view(-1, 2)
            pairs = np.zeros((max(total.value, 1), 2), dtype=np.int32)
            check(self._ctx, L.dcd_hits_export_pairs32(h, page_base, None, _ptr(pairs), total.value, None))
            return counts, pairs[:total.value]
        finally:
            L.dcd_hits_free(h)

    def _result(self, h, q):
        L = lib()
        r = SearchResult()
        st = L.dcd_hits_status(h, q)
        if st == -8:
            r.Success, r.Error = False, "Syntax Error in search request"   # Search.cs:505-507
            return r
        if st == -6:
            r.Success, r.Error = False, "Index is not built"               # Search.cs:445
            return r
        if st != 0:
            r.Success, r.Error = False, "Error: %d" % st
            return r
        for k in range(L.dcd_hits_num_pages(h, q)):
            o, rank, pp, npos = C.c_int64(), C.c_float(), C.c_void_p(), C.c_int32()
            L.dcd_hits_page(h, q, k, C.byref(o), C.byref(rank), C.byref(pp), C.byref(npos))
            pos = list(np.frombuffer(C.string_at(pp, 4 * npos.value), dtype=np.int32)) if npos.value else []
            pid = self.page_info(o.value)[0]
            r.foundPages.append(ResultDocPage(pid, [int(x) for x in pos], rank.value, o.value))
        for k in range(L.dcd_hits_num_docs(h, q)):
            d, rank, fp, npg = C.c_int64(), C.c_float(), C.c_int32(), C.c_int32()
            L.dcd_hits_doc(h, q, k, C.byref(d), C.byref(rank), C.byref(fp), C.byref(npg))
            doc = ResultDocument(self.doc_name(d.value), rank.value, r.foundPages[fp.value:fp.value + npg.value])
            doc.first_page = fp.value
            r.foundDocs.append(doc)
        for k in range(L.dcd_hits_num_words(h, q)):
            wp, n, nf = C.c_void_p(), C.c_int32(), C.c_int32()
            L.dcd_hits_word(h, q, k, C.byref(wp), C.byref(n), C.byref(nf))
            r.words.append({"Word": _str_at(wp, n.value), "nFound": nf.value})
        return r

    def Search(self, req, opt=None):   # Search.cs:437-617
        return self.SearchBatch([req], [opt])[0]

    # ---- synthetic corpus (benchmark utility) ----
    def synth_generate(self, words, zipf_s, seed, ndocs, pages_per_doc=20, min_page_chars=200, first_doc=0):
        enc = [w.encode("ascii") for w in words]
        offs = np.zeros(len(enc) + 1, dtype=np.uint32)
        offs[1:] = np.cumsum([len(w) for w in enc])
        blob = np.frombuffer(b"".join(enc), dtype=np.uint8).copy()
        out = dcd_synth()
        check(self._ctx, lib().dcd_synth_generate(self._ctx, _ptr(blob), _ptr(offs), len(enc), zipf_s, seed, ndocs, pages_per_doc,
                                                  min_page_chars, first_doc, C.byref(out)))
        return out

    def synth_free(self, s):
        lib().dcd_synth_free(self._ctx, C.byref(s))

    def d2h(self, dptr, nbytes, dtype=np.uint8):
        out = np.zeros(nbytes // np.dtype(dtype).itemsize, dtype=dtype)
        check(self._ctx, lib().dcd_memcpy_d2h(self._ctx, _ptr(out), dptr, nbytes))
        return out

    def download_raw(self):
        k = self.counters
        arena = np.zeros(max(int(k.encoded_u16), 1), dtype=np.uint16)
        offs = np.zeros(int(k.codes) + 1, dtype=np.uint64)
        pages = np.zeros(max(int(k.pages), 1), dtype=np.uint64)
        check(self._ctx, lib().dcd_index_download_raw(self._ctx, _ptr(arena), len(arena), _ptr(offs), _ptr(pages)))
        return arena[:int(k.encoded_u16)], offs, pages[:int(k.pages)]


class ShardComm:
    """One rank's handle of a shard group (dcd_comm): NCCL across processes / GPUs, or an in-process loopback group."""

    def __init__(self, handle, rank, nranks, transport):
        self._h, self.rank, self.nranks, self.transport = handle, rank, nranks, transport

    @staticmethod
    def unique_id():
        buf = (C.c_uint8 * 128)()
        check(None, lib().dcd_comm_unique_id(buf))
        return bytes(buf)

    @staticmethod
    def nccl(index, unique_id, rank, nranks):
        h = C.c_void_p()
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        check(index._ctx, lib().dcd_comm_init_nccl(index._ctx, buf, rank, nranks, C.byref(h)))
        return ShardComm(h, rank, nranks, "nccl")

    @staticmethod
    def local(nranks):
        arr = (C.c_void_p * nranks)()
        check(None, lib().dcd_comm_init_local(nranks, arr))
        return [ShardComm(C.c_void_p(arr[r]), r, nranks, "local") for r in range(nranks)]

    def destroy(self):
        if self._h:
            lib().dcd_comm_destroy(self._h)
            self._h = None


def shard_ranges(n_items, nranks):
    """Contiguous document ranges of near-equal size: [(first, count)] per rank (SURVEY.md 8e build partitioning)."""
    base, rem = divmod(n_items, nranks)
    out, first = [], 0
    for r in range(nranks):
        cnt = base + (1 if r < rem else 0)
        out.append((first, cnt))
        first += cnt
    return out


def coord_bases(text_lengths):
    """Global coordinate of each shard's first character = sum of the exact text lengths of the shards before it."""
    out, acc = [], 0
    for n in text_lengths:
        out.append(acc)
        acc += int(n)
    return out
