"""ctypes binding of libdocodo_b200.so -- the same entry points a C# [DllImport] shim binds."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdocodo_b200.so")
_LIB = None


class DocodoError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("docodo_b200 error %d: %s" % (code, msg))
        self.code = code


class dcd_options(C.Structure):
    _fields_ = [("min_word_length", C.c_int32), ("max_word_length", C.c_int32), ("max_found_pages", C.c_int32),
                ("max_found_docs", C.c_int32), ("keep_forms", C.c_int32), ("stemmers_mask", C.c_int32),
                ("dict_capacity_log2", C.c_uint32), ("reserved", C.c_uint32)]


class dcd_counters(C.Structure):
    _fields_ = [("text_units", C.c_uint64), ("tokens", C.c_uint64), ("postings", C.c_uint64), ("codes", C.c_uint64),
                ("distinct_words", C.c_uint64), ("encoded_u16", C.c_uint64), ("max_coord", C.c_uint64), ("pages", C.c_uint64),
                ("build_ms_device", C.c_double), ("stage_ms", C.c_double * 8), ("kernel_launches", C.c_uint64),
                ("sort_passes", C.c_uint32), ("code_bits", C.c_uint32), ("search_ms_device", C.c_double),
                ("search_hits", C.c_uint64), ("build_ms_wall", C.c_double), ("search_bytes", C.c_uint64),
                ("shard_ms", C.c_double * 4), ("shard_ranks", C.c_uint32), ("shard_rank", C.c_uint32),
                ("merged_u16", C.c_uint64), ("exchange_bytes", C.c_uint64), ("merged_code_begin", C.c_uint32),
                ("merged_code_end", C.c_uint32), ("shard_p2p", C.c_uint32), ("reserved2", C.c_uint32)]


class dcd_synth(C.Structure):
    _fields_ = [("d_text", C.c_void_p), ("d_page_offs", C.c_void_p), ("d_doc_first_page", C.c_void_p),
                ("text_bytes", C.c_uint64), ("npages", C.c_uint64), ("ndocs", C.c_uint64)]


P = C.POINTER
_vp, _i32, _i64, _u64, _u32, _u16 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_uint32, C.c_uint16

# name -> (restype, argtypes); every symbol declared in include/docodo_b200.h
SIGNATURES = {
    "dcd_ctx_create": (_i32, [_i32, P(_vp)]),
    "dcd_ctx_destroy": (None, [_vp]),
    "dcd_ctx_set_stream": (_i32, [_vp, _vp]),
    "dcd_last_error": (C.c_char_p, [_vp]),
    "dcd_options_default": (_i32, [P(dcd_options)]),
    "dcd_options_set": (_i32, [_vp, P(dcd_options)]),
    "dcd_counters_get": (_i32, [_vp, P(dcd_counters)]),
    "dcd_vocab_load": (_i32, [_vp, C.c_char_p, C.c_size_t, _i32]),
    "dcd_vocab_add_pairs": (_i32, [_vp, _vp, _vp, _vp, _i32, _u16, _u16, _i32]),
    "dcd_stopwords_set": (_i32, [_vp, _vp, _vp, _i32]),
    "dcd_build_begin": (_i32, [_vp, P(_vp)]),
    "dcd_build_abort": (None, [_vp]),
    "dcd_build_set_origin": (_i32, [_vp, _u64, _u64]),
    "dcd_build_add_doc": (_i32, [_vp, _vp, _i32, _vp, _i32]),
    "dcd_build_add_page": (_i32, [_vp, _vp, _i32, _vp, C.c_size_t]),
    "dcd_build_add_pages_bulk": (_i32, [_vp, _vp, _i32, _vp, _i64, _vp, _i64, _vp, _i32, _i32]),
    "dcd_build_add_word": (_i32, [_vp, _vp, _i32, _u64]),
    "dcd_build_add_posting": (_i32, [_vp, _vp, _i32, _u64]),
    "dcd_build_end_page": (_i32, [_vp, _vp, _i32, _i32, _u64]),
    "dcd_build_finish": (_i32, [_vp]),
    "dcd_comm_unique_id": (_i32, [_vp]),
    "dcd_comm_init_nccl": (_i32, [_vp, _vp, _i32, _i32, P(_vp)]),
    "dcd_comm_init_local": (_i32, [_i32, P(_vp)]),
    "dcd_comm_destroy": (None, [_vp]),
    "dcd_build_finish_sharded": (_i32, [_vp, _vp]),
    "dcd_merged_serialized_size": (_i32, [_vp, P(_u64)]),
    "dcd_merged_serialize": (_i32, [_vp, _vp, _u64]),
    "dcd_merged_download_raw": (_i32, [_vp, _vp, _u64, _vp, _u64]),
    "dcd_index_serialized_size": (_i32, [_vp, P(_u64), P(_u64)]),
    "dcd_index_serialize": (_i32, [_vp, _vp, _u64, _vp, _u64]),
    "dcd_index_save": (_i32, [_vp, C.c_char_p]),
    "dcd_index_load_bytes": (_i32, [_vp, C.c_char_p, _u64, C.c_char_p, _u64]),
    "dcd_index_load": (_i32, [_vp, C.c_char_p]),
    "dcd_index_sequence": (_i32, [_vp, _vp, _i32, P(_vp), P(_i32)]),
    "dcd_index_download_raw": (_i32, [_vp, _vp, _u64, _vp, _vp]),
    "dcd_word_codes": (_i32, [_vp, _vp, _i32, _vp, _i32, P(_i32), P(_i32)]),
    "dcd_search_batch": (_i32, [_vp, _vp, _vp, _i32, _vp, _vp, P(_vp)]),
    "dcd_hits_free": (None, [_vp]),
    "dcd_hits_status": (_i32, [_vp, _i32]),
    "dcd_hits_num_pages": (_i32, [_vp, _i32]),
    "dcd_hits_page": (_i32, [_vp, _i32, _i32, P(_i64), P(C.c_float), P(_vp), P(_i32)]),
    "dcd_hits_export_pages": (_i32, [_vp, _vp, _vp, _vp, _vp, _u64, P(_u64)]),
    "dcd_hits_export_pairs32": (_i32, [_vp, C.c_int64, _vp, _vp, _u64, P(_u64)]),
    "dcd_search_batch_pages": (_i32, [_vp, _vp, _vp, _i32, _vp, _vp, C.c_int64, _vp, P(_u64)]),
    "dcd_search_pages_fetch": (_i32, [_vp, _vp, _u64, _i32]),
    "dcd_search_batch_sharded": (_i32, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, P(_vp)]),
    "dcd_search_batch_pages_sharded": (_i32, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, C.c_int64, _vp, P(_u64)]),
    "dcd_hits_export_positions": (_i32, [_vp, _i32, _vp, _u64, P(_u64)]),
    "dcd_hits_num_docs": (_i32, [_vp, _i32]),
    "dcd_hits_doc": (_i32, [_vp, _i32, _i32, P(_i64), P(C.c_float), P(_i32), P(_i32)]),
    "dcd_hits_num_words": (_i32, [_vp, _i32]),
    "dcd_hits_word": (_i32, [_vp, _i32, _i32, P(_vp), P(_i32), P(_i32)]),
    "dcd_index_page_info": (_i32, [_vp, _i64, P(_vp), P(_i32), P(_i64), P(_u64), P(_u64)]),
    "dcd_index_doc_name": (_i32, [_vp, _i64, P(_vp), P(_i32)]),
    "dcd_like_words": (_i32, [_vp, _vp, _i32, _vp, _i32, P(_i32), P(_i32)]),
    "dcd_close_words": (_i32, [_vp, _vp, _i32, _vp, _i32, P(_i32), P(_i32)]),
    "dcd_suggestions": (_i32, [_vp, _vp, _i32, _i32, _vp, _i32, P(_i32), P(_i32)]),
    "dcd_histogram": (_i32, [_vp, _i32, _vp, _i32, P(_i32), _vp, P(_i32)]),
    "dcd_seq_encode": (_i32, [_vp, _vp, _u64, _vp, _u64, P(_u64)]),
    "dcd_seq_encode_noalloc": (_i32, [_vp, _vp, _u64, _vp, _u64, _vp]),
    "dcd_seq_decode": (_i32, [_vp, _vp, _u64, _vp, _u64, P(_u64)]),
    "dcd_seq_op": (_i32, [_vp, _i32, _vp, _u64, _i32, _vp, _u64, _i32, _vp, _u64, P(_u64), P(_i32)]),
    "dcd_synth_generate": (_i32, [_vp, _vp, _vp, _i32, C.c_double, _u64, _i64, _i32, _i32, _i64, P(dcd_synth)]),
    "dcd_synth_free": (_i32, [_vp, P(dcd_synth)]),
    "dcd_memcpy_d2h": (_i32, [_vp, _vp, _vp, _u64]),
    "dcd_stem_host": (_i32, [_i32, _vp, _i32, _vp, _i32, P(_i32)]),
}


def lib():
    """Load libdocodo_b200.so.  Fails loudly when the CUDA extension has not been built (no fallback exists)."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` or "
                              "`make -C docodo_b200/csrc` -- docodo_b200 has no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _LIB = L
    return _LIB


def check(ctx, status):
    if status != 0:
        msg = lib().dcd_last_error(ctx)
        raise DocodoError(status, msg.decode("utf-8", "replace") if msg else "")
