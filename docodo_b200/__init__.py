"""docodo_b200 -- B200-native (sm_100a) inverted-index build and posting-list query engine that drops in behind
Docodo.NET's ``Index`` API for that path (see DESIGN.md / INTEGRATION.md).

The package is a thin host mirror of the reference's C# surface (``Index``, ``Index.Builder``, ``IndexSequence``,
``Vocab``, ``SearchOptions``) over the C ABI of ``libdocodo_b200.so`` (include/docodo_b200.h).  There is no CPU
fallback: importing works without a GPU (so that build checks and symbol tests can run), creating an ``Index`` does not.
"""
from .capi import LIB_PATH, DocodoError, lib  # noqa: F401
from .index import (Index, IndexSequence, SearchOptions, SearchResult, ResultDocument, ResultDocPage, Vocab,  # noqa: F401
                    STEM_EN, STEM_NONE, STEM_RU, ENC_LATIN1, ENC_UTF16, ENC_UTF8, stem_host, ShardComm, shard_ranges, coord_bases)

__all__ = ["Index", "IndexSequence", "SearchOptions", "SearchResult", "ResultDocument", "ResultDocPage", "Vocab", "DocodoError",
           "lib", "LIB_PATH", "STEM_EN", "STEM_NONE", "STEM_RU", "ENC_LATIN1", "ENC_UTF16", "ENC_UTF8", "stem_host", "ShardComm", "shard_ranges", "coord_bases"]
