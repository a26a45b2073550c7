// DocodoB200.cs -- the reference-side binding of libdocodo_b200.so: what a Docodo maintainer adds to Docodo.NET so that
// `Docodo.Index` builds and queries on a B200 while every caller (Program.cs, server.cs, the xUnit tests, NuGet users)
// stays unchanged.  One [DllImport] per entry point of include/docodo_b200.h, 1:1, plus a partial-class body that
// reroutes the hot-path members of `Index` / `Index.Builder` (Index.cs, Build.cs, Search.cs of the reference).
//
// This file cannot be compiled in the build container (no .NET SDK); it is kept mechanical.  The Python mirror
// docodo_b200/index.py drives the very same ABI in the same order and is what the parity tests exercise.
//
// Build: add to Docodo.NET.csproj  <Compile Include="DocodoB200.cs" />  and define DOCODO_B200; ship libdocodo_b200.so
// next to Docodo.NET.dll (or on LD_LIBRARY_PATH).
#if DOCODO_B200
using System;
using System.Collections.Generic;
using System.Linq;
using System.Runtime.InteropServices;
using System.Text;

namespace Docodo
{
    internal static class B200
    {
        const string LIB = "docodo_b200";   // -> libdocodo_b200.so

        [StructLayout(LayoutKind.Sequential)]
        public struct Options
        {
            public int min_word_length, max_word_length, max_found_pages, max_found_docs, keep_forms, stemmers_mask;
            public uint dict_capacity_log2, reserved;
        }

        [StructLayout(LayoutKind.Sequential)]
        public unsafe struct Counters
        {
            public ulong text_units, tokens, postings, codes, distinct_words, encoded_u16, max_coord, pages;
            public double build_ms_device;
            public fixed double stage_ms[8];
            public ulong kernel_launches;
            public uint sort_passes, code_bits;
            public double search_ms_device;
            public ulong search_hits;
            public double build_ms_wall;
            public ulong search_bytes;
            public fixed double shard_ms[4];
            public uint shard_ranks, shard_rank;
            public ulong merged_u16, exchange_bytes;
            public uint merged_code_begin, merged_code_end, shard_p2p, reserved2;
        }

        // ---- context (Index ctor / Dispose: Index.cs:124-140, 769-786)
        [DllImport(LIB)] public static extern int dcd_ctx_create(int device, out IntPtr ctx);
        [DllImport(LIB)] public static extern void dcd_ctx_destroy(IntPtr ctx);
        [DllImport(LIB)] public static extern int dcd_ctx_set_stream(IntPtr ctx, IntPtr cudaStream);
        [DllImport(LIB)] public static extern IntPtr dcd_last_error(IntPtr ctx);
        [DllImport(LIB)] public static extern int dcd_options_default(out Options o);
        [DllImport(LIB)] public static extern int dcd_options_set(IntPtr ctx, ref Options o);
        [DllImport(LIB)] public static extern int dcd_counters_get(IntPtr ctx, out Counters k);
        // ---- vocabulary / stop list (Index.AddVoc Index.cs:221; Vocab.Load Dict.cs:71-95; LoadStopWords Index.cs:226-235)
        [DllImport(LIB)] public static extern int dcd_vocab_load(IntPtr ctx, byte[] voc, UIntPtr n, int stemmer);
        [DllImport(LIB)] public static extern int dcd_vocab_add_pairs(IntPtr ctx, char[] chars, uint[] offs, int[] groups, int n,
                                                                       ushort rangeBegin, ushort rangeEnd, int stemmer);
        [DllImport(LIB)] public static extern int dcd_stopwords_set(IntPtr ctx, char[] chars, uint[] offs, int n);
        // ---- build (CreateAsync / IndexTask Index.cs:424-534, Build.cs:445-564; Index.Builder Build.cs:258-437)
        // ---- sharded build across GPUs (replaces MergeAll / MergeFiles(shift_coords) / MergeIndexes, Index.cs:537-752)
        [DllImport(LIB)] public static extern int dcd_comm_unique_id(byte[] out128);
        [DllImport(LIB)] public static extern int dcd_comm_init_nccl(IntPtr ctx, byte[] id128, int rank, int nranks, out IntPtr comm);
        [DllImport(LIB)] public static extern int dcd_comm_init_local(int nranks, IntPtr[] comms);
        [DllImport(LIB)] public static extern void dcd_comm_destroy(IntPtr comm);
        [DllImport(LIB)] public static extern int dcd_build_finish_sharded(IntPtr b, IntPtr comm);
        [DllImport(LIB)] public static extern int dcd_merged_serialized_size(IntPtr ctx, out ulong indexBytes);
        [DllImport(LIB)] public static extern int dcd_merged_serialize(IntPtr ctx, byte[] indexOut, ulong indexCap);
        [DllImport(LIB)] public static extern int dcd_merged_download_raw(IntPtr ctx, ushort[] arena, ulong arenaCap, ulong[] codeOffs, ulong offsCap);
        [DllImport(LIB)] public static extern int dcd_build_begin(IntPtr ctx, out IntPtr b);
        [DllImport(LIB)] public static extern void dcd_build_abort(IntPtr b);
        [DllImport(LIB)] public static extern int dcd_build_set_origin(IntPtr b, ulong coordBase, ulong docNumberBase);
        [DllImport(LIB, CharSet = CharSet.Unicode)]
        public static extern int dcd_build_add_doc(IntPtr b, string source, int ns, string name, int nn);
        [DllImport(LIB, CharSet = CharSet.Unicode)]
        public static extern int dcd_build_add_page(IntPtr b, string id, int nid, string text, UIntPtr n);
        [DllImport(LIB, CharSet = CharSet.Unicode)]
        public static extern int dcd_build_add_pages_bulk(IntPtr b, IntPtr text, int encoding, ulong[] pageOffs, long npages,
                                                          ulong[] docFirstPage, long ndocs, string source, int ns, int onDevice);
        [DllImport(LIB, CharSet = CharSet.Unicode)]
        public static extern int dcd_build_add_word(IntPtr b, string word, int n, ulong coord);
        [DllImport(LIB, CharSet = CharSet.Unicode)]
        public static extern int dcd_build_add_posting(IntPtr b, string code, int n, ulong coord);
        [DllImport(LIB, CharSet = CharSet.Unicode)]
        public static extern int dcd_build_end_page(IntPtr b, string id, int nid, int useCoord, ulong coord);
        [DllImport(LIB)] public static extern int dcd_build_finish(IntPtr b);
        // ---- persistence (.index / .index.list: Index.cs:312-380, Build.cs:99-148,376-386)
        [DllImport(LIB)] public static extern int dcd_index_serialized_size(IntPtr ctx, out ulong indexBytes, out ulong listBytes);
        [DllImport(LIB)] public static extern int dcd_index_serialize(IntPtr ctx, byte[] index, ulong icap, byte[] list, ulong lcap);
        [DllImport(LIB)] public static extern int dcd_index_save(IntPtr ctx, byte[] dirUtf8Z);
        [DllImport(LIB)] public static extern int dcd_index_load(IntPtr ctx, byte[] dirUtf8Z);
        [DllImport(LIB)] public static extern int dcd_index_load_bytes(IntPtr ctx, byte[] index, ulong ni, byte[] list, ulong nl);
        [DllImport(LIB, CharSet = CharSet.Unicode)]
        public static extern int dcd_index_sequence(IntPtr ctx, string code, int n, out IntPtr data, out int count);
        [DllImport(LIB)] public static extern int dcd_index_download_raw(IntPtr ctx, ushort[] arena, ulong cap, ulong[] codeOffs, ulong[] pageEnds);
        // ---- query (Index.Search Search.cs:437-617; GetWordCodes Build.cs:175-247)
        [DllImport(LIB, CharSet = CharSet.Unicode)]
        public static extern int dcd_word_codes(IntPtr ctx, string word, int n, char[] outChars, int cap, out int units, out int ncodes);
        [DllImport(LIB)] public static extern int dcd_search_batch(IntPtr ctx, char[] reqs, uint[] reqOffs, int nreq, int[] hasOpt,
                                                                    int[] dist, out IntPtr hits);
        [DllImport(LIB)] public static extern void dcd_hits_free(IntPtr h);
        [DllImport(LIB)] public static extern int dcd_hits_status(IntPtr h, int q);
        [DllImport(LIB)] public static extern int dcd_hits_num_pages(IntPtr h, int q);
        [DllImport(LIB)] public static extern int dcd_hits_page(IntPtr h, int q, int k, out long pageOrdinal, out float rank,
                                                                 out IntPtr pos, out int npos);
        [DllImport(LIB)] public static extern int dcd_hits_export_pages(IntPtr hits, int[] pageCounts, long[] ordinals, float[] ranks, int[] npos, ulong cap, out ulong total);
        // benchmark / test helpers (no Docodo counterpart): device-side synthetic corpus, device -> host copy, host stemmer
        [StructLayout(LayoutKind.Sequential)] public struct dcd_synth { public IntPtr d_text, d_page_offs, d_doc_first_page; public ulong text_bytes, npages, ndocs; }
        [DllImport(LIB)] public static extern int dcd_synth_generate(IntPtr ctx, byte[] words, uint[] wordOffs, int nwords, double zipfS, ulong seed, long ndocs, int pagesPerDoc, int minPageChars, long firstDoc, out dcd_synth syn);
        [DllImport(LIB)] public static extern int dcd_synth_free(IntPtr ctx, ref dcd_synth syn);
        [DllImport(LIB)] public static extern int dcd_memcpy_d2h(IntPtr ctx, IntPtr dst, IntPtr dSrc, ulong n);
        [DllImport(LIB)] public static extern int dcd_stem_host(int stemmer, ushort[] word, int n, ushort[] outWord, int cap, out int outN);
        [DllImport(LIB)] public static extern int dcd_search_batch_pages(IntPtr ctx, ushort[] reqs, uint[] reqOffs, int nreq, int[] hasOpt, int[] dist, long pageBase, int[] pageCounts, out ulong total);
        [DllImport(LIB)] public static extern int dcd_search_batch_sharded(IntPtr ctx, IntPtr comm, ushort[] reqs, uint[] reqOffs, int nreq, int[] hasOpt, int[] dist, out IntPtr hits);
        [DllImport(LIB)] public static extern int dcd_search_batch_pages_sharded(IntPtr ctx, IntPtr comm, ushort[] reqs, uint[] reqOffs, int nreq, int[] hasOpt, int[] dist, long pageBase, int[] pageCounts, out ulong total);
        [DllImport(LIB)] public static extern int dcd_search_pages_fetch(IntPtr ctx, IntPtr pairs, ulong capPairs, int onDevice);
        [DllImport(LIB)] public static extern int dcd_hits_export_pairs32(IntPtr hits, long pageBase, int[] pageCounts, IntPtr pairs, ulong cap, out ulong total);
        [DllImport(LIB)] public static extern int dcd_hits_export_positions(IntPtr hits, int q, int[] pos, ulong cap, out ulong total);
        [DllImport(LIB)] public static extern int dcd_like_words(IntPtr ctx, ushort[] word, int n, ushort[] outChars, int cap, out int outUnits, out int count);
        [DllImport(LIB)] public static extern int dcd_close_words(IntPtr ctx, ushort[] word, int n, ushort[] outChars, int cap, out int outUnits, out int count);
        [DllImport(LIB)] public static extern int dcd_suggestions(IntPtr ctx, ushort[] req, int n, int maxWords, ushort[] outChars, int cap, out int outUnits, out int count);
        [DllImport(LIB)] public static extern int dcd_histogram(IntPtr ctx, int top, ushort[] outChars, int cap, out int outUnits, ulong[] counts, out int count);
        [DllImport(LIB)] public static extern int dcd_hits_num_docs(IntPtr h, int q);
        [DllImport(LIB)] public static extern int dcd_hits_doc(IntPtr h, int q, int k, out long docOrdinal, out float rank,
                                                                out int firstPage, out int npages);
        [DllImport(LIB)] public static extern int dcd_hits_num_words(IntPtr h, int q);
        [DllImport(LIB)] public static extern int dcd_hits_word(IntPtr h, int q, int k, out IntPtr word, out int n, out int nfound);
        [DllImport(LIB)] public static extern int dcd_index_page_info(IntPtr ctx, long pageOrdinal, out IntPtr id, out int nid,
                                                                      out long docOrdinal, out ulong start, out ulong end);
        [DllImport(LIB)] public static extern int dcd_index_doc_name(IntPtr ctx, long docOrdinal, out IntPtr name, out int n);
        // ---- IndexSequence algebra (IndexSequence.cs:63-84,114-136,218-322)
        [DllImport(LIB)] public static extern int dcd_seq_encode(IntPtr ctx, ulong[] values, ulong n, ushort[] o, ulong cap, out ulong on);
        // no counterpart in the reference: device pointers, nothing allocated, n <= 2^24
        [DllImport(LIB)] public static extern int dcd_seq_encode_noalloc(IntPtr ctx, IntPtr dValues, ulong n, IntPtr dOut, ulong cap, IntPtr dOutN);
        [DllImport(LIB)] public static extern int dcd_seq_decode(IntPtr ctx, ushort[] words, ulong n, ulong[] o, ulong cap, out ulong on);
        [DllImport(LIB)] public static extern int dcd_seq_op(IntPtr ctx, int op, ushort[] a, ulong na, int Ra, ushort[] b, ulong nb, int Rb,
                                                              ushort[] o, ulong cap, out ulong on, out int Rout);

        public static void Check(IntPtr ctx, int status)
        {
            if (status == 0) return;
            IntPtr p = dcd_last_error(ctx);
            throw new InvalidOperationException("docodo_b200 " + status + ": " + (p == IntPtr.Zero ? "" : Marshal.PtrToStringAnsi(p)));
        }
        public static string Str(IntPtr p, int n) => n == 0 ? "" : Marshal.PtrToStringUni(p, n);
        public static (char[], uint[]) Pack(IEnumerable<string> strs)
        {
            var sb = new StringBuilder();
            var offs = new List<uint> { 0 };
            foreach (string s in strs) { sb.Append(s); offs.Add((uint)sb.Length); }
            return (sb.Length == 0 ? new char[1] : sb.ToString().ToCharArray(), offs.ToArray());
        }
    }

    public partial class Index
    {
        IntPtr b200 = IntPtr.Zero;

        IntPtr Ctx()
        {
            if (b200 != IntPtr.Zero) return b200;
            int st = B200.dcd_ctx_create(0, out b200);
            if (st != 0) B200.Check(IntPtr.Zero, st);     // no CUDA device: loud failure, there is no CPU fallback
            B200.dcd_options_default(out B200.Options o);
            o.min_word_length = MIN_WORD_LENGTH; o.max_word_length = MAX_WORD_LENGTH;
            o.max_found_pages = MAX_FOUND_PAGES; o.max_found_docs = MAX_FOUND_DOCS; o.keep_forms = bKeepForms ? 1 : 0;
            o.stemmers_mask = 0;
            for (int k = 0; k < KnownStemmers.Length; k++)
                if (Stemmers.Any(s => s.lang == KnownStemmers[k].lang)) o.stemmers_mask |= 1 << k;
            B200.Check(b200, B200.dcd_options_set(b200, ref o));
            int slot = 0;
            foreach (Vocab v in vocs)                      // Index.AddVoc order == "#" + HEX(nVoc << 24 | group)
            {
                var (chars, offs) = B200.Pack(v.Keys);
                int stem = v.Name == "en" ? 1 : v.Name == "ru" ? 2 : 0;
                B200.Check(b200, B200.dcd_vocab_add_pairs(b200, chars, offs, v.Values.ToArray(), v.Count, v.Range.begin, v.Range.end, stem));
                slot++;
            }
            var (sc, so) = B200.Pack(stopWords);
            B200.Check(b200, B200.dcd_stopwords_set(b200, sc, so, stopWords.Count));
            return b200;
        }

        // Replaces the body of IndexTask (Build.cs:445-564): pages go to the device instead of through Regex/AddWord.
        void IndexTaskB200(IIndexDataSource source)
        {
            IntPtr ctx = Ctx();
            B200.Check(ctx, B200.dcd_build_begin(ctx, out IntPtr b));
            IIndexDocument doc;
            while ((doc = source.Next()) != null)
            {
                B200.Check(ctx, B200.dcd_build_add_doc(b, source.Name, source.Name.Length, doc.Name, doc.Name.Length));
                foreach (IndexPage page in doc)
                    B200.Check(ctx, B200.dcd_build_add_page(b, page.id, page.id.Length, page.text, (UIntPtr)page.text.Length));
                doc.Dispose();
            }
            B200.Check(ctx, B200.dcd_build_finish(b));     // Builder.Save + MergeIndexes + Load (Build.cs:370-434, Index.cs:488-512)
            byte[] dir = Encoding.UTF8.GetBytes(WorkPath + "\0");
            B200.Check(ctx, B200.dcd_index_save(ctx, dir));  // byte-identical .index / .index.list for stock Docodo readers
            CanSearch = true;
        }

        // Replaces Search(string, SearchOptions) (Search.cs:437-617) up to and including ranks; snippets stay managed.
        public SearchResult SearchB200(string req, SearchOptions opt = null)
        {
            IntPtr ctx = Ctx();
            var (chars, offs) = B200.Pack(new[] { req });
            int[] has = { opt == null ? 0 : 1 }, dist = { opt == null ? 0 : opt.dist };
            int st = B200.dcd_search_batch(ctx, chars, offs, 1, has, dist, out IntPtr h);
            if (st != 0) return new ErrorSearchResult("Error: " + st);
            try
            {
                int qs = B200.dcd_hits_status(h, 0);
                if (qs == -8) return new ErrorSearchResult("Syntax Error in search request");   // Search.cs:505-507
                if (qs == -6) return new ErrorSearchResult("Index is not built");                // Search.cs:445
                var res = new SearchResult();
                int np = B200.dcd_hits_num_pages(h, 0);
                for (int k = 0; k < np; k++)
                {
                    B200.dcd_hits_page(h, 0, k, out long ord, out float rank, out IntPtr pos, out int npos);
                    B200.dcd_index_page_info(ctx, ord, out IntPtr id, out int nid, out long docOrd, out ulong s, out ulong e);
                    var page = new ResultDocPage(new DocPage(B200.Str(id, nid)));
                    int[] p = new int[npos];
                    if (npos > 0) Marshal.Copy(pos, p, 0, npos);
                    page.pos.AddRange(p);
                    res.foundPages.Add(page);
                }
                int nd = B200.dcd_hits_num_docs(h, 0);
                for (int k = 0; k < nd; k++)
                {
                    B200.dcd_hits_doc(h, 0, k, out long docOrd, out float rank, out int first, out int npg);
                    B200.dcd_index_doc_name(ctx, docOrd, out IntPtr nm, out int nn);
                    var d = new ResultDocument(B200.Str(nm, nn)) { rank = rank };
                    for (int q = first; q < first + npg; q++) d.pages.Add(res.foundPages[q]);
                    res.foundDocs.Add(d);
                }
                int nw = B200.dcd_hits_num_words(h, 0);
                res.words = new SearchResult.WordInfo[nw];
                for (int k = 0; k < nw; k++)
                {
                    B200.dcd_hits_word(h, 0, k, out IntPtr w, out int n, out int nf);
                    res.words[k] = new SearchResult.WordInfo { Word = B200.Str(w, n), nFound = nf };
                }
                return res;
            }
            finally { B200.dcd_hits_free(h); }
        }

        void DisposeB200() { if (b200 != IntPtr.Zero) { B200.dcd_ctx_destroy(b200); b200 = IntPtr.Zero; } }
    }
}
#endif
