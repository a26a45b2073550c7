#include "query_kernels.cuh"

namespace dcd {

constexpr int QT = 512;          // threads per CTA: one program runs on ONE CTA, so wide CTAs / big tiles bound the longest program
constexpr int QI = 8;            // items per thread per tile
constexpr int QTILE = QT * QI;
constexpr int QSTACK = 16;

constexpr int QMERGE = QTILE + 2;   // merged tile + one halo element on each side
struct QSmem {
    u64 s64[QT / 32 + 1];
    u32 s32[QT / 32 + 1];
    u64* st_ptr[QSTACK];
    u32 st_n[QSTACK];
    u32 bcast;
    u32 cut[QT + 1];       // merge-path cuts of the tile boundaries of the current chunk of QT tiles
    // The three tile arrays are indexed through qpad(): thread t works on elements t * QI + k, so one padding slot per QI
    // elements turns the lane stride QI (16-way bank conflicts for 8-byte elements) into the odd stride QI + 1.
    u64 in[QMERGE + QMERGE / QI + 8];        // A part followed by B part
    u64 mv[QMERGE + QMERGE / QI + 8];        // merged values
    u8 mt[QMERGE + QMERGE / QI + 16];        // source list of each merged value (1 / 2)
};
DCD_D u32 qpad(u32 i) { return i + i / QI; }

// ---- decode (SeqEnumerator.MoveNext, IndexSequence.cs:114-136) -------------------------------------
// coordinate at a terminator word == inclusive prefix sum of ((w & 0x7FFF) << 15*g) over the whole stream, where g is
// the number of continuation words directly in front of the word; the value index is the count of terminators before it.
// first: the value Last starts from (0; a document-range shard of a sharded build: the code's first-delta base).
// limit / cap (early termination): decoding stops after the tile in which the running coordinate reaches `limit`; it stops
// BEFORE a tile that could overflow `cap` output elements, and *cut_short is set (the list is then incomplete below limit).
DCD_D u32 blk_decode(const u16* __restrict__ w, u64 n, u64* __restrict__ out, QSmem& sm, u64 first = 0, u64 limit = ~0ull,
                     u64 cap = ~0ull, bool* cut_short = nullptr) {
    const int tid = threadIdx.x;
    u64 carry_sum = first;
    u32 carry_cnt = 0;
    for (u64 base = 0; base < n; base += QTILE) {
        if (carry_sum >= limit) break;
        if ((u64)carry_cnt + min((u64)QTILE, n - base) > cap) { if (cut_short) *cut_short = true; break; }
        const u64 j0 = base + (u64)tid * QI;
        u64 c[QI];
        u32 t[QI];
        u64 lsum = 0;
        u32 lcnt = 0;
#pragma unroll
        for (int k = 0; k < QI; k++) {
            const u64 j = j0 + k;
            c[k] = 0;
            t[k] = 0;
            if (j < n) {
                const u16 v = w[j];
                int g = 0;
                while (g < 4 && j > (u64)g && (w[j - 1 - g] & 0x8000)) g++;
                c[k] = (u64)(v & 0x7FFF) << (15 * g);
                t[k] = (v & 0x8000) ? 0u : 1u;
            }
            lsum += c[k];
            lcnt += t[k];
        }
        u64 tot_sum;
        u32 tot_cnt;
        u64 run = carry_sum + block_excl_scan<u64, QT>(lsum, sm.s64, &tot_sum);
        u32 idx = carry_cnt + block_excl_scan<u32, QT>(lcnt, sm.s32, &tot_cnt);
#pragma unroll
        for (int k = 0; k < QI; k++) {
            run += c[k];
            if (t[k]) out[idx++] = run;
        }
        carry_sum += tot_sum;
        carry_cnt += tot_cnt;
        __syncthreads();
    }
    return carry_cnt;
}

// ---- stable merge with source tags -------------------------------------------------------------------
// Merge-path cut of diagonal d: number of A elements among the first d elements of the stable merge (A before B on
// ties), by one thread with a plain binary search: QT tile boundaries are cut at once, one boundary per thread.
DCD_D u32 thread_merge_cut(const u64* __restrict__ A, u32 nA, const u64* __restrict__ B, u32 nB, u32 d) {
    u32 lo = d > nB ? d - nB : 0, hi = min(d, nA);
    while (lo < hi) {
        const u32 mid = (lo + hi) >> 1;
        if (A[mid] <= B[d - 1 - mid]) lo = mid + 1; else hi = mid;
    }
    return lo;
}
// cut[i] = A-count of diagonal min(m, (tile0 + i) * QTILE) for i = 0 .. QT (the boundaries of tiles tile0 .. tile0 + QT - 1)
DCD_D void chunk_cuts(const u64* __restrict__ A, u32 nA, const u64* __restrict__ B, u32 nB, u32 m, u32 tile0, QSmem& sm) {
    for (u32 i = threadIdx.x; i <= QT; i += QT) {
        const u64 d = (u64)(tile0 + i) * QTILE;
        sm.cut[i] = thread_merge_cut(A, nA, B, nB, (u32)min(d, (u64)m));
    }
    __syncthreads();
}

// ---- fused tile pipeline of the binary operators -------------------------------------------------------------------
// The union of two lists is never written to memory: every pass re-creates it tile by tile in shared memory.  A tile is
// QTILE consecutive elements of the stable merge (A before B on ties); its input ranges follow from the merge-path cut of
// the tile's end diagonal (the start cut is the previous tile's end cut), both ranges are staged with coalesced loads
// and every element finds its rank with a shared-memory binary search.  mv[1 .. cnt] / mt[1 .. cnt] hold the merged
// values and their source list; mv[0] is the element in front of the tile, mv[cnt + 1] / mt[cnt + 1] the one behind it.
struct TileCtx {
    u32 d0, cnt;      // first union index of the tile, elements in it
    u32 a0, a1;       // A range [a0, a1); B range is [d0 - a0, d0 + cnt - a1)
};
// Asynchronous staging (cp.async, 8 bytes per element): the inputs of a tile -- and the four elements its halo is made of --
// travel to shared memory while the previous tile is still being processed; tile_merge waits for them.
constexpr u32 Q_HALO = QTILE + QTILE / QI + 4;   // in[Q_HALO .. +3]: A[a0 - 1], B[b0 - 1], A[a1], B[b1]
DCD_D void cp_async8(u64* smem_dst, const u64* gsrc) {
    const u32 d = (u32)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gsrc) : "memory");
}
DCD_D void tile_prefetch(const u64* __restrict__ A, u32 nA, const u64* __restrict__ B, u32 nB, const TileCtx& t, QSmem& sm) {
    const int tid = threadIdx.x;
    const u32 b0 = t.d0 - t.a0, b1 = t.d0 + t.cnt - t.a1;
    const u32 na = t.a1 - t.a0, nb = b1 - b0;
    for (u32 i = tid; i < na; i += QT) cp_async8(&sm.in[qpad(i)], A + t.a0 + i);
    for (u32 j = tid; j < nb; j += QT) cp_async8(&sm.in[qpad(na + j)], B + b0 + j);
    if (tid == 0 && t.a0 > 0) cp_async8(&sm.in[Q_HALO + 0], A + t.a0 - 1);
    if (tid == 1 && b0 > 0) cp_async8(&sm.in[Q_HALO + 1], B + b0 - 1);
    if (tid == 2 && t.a1 < nA) cp_async8(&sm.in[Q_HALO + 2], A + t.a1);
    if (tid == 3 && b1 < nB) cp_async8(&sm.in[Q_HALO + 3], B + b1);
    asm volatile("cp.async.commit_group;" ::: "memory");
}
// `staged`: tile_prefetch(t) has already been issued (after the previous tile's merge); otherwise it is issued here.
DCD_D void tile_merge(const u64* __restrict__ A, u32 nA, const u64* __restrict__ B, u32 nB, const TileCtx& t, QSmem& sm,
                      bool staged = false) {
    const int tid = threadIdx.x;
    const u32 b0 = t.d0 - t.a0, b1 = t.d0 + t.cnt - t.a1;
    const u32 na = t.a1 - t.a0, nb = b1 - b0;
    if (!staged) tile_prefetch(A, nA, B, nB, t, sm);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (tid == 0) {          // element in front of the tile: the later of A[a0 - 1], B[b0 - 1]
        u64 pv = 0;
        if (t.a0 > 0) pv = sm.in[Q_HALO + 0];
        if (b0 > 0) { const u64 bv = sm.in[Q_HALO + 1]; if (t.a0 == 0 || bv >= pv) pv = bv; }
        sm.mv[qpad(0)] = pv;
    }
    if (tid == 32) {         // element behind the tile: the head of the remaining merge
        u64 nv = 0;
        u8 nt = 0;
        const bool ha = t.a1 < nA, hb = b1 < nB;
        const u64 av = ha ? sm.in[Q_HALO + 2] : 0, bv = hb ? sm.in[Q_HALO + 3] : 0;
        if (ha && (!hb || av <= bv)) { nv = av; nt = 1; }
        else if (hb) { nv = bv; nt = 2; }
        sm.mv[qpad(t.cnt + 1)] = nv;
        sm.mt[qpad(t.cnt + 1)] = nt;
    }
    auto sa = [&](u32 i) { return sm.in[qpad(i)]; };
    auto sb = [&](u32 j) { return sm.in[qpad(na + j)]; };
    {   // thread tid merges outputs [tid * QI, tid * QI + QI) of the tile: merge-path cut in shared memory, then a serial merge
        const u32 dd = min((u32)tid * QI, t.cnt);
        u32 lo = dd > nb ? dd - nb : 0, hi = min(dd, na);
        while (lo < hi) {
            const u32 mid = (lo + hi) >> 1;
            if (sa(mid) <= sb(dd - 1 - mid)) lo = mid + 1; else hi = mid;
        }
        u32 ai = lo, bi = dd - lo;
#pragma unroll
        for (int k = 0; k < QI; k++) {
            const u32 o = dd + k;
            if (o < t.cnt) {
                const u64 va = ai < na ? sa(ai) : 0, vb = bi < nb ? sb(bi) : 0;
                const bool ta = ai < na && (bi >= nb || va <= vb);
                sm.mv[qpad(o + 1)] = ta ? va : vb;
                sm.mt[qpad(o + 1)] = ta ? 1 : 2;
                if (ta) ai++; else bi++;
            }
        }
    }
    __syncthreads();
}
// element k (0-based) of the current tile: value, lists carrying it (1, 2, 3 = both), validity (the second copy of a
// value present in both lists is dropped) -- the reference's merge loop emits one element per distinct coordinate pair
DCD_D void tile_element(const QSmem& sm, const TileCtx& t, u32 m, u32 k, u64& v, u32& tag, bool& valid) {
    const u32 q = k + 1, p = t.d0 + k;
    v = sm.mv[qpad(q)];
    const u32 src = sm.mt[qpad(q)];
    valid = !(src == 2 && p > 0 && sm.mv[qpad(q - 1)] == v);
    const bool both = src == 1 && p + 1 < m && sm.mt[qpad(q + 1)] == 2 && sm.mv[qpad(q + 1)] == v;
    tag = both ? 3u : src;
}

// Segmented scan operator.  A summary is: bit 7 = the range contains a run start; bits 0-2 describe the elements after
// the last start: a tag-1 element seen, a tag-2 element seen, a tag-2 element seen at or after the first tag-1 element.
DCD_D u32 seg_comb(u32 L, u32 R) { return (R & 0x80u) ? R : (L | R | (((L & 1u) & ((R >> 1) & 1u)) << 2)); }
// Block-wide EXCLUSIVE scan of one summary per thread under seg_comb; *total = reduction.  smem: QT/32+1 entries.
DCD_D u32 block_excl_scan_seg(u32 v, u32* smem, u32* total) {
    u32 inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u32 t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane_id() >= (u32)o) inc = seg_comb(t, inc);
    }
    if (lane_id() == 31) smem[warp_id()] = inc;
    __syncthreads();
    if (warp_id() == 0) {
        const u32 w = lane_id() < QT / 32 ? smem[lane_id()] : 0u;
        u32 wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const u32 t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane_id() >= (u32)o) wi = seg_comb(t, wi);
        }
        u32 wex = __shfl_up_sync(0xffffffffu, wi, 1);
        if (lane_id() == 0) wex = 0;
        if (lane_id() < QT / 32) smem[lane_id()] = wex;
        if (lane_id() == QT / 32 - 1) smem[QT / 32] = wi;
    }
    __syncthreads();
    u32 lane_ex = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane_id() == 0) lane_ex = 0;
    const u32 r = seg_comb(smem[warp_id()], lane_ex);
    *total = smem[QT / 32];
    __syncthreads();
    return r;
}

// operator + : sorted union with duplicates across the lists removed (IndexSequence.cs:286-322).  One forward pass.
DCD_D u32 blk_add(const u64* __restrict__ A, u32 nA, const u64* __restrict__ B, u32 nB, u64* __restrict__ out, QSmem& sm) {
    const u32 m = nA + nB;
    const int tid = threadIdx.x;
    u32 carry = 0;
    TileCtx t{};
    bool staged = false;
    for (u32 d0 = 0; d0 < m; d0 += QTILE) {
        t.d0 = d0;
        t.cnt = min(m - d0, (u32)QTILE);
        const u32 ti = d0 / QTILE;
        if (ti % QT == 0) chunk_cuts(A, nA, B, nB, m, ti, sm);
        t.a0 = sm.cut[ti % QT];
        t.a1 = sm.cut[ti % QT + 1];
        tile_merge(A, nA, B, nB, t, sm, staged);
        staged = (ti + 1) % QT != 0 && d0 + QTILE < m;   // next tile of the same cut chunk: stage it behind this tile's work
        if (staged) {
            TileCtx nx;
            nx.d0 = d0 + QTILE;
            nx.cnt = min(m - nx.d0, (u32)QTILE);
            nx.a0 = sm.cut[(ti + 1) % QT];
            nx.a1 = sm.cut[(ti + 1) % QT + 1];
            tile_prefetch(A, nA, B, nB, nx, sm);
        }
        u64 v[QI];
        u32 keep = 0, cnt = 0;
#pragma unroll
        for (int k = 0; k < QI; k++) {
            const u32 e = tid * QI + k;
            u32 tag;
            bool valid = false;
            v[k] = 0;
            if (e < t.cnt) tile_element(sm, t, m, e, v[k], tag, valid);
            if (valid) { keep |= 1u << k; cnt++; }
        }
        u32 tot;
        u32 idx = carry + block_excl_scan<u32, QT>(cnt, sm.s32, &tot);
#pragma unroll
        for (int k = 0; k < QI; k++)
            if (keep >> k & 1u) out[idx++] = v[k];
        carry += tot;
        __syncthreads();
    }
    return carry;
}

// operator * (IndexSequence.cs:218-284) in data-parallel form:
//   runs   = maximal stretches of the tagged union whose consecutive gaps are <= absR (absR == 0: one run)
//   output = unordered: every element of every run that contains tag 1 and tag 2;
//            ordered (both operands "exact", R < 0): the elements of a run from its first tag-1 element on, provided a
//            tag-2 element exists at or behind that element (the elements in front of it are dropped).
// An element is kept iff  prefix-within-run | suffix-within-run  says so: a BACKWARD pass over the tiles stores the
// suffix flags (1 byte per element: tag 1 / tag 2 seen at or behind the element), the FORWARD pass carries the prefix
// flags, decides, compacts and writes the output.  Both passes are single-pass segmented scans with a carry between tiles.
DCD_D u32 blk_mul(const u64* __restrict__ A, u32 nA, const u64* __restrict__ B, u32 nB, u32 absR, u32 ordered, u8* __restrict__ sfx,
                  u64* __restrict__ out, QSmem& sm) {
    const u32 m = nA + nB;
    if (!m) return 0;
    const int tid = threadIdx.x;
    const u32 ntiles = (m + QTILE - 1) / QTILE;
    TileCtx t{};
    // ---- backward pass: suffix flags -----------------------------------------------------------------------------
    {
        u32 carry = 0;       // summary of everything behind the current tile (in reverse processing order)
        bool staged = false;
        for (u32 ti = ntiles; ti-- > 0;) {
            t.d0 = ti * QTILE;
            t.cnt = min(m - t.d0, (u32)QTILE);
            if (ti % QT == QT - 1 || ti == ntiles - 1) chunk_cuts(A, nA, B, nB, m, ti - ti % QT, sm);
            t.a0 = sm.cut[ti % QT];
            t.a1 = sm.cut[ti % QT + 1];
            tile_merge(A, nA, B, nB, t, sm, staged);
            staged = ti % QT != 0;   // the tile in front belongs to the same cut chunk: stage it behind this tile's work
            if (staged) {
                TileCtx nx;
                nx.d0 = (ti - 1) * QTILE;
                nx.cnt = QTILE;
                nx.a0 = sm.cut[(ti - 1) % QT];
                nx.a1 = sm.cut[(ti - 1) % QT + 1];
                tile_prefetch(A, nA, B, nB, nx, sm);
            }
            // reversed order: thread tid owns reversed positions tid*QI + k  <->  element e = cnt - 1 - (tid*QI + k)
            u32 el[QI], x[QI];
            u32 loc = 0;
#pragma unroll
            for (int k = 0; k < QI; k++) {
                const u32 rp = tid * QI + k;
                el[k] = 0;
                x[k] = 0x100u;   // bit 8: not an element
                if (rp < t.cnt) {
                    const u32 e = t.cnt - 1 - rp;
                    u64 v;
                    u32 tag;
                    bool valid;
                    tile_element(sm, t, m, e, v, tag, valid);
                    if (valid) {
                        const u32 p = t.d0 + e;
                        const bool start = p == 0 || (absR != 0 && v - sm.mv[qpad(e)] > (u64)absR);   // mv[e] = element in front
                        x[k] = tag;
                        el[k] = start ? 0x80u : tag;     // the flags of a run start do not flow to the elements in front of it
                    }
                }
                loc = seg_comb(loc, el[k]);
            }
            u32 tot;
            u32 run = seg_comb(carry, block_excl_scan_seg(loc, sm.s32, &tot));
#pragma unroll
            for (int k = 0; k < QI; k++) {
                const u32 rp = tid * QI + k;
                if (rp < t.cnt) {
                    const u32 e = t.cnt - 1 - rp;
                    if (!(x[k] & 0x100u)) sfx[t.d0 + e] = (u8)((x[k] | run) & 3u);   // inclusive suffix within the run
                    run = seg_comb(run, el[k]);
                }
            }
            carry = seg_comb(carry, tot);
            __syncthreads();
        }
    }
    // ---- forward pass: prefix flags, decision, compaction -----------------------------------------------------------------
    u32 carry_n = 0, carry_s = 0;
    bool staged = false;
    for (u32 ti = 0; ti < ntiles; ti++) {
        t.d0 = ti * QTILE;
        t.cnt = min(m - t.d0, (u32)QTILE);
        if (ti % QT == 0) chunk_cuts(A, nA, B, nB, m, ti, sm);
        t.a0 = sm.cut[ti % QT];
        t.a1 = sm.cut[ti % QT + 1];
        tile_merge(A, nA, B, nB, t, sm, staged);
        staged = (ti + 1) % QT != 0 && ti + 1 < ntiles;
        if (staged) {
            TileCtx nx;
            nx.d0 = (ti + 1) * QTILE;
            nx.cnt = min(m - nx.d0, (u32)QTILE);
            nx.a0 = sm.cut[(ti + 1) % QT];
            nx.a1 = sm.cut[(ti + 1) % QT + 1];
            tile_prefetch(A, nA, B, nB, nx, sm);
        }
        u64 v[QI];
        u32 el[QI];
        u32 loc = 0, validm = 0;
#pragma unroll
        for (int k = 0; k < QI; k++) {
            const u32 e = tid * QI + k;
            el[k] = 0;
            v[k] = 0;
            if (e < t.cnt) {
                u32 tag;
                bool valid;
                tile_element(sm, t, m, e, v[k], tag, valid);
                if (valid) {
                    const u32 p = t.d0 + e;
                    const bool start = p == 0 || (absR != 0 && v[k] - sm.mv[qpad(e)] > (u64)absR);
                    el[k] = (start ? 0x80u : 0u) | (tag & 3u) | ((tag == 3u) ? 4u : 0u);
                    validm |= 1u << k;
                }
            }
            loc = seg_comb(loc, el[k]);
        }
        u32 tot_s;
        u32 run = seg_comb(carry_s, block_excl_scan_seg(loc, sm.s32, &tot_s));
        u32 keep = 0, cnt = 0;
#pragma unroll
        for (int k = 0; k < QI; k++) {
            if (validm >> k & 1u) {
                run = seg_comb(run, el[k]);
                const u32 sf = sfx[t.d0 + tid * QI + k];
                const bool kp = ordered ? ((run & 1u) && ((sf & 2u) || (run & 4u))) : (((run | sf) & 3u) == 3u);
                if (kp) { keep |= 1u << k; cnt++; }
            }
        }
        u32 tot;
        u32 idx = carry_n + block_excl_scan<u32, QT>(cnt, sm.s32, &tot);
#pragma unroll
        for (int k = 0; k < QI; k++)
            if (keep >> k & 1u) out[idx++] = v[k];
        carry_n += tot;
        carry_s = seg_comb(carry_s, tot_s);
        __syncthreads();
    }
    return carry_n;
}

// first index with v[i] >= x / > x
DCD_D u32 lower_bound_u64(const u64* __restrict__ v, u32 n, u64 x) {
    u32 lo = 0, hi = n;
    while (lo < hi) { const u32 mid = (lo + hi) >> 1; if (v[mid] < x) lo = mid + 1; else hi = mid; }
    return lo;
}
DCD_D u32 upper_bound_u64(const u64* __restrict__ v, u32 n, u64 x) {
    u32 lo = 0, hi = n;
    while (lo < hi) { const u32 mid = (lo + hi) >> 1; if (v[mid] <= x) lo = mid + 1; else hi = mid; }
    return lo;
}

// operator * when one list is much shorter than the other (galloping): same result as blk_mul, without touching the long
// list G except around the elements of the short list S.  Only runs that hold an S element can be emitted, so:
//   1. per S element: binary search in G, then the stretch of G chained to it (gaps <= absR) on either side ->
//      first / last coordinate of "S element + its G neighbourhood"
//   2. consecutive S elements whose neighbourhoods touch (gap <= absR) are one run; a run's G elements are exactly the G
//      elements between the run's first and last coordinate
//   3. per run: size (unordered: everything when G elements exist; ordered: from the first tag-1 element on, provided a
//      tag-2 element is at or behind it), block scan, serial merge of the two small slices
// s_is_first: S is the operator's first operand (tag 1).  Returns false (nothing written) when a neighbourhood walk exceeds
// GALLOP_WALK elements -- G is dense against absR there and the merge path is the right tool; the caller falls back to it.
// scratch64: 3 * nS u64 (first, last, run starts); scratch8: nS bytes.
constexpr u32 GALLOP_WALK = 192;
DCD_D bool blk_mul_gallop(const u64* __restrict__ S, u32 nS, const u64* __restrict__ G, u32 nG, u32 absR, u32 ordered, bool s_is_first,
                          u64* __restrict__ scratch64, u8* __restrict__ scratch8, u64* __restrict__ out, u32* n_out, QSmem& sm) {
    const int tid = threadIdx.x;
    u64* firstE = scratch64;
    u64* lastE = scratch64 + nS;
    u32* run_first = reinterpret_cast<u32*>(scratch64 + 2 * (u64)nS);
    // ---- 1. neighbourhoods ---------------------------------------------------------------------------------------------
    int over = 0;
    for (u32 i = tid; i < nS; i += QT) {
        const u64 s = S[i];
        const u32 p = lower_bound_u64(G, nG, s);
        u64 fe = s, le = s;
        u32 steps = 0;
        for (u32 j = p; j > 0 && fe - G[j - 1] <= (u64)absR; j--) { fe = G[j - 1]; if (++steps > GALLOP_WALK) { over = 1; break; } }
        steps = 0;
        for (u32 j = p; j < nG && G[j] - le <= (u64)absR; j++) { le = G[j]; if (++steps > GALLOP_WALK) { over = 1; break; } }
        firstE[i] = fe;
        lastE[i] = le;
        scratch8[i] = (p < nG && G[p] == s) ? 1 : 0;   // the coordinate is in both lists: emitted once
    }
    if (__syncthreads_or(over)) return false;
    // ---- 2. run starts ----------------------------------------------------------------------------------------------------
    u32 nruns = 0;
    for (u32 base = 0; base < nS; base += QT) {
        const u32 i = base + tid;
        const u32 f = (i < nS && (i == 0 || firstE[i] > lastE[i - 1] + (u64)absR)) ? 1u : 0u;
        u32 tot;
        const u32 idx = nruns + block_excl_scan<u32, QT>(f, sm.s32, &tot);
        if (f) run_first[idx] = i;
        nruns += tot;
        __syncthreads();
    }
    // ---- 3. sizes, offsets, output ---------------------------------------------------------------------------------------------
    u32 carry = 0;
    for (u32 base = 0; base < nruns; base += QT) {
        const u32 r = base + tid;
        u32 i0 = 0, i1 = 0, j0 = 0, j1 = 0, size = 0;   // S range [i0, i1), G range [j0, j1) to emit
        if (r < nruns) {
            i0 = run_first[r];
            i1 = r + 1 < nruns ? run_first[r + 1] : nS;
            j0 = lower_bound_u64(G, nG, firstE[i0]);
            j1 = upper_bound_u64(G, nG, lastE[i1 - 1]);
            bool emit = j1 > j0;                         // the run holds elements of both lists
            if (emit && ordered) {
                if (s_is_first) {                        // first tag-1 element = S[i0]; G elements in front of it are dropped
                    j0 = lower_bound_u64(G, nG, S[i0]);
                    emit = j1 > j0;                      // a tag-2 element at or behind it
                } else {                                 // first tag-1 element = G[j0]; S elements in front of it are dropped
                    i0 += lower_bound_u64(S + i0, i1 - i0, G[j0]);
                    emit = i1 > i0;
                }
            }
            if (emit) {
                u32 dups = 0;
                for (u32 i = i0; i < i1; i++) dups += scratch8[i];
                size = (i1 - i0) + (j1 - j0) - dups;
            }
        }
        u32 tot;
        u32 o = carry + block_excl_scan<u32, QT>(size, sm.s32, &tot);
        if (size) {
            u32 i = i0, j = j0;
            while (i < i1 || j < j1) {
                const u64 sv = i < i1 ? S[i] : ~0ull, gv = j < j1 ? G[j] : ~0ull;
                if (sv == gv) { out[o++] = sv; i++; j++; }
                else if (sv < gv) { out[o++] = sv; i++; }
                else { out[o++] = gv; j++; }
            }
        }
        carry += tot;
        __syncthreads();
    }
    *n_out = carry;
    return true;
}

// ---- hits -> pages (K9) --------------------------------------------------------------------------
DCD_D void blk_hits(const u64* __restrict__ R, u32 n, const u64* __restrict__ page_end, u64 npages, u64 origin, u32* __restrict__ hit_page,
                    int* __restrict__ hit_pos, u32* __restrict__ page_first, float* __restrict__ page_rank, u32* n_pages_out,
                    u32* n_hits_out, u32 max_pages, QSmem& sm) {
    const int tid = threadIdx.x;
    for (u32 i = tid; i < n; i += QT) {
        const u64 c = R[i];
        u64 lo = 0, hi = npages;   // first page with end > c   (IndexPageList.GetPage, Build.cs:86-94)
        while (lo < hi) { u64 mid = (lo + hi) >> 1; if (page_end[mid] > c) hi = mid; else lo = mid + 1; }
        if (lo >= npages) { hit_page[i] = 0xFFFFFFFFu; hit_pos[i] = 0; }
        else {
            const u64 start = lo ? page_end[lo - 1] : origin;
            hit_page[i] = (u32)lo;
            hit_pos[i] = (int)(c - start);                  // (int)(coord - _p.coord), Search.cs:381,407
        }
    }
    __syncthreads();
    // page starts
    u32 carry = 0;
    for (u32 base = 0; base < n; base += QT) {
        const u32 i = base + tid;
        const u32 f = (i < n && (i == 0 || hit_page[i] != hit_page[i - 1])) ? 1u : 0u;
        u32 tot;
        const u32 idx = carry + block_excl_scan<u32, QT>(f, sm.s32, &tot);
        if (f) page_first[idx] = i;
        carry += tot;
        __syncthreads();
    }
    u32 np = carry;
    // the reference stops shaping right after page max_pages + 1 (only its first position is kept, Search.cs:409)
    if (np > max_pages + 1) {
        const u32 cut = page_first[max_pages + 1];
        np = max_pages + 1;
        n = cut;
        if (tid == 0) *n_hits_out = cut;
    }
    // ResultDocPage.rank, Search.cs:99-111: 1 + sum (float)(30 / max(5, dpos)) [integer division] + (float)ln(count)
    for (u32 k = tid; k < np; k += QT) {
        const u32 a = page_first[k], e = k + 1 < np ? page_first[k + 1] : n;
        float bonus = 0.f;
        for (u32 q = a + 1; q < e; q++) bonus += (float)(30 / max(5, hit_pos[q] - hit_pos[q - 1]));
        page_rank[k] = 1.f + bonus + (float)log((double)(e - a));
    }
    if (tid == 0) *n_pages_out = np;
    __syncthreads();
}

// LOAD: OR (operator+) of the decoded lists of in.b codes (SearchWord, Search.cs:235-245).  The region at `out` holds one half
// of in.cap elements (two when in.b > 1: ping-pong of the fold); every list is written H elements into its half, so that the
// caller can put a halo of H elements in front of the result.  Returns false after flagging a capacity error.
DCD_D bool blk_load(const QBatch& qb, const QInstr& in, u64* out, u32 H, u64* tmp_dec, u64 tmp_cap, QSmem& sm, u64*& cur, u32& n,
                    u64 limit = ~0ull, bool* cut_short = nullptr) {
    out += H;
    cur = out;
    n = 0;
    if (in.op == QOP_LOADC) {   // decoded once per index (search.cu): no decode; copied only when a halo goes in front of it
        const u64* src = reinterpret_cast<const u64*>(in.src);
        n = in.cap;
        if (limit != ~0ull) {
            n = lower_bound_u64(src, n, limit);
            if (n > in.lim_cap) { n = 0; if (cut_short) *cut_short = true; }   // more than the planned outputs hold: run again wider
        }
        if (H) for (u32 i = threadIdx.x; i < n; i += QT) cur[i] = src[i];
        else cur = const_cast<u64*>(src);
        __syncthreads();
        return true;
    }
    if (in.b >= 1) {
        const u32 c0 = qb.code_ids[in.a];
        const u64 a0 = qb.code_off[c0], e0 = qb.code_off[c0 + 1];
        if (limit != ~0ull) {   // limited programs hold single-code LOADs only (search.cu)
            n = blk_decode(qb.arena + a0, e0 - a0, cur, sm, qb.code_base ? qb.code_base[c0] : 0ull, limit, in.cap, cut_short);
            return true;
        }
        n = blk_decode(qb.arena + a0, e0 - a0, cur, sm, qb.code_base ? qb.code_base[c0] : 0ull);
        for (u32 k = 1; k < in.b; k++) {
            const u32 ck = qb.code_ids[in.a + k];
            const u64 ak = qb.code_off[ck], ek = qb.code_off[ck + 1];
            const u32 nk = blk_decode(qb.arena + ak, ek - ak, tmp_dec, sm, qb.code_base ? qb.code_base[ck] : 0ull);
            if ((u64)n + nk > tmp_cap) { if (threadIdx.x == 0) atomicExch(qb.error, 1u); return false; }
            u64* nxt = (cur == out) ? out + in.cap : out;
            n = blk_add(cur, n, tmp_dec, nk, nxt, sm);
            cur = nxt;
        }
    }
    return true;
}

// Start of the chain of U (union of the nl sorted lists, consecutive gaps <= R) that ends at the last element of U: the largest
// element whose predecessor in U is more than R away (or absent).  0 when U is empty.  red: QT / 32 + 1 shared words.
constexpr int QMAX_LOADS = 64;
DCD_D u64 blk_chain_start(u64* const* l_ptr, const u32* l_n, u32 nl, u64 R, u64* red) {
    const int tid = threadIdx.x;
    u64 first = ~0ull, last = 0;
    for (u32 i = 0; i < nl; i++)
        if (l_n[i]) { first = min(first, l_ptr[i][0]); last = max(last, l_ptr[i][l_n[i] - 1]); }
    if (first == ~0ull) return 0;
    // Only the LAST break matters, so the lists are examined from their ends: the elements at or above x = last - span; a break
    // found there is the largest one (every element above it was looked at).  The span grows until a break is found or the
    // window holds everything (then the chain starts at the first element of U).
    u64 span = max(R, (u64)1) * 64;
    while (true) {
        const u64 x = (span >= last - first) ? first : last - span;
        u64 lo = 0;
        for (u32 i = 0; i < nl; i++) {
            const u64* L = l_ptr[i];
            const u32 n = l_n[i];
            const u32 j0 = x == first ? 0u : lower_bound_u64(L, n, x);
            for (u32 j = j0 + tid; j < n; j += QT) {
                const u64 e = L[j];
                if (e <= lo) continue;
                bool has = j > 0;
                u64 pred = has ? L[j - 1] : 0;
                if (has && e - pred <= R) continue;
                for (u32 i2 = 0; i2 < nl; i2++) {
                    if (i2 == i) continue;
                    const u32 k = lower_bound_u64(l_ptr[i2], l_n[i2], e);   // elements < e
                    if (k) { const u64 p2 = l_ptr[i2][k - 1]; if (!has || p2 > pred) pred = p2; has = true; }
                }
                if (!has || e - pred > R) lo = max(lo, e);
            }
        }
        for (int o = 16; o > 0; o >>= 1) lo = max(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        __syncthreads();
        if (lane_id() == 0) red[warp_id()] = lo;
        __syncthreads();
        if (tid == 0) {
            u64 m = 0;
            for (int w = 0; w < QT / 32; w++) m = max(m, red[w]);
            red[QT / 32] = m;
        }
        __syncthreads();
        const u64 found = red[QT / 32];
        if (found || x == first) return found;   // x == first: every element was examined (the first element of U is a break)
        span *= 8;
    }
}

// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(QT, 2) k_query_batch(QBatch qb) {
    extern __shared__ __align__(16) unsigned char q_smem_raw[];
    QSmem& sm = *reinterpret_cast<QSmem*>(q_smem_raw);
    const u32 q = blockIdx.x;
    if (q >= qb.nqueries) return;
    unsigned long long t_begin = 0;
    if (qb.cta_ns && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_begin));
    const QDesc d = qb.desc[q];
    u64* lists = qb.lists + d.list_off;
    u64* tmp_dec = qb.tmp_dec + d.tmp_off;
    u8* tmp_tag = qb.tmp_tag + d.tmp_off;
    __shared__ u64* l_ptr[QMAX_LOADS];
    __shared__ u32 l_n[QMAX_LOADS];
    __shared__ u64 l_red[QT / 32 + 1];
    const bool limited = d.limit != ~0ull;
    bool cut_short = false, any_cut = false;   // a LOAD ran out of room / some list goes on beyond the limit
    int sp = 0;
    u32 nload = 0;
    for (u32 pc = d.instr_begin; pc < d.instr_end; pc++) {
        const QInstr in = qb.instr[pc];
        if (in.op == QOP_LOAD || in.op == QOP_LOADC) {
            u32 H = 0;
            const u64* halo = nullptr;
            if (qb.halo_cnt) {
                const u32 slot = qb.load_base[q] + nload;
                H = qb.halo_cnt[slot];
                halo = qb.halo_vals + qb.halo_off[slot];
            }
            nload++;
            u64* cur;
            u32 n;
            if (!blk_load(qb, in, lists + in.out_off, H, tmp_dec, d.tmp_cap, sm, cur, n, d.limit, &cut_short)) return;
            if (limited) {   // keep what lies in front of the limit; remember the leaf list for the trust bound
                const u32 full = in.op == QOP_LOADC ? in.cap : n;
                if (in.op != QOP_LOADC) n = lower_bound_u64(cur, n, d.limit);
                // the list goes on iff words were left undecoded or its values reach the limit
                if (in.b >= 1) any_cut = any_cut || cut_short || n < full || (in.op != QOP_LOADC && full && cur[full - 1] >= d.limit);
                if (nload - 1 < (u32)QMAX_LOADS && threadIdx.x == 0) { l_ptr[nload - 1] = cur; l_n[nload - 1] = n; }
            }
            if (H) {   // the halo (all its coordinates lie left of this shard) goes in front of the decoded list
                cur -= H;
                for (u32 i = threadIdx.x; i < H; i += QT) cur[i] = halo[i];
                n += H;
            }
            if (sp >= QSTACK) { if (threadIdx.x == 0) atomicExch(qb.error, 1u); return; }
            if (threadIdx.x == 0) { sm.st_ptr[sp] = cur; sm.st_n[sp] = n; }
            sp++;
            __syncthreads();
        } else if (in.op == QOP_MUL || in.op == QOP_ADD) {
            if (sp < 2) { if (threadIdx.x == 0) atomicExch(qb.error, 1u); return; }
            const u64* B = sm.st_ptr[sp - 1];
            const u32 nB = sm.st_n[sp - 1];
            const u64* A = sm.st_ptr[sp - 2];
            const u32 nA = sm.st_n[sp - 2];
            u64* out = lists + in.out_off;
            u32 n = 0;
            if ((u64)nA + nB > d.tmp_cap) { if (threadIdx.x == 0) atomicExch(qb.error, 1u); return; }
            if (in.op == QOP_ADD) n = blk_add(A, nA, B, nB, out, sm);
            else {
                // one list much shorter than the other: work only around the short list's elements (galloping); falls back to
                // the merge path where the long list is dense against absR
                bool done = false;
                const u32 small = min(nA, nB), big = max(nA, nB);
                if (qb.gallop_ratio && in.a != 0 && small && (u64)small * qb.gallop_ratio <= big)
                    done = nA <= nB ? blk_mul_gallop(A, nA, B, nB, in.a, in.b, true, tmp_dec, tmp_tag, out, &n, sm)
                                    : blk_mul_gallop(B, nB, A, nA, in.a, in.b, false, tmp_dec, tmp_tag, out, &n, sm);
                if (!done) n = small ? blk_mul(A, nA, B, nB, in.a, in.b, tmp_tag, out, sm) : 0;
            }
            sp -= 2;
            __syncthreads();
            if (threadIdx.x == 0) { sm.st_ptr[sp] = out; sm.st_n[sp] = n; }
            sp++;
            __syncthreads();
        } else if (in.op == QOP_COUNT) {
            // wordInfo.nFound; a shard with a halo counts its own range only (the halo elements are counted by their own shard)
            if (sp >= 1 && threadIdx.x == 0 && !limited) {   // limited programs: the host takes nFound from the per-code counts
                u32 cnt = sm.st_n[sp - 1];
                if (qb.halo_cnt) cnt -= lower_bound_u64(sm.st_ptr[sp - 1], cnt, qb.coord_origin);
                qb.nfound[in.a] = cnt;
            }
        }
    }
    u32 n = 0;
    const u64* R = nullptr;
    if (sp >= 1) { R = sm.st_ptr[sp - 1]; n = sm.st_n[sp - 1]; }
    bool trust_all = true;
    if (limited) {
        // Was every list complete in front of the limit?  A list is complete when its decode ran to the end of its words.
        // any_cut is conservative: it is also set when the list merely has elements at or beyond the limit.
        __syncthreads();
        if (any_cut) {
            trust_all = false;
            const u64 S = blk_chain_start(l_ptr, l_n, min(nload, (u32)QMAX_LOADS), d.rmax, l_red);
            n = lower_bound_u64(R, n, S);   // elements from the last chain's start on may belong to a run that goes on
        }
    }
    if (qb.defer_pages) {   // exact sharded serving: the list goes through the exchange first (k_query_pages)
        if (threadIdx.x == 0) {
            const u32 keep = qb.keep_below ? lower_bound_u64(R, n, qb.keep_below[q]) : n;
            qb.fin_ptr[q] = const_cast<u64*>(R);
            qb.fin_n[q] = keep;
            qb.fin_foreign[q] = min(keep, lower_bound_u64(R, n, qb.coord_origin));
        }
        return;
    }
    if (n > d.hit_cap) { if (threadIdx.x == 0) atomicExch(qb.error, 1u); n = (u32)d.hit_cap; }
    if (threadIdx.x == 0) qb.nhits[q] = n;
    blk_hits(R, n, qb.page_end, qb.npages, qb.coord_origin, qb.hit_page + d.hit_off, qb.hit_pos + d.hit_off, qb.page_first + d.hit_off,
             qb.page_rank + d.hit_off, qb.npages_found + q, qb.nhits + q, qb.max_pages, sm);
    if (qb.incomplete && threadIdx.x == 0) {   // not enough trusted pages (or a LOAD ran out of room): run again with a larger limit
        const bool inc = limited && (cut_short || (!trust_all && qb.npages_found[q] <= qb.max_pages));
        qb.incomplete[q] = inc ? 1u : 0u;
        if (inc) { qb.nhits[q] = 0; qb.npages_found[q] = 0; }   // nothing of this run is packed or shaped
    }
    if (qb.cta_ns && threadIdx.x == 0) {
        unsigned long long t_end;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
        qb.cta_ns[q] = t_end - t_begin;
    }
}

void launch_query_batch(const QBatch& qb, cudaStream_t s) {
    if (!qb.nqueries) return;
    cudaFuncSetAttribute(k_query_batch, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(QSmem));
    k_query_batch<<<qb.nqueries, QT, sizeof(QSmem), s>>>(qb);
}

// ---- exact serving from document-range shards: halo pass, final-list packing, hits -> pages after the exchange -----------
__global__ void __launch_bounds__(QT, 2) k_query_halo(QBatch qb, HaloPass hp) {
    extern __shared__ __align__(16) unsigned char q_smem_raw[];
    QSmem& sm = *reinterpret_cast<QSmem*>(q_smem_raw);
    __shared__ u64* l_ptr[QMAX_LOADS];
    __shared__ u32 l_n[QMAX_LOADS];
    __shared__ u64 red[QT / 32 + 1];
    __shared__ u64 s_base;
    const u32 q = blockIdx.x;
    if (q >= qb.nqueries) return;
    const int tid = threadIdx.x;
    const QDesc d = qb.desc[q];
    u64* lists = qb.lists + d.list_off;
    u64* tmp_dec = qb.tmp_dec + d.tmp_off;
    const u32 slot0 = qb.load_base[q];
    u32 nl = 0;
    u64 run = 0, next_first = ~0ull;
    for (u32 pc = d.instr_begin; pc < d.instr_end; pc++) {
        const QInstr in = qb.instr[pc];
        if (in.op != QOP_LOAD && in.op != QOP_LOADC) continue;
        if (nl >= (u32)QMAX_LOADS) { if (tid == 0) atomicExch(qb.error, 2u); return; }
        u64* cur;
        u32 n;
        if (!blk_load(qb, in, lists + run, 0, tmp_dec, d.tmp_cap, sm, cur, n)) return;
        if (in.op == QOP_LOAD) run += (u64)(in.b > 1 ? 2 : 1) * in.cap;
        for (u32 k = 0; k < in.b; k++) next_first = min(next_first, hp.next_first[qb.code_ids[in.a + k]]);
        if (tid == 0) { l_ptr[nl] = cur; l_n[nl] = n; }
        nl++;
        __syncthreads();
    }
    const u64 R = hp.rmax[q];
    // first / last element of U; chain start = the largest element of U whose predecessor in U is more than R away (or absent)
    u64 lo = 0, first = ~0ull, last = 0;
    bool any = false;
    for (u32 i = 0; i < nl; i++)
        if (l_n[i]) { any = true; first = min(first, l_ptr[i][0]); last = max(last, l_ptr[i][l_n[i] - 1]); }
    if (any) lo = blk_chain_start(l_ptr, l_n, nl, R, red);
    const bool bridged = any && next_first != ~0ull && next_first >= last && next_first - last <= R;
    if (tid == 0) {
        HaloDesc hd;
        hd.lo = lo;
        hd.first = first;
        hd.last = last;
        hd.flags = (any ? 1u : 0u) | (bridged ? 2u : 0u) | ((any && lo == first) ? 4u : 0u);
        hd.pad = 0;
        hp.desc[q] = hd;
    }
    // tails [lo, ..) of every leaf list: the halo the next non-empty shard prepends
    u32 total = 0;
    for (u32 i = 0; i < nl; i++) {
        u32 cnt = 0;
        if (bridged) cnt = l_n[i] - lower_bound_u64(l_ptr[i], l_n[i], lo);
        total += cnt;
    }
    if (tid == 0) s_base = total ? atomicAdd(hp.cursor, (unsigned long long)total) : 0ull;
    __syncthreads();
    u64 at = s_base;
    const bool fits = at + total <= hp.vals_cap;
    if (!fits && tid == 0) atomicExch(qb.error, 3u);
    for (u32 i = 0; i < nl; i++) {
        u32 cnt = 0, t0 = 0;
        if (bridged) { t0 = lower_bound_u64(l_ptr[i], l_n[i], lo); cnt = l_n[i] - t0; }
        if (tid == 0) { hp.slot_cnt[slot0 + i] = fits ? cnt : 0u; hp.slot_off[slot0 + i] = at; }
        if (fits) for (u32 j = tid; j < cnt; j += QT) hp.vals[at + j] = l_ptr[i][t0 + j];
        at += cnt;
    }
}
void launch_query_halo(const QBatch& qb, const HaloPass& hp, cudaStream_t s) {
    if (!qb.nqueries) return;
    cudaFuncSetAttribute(k_query_halo, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(QSmem));
    k_query_halo<<<qb.nqueries, QT, sizeof(QSmem), s>>>(qb, hp);
}

__global__ void k_pack_final(QBatch qb, const u64* __restrict__ own_off, const u64* __restrict__ foreign_off, u64* __restrict__ own_vals,
                             u64* __restrict__ foreign_vals) {
    const u32 q = blockIdx.x;
    const u64* R = qb.fin_ptr[q];
    const u32 n = qb.fin_n[q], nf = qb.fin_foreign[q];
    for (u32 i = threadIdx.x; i < nf; i += blockDim.x) foreign_vals[foreign_off[q] + i] = R[i];
    for (u32 i = nf + threadIdx.x; i < n; i += blockDim.x) own_vals[own_off[q] + (i - nf)] = R[i];
}
void launch_pack_final(const QBatch& qb, const u64* own_off, const u64* foreign_off, u64* own_vals, u64* foreign_vals, cudaStream_t s) {
    if (qb.nqueries) k_pack_final<<<qb.nqueries, 256, 0, s>>>(qb, own_off, foreign_off, own_vals, foreign_vals);
}

__global__ void __launch_bounds__(QT, 2) k_query_pages(QBatch qb, PagesPass pp) {
    extern __shared__ __align__(16) unsigned char q_smem_raw[];
    QSmem& sm = *reinterpret_cast<QSmem*>(q_smem_raw);
    __shared__ u32 s_a, s_b;
    const u32 q = blockIdx.x;
    if (q >= qb.nqueries) return;
    const int tid = threadIdx.x;
    const QDesc d = qb.desc[q];
    u64* W = pp.work + d.list_off;
    u32 n = pp.own_n[q];
    const u64* own = pp.own_ptr[q];
    for (u32 i = tid; i < n; i += QT) W[i] = own[i];
    // later ranks in rank order: their valid intervals ascend, so the pieces inside this shard's range concatenate in order
    for (int r = pp.rank + 1; r < pp.nranks; r++) {
        const u32 cnt = pp.f_cnt[r][q];
        if (!cnt) continue;
        const u64* F = pp.f_vals[r] + pp.f_off[r][q];
        if (tid == 0) { s_a = lower_bound_u64(F, cnt, qb.coord_origin); s_b = lower_bound_u64(F, cnt, pp.range_end); }
        __syncthreads();
        const u32 a = s_a, b = s_b;
        if ((u64)n + (b - a) > d.hit_cap) { if (tid == 0) atomicExch(qb.error, 1u); return; }
        for (u32 i = tid; i < b - a; i += QT) W[n + i] = F[a + i];
        n += b - a;
        __syncthreads();
    }
    __syncthreads();
    if (tid == 0) qb.nhits[q] = n;
    blk_hits(W, n, qb.page_end, qb.npages, qb.coord_origin, qb.hit_page + d.hit_off, qb.hit_pos + d.hit_off, qb.page_first + d.hit_off,
             qb.page_rank + d.hit_off, qb.npages_found + q, qb.nhits + q, qb.max_pages, sm);
}
void launch_query_pages(const QBatch& qb, const PagesPass& pp, cudaStream_t s) {
    if (!qb.nqueries) return;
    cudaFuncSetAttribute(k_query_pages, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(QSmem));
    k_query_pages<<<qb.nqueries, QT, sizeof(QSmem), s>>>(qb, pp);
}

__global__ void k_copy_segments(const u64* __restrict__ src, u64* __restrict__ dst, const u64* __restrict__ seg_src,
                                const u64* __restrict__ seg_dst, const u32* __restrict__ seg_cnt) {
    const u32 i = blockIdx.x;
    const u64* a = src + seg_src[i];
    u64* b = dst + seg_dst[i];
    for (u32 k = threadIdx.x; k < seg_cnt[i]; k += blockDim.x) b[k] = a[k];
}
void launch_copy_segments(const u64* src, u64* dst, const u64* seg_src, const u64* seg_dst, const u32* seg_cnt, u32 nseg, cudaStream_t s) {
    if (nseg) k_copy_segments<<<nseg, 256, 0, s>>>(src, dst, seg_src, seg_dst, seg_cnt);
}
__global__ void k_first_coords(const u16* __restrict__ arena, const u64* __restrict__ code_off, const u64* __restrict__ code_base, u32 ncodes,
                               u64* __restrict__ first) {
    const u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncodes) return;
    const u64 a = code_off[c], e = code_off[c + 1];
    u64 v = ~0ull;
    if (e > a) {   // SeqEnumerator.MoveNext on the first value (IndexSequence.cs:114-136)
        v = code_base ? code_base[c] : 0ull;
        int sh = 0;
        for (u64 j = a; j < e; j++) {
            const u16 w = arena[j];
            v += (u64)(w & 0x7FFF) << sh;
            sh += 15;
            if (!(w & 0x8000)) break;
        }
    }
    first[c] = v;
}
void launch_first_coords(const u16* arena, const u64* code_off, const u64* code_base, u32 ncodes, u64* first, cudaStream_t s) {
    if (ncodes) k_first_coords<<<(ncodes + 255) / 256, 256, 0, s>>>(arena, code_off, code_base, ncodes, first);
}
__global__ void k_next_first(const u64* __restrict__ all, int nranks, int me, u32 row, u32 ncodes, u64* __restrict__ out) {
    const u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncodes) return;
    u64 m = ~0ull;
    for (int r = me + 1; r < nranks; r++) m = min(m, all[(size_t)r * row + c]);
    out[c] = m;
}
void launch_next_first(const u64* all, int nranks, int me, u32 row, u32 ncodes, u64* out, cudaStream_t s) {
    if (ncodes) k_next_first<<<(ncodes + 255) / 256, 256, 0, s>>>(all, nranks, me, row, ncodes, out);
}

// ---- decoded-list cache: chunk-parallel decode ---------------------------------------------------------------------------
constexpr int DJ_THREADS = 256, DJ_ITEMS = 16, DJ_CHUNK = DJ_THREADS * DJ_ITEMS;
DCD_D void dj_contrib(const u16* __restrict__ w, u64 n, u64 j, u64& c, u32& t) {   // as in blk_decode
    c = 0;
    t = 0;
    if (j < n) {
        const u16 v = w[j];
        int g = 0;
        while (g < 4 && j > (u64)g && (w[j - 1 - g] & 0x8000)) g++;
        c = (u64)(v & 0x7FFF) << (15 * g);
        t = (v & 0x8000) ? 0u : 1u;
    }
}
__global__ void __launch_bounds__(DJ_THREADS) k_dj_partial(const u16* __restrict__ arena, DecodeJobs dj) {
    __shared__ u64 s64[DJ_THREADS / 32 + 1];
    __shared__ u32 s32[DJ_THREADS / 32 + 1];
    const u32 ch = blockIdx.x;
    const u32 job = dj.chunk_job[ch];
    const u16* w = arena + dj.word_off[job];
    const u64 n = dj.nwords[job];
    const u64 j0 = (u64)dj.chunk_idx[ch] * DJ_CHUNK + (u64)threadIdx.x * DJ_ITEMS;
    u64 sum = 0;
    u32 cnt = 0;
#pragma unroll 4
    for (int k = 0; k < DJ_ITEMS; k++) {
        u64 c;
        u32 t;
        dj_contrib(w, n, j0 + k, c, t);
        sum += c;
        cnt += t;
    }
    u64 ts;
    u32 tc;
    block_excl_scan<u64, DJ_THREADS>(sum, s64, &ts);
    block_excl_scan<u32, DJ_THREADS>(cnt, s32, &tc);
    if (threadIdx.x == 0) { dj.psum[ch] = ts; dj.pcnt[ch] = tc; }
}
// one warp per job: exclusive scan of its chunk totals, in place
__global__ void k_dj_scan(DecodeJobs dj) {
    const u32 job = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (job >= dj.njobs) return;
    const u32 c0 = dj.job_first_chunk[job], c1 = dj.job_first_chunk[job + 1];
    u64 rs = 0;
    u32 rc = 0;
    for (u32 b = c0; b < c1; b += 32) {
        const u32 c = b + lane_id();
        const u64 v = c < c1 ? dj.psum[c] : 0;
        const u32 t = c < c1 ? dj.pcnt[c] : 0;
        const u64 iv = warp_incl_scan(v);
        const u32 it = warp_incl_scan(t);
        if (c < c1) { dj.psum[c] = rs + iv - v; dj.pcnt[c] = rc + it - t; }
        rs += __shfl_sync(0xffffffffu, iv, 31);
        rc += __shfl_sync(0xffffffffu, it, 31);
    }
}
__global__ void __launch_bounds__(DJ_THREADS) k_dj_write(const u16* __restrict__ arena, DecodeJobs dj) {
    __shared__ u64 s64[DJ_THREADS / 32 + 1];
    __shared__ u32 s32[DJ_THREADS / 32 + 1];
    const u32 ch = blockIdx.x;
    const u32 job = dj.chunk_job[ch];
    const u16* w = arena + dj.word_off[job];
    const u64 n = dj.nwords[job];
    const u64 j0 = (u64)dj.chunk_idx[ch] * DJ_CHUNK + (u64)threadIdx.x * DJ_ITEMS;
    u64 c[DJ_ITEMS];
    u32 t[DJ_ITEMS];
    u64 sum = 0;
    u32 cnt = 0;
#pragma unroll
    for (int k = 0; k < DJ_ITEMS; k++) {
        dj_contrib(w, n, j0 + k, c[k], t[k]);
        sum += c[k];
        cnt += t[k];
    }
    u64 ts;
    u32 tc;
    u64 run = dj.base[job] + dj.psum[ch] + block_excl_scan<u64, DJ_THREADS>(sum, s64, &ts);
    u32 idx = dj.pcnt[ch] + block_excl_scan<u32, DJ_THREADS>(cnt, s32, &tc);
    u64* out = dj.out[job];
#pragma unroll
    for (int k = 0; k < DJ_ITEMS; k++) {
        run += c[k];
        if (t[k]) out[idx++] = run;
    }
}
void launch_decode_jobs(const u16* arena, const DecodeJobs& dj, cudaStream_t s) {
    if (!dj.nchunks) return;
    k_dj_partial<<<dj.nchunks, DJ_THREADS, 0, s>>>(arena, dj);
    k_dj_scan<<<(dj.njobs * 32 + 127) / 128, 128, 0, s>>>(dj);
    k_dj_write<<<dj.nchunks, DJ_THREADS, 0, s>>>(arena, dj);
}

// one warp per code: count the terminator words of its stream
__global__ void k_code_nvals(const u16* __restrict__ arena, const u64* __restrict__ code_off, u32 ncodes, u32* __restrict__ nvals) {
    const u32 c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (c >= ncodes) return;
    const u64 a = code_off[c], e = code_off[c + 1];
    u32 cnt = 0;
    for (u64 j = a + lane_id(); j < e; j += 32) cnt += (arena[j] & 0x8000u) ? 0u : 1u;
    cnt = warp_sum(cnt);
    if (lane_id() == 0) nvals[c] = cnt;
}
void launch_code_nvals(const u16* arena, const u64* code_off, u32 ncodes, u32* nvals, cudaStream_t s) {
    if (ncodes) k_code_nvals<<<(unsigned)(((u64)ncodes * 32 + 255) / 256), 256, 0, s>>>(arena, code_off, ncodes, nvals);
}

__global__ void k_pack_results(QBatch qb, const u64* __restrict__ hoff, const u64* __restrict__ poff, u32* __restrict__ o_page,
                               int* __restrict__ o_pos, u32* __restrict__ o_first, float* __restrict__ o_rank) {
    const u32 q = blockIdx.x;
    const QDesc d = qb.desc[q];
    const u32 nh = qb.nhits[q], np = qb.npages_found[q];
    const u64 ho = hoff[q], po = poff[q];
    for (u32 i = threadIdx.x; i < nh; i += blockDim.x) {
        o_page[ho + i] = qb.hit_page[d.hit_off + i];
        o_pos[ho + i] = qb.hit_pos[d.hit_off + i];
    }
    for (u32 i = threadIdx.x; i < np; i += blockDim.x) {
        o_first[po + i] = qb.page_first[d.hit_off + i];
        o_rank[po + i] = qb.page_rank[d.hit_off + i];
    }
}
__global__ void k_pack_pairs(QBatch qb, const u64* __restrict__ poff, long long page_base, int* __restrict__ out) {
    const u32 q = blockIdx.x;
    const QDesc d = qb.desc[q];
    const u32 nh = qb.nhits[q], np = qb.npages_found[q];
    const u64 po = poff[q];
    for (u32 k = threadIdx.x; k < np; k += blockDim.x) {
        const u32 a = qb.page_first[d.hit_off + k], e = k + 1 < np ? qb.page_first[d.hit_off + k + 1] : nh;
        out[2 * (po + k)] = (int)((long long)qb.hit_page[d.hit_off + a] + page_base);
        out[2 * (po + k) + 1] = k >= qb.max_pages ? 1 : (int)(e - a);
    }
}
void launch_pack_pairs(const QBatch& qb, const u64* page_dense_off, long long page_base, int* out_pairs, cudaStream_t s) {
    if (qb.nqueries) k_pack_pairs<<<qb.nqueries, 256, 0, s>>>(qb, page_dense_off, page_base, out_pairs);
}
__global__ void k_gather_pairs(const u64* __restrict__ src, const u64* __restrict__ dst_off, const u32* __restrict__ cnt, int* __restrict__ out) {
    const u32 r = blockIdx.x;
    const int2* sp = reinterpret_cast<const int2*>(src[r]);
    int2* dp = reinterpret_cast<int2*>(out) + dst_off[r];
    for (u32 k = threadIdx.x; k < cnt[r]; k += blockDim.x) dp[k] = sp[k];
}
void launch_gather_pairs(const u64* src, const u64* dst_off, const u32* cnt, u32 nreq, int* out, cudaStream_t s) {
    if (nreq) k_gather_pairs<<<nreq, 256, 0, s>>>(src, dst_off, cnt, out);
}
void launch_pack_results(const QBatch& qb, const u64* hoff, const u64* poff, u32* o_page, int* o_pos, u32* o_first, float* o_rank,
                         cudaStream_t s) {
    if (!qb.nqueries) return;
    k_pack_results<<<qb.nqueries, 256, 0, s>>>(qb, hoff, poff, o_page, o_pos, o_first, o_rank);
}

// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(QT) k_seq_op(int op, const u16* a, u64 na, const u16* b, u64 nb, u32 absR, u32 ordered, u64* va,
                                               u64* vb, u8* tmp_tag, u64* scratch, u32 gallop_ratio, u64* out, u32* out_n) {
    extern __shared__ __align__(16) unsigned char q_smem_raw[];
    QSmem& sm = *reinterpret_cast<QSmem*>(q_smem_raw);
    const u32 nA = blk_decode(a, na, va, sm);
    const u32 nB = blk_decode(b, nb, vb, sm);
    u32 n = 0;
    if (op == 1) n = blk_add(va, nA, vb, nB, out, sm);
    else {   // same dispatch as k_query_batch
        bool done = false;
        const u32 small = min(nA, nB), big = max(nA, nB);
        if (gallop_ratio && absR != 0 && small && (u64)small * gallop_ratio <= big)
            done = nA <= nB ? blk_mul_gallop(va, nA, vb, nB, absR, ordered, true, scratch, tmp_tag, out, &n, sm)
                            : blk_mul_gallop(vb, nB, va, nA, absR, ordered, false, scratch, tmp_tag, out, &n, sm);
        if (!done) n = blk_mul(va, nA, vb, nB, absR, ordered, tmp_tag, out, sm);
    }
    if (threadIdx.x == 0) *out_n = n;
}
void launch_seq_op(int op, const u16* a, u64 na, const u16* b, u64 nb, u32 absR, u32 ordered, u64* va, u64* vb, u8* tmp_tag,
                   u64* scratch, u32 gallop_ratio, u64* out, u32* out_n, cudaStream_t s) {
    cudaFuncSetAttribute(k_seq_op, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(QSmem));
    k_seq_op<<<1, QT, sizeof(QSmem), s>>>(op, a, na, b, nb, absR, ordered, va, vb, tmp_tag, scratch, gallop_ratio, out, out_n);
}
__global__ void __launch_bounds__(QT) k_seq_decode(const u16* w, u64 n, u64* out, u32* out_n) {
    extern __shared__ __align__(16) unsigned char q_smem_raw[];
    QSmem& sm = *reinterpret_cast<QSmem*>(q_smem_raw);
    const u32 c = blk_decode(w, n, out, sm);
    if (threadIdx.x == 0) *out_n = c;
}
void launch_seq_decode(const u16* w, u64 n, u64* out, u32* out_n, cudaStream_t s) {
    cudaFuncSetAttribute(k_seq_decode, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(QSmem));
    k_seq_decode<<<1, QT, sizeof(QSmem), s>>>(w, n, out, out_n);
}

}  // namespace dcd
