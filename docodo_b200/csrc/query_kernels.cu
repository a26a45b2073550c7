#include "query_kernels.cuh"

namespace dcd {

constexpr int QT = 256;          // threads per CTA
constexpr int QI = 4;            // items per thread per tile
constexpr int QTILE = QT * QI;
constexpr int QSTACK = 16;

struct QSmem {
    u64 s64[QT / 32 + 1];
    u32 s32[QT / 32 + 1];
    u64 smax[QT / 32 + 1];
    u64* st_ptr[QSTACK];
    u32 st_n[QSTACK];
    u32 bcast;
};

// block-wide inclusive scan with componentwise max over the two 32-bit halves of a u64
DCD_D u64 max2(u64 a, u64 b) {
    u32 ah = (u32)(a >> 32), al = (u32)a, bh = (u32)(b >> 32), bl = (u32)b;
    return ((u64)max(ah, bh) << 32) | (u64)max(al, bl);
}
// Block-wide EXCLUSIVE scan under max2; *total = reduction over the block.  smem: QT/32+1 entries.
DCD_D u64 block_excl_scan_max2(u64 v, u64* smem, u64* total) {
    u64 inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        u64 t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane_id() >= (u32)o) inc = max2(inc, t);
    }
    if (lane_id() == 31) smem[warp_id()] = inc;
    __syncthreads();
    if (warp_id() == 0) {
        u64 w = lane_id() < QT / 32 ? smem[lane_id()] : 0;
        u64 wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            u64 t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane_id() >= (u32)o) wi = max2(wi, t);
        }
        u64 wex = __shfl_up_sync(0xffffffffu, wi, 1);
        if (lane_id() == 0) wex = 0;
        if (lane_id() < QT / 32) smem[lane_id()] = wex;
        if (lane_id() == QT / 32 - 1) smem[QT / 32] = wi;
    }
    __syncthreads();
    u64 lane_ex = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane_id() == 0) lane_ex = 0;
    const u64 r = max2(lane_ex, smem[warp_id()]);
    *total = smem[QT / 32];
    __syncthreads();
    return r;
}

// ---- decode (SeqEnumerator.MoveNext, IndexSequence.cs:114-136) -------------------------------------
// coordinate at a terminator word == inclusive prefix sum of ((w & 0x7FFF) << 15*g) over the whole stream, where g is
// the number of continuation words directly in front of the word; the value index is the count of terminators before it.
DCD_D u32 blk_decode(const u16* __restrict__ w, u64 n, u64* __restrict__ out, QSmem& sm) {
    const int tid = threadIdx.x;
    u64 carry_sum = 0;
    u32 carry_cnt = 0;
    for (u64 base = 0; base < n; base += QTILE) {
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

// ---- stable merge with source tags (1 = first list, 2 = second) --------------------------------------
DCD_D void blk_merge_tags(const u64* __restrict__ A, u32 nA, const u64* __restrict__ B, u32 nB, u64* __restrict__ M,
                          u8* __restrict__ T) {
    for (u32 i = threadIdx.x; i < nA; i += QT) {
        const u64 v = A[i];
        u32 lo = 0, hi = nB;   // # of B elements < v
        while (lo < hi) { u32 mid = (lo + hi) >> 1; if (B[mid] < v) lo = mid + 1; else hi = mid; }
        M[i + lo] = v;
        T[i + lo] = 1;
    }
    for (u32 j = threadIdx.x; j < nB; j += QT) {
        const u64 v = B[j];
        u32 lo = 0, hi = nA;   // # of A elements <= v
        while (lo < hi) { u32 mid = (lo + hi) >> 1; if (A[mid] <= v) lo = mid + 1; else hi = mid; }
        M[j + lo] = v;
        T[j + lo] = 2;
    }
    __syncthreads();
}

// an element of the second list that repeats the value in front of it is the same coordinate (tag "both")
DCD_D bool el_valid(const u64* M, const u8* T, u32 p) { return !(T[p] == 2 && p > 0 && M[p - 1] == M[p]); }
DCD_D u32 el_tag(const u64* M, const u8* T, u32 p, u32 m) {
    return (T[p] == 1 && p + 1 < m && T[p + 1] == 2 && M[p + 1] == M[p]) ? 3u : (u32)T[p];
}

// operator + : sorted union with duplicates across the lists removed (IndexSequence.cs:286-322)
DCD_D u32 blk_add_finish(const u64* __restrict__ M, const u8* __restrict__ T, u32 m, u64* __restrict__ out, QSmem& sm) {
    const int tid = threadIdx.x;
    u32 carry = 0;
    for (u32 base = 0; base < m; base += QTILE) {
        const u32 p0 = base + tid * QI;
        u32 keep[QI];
        u32 cnt = 0;
#pragma unroll
        for (int k = 0; k < QI; k++) {
            const u32 p = p0 + k;
            keep[k] = (p < m && el_valid(M, T, p)) ? 1u : 0u;
            cnt += keep[k];
        }
        u32 tot;
        u32 idx = carry + block_excl_scan<u32, QT>(cnt, sm.s32, &tot);
#pragma unroll
        for (int k = 0; k < QI; k++)
            if (keep[k]) out[idx++] = M[p0 + k];
        carry += tot;
        __syncthreads();
    }
    return carry;
}

// operator * (IndexSequence.cs:218-284) in data-parallel form:
//   runs   = maximal stretches of the tagged union whose consecutive gaps are <= absR (absR == 0: one run)
//   groups = runs, except that in ordered mode (both operands "exact", R < 0) the elements of a run in front of its
//            first element carrying tag 1 are split off as their own (tag-2-only, hence dropped) group
//   output = every element of every group that contains tag 1 and tag 2
DCD_D u32 blk_mul_finish(const u64* __restrict__ M, const u8* __restrict__ T, u32 m, u32 absR, u32 ordered, u32* __restrict__ gid,
                         u32* __restrict__ gflag, u64* __restrict__ out, QSmem& sm) {
    const int tid = threadIdx.x;
    for (u32 p = tid; p < m; p += QT) gflag[p] = 0;
    __syncthreads();
    // pass 1: group ids + group flags
    u32 carry_g = 0;      // groups started so far
    u64 carry_max = 0;    // hi: (index+1) of the last run start, lo: (index+1) of the last tag-1 element
    for (u32 base = 0; base < m; base += QTILE) {
        const u32 p0 = base + tid * QI;
        u32 valid[QI], tag[QI], start[QI];
        u64 loc = 0;      // running max2 within the thread (inclusive)
        u64 pre[QI];      // thread-local EXCLUSIVE max2 in front of item k
#pragma unroll
        for (int k = 0; k < QI; k++) {
            const u32 p = p0 + k;
            valid[k] = tag[k] = start[k] = 0;
            pre[k] = loc;
            if (p < m && el_valid(M, T, p)) {
                valid[k] = 1;
                tag[k] = el_tag(M, T, p, m);
                start[k] = (p == 0 || (absR != 0 && M[p] - M[p - 1] > (u64)absR)) ? 1u : 0u;
                u64 mine = (start[k] ? ((u64)(p + 1) << 32) : 0ull) | ((tag[k] & 1u) ? (u64)(p + 1) : 0ull);
                loc = max2(loc, mine);
            }
        }
        u64 tot_max;
        const u64 excl_thread = max2(carry_max, block_excl_scan_max2(loc, sm.smax, &tot_max));
        u32 gs[QI];
        u32 cnt = 0;
#pragma unroll
        for (int k = 0; k < QI; k++) {
            gs[k] = 0;
            if (valid[k]) {
                const u32 p = p0 + k;
                const u64 before = max2(excl_thread, pre[k]);              // state in front of p
                const u32 last_t1 = (u32)before;                            // (index+1) of the last tag-1 element before p
                const u32 run_start = start[k] ? p + 1 : (u32)(before >> 32);   // (index+1) of p's run start
                gs[k] = start[k] | ((ordered && (tag[k] & 1u) && last_t1 < run_start) ? 1u : 0u);
                cnt += gs[k];
            }
        }
        u32 tot;
        u32 g = carry_g + block_excl_scan<u32, QT>(cnt, sm.s32, &tot);
#pragma unroll
        for (int k = 0; k < QI; k++) {
            if (valid[k]) {
                g += gs[k];
                const u32 p = p0 + k;
                gid[p] = g - 1;
                // warp-aggregated OR into the group's flag word
                u32 bits = tag[k];
                if ((__ldcg(&gflag[g - 1]) & bits) != bits) atomicOr(&gflag[g - 1], bits);
            }
        }
        carry_g += tot;
        carry_max = max2(carry_max, tot_max);
        __syncthreads();
    }
    __threadfence_block();
    __syncthreads();
    // pass 2: keep the elements of complete groups
    u32 carry = 0;
    for (u32 base = 0; base < m; base += QTILE) {
        const u32 p0 = base + tid * QI;
        u32 keep[QI];
        u32 cnt = 0;
#pragma unroll
        for (int k = 0; k < QI; k++) {
            const u32 p = p0 + k;
            keep[k] = (p < m && el_valid(M, T, p) && __ldcg(&gflag[gid[p]]) == 3u) ? 1u : 0u;
            cnt += keep[k];
        }
        u32 tot;
        u32 idx = carry + block_excl_scan<u32, QT>(cnt, sm.s32, &tot);
#pragma unroll
        for (int k = 0; k < QI; k++)
            if (keep[k]) out[idx++] = M[p0 + k];
        carry += tot;
        __syncthreads();
    }
    return carry;
}

// ---- hits -> pages (K9) --------------------------------------------------------------------------
DCD_D void blk_hits(const u64* __restrict__ R, u32 n, const u64* __restrict__ page_end, u64 npages, u64 origin, u32* __restrict__ hit_page,
                    int* __restrict__ hit_pos, u32* __restrict__ page_first, float* __restrict__ page_rank, u32* n_pages_out,
                    QSmem& sm) {
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
    const u32 np = carry;
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

// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(QT) k_query_batch(QBatch qb) {
    __shared__ QSmem sm;
    const u32 q = blockIdx.x;
    if (q >= qb.nqueries) return;
    const QDesc d = qb.desc[q];
    u64* lists = qb.lists + d.list_off;
    u64* tmp_val = qb.tmp_val + d.tmp_off;
    u64* tmp_dec = qb.tmp_dec + d.tmp_off;
    u32* tmp_gid = qb.tmp_gid + d.tmp_off;
    u32* tmp_gflag = qb.tmp_gflag + d.tmp_off;
    u8* tmp_tag = qb.tmp_tag + d.tmp_off;
    int sp = 0;
    for (u32 pc = d.instr_begin; pc < d.instr_end; pc++) {
        const QInstr in = qb.instr[pc];
        if (in.op == QOP_LOAD) {
            u64* out = lists + in.out_off;
            u64* cur = out;
            u32 n = 0;
            if (in.b >= 1) {
                const u32 c0 = qb.code_ids[in.a];
                const u64 a0 = qb.code_off[c0], e0 = qb.code_off[c0 + 1];
                n = blk_decode(qb.arena + a0, e0 - a0, cur, sm);
                // further codes: fold with operator+ ; the LOAD region holds two halves of `cap` elements each
                for (u32 k = 1; k < in.b; k++) {
                    const u32 ck = qb.code_ids[in.a + k];
                    const u64 ak = qb.code_off[ck], ek = qb.code_off[ck + 1];
                    const u32 nk = blk_decode(qb.arena + ak, ek - ak, tmp_dec, sm);
                    if ((u64)n + nk > d.tmp_cap) { if (threadIdx.x == 0) atomicExch(qb.error, 1u); break; }
                    blk_merge_tags(cur, n, tmp_dec, nk, tmp_val, tmp_tag);
                    u64* nxt = (cur == out) ? out + in.cap : out;
                    n = blk_add_finish(tmp_val, tmp_tag, n + nk, nxt, sm);
                    cur = nxt;
                }
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
            blk_merge_tags(A, nA, B, nB, tmp_val, tmp_tag);
            if (in.op == QOP_ADD) n = blk_add_finish(tmp_val, tmp_tag, nA + nB, out, sm);
            else n = blk_mul_finish(tmp_val, tmp_tag, nA + nB, in.a, in.b, tmp_gid, tmp_gflag, out, sm);
            sp -= 2;
            __syncthreads();
            if (threadIdx.x == 0) { sm.st_ptr[sp] = out; sm.st_n[sp] = n; }
            sp++;
            __syncthreads();
        } else if (in.op == QOP_COUNT) {
            if (sp >= 1 && threadIdx.x == 0) qb.nfound[in.a] = sm.st_n[sp - 1];
        }
    }
    u32 n = 0;
    const u64* R = nullptr;
    if (sp >= 1) { R = sm.st_ptr[sp - 1]; n = sm.st_n[sp - 1]; }
    if (n > d.hit_cap) { if (threadIdx.x == 0) atomicExch(qb.error, 1u); n = (u32)d.hit_cap; }
    if (threadIdx.x == 0) qb.nhits[q] = n;
    blk_hits(R, n, qb.page_end, qb.npages, qb.coord_origin, qb.hit_page + d.hit_off, qb.hit_pos + d.hit_off, qb.page_first + d.hit_off,
             qb.page_rank + d.hit_off, qb.npages_found + q, sm);
}

void launch_query_batch(const QBatch& qb, cudaStream_t s) {
    if (!qb.nqueries) return;
    k_query_batch<<<qb.nqueries, QT, 0, s>>>(qb);
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
void launch_pack_results(const QBatch& qb, const u64* hoff, const u64* poff, u32* o_page, int* o_pos, u32* o_first, float* o_rank,
                         cudaStream_t s) {
    if (!qb.nqueries) return;
    k_pack_results<<<qb.nqueries, 256, 0, s>>>(qb, hoff, poff, o_page, o_pos, o_first, o_rank);
}

// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(QT) k_seq_op(int op, const u16* a, u64 na, const u16* b, u64 nb, u32 absR, u32 ordered, u64* va,
                                               u64* vb, u64* tmp_val, u32* tmp_gid, u32* tmp_gflag, u8* tmp_tag, u64* out,
                                               u32* out_n) {
    __shared__ QSmem sm;
    const u32 nA = blk_decode(a, na, va, sm);
    const u32 nB = blk_decode(b, nb, vb, sm);
    blk_merge_tags(va, nA, vb, nB, tmp_val, tmp_tag);
    u32 n;
    if (op == 1) n = blk_add_finish(tmp_val, tmp_tag, nA + nB, out, sm);
    else n = blk_mul_finish(tmp_val, tmp_tag, nA + nB, absR, ordered, tmp_gid, tmp_gflag, out, sm);
    if (threadIdx.x == 0) *out_n = n;
}
void launch_seq_op(int op, const u16* a, u64 na, const u16* b, u64 nb, u32 absR, u32 ordered, u64* va, u64* vb, u64* tmp_val,
                   u32* tmp_gid, u32* tmp_gflag, u8* tmp_tag, u64* out, u32* out_n, cudaStream_t s) {
    k_seq_op<<<1, QT, 0, s>>>(op, a, na, b, nb, absR, ordered, va, vb, tmp_val, tmp_gid, tmp_gflag, tmp_tag, out, out_n);
}
__global__ void __launch_bounds__(QT) k_seq_decode(const u16* w, u64 n, u64* out, u32* out_n) {
    __shared__ QSmem sm;
    const u32 c = blk_decode(w, n, out, sm);
    if (threadIdx.x == 0) *out_n = c;
}
void launch_seq_decode(const u16* w, u64 n, u64* out, u32* out_n, cudaStream_t s) { k_seq_decode<<<1, QT, 0, s>>>(w, n, out, out_n); }

}  // namespace dcd
