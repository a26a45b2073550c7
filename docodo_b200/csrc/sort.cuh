// K5: stable onesweep LSD radix sort (hand-written; replaces the SortedDictionary accumulation of Build.cs:324-341
// and the k-way MergeFiles rounds of Index.cs:537-689 with one device sort of packed keys).
#pragma once
#include "common.cuh"

namespace dcd {

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 16;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;   // 4096 keys per tile

struct SortPlan {      // digit passes covering bits [begin_bit, end_bit)
    int npasses;
    int shift[8];
    int bits[8];
};
SortPlan make_sort_plan(int begin_bit, int end_bit);

// scratch needed by radix_sort_* for n keys: bytes
size_t sort_scratch_bytes(u64 n);

// Histogram of every pass digit in one read of the keys: hist[pass*256 + digit] (u64), must be zeroed by the caller.
void launch_histogram(const u64* keys, u64 n, const SortPlan& plan, u64* hist, cudaStream_t s);

// Runs the passes of `plan` ping-ponging between (a, va) and (b, vb); values may be null (keys only).
// `hist` holds the per-pass digit histograms (device, from launch_histogram or the fused emit kernel).
// If host_hist is non-null it is used to skip passes whose digit is uniform.  Returns 0 if the result is in a, 1 if in b.
int radix_sort_run(u64* a, u64* b, u32* va, u32* vb, u64 n, const SortPlan& plan, u64* hist, const u64* host_hist,
                   void* scratch, cudaStream_t s, u64* launches);

// ---- wide-digit variant for the posting sort: digits of up to 11 bits (2048 bins), 8192-key tiles, keys only.
// 20 code-id bits are two passes instead of three.  hist layout: [pass][WIDE_HIST_STRIDE] u64.
constexpr int WIDE_MAX_BITS = 11;
constexpr int WIDE_HIST_STRIDE = 1 << WIDE_MAX_BITS;
constexpr int WIDE_THREADS = 512;
constexpr int WIDE_ITEMS = 16;
constexpr int WIDE_TILE = WIDE_THREADS * WIDE_ITEMS;   // 8192 keys per tile
SortPlan make_wide_plan(int begin_bit, int end_bit);   // ceil(bits / 11) balanced passes
size_t wide_scratch_bytes(u64 n, int npasses);
void launch_histogram_wide(const u64* keys, u64 n, const SortPlan& plan, u64* hist, cudaStream_t s);
// Runs every pass of `plan` ping-ponging between a and b (stable).  Returns 0 if the result is in a, 1 if in b.
int radix_sort_wide(u64* a, u64* b, u64 n, const SortPlan& plan, const u64* hist, void* scratch, cudaStream_t s, u64* launches);

// out = merge of two arrays sorted by full 64-bit key (b is small): used to fold host-supplied postings in.
void launch_merge_small(const u64* a, u64 na, const u64* b, u64 nb, u64* out, cudaStream_t s);

}  // namespace dcd
