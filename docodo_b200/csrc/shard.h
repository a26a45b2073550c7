// Document-range sharded build across GPUs (BASELINE.json north_star: "each GPU sorts its shard, and
// per-word posting segments are exchanged and merged with NCCL over NVLink").  Replaces the partial-index stitch of
// Index.MergeAll / MergeFiles(shift_coords) / MergeIndexes (Index.cs:537-752) and IndexSequence.Shift
// (IndexSequence.cs:191-202): every GPU indexes a contiguous range of documents with GLOBAL coordinates, the code
// dictionaries are unified, the first delta of every posting segment is taken against the previous shard's last
// coordinate of that code, and the encoded segments travel to the rank owning the code range.
#pragma once
#include <cstddef>
#include <memory>
#include <vector>

#include "common.cuh"

namespace dcd {

// Transport of the collectives.  Two implementations (shard.cu): NCCL (one process per GPU; NVLink / NVSwitch) and an
// in-process loopback group (several contexts of one process on one device exchanging through device-to-device
// copies; every member runs on its own host thread).  All calls are collective and ordered on the caller's stream.
struct ShardComm {
    int rank = 0, nranks = 1;
    unsigned long long bytes_moved = 0;   // bytes received from peers since the last reset (reporting only)
    virtual ~ShardComm() {}
    // recv[r * bytes .. (r+1) * bytes) = rank r's send[0 .. bytes)
    virtual void allgather(const void* send, void* recv, size_t bytes, cudaStream_t s) = 0;
    // recv[offs[r] .. offs[r] + counts[r]) = rank r's send[0 .. counts[r])
    virtual void allgatherv(const void* send, void* recv, const size_t* counts, const size_t* offs, cudaStream_t s) = 0;
    // send[send_off[d] .. +send_n[d]) goes to rank d; recv[recv_off[r] .. +recv_n[r]) comes from rank r (bytes)
    virtual void alltoallv(const void* send, const size_t* send_off, const size_t* send_n, void* recv, const size_t* recv_off,
                           const size_t* recv_n, cudaStream_t s) = 0;
    virtual const char* transport() const = 0;
    virtual void abort() {}   // a member failed: release peers blocked in a collective (loopback group)
    // Peer memory: every rank offers the device buffer `mine` (inside the cudaMalloc allocation that starts at
    // alloc_base); all[r] receives a pointer THIS rank's kernels can store through into rank r's buffer (NVLink peer
    // mapping via CUDA IPC across processes, the plain pointer inside one process).  Returns false when peer
    // mappings are unavailable (the caller then uses the all-to-all path).  Collective.
    virtual bool exchange_ptrs(void* mine, void* alloc_base, void** all, cudaStream_t s) = 0;
    // Every rank's work enqueued on its stream before the call has completed when this returns.  Collective.
    virtual void barrier(cudaStream_t s) = 0;
};

// ncclUniqueId bytes (128) for the rendezvous; the host side carries them to every rank (torch.distributed / MPI / TCP).
void nccl_unique_id(unsigned char* out128);
ShardComm* make_nccl_comm(int device, const unsigned char* id128, int rank, int nranks);
// nranks members of one in-process group (out[0..nranks))
void make_local_comms(int nranks, ShardComm** out);

// ---- device kernels of the stitch (shard_kernels.cu) ---------------------------------------------------------
constexpr u64 COORD_NONE = ~0ull;
// keys sorted by (code, coord): last[c] = coordinate of the last posting of code c, COORD_NONE when c has none.
// local_of_glob (nullable): the keys carry dense local code ids; last[] is indexed by the global id either way.
void launch_code_last(const u64* keys, u64 n, u32 ncodes, const u32* local_of_glob, u64* last, cudaStream_t s);
// all_last[r][c] of every rank -> base[c] = last coordinate of code c on the nearest rank < my_rank that has one, else 0
// (the value IndexSequence.Builder.Last holds when this shard's first posting of c is appended, IndexSequence.cs:63-84).
void launch_first_base(const u64* all_last, int nranks, int my_rank, u32 ncodes, u64* base, cudaStream_t s);
// all_off[r][0..ncodes]: per-rank arena offsets.  tot[c] = sum over ranks of the segment lengths of code c (ushorts).
void launch_total_len(const u64* all_off, int nranks, u32 ncodes, u64* tot, cudaStream_t s);
// splits[o] (o = 0..nranks) = first code of owner o: code ranges of near-equal merged size from the inclusive prefix.
void launch_find_splits(const u64* tot_incl, u32 ncodes, int nranks, u32* splits, cudaStream_t s);
// cut[r * (nranks + 1) + o] = all_off[r][splits[o]]
void launch_gather_cuts(const u64* all_off, u32 ncodes, int nranks, const u32* splits, u64* cut, cudaStream_t s);
// merged_off[i] (i = 0..n_own) = tot_excl[c_begin + i] - tot_excl[c_begin]
void launch_merged_off(const u64* tot_incl, u32 c_begin, u32 n_own, u64* merged_off, cudaStream_t s);
// Fused exchange + merge over peer memory: every rank stores each of its encoded segments straight into its final place
// in the owner's merged arena (peer[o] = mapping of owner o's arena): destination of rank `me`'s segment of code c =
// peer[owner(c)] + (tot_excl[c] - tot_excl[splits[owner]]) + sum of the lengths of c's segments on ranks < me.
// A launch takes up to 16 source ranges [u0[i], u1[i]) of the arena -- normally one per owner: cut(me, o) .. cut(me, o + 1), or a
// part of each while the rest is still being encoded -- and interleaves their tiles, so the stores of a rank go to every
// owner at once (no incast however the ranks are skewed).  Pieces are clipped to their range.
// scratch: scatter_scratch_words(my_units) u32 words (first code of every source tile), filled by launch_scatter_tile_codes
// from this rank's offsets (possible before any ushort exists).
u64 scatter_scratch_words(u64 my_units);
void launch_scatter_tile_codes(const u64* my_off, u32 ncodes, u64 my_units, u32* scratch, cudaStream_t s);
void launch_scatter_ranges(const u16* my_arena, u64 my_units, const u64* all_off, const u64* tot_incl, u32 ncodes, int nranks, int me,
                           const u32* splits, u16* const* peer, const u64* u0, const u64* u1, int n, u32 tile_rot, const u32* scratch,
                           cudaStream_t s, u64 ce_min = 0 /* > 0: segments of at least ce_min ushorts are left to the copy engines */);
// list[0] = number of segments of this rank with at least ce_min ushorts, then {arena offset, length, owner, offset in the
// owner's merged arena} (u64 each) per segment: what the host hands to cudaMemcpyAsync.  cap entries.
void launch_big_segments(const u64* all_off, const u64* tot_incl, u32 ncodes, int nranks, int me, const u32* splits, u64 ce_min, u64* list,
                         u32 cap, cudaStream_t s);
// key_begin[o] (o = 0..nranks) = index of the first sorted key whose (global) code id is >= splits[o]; key_begin[nranks] = n.
// local_of_glob (nullable): the keys carry dense local ids (order preserving), absent codes map to 0xFFFFFFFF.
void launch_key_begins(const u64* keys, u64 n, const u32* splits, int nranks, const u32* glob, u64* key_begin, cudaStream_t s);
// Concatenate, per owned code, the segments received from ranks 0..R-1 (rank order == coordinate order).
// recv holds rank r's slice [all_off[r][c_begin], all_off[r][c_begin + n_own]) at unit offset slab[r].
void launch_concat_segments(const u64* all_off, u32 ncodes, int nranks, u32 c_begin, u32 n_own, const u64* merged_off,
                            const u16* recv, const u64* slab, u16* out, u64 total_units, cudaStream_t s);

}  // namespace dcd
