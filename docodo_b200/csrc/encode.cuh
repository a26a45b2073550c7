// K6/K7: IndexSequence.Builder.Add for every posting at once (IndexSequence.cs:63-84): segmented adjacent
// difference over the sorted keys, 15-bit-group varint lengths, single-pass chained scan for the output offsets,
// shared-memory staged coalesced write of the ushort stream.  Also a generic u64 scan used for page coordinates.
#pragma once
#include "common.cuh"

namespace dcd {

constexpr int ENC_THREADS = 256;
constexpr int ENC_ITEMS = 8;
constexpr int ENC_TILE = ENC_THREADS * ENC_ITEMS;   // postings per tile

// keys sorted by (code, coord).  arena receives the concatenated ushort streams in code order;
// code_off[c] = offset (in ushorts) of code c's stream, code_off[ncodes] = total.  first_delta_base (nullable):
// per-code value subtracted from the first coordinate of the code (multi-GPU stitch == IndexSequence.Shift).
// lookback: [ntiles] u64 zeroed; ticket zeroed; total_out[0] = total ushorts.
// max_coord_out (nullable, zeroed): receives the largest coordinate (Builder.maxCoord, Build.cs:326).
// glob (nullable): the keys carry dense LOCAL code ids (sharded build) and glob[local] is the global ordinal id that indexes
// code_off / first_delta_base.
void launch_encode(const u64* keys, u64 n, u32 ncodes, u16* arena, u64 arena_cap, u64* code_off, const u64* first_delta_base,
                   const u32* glob, u64* lookback, u32* ticket, u64* total_out, u64* max_coord_out, u32* error, cudaStream_t s,
                   const u64* key_base = nullptr /* first_delta_base re-indexed by the keys' local ids: key_base[l] = base[glob[l]] */,
                   u32 n_local = 0 /* entries of glob (needed with glob + first_delta_base: absent codes are filled per present code) */);
size_t encode_scratch_tiles(u64 n);
// The same encoder in two steps (sharded build: the list lengths feed the owner split before any ushort is stored).
// launch_encode_lengths: code_off / total / max coordinate as launch_encode, nothing stored; `lookback`
// (encode_lengths_scratch(n) words, zeroed) keeps every tile's inclusive ushort prefix in its first ntiles words.  launch_encode_placed: the ushorts of the key tiles [tile_first, tile_end) (ENC_TILE
// keys each) at the places those prefixes give; tile ranges may be encoded in any order and on any stream.
constexpr int ENC_LENGTHS = 1, ENC_PLACED = 2;
size_t encode_lengths_scratch(u64 n);   // u64 words of `lookback` launch_encode_lengths needs (zeroed)
void launch_encode_lengths(const u64* keys, u64 n, u32 ncodes, u64* code_off, const u64* first_delta_base, const u32* glob,
                           u64* lookback, u32* ticket, u64* total_out, u64* max_coord_out, cudaStream_t s, const u64* key_base = nullptr,
                           u32 n_local = 0);
void launch_encode_placed(const u64* keys, u64 n, u64 tile_first, u64 tile_end, u16* arena, u64 arena_cap, const u64* first_delta_base,
                          const u32* glob, const u64* lookback, u32* error, cudaStream_t s, const u64* key_base = nullptr);

// out[i] = base + inclusive_sum(in[0..i]); single pass, decoupled look-back.  lookback: [ceil(n/2048)] zeroed.
void launch_inclusive_scan_u64(const u64* in, u64* out, u64 n, u64 base, u64* lookback, u32* ticket, cudaStream_t s);

// Decode one or many encoded lists (SeqEnumerator.MoveNext, IndexSequence.cs:114-136): see query.cu for the batched
// form; this one decodes a single list of n ushorts into at most n values.  count_out[0] = number of values.
void launch_decode_single(const u16* words, u64 n, u64* out, u64* lookback, u32* ticket, u64* count_out, cudaStream_t s);

}  // namespace dcd
