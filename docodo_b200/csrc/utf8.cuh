// UTF-8 page input (DCD_ENC_UTF8).  The reference hands the path decoded .NET strings (DataSources.cs:357-379) and counts
// coordinates in UTF-16 code units (`coord += c.Length`, Build.cs:529-531), so UTF-8 text is brought to that unit axis first:
//
//   k_utf8_scan       one warp per 1024-byte block: UTF-16 units the block decodes to (a non-continuation byte is one unit,
//                     a 4-byte lead two: a surrogate pair), plus "has a byte >= 0x80" / "has a code point >= U+0100" flags
//   (scan)            block unit bases
//   k_utf8_transcode  same blocks: every lead byte decodes its code point and stores it at base + in-block unit prefix,
//                     as Latin-1 bytes when no code point reaches U+0100, else as UTF-16LE units
//   k_utf8_page_units page byte offsets -> unit offsets (block base + the units of the block's bytes before the offset)
//
// Pure-ASCII text needs none of the copies: its bytes ARE DCD_ENC_LATIN1 text with identical offsets, so only the scan runs.
// The text must be well-formed UTF-8 (what Encoding.UTF8.GetBytes produces; lone surrogates arrive as U+FFFD = EF BF BD).
// A stray continuation byte decodes to nothing and a truncated sequence to whatever bits it has: never out of bounds, but the
// coordinates of malformed text are not those of .NET's replacement-character decoding.
#pragma once
#include "common.cuh"

namespace dcd {

constexpr int U8_BLOCK = 1024;   // bytes per warp iteration (32 per lane, two 128-bit loads)

// flags[0] |= 1: some byte >= 0x80; |= 2: some code point >= U+0100 (UTF-16 output needed).  block_units[nblocks].
void launch_utf8_scan(const u8* text, u64 nbytes, u64* block_units, u32* flags, int nsm, cudaStream_t s);
// block_base = EXCLUSIVE prefix of block_units.  wide != 0: out is u16 units, else u8 Latin-1 bytes.
void launch_utf8_transcode(const u8* text, u64 nbytes, const u64* block_base, void* out, int wide, int nsm, cudaStream_t s);
// unit_off[p] for the npages + 1 byte offsets of page_off (page_off[npages] == nbytes -> total units)
void launch_utf8_page_units(const u8* text, u64 nbytes, const u64* block_base, u64 total_units, const u64* page_off, u64 n,
                            u64* unit_off, cudaStream_t s);

}  // namespace dcd
