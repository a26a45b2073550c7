#include "synth.cuh"

namespace dcd {

struct Rng {   // xoshiro256** seeded by SplitMix64
    u64 s[4];
    DCD_D static u64 splitmix(u64& x) {
        u64 z = (x += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    DCD_D void seed(u64 x) { for (int i = 0; i < 4; i++) s[i] = splitmix(x); }
    DCD_D static u64 rotl(u64 x, int k) { return (x << k) | (x >> (64 - k)); }
    DCD_D u64 next() {
        const u64 r = rotl(s[1] * 5, 7) * 9, t = s[1] << 17;
        s[2] ^= s[0]; s[3] ^= s[1]; s[1] ^= s[2]; s[0] ^= s[3]; s[2] ^= t; s[3] = rotl(s[3], 45);
        return r;
    }
    DCD_D double uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
};

// One document = pages_per_doc pages; a page is words until >= min_page_chars characters.  Words are separated by one
// space; a sentence ends (". ") after Geometric(mean 12) words and the next word is capitalised; with p = 0.01 a token is
// a 1-6 digit number.  WRITE == false only measures page lengths (identical random sequence).
template <bool WRITE>
__global__ void k_synth(SynthParams sp, u64* __restrict__ page_len, const u64* __restrict__ page_offs, u8* __restrict__ text) {
    const u64 d = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= sp.ndocs) return;
    Rng rng;
    rng.seed(sp.seed ^ ((sp.first_doc + d) * 0xD6E8FEB86659FD93ull));
    bool cap = true;
    for (u32 p = 0; p < sp.pages_per_doc; p++) {
        const u64 pg = d * sp.pages_per_doc + p;
        u8* out = WRITE ? text + page_offs[pg] : nullptr;
        u64 len = 0;
        while (len < sp.min_page_chars) {
            const double u = rng.uniform();
            if (u < 0.01) {
                const u32 nd = 1 + (u32)(rng.next() % 6);
                u64 v = rng.next();
                for (u32 k = 0; k < nd; k++) { if (WRITE) out[len] = (u8)('0' + v % 10); v /= 10; len++; }
            } else {
                const double x = rng.uniform();
                u32 lo = 0, hi = sp.nwords - 1;   // first rank with cdf >= x
                while (lo < hi) { u32 mid = (lo + hi) >> 1; if (sp.cdf[mid] >= x) hi = mid; else lo = mid + 1; }
                const u32 a = sp.word_offs[lo], e = sp.word_offs[lo + 1];
                for (u32 k = a; k < e; k++) {
                    if (WRITE) { u8 ch = sp.words[k]; if (cap && k == a && ch >= 'a' && ch <= 'z') ch -= 32; out[len] = ch; }
                    len++;
                }
            }
            cap = false;
            if (rng.uniform() < 1.0 / 12.0) { if (WRITE) out[len] = '.'; len++; cap = true; }
            if (WRITE) out[len] = ' ';
            len++;
        }
        if (!WRITE) page_len[pg] = len;
    }
}
void launch_synth_lengths(const SynthParams& sp, u64* page_len, cudaStream_t s) {
    if (!sp.ndocs) return;
    k_synth<false><<<(unsigned)((sp.ndocs + 127) / 128), 128, 0, s>>>(sp, page_len, nullptr, nullptr);
}
void launch_synth_write(const SynthParams& sp, const u64* page_offs, u8* text, cudaStream_t s) {
    if (!sp.ndocs) return;
    k_synth<true><<<(unsigned)((sp.ndocs + 127) / 128), 128, 0, s>>>(sp, nullptr, page_offs, text);
}
__global__ void k_doc_pages(u64 ndocs, u32 ppd, u64* __restrict__ out) {
    const u64 d = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (d <= ndocs) out[d] = d * ppd;
}
void launch_synth_doc_pages(u64 ndocs, u32 ppd, u64* out, cudaStream_t s) {
    k_doc_pages<<<(unsigned)((ndocs + 1 + 255) / 256), 256, 0, s>>>(ndocs, ppd, out);
}

}  // namespace dcd
