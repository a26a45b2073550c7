// Transports and kernels of the document-range sharded build (see shard.h).
#include "shard.h"

#include <dlfcn.h>

#include <condition_variable>
#include <cstring>
#include <map>
#include <mutex>
#include <string>

namespace dcd {

// =====================================================================================================
// NCCL transport: libnccl.so.2 is bound at run time (the copy torch already loaded when there is one, else the
// system library), so libdocodo_b200.so itself carries no link-time dependency on it.
// =====================================================================================================
namespace {
typedef struct ncclComm* ncclComm_t;
struct ncclUniqueIdT { char internal[128]; };
enum { NCCL_UINT8 = 1 };
struct NcclApi {
    void* h = nullptr;
    int (*GetUniqueId)(ncclUniqueIdT*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueIdT, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return;
        api.h = h;
#define BIND(field, sym) api.field = reinterpret_cast<decltype(api.field)>(dlsym(h, sym))
        BIND(GetUniqueId, "ncclGetUniqueId");
        BIND(CommInitRank, "ncclCommInitRank");
        BIND(CommDestroy, "ncclCommDestroy");
        BIND(AllGather, "ncclAllGather");
        BIND(Send, "ncclSend");
        BIND(Recv, "ncclRecv");
        BIND(GroupStart, "ncclGroupStart");
        BIND(GroupEnd, "ncclGroupEnd");
        BIND(GetErrorString, "ncclGetErrorString");
#undef BIND
    });
    if (!api.h || !api.GetUniqueId || !api.CommInitRank || !api.AllGather || !api.Send || !api.Recv || !api.GroupStart || !api.GroupEnd)
        throw StatusError(-9 /* DCD_E_UNSUPPORTED */, "libnccl.so.2 could not be loaded: the sharded build needs NCCL");
    return api;
}
void nccl_check(int rc, const char* what) {
    if (rc == 0) return;
    NcclApi& a = nccl();
    throw StatusError(-3 /* DCD_E_CUDA */, std::string(what) + " failed: " + (a.GetErrorString ? a.GetErrorString(rc) : "NCCL error"));
}

struct NcclComm : ShardComm {
    ncclComm_t comm = nullptr;
    int device = 0;
    bool p2p_ok = true;
    std::map<std::string, void*> opened;   // IPC handle bytes -> mapped base (peer slabs are recycled across builds)
    void* d_scratch = nullptr;              // [nranks + 1] x 128 bytes
    ~NcclComm() override {
        for (auto& kv : opened) cudaIpcCloseMemHandle(kv.second);
        if (d_scratch) cudaFree(d_scratch);
        if (comm && nccl().CommDestroy) nccl().CommDestroy(comm);
    }
    struct PtrMsg { cudaIpcMemHandle_t h; unsigned long long offset; int ok; int pad; char fill[128 - sizeof(cudaIpcMemHandle_t) - 16]; };
    bool exchange_ptrs(void* mine, void* alloc_base, void** all, cudaStream_t s) override {
        static_assert(sizeof(PtrMsg) == 128, "PtrMsg layout");
        if (!d_scratch) DCD_CUDA(cudaMalloc(&d_scratch, (size_t)(nranks + 1) * 128));
        PtrMsg me{};
        me.ok = 0;
        if (p2p_ok && mine && alloc_base) {
            if (cudaIpcGetMemHandle(&me.h, alloc_base) == cudaSuccess) { me.ok = 1; me.offset = (unsigned long long)((char*)mine - (char*)alloc_base); }
            else cudaGetLastError();
        } else if (p2p_ok && !mine) {
            me.ok = 2;   // nothing to offer (empty owner range): peers never store through it
        }
        char* d = (char*)d_scratch;
        DCD_CUDA(cudaMemcpyAsync(d, &me, 128, cudaMemcpyHostToDevice, s));
        allgather(d, d + 128, 128, s);
        std::vector<PtrMsg> msgs(nranks);
        DCD_CUDA(cudaMemcpyAsync(msgs.data(), d + 128, (size_t)nranks * 128, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        bool ok = true;
        for (int r = 0; r < nranks; r++) ok = ok && msgs[r].ok != 0;
        if (ok) {
            for (int r = 0; r < nranks && ok; r++) {
                if (r == rank) { all[r] = mine; continue; }
                if (msgs[r].ok == 2) { all[r] = nullptr; continue; }
                const std::string key((const char*)&msgs[r].h, sizeof(cudaIpcMemHandle_t));
                auto it = opened.find(key);
                void* base = nullptr;
                if (it != opened.end()) base = it->second;
                else {
                    if (cudaIpcOpenMemHandle(&base, msgs[r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
                    opened[key] = base;
                }
                all[r] = (char*)base + msgs[r].offset;
            }
        }
        // the decision must be the same everywhere: agree through one more tiny all-gather
        int flag = ok ? 1 : 0;
        DCD_CUDA(cudaMemcpyAsync(d, &flag, 4, cudaMemcpyHostToDevice, s));
        allgather(d, d + 128, 4, s);
        std::vector<int> flags(nranks);
        DCD_CUDA(cudaMemcpyAsync(flags.data(), d + 128, (size_t)nranks * 4, cudaMemcpyDeviceToHost, s));
        DCD_CUDA(cudaStreamSynchronize(s));
        for (int r = 0; r < nranks; r++) ok = ok && flags[r] == 1;
        if (!ok) p2p_ok = false;
        return ok;
    }
    void barrier(cudaStream_t s) override {
        if (!d_scratch) DCD_CUDA(cudaMalloc(&d_scratch, (size_t)(nranks + 1) * 128));
        char* d = (char*)d_scratch;
        allgather(d, d + 128, 4, s);   // ordered after this rank's work on s; completes only when every rank has reached it
        DCD_CUDA(cudaStreamSynchronize(s));
    }
    const char* transport() const override { return "nccl"; }
    void allgather(const void* send, void* recv, size_t bytes, cudaStream_t s) override {
        if (!bytes) return;
        nccl_check(nccl().AllGather(send, recv, bytes, NCCL_UINT8, comm, s), "ncclAllGather");
        bytes_moved += bytes * (size_t)(nranks - 1);
    }
    void allgatherv(const void* send, void* recv, const size_t* counts, const size_t* offs, cudaStream_t s) override {
        NcclApi& a = nccl();
        if (counts[rank]) DCD_CUDA(cudaMemcpyAsync((char*)recv + offs[rank], send, counts[rank], cudaMemcpyDeviceToDevice, s));
        nccl_check(a.GroupStart(), "ncclGroupStart");
        for (int p = 0; p < nranks; p++) {
            if (p == rank) continue;
            if (counts[rank]) nccl_check(a.Send(send, counts[rank], NCCL_UINT8, p, comm, s), "ncclSend");
            if (counts[p]) nccl_check(a.Recv((char*)recv + offs[p], counts[p], NCCL_UINT8, p, comm, s), "ncclRecv");
            bytes_moved += counts[p];
        }
        nccl_check(a.GroupEnd(), "ncclGroupEnd");
    }
    void alltoallv(const void* send, const size_t* send_off, const size_t* send_n, void* recv, const size_t* recv_off,
                   const size_t* recv_n, cudaStream_t s) override {
        NcclApi& a = nccl();
        if (send_n[rank])
            DCD_CUDA(cudaMemcpyAsync((char*)recv + recv_off[rank], (const char*)send + send_off[rank], send_n[rank], cudaMemcpyDeviceToDevice, s));
        nccl_check(a.GroupStart(), "ncclGroupStart");
        for (int p = 0; p < nranks; p++) {
            if (p == rank) continue;
            if (send_n[p]) nccl_check(a.Send((const char*)send + send_off[p], send_n[p], NCCL_UINT8, p, comm, s), "ncclSend");
            if (recv_n[p]) nccl_check(a.Recv((char*)recv + recv_off[p], recv_n[p], NCCL_UINT8, p, comm, s), "ncclRecv");
            bytes_moved += recv_n[p];
        }
        nccl_check(a.GroupEnd(), "ncclGroupEnd");
    }
};

// =====================================================================================================
// in-process loopback group: members post their send pointers, meet at a barrier, and copy from each other
// =====================================================================================================
struct LocalShared {
    int n = 0;
    std::mutex m;
    std::condition_variable cv;
    int waiting = 0;
    unsigned long long generation = 0;
    bool poisoned = false;
    std::vector<const void*> send;
    std::vector<const size_t*> send_off, send_n;
    void barrier() {
        std::unique_lock<std::mutex> lk(m);
        const unsigned long long g = generation;
        if (poisoned) throw StatusError(-6, "a peer of the shard group failed");
        if (++waiting == n) {
            waiting = 0;
            generation++;
            cv.notify_all();
        } else {
            cv.wait(lk, [&] { return generation != g || poisoned; });
            if (generation == g) throw StatusError(-6, "a peer of the shard group failed");
        }
    }
    void poison() {
        std::lock_guard<std::mutex> lk(m);
        poisoned = true;
        cv.notify_all();
    }
};
struct LocalComm : ShardComm {
    std::shared_ptr<LocalShared> sh;
    const char* transport() const override { return "local"; }
    void abort() override { sh->poison(); }
    void post(const void* send, const size_t* so, const size_t* sn, cudaStream_t s) {
        DCD_CUDA(cudaStreamSynchronize(s));   // my send data is complete before peers read it
        sh->send[rank] = send;
        sh->send_off[rank] = so;
        sh->send_n[rank] = sn;
        sh->barrier();
    }
    void done(cudaStream_t s) {
        DCD_CUDA(cudaStreamSynchronize(s));
        sh->barrier();                         // nobody re-uses a send buffer before every copy has finished
    }
    bool exchange_ptrs(void* mine, void*, void** all, cudaStream_t s) override {
        post(mine, nullptr, nullptr, s);
        for (int r = 0; r < nranks; r++) all[r] = const_cast<void*>(sh->send[r]);
        done(s);
        return true;
    }
    void barrier(cudaStream_t s) override {
        DCD_CUDA(cudaStreamSynchronize(s));
        sh->barrier();
    }
    void allgather(const void* send, void* recv, size_t bytes, cudaStream_t s) override {
        post(send, nullptr, nullptr, s);
        for (int r = 0; r < nranks; r++)
            if (bytes) DCD_CUDA(cudaMemcpyAsync((char*)recv + (size_t)r * bytes, sh->send[r], bytes, cudaMemcpyDeviceToDevice, s));
        bytes_moved += bytes * (size_t)(nranks - 1);
        done(s);
    }
    void allgatherv(const void* send, void* recv, const size_t* counts, const size_t* offs, cudaStream_t s) override {
        post(send, nullptr, nullptr, s);
        for (int r = 0; r < nranks; r++) {
            if (counts[r]) DCD_CUDA(cudaMemcpyAsync((char*)recv + offs[r], sh->send[r], counts[r], cudaMemcpyDeviceToDevice, s));
            if (r != rank) bytes_moved += counts[r];
        }
        done(s);
    }
    void alltoallv(const void* send, const size_t* send_off, const size_t* send_n, void* recv, const size_t* recv_off,
                   const size_t* recv_n, cudaStream_t s) override {
        post(send, send_off, send_n, s);
        for (int r = 0; r < nranks; r++) {
            const size_t n = sh->send_n[r][rank];
            if (n != recv_n[r]) throw StatusError(-1, "alltoallv: send / receive sizes disagree");
            if (n) DCD_CUDA(cudaMemcpyAsync((char*)recv + recv_off[r], (const char*)sh->send[r] + sh->send_off[r][rank], n, cudaMemcpyDeviceToDevice, s));
            if (r != rank) bytes_moved += n;
        }
        done(s);
    }
};
}  // namespace

void nccl_unique_id(unsigned char* out128) {
    ncclUniqueIdT id;
    nccl_check(nccl().GetUniqueId(&id), "ncclGetUniqueId");
    memcpy(out128, id.internal, 128);
}
ShardComm* make_nccl_comm(int device, const unsigned char* id128, int rank, int nranks) {
    if (nranks < 1 || rank < 0 || rank >= nranks) throw StatusError(-1, "bad rank / nranks");
    DCD_CUDA(cudaSetDevice(device));
    ncclUniqueIdT id;
    memcpy(id.internal, id128, 128);
    std::unique_ptr<NcclComm> c(new NcclComm);
    c->rank = rank;
    c->nranks = nranks;
    c->device = device;
    nccl_check(nccl().CommInitRank(&c->comm, nranks, id, rank), "ncclCommInitRank");
    return c.release();
}
void make_local_comms(int nranks, ShardComm** out) {
    if (nranks < 1) throw StatusError(-1, "bad nranks");
    auto sh = std::make_shared<LocalShared>();
    sh->n = nranks;
    sh->send.assign(nranks, nullptr);
    sh->send_off.assign(nranks, nullptr);
    sh->send_n.assign(nranks, nullptr);
    for (int r = 0; r < nranks; r++) {
        LocalComm* c = new LocalComm;
        c->rank = r;
        c->nranks = nranks;
        c->sh = sh;
        out[r] = c;
    }
}

// =====================================================================================================
// kernels
// =====================================================================================================
static inline unsigned grid_for(u64 n, int block = 256) { return (unsigned)((n + block - 1) / block); }

// One thread per code: upper bound of the code's segment in the (code, coord)-sorted keys by binary search (the top
// of the search tree stays in L1/L2: ~30 dependent cached loads per code instead of another pass over the keys).
__global__ void k_code_last(const u64* __restrict__ keys, u64 n, u32 ncodes, const u32* __restrict__ local_of_glob, u64* __restrict__ last) {
    const u32 g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ncodes) return;
    u32 c = g;
    if (local_of_glob) {
        c = local_of_glob[g];
        if (c == 0xFFFFFFFFu) { last[g] = COORD_NONE; return; }
    }
    const u64 bound = (u64)(c + 1) << COORD_BITS;   // first key of the next code
    u64 lo = 0, hi = n;                             // first index with keys[i] >= bound
    while (lo < hi) {
        const u64 mid = (lo + hi) >> 1;
        if (keys[mid] < bound) lo = mid + 1; else hi = mid;
    }
    u64 v = COORD_NONE;
    if (lo > 0) {
        const u64 k = keys[lo - 1];
        if ((u32)(k >> COORD_BITS) == c) v = k & COORD_MASK;
    }
    last[g] = v;
}
void launch_code_last(const u64* keys, u64 n, u32 ncodes, const u32* local_of_glob, u64* last, cudaStream_t s) {
    if (ncodes) k_code_last<<<grid_for(ncodes), 256, 0, s>>>(keys, n, ncodes, local_of_glob, last);
}

__global__ void k_key_begins(const u64* __restrict__ keys, u64 n, const u32* __restrict__ splits, int nranks, const u32* __restrict__ glob,
                             u64* __restrict__ key_begin) {
    const int o = threadIdx.x;
    if (o > nranks) return;
    if (o == nranks) { key_begin[o] = n; return; }
    const u32 want = splits[o];
    u64 lo = 0, hi = n;   // first index whose global code id is >= want
    while (lo < hi) {
        const u64 mid = (lo + hi) >> 1;
        const u32 c = (u32)(keys[mid] >> COORD_BITS);
        if ((glob ? glob[c] : c) < want) lo = mid + 1; else hi = mid;
    }
    key_begin[o] = lo;
}
void launch_key_begins(const u64* keys, u64 n, const u32* splits, int nranks, const u32* glob, u64* key_begin, cudaStream_t s) {
    k_key_begins<<<1, 32, 0, s>>>(keys, n, splits, nranks, glob, key_begin);
}

__global__ void k_first_base(const u64* __restrict__ all_last, int my_rank, u32 ncodes, u64* __restrict__ base) {
    const u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncodes) return;
    u64 b = 0;
    for (int r = my_rank - 1; r >= 0; r--) {
        const u64 v = all_last[(size_t)r * ncodes + c];
        if (v != COORD_NONE) { b = v; break; }
    }
    base[c] = b;
}
void launch_first_base(const u64* all_last, int nranks, int my_rank, u32 ncodes, u64* base, cudaStream_t s) {
    (void)nranks;
    if (ncodes) k_first_base<<<grid_for(ncodes), 256, 0, s>>>(all_last, my_rank, ncodes, base);
}

__global__ void k_total_len(const u64* __restrict__ all_off, int nranks, u32 ncodes, u64* __restrict__ tot) {
    const u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncodes) return;
    u64 t = 0;
    for (int r = 0; r < nranks; r++) {
        const u64* o = all_off + (size_t)r * (ncodes + 1);
        t += o[c + 1] - o[c];
    }
    tot[c] = t;
}
void launch_total_len(const u64* all_off, int nranks, u32 ncodes, u64* tot, cudaStream_t s) {
    if (ncodes) k_total_len<<<grid_for(ncodes), 256, 0, s>>>(all_off, nranks, ncodes, tot);
}

// splits[o] = number of codes whose inclusive prefix is <= o * total / nranks  (so every code belongs to exactly one owner)
__global__ void k_find_splits(const u64* __restrict__ tot_incl, u32 ncodes, int nranks, u32* __restrict__ splits) {
    const int o = threadIdx.x;
    if (o > nranks) return;
    if (o == 0) { splits[0] = 0; return; }
    if (o == nranks) { splits[o] = ncodes; return; }
    const u64 total = ncodes ? tot_incl[ncodes - 1] : 0;
    const u64 target = (u64)(((unsigned __int128)total * (unsigned)o) / (unsigned)nranks);
    u32 lo = 0, hi = ncodes;   // first c with tot_incl[c] > target
    while (lo < hi) {
        const u32 mid = (lo + hi) >> 1;
        if (tot_incl[mid] <= target) lo = mid + 1; else hi = mid;
    }
    splits[o] = lo;
}
void launch_find_splits(const u64* tot_incl, u32 ncodes, int nranks, u32* splits, cudaStream_t s) {
    k_find_splits<<<1, 64, 0, s>>>(tot_incl, ncodes, nranks, splits);
}

__global__ void k_gather_cuts(const u64* __restrict__ all_off, u32 ncodes, int nranks, const u32* __restrict__ splits, u64* __restrict__ cut) {
    const int i = threadIdx.x;
    if (i >= nranks * (nranks + 1)) return;
    const int r = i / (nranks + 1), o = i % (nranks + 1);
    cut[i] = all_off[(size_t)r * (ncodes + 1) + splits[o]];
}
void launch_gather_cuts(const u64* all_off, u32 ncodes, int nranks, const u32* splits, u64* cut, cudaStream_t s) {
    k_gather_cuts<<<1, 1024, 0, s>>>(all_off, ncodes, nranks, splits, cut);
}

__global__ void k_merged_off(const u64* __restrict__ tot_incl, u32 c_begin, u32 n_own, u64* __restrict__ merged_off) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n_own) return;
    const u64 base = c_begin ? tot_incl[c_begin - 1] : 0;
    const u32 c = c_begin + i;
    merged_off[i] = (c ? tot_incl[c - 1] : 0) - base;
}
void launch_merged_off(const u64* tot_incl, u32 c_begin, u32 n_own, u64* merged_off, cudaStream_t s) {
    k_merged_off<<<grid_for((u64)n_own + 1), 256, 0, s>>>(tot_incl, c_begin, n_own, merged_off);
}

// ---- fused exchange + merge: store every local segment at its final position in the owner's arena (peer memory) ----------
// One CTA per tile of SC_TILE source ushorts.  Lanes take consecutive ushorts, so every warp instruction reads and writes
// one contiguous 64-byte stretch (a segment boundary splits it): local reads stay coalesced and the remote stores leave
// as full sectors.  The (code, destination) of a unit is cached in registers while the lane stays inside the same segment.
constexpr int SC_THREADS = 256;
constexpr int SC_TILE = 4096;
constexpr int SC_SHORT = 15;      // pieces up to this many ushorts are stored by one thread
constexpr int SC_OFFS = 1024;     // segment offsets of the tile kept in shared memory when its code range is this small
// tile_code[t] = code whose segment holds source unit t * SC_TILE (one thread per tile: the ~21 dependent loads of the binary
// search overlap across a quarter of a million threads instead of stalling every scatter CTA for ~10 us at its start)
__global__ void k_tile_codes(const u64* __restrict__ off, u32 ncodes, u32 ntiles, u32* __restrict__ tile_code) {
    const u32 t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    const u64 pos = (u64)t * SC_TILE;
    u32 lo = 0, hi = ncodes;   // first c with off[c] > pos
    while (lo < hi) { const u32 mid = (lo + hi) >> 1; if (off[mid] <= pos) lo = mid + 1; else hi = mid; }
    tile_code[t] = lo - 1;
}
constexpr u32 SC_BIG = 1024;
struct ScatterRanges { u64 u0[16], u1[16]; u32 tile_first[16], ntiles[16]; int n; };
// len ushorts (a piece's unaligned head or tail, or a whole short piece) by ONE thread with the widest stores the destination
// alignment allows: 2, 4, 8 or 16 bytes each instead of one store per ushort -- over NVLink every store is a packet.
DCD_D void store_span(u16* dst, const u16* src, u32 len) {
    while (len) {
        const u32 al = (u32)((size_t)dst >> 1) & 7u;            // position inside the 16-byte line, in ushorts
        u32 w = al ? (al & (0u - al)) : 8u;                     // largest power of two the address is aligned to (<= 8 ushorts)
        while (w > len) w >>= 1;
        if (w == 8) {
            uint4 q;
            q.x = src[0] | ((u32)src[1] << 16); q.y = src[2] | ((u32)src[3] << 16);
            q.z = src[4] | ((u32)src[5] << 16); q.w = src[6] | ((u32)src[7] << 16);
            *reinterpret_cast<uint4*>(dst) = q;
        } else if (w == 4) {
            uint2 q;
            q.x = src[0] | ((u32)src[1] << 16); q.y = src[2] | ((u32)src[3] << 16);
            *reinterpret_cast<uint2*>(dst) = q;
        } else if (w == 2) {
            *reinterpret_cast<u32*>(dst) = src[0] | ((u32)src[1] << 16);
        } else {
            dst[0] = src[0];
        }
        dst += w; src += w; len -= w;
    }
}
// len ushorts from shared memory to dst with destination-aligned 16-byte stores (full sectors for the local L2 and for NVLink);
// t = 0 .. nt - 1 are the cooperating threads (nt >= 32).  The source phase is arbitrary: five aligned 32-bit words, shifted.
DCD_D void copy_piece(u16* dst, const u16* src, u32 len, u32 t, u32 nt) {
    const u32 head = min(len, (u32)(((16u - (u32)((size_t)dst & 15u)) & 15u) >> 1));
    const u32 nvec = (len - head) >> 3;
    const u32 tail0 = head + (nvec << 3);
    const u16* s0 = src + head;
    const u32 sh = ((u32)((size_t)s0 >> 1) & 1u) * 16u;   // 16 when s0 sits in the upper half of a 32-bit word
    const u32* w = reinterpret_cast<const u32*>(reinterpret_cast<size_t>(s0) & ~size_t(3));
    for (u32 v = t; v < nvec; v += nt) {
        const u32* p = w + (v << 2);
        const u32 a0 = p[0], a1 = p[1], a2 = p[2], a3 = p[3], a4 = sh ? p[4] : 0u;
        uint4 q;
        q.x = __funnelshift_r(a0, a1, sh);
        q.y = __funnelshift_r(a1, a2, sh);
        q.z = __funnelshift_r(a2, a3, sh);
        q.w = __funnelshift_r(a3, a4, sh);
        *reinterpret_cast<uint4*>(dst + head + (v << 3)) = q;
    }
    if (t == 0 && head) store_span(dst, src, head);
    if (t == nt - 1u && len > tail0) store_span(dst + tail0, src + tail0, len - tail0);   // the tail goes to the other end of the thread range
}
__global__ void __launch_bounds__(SC_THREADS) k_scatter_segments(const u16* __restrict__ arena, u64 units, const u64* __restrict__ all_off,
                                                                 const u64* __restrict__ tot_incl, u32 ncodes, int nranks, int me,
                                                                 const u32* __restrict__ splits, u16* const* __restrict__ peer,
                                                                 u32 tile_rot, const u32* __restrict__ tile_code, ScatterRanges rg,
                                                                 u32 ntiles_all, u64 ce_min) {
    __shared__ __align__(16) u16 s_tile[SC_TILE];
    __shared__ u64 s_off[SC_OFFS + 1];
    __shared__ u32 s_lo, s_hi, s_nlong;
    __shared__ u32 s_split[18];
    __shared__ u16* s_peer[17];
    __shared__ u32 s_long_src[SC_TILE / (SC_SHORT + 1) + 1], s_long_len[SC_TILE / (SC_SHORT + 1) + 1];
    __shared__ u16* s_long_dst[SC_TILE / (SC_SHORT + 1) + 1];
    const u64* off = all_off + (size_t)me * (ncodes + 1);
    // tiles are visited in an order rotated per rank: the arena is in code order == owner order, so without the rotation
    // every rank would store to owner 0 first, then owner 1, ... (incast on one GPU's NVLink ingress at a time)
    // A launch covers up to 16 source ranges [u0, u1) (one per owner), pieces clipped to their range.  Consecutive CTAs take
    // tiles of DIFFERENT ranges, so at any moment the stores of a rank go to all owners at once: every link of the fabric
    // is busy whatever the skew between the ranks (an owner-by-owner order needs the ranks in lock step to avoid incast).
    const u32 ri = blockIdx.x % (u32)rg.n, ti = blockIdx.x / (u32)rg.n;
    if (ti >= rg.ntiles[ri]) return;
    const u64 u0 = rg.u0[ri], u1 = rg.u1[ri];
    const u32 tile = rg.tile_first[ri] + (ti + tile_rot) % rg.ntiles[ri];
    const u64 t0 = (u64)tile * SC_TILE;
    const u64 t1 = min(units, t0 + SC_TILE);
    const u64 k0 = max(t0, u0), k1 = min(t1, u1);
    if (threadIdx.x == 0) {   // codes that can have units in the tile (the loop below clips every segment to [k0, k1))
        s_lo = tile_code[tile];
        s_hi = tile + 1 < ntiles_all ? tile_code[tile + 1] : ncodes - 1;
    }
    if (threadIdx.x <= (unsigned)nranks) s_split[threadIdx.x] = splits[threadIdx.x];
    if (threadIdx.x < (unsigned)nranks) s_peer[threadIdx.x] = peer[threadIdx.x];
    // the tile itself: 128-bit coalesced loads (the arena is 16-byte aligned and padded past `units`)
    {
        const uint4* src = reinterpret_cast<const uint4*>(arena + t0);
        uint4* dst = reinterpret_cast<uint4*>(s_tile);
        const u32 nvec = (u32)((t1 - t0 + 7) / 8);
        for (u32 i = threadIdx.x; i < nvec; i += SC_THREADS) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const u32 c_lo = s_lo, c_hi = s_hi;
    const bool offs_in_smem = c_hi - c_lo + 2 <= SC_OFFS + 1;
    if (offs_in_smem) {
        for (u32 i = threadIdx.x; i <= c_hi - c_lo + 1; i += SC_THREADS) s_off[i] = off[c_lo + i];
        __syncthreads();
    }
    // Segments of the tile, one per thread at a time.  Short pieces are stored by their thread; long ones go to a list and
    // are copied by the whole CTA with destination-aligned 16-byte stores (full sectors for the local L2 and for NVLink).
    if (threadIdx.x == 0) s_nlong = 0;
    __syncthreads();
    for (u32 c = c_lo + threadIdx.x; c <= c_hi; c += SC_THREADS) {
        const u64 a = offs_in_smem ? s_off[c - c_lo] : off[c];
        const u64 e = offs_in_smem ? s_off[c + 1 - c_lo] : off[c + 1];
        const u64 ca = max(a, k0), ce = min(e, k1);
        if (ca >= ce) continue;          // code absent on this rank, or outside the tile / the launch's range
        if (e - a >= ce_min) continue;   // a big segment: it travels by copy engine (k_big_segments)
        int o = 0;
        while (o + 1 < nranks && s_split[o + 1] <= c) o++;
        const u32 cb = s_split[o];
        u64 d = (c ? tot_incl[c - 1] : 0) - (cb ? tot_incl[cb - 1] : 0);
        for (int r = 0; r < me; r++) {
            const u64* orow = all_off + (size_t)r * (ncodes + 1);
            d += orow[c + 1] - orow[c];
        }
        u16* dst = s_peer[o] + d + (ca - a);
        const u32 src = (u32)(ca - t0), len = (u32)(ce - ca);
        if (len <= SC_SHORT) {
            store_span(dst, s_tile + src, len);
        } else {
            const u32 slot = atomicAdd(&s_nlong, 1u);   // at most SC_TILE / (SC_SHORT + 1) long pieces fit in a tile
            s_long_src[slot] = src;
            s_long_len[slot] = len;
            s_long_dst[slot] = dst;
        }
    }
    __syncthreads();
    // Long pieces: a warp per piece (the cold part of the arena has ~100 pieces per tile: eight go at a time instead of the whole
    // CTA walking them one by one), the whole CTA for the few pieces of SC_BIG ushorts or more.
    const u32 nlong = s_nlong;
    const u32 warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    for (u32 li = warp; li < nlong; li += SC_THREADS / 32) {
        if (s_long_len[li] < SC_BIG) copy_piece(s_long_dst[li], s_tile + s_long_src[li], s_long_len[li], lane, 32u);
    }
    for (u32 li = 0; li < nlong; li++) {
        if (s_long_len[li] >= SC_BIG) copy_piece(s_long_dst[li], s_tile + s_long_src[li], s_long_len[li], threadIdx.x, (u32)SC_THREADS);
    }
}
u64 scatter_scratch_words(u64 my_units) { return (my_units + SC_TILE - 1) / SC_TILE + 1; }
void launch_scatter_ranges(const u16* my_arena, u64 my_units, const u64* all_off, const u64* tot_incl, u32 ncodes, int nranks, int me,
                           const u32* splits, u16* const* peer, const u64* u0, const u64* u1, int n, u32 tile_rot, const u32* scratch,
                           cudaStream_t s, u64 ce_min) {
    if (nranks > 16 || n > 16) throw StatusError(-1, "more than 16 shards");
    ScatterRanges rg{};
    u32 most = 0;
    for (int i = 0; i < n; i++) {
        if (u1[i] <= u0[i]) continue;
        rg.u0[rg.n] = u0[i];
        rg.u1[rg.n] = u1[i];
        rg.tile_first[rg.n] = (u32)(u0[i] / SC_TILE);
        rg.ntiles[rg.n] = (u32)((u1[i] - 1) / SC_TILE) - rg.tile_first[rg.n] + 1;
        most = std::max(most, rg.ntiles[rg.n]);
        rg.n++;
    }
    if (!rg.n) return;
    k_scatter_segments<<<most * (u32)rg.n, SC_THREADS, 0, s>>>(my_arena, my_units, all_off, tot_incl, ncodes, nranks, me, splits, peer,
                                                              tile_rot, scratch, rg, (u32)grid_for(my_units, SC_TILE), ce_min ? ce_min : ~0ull);
}
// Segments of at least ce_min ushorts: {source offset, length, owner, destination offset} for the host, which sends them with
// cudaMemcpyAsync (copy engines: no SM time, full link rate); k_scatter_segments skips them.  list[0] = count, entries follow.
__global__ void k_big_segments(const u64* __restrict__ all_off, const u64* __restrict__ tot_incl, u32 ncodes, int nranks, int me,
                               const u32* __restrict__ splits, u64 ce_min, u64* __restrict__ list, u32 cap) {
    const u32 c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncodes) return;
    const u64* off = all_off + (size_t)me * (ncodes + 1);
    const u64 a = off[c], e = off[c + 1];
    if (e - a < ce_min) return;
    int o = 0;
    while (o + 1 < nranks && splits[o + 1] <= c) o++;
    const u32 cb = splits[o];
    u64 d = (c ? tot_incl[c - 1] : 0) - (cb ? tot_incl[cb - 1] : 0);
    for (int r = 0; r < me; r++) {
        const u64* orow = all_off + (size_t)r * (ncodes + 1);
        d += orow[c + 1] - orow[c];
    }
    const u32 slot = (u32)atomicAdd(reinterpret_cast<unsigned long long*>(list), 1ull);
    if (slot >= cap) return;   // (cannot happen: cap >= arena units / ce_min)
    u64* ent = list + 1 + (size_t)slot * 4;
    ent[0] = a; ent[1] = e - a; ent[2] = (u64)o; ent[3] = d;
}
void launch_big_segments(const u64* all_off, const u64* tot_incl, u32 ncodes, int nranks, int me, const u32* splits, u64 ce_min, u64* list,
                         u32 cap, cudaStream_t s) {
    cudaMemsetAsync(list, 0, 8, s);
    if (ncodes) k_big_segments<<<grid_for(ncodes), 256, 0, s>>>(all_off, tot_incl, ncodes, nranks, me, splits, ce_min, list, cap);
}
void launch_scatter_tile_codes(const u64* my_off, u32 ncodes, u64 my_units, u32* scratch, cudaStream_t s) {
    if (!my_units) return;
    const unsigned grid = grid_for(my_units, SC_TILE);
    k_tile_codes<<<grid_for(grid), 256, 0, s>>>(my_off, ncodes, grid, scratch);
}

// Each thread produces 8 consecutive ushorts of the merged arena (one 16-byte store).  The CTA first narrows the code
// range of its 2048-unit tile, each thread then finds its (code, rank) piece and walks forward across piece ends.
constexpr int CC_THREADS = 256;
constexpr int CC_PER = 8;
__global__ void __launch_bounds__(CC_THREADS) k_concat_segments(const u64* __restrict__ all_off, u32 ncodes, int nranks, u32 c_begin,
                                                                 u32 n_own, const u64* __restrict__ merged_off,
                                                                 const u16* __restrict__ recv, const u64* __restrict__ slab,
                                                                 u16* __restrict__ out, u64 total) {
    __shared__ u32 s_lo, s_hi;
    __shared__ u64 s_slab[16];
    const u64 tile0 = (u64)blockIdx.x * (CC_THREADS * CC_PER);
    if (threadIdx.x < 2) {
        // last code index i with merged_off[i] <= pos
        const u64 pos = threadIdx.x == 0 ? tile0 : min(total - 1, tile0 + CC_THREADS * CC_PER - 1);
        u32 lo = 0, hi = n_own;   // first i with merged_off[i] > pos
        while (lo < hi) {
            const u32 mid = (lo + hi) >> 1;
            if (merged_off[mid] <= pos) lo = mid + 1; else hi = mid;
        }
        if (threadIdx.x == 0) s_lo = lo - 1; else s_hi = lo - 1;
    }
    if (threadIdx.x < nranks) s_slab[threadIdx.x] = slab[threadIdx.x];
    __syncthreads();
    const u64 pos0 = tile0 + (u64)threadIdx.x * CC_PER;
    if (pos0 >= total) return;
    u32 lo = s_lo, hi = s_hi + 1;   // first i in [lo, hi] with merged_off[i] > pos0
    while (lo < hi) {
        const u32 mid = (lo + hi) >> 1;
        if (merged_off[mid] <= pos0) lo = mid + 1; else hi = mid;
    }
    u32 i = lo - 1;
    u64 k = pos0 - merged_off[i];   // offset inside code i's merged list
    int r = 0;
    u64 len = 0;
    const u16* src = nullptr;
    auto open_piece = [&]() {       // piece (i, r): length and source pointer
        const u64* o = all_off + (size_t)r * (ncodes + 1);
        const u64 a = o[c_begin + i];
        len = o[c_begin + i + 1] - a;
        src = recv + s_slab[r] + (a - o[c_begin]);
    };
    open_piece();
    while (k >= len) {              // codes in [c_begin, c_begin + n_own) cover [0, total): terminates
        k -= len;
        if (++r == nranks) { r = 0; i++; }
        open_piece();
    }
    u16 v[CC_PER];
    const int cnt = (int)min((u64)CC_PER, total - pos0);
    for (int j = 0; j < cnt; j++) {
        while (k >= len) {
            k = 0;
            if (++r == nranks) { r = 0; i++; }
            open_piece();
        }
        v[j] = src[k++];
    }
    if (cnt == CC_PER) {
        uint4 q;
        q.x = v[0] | ((u32)v[1] << 16);
        q.y = v[2] | ((u32)v[3] << 16);
        q.z = v[4] | ((u32)v[5] << 16);
        q.w = v[6] | ((u32)v[7] << 16);
        *reinterpret_cast<uint4*>(out + pos0) = q;
    } else {
        for (int j = 0; j < cnt; j++) out[pos0 + j] = v[j];
    }
}
void launch_concat_segments(const u64* all_off, u32 ncodes, int nranks, u32 c_begin, u32 n_own, const u64* merged_off,
                            const u16* recv, const u64* slab, u16* out, u64 total_units, cudaStream_t s) {
    if (!total_units) return;
    if (nranks > 16) throw StatusError(-1, "more than 16 shards");
    k_concat_segments<<<grid_for(total_units, CC_THREADS * CC_PER), CC_THREADS, 0, s>>>(all_off, ncodes, nranks, c_begin, n_own,
                                                                                        merged_off, recv, slab, out, total_units);
}

}  // namespace dcd
