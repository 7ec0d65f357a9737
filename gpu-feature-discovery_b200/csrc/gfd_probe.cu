// Opt-in capability probes for libgfd_b200.so: hand-written sm_100a streaming kernels.
//
// NEW capability -- the reference (GFD v0.7.0) launches no GPU work; nothing here replaces reference code
// and nothing here runs unless a probe flag / entry point asks for it (BASELINE.json north_star).
//
//   hbm_read_fold   stream N bytes from HBM with coalesced 128/256-bit loads, fold every 64-bit word into
//                   (xor, sum) so the reads cannot be elided and the result is checkable against the CPU
//                   oracle bit for bit.  Algorithmic traffic: N bytes read.       Roofline: HBM3e.
//   hbm_copy        dst[i] = src[i] with the same access shape.  Traffic 2N.     Roofline: HBM3e (or
//                   NVLink 5 when one side is a peer mapping -- that is the P2P probe's kernel).
//   fill_pattern    word[i] = 0x9E3779B97F4A7C15 * (i + seed): deterministic, seedless probe input.
//
// Shape (constants chosen from the on-B200 sweep in profiles/r1_hbm_variant_sweep*.txt):
//   read: persistent grid = SMs x kBlocksPerSM CTAs of kThreads threads that PULL 64 KiB tiles from a global
//         counter (a static tile->CTA stride leaves the slower SMs as a tail: 7.2 TB/s; pulling: 7.46 TB/s);
//         inside a tile every thread issues kUnroll independent 16-byte loads before consuming any
//         (128 KiB in flight per SM, >> HBM latency x bandwidth / 148).  Loads bypass L1 (.nc.L1::no_allocate)
//         and are evict-first in L2: each byte is touched once.  256-bit loads measured identical; not used.
//   copy: one 32 KiB tile per CTA, 1024 threads x 2 vectors, as many CTAs as tiles: the hardware CTA scheduler
//         does the balancing and short CTAs interleave reads and writes finely (7.06 TB/s r+w vs 5.9 persistent).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "gfd_internal.h"

namespace {

constexpr int kThreads = 512;
constexpr int kUnroll = 8;             // 16-byte loads in flight per thread
constexpr int kBlocksPerSM = 2;
constexpr int kMaxBlocks = 4095;       // scratch = one 16-byte header (tile counter) + this many 16-byte partials
constexpr int kCopyThreads = 1024;
constexpr int kCopyUnroll = 2;         // 32 KiB per CTA
constexpr uint64_t kGolden = 0x9E3779B97F4A7C15ull;
constexpr int kMaxDevices = 64;

// A failed runtime call also parks its code in the runtime's "last error" slot; take it out again so that the next
// launch's cudaGetLastError() check (e.g. the retry after an out-of-memory probe size) does not trip over it.
#define CUDA_TRY(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) { (void)cudaGetLastError(); return GFD_ERR_CUDA + (int)e__; } } while (0)

struct U2 { uint64_t a, b; };

// 128-bit streaming load: no L1 allocation, evict-first in L2 through a cache-policy operand (the bare
// .L2::evict_first qualifier exists only on the 256-bit ld.global.v4.u64 form).
__device__ __forceinline__ uint64_t evict_first_policy() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// kLocal = true: the source is this GPU's HBM -- no L1 allocation, evict-first in L2 (each byte is touched once).
// kLocal = false: the source is a peer mapping (NVLink).  There the plain form is the right one: with L1::no_allocate
// a 128-bit load costs link bandwidth as soon as the link carries traffic in both directions -- 643 vs 671 GB/s per
// direction, same 780 one way (profiles/r1c_p2p_sweep_2xB200.txt isolates each qualifier).
template <bool kLocal>
__device__ __forceinline__ U2 ld_stream16(const void *p, uint64_t pol) {
    U2 r;
    if (kLocal)
        asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u64 {%0,%1}, [%2], %3;"
                     : "=l"(r.a), "=l"(r.b) : "l"(p), "l"(pol));
    else
        asm volatile("ld.global.nc.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream16(void *p, U2 v) {
    asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(v.a), "l"(v.b) : "memory");
}

struct Fold {
    uint64_t x = 0, s = 0;
    __device__ __forceinline__ void add(uint64_t w) { x ^= w; s += w; }
    __device__ __forceinline__ void add(U2 v) { add(v.a); add(v.b); }
};

// Scratch header (16 bytes, in front of the per-CTA partials).  BOTH words are zero whenever no launch is in flight: the
// scratch is zeroed once when it is created (gfd_hbm_scratch_init; the no-alloc array by its static initialiser) and the
// last CTA of every launch puts the zeros back, so a checksum is ONE kernel launch -- no memset, no finishing kernel.
struct ScratchHeader {
    unsigned long long tile_counter;  // next 64 KiB tile to hand out (hbm_read_fold_kernel)
    unsigned int done;                // CTAs that have published their partial
    unsigned int pad;
};
static_assert(sizeof(ScratchHeader) == 16, "header is one 16-byte slot");

// CTA-wide fold of `f`; the result is valid in thread 0 only.
__device__ __forceinline__ Fold block_fold(Fold f) {
    for (int o = 16; o > 0; o >>= 1) {
        f.x ^= __shfl_xor_sync(0xffffffffu, f.x, o);
        f.s += __shfl_xor_sync(0xffffffffu, f.s, o);
    }
    __shared__ uint64_t sx[kThreads / 32], ss[kThreads / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();  // the arrays may still be read by a previous block_fold of this CTA
    if (lane == 0) { sx[warp] = f.x; ss[warp] = f.s; }
    __syncthreads();
    Fold r;
    if (warp == 0) {
        r.x = lane < kThreads / 32 ? sx[lane] : 0;
        r.s = lane < kThreads / 32 ? ss[lane] : 0;
        for (int o = 16; o > 0; o >>= 1) {
            r.x ^= __shfl_xor_sync(0xffffffffu, r.x, o);
            r.s += __shfl_xor_sync(0xffffffffu, r.s, o);
        }
    }
    return r;
}

// The < 16 tail bytes as (up to) two zero-padded little-endian 64-bit words.
__device__ __forceinline__ Fold fold_tail(const unsigned char *__restrict__ tail, int tail_bytes) {
    Fold f;
    if (tail_bytes > 0) {
        uint64_t w0 = 0, w1 = 0;
        for (int b = 0; b < tail_bytes; b++) {
            const uint64_t v = (uint64_t)tail[b] << (8 * (b & 7));
            if (b < 8) w0 |= v; else w1 |= v;
        }
        f.add(w0);
        if (tail_bytes > 8) f.add(w1);
    }
    return f;
}

// Publish this CTA's partial; the CTA that publishes LAST folds all of them, thread 0's `extra` and the `tail_bytes` (< 16)
// bytes at `tail` into *out and re-arms the header for the next launch.  Classic last-block reduction: partial store ->
// __threadfence -> ticket.
__device__ __forceinline__ void publish_and_finish(Fold f, ScratchHeader *hdr, ulonglong2 *partials, Fold extra,
                                                   const unsigned char *__restrict__ tail, int tail_bytes,
                                                   gfd_checksum *__restrict__ out) {
    Fold mine = block_fold(f);
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = make_ulonglong2(mine.x, mine.s);
        __threadfence();                                   // the partial is visible device-wide before the ticket is
        const unsigned ticket = atomicAdd(&hdr->done, 1u);
        is_last = (ticket == gridDim.x - 1);
        if (is_last) __threadfence();                      // and every other CTA's partial is visible to us after it
    }
    __syncthreads();
    if (!is_last) return;
    Fold g;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
        const ulonglong2 p = __ldcg(&partials[i]);         // L2 is the point of coherence; never a stale L1 line
        g.x ^= p.x;
        g.s += p.y;
    }
    g = block_fold(g);
    if (threadIdx.x == 0) {
        const Fold t = fold_tail(tail, tail_bytes);
        out->xor64 = g.x ^ extra.x ^ t.x;
        out->sum64 = g.s + extra.s + t.s;
        hdr->tile_counter = 0ull;                          // every CTA has drawn its last tile before taking a ticket
        hdr->done = 0u;
    }
}

// n16 = number of whole 16-byte vectors; `tail_bytes` (< 16) follow them.  hdr must be all-zero at launch (it is, unless a
// launch sharing the scratch is still running) and is all-zero again when the kernel ends.
template <bool kLocal>
__global__ void __launch_bounds__(kThreads, kBlocksPerSM)
hbm_read_fold_kernel(const char *__restrict__ src, size_t n16, int tail_bytes, ScratchHeader *__restrict__ hdr,
                     ulonglong2 *__restrict__ partials, gfd_checksum *__restrict__ out) {
    constexpr size_t kTile = (size_t)kThreads * kUnroll;  // 16-byte vectors per tile (64 KiB)
    const size_t full_tiles = n16 / kTile;
    Fold f;
    const uint64_t pol = evict_first_policy();
    // Dynamic tile scheduler: thread 0 draws the NEXT tile index while the CTA streams the current one, so the
    // atomic's round trip hides behind 64 KiB of loads; the two-slot mailbox costs one barrier per tile.
    __shared__ unsigned long long next_tile[2];
    if (threadIdx.x == 0) next_tile[0] = atomicAdd(&hdr->tile_counter, 1ull);
    __syncthreads();
    for (int slot = 0;; slot ^= 1) {
        const unsigned long long t = next_tile[slot];
        if (t >= full_tiles) break;
        if (threadIdx.x == 0) next_tile[slot ^ 1] = atomicAdd(&hdr->tile_counter, 1ull);
        const char *base = src + (size_t)t * kTile * 16;
        U2 v[kUnroll];
#pragma unroll
        for (int j = 0; j < kUnroll; j++) v[j] = ld_stream16<kLocal>(base + ((size_t)j * kThreads + threadIdx.x) * 16, pol);
#pragma unroll
        for (int j = 0; j < kUnroll; j++) f.add(v[j]);
        __syncthreads();
    }
    // ragged remainder (< one tile), bounds-checked, spread over the whole grid
    for (size_t i = full_tiles * kTile + (size_t)blockIdx.x * kThreads + threadIdx.x; i < n16;
         i += (size_t)gridDim.x * kThreads)
        f.add(ld_stream16<kLocal>(src + i * 16, pol));
    publish_and_finish(f, hdr, partials, Fold(), reinterpret_cast<const unsigned char *>(src) + n16 * 16, tail_bytes, out);
}

// One 32 KiB tile per CTA; the last CTA's tile may be partial; CTA 0 also moves the < 16 tail bytes.
template <bool kLocal>
__global__ void __launch_bounds__(kCopyThreads, 2)
hbm_copy_kernel(char *__restrict__ dst, const char *__restrict__ src, size_t n16, int tail_bytes) {
    constexpr size_t kTile = (size_t)kCopyThreads * kCopyUnroll;
    const uint64_t pol = evict_first_policy();
    const size_t first = (size_t)blockIdx.x * kTile;
    if (first + kTile <= n16) {
        U2 v[kCopyUnroll];
#pragma unroll
        for (int j = 0; j < kCopyUnroll; j++) v[j] = ld_stream16<kLocal>(src + (first + (size_t)j * kCopyThreads + threadIdx.x) * 16, pol);
#pragma unroll
        for (int j = 0; j < kCopyUnroll; j++) st_stream16(dst + (first + (size_t)j * kCopyThreads + threadIdx.x) * 16, v[j]);
    } else {
        for (size_t i = first + threadIdx.x; i < n16; i += kCopyThreads) st_stream16(dst + i * 16, ld_stream16<kLocal>(src + i * 16, pol));
    }
    if (blockIdx.x == 0 && threadIdx.x < tail_bytes) dst[n16 * 16 + threadIdx.x] = src[n16 * 16 + threadIdx.x];
}

// One 32 KiB tile per CTA like the copy (the write probe times this kernel); the last tile may be partial.
__global__ void __launch_bounds__(kCopyThreads, 2)
fill_pattern_kernel(char *__restrict__ dst, size_t n16, int tail_bytes, uint64_t seed) {
    constexpr size_t kTile = (size_t)kCopyThreads * kCopyUnroll;
    const size_t first = (size_t)blockIdx.x * kTile;
    auto put = [&](size_t i) {
        U2 v;
        v.a = kGolden * (2 * i + seed);
        v.b = kGolden * (2 * i + 1 + seed);
        st_stream16(dst + i * 16, v);
    };
    if (first + kTile <= n16) {
#pragma unroll
        for (int j = 0; j < kCopyUnroll; j++) put(first + (size_t)j * kCopyThreads + threadIdx.x);
    } else {
        for (size_t i = first + threadIdx.x; i < n16; i += kCopyThreads) put(i);
    }
    if (blockIdx.x == 0 && threadIdx.x < tail_bytes) {
        uint64_t w = kGolden * (2 * n16 + (threadIdx.x >> 3) + seed);
        dst[n16 * 16 + threadIdx.x] = (char)(w >> (8 * (threadIdx.x & 7)));
    }
}

// What a buffer filled by fill_pattern_kernel MUST fold to, computed from the definition alone -- no memory is read.
// Comparing it with what hbm_read_fold_kernel found in HBM (or behind NVLink) turns a bandwidth probe into a
// write -> read verification of every byte it touched.
__global__ void __launch_bounds__(kThreads, kBlocksPerSM)
pattern_fold_kernel(size_t n16, int tail_bytes, uint64_t seed, ScratchHeader *__restrict__ hdr, ulonglong2 *__restrict__ partials,
                    gfd_checksum *__restrict__ out) {
    Fold f;
    for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < n16; i += (size_t)gridDim.x * kThreads) {
        f.add(kGolden * (2 * i + seed));
        f.add(kGolden * (2 * i + 1 + seed));
    }
    Fold tail;
    if (threadIdx.x == 0 && tail_bytes > 0) {  // the ragged tail keeps the LOW bytes of its words
        const int first = tail_bytes < 8 ? tail_bytes : 8, second = tail_bytes - first;
        const uint64_t w0 = kGolden * (2 * n16 + seed), w1 = kGolden * (2 * n16 + 1 + seed);
        tail.add(first == 8 ? w0 : (w0 & ((1ull << (8 * first)) - 1)));
        if (second > 0) tail.add(w1 & ((1ull << (8 * second)) - 1));
    }
    publish_and_finish(f, hdr, partials, tail, nullptr, 0, out);
}

// ---- host side --------------------------------------------------------------------------------------
struct DevInfo { int sms = 0; bool ok = false; };
DevInfo g_dev[kMaxDevices];
std::mutex g_mu;

int device_info(int dev, DevInfo *out) {
    if (dev < 0 || dev >= kMaxDevices) return GFD_ERR_BAD_ARGUMENT;
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_dev[dev].ok) {
        int sms = 0;
        CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        g_dev[dev].sms = sms;
        g_dev[dev].ok = true;
    }
    *out = g_dev[dev];
    return 0;
}

int device_count_checked(int *n) {
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess || c <= 0) {
        (void)cudaGetLastError();
        return GFD_ERR_NO_GPU;  // no CPU fallback for GPU work
    }
    *n = c;
    return 0;
}

int device_of_pointer(const void *p, int *dev) {
    cudaPointerAttributes a;
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess) { (void)cudaGetLastError(); return GFD_ERR_CUDA + (int)e; }
    if (a.type != cudaMemoryTypeDevice && a.type != cudaMemoryTypeManaged) return GFD_ERR_BAD_ARGUMENT;
    *dev = a.device;
    return 0;
}

// `dev` (current) is about to dereference memory of `other`: make sure the peer mapping exists.  Enabling it twice is fine.
int ensure_peer_mapping(int dev, int other) {
    if (dev == other) return 0;
    int can = 0;
    CUDA_TRY(cudaDeviceCanAccessPeer(&can, dev, other));
    if (!can) return GFD_ERR_PEER_ACCESS;
    cudaError_t e = cudaDeviceEnablePeerAccess(other, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { (void)cudaGetLastError(); return 0; }
    if (e != cudaSuccess) { (void)cudaGetLastError(); return GFD_ERR_CUDA + (int)e; }
    return 0;
}

struct DeviceScope {  // this library's static CUDA runtime keeps its own current device; set + restore it
    int prev = -1;
    int rc = 0;
    explicit DeviceScope(int dev) {
        cudaError_t e = cudaGetDevice(&prev);
        if (e != cudaSuccess) { prev = -1; (void)cudaGetLastError(); }
        e = cudaSetDevice(dev);
        if (e != cudaSuccess) rc = GFD_ERR_CUDA + (int)e;
    }
    ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
};

int grid_for(int sms, size_t n16) {
    constexpr size_t kTile = (size_t)kThreads * kUnroll;
    size_t want = (n16 + kTile - 1) / kTile;
    size_t cap = (size_t)sms * kBlocksPerSM;
    if (cap > (size_t)kMaxBlocks) cap = kMaxBlocks;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

// Scratch layout: [0,16) ScratchHeader, then one ulonglong2 partial per CTA.  One launch, nothing else on the stream.
// src_is_peer: d_buf is another GPU's memory reached through a peer mapping (the NVLink probe)
int launch_read(int dev, const void *d_buf, size_t bytes, gfd_checksum *d_out, void *d_scratch, cudaStream_t st,
                int *grid_out, bool src_is_peer = false) {
    DevInfo di;
    int rc = device_info(dev, &di);
    if (rc) return rc;
    size_t n16 = bytes / 16;
    int tail = (int)(bytes % 16);
    int grid = grid_for(di.sms, n16);
    auto *hdr = static_cast<ScratchHeader *>(d_scratch);
    auto *partials = reinterpret_cast<ulonglong2 *>(static_cast<char *>(d_scratch) + sizeof(ScratchHeader));
    if (src_is_peer) hbm_read_fold_kernel<false><<<grid, kThreads, 0, st>>>(static_cast<const char *>(d_buf), n16, tail, hdr, partials, d_out);
    else hbm_read_fold_kernel<true><<<grid, kThreads, 0, st>>>(static_cast<const char *>(d_buf), n16, tail, hdr, partials, d_out);
    CUDA_TRY(cudaGetLastError());
    if (grid_out) *grid_out = grid;
    return 0;
}

int launch_copy(int dev, void *d_dst, const void *d_src, size_t bytes, cudaStream_t st, int *grid_out, bool src_is_peer = false) {
    (void)dev;
    constexpr size_t kTile = (size_t)kCopyThreads * kCopyUnroll;
    size_t n16 = bytes / 16;
    size_t tiles = (n16 + kTile - 1) / kTile;
    if (tiles < 1) tiles = 1;
    if (tiles > 0x7fffffffull) return GFD_ERR_SIZE_LIMIT;  // 64 TiB
    if (src_is_peer) hbm_copy_kernel<false><<<(unsigned)tiles, kCopyThreads, 0, st>>>(static_cast<char *>(d_dst), static_cast<const char *>(d_src), n16, (int)(bytes % 16));
    else hbm_copy_kernel<true><<<(unsigned)tiles, kCopyThreads, 0, st>>>(static_cast<char *>(d_dst), static_cast<const char *>(d_src), n16, (int)(bytes % 16));
    CUDA_TRY(cudaGetLastError());
    if (grid_out) *grid_out = (int)tiles;
    return 0;
}

int launch_fill(int dev, void *d_buf, size_t bytes, uint64_t seed, cudaStream_t st, int *grid_out = nullptr) {
    (void)dev;
    constexpr size_t kTile = (size_t)kCopyThreads * kCopyUnroll;
    size_t n16 = bytes / 16;
    size_t tiles = (n16 + kTile - 1) / kTile;
    if (tiles < 1) tiles = 1;
    if (tiles > 0x7fffffffull) return GFD_ERR_SIZE_LIMIT;  // 64 TiB
    fill_pattern_kernel<<<(unsigned)tiles, kCopyThreads, 0, st>>>(static_cast<char *>(d_buf), n16, (int)(bytes % 16), seed);
    CUDA_TRY(cudaGetLastError());
    if (grid_out) *grid_out = (int)tiles;
    return 0;
}

// expected checksum of a pattern-filled buffer; partials go to d_scratch + 16 like launch_read's
int launch_pattern(int dev, size_t bytes, uint64_t seed, gfd_checksum *d_out, void *d_scratch, cudaStream_t st) {
    DevInfo di;
    int rc = device_info(dev, &di);
    if (rc) return rc;
    size_t n16 = bytes / 16;
    size_t want = (n16 + kThreads - 1) / kThreads;
    size_t cap = (size_t)di.sms * kBlocksPerSM;
    if (cap > (size_t)kMaxBlocks) cap = kMaxBlocks;
    int grid = (int)(want < 1 ? 1 : (want < cap ? want : cap));
    auto *hdr = static_cast<ScratchHeader *>(d_scratch);
    auto *partials = reinterpret_cast<ulonglong2 *>(static_cast<char *>(d_scratch) + sizeof(ScratchHeader));
    pattern_fold_kernel<<<grid, kThreads, 0, st>>>(n16, (int)(bytes % 16), seed, hdr, partials, d_out);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// a scratch this library allocates itself: zero the header once, right after cudaMalloc
int alloc_scratch(void **out) {
    CUDA_TRY(cudaMalloc(out, sizeof(ScratchHeader) + sizeof(ulonglong2) * (size_t)kMaxBlocks));
    CUDA_TRY(cudaMemset(*out, 0, sizeof(ScratchHeader)));
    CUDA_TRY(cudaDeviceSynchronize());  // the users are non-blocking streams, which do not order after the memset's stream
    return 0;
}

}  // namespace

GFD_API size_t gfd_hbm_scratch_bytes(void) { return sizeof(ScratchHeader) + sizeof(ulonglong2) * (size_t)kMaxBlocks; }

GFD_API int gfd_hbm_scratch_init(void *d_scratch, size_t scratch_bytes, void *stream) {
    int n, dev;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (!d_scratch) return GFD_ERR_BAD_ARGUMENT;
    if (scratch_bytes < gfd_hbm_scratch_bytes()) return GFD_ERR_BUFFER_TOO_SMALL;
    if ((uintptr_t)d_scratch % 16) return GFD_ERR_ALIGNMENT;
    if ((rc = device_of_pointer(d_scratch, &dev))) return rc;
    DeviceScope scope(dev);
    if (scope.rc) return scope.rc;
    CUDA_TRY(cudaMemsetAsync(d_scratch, 0, sizeof(ScratchHeader), static_cast<cudaStream_t>(stream)));
    return 0;
}

GFD_API int gfd_cuda_device_count(int *count) {
    if (!count) return GFD_ERR_BAD_ARGUMENT;
    return device_count_checked(count);
}

GFD_API int gfd_hbm_fill_pattern(void *d_buf, size_t bytes, uint64_t seed, void *stream) {
    int n, dev;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (bytes == 0) return 0;
    if (!d_buf) return GFD_ERR_BAD_ARGUMENT;
    if ((uintptr_t)d_buf % 16) return GFD_ERR_ALIGNMENT;
    if ((rc = device_of_pointer(d_buf, &dev))) return rc;
    DeviceScope scope(dev);
    if (scope.rc) return scope.rc;
    return launch_fill(dev, d_buf, bytes, seed, static_cast<cudaStream_t>(stream));
}

GFD_API int gfd_hbm_pattern_checksum(size_t bytes, uint64_t seed, gfd_checksum *d_out, void *d_scratch, size_t scratch_bytes,
                                     void *stream) {
    int n, dev;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (!d_out || !d_scratch) return GFD_ERR_BAD_ARGUMENT;
    if (scratch_bytes < gfd_hbm_scratch_bytes()) return GFD_ERR_BUFFER_TOO_SMALL;
    if ((uintptr_t)d_scratch % 16 || (uintptr_t)d_out % 8) return GFD_ERR_ALIGNMENT;
    if ((rc = device_of_pointer(d_out, &dev))) return rc;
    DeviceScope scope(dev);
    if (scope.rc) return scope.rc;
    return launch_pattern(dev, bytes, seed, d_out, d_scratch, static_cast<cudaStream_t>(stream));
}

GFD_API int gfd_hbm_read_checksum(const void *d_buf, size_t bytes, gfd_checksum *d_out, void *d_scratch,
                                  size_t scratch_bytes, void *stream) {
    int n, dev;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (!d_out || !d_scratch || (bytes && !d_buf)) return GFD_ERR_BAD_ARGUMENT;
    if (scratch_bytes < gfd_hbm_scratch_bytes()) return GFD_ERR_BUFFER_TOO_SMALL;
    if ((uintptr_t)d_buf % 16 || (uintptr_t)d_scratch % 16 || (uintptr_t)d_out % 8) return GFD_ERR_ALIGNMENT;
    if ((rc = device_of_pointer(d_out, &dev))) return rc;
    int src_dev = dev;
    if (bytes && (rc = device_of_pointer(d_buf, &src_dev))) return rc;
    DeviceScope scope(dev);
    if (scope.rc) return scope.rc;
    if ((rc = ensure_peer_mapping(dev, src_dev))) return rc;  // a source on another GPU is read over NVLink, not faulted on
    return launch_read(dev, d_buf, bytes, d_out, d_scratch, static_cast<cudaStream_t>(stream), nullptr, src_dev != dev);
}

GFD_API int gfd_hbm_copy(void *d_dst, const void *d_src, size_t bytes, void *stream) {
    int n, dev;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (bytes == 0) return 0;
    if (!d_dst || !d_src) return GFD_ERR_BAD_ARGUMENT;
    if ((uintptr_t)d_dst % 16 || (uintptr_t)d_src % 16) return GFD_ERR_ALIGNMENT;
    if ((rc = device_of_pointer(d_dst, &dev))) return rc;
    int src_dev = dev;
    if ((rc = device_of_pointer(d_src, &src_dev))) return rc;
    DeviceScope scope(dev);
    if (scope.rc) return scope.rc;
    if ((rc = ensure_peer_mapping(dev, src_dev))) return rc;
    return launch_copy(dev, d_dst, d_src, bytes, static_cast<cudaStream_t>(stream), nullptr, src_dev != dev);
}

// ---- allocation-free form of the read checksum -----------------------------------------------------------
//
// Who this is for.  gfd_hbm_read_checksum needs a scratch buffer, so somebody has to own a device allocation.
// Some users of this library cannot make one at the moment they want a checksum:
//   * a health-check or telemetry agent that shares the GPU with a job which has already claimed (nearly) all
//     of HBM -- a cudaMalloc from the agent can fail, or can push the job itself into an out-of-memory error;
//   * code that runs inside somebody else's memory pool discipline (a framework that forbids raw cudaMalloc
//     because it would fragment or bypass its caching allocator).
// For them the per-CTA partial sums live in the fixed array below.  It is part of this library's device image:
// every GPU on which the library's kernels get loaded carries these 16 KiB from then until the process exits,
// whether or not this entry point is ever called.
//
// Limits and rules (also in include/gfd_b200.h):
//   * at most GFD_NOALLOC_MAX_BYTES (16 GiB) per call; larger buffers are refused with GFD_ERR_SIZE_LIMIT --
//     checksum them piecewise (whole 8-byte words per piece) and combine: xor the xor64s, add the sum64s;
//   * the array has room for 1023 CTAs' partials (plus a 16-byte header); a GPU whose launch shape would need
//     more is refused the same way;
//   * the FIRST call on each GPU looks the array's device address up (cudaGetSymbolAddress), which may make the
//     driver load this library's kernels on that GPU.  Make that first call early, before any stream capture;
//   * all calls on one GPU share the one array and it keeps what the last call left there.  The only part that
//     must start at zero is the 16-byte header (the tile counter the CTAs draw work from and the count of CTAs
//     that have finished).  It is zero when the library's device image is loaded (static initialiser), and the
//     kernel itself puts the zeros back as its last act, so every call finds it cleared without any memset on
//     the caller's stream.  The partial-sum slots need no clearing: each launched CTA overwrites its own slot
//     and the last CTA reads exactly the slots of that launch.  But two calls that run on the GPU at the same
//     time (different streams, or different host threads) share the counters and the slots and corrupt each
//     other's result -- and can leave the header non-zero for later calls: keep calls on one stream, or order them.
namespace {
constexpr int kNoAllocMaxBlocks = 1024;
constexpr size_t kNoAllocMaxBytes = 16ull << 30;
static __device__ __align__(256) ulonglong2 g_noalloc_partials[kNoAllocMaxBlocks] = {};  // slot 0 = ScratchHeader, zero at load

// In host code `g_noalloc_partials` is only a handle for the runtime, not an address a kernel can use.
// The real per-device address is resolved once per GPU, with that GPU current, and cached here.
void *g_noalloc_addr[kMaxDevices];
bool g_noalloc_ready[kMaxDevices];
std::mutex g_noalloc_mu;

int noalloc_scratch(int dev, void **out) {
    if (dev < 0 || dev >= kMaxDevices) return GFD_ERR_BAD_ARGUMENT;
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ready[dev]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_partials);  // the caller has made `dev` current
        if (e != cudaSuccess) { (void)cudaGetLastError(); return GFD_ERR_CUDA + (int)e; }
        if (!p) return GFD_ERR_CUDA + (int)cudaErrorInvalidSymbol;
        g_noalloc_addr[dev] = p;
        g_noalloc_ready[dev] = true;
    }
    *out = g_noalloc_addr[dev];
    return 0;
}
}  // namespace

GFD_API int gfd_hbm_read_checksum_noalloc(const void *d_buf, size_t bytes, gfd_checksum *d_out, void *stream) {
    int n, dev;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (!d_out || (bytes && !d_buf)) return GFD_ERR_BAD_ARGUMENT;
    if (bytes > kNoAllocMaxBytes) return GFD_ERR_SIZE_LIMIT;
    if ((uintptr_t)d_buf % 16 || (uintptr_t)d_out % 8) return GFD_ERR_ALIGNMENT;
    if ((rc = device_of_pointer(d_out, &dev))) return rc;
    int src_dev = dev;
    if (bytes && (rc = device_of_pointer(d_buf, &src_dev))) return rc;  // a host or stale pointer is refused here, not faulted on
    // Reading another GPU's memory needs a peer mapping, and setting one up is exactly the kind of driver-side resource
    // grab this entry point promises not to make (nor can it tell whether the caller already has one): same GPU only.
    if (src_dev != dev) return GFD_ERR_BAD_ARGUMENT;
    DeviceScope scope(dev);
    if (scope.rc) return scope.rc;
    DevInfo di;
    if ((rc = device_info(dev, &di))) return rc;
    if (grid_for(di.sms, bytes / 16) > kNoAllocMaxBlocks - 1) return GFD_ERR_SIZE_LIMIT;  // slot 0 is the header
    void *scratch = nullptr;
    if ((rc = noalloc_scratch(dev, &scratch))) return rc;
    return launch_read(dev, d_buf, bytes, d_out, scratch, static_cast<cudaStream_t>(stream), nullptr);
}

// ---- host-buffer form ---------------------------------------------------------------------------------
namespace {
struct Staging {
    void *d_buf = nullptr;
    size_t cap = 0;
    void *d_scratch = nullptr;
    gfd_checksum *d_out = nullptr;
    cudaStream_t st = nullptr;
};
Staging g_staging[kMaxDevices];
std::mutex g_staging_mu[kMaxDevices];  // one host-path call at a time PER DEVICE (the staging buffer is shared per device)
constexpr size_t kHostPathMax = 16ull << 30;  // 16 GiB per call
}  // namespace

GFD_API int gfd_checksum_host(int cuda_device, const void *host_buf, size_t bytes, gfd_checksum *out) {
    int n;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (!out || (bytes && !host_buf) || cuda_device < 0 || cuda_device >= n || cuda_device >= kMaxDevices) return GFD_ERR_BAD_ARGUMENT;
    if (bytes > kHostPathMax) return GFD_ERR_SIZE_LIMIT;
    DeviceScope scope(cuda_device);
    if (scope.rc) return scope.rc;
    std::lock_guard<std::mutex> lk(g_staging_mu[cuda_device]);
    Staging &s = g_staging[cuda_device];
    // each piece is created on its own so that a call that failed half-way (out of memory) is completed by the next one
    if (!s.st) CUDA_TRY(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
    if (!s.d_scratch && (rc = alloc_scratch(&s.d_scratch))) return rc;
    if (!s.d_out) CUDA_TRY(cudaMalloc((void **)&s.d_out, sizeof(gfd_checksum)));
    size_t need = (bytes + 255) / 256 * 256;
    if (need > s.cap) {
        if (s.d_buf) CUDA_TRY(cudaFree(s.d_buf));
        s.d_buf = nullptr;
        s.cap = 0;
        CUDA_TRY(cudaMalloc(&s.d_buf, need));
        s.cap = need;
    }
    if (bytes) CUDA_TRY(cudaMemcpyAsync(s.d_buf, host_buf, bytes, cudaMemcpyHostToDevice, s.st));
    rc = launch_read(cuda_device, s.d_buf, bytes, s.d_out, s.d_scratch, s.st, nullptr);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(out, s.d_out, sizeof(gfd_checksum), cudaMemcpyDeviceToHost, s.st));
    CUDA_TRY(cudaStreamSynchronize(s.st));
    return 0;
}

// ---- self-contained probes ----------------------------------------------------------------------------
namespace {
struct Buf {
    void *p = nullptr;
    ~Buf() { if (p) cudaFree(p); }
};
struct Ev {
    cudaEvent_t e = nullptr;
    ~Ev() { if (e) cudaEventDestroy(e); }
};
struct Stream {
    cudaStream_t s = nullptr;
    ~Stream() { if (s) cudaStreamDestroy(s); }
};

// Time `iters` launches of the read/copy kernel running on `run_dev`; src may live on another device.
// Every mode ends by checking the bytes that arrived (READ: as folded; COPY / PUSH: the destination, re-read) against the
// pattern's definition (the source was filled with seed `seed`).
int time_kernel(int run_dev, void *dst, const void *src, size_t bytes, int warmup, int iters, int mode,
                gfd_probe_result *out, bool src_is_peer = false, uint64_t seed = 0) {
    DeviceScope scope(run_dev);
    if (scope.rc) return scope.rc;
    DevInfo di;
    int rc = device_info(run_dev, &di);
    if (rc) return rc;
    Stream st;
    CUDA_TRY(cudaStreamCreateWithFlags(&st.s, cudaStreamNonBlocking));
    Buf scratch, d_out;
    if ((rc = alloc_scratch(&scratch.p))) return rc;
    CUDA_TRY(cudaMalloc(&d_out.p, sizeof(gfd_checksum)));
    Ev e0, e1;
    CUDA_TRY(cudaEventCreate(&e0.e));
    CUDA_TRY(cudaEventCreate(&e1.e));
    int grid = 0;
    auto once = [&]() -> int {
        if (mode == GFD_PROBE_WRITE) return launch_fill(run_dev, dst, bytes, seed, st.s, &grid);
        if (mode == GFD_PROBE_DMA) {  // the copy engine, for comparison: no kernel of ours moves these bytes
            CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, st.s));
            return 0;
        }
        if (mode == GFD_PROBE_COPY || mode == GFD_PROBE_PUSH) return launch_copy(run_dev, dst, src, bytes, st.s, &grid, src_is_peer);
        return launch_read(run_dev, src, bytes, static_cast<gfd_checksum *>(d_out.p), scratch.p, st.s, &grid, src_is_peer);
    };
    for (int i = 0; i < warmup; i++)
        if ((rc = once())) return rc;
    CUDA_TRY(cudaStreamSynchronize(st.s));
    CUDA_TRY(cudaEventRecord(e0.e, st.s));
    for (int i = 0; i < iters; i++)
        if ((rc = once())) return rc;
    CUDA_TRY(cudaEventRecord(e1.e, st.s));
    CUDA_TRY(cudaEventSynchronize(e1.e));
    float ms = 0;
    CUDA_TRY(cudaEventElapsedTime(&ms, e0.e, e1.e));
    memset(out, 0, sizeof *out);
    out->bytes = bytes;
    out->iters = iters;
    out->mode = mode;
    out->sm_count = di.sms;
    out->grid = grid;
    out->block = kThreads;
    out->ms_per_launch = (double)ms / iters;
    out->gbps = (double)bytes * iters / ((double)ms * 1e-3) / 1e9;
    {
        // READ: d_out holds the fold of the last timed launch.  COPY / PUSH: fold the destination now (PUSH: it is the peer's)
        if (mode != GFD_PROBE_READ) {
            if ((rc = launch_read(run_dev, dst, bytes, static_cast<gfd_checksum *>(d_out.p), scratch.p, st.s, nullptr, mode == GFD_PROBE_PUSH))) return rc;
            CUDA_TRY(cudaStreamSynchronize(st.s));
        }
        CUDA_TRY(cudaMemcpy(&out->checksum, d_out.p, sizeof(gfd_checksum), cudaMemcpyDeviceToHost));
        if ((rc = launch_pattern(run_dev, bytes, seed, static_cast<gfd_checksum *>(d_out.p), scratch.p, st.s))) return rc;
        CUDA_TRY(cudaStreamSynchronize(st.s));
        CUDA_TRY(cudaMemcpy(&out->expected, d_out.p, sizeof(gfd_checksum), cudaMemcpyDeviceToHost));
        out->verified = (out->checksum.xor64 == out->expected.xor64 && out->checksum.sum64 == out->expected.sum64) ? 1 : 0;
        if (!out->verified) return GFD_ERR_DATA_MISMATCH;
    }
    return 0;
}
}  // namespace

GFD_API int gfd_probe_hbm(int cuda_device, size_t bytes, int warmup, int iters, int mode, gfd_probe_result *out) {
    int n;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (!out || cuda_device < 0 || cuda_device >= n || bytes == 0 || iters <= 0 || warmup < 0 ||
        (mode != GFD_PROBE_READ && mode != GFD_PROBE_COPY && mode != GFD_PROBE_WRITE))
        return GFD_ERR_BAD_ARGUMENT;
    DeviceScope scope(cuda_device);
    if (scope.rc) return scope.rc;
    Buf src, dst;
    if (mode != GFD_PROBE_WRITE) {
        CUDA_TRY(cudaMalloc(&src.p, bytes));
        if ((rc = launch_fill(cuda_device, src.p, bytes, 0, nullptr))) return rc;
    }
    if (mode != GFD_PROBE_READ) CUDA_TRY(cudaMalloc(&dst.p, bytes));
    CUDA_TRY(cudaDeviceSynchronize());
    rc = time_kernel(cuda_device, dst.p, src.p, bytes, warmup, iters, mode, out);
    if (rc) return rc;
    if (mode == GFD_PROBE_COPY) out->gbps *= 2.0;  // read + write both cross HBM
    return 0;
}

GFD_API int gfd_probe_p2p(int src_device, int dst_device, size_t bytes, int warmup, int iters, int mode,
                          gfd_probe_result *out) {
    int n;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (!out || src_device < 0 || dst_device < 0 || src_device >= n || dst_device >= n || src_device == dst_device ||
        bytes == 0 || iters <= 0 || warmup < 0 ||
        (mode != GFD_PROBE_READ && mode != GFD_PROBE_COPY && mode != GFD_PROBE_PUSH && mode != GFD_PROBE_DMA))
        return GFD_ERR_BAD_ARGUMENT;
    // pull modes run on the destination GPU and load from the source's memory; push runs on the source GPU and stores
    // into the destination's memory (posted writes: no request/response round trip on the link)
    const int run_device = (mode == GFD_PROBE_PUSH) ? src_device : dst_device;
    const int far_device = (mode == GFD_PROBE_PUSH) ? dst_device : src_device;
    int can = 0;
    CUDA_TRY(cudaDeviceCanAccessPeer(&can, run_device, far_device));
    if (!can) return GFD_ERR_PEER_ACCESS;
    Buf src, dst;
    {
        DeviceScope s(src_device);
        if (s.rc) return s.rc;
        CUDA_TRY(cudaMalloc(&src.p, bytes));
        if ((rc = launch_fill(src_device, src.p, bytes, 0, nullptr))) return rc;
        CUDA_TRY(cudaDeviceSynchronize());
    }
    {
        DeviceScope s(run_device);
        if (s.rc) return s.rc;
        cudaError_t e = cudaDeviceEnablePeerAccess(far_device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
        else if (e != cudaSuccess) return GFD_ERR_CUDA + (int)e;
    }
    if (mode != GFD_PROBE_READ) {
        DeviceScope s(dst_device);
        if (s.rc) return s.rc;
        CUDA_TRY(cudaMalloc(&dst.p, bytes));
    }
    rc = time_kernel(run_device, dst.p, src.p, bytes, warmup, iters, mode, out, /*src_is_peer=*/mode != GFD_PROBE_PUSH);
    {
        DeviceScope s(src_device);  // free on the owning device
        if (src.p) { cudaFree(src.p); src.p = nullptr; }
    }
    {
        DeviceScope s(dst_device);
        if (dst.p) { cudaFree(dst.p); dst.p = nullptr; }
    }
    return rc;
}

// ---- all pairs at once: the round-robin tournament of SURVEY.md 8(e) ---------------------------------------
// n GPUs (n even): n-1 rounds of n/2 disjoint pairs (circle method); in every round each pair runs BOTH directions
// at the same time, so all n GPUs pull from a distinct peer concurrently -- 2 * (n/2) = n kernels in flight, each
// timed with its own events on its own stream.  Then one "shift" round per k = 1..n-1 is NOT needed: the tournament
// already covers every directed pair exactly once.  Finally an all-at-once ring round (d pulls from d+1) is repeated
// to report the aggregate.  No NCCL, no collective: nothing is reduced.
namespace {
struct P2PDev {
    void *src = nullptr, *dst = nullptr, *scratch = nullptr, *out = nullptr;
    cudaStream_t st = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
};

int p2p_launch(int run_dev, P2PDev &me, const void *peer_src, size_t bytes, int mode, int reps, bool timed) {
    DeviceScope s(run_dev);
    if (s.rc) return s.rc;
    int rc;
    if (timed) CUDA_TRY(cudaEventRecord(me.e0, me.st));
    for (int i = 0; i < reps; i++) {
        if (mode == GFD_PROBE_DMA) {
            CUDA_TRY(cudaMemcpyAsync(me.dst, peer_src, bytes, cudaMemcpyDeviceToDevice, me.st));
            rc = 0;
        } else if (mode == GFD_PROBE_COPY) rc = launch_copy(run_dev, me.dst, peer_src, bytes, me.st, nullptr, true);
        else if (mode == GFD_PROBE_PUSH) rc = launch_copy(run_dev, const_cast<void *>(peer_src), me.src, bytes, me.st, nullptr);  // peer_src = the partner's dst
        else rc = launch_read(run_dev, peer_src, bytes, static_cast<gfd_checksum *>(me.out), me.scratch, me.st, nullptr, true);
        if (rc) return rc;
    }
    if (timed) CUDA_TRY(cudaEventRecord(me.e1, me.st));
    return 0;
}

int p2p_elapsed(int dev, P2PDev &me, float *ms) {
    DeviceScope s(dev);
    if (s.rc) return s.rc;
    CUDA_TRY(cudaEventSynchronize(me.e1));
    CUDA_TRY(cudaEventElapsedTime(ms, me.e0, me.e1));
    return 0;
}
}  // namespace

GFD_API int gfd_probe_p2p_matrix(int n_devices, size_t bytes, int warmup, int iters, int mode, double *gbps_matrix,
                                 double *ring_aggregate_gbps) {
    int n;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (!gbps_matrix || n_devices < 2 || n_devices > n || n_devices > kMaxDevices || bytes == 0 || iters <= 0 || warmup < 0 ||
        (mode != GFD_PROBE_READ && mode != GFD_PROBE_COPY && mode != GFD_PROBE_PUSH && mode != GFD_PROBE_DMA))
        return GFD_ERR_BAD_ARGUMENT;
    const int N = n_devices;
    for (int i = 0; i < N * N; i++) gbps_matrix[i] = 0.0;
    std::vector<P2PDev> dev((size_t)N);
    std::vector<gfd_checksum> expected((size_t)N);
    auto cleanup = [&]() {
        for (int d = 0; d < N; d++) {
            DeviceScope s(d);
            if (dev[d].st) cudaStreamSynchronize(dev[d].st);
            if (dev[d].src) cudaFree(dev[d].src);
            if (dev[d].dst) cudaFree(dev[d].dst);
            if (dev[d].scratch) cudaFree(dev[d].scratch);
            if (dev[d].out) cudaFree(dev[d].out);
            if (dev[d].e0) cudaEventDestroy(dev[d].e0);
            if (dev[d].e1) cudaEventDestroy(dev[d].e1);
            if (dev[d].st) cudaStreamDestroy(dev[d].st);
        }
    };
    auto body = [&]() -> int {
        for (int d = 0; d < N; d++) {
            DeviceScope s(d);
            if (s.rc) return s.rc;
            for (int p = 0; p < N; p++) {
                if (p == d) continue;
                int can = 0;
                CUDA_TRY(cudaDeviceCanAccessPeer(&can, d, p));
                if (!can) return GFD_ERR_PEER_ACCESS;
                cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
                else if (e != cudaSuccess) return GFD_ERR_CUDA + (int)e;
            }
            CUDA_TRY(cudaMalloc(&dev[d].src, bytes));
            if (mode != GFD_PROBE_READ) CUDA_TRY(cudaMalloc(&dev[d].dst, bytes));
            { int arc = alloc_scratch(&dev[d].scratch); if (arc) return arc; }
            CUDA_TRY(cudaMalloc(&dev[d].out, sizeof(gfd_checksum)));
            CUDA_TRY(cudaStreamCreateWithFlags(&dev[d].st, cudaStreamNonBlocking));
            CUDA_TRY(cudaEventCreate(&dev[d].e0));
            CUDA_TRY(cudaEventCreate(&dev[d].e1));
            int frc = launch_fill(d, dev[d].src, bytes, (uint64_t)d << 40, dev[d].st);
            if (frc) return frc;
            // what any reader of this GPU's buffer must find (READ mode checks it after every timed round)
            if ((frc = launch_pattern(d, bytes, (uint64_t)d << 40, static_cast<gfd_checksum *>(dev[d].out), dev[d].scratch, dev[d].st))) return frc;
            CUDA_TRY(cudaMemcpyAsync(&expected[(size_t)d], dev[d].out, sizeof(gfd_checksum), cudaMemcpyDeviceToHost, dev[d].st));
            CUDA_TRY(cudaStreamSynchronize(dev[d].st));
        }
        // GPU d has just received GPU `origin`'s bytes: folded into dev[d].out (READ) or stored in dev[d].dst (COPY: pulled by
        // d; PUSH: stored there by `origin`).  Call only after EVERY stream of the round has been synchronised.
        auto verify = [&](int d, int origin) -> int {
            DeviceScope s(d);
            if (s.rc) return s.rc;
            if (mode != GFD_PROBE_READ) {
                int vrc = launch_read(d, dev[d].dst, bytes, static_cast<gfd_checksum *>(dev[d].out), dev[d].scratch, dev[d].st, nullptr);
                if (vrc) return vrc;
                CUDA_TRY(cudaStreamSynchronize(dev[d].st));
            }
            gfd_checksum got;
            CUDA_TRY(cudaMemcpy(&got, dev[d].out, sizeof got, cudaMemcpyDeviceToHost));
            return (got.xor64 == expected[(size_t)origin].xor64 && got.sum64 == expected[(size_t)origin].sum64) ? 0 : GFD_ERR_DATA_MISMATCH;
        };
        // circle method; an odd N gets a bye (partner == N)
        const int M = (N % 2 == 0) ? N : N + 1;
        std::vector<int> partner((size_t)M);
        for (int round = 0; round < M - 1; round++) {
            for (int i = 0; i < M; i++) partner[(size_t)i] = -1;
            auto seat = [&](int k) { return k == 0 ? M - 1 : (round + k - 1) % (M - 1); };  // fixed seat + rotating ring
            for (int k = 0; k < M / 2; k++) {
                int a = seat(k), b = seat(M - 1 - k);
                partner[(size_t)a] = b;
                partner[(size_t)b] = a;
            }
            for (int pass = 0; pass < 2; pass++) {  // pass 0 = warm-up launches, pass 1 = timed
                int reps = pass == 0 ? warmup : iters;
                if (reps == 0) continue;
                for (int d = 0; d < N; d++) {
                    int p = partner[(size_t)d];
                    if (p < 0 || p >= N) continue;
                    int lrc = p2p_launch(d, dev[d], mode == GFD_PROBE_PUSH ? dev[p].dst : dev[p].src, bytes, mode, reps, pass == 1);
                    if (lrc) return lrc;
                }
                for (int d = 0; d < N; d++) {
                    int p = partner[(size_t)d];
                    if (p < 0 || p >= N) continue;
                    DeviceScope s(d);
                    CUDA_TRY(cudaStreamSynchronize(dev[d].st));
                    if (pass == 1) {
                        float ms = 0;
                        int erc = p2p_elapsed(d, dev[d], &ms);
                        if (erc) return erc;
                        gbps_matrix[d * N + p] = (double)bytes * iters / ((double)ms * 1e-3) / 1e9;  // row = puller, column = source
                    }
                }
                if (pass == 1)
                    for (int d = 0; d < N; d++) {
                        int p = partner[(size_t)d], vrc;
                        if (p >= 0 && p < N && (vrc = verify(d, p))) return vrc;
                    }
            }
        }
        if (ring_aggregate_gbps) {  // every GPU pulls from its right-hand neighbour at once
            *ring_aggregate_gbps = 0.0;
            for (int pass = 0; pass < 2; pass++) {
                int reps = pass == 0 ? warmup : iters;
                if (reps == 0) continue;
                for (int d = 0; d < N; d++) {
                    int lrc = p2p_launch(d, dev[d], mode == GFD_PROBE_PUSH ? dev[(d + 1) % N].dst : dev[(d + 1) % N].src, bytes, mode, reps, pass == 1);
                    if (lrc) return lrc;
                }
                for (int d = 0; d < N; d++) {
                    DeviceScope s(d);
                    CUDA_TRY(cudaStreamSynchronize(dev[d].st));
                    if (pass == 1) {
                        float ms = 0;
                        int erc = p2p_elapsed(d, dev[d], &ms);
                        if (erc) return erc;
                        *ring_aggregate_gbps += (double)bytes * iters / ((double)ms * 1e-3) / 1e9;
                    }
                }
                if (pass == 1)
                    for (int d = 0; d < N; d++) {
                        int vrc = verify(d, mode == GFD_PROBE_PUSH ? (d + N - 1) % N : (d + 1) % N);
                        if (vrc) return vrc;
                    }
            }
        }
        return 0;
    };
    rc = body();
    cleanup();
    return rc;
}

GFD_API int gfd_cuda_device_by_uuid(const char *uuid, int *ordinal) {
    int n;
    int rc = device_count_checked(&n);
    if (rc) return rc;
    if (!uuid || !ordinal) return GFD_ERR_BAD_ARGUMENT;
    for (int d = 0; d < n; d++) {
        cudaDeviceProp prop;
        CUDA_TRY(cudaGetDeviceProperties(&prop, d));
        const unsigned char *b = reinterpret_cast<const unsigned char *>(prop.uuid.bytes);
        char s[64];
        snprintf(s, sizeof s, "GPU-%02x%02x%02x%02x-%02x%02x-%02x%02x-%02x%02x-%02x%02x%02x%02x%02x%02x", b[0], b[1], b[2],
                 b[3], b[4], b[5], b[6], b[7], b[8], b[9], b[10], b[11], b[12], b[13], b[14], b[15]);
        if (strcmp(s, uuid) == 0) { *ordinal = d; return 0; }
    }
    return GFD_ERR_BAD_ARGUMENT;
}
