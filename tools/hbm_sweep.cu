// Experiment harness (not product, not a bench): sweep launch shapes / load paths for the HBM read and copy
// probes on one B200 and print GB/s per variant, so the product kernel's constants come from measurement.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/hbm_sweep tools/hbm_sweep.cu
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

struct U2 { uint64_t a, b; };

__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
template <int HINT>  // 0 plain nc, 1 nc + L1::no_allocate, 2 + L2 evict_first policy
__device__ __forceinline__ U2 ld16(const void *p, uint64_t pol) {
    U2 r;
    if (HINT == 0) asm volatile("ld.global.nc.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
    else if (HINT == 1) asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
    else asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u64 {%0,%1}, [%2], %3;" : "=l"(r.a), "=l"(r.b) : "l"(p), "l"(pol));
    return r;
}
template <int HINT>
__device__ __forceinline__ void st16(void *p, U2 v, uint64_t pol) {
    if (HINT == 0) asm volatile("st.global.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(v.a), "l"(v.b) : "memory");
    else if (HINT == 1) asm volatile("st.global.L1::no_allocate.v2.u64 [%0], {%1,%2};" ::"l"(p), "l"(v.a), "l"(v.b) : "memory");
    else asm volatile("st.global.L1::no_allocate.L2::cache_hint.v2.u64 [%0], {%1,%2}, %3;" ::"l"(p), "l"(v.a), "l"(v.b), "l"(pol) : "memory");
}

template <int T, int U, int BPS, int HINT>
__global__ void __launch_bounds__(T, BPS) read_ldg(const char *__restrict__ src, size_t n16, ulonglong2 *out) {
    constexpr size_t kTile = (size_t)T * U;
    const size_t tiles = n16 / kTile;
    uint64_t x = 0, s = 0;
    const uint64_t pol = policy_evict_first();
    for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        const char *base = src + t * kTile * 16;
        U2 v[U];
#pragma unroll
        for (int j = 0; j < U; j++) v[j] = ld16<HINT>(base + ((size_t)j * T + threadIdx.x) * 16, pol);
#pragma unroll
        for (int j = 0; j < U; j++) { x ^= v[j].a ^ v[j].b; s += v[j].a + v[j].b; }
    }
    if (x == 0x1234567 && s == 0x7654321) out[blockIdx.x * T + threadIdx.x] = make_ulonglong2(x, s);  // never true for the pattern
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = make_ulonglong2(x, s);
}

template <int T, int U, int BPS, int LH, int SH>
__global__ void __launch_bounds__(T, BPS) copy_ldg(char *__restrict__ dst, const char *__restrict__ src, size_t n16) {
    constexpr size_t kTile = (size_t)T * U;
    const size_t tiles = n16 / kTile;
    const uint64_t pol = policy_evict_first();
    for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        const size_t off = t * kTile * 16;
        U2 v[U];
#pragma unroll
        for (int j = 0; j < U; j++) v[j] = ld16<LH>(src + off + ((size_t)j * T + threadIdx.x) * 16, pol);
#pragma unroll
        for (int j = 0; j < U; j++) st16<SH>(dst + off + ((size_t)j * T + threadIdx.x) * 16, v[j], pol);
    }
}

// one tile per CTA (grid = tiles), block-reduce, then atomics into 1024 spread slots (must be zeroed beforehand)
template <int T, int U, int BPS>
__global__ void __launch_bounds__(T, BPS) read_onetile(const char *__restrict__ src, size_t n16, unsigned long long *slots) {
    constexpr size_t kTile = (size_t)T * U;
    const uint64_t pol = policy_evict_first();
    const char *base = src + (size_t)blockIdx.x * kTile * 16;
    U2 v[U];
#pragma unroll
    for (int j = 0; j < U; j++) v[j] = ld16<2>(base + ((size_t)j * T + threadIdx.x) * 16, pol);
    uint64_t x = 0, s = 0;
#pragma unroll
    for (int j = 0; j < U; j++) { x ^= v[j].a ^ v[j].b; s += v[j].a + v[j].b; }
    for (int o = 16; o > 0; o >>= 1) { x ^= __shfl_xor_sync(~0u, x, o); s += __shfl_xor_sync(~0u, s, o); }
    __shared__ uint64_t sx[T / 32], ss[T / 32];
    if ((threadIdx.x & 31) == 0) { sx[threadIdx.x >> 5] = x; ss[threadIdx.x >> 5] = s; }
    __syncthreads();
    if (threadIdx.x < 32) {
        x = threadIdx.x < T / 32 ? sx[threadIdx.x] : 0;
        s = threadIdx.x < T / 32 ? ss[threadIdx.x] : 0;
        for (int o = 16; o > 0; o >>= 1) { x ^= __shfl_xor_sync(~0u, x, o); s += __shfl_xor_sync(~0u, s, o); }
        if (threadIdx.x == 0) {
            unsigned slot = blockIdx.x & 1023;
            atomicXor(&slots[2 * slot], (unsigned long long)x);
            atomicAdd(&slots[2 * slot + 1], (unsigned long long)s);
        }
    }
}

// persistent CTAs pulling tiles from a global counter (dynamic load balance), next index prefetched one tile ahead
template <int T, int U, int BPS>
__global__ void __launch_bounds__(T, BPS) read_dyn(const char *__restrict__ src, size_t n16, ulonglong2 *out, unsigned long long *counter) {
    constexpr size_t kTile = (size_t)T * U;
    const size_t tiles = n16 / kTile;
    const uint64_t pol = policy_evict_first();
    __shared__ unsigned long long next[2];
    uint64_t x = 0, s = 0;
    if (threadIdx.x == 0) next[0] = atomicAdd(counter, 1ull);
    __syncthreads();
    int buf = 0;
    while (true) {
        unsigned long long t = next[buf];
        if (t >= tiles) break;
        if (threadIdx.x == 0) next[buf ^ 1] = atomicAdd(counter, 1ull);
        const char *base = src + t * kTile * 16;
        U2 v[U];
#pragma unroll
        for (int j = 0; j < U; j++) v[j] = ld16<2>(base + ((size_t)j * T + threadIdx.x) * 16, pol);
#pragma unroll
        for (int j = 0; j < U; j++) { x ^= v[j].a ^ v[j].b; s += v[j].a + v[j].b; }
        __syncthreads();
        buf ^= 1;
    }
    if (x == 0x1234567 && s == 0x7654321) out[blockIdx.x * T + threadIdx.x] = make_ulonglong2(x, s);
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = make_ulonglong2(x, s);
}

template <int T, int U, int BPS>
__global__ void __launch_bounds__(T, BPS) copy_onetile(char *__restrict__ dst, const char *__restrict__ src) {
    constexpr size_t kTile = (size_t)T * U;
    const uint64_t pol = policy_evict_first();
    const size_t off = (size_t)blockIdx.x * kTile * 16;
    U2 v[U];
#pragma unroll
    for (int j = 0; j < U; j++) v[j] = ld16<2>(src + off + ((size_t)j * T + threadIdx.x) * 16, pol);
#pragma unroll
    for (int j = 0; j < U; j++) st16<1>(dst + off + ((size_t)j * T + threadIdx.x) * 16, v[j], pol);
}

// ---- TMA bulk (cp.async.bulk, UBLKCP) ring: 1 producer warp + CW consumer warps ------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *b) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nWAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\nbra WAIT_LOOP;\nDONE:\n}\n" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void *gdst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}

template <int STAGES, int CHUNK, int CW>
__global__ void __launch_bounds__((CW + 1) * 32, 1) read_tma(const char *__restrict__ src, size_t n_chunks, ulonglong2 *out) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t full[STAGES], empty[STAGES];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(&full[s], 1); mbar_init(&empty[s], CW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const size_t mine = n_chunks > blockIdx.x ? (n_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    uint64_t x = 0, s64 = 0;
    if (warp == CW) {
        if (lane == 0) {
            const uint64_t pol = policy_evict_first();
            for (size_t k = 0; k < mine; k++) {
                int st = (int)(k % STAGES);
                uint32_t ph = (uint32_t)((k / STAGES) & 1);
                if (k >= STAGES) mbar_wait(&empty[st], ph ^ 1);
                mbar_expect_tx(&full[st], CHUNK);
                bulk_g2s(smem + (size_t)st * CHUNK, src + (blockIdx.x + k * gridDim.x) * (size_t)CHUNK, CHUNK, &full[st], pol);
            }
        }
    } else {
        const int ctid = threadIdx.x;  // 0 .. CW*32-1
        for (size_t k = 0; k < mine; k++) {
            int st = (int)(k % STAGES);
            uint32_t ph = (uint32_t)((k / STAGES) & 1);
            mbar_wait(&full[st], ph);
            const ulonglong2 *p = reinterpret_cast<const ulonglong2 *>(smem + (size_t)st * CHUNK);
#pragma unroll 8
            for (int i = ctid; i < CHUNK / 16; i += CW * 32) { ulonglong2 v = p[i]; x ^= v.x ^ v.y; s64 += v.x + v.y; }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[st]);
        }
    }
    if (x == 0x1234567 && s64 == 0x7654321) out[blockIdx.x * blockDim.x + threadIdx.x] = make_ulonglong2(x, s64);
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = make_ulonglong2(x, s64);
}

// copy via TMA: bulk load to smem, bulk store from smem (no register traffic at all)
template <int STAGES, int CHUNK>
__global__ void __launch_bounds__(64, 1) copy_tma(char *__restrict__ dst, const char *__restrict__ src, size_t n_chunks) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t full[STAGES];
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    const size_t mine = n_chunks > blockIdx.x ? (n_chunks - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const uint64_t pol = policy_evict_first();
    // prologue: fill the ring
    for (size_t k = 0; k < mine && k < (size_t)STAGES; k++) {
        mbar_expect_tx(&full[k], CHUNK);
        bulk_g2s(smem + k * CHUNK, src + (blockIdx.x + k * gridDim.x) * (size_t)CHUNK, CHUNK, &full[k], pol);
    }
    for (size_t k = 0; k < mine; k++) {
        int st = (int)(k % STAGES);
        uint32_t ph = (uint32_t)((k / STAGES) & 1);
        mbar_wait(&full[st], ph);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        bulk_s2g(dst + (blockIdx.x + k * gridDim.x) * (size_t)CHUNK, smem + (size_t)st * CHUNK, CHUNK);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        size_t nk = k + STAGES;
        if (nk < mine) {
            // the stage can be refilled once its store has finished READING smem
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            mbar_expect_tx(&full[st], CHUNK);
            bulk_g2s(smem + (size_t)st * CHUNK, src + (blockIdx.x + nk * gridDim.x) * (size_t)CHUNK, CHUNK, &full[st], pol);
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <int T, int U, int BPS, int LH, int SH>
__global__ void __launch_bounds__(T, BPS) copy_onetile_h(char *__restrict__ dst, const char *__restrict__ src) {
    constexpr size_t kTile = (size_t)T * U;
    const uint64_t pol = policy_evict_first();
    const size_t off = (size_t)blockIdx.x * kTile * 16;
    U2 v[U];
#pragma unroll
    for (int j = 0; j < U; j++) v[j] = ld16<LH>(src + off + ((size_t)j * T + threadIdx.x) * 16, pol);
#pragma unroll
    for (int j = 0; j < U; j++) st16<SH>(dst + off + ((size_t)j * T + threadIdx.x) * 16, v[j], pol);
}
// 256-bit loads and stores, one tile per CTA
template <int T, int U, int BPS>
__global__ void __launch_bounds__(T, BPS) copy_onetile_256(char *__restrict__ dst, const char *__restrict__ src) {
    constexpr size_t kTile = (size_t)T * U;
    const size_t off = (size_t)blockIdx.x * kTile * 32;
    uint64_t a[U], b[U], c[U], d[U];
#pragma unroll
    for (int j = 0; j < U; j++)
        asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(a[j]), "=l"(b[j]), "=l"(c[j]), "=l"(d[j]) : "l"(src + off + ((size_t)j * T + threadIdx.x) * 32));
#pragma unroll
    for (int j = 0; j < U; j++)
        asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(dst + off + ((size_t)j * T + threadIdx.x) * 32), "l"(a[j]), "l"(b[j]), "l"(c[j]), "l"(d[j]) : "memory");
}
// write-only: one tile per CTA
template <int T, int U, int BPS, int SH>
__global__ void __launch_bounds__(T, BPS) write_onetile(char *__restrict__ dst) {
    constexpr size_t kTile = (size_t)T * U;
    const uint64_t pol = policy_evict_first();
    const size_t first = (size_t)blockIdx.x * kTile;
#pragma unroll
    for (int j = 0; j < U; j++) {
        size_t i = first + (size_t)j * T + threadIdx.x;
        U2 v{0x9E3779B97F4A7C15ull * (2 * i), 0x9E3779B97F4A7C15ull * (2 * i + 1)};
        st16<SH>(dst + i * 16, v, pol);
    }
}

#ifndef HBM_SWEEP_NO_MAIN
static cudaEvent_t e0, e1;
template <class F>
double time_ms(F launch, int warm = 3, int iters = 10) {
    for (int i = 0; i < warm; i++) launch();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    for (int i = 0; i < iters; i++) launch();
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaGetLastError());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    return ms / iters;
}

__global__ void fill(uint64_t *p, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = 0x9E3779B97F4A7C15ull * i;
}

int main(int argc, char **argv) {
    size_t bytes = (argc > 1 ? strtoull(argv[1], nullptr, 0) : 4096) << 20;
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    char *src, *dst;
    ulonglong2 *out;
    CK(cudaMalloc(&src, bytes));
    CK(cudaMalloc(&dst, bytes));
    CK(cudaMalloc(&out, sizeof(ulonglong2) * 1024 * 4096));
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    fill<<<sms * 8, 512>>>((uint64_t *)src, bytes / 8);
    CK(cudaDeviceSynchronize());
    const size_t n16 = bytes / 16;
    const double gb = bytes / 1e9;
    printf("sms=%d bytes=%zu\n", sms, bytes);

#define COPY(T, U, BPS, LH, SH, MULT)                                                                            \
    {                                                                                                            \
        double ms = time_ms([&] { copy_ldg<T, U, BPS, LH, SH><<<sms * BPS * MULT, T>>>(dst, src, n16); });      \
        printf("copy  ldg T=%4d U=%2d BPS=%d lh=%d sh=%d grid=%5d : %8.1f GB/s (r+w) %.4f ms\n", T, U, BPS, LH, SH, sms * BPS * MULT, 2 * gb / ms * 1e3, ms); \
    }
#define READ(T, U, BPS, HINT, MULT)                                                                              \
    {                                                                                                            \
        double ms = time_ms([&] { read_ldg<T, U, BPS, HINT><<<sms * BPS * MULT, T>>>(src, n16, out); });        \
        printf("read  ldg T=%4d U=%2d BPS=%d hint=%d grid=%5d : %8.1f GB/s  %.4f ms\n", T, U, BPS, HINT, sms * BPS * MULT, gb / ms * 1e3, ms); \
    }
    if (argc > 2 && argv[2][0] == '4') {  // fourth stage: cache qualifiers and widths on the one-tile copy; write-only shapes
#define CH(T, U, BPS, LH, SH)                                                                                    \
    {                                                                                                            \
        size_t tiles = n16 / ((size_t)T * U);                                                                    \
        double ms = time_ms([&] { copy_onetile_h<T, U, BPS, LH, SH><<<(unsigned)tiles, T>>>(dst, src); });       \
        printf("copy  onetile T=%4d U=%2d BPS=%d lh=%d sh=%d : %8.1f GB/s (r+w) %.4f ms\n", T, U, BPS, LH, SH, 2 * gb / ms * 1e3, ms); \
    }
        for (int rep = 0; rep < 2; rep++) {
            CH(1024, 2, 2, 2, 1) CH(1024, 2, 2, 0, 0) CH(1024, 2, 2, 1, 1) CH(1024, 2, 2, 0, 1) CH(1024, 2, 2, 2, 0) CH(1024, 2, 2, 2, 2) CH(1024, 2, 2, 1, 0) CH(1024, 2, 2, 1, 2)
            CH(1024, 1, 2, 2, 1) CH(512, 2, 4, 2, 1) CH(512, 1, 4, 2, 1) CH(256, 2, 8, 2, 1) CH(1024, 4, 2, 2, 1)
#define C256(T, U, BPS)                                                                                          \
    {                                                                                                            \
        size_t tiles = (bytes / 32) / ((size_t)T * U);                                                           \
        double ms = time_ms([&] { copy_onetile_256<T, U, BPS><<<(unsigned)tiles, T>>>(dst, src); });             \
        printf("copy  onetile256 T=%4d U=%2d BPS=%d       : %8.1f GB/s (r+w) %.4f ms\n", T, U, BPS, 2 * gb / ms * 1e3, ms); \
    }
            C256(1024, 1, 2) C256(512, 2, 4) C256(1024, 2, 2) C256(512, 1, 4) C256(256, 2, 8)
#define WR(T, U, BPS, SH)                                                                                        \
    {                                                                                                            \
        size_t tiles = n16 / ((size_t)T * U);                                                                    \
        double ms = time_ms([&] { write_onetile<T, U, BPS, SH><<<(unsigned)tiles, T>>>(dst); });                 \
        printf("write onetile T=%4d U=%2d BPS=%d sh=%d      : %8.1f GB/s %.4f ms\n", T, U, BPS, SH, gb / ms * 1e3, ms); \
    }
            WR(1024, 2, 2, 1) WR(1024, 2, 2, 0) WR(1024, 2, 2, 2) WR(512, 4, 4, 1) WR(1024, 4, 2, 1) WR(1024, 1, 2, 1) WR(256, 8, 8, 1)
        }
        return 0;
    }
    if (argc > 2 && argv[2][0] == '3') {  // third stage: one tile per CTA, and dynamic tile counter
        unsigned long long *slots;
        CK(cudaMalloc(&slots, 2048 * 8 + 64));
        unsigned long long *counter = slots + 2048;
#define ONE(T, U, BPS)                                                                                           \
    {                                                                                                            \
        size_t tiles = n16 / ((size_t)T * U);                                                                    \
        double ms = time_ms([&] { CK(cudaMemsetAsync(slots, 0, 2048 * 8)); read_onetile<T, U, BPS><<<(unsigned)tiles, T>>>(src, n16, slots); }); \
        printf("read  onetile T=%4d U=%2d BPS=%d grid=%7zu (+memset+atomics) : %8.1f GB/s  %.4f ms\n", T, U, BPS, tiles, gb / ms * 1e3, ms); \
    }
        ONE(512, 8, 2) ONE(512, 4, 4) ONE(256, 8, 4) ONE(256, 16, 2) ONE(1024, 4, 2) ONE(512, 16, 1) ONE(256, 4, 8) ONE(128, 8, 8) ONE(1024, 8, 1) ONE(512, 2, 4)
#define DYN(T, U, BPS, MULT)                                                                                     \
    {                                                                                                            \
        double ms = time_ms([&] { CK(cudaMemsetAsync(counter, 0, 8)); read_dyn<T, U, BPS><<<sms * BPS * MULT, T>>>(src, n16, out, counter); }); \
        printf("read  dyn     T=%4d U=%2d BPS=%d grid=%7d (+memset)         : %8.1f GB/s  %.4f ms\n", T, U, BPS, sms * BPS * MULT, gb / ms * 1e3, ms); \
    }
        DYN(512, 8, 2, 1) DYN(512, 4, 4, 1) DYN(256, 8, 4, 1) DYN(1024, 4, 2, 1) DYN(512, 8, 2, 2) DYN(256, 16, 2, 1) DYN(512, 4, 2, 1) DYN(256, 4, 8, 1)
#define CONE(T, U, BPS)                                                                                          \
    {                                                                                                            \
        size_t tiles = n16 / ((size_t)T * U);                                                                    \
        double ms = time_ms([&] { copy_onetile<T, U, BPS><<<(unsigned)tiles, T>>>(dst, src); });                 \
        printf("copy  onetile T=%4d U=%2d BPS=%d grid=%7zu : %8.1f GB/s (r+w) %.4f ms\n", T, U, BPS, tiles, 2 * gb / ms * 1e3, ms); \
    }
        CONE(512, 8, 2) CONE(512, 4, 4) CONE(256, 8, 4) CONE(256, 4, 8) CONE(1024, 4, 2) CONE(512, 2, 4) CONE(128, 8, 8) CONE(256, 16, 2) CONE(1024, 2, 2) CONE(256, 2, 8)
        for (size_t mb : {64, 256, 1024}) {
            size_t b2 = mb << 20, m16 = b2 / 16;
            double ms = time_ms([&] { CK(cudaMemsetAsync(slots, 0, 2048 * 8)); read_onetile<512, 8, 2><<<(unsigned)(m16 / 4096), 512>>>(src, m16, slots); });
            printf("read  onetile %5zu MiB: %8.1f GB/s %.4f ms\n", mb, b2 / 1e9 / ms * 1e3, ms);
            ms = time_ms([&] { read_ldg<512, 8, 2, 2><<<sms * 2, 512>>>(src, m16, out); });
            printf("read  persist %5zu MiB: %8.1f GB/s %.4f ms\n", mb, b2 / 1e9 / ms * 1e3, ms);
        }
        return 0;
    }
    if (argc > 2) {  // second-stage sweep: grid oversubscription
        for (int mult : {1, 2, 4, 8, 16, 32, 64, 128}) {
            READ(512, 8, 2, 2, mult)
            READ(512, 8, 2, 0, mult)
            READ(256, 16, 2, 2, mult)
            READ(256, 8, 4, 2, mult)
            READ(1024, 4, 2, 2, mult)
        }
        for (int mult : {1, 2, 4, 8, 16, 32, 64, 128}) {
            COPY(512, 8, 2, 2, 1, mult)
            COPY(512, 8, 2, 0, 0, mult)
            COPY(512, 4, 4, 2, 1, mult)
            COPY(256, 8, 4, 2, 1, mult)
            COPY(512, 8, 2, 2, 2, mult)
        }
        return 0;
    }
    READ(512, 8, 2, 2, 1)
    READ(512, 8, 2, 2, 16)

#define RTMA(ST, CH, CW)                                                                                         \
    {                                                                                                            \
        size_t smem = (size_t)ST * CH;                                                                           \
        CK(cudaFuncSetAttribute(read_tma<ST, CH, CW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        double ms = time_ms([&] { read_tma<ST, CH, CW><<<sms, (CW + 1) * 32, smem>>>(src, bytes / CH, out); }); \
        printf("read  tma stages=%d chunk=%6d cw=%2d smem=%6zu        : %8.1f GB/s  %.4f ms\n", ST, CH, CW, smem, gb / ms * 1e3, ms); \
    }
    RTMA(4, 32768, 8)
    RTMA(6, 32768, 8)
    RTMA(8, 16384, 8)
    RTMA(12, 16384, 8)
    RTMA(3, 65536, 8)
    RTMA(6, 32768, 16)
    RTMA(13, 16384, 4)
    RTMA(24, 8192, 8)

    COPY(512, 8, 2, 2, 1, 1)
    COPY(512, 8, 2, 2, 0, 1)
    COPY(512, 8, 2, 2, 2, 1)
    COPY(512, 8, 2, 0, 0, 1)
    COPY(1024, 4, 2, 2, 1, 1)
    COPY(512, 4, 4, 2, 1, 1)
    COPY(256, 8, 4, 2, 1, 1)
    COPY(512, 4, 2, 2, 1, 1)
    COPY(256, 4, 8, 2, 1, 1)
    COPY(512, 8, 2, 2, 1, 8)
    COPY(512, 2, 4, 2, 1, 1)
    COPY(1024, 2, 2, 2, 1, 1)

#define CTMA(ST, CH)                                                                                             \
    {                                                                                                            \
        size_t smem = (size_t)ST * CH;                                                                           \
        CK(cudaFuncSetAttribute(copy_tma<ST, CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));     \
        double ms = time_ms([&] { copy_tma<ST, CH><<<sms, 64, smem>>>(dst, src, bytes / CH); });                \
        printf("copy  tma stages=%d chunk=%6d smem=%6zu              : %8.1f GB/s (r+w) %.4f ms\n", ST, CH, smem, 2 * gb / ms * 1e3, ms); \
    }
    CTMA(4, 32768)
    CTMA(6, 32768)
    CTMA(8, 16384)
    CTMA(12, 16384)
    CTMA(3, 65536)
    {
        double ms = time_ms([&] { CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice)); });
        printf("copy  cudaMemcpyAsync D2D                              : %8.1f GB/s (r+w) %.4f ms\n", 2 * gb / ms * 1e3, ms);
    }
    // verify the TMA copy actually copied
    CK(cudaMemset(dst, 0, bytes));
    {
        size_t smem = (size_t)6 * 32768;
        copy_tma<6, 32768><<<sms, 64, smem>>>(dst, src, bytes / 32768);
        CK(cudaDeviceSynchronize());
        uint64_t a[4], b[4];
        size_t offs[2] = {0, bytes - 32};
        int ok = 1;
        for (size_t o : offs) {
            CK(cudaMemcpy(a, src + o, 32, cudaMemcpyDeviceToHost));
            CK(cudaMemcpy(b, dst + o, 32, cudaMemcpyDeviceToHost));
            for (int i = 0; i < 4; i++) ok &= a[i] == b[i];
        }
        printf("tma copy verify: %s\n", ok ? "ok" : "MISMATCH");
    }
    return 0;
}
#endif  // HBM_SWEEP_NO_MAIN
