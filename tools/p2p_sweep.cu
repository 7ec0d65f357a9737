// Experiment harness (not product, not a bench): which access path moves bytes over NVLink fastest between two B200s?
// GPU A runs the kernel; the far buffer lives on GPU B and is reached through a peer mapping.  Variants: the product's
// kernels (LDG.128 pull with a dynamic tile counter, one-tile-per-CTA copy), persistent LDG, 256-bit loads, TMA bulk
// copies (cp.async.bulk) as the puller and as the pusher, and the copy engine (cudaMemcpyPeerAsync) for reference.
// Each is timed one way (A only) and both ways at once (A and B each run it against the other).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/p2p_sweep tools/p2p_sweep.cu
#define HBM_SWEEP_NO_MAIN
#include "hbm_sweep.cu"

struct U4 { uint64_t a, b, c, d; };
__device__ __forceinline__ U4 ld32(const void *p) {
    U4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r.a), "=l"(r.b), "=l"(r.c), "=l"(r.d) : "l"(p));
    return r;
}
__device__ __forceinline__ void st32(void *p, U4 v) {
    asm volatile("st.global.L1::no_allocate.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(v.a), "l"(v.b), "l"(v.c), "l"(v.d) : "memory");
}
// persistent, 256-bit loads: T threads x U vectors of 32 bytes per tile
template <int T, int U, int BPS>
__global__ void __launch_bounds__(T, BPS) read_ldg256(const char *__restrict__ src, size_t n32, ulonglong2 *out) {
    constexpr size_t kTile = (size_t)T * U;
    const size_t tiles = n32 / kTile;
    uint64_t x = 0, s = 0;
    for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        const char *base = src + t * kTile * 32;
        U4 v[U];
#pragma unroll
        for (int j = 0; j < U; j++) v[j] = ld32(base + ((size_t)j * T + threadIdx.x) * 32);
#pragma unroll
        for (int j = 0; j < U; j++) { x ^= v[j].a ^ v[j].b ^ v[j].c ^ v[j].d; s += v[j].a + v[j].b + v[j].c + v[j].d; }
    }
    if (x == 0x1234567 && s == 0x7654321) out[blockIdx.x * T + threadIdx.x] = make_ulonglong2(x, s);
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = make_ulonglong2(x, s);
}
template <int T, int U, int BPS>
__global__ void __launch_bounds__(T, BPS) copy_onetile256(char *__restrict__ dst, const char *__restrict__ src) {
    constexpr size_t kTile = (size_t)T * U;
    const size_t off = (size_t)blockIdx.x * kTile * 32;
    U4 v[U];
#pragma unroll
    for (int j = 0; j < U; j++) v[j] = ld32(src + off + ((size_t)j * T + threadIdx.x) * 32);
#pragma unroll
    for (int j = 0; j < U; j++) st32(dst + off + ((size_t)j * T + threadIdx.x) * 32, v[j]);
}

__global__ void fill(uint64_t *p, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = 0x9E3779B97F4A7C15ull * i;
}

struct Side {
    int dev = 0, sms = 0;
    cudaStream_t st = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    char *src = nullptr, *dst = nullptr;
    ulonglong2 *out = nullptr;
    unsigned long long *counter = nullptr;
};
static Side g[2];
static size_t g_bytes;

// launch(me, peer): enqueue ONE pass of the variant on me.st (me.dev is current)
template <class F>
void run(const char *name, F launch, int warm = 2, int iters = 10) {
    double uni = 0, bidi[2] = {0, 0};
    for (int mode = 0; mode < 2; mode++) {  // 0: GPU 0 alone, 1: both at once
        int nside = mode == 0 ? 1 : 2;
        for (int pass = 0; pass < 2; pass++) {
            int reps = pass == 0 ? warm : iters;
            for (int a = 0; a < nside; a++) {
                CK(cudaSetDevice(g[a].dev));
                if (pass == 1) CK(cudaEventRecord(g[a].e0, g[a].st));
                for (int i = 0; i < reps; i++) launch(g[a], g[a ^ 1]);
                if (pass == 1) CK(cudaEventRecord(g[a].e1, g[a].st));
            }
            for (int a = 0; a < nside; a++) {
                CK(cudaSetDevice(g[a].dev));
                CK(cudaStreamSynchronize(g[a].st));
                CK(cudaGetLastError());
            }
        }
        for (int a = 0; a < nside; a++) {
            float ms;
            CK(cudaEventElapsedTime(&ms, g[a].e0, g[a].e1));
            double gbps = (double)g_bytes * iters / (ms * 1e-3) / 1e9;
            if (mode == 0) uni = gbps; else bidi[a] = gbps;
        }
    }
    printf("%-58s one way %7.1f GB/s | both ways %7.1f + %7.1f GB/s\n", name, uni, bidi[0], bidi[1]);
    fflush(stdout);
}

int main(int argc, char **argv) {
    g_bytes = (argc > 1 ? strtoull(argv[1], nullptr, 0) : 1024) << 20;
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("needs 2 GPUs\n"); return 1; }
    for (int a = 0; a < 2; a++) {
        g[a].dev = a;
        CK(cudaSetDevice(a));
        int can = 0;
        CK(cudaDeviceCanAccessPeer(&can, a, a ^ 1));
        if (!can) { printf("no peer access %d -> %d\n", a, a ^ 1); return 1; }
        CK(cudaDeviceEnablePeerAccess(a ^ 1, 0));
        CK(cudaDeviceGetAttribute(&g[a].sms, cudaDevAttrMultiProcessorCount, a));
        CK(cudaStreamCreateWithFlags(&g[a].st, cudaStreamNonBlocking));
        CK(cudaEventCreate(&g[a].e0));
        CK(cudaEventCreate(&g[a].e1));
        CK(cudaMalloc(&g[a].src, g_bytes));
        CK(cudaMalloc(&g[a].dst, g_bytes));
        CK(cudaMalloc(&g[a].out, sizeof(ulonglong2) * 1024 * 1200));
        CK(cudaMalloc(&g[a].counter, 64));
        fill<<<g[a].sms * 8, 512>>>((uint64_t *)g[a].src, g_bytes / 8);
        CK(cudaDeviceSynchronize());
    }
    const size_t n16 = g_bytes / 16, n32 = g_bytes / 32;
    printf("bytes per pass per direction = %zu, sms = %d\n", g_bytes, g[0].sms);

    // ---- pull, read only (the far GPU's memory is loaded and folded) ----
    run("pull read  dyn LDG.128 T512 U8 x2/SM   [product]", [&](Side &me, Side &peer) {
        CK(cudaMemsetAsync(me.counter, 0, 8, me.st));
        read_dyn<512, 8, 2><<<me.sms * 2, 512, 0, me.st>>>(peer.src, n16, me.out, me.counter);
    });
    run("pull read  persistent LDG.128 T512 U8 x2/SM", [&](Side &me, Side &peer) { read_ldg<512, 8, 2, 2><<<me.sms * 2, 512, 0, me.st>>>(peer.src, n16, me.out); });
    run("pull read  persistent LDG.128 T512 U8 x2/SM, no hints", [&](Side &me, Side &peer) { read_ldg<512, 8, 2, 0><<<me.sms * 2, 512, 0, me.st>>>(peer.src, n16, me.out); });
    run("pull read  persistent LDG.128 T512 U8 x2/SM, L1::no_allocate only", [&](Side &me, Side &peer) { read_ldg<512, 8, 2, 1><<<me.sms * 2, 512, 0, me.st>>>(peer.src, n16, me.out); });
    run("pull read  persistent LDG.128 T512 U8 x2/SM, no hints (again)", [&](Side &me, Side &peer) { read_ldg<512, 8, 2, 0><<<me.sms * 2, 512, 0, me.st>>>(peer.src, n16, me.out); });
    run("pull read  persistent LDG.128 T512 U8 x2/SM, all hints (again)", [&](Side &me, Side &peer) { read_ldg<512, 8, 2, 2><<<me.sms * 2, 512, 0, me.st>>>(peer.src, n16, me.out); });
    run("pull read  persistent LDG.128 T1024 U4 x2/SM", [&](Side &me, Side &peer) { read_ldg<1024, 4, 2, 2><<<me.sms * 2, 1024, 0, me.st>>>(peer.src, n16, me.out); });
    run("pull read  persistent LDG.128 T512 U16 x1/SM", [&](Side &me, Side &peer) { read_ldg<512, 16, 1, 2><<<me.sms, 512, 0, me.st>>>(peer.src, n16, me.out); });
    run("pull read  persistent LDG.128 T256 U8 x2/SM (half the loads in flight)", [&](Side &me, Side &peer) { read_ldg<256, 8, 2, 2><<<me.sms * 2, 256, 0, me.st>>>(peer.src, n16, me.out); });
    run("pull read  persistent LDG.256 T512 U4 x2/SM", [&](Side &me, Side &peer) { read_ldg256<512, 4, 2><<<me.sms * 2, 512, 0, me.st>>>(peer.src, n32, me.out); });
    run("pull read  persistent LDG.256 T512 U8 x2/SM", [&](Side &me, Side &peer) { read_ldg256<512, 8, 2><<<me.sms * 2, 512, 0, me.st>>>(peer.src, n32, me.out); });
#define RTMA(ST, CH, CW)                                                                                          \
    {                                                                                                             \
        size_t smem = (size_t)ST * CH;                                                                            \
        for (int a = 0; a < 2; a++) { CK(cudaSetDevice(a)); CK(cudaFuncSetAttribute(read_tma<ST, CH, CW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); } \
        run("pull read  TMA bulk stages=" #ST " chunk=" #CH " consumers=" #CW, [&](Side &me, Side &peer) {       \
            read_tma<ST, CH, CW><<<me.sms, (CW + 1) * 32, smem, me.st>>>(peer.src, g_bytes / CH, me.out);        \
        });                                                                                                       \
    }
    RTMA(6, 32768, 8)
    RTMA(12, 16384, 8)
    RTMA(24, 8192, 8)
    RTMA(48, 4096, 8)

    // ---- pull copy (far -> near HBM) ----
    run("pull copy  one tile/CTA LDG.128 T1024 U2     [product]", [&](Side &me, Side &peer) { copy_onetile<1024, 2, 2><<<(unsigned)(n16 / 2048), 1024, 0, me.st>>>(me.dst, peer.src); });
    run("pull copy  one tile/CTA LDG.128 T512 U8", [&](Side &me, Side &peer) { copy_onetile<512, 8, 2><<<(unsigned)(n16 / 4096), 512, 0, me.st>>>(me.dst, peer.src); });
    run("pull copy  one tile/CTA LDG.256 T512 U4", [&](Side &me, Side &peer) { copy_onetile256<512, 4, 2><<<(unsigned)(n32 / 2048), 512, 0, me.st>>>(me.dst, peer.src); });
    run("pull copy  persistent LDG.128 T512 U8 x2/SM, loads without hints", [&](Side &me, Side &peer) { copy_ldg<512, 8, 2, 0, 1><<<me.sms * 2, 512, 0, me.st>>>(me.dst, peer.src, n16); });
    run("pull copy  persistent LDG.128 T512 U8 x2/SM, loads L1::no_allocate only", [&](Side &me, Side &peer) { copy_ldg<512, 8, 2, 1, 1><<<me.sms * 2, 512, 0, me.st>>>(me.dst, peer.src, n16); });
    run("pull copy  persistent LDG.128 T512 U8 x2/SM", [&](Side &me, Side &peer) { copy_ldg<512, 8, 2, 2, 1><<<me.sms * 2, 512, 0, me.st>>>(me.dst, peer.src, n16); });
#define CTMA(ST, CH, WHAT, DST, SRC)                                                                              \
    {                                                                                                             \
        size_t smem = (size_t)ST * CH;                                                                            \
        for (int a = 0; a < 2; a++) { CK(cudaSetDevice(a)); CK(cudaFuncSetAttribute(copy_tma<ST, CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); } \
        run(WHAT " TMA bulk load+store stages=" #ST " chunk=" #CH, [&](Side &me, Side &peer) {                   \
            copy_tma<ST, CH><<<me.sms, 64, smem, me.st>>>(DST, SRC, g_bytes / CH);                               \
        });                                                                                                       \
    }
    CTMA(6, 32768, "pull copy ", me.dst, peer.src)
    CTMA(12, 16384, "pull copy ", me.dst, peer.src)
    CTMA(24, 8192, "pull copy ", me.dst, peer.src)
    run("pull copy  cudaMemcpyPeerAsync (copy engine)", [&](Side &me, Side &peer) { CK(cudaMemcpyPeerAsync(me.dst, me.dev, peer.src, peer.dev, g_bytes, me.st)); });

    // ---- push (near HBM -> far) ----
    run("push       one tile/CTA LDG/STG.128 T1024 U2 [product]", [&](Side &me, Side &peer) { copy_onetile<1024, 2, 2><<<(unsigned)(n16 / 2048), 1024, 0, me.st>>>(peer.dst, me.src); });
    run("push       one tile/CTA LDG/STG.128 T512 U8", [&](Side &me, Side &peer) { copy_onetile<512, 8, 2><<<(unsigned)(n16 / 4096), 512, 0, me.st>>>(peer.dst, me.src); });
    run("push       one tile/CTA LDG/STG.256 T512 U4", [&](Side &me, Side &peer) { copy_onetile256<512, 4, 2><<<(unsigned)(n32 / 2048), 512, 0, me.st>>>(peer.dst, me.src); });
    run("push       persistent LDG/STG.128 T512 U8 x2/SM", [&](Side &me, Side &peer) { copy_ldg<512, 8, 2, 2, 1><<<me.sms * 2, 512, 0, me.st>>>(peer.dst, me.src, n16); });
    CTMA(6, 32768, "push      ", peer.dst, me.src)
    CTMA(12, 16384, "push      ", peer.dst, me.src)
    run("push       cudaMemcpyPeerAsync (copy engine)", [&](Side &me, Side &peer) { CK(cudaMemcpyPeerAsync(peer.dst, peer.dev, me.src, me.dev, g_bytes, me.st)); });

    // did the last push really land?  compare the two ends of GPU 1's dst with GPU 0's src
    uint64_t a[4], b[4];
    int ok = 1;
    for (size_t o : {(size_t)0, g_bytes - 32}) {
        CK(cudaMemcpy(a, g[0].src + o, 32, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(b, g[1].dst + o, 32, cudaMemcpyDeviceToHost));
        for (int i = 0; i < 4; i++) ok &= a[i] == b[i];
    }
    printf("push verify: %s\n", ok ? "ok" : "MISMATCH");
    return 0;
}
