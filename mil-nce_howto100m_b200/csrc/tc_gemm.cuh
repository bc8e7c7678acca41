// tcgen05 / TMEM / TMA GEMM core for sm_100a, with pluggable epilogues.
//
//   C[m, n] = sum_k A[m, k] * Bm[n, k]        A (M, Kd), Bm (N, Kd) fp16, K-major
//
// One persistent CTA per SM, 192 threads:
//   warp 0      TMA producer   (cp.async.bulk.tensor -> 128B-swizzled smem ring)
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer
//   warps 2..5  epilogue: tcgen05.ld the fp32 accumulator (one TMEM lane = one
//               output row per thread) into registers, then run the epilogue functor
// Accumulators are double-buffered in TMEM (2 x BLOCK_N columns) and drained to
// registers per K-chunk, so the epilogue of tile t overlaps the MMAs of tile t+1
// and long K loops are summed in fp32 registers (round-to-nearest) rather than
// inside the tensor core (which truncates).
//
// The loss kernels never write C: their epilogues reduce the tile on the fly
// (row log-sum-exp for the MIL-NCE forward, the softmax-gradient tile G for the
// backward).  This replaces reference loss.py:11 (th.matmul) + :13-18.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tc {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // fp16 elements = 128 bytes = one SWIZZLE_128B row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 192;
constexpr int NUM_EPI_THREADS = 128;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

// ----------------------------------------------------------------------------
// PTX wrappers
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ uint64_t global_timer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// Bounded wait: a protocol bug traps (launch failure) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    uint64_t t0 = global_timer_ns();
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++spins & 1023u) == 0 && global_timer_ns() - t0 > 4000000000ull) __trap();
    }
}

__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void epi_bar_sync() {  // named barrier 1: the 128 epilogue threads
    asm volatile("bar.sync 1, 128;" ::: "memory");
}

__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, uint64_t *bar,
                                            int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

__device__ __forceinline__ void tmem_alloc(uint32_t *slot_in_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(slot_in_smem)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                 : "memory");
}

// K-major operand tile in smem, rows of 128 bytes, SWIZZLE_128B (what TMA wrote):
// 8-row groups are 1024 bytes apart (SBO); LBO is unused for swizzled K-major.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);  // start address  [0,14)
    d |= (uint64_t)1 << 16;                    // LBO (ignored)  [16,30)
    d |= (uint64_t)(1024 >> 4) << 32;          // SBO = 1024 B   [32,46)
    d |= (uint64_t)1 << 46;                    // descriptor version 1 (sm_100)
    d |= (uint64_t)2 << 61;                    // SWIZZLE_128B
    return d;
}
// kind::f16 instruction descriptor: fp16 x fp16 -> fp32, both operands K-major.
__host__ __device__ constexpr uint32_t make_idesc_f16(int m, int n) {
    return (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                         uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Arrives on `bar` once all previously issued MMAs of this thread have completed.
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile(
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
            smem_u32(bar))
        : "memory");
}

// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (lane = row).
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        "tcgen05.wait::ld.sync.aligned;\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
          "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
          "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
          "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
          "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ----------------------------------------------------------------------------
// Epilogue context handed to the functors
// ----------------------------------------------------------------------------
struct EpiCtx {
    int m0, n0;        // tile origin
    int m_blk, n_blk;  // tile coordinates
    int M, N;          // problem extent
    int tid;           // 0..127 within the epilogue warps; row in tile = tid
    int par;           // tile parity, for double-buffered smem scratch
    float *scratch;    // 2 * BLOCK_N floats of smem private to the epilogue warps
};
// Every functor receives the finished fp32 accumulator row of its thread in
// registers (acc[BLOCK_N], compile-time indexed): the kernel has already drained
// TMEM, so the tensor core is working on the next tile while the functor runs.

// ---- plain store: C = alpha * acc (dvideo / dtext GEMMs, and the test hook) ----
struct EpiStore {
    struct Params {
        float *C;
        long long ldc;
        float alpha;
        const float *alpha_dev;  // optional device scalar multiplied into alpha
    };
    template <int BLOCK_N>
    __device__ static void run(const Params &p, const EpiCtx &c, float (&acc)[BLOCK_N]) {
        float alpha = p.alpha;
        if (p.alpha_dev) alpha *= __ldg(p.alpha_dev);
        const int row = c.m0 + c.tid;
        if (row >= c.M) return;
#pragma unroll
        for (int ch = 0; ch < BLOCK_N / 32; ++ch) {
            const int col0 = c.n0 + ch * 32;
            if (col0 >= c.N) break;
            float *dst = p.C + (long long)row * p.ldc + col0;
            if (col0 + 32 <= c.N && ((((uintptr_t)dst) & 15) == 0)) {
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    const int e = ch * 32 + i;
                    float4 o = make_float4(acc[e] * alpha, acc[e + 1] * alpha, acc[e + 2] * alpha,
                                           acc[e + 3] * alpha);
                    *reinterpret_cast<float4 *>(dst + i) = o;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    if (col0 + i < c.N) dst[i] = acc[ch * 32 + i] * alpha;
            }
        }
    }
};

// ---- row log-sum-exp of the tile (MIL-NCE forward; reference loss.py:16-17) ----
// part[n_blk][row] = log sum_{n in tile} exp(scale * acc[row, n])      (natural log)
struct EpiRowLse {
    struct Params {
        float *part;
        long long ldp;           // rows per n-block slab
        const float *scale_dev;  // device scalar: un-does the fp16 staging scale
    };
    template <int BLOCK_N>
    __device__ static void run(const Params &p, const EpiCtx &c, float (&acc)[BLOCK_N]) {
        const float sc = __ldg(p.scale_dev) * LOG2E;  // work in the log2 domain
        float mx = -INFINITY;
#pragma unroll
        for (int i = 0; i < BLOCK_N; ++i) {
            acc[i] = (c.n0 + i < c.N) ? acc[i] * sc : -INFINITY;
            mx = fmaxf(mx, acc[i]);
        }
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
        for (int i = 0; i < BLOCK_N; i += 4) {
            s0 += fast_exp2(acc[i] - mx);
            s1 += fast_exp2(acc[i + 1] - mx);
            s2 += fast_exp2(acc[i + 2] - mx);
            s3 += fast_exp2(acc[i + 3] - mx);
        }
        const int row = c.m0 + c.tid;
        if (row < c.M)
            p.part[(long long)c.n_blk * p.ldp + row] = (mx + log2f((s0 + s1) + (s2 + s3))) * LN2;
    }
};

// ---- softmax-gradient tile (MIL-NCE backward; autograd of loss.py:11-18) ----
// G[r, n] = 2^13 * ( e^{s - den[rs]} + e^{s - den[cs]} - [rs == cs] e^{s - nom[rs]} ),
// s = scale * acc, rs = (row_off + r) / rdiv, cs = n / cdiv the sample indices of the
// row and the column.  Written as fp16 (planes == 1) or as hi/lo fp16 planes laid
// out [hi | lo | hi] along the row (planes == 3) for the fp32-accurate path.
constexpr float G_SCALE = 8192.f;
struct EpiG {
    struct Params {
        const float *den;  // all B entries
        const float *nom;
        __half *out;
        long long ldo;
        long long plane_stride;  // = N (columns per plane)
        int planes;              // 1 or 3
        int row_off, rdiv, cdiv;
        const float *scale_dev;
    };
    template <int BLOCK_N>
    __device__ static void run(const Params &p, const EpiCtx &c, float (&acc)[BLOCK_N]) {
        static_assert(BLOCK_N <= NUM_EPI_THREADS, "one table entry per epilogue thread");
        float *tab = c.scratch + c.par * BLOCK_N;  // den of the column's sample, log2 domain
        if (c.tid < BLOCK_N) {
            const int n = c.n0 + c.tid;
            tab[c.tid] = (n < c.N) ? __ldg(p.den + n / p.cdiv) * LOG2E : INFINITY;
        }
        epi_bar_sync();
        const float sc = __ldg(p.scale_dev) * LOG2E;
        const int row = c.m0 + c.tid;
        const int rs = (p.row_off + min(row, c.M - 1)) / p.rdiv;
        const float dr = __ldg(p.den + rs) * LOG2E;
        const float nr = __ldg(p.nom + rs) * LOG2E;
        // does this tile touch the positives (row sample == column sample)?
        const int rs_lo = (p.row_off + c.m0) / p.rdiv;
        const int rs_hi = (p.row_off + min(c.m0 + BLOCK_M, c.M) - 1) / p.rdiv;
        const int cs_lo = c.n0 / p.cdiv;
        const int cs_hi = (min(c.n0 + BLOCK_N, c.N) - 1) / p.cdiv;
        const bool diag_tile = !(rs_hi < cs_lo || cs_hi < rs_lo);
        if (!diag_tile) {
#pragma unroll
            for (int i = 0; i < BLOCK_N; ++i) {
                const float y = acc[i] * sc;
                acc[i] = (fast_exp2(y - dr) + fast_exp2(y - tab[i])) * G_SCALE;
            }
        } else {
#pragma unroll
            for (int i = 0; i < BLOCK_N; ++i) {
                const float y = acc[i] * sc;
                float g = fast_exp2(y - dr) + fast_exp2(y - tab[i]);
                if ((c.n0 + i) / p.cdiv == rs) g -= fast_exp2(y - nr);  // the positives
                acc[i] = g * G_SCALE;
            }
        }
        if (row >= c.M) return;
#pragma unroll
        for (int ch = 0; ch < BLOCK_N / 32; ++ch) {
            const int col0 = c.n0 + ch * 32;
            if (col0 >= c.N) break;
            __half *dst = p.out + (long long)row * p.ldo + col0;
            const bool full = (col0 + 32 <= c.N);
            if (p.planes == 1) {
                if (full && ((((uintptr_t)dst) & 15) == 0)) {
#pragma unroll
                    for (int i = 0; i < 32; i += 8) {
                        const int e = ch * 32 + i;
                        __half2 h0 = __floats2half2_rn(acc[e], acc[e + 1]);
                        __half2 h1 = __floats2half2_rn(acc[e + 2], acc[e + 3]);
                        __half2 h2 = __floats2half2_rn(acc[e + 4], acc[e + 5]);
                        __half2 h3 = __floats2half2_rn(acc[e + 6], acc[e + 7]);
                        uint4 o;
                        o.x = *reinterpret_cast<uint32_t *>(&h0);
                        o.y = *reinterpret_cast<uint32_t *>(&h1);
                        o.z = *reinterpret_cast<uint32_t *>(&h2);
                        o.w = *reinterpret_cast<uint32_t *>(&h3);
                        *reinterpret_cast<uint4 *>(dst + i) = o;
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (col0 + i < c.N) dst[i] = __float2half_rn(acc[ch * 32 + i]);
                }
            } else {
                // hi/lo split: g = hi + lo to ~22 bits; planes [hi | lo | hi]
                __half *d_lo = dst + p.plane_stride;
                __half *d_h2 = dst + 2 * p.plane_stride;
                const bool al = full && ((((uintptr_t)dst) & 15) == 0) &&
                                ((((uintptr_t)d_lo) & 15) == 0) && ((((uintptr_t)d_h2) & 15) == 0);
                if (al) {
#pragma unroll
                    for (int i = 0; i < 32; i += 8) {
                        uint32_t hi[4], lo[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float a = acc[ch * 32 + i + 2 * j], b = acc[ch * 32 + i + 2 * j + 1];
                            __half2 h = __floats2half2_rn(a, b);
                            float2 hf = __half22float2(h);
                            __half2 l = __floats2half2_rn(a - hf.x, b - hf.y);
                            hi[j] = *reinterpret_cast<uint32_t *>(&h);
                            lo[j] = *reinterpret_cast<uint32_t *>(&l);
                        }
                        uint4 oh = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                        uint4 ol = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                        *reinterpret_cast<uint4 *>(dst + i) = oh;
                        *reinterpret_cast<uint4 *>(d_lo + i) = ol;
                        *reinterpret_cast<uint4 *>(d_h2 + i) = oh;
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (col0 + i < c.N) {
                            const float a = acc[ch * 32 + i];
                            const __half h = __float2half_rn(a);
                            const __half l = __float2half_rn(a - __half2float(h));
                            dst[i] = h;
                            d_lo[i] = l;
                            d_h2[i] = h;
                        }
                }
            }
        }
    }
};

// ----------------------------------------------------------------------------
// The kernel
// ----------------------------------------------------------------------------
struct GemmShape {
    int M, N, Kd;
    // k-blocks (of 64) accumulated inside the tensor core before the partial sum is
    // moved to fp32 registers and added there with round-to-nearest.  The tensor
    // core's fp32 accumulation truncates, which biases long K loops (measured:
    // 1e-4 relative after 6144 MMA steps); short chunks keep it below 1e-6.
    int chunk_iters;
};

template <int BLOCK_N>
struct SmemLayout {
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
    static constexpr int B_BYTES = BLOCK_N * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (BLOCK_N <= 128) ? 6 : 4;
    static constexpr int TAIL_BYTES = 4096;  // barriers, tmem slot, epilogue scratch
    static constexpr int TOTAL = STAGES * STAGE_BYTES + TAIL_BYTES + 1024 /* alignment slack */;
    static_assert((2 * STAGES + 4) * 8 + 16 + 2 * BLOCK_N * 4 <= TAIL_BYTES, "tail");
};

template <int BLOCK_N, class Epi>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tn_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const GemmShape shape, const typename Epi::Params ep) {
    using L = SmemLayout<BLOCK_N>;
    constexpr int STAGES = L::STAGES;
    constexpr uint32_t TMEM_COLS = (2 * BLOCK_N <= 32) ? 32 : (2 * BLOCK_N <= 64) ? 64
                                   : (2 * BLOCK_N <= 128) ? 128 : (2 * BLOCK_N <= 256) ? 256 : 512;
    static_assert(2 * BLOCK_N <= 512, "two accumulators must fit in TMEM");
    static_assert(BLOCK_N % 32 == 0 && BLOCK_N >= 32 && BLOCK_N <= 256, "BLOCK_N");

    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                ~(uintptr_t)1023);
    uint8_t *tail = smem + STAGES * L::STAGE_BYTES;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tfull_bar = empty_bar + STAGES;  // [2] accumulator ready
    uint64_t *tempty_bar = tfull_bar + 2;      // [2] accumulator drained
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty_bar + 2);
    float *epi_scratch = reinterpret_cast<float *>(tmem_slot + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const int num_m = (shape.M + BLOCK_M - 1) / BLOCK_M;
    const int num_n = (shape.N + BLOCK_N - 1) / BLOCK_N;
    const int num_tiles = num_m * num_n;
    const int k_iters = (shape.Kd + BLOCK_K - 1) / BLOCK_K;
    const int chunk_iters = (shape.chunk_iters > 0 && shape.chunk_iters < k_iters) ? shape.chunk_iters
                                                                                   : k_iters;
    const int n_chunks = (k_iters + chunk_iters - 1) / chunk_iters;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull_bar[b], 1);
            mbar_init(&tempty_bar[b], NUM_EPI_THREADS / 32);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m_blk = t % num_m, n_blk = t / num_m;
                for (int kb = 0; kb < k_iters; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    mbar_expect_tx(&full_bar[stage], L::STAGE_BYTES);
                    uint8_t *sa = smem + stage * L::STAGE_BYTES;
                    tma_load_2d(sa, &map_a, &full_bar[stage], kb * BLOCK_K, m_blk * BLOCK_M);
                    tma_load_2d(sa + L::A_BYTES, &map_b, &full_bar[stage], kb * BLOCK_K,
                                n_blk * BLOCK_N);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (one thread) =====================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_f16(BLOCK_M, BLOCK_N);
            int stage = 0;
            uint32_t phase = 0;
            uint32_t cc = 0;  // accumulator-chunk counter: buffer = cc & 1
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                for (int ck = 0; ck < n_chunks; ++ck, ++cc) {
                    const uint32_t buf = cc & 1;
                    mbar_wait(&tempty_bar[buf], ((cc >> 1) & 1) ^ 1);
                    tcgen05_fence_after();
                    const uint32_t tmem_d = tmem_base + buf * BLOCK_N;
                    const int kb0 = ck * chunk_iters;
                    const int kb1 = min(k_iters, kb0 + chunk_iters);
                    for (int kb = kb0; kb < kb1; ++kb) {
                        mbar_wait(&full_bar[stage], phase);
                        tcgen05_fence_after();
                        const uint32_t sa = smem_u32(smem + stage * L::STAGE_BYTES);
                        const uint64_t adesc = make_smem_desc(sa);
                        const uint64_t bdesc = make_smem_desc(sa + L::A_BYTES);
#pragma unroll
                        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                            // +32 bytes per UMMA_K step inside the swizzled 128-byte row
                            umma_f16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2),
                                     idesc, (uint32_t)((kb > kb0) || (k > 0)));
                        }
                        umma_commit(&empty_bar[stage]);  // frees the smem slot when the MMAs retire
                        if (++stage == STAGES) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                    umma_commit(&tfull_bar[buf]);  // this chunk's partial sum is complete
                }
            }
        }
    } else {
        // ===================== epilogue warps 2..5 =====================
        const int q = warp & 3;  // TMEM lane quarter this warp may access
        EpiCtx c;
        c.M = shape.M;
        c.N = shape.N;
        c.tid = q * 32 + lane;
        c.scratch = epi_scratch;
        uint32_t cc = 0;
        int lt = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++lt) {
            c.m_blk = t % num_m;
            c.n_blk = t / num_m;
            c.m0 = c.m_blk * BLOCK_M;
            c.n0 = c.n_blk * BLOCK_N;
            c.par = lt & 1;
            float acc[BLOCK_N];
            for (int ck = 0; ck < n_chunks; ++ck, ++cc) {
                const uint32_t buf = cc & 1;
                const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * BLOCK_N;
                mbar_wait(&tfull_bar[buf], (cc >> 1) & 1);
                tcgen05_fence_after();
                if (ck == 0) {
#pragma unroll
                    for (int ch = 0; ch < BLOCK_N / 32; ++ch) {
                        float v[32];
                        tmem_ld32(taddr + ch * 32, v);
#pragma unroll
                        for (int i = 0; i < 32; ++i) acc[ch * 32 + i] = v[i];
                    }
                } else {
#pragma unroll
                    for (int ch = 0; ch < BLOCK_N / 32; ++ch) {
                        float v[32];
                        tmem_ld32(taddr + ch * 32, v);
#pragma unroll
                        for (int i = 0; i < 32; ++i) acc[ch * 32 + i] += v[i];
                    }
                }
                tcgen05_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty_bar[buf]);  // TMEM buffer is free again
            }
            Epi::template run<BLOCK_N>(ep, c, acc);
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

}  // namespace tc
