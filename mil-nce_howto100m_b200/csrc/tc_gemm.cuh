// tcgen05 / TMEM / TMA GEMM core for sm_100a, with pluggable epilogues.
//
//   C[m, n] = sum_k A[m, k] * Bm[n, k]        A (M, Kd), Bm (N, Kd) fp16, K-major
//
// One persistent CTA per SM, 64 + 32 * (epilogue warps) threads:
//   warp 0      TMA producer   (cp.async.bulk.tensor -> 128B-swizzled smem ring)
//   warp 1      TMEM allocator + tcgen05.mma issuer
//               (both run convergent: every lane polls the mbarrier, elect.sync picks the
//               lane that issues - see mbar_wait_warp for why)
//   warps 2..   epilogue, 8 or 16 warps chosen by the functor: tcgen05.ld the fp32
//               accumulator (one TMEM lane = one output row, a 32- or 64-column part per
//               thread) into registers, then run the epilogue functor
// Accumulators are double-buffered in TMEM (2 x BLOCK_N columns) and drained to
// registers per K-chunk, so the epilogue of tile t overlaps the MMAs of tile t+1
// and long K loops are summed in fp32 registers (round-to-nearest) rather than
// inside the tensor core (which truncates).
//
// The loss kernels never write C: their epilogues reduce the tile on the fly
// (row and column log-sum-exp for the MIL-NCE forward, the softmax-gradient tile G for
// the backward).  This replaces reference loss.py:11 (th.matmul) + :13-18.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace tc {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // fp16 elements = 128 bytes = one SWIZZLE_128B row
constexpr int UMMA_K = 16;
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

// Warp-uniform wait: all 32 lanes poll and leave together, so the code after it stays convergent
// (the single-lane form makes ptxas wrap every tcgen05/TMA issue in an elect + R2UR + retry loop,
// ~25 instructions per MMA, and the issuing thread then cannot keep the tensor core fed while the
// epilogue warps compete for its scheduler: measured 5100-6400 instead of 4250 cycles per tile).
__device__ __forceinline__ void mbar_wait_warp(uint64_t *bar, uint32_t parity) {
    if (__all_sync(0xffffffffu, mbar_try_wait(bar, parity))) return;
    uint64_t t0 = global_timer_ns();
    uint32_t spins = 0;
    while (!__all_sync(0xffffffffu, mbar_try_wait(bar, parity))) {
        if ((++spins & 1023u) == 0 && global_timer_ns() - t0 > 4000000000ull) __trap();
    }
}
// true in exactly one lane of a fully active warp
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(pred));
    return pred != 0;
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
// MN-major operand tile (the M/N index is the contiguous one): the tile is stored
// as 64-element-wide slabs, each slab = BLOCK_K rows (K index) of 128 bytes,
// SWIZZLE_128B.  Canonical layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units:
// LBO = distance between slabs (next 64 M/N elements), SBO = distance between
// 8-row K groups (1024 B).  `lbo_bytes` is the slab size.
__device__ __forceinline__ uint64_t make_smem_desc_mn(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// The same descriptors as (low word, high word): the high word is a constant, so stepping through
// K is one 32-bit add on the low word.
constexpr uint32_t SMEM_DESC_HI = (1024u >> 4) | (1u << 14) | (2u << 29);  // SBO, version, SWIZZLE_128B
__device__ __forceinline__ uint32_t smem_desc_lo(uint32_t saddr) { return ((saddr & 0x3FFFFu) >> 4) | (1u << 16); }
__device__ __forceinline__ uint32_t smem_desc_mn_lo(uint32_t saddr, uint32_t lbo_bytes) {
    return ((saddr & 0x3FFFFu) >> 4) | (((lbo_bytes >> 4) & 0x3FFFu) << 16);
}

// kind::f16 instruction descriptor: fp16 x fp16 -> fp32; a_mn / b_mn = 1 select an
// MN-major A / B operand (bits 15 / 16).
__host__ __device__ constexpr uint32_t make_idesc_f16(int m, int n, int a_mn = 0, int b_mn = 0) {
    return (1u << 4) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(n >> 3) << 17) |
           ((uint32_t)(m >> 4) << 24);
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
// the same with the descriptors given as low words (see smem_desc_lo)
__device__ __forceinline__ void umma_f16_lo(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "mov.b64 da, {%1, %5};\n"
        "mov.b64 db, {%2, %5};\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate), "r"(SMEM_DESC_HI)
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

// programmatic dependent launch (see b200::launch_pdl): wait for the previous grid of the stream,
// and allow the next one to be scheduled early.  No-ops under a plain launch.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Development instrumentation (variant builds with -DB200_EPI_STAMP only): clock64 stamps of CTA 0's
// warps for the first 64 tiles, read back with b200head_debug_stamps (profiles/microbench/epi_timeline.py).
#ifdef B200_EPI_STAMP
constexpr int STAMP_KINDS = 3, STAMP_WARPS = 20, STAMP_TILES = 64, STAMP_SLOTS = 16;
__device__ long long g_stamp[STAMP_KINDS * STAMP_WARPS * STAMP_TILES * STAMP_SLOTS];
__device__ __forceinline__ void epi_stamp(int kind, int tile, int slot) {
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0 && tile < STAMP_TILES)
        g_stamp[((kind * STAMP_WARPS + (threadIdx.x >> 5)) * STAMP_TILES + tile) * STAMP_SLOTS + slot] = clock64();
}
#else
__device__ __forceinline__ void epi_stamp(int, int, int) {}
#endif

__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_log2(float x) {  // one MUFU; log2(0) = -inf; |error| ~ 1e-7
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Division of a non-negative 32-bit integer by a runtime constant without the ~20-instruction
// software divide (the epilogues divide row/column indices by K several times per 32 columns).
// Granlund-Montgomery round-up form, exact for every 32-bit numerator and d >= 1.
struct FastDiv {
    uint32_t mul, sh1, sh2;
    int d;
    __device__ __forceinline__ int div(int n) const {
        const uint32_t t = __umulhi(mul, (uint32_t)n);
        return (int)((t + (((uint32_t)n - t) >> sh1)) >> sh2);
    }
};
inline FastDiv make_fastdiv(int d) {
    uint32_t l = 0;
    while ((1ull << l) < (unsigned long long)d) ++l;
    FastDiv f;
    f.d = d;
    f.mul = (uint32_t)((((1ull << l) - (unsigned long long)d) << 32) / (unsigned long long)d + 1ull);
    f.sh1 = l < 1 ? l : 1;
    f.sh2 = l > 0 ? l - 1 : 0;
    return f;
}

// A load the compiler must leave where it is written (asm volatile keeps its order relative to the
// other volatile statements - TMEM loads, barrier waits): used to issue per-tile lookups one tile
// ahead.  A plain __ldg is sunk down to its first use under register pressure, which puts the whole
// L2 round trip back on the critical path.
__device__ __forceinline__ float ldg_pinned(const float *p) {
    float v;
    asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

// explicit shared-space accesses: pointers that travel through EpiCtx are generic to the
// compiler, which then emits generic LD/ST for them
__device__ __forceinline__ void sts128(uint32_t addr, const uint4 &v) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "r"(addr)
                 : "memory");
    return v;
}
__device__ __forceinline__ void sts32f(uint32_t addr, float v) {
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ float lds32f(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ float4 lds128f(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "r"(addr)
                 : "memory");
    return v;
}

// ----------------------------------------------------------------------------
// Epilogue context handed to the functors
// ----------------------------------------------------------------------------
// 4 * PARTS epilogue warps: warp (q, h) owns TMEM lanes [32q, 32q+32) (= 32 output rows)
// and column part h of each 128-wide block, i.e. HN = 128 / PARTS columns per thread.
// The epilogues are latency-bound straight-line code (ncu: "wait" and "long scoreboard"
// stalls with 2 warps per scheduler), so more, shorter warps finish a tile sooner.
// Each functor picks its own PARTS: the gradient-tile epilogue (EpiG) is fastest with 16 warps
// (353 -> 295 us), the row+column statistics (warp-wide butterflies, fixed cost per warp) with 8.
struct EpiCtx {
    int m0, n0;        // origin of this thread's sub-tile: row block start, FIRST COLUMN OF ITS HALF
    int m_blk, n_blk;  // tile coordinates
    int half;          // 0 .. PARTS-1: which column part of the 128-wide block
    int M, N;          // problem extent
    int tid;           // row within the tile, 0..127
    int lane, q;       // lane, lane quarter
    float *scratch;    // SCRATCH_FLOATS floats of smem shared by the epilogue warps
    uint8_t *stage;    // Epi::STAGE_BYTES bytes of smem (this warp's slice starts at warp_slot)
    int warp_slot;     // 0 .. 4*PARTS-1
    int sub, nsub;     // column sub-block of the tile this run covers, and how many a tile has (1 or 2)
    int dbg_tile;      // tile counter of this CTA (instrumented builds)
    int slice;         // reduction slice of this work unit (GemmShape::k_splits), else 0
};
constexpr int SCRATCH_FLOATS = 2 * 4 * 128;
// Every functor receives the finished fp32 accumulator (its row, its column half) in
// registers (acc[HN], compile-time indexed): the kernel has already drained TMEM, so
// the tensor core is working on the next tile while the functor runs.

// ---- plain store: C = alpha * acc (dvideo / dtext GEMMs, and the test hook) ----
struct EpiStore {
    static constexpr int STAMP_KIND = 0;
    static constexpr int STAGE_BYTES = 0;
    static constexpr int PARTS = 2;
    struct Params {
        void *C;  // float32, or bfloat16 when out_bf16 != 0
        long long ldc;
        float alpha;
        const float *alpha_dev;      // optional device scalar multiplied into alpha
        const float *alpha_div_dev;  // optional device scalar alpha is divided by
        int out_bf16;
        long long split_stride;      // elements between the output planes of consecutive reduction slices (0: none)
        // Row-segmented output (0: off): rows [q * seg_rows, (q + 1) * seg_rows) are written to seg[q] (row index
        // relative to the segment) instead of C.  seg_rows is a multiple of 128, so a CTA's rows share one
        // segment.  The segments may be buffers in other GPUs' memory (peer-mapped): this is how the partial text
        // gradient of the rows-only scheme is delivered to the ranks that own the rows, from the GEMM itself.
        int seg_rows;
        void *seg[16];
    };
    struct Tile {
        float alpha;
    };
    struct Pre {};
    struct Consts {
        float alpha;
    };
    // kernel-lifetime scalars, loaded once per thread before the tile loop
    __device__ static Consts load_consts(const Params &p) {
        Consts k;
        k.alpha = p.alpha;
        if (p.alpha_dev) k.alpha *= __ldg(p.alpha_dev);
        if (p.alpha_div_dev) k.alpha /= __ldg(p.alpha_div_dev);
        return k;
    }
    __device__ static Pre preload(const Params &, const EpiCtx &, int, int, bool) { return Pre{}; }
    // new_row: the tile's row block differs from the previous tile's (per-row state must be rebuilt)
    __device__ static void begin_tile(const Params &, const EpiCtx &, const Pre &, const Consts &k, Tile &t, bool) {
        t.alpha = k.alpha;
    }
    template <int HN>
    __device__ static void run(const Params &p, const EpiCtx &c, const Tile &t, float (&acc)[HN]) {
        const float alpha = t.alpha;
        int row = c.m0 + c.tid;
        if (row >= c.M) return;
        void *cbase = p.C;
        if (p.seg_rows > 0) {
            const int q = c.m0 / p.seg_rows;
            cbase = p.seg[q];
            row -= q * p.seg_rows;
        }
#pragma unroll
        for (int ch = 0; ch < HN / 32; ++ch) {
            const int col0 = c.n0 + ch * 32;
            if (col0 >= c.N) break;
            if (p.out_bf16) {
                __nv_bfloat16 *dst = reinterpret_cast<__nv_bfloat16 *>(cbase) + c.slice * p.split_stride +
                                     (long long)row * p.ldc + col0;
                if (col0 + 32 <= c.N && ((((uintptr_t)dst) & 15) == 0)) {
#pragma unroll
                    for (int i = 0; i < 32; i += 8) {
                        const int e = ch * 32 + i;
                        __nv_bfloat162 h[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            h[u] = __floats2bfloat162_rn(acc[e + 2 * u] * alpha, acc[e + 2 * u + 1] * alpha);
                        *reinterpret_cast<uint4 *>(dst + i) = *reinterpret_cast<uint4 *>(h);
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (col0 + i < c.N) dst[i] = __float2bfloat16_rn(acc[ch * 32 + i] * alpha);
                }
                continue;
            }
            float *dst = reinterpret_cast<float *>(cbase) + c.slice * p.split_stride + (long long)row * p.ldc + col0;
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

// ---- row AND column log-sum-exp of the tile (MIL-NCE forward; loss.py:16-17) ----
// part_row[PARTS*n_blk + part][row] = log sum_{n in part} exp(s[row, n])   s = scale * acc
// part_col[m_blk][col]          = log sum_{m in tile} exp(s[m, col])          (natural logs)
// One exp per element serves both statistics: e = 2^(y - wmax) with wmax the maximum over
// the warp's 32 x HN sub-tile.  Row sums are thread-local; column sums need a reduction
// across the 32 lanes, done with a transposing butterfly (31 shuffles per 32 columns).
// A row or column whose sum underflows under the shared shift (everything in it more than
// 2^100 below wmax) is recomputed exactly with its own maximum - rare, so the common case
// pays nothing for arbitrary logit ranges (embeddings are not normalised).
template <int H, int N>
__device__ __forceinline__ void bfly_step(float (&a)[N], int lane) {
    const bool up = (lane & H) != 0;
#pragma unroll
    for (int j = 0; j < H; ++j) {
        const float send = up ? a[j] : a[j + H];
        const float keep = up ? a[j + H] : a[j];
        a[j] = keep + __shfl_xor_sync(0xffffffffu, send, H);
    }
}
// a[0..N) per lane, N a power of two <= 32: on return a[0] of lane l holds the sum over all 32
// lanes of the original a[l % N] (transposing butterfly, then plain xor-sums over the lanes that
// share l % N).  N - 1 + log2(32 / N) shuffles.
template <int N>
__device__ __forceinline__ float warp_transpose_sum(float (&a)[N], int lane) {
    if constexpr (N >= 32) bfly_step<16, N>(a, lane);
    if constexpr (N >= 16) bfly_step<8, N>(a, lane);
    if constexpr (N >= 8) bfly_step<4, N>(a, lane);
    if constexpr (N >= 4) bfly_step<2, N>(a, lane);
    if constexpr (N >= 2) bfly_step<1, N>(a, lane);
    float v = a[0];
#pragma unroll
    for (int o = N; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
constexpr float TINY_SUM = 7.8886090522101181e-31f;  // 2^-100

// G = column group: the column statistic is wanted per group of G adjacent columns (the K
// candidates of one sample, loss.py:16, when K is a power of two <= 32), so the G exponentials
// are added inside the thread before the cross-lane reduction: 32/G - 1 + log2(G) shuffles per 32
// columns instead of 31.  G = 1 keeps one statistic per column (any K).
//   part_row[PARTS*(n_blk/nsub) + part][row] = log sum_{n in part of the tile} exp(s[row, n])
//   part_col[4*m_blk + q][col / G]           = log sum_{m in 32-row block q of the tile, n in group} exp(s[m, n])
template <int G>
struct EpiRowColT {
    static_assert(G == 1 || G == 2 || G == 4 || G == 8 || G == 16 || G == 32, "group size");
    static constexpr int STAMP_KIND = 1;
    static constexpr int STAGE_BYTES = 0;
#ifndef B200_ROWCOL_PARTS
#define B200_ROWCOL_PARTS 2
#endif
    static constexpr int PARTS = B200_ROWCOL_PARTS;
    static constexpr int NG = 32 / G;  // statistics per 32 columns
    struct Params {
        float *part_row;  // may be null: skip the row statistic
        long long ldr;
        float *part_col;  // may be null: skip the column statistic
        long long ldc;
        const float *scale_dev;
    };
    struct Tile {
        float sc;
        float row_m, row_s;  // row statistic carried across the column sub-blocks of one tile: 2^row_m * row_s
    };
    struct Pre {};
    struct Consts {
        float sc;
    };
    __device__ static Consts load_consts(const Params &p) {
        Consts k;
        k.sc = (p.scale_dev ? __ldg(p.scale_dev) : 1.f) * LOG2E;  // log2 domain; null: operands are unscaled
        return k;
    }
    __device__ static Pre preload(const Params &, const EpiCtx &, int, int, bool) { return Pre{}; }
    __device__ static void begin_tile(const Params &, const EpiCtx &, const Pre &, const Consts &k, Tile &t, bool) {
        t.sc = k.sc;
        t.row_m = -INFINITY;
        t.row_s = 0.f;
    }
    template <int HN>
    __device__ static void run(const Params &p, const EpiCtx &c, Tile &t, float (&acc)[HN]) {
        static_assert(HN * PARTS == 128, "scratch layout assumes 128-wide column blocks");
        const float sc = t.sc;
        const int lane = c.lane;
        const int row = c.m0 + c.tid;
        const bool row_ok = row < c.M;
        const bool want_col = (p.part_col != nullptr);
        if (c.n0 + HN > c.N) {  // ragged right edge (warp-uniform): mask the columns past N
#pragma unroll
            for (int i = 0; i < HN; ++i)
                if (c.n0 + i >= c.N) acc[i] = -INFINITY;
        }
        // (four independent chains: with two or four warps per scheduler a 64-long dependent chain
        // of 4-cycle operations is not hidden)
        float tm[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int i = 0; i < HN; ++i) tm[i & 3] = fmaxf(tm[i & 3], acc[i]);
        float tmax = fmaxf(fmaxf(tm[0], tm[1]), fmaxf(tm[2], tm[3]));
        tmax *= sc;  // sc > 0
        const float wmax = warp_max_f(row_ok ? tmax : -INFINITY);
        float wl[HN / 32];
        float pm = -INFINITY, ps = 0.f;  // this run's row statistic as (shift, sum): log2 domain
        if (wmax > -INFINITY) {          // warp-uniform; false only for fully masked sub-tiles
            // Phase 1, all independent work: the exponentials and the in-thread group sums.
            // Phase 2: the cross-lane butterflies of all 32-column chunks (independent chains).
            // Phase 3: logarithms and ONE vote for the rare exact path.  (A vote + branch after
            // every chunk put a ~450-cycle serial chain - butterfly, log2f, vote - between chunks
            // that two warps per scheduler could not hide.)
            float rsum[4] = {0.f, 0.f, 0.f, 0.f};
            float g[HN / 32][NG];
#pragma unroll
            for (int ch = 0; ch < HN / 32; ++ch) {
#pragma unroll
                for (int u = 0; u < NG; ++u) {
                    float e[G];
#pragma unroll
                    for (int v = 0; v < G; ++v)
                        e[v] = row_ok ? fast_exp2(fmaf(acc[ch * 32 + u * G + v], sc, -wmax)) : 0.f;
#pragma unroll
                    for (int w = G / 2; w > 0; w >>= 1)  // pairwise tree
#pragma unroll
                        for (int v = 0; v < w; ++v) e[v] += e[v + w];
                    g[ch][u] = e[0];
                    rsum[u & 3] += e[0];
                }
            }
            bool bad = false;
            if (want_col) {
                float cs[HN / 32];
#pragma unroll
                for (int ch = 0; ch < HN / 32; ++ch) cs[ch] = warp_transpose_sum<NG>(g[ch], lane);  // lane l: group l % NG
#pragma unroll
                for (int ch = 0; ch < HN / 32; ++ch) {
                    wl[ch] = wmax + fast_log2(cs[ch]);
                    bad |= ((c.n0 + ch * 32 + (lane % NG) * G) < c.N) && cs[ch] < TINY_SUM;
                }
            } else {
#pragma unroll
                for (int ch = 0; ch < HN / 32; ++ch) wl[ch] = -INFINITY;
            }
            if (__any_sync(0xffffffffu, bad)) {
                // exact path: every column with its own maximum, the G columns of a group merged
                // on the fly (uniform across lanes)
#pragma unroll
                for (int ch = 0; ch < HN / 32; ++ch) {
                    float gm = -INFINITY, gsum = 0.f;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const float y = row_ok ? acc[ch * 32 + j] * sc : -INFINITY;
                        const float cm = warp_max_f(y);
                        const float e = (cm > -INFINITY) ? fast_exp2(y - cm) : 0.f;
                        const float s2 = warp_sum_f(e);
                        if (cm > gm) {
                            gsum = gsum * fast_exp2(gm - cm) + s2;  // gm = -inf: 2^-inf = 0
                            gm = cm;
                        } else if (cm > -INFINITY) {
                            gsum += s2 * fast_exp2(cm - gm);
                        }
                        if ((j % G) == G - 1) {
                            if ((lane % NG) == j / G) wl[ch] = (gm > -INFINITY) ? gm + log2f(gsum) : -INFINITY;
                            gm = -INFINITY;
                            gsum = 0.f;
                        }
                    }
                }
            }
            pm = wmax;
            ps = (rsum[0] + rsum[1]) + (rsum[2] + rsum[3]);
            if (ps < TINY_SUM && tmax > -INFINITY) {  // exact per-row path
                float r2 = 0.f;
#pragma unroll
                for (int i = 0; i < HN; ++i) r2 += fast_exp2(fmaf(acc[i], sc, -tmax));
                pm = tmax;
                ps = r2;
            }
        } else {
#pragma unroll
            for (int ch = 0; ch < HN / 32; ++ch) wl[ch] = -INFINITY;
        }
        if (p.part_row != nullptr) {
            // one row partial per (tile, column part): the tile's column sub-blocks are merged as
            // (shift, sum) pairs, one logarithm at the end
            float m = pm, sum = ps;
            if (c.sub > 0) {
                m = fmaxf(t.row_m, pm);
                sum = (m > -INFINITY) ? t.row_s * fast_exp2(t.row_m - m) + ps * fast_exp2(pm - m) : 0.f;
            }
            t.row_m = m;
            t.row_s = sum;
            if (c.sub == c.nsub - 1 && row_ok)
                p.part_row[(long long)((c.n_blk >> (c.nsub - 1)) * PARTS + c.half) * p.ldr + row] =
                    (m + fast_log2(sum)) * LN2;  // sum = 0 -> -inf
        }
        if (want_col && lane < NG) {
            // one column partial per (32-row block, column group): rows 4*m_blk + q of part_col.
            // (merging the four lane quarters of the CTA here needed a barrier across all epilogue
            // warps per run; the merge kernel reads four times as many rows instead)
            constexpr int NGB = 128 / G;  // statistics per 128-column block
#pragma unroll
            for (int ch = 0; ch < HN / 32; ++ch) {
                const int gi = (c.half * HN + ch * 32) / G + lane;  // group within the 128-column block
                if (c.n_blk * 128 + gi * G < c.N)
                    p.part_col[(long long)(c.m_blk * 4 + c.q) * p.ldc + c.n_blk * NGB + gi] = wl[ch] * LN2;
            }
        }
    }
};
using EpiRowCol = EpiRowColT<1>;

// ---- softmax-gradient tile (MIL-NCE backward; autograd of loss.py:11-18) ----
// G[r, n] = 2^13 * ( e^{s - den[rs]} + e^{s - den[cs]} - [rs == cs] e^{s - nom[rs]} ),
// s = scale * acc, rs = (row_off + r) / rdiv, cs = (col_off + n) / cdiv the sample indices
// of the row and the column.  Written as fp16 (planes == 1) or as hi/lo fp16 planes laid
// out [hi | lo | hi] along the row (planes == 3) for the fp32-accurate path.
// Fast path (den range < 60 nats, decided once per call on the device): one exp per
// element, e^{s-den_r} (1 + 2^{den_r - mid} 2^{mid - den_c}); otherwise two exps.
// The fp16 sub-tile of a warp (32 rows x 64 columns) is staged in that warp's 4 KB of
// shared memory (XOR-swizzled 16-byte chunks, conflict-free) and written out as full
// 128-byte row segments; row-per-thread stores straight from registers left half-filled
// 32-byte sectors 64 KB apart and ran the G pass at under 1 TB/s.
constexpr float G_SCALE = 8192.f;
struct GFast {
    const int *flag_dev;
    const float *mid_log2_dev;
};
struct EpiG {
    static constexpr int STAMP_KIND = 2;
    static constexpr int STAGE_BYTES = 4 * 32 * 128 * 2;  // 32 KB: a 128 x 128 fp16 block
#ifndef B200_EPIG_PARTS
#define B200_EPIG_PARTS 4
#endif
    static constexpr int PARTS = B200_EPIG_PARTS;
    struct Params {
        const float *den;  // all B entries
        const float *nom;
        __half *out;
        long long ldo;
        long long plane_stride;  // = N (columns per plane)
        int planes;              // 1 or 3
        int row_off;
        FastDiv rdiv;
        int col_off;
        FastDiv cdiv;
        const float *scale_dev;
        GFast fast;
    };
    // write this warp's 32 x HN fp16 sub-tile (already scaled, packed) to one plane of `out`
    template <int HN>
    __device__ static void store_plane(const EpiCtx &c, uint32_t wstage, const uint32_t (&packed)[HN / 2],
                                       __half *out, long long ldo) {
        constexpr int CH = HN / 8;        // 16-byte chunks per row (8 or 4)
        constexpr int ROWB = HN * 2;      // bytes per staged row (128 or 64)
        constexpr int RPI = 32 / CH;      // rows covered by one write-out instruction (4 or 8)
        // stage: row r = lane; chunk k is stored at a rotated position so that the eight lanes of a
        // shared-memory phase hit different banks (rows 128 B apart: k ^ (r & 7); 64 B: k ^ ((r>>1)&3))
        const int r = c.lane;
        const int rot = (CH == 8) ? (r & 7) : ((r >> 1) & 3);
#pragma unroll
        for (int k = 0; k < CH; ++k)
            sts128(wstage + r * ROWB + ((k ^ rot) << 4),
                   make_uint4(packed[4 * k], packed[4 * k + 1], packed[4 * k + 2], packed[4 * k + 3]));
        __syncwarp();
        // write-out: lane -> (row = it*RPI + lane/CH, chunk = lane%CH): full 16-byte chunks, RPI rows
        const int sub = c.lane / CH, k = c.lane % CH;
        const int col = c.n0 + k * 8;
        const int row0 = c.m0 + c.q * 32;
        __half *dst0 = out + (long long)(row0 + sub) * ldo + col;
        // interior sub-tile with 16-byte aligned rows (warp-uniform): no per-chunk checks
        const bool interior = (row0 + 32 <= c.M) && (c.n0 + HN <= c.N) && ((ldo & 7) == 0) &&
                              ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
        if (interior) {
            const long long step = (long long)RPI * ldo;
#pragma unroll
            for (int it = 0; it < CH; ++it) {
                const int rr = it * RPI + sub;
                const int rrot = (CH == 8) ? (rr & 7) : ((rr >> 1) & 3);
                *reinterpret_cast<uint4 *>(dst0 + it * step) = lds128(wstage + rr * ROWB + ((k ^ rrot) << 4));
            }
        } else {
#pragma unroll 1
            for (int it = 0; it < CH; ++it) {
                const int rr = it * RPI + sub;
                const int grow = row0 + rr;
                const int rrot = (CH == 8) ? (rr & 7) : ((rr >> 1) & 3);
                const uint4 v = lds128(wstage + rr * ROWB + ((k ^ rrot) << 4));
                if (grow < c.M) {
                    __half *dst = out + (long long)grow * ldo + col;
                    if (col + 8 <= c.N && ((((uintptr_t)dst) & 15) == 0)) {
                        *reinterpret_cast<uint4 *>(dst) = v;
                    } else {
                        const __half *hv = reinterpret_cast<const __half *>(&v);
                        for (int e = 0; e < 8; ++e)
                            if (col + e < c.N) dst[e] = hv[e];
                    }
                }
            }
        }
        __syncwarp();
    }
    // per-tile state: everything that depends only on the row block (c.m0, c.tid), loaded once per
    // tile instead of once per 32-column run.  Kept to what the common path needs (the kernel
    // runs at 96 registers per thread); the rare paths re-derive the rest.
    struct Tile {
        bool fast;
        float mid, sc, dr13, ar;
        int n_lo, n_hi;  // columns whose sample is one of this row block's samples: [n_lo, n_hi)
        float cf[4];     // per-lane column entries of the factor table, [sub * HN/32 + h]
    };
    // Raw per-sample values of one tile, loaded one tile AHEAD (preload for tile t+1 is issued
    // before the runs of tile t): the den lookups were the top stall of this epilogue (ncu source
    // page: long-scoreboard waits at the table build and at begin_tile).
    struct Pre {
        float den_r;
        float den_c[4];
    };
    // m0: first row of the 128-row block; n_blk0: first 128-column block of the tile
    // new_row: that tile starts a new row block (else den_r is not needed again)
    __device__ static Pre preload(const Params &p, const EpiCtx &c, int m0, int n_blk0, bool new_row) {
        Pre q;
        const int HN = 128 / PARTS;
        q.den_r = 0.f;
        if (new_row) q.den_r = ldg_pinned(p.den + p.rdiv.div(p.row_off + min(m0 + c.tid, c.M - 1)));
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
            for (int h = 0; h < HN / 32; ++h) {
                const int n = (n_blk0 + s) * 128 + c.half * HN + h * 32 + c.lane;
                q.den_c[s * (HN / 32) + h] =
                    (s < c.nsub && n < c.N) ? ldg_pinned(p.den + p.cdiv.div(p.col_off + n)) : INFINITY;
            }
        return q;
    }
    // (loading these three at every tile start exposed an L2 round trip per tile: ~1000 cycles
    // between a warp's last run and its next wait in the instrumented timeline)
    struct Consts {
        bool fast;
        float mid, sc;
    };
    __device__ static Consts load_consts(const Params &p) {
        Consts k;
        k.fast = __ldg(p.fast.flag_dev) != 0;
        k.mid = __ldg(p.fast.mid_log2_dev);
        k.sc = (p.scale_dev ? __ldg(p.scale_dev) : 1.f) * LOG2E;
        return k;
    }
    __device__ static void begin_tile(const Params &p, const EpiCtx &c, const Pre &q, const Consts &k, Tile &t,
                                      bool new_row) {
        t.fast = k.fast;
        t.mid = k.mid;
        t.sc = k.sc;
        if (new_row) {  // (warp-uniform) everything that depends only on the row block
            const float dr = q.den_r * LOG2E;
            t.dr13 = dr - 13.f;  // folds G_SCALE = 2^13 into the exponent
            t.ar = t.fast ? fast_exp2(dr - t.mid) : 0.f;
            const int rs_lo = p.rdiv.div(p.row_off + c.m0);
            const int rs_hi = p.rdiv.div(p.row_off + min(c.m0 + BLOCK_M, c.M) - 1);
            t.n_lo = rs_lo * p.cdiv.d - p.col_off;
            t.n_hi = (rs_hi + 1) * p.cdiv.d - p.col_off;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float dc = q.den_c[i] * LOG2E;  // masked column: +inf -> factor 0 / exponent -inf
            t.cf[i] = t.fast ? fast_exp2(t.mid - dc) : dc;
        }
    }
    template <int HN>
    __device__ static void run(const Params &p, const EpiCtx &c, const Tile &t, float (&acc)[HN]) {
        static_assert(HN == 64 || HN == 32, "staging layout: 64 or 32 columns per warp");
        static_assert(HN == 128 / PARTS, "preload assumes the kernel's column split");
        const bool fast = t.fast;
        // per-column factors of this warp's columns, in the warp's own slice of smem: no CTA-wide
        // barrier (a shared table made all warps wait for the slowest)
        const uint32_t tab = smem_u32(c.scratch) + (uint32_t)(c.warp_slot * HN * 4);
#pragma unroll
        for (int h = 0; h < HN / 32; ++h)
            sts32f(tab + (uint32_t)((h * 32 + c.lane) * 4), c.sub == 0 ? t.cf[h] : t.cf[HN / 32 + h]);
        __syncwarp();
        epi_stamp(STAMP_KIND, c.dbg_tile, 8 + 2 * c.sub);  // table ready
        const float sc = t.sc, dr13 = t.dr13, ar = t.ar;
        // does this run touch the positives (row sample == column sample)?
        const bool diag_tile = (c.n0 < c.N) && (c.n0 < t.n_hi) && (c.n0 + HN > t.n_lo);
        if (!diag_tile) {
            if (fast) {
#pragma unroll
                for (int i = 0; i < HN; i += 4) {
                    const float4 t4 = lds128f(tab + (uint32_t)(i * 4));
                    const float tv[4] = {t4.x, t4.y, t4.z, t4.w};
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const float e = fast_exp2(fmaf(acc[i + u], sc, -dr13));
                        acc[i + u] = fmaf(e, ar * tv[u], e);
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < HN; i += 4) {
                    const float4 t4 = lds128f(tab + (uint32_t)(i * 4));
                    const float tv[4] = {t4.x, t4.y, t4.z, t4.w};
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const float y = acc[i + u] * sc;
                        acc[i + u] = (fast_exp2(y - dr13) + fast_exp2(y - tv[u] + 13.f));
                    }
                }
            }
        } else {
            const int rs = p.rdiv.div(p.row_off + min(c.m0 + c.tid, c.M - 1));
            const float dr = dr13 + 13.f;
            const float nr = __ldg(p.nom + rs) * LOG2E;
#pragma unroll
            for (int i = 0; i < HN; ++i) {
                const float y = acc[i] * sc;
                const int cs = p.cdiv.div(p.col_off + c.n0 + i);
                const float dc = (c.n0 + i < c.N) ? __ldg(p.den + cs) * LOG2E : INFINITY;
                float g = fast_exp2(y - dr) + fast_exp2(y - dc);
                if (cs == rs) g -= fast_exp2(y - nr);  // the positives
                acc[i] = g * G_SCALE;
            }
        }
        __syncwarp();             // all lanes are done with this run's table
        epi_stamp(STAMP_KIND, c.dbg_tile, 9 + 2 * c.sub);  // math done
        if (c.n0 >= c.N) return;  // warp-uniform: nothing of this half is inside the matrix
        const uint32_t wstage = smem_u32(c.stage) + (uint32_t)(c.warp_slot * (32 * HN * 2));
        uint32_t hi[HN / 2];
#pragma unroll
        for (int i = 0; i < HN / 2; ++i) {
            __half2 h = __floats2half2_rn(acc[2 * i], acc[2 * i + 1]);
            hi[i] = *reinterpret_cast<uint32_t *>(&h);
        }
        store_plane<HN>(c, wstage, hi, p.out, p.ldo);
        if (p.planes == 3) {
            // hi/lo split: g = hi + lo to ~22 bits; planes [hi | lo | hi]
            store_plane<HN>(c, wstage, hi, p.out + 2 * p.plane_stride, p.ldo);
            uint32_t lo[HN / 2];
#pragma unroll
            for (int i = 0; i < HN / 2; ++i) {
                __half2 h = *reinterpret_cast<__half2 *>(&hi[i]);
                float2 hf = __half22float2(h);
                __half2 l = __floats2half2_rn(acc[2 * i] - hf.x, acc[2 * i + 1] - hf.y);
                lo[i] = *reinterpret_cast<uint32_t *>(&l);
            }
            store_plane<HN>(c, wstage, lo, p.out + p.plane_stride, p.ldo);
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
    // MN-major A made of K-planes laid side by side (the fp32-accurate gradient tile matrix,
    // planes [hi | lo | hi] along its columns): k-row kr of the reduction lives at row
    // kr % a_plane_rows, column offset (kr / a_plane_rows) * a_plane_cols.  0 = one plane.
    // (a_plane_rows must be a multiple of BLOCK_K.)
    int a_plane_rows, a_plane_cols;
    // OR-ed into the instruction descriptor (CTA-pair kernels): IDESC_A_BF16 | IDESC_B_BF16 makes the
    // tensor core read both operands as bfloat16 instead of float16.  Both or neither: an instruction
    // with two different operand formats is rejected by the hardware (illegal instruction, measured).
    uint32_t idesc_or;
    // CTA-pair streaming kernel: split the reduction into k_splits slices (0 or 1 = none).  A work unit is
    // (slice, tile); the epilogue context carries the slice so that EpiStore can write each slice's partial
    // result to its own plane (Params::split_stride), to be summed afterwards.  For GEMMs whose tile count is
    // below the number of CTA pairs (the sharded dvideo: n/256 x D/256 tiles against a reduction of B*K).
    int k_splits;
};
constexpr uint32_t IDESC_A_BF16 = 1u << 7, IDESC_B_BF16 = 1u << 10;

// threads per CTA: TMA warp + MMA warp + 4 * PARTS epilogue warps
template <class Epi>
constexpr int num_threads() { return 64 + 4 * Epi::PARTS * 32; }

template <int BLOCK_N, int EPI_STAGE_BYTES, int MB>
struct SmemLayout {
    static constexpr int A_BYTES = MB * BLOCK_M * BLOCK_K * 2;  // MB row blocks of 128
    static constexpr int B_BYTES = BLOCK_N * BLOCK_K * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TAIL_BYTES = 8192;  // barriers, tmem slot, epilogue scratch (SCRATCH_FLOATS)
    static constexpr int MAX_SMEM = 232448;  // 227 KB
    static constexpr int STAGES_FIT = (MAX_SMEM - 1024 - TAIL_BYTES - EPI_STAGE_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_FIT > 6 ? 6 : STAGES_FIT;
    static constexpr int TOTAL = STAGES * STAGE_BYTES + EPI_STAGE_BYTES + TAIL_BYTES + 1024;
    static_assert(STAGES >= 3, "pipeline too shallow");
    static_assert((2 * 6 + 4) * 8 + 16 + SCRATCH_FLOATS * 4 <= TAIL_BYTES, "tail");
};

// A_MN = false: A is (M, Kd) row-major (K-major), map_a box = 64 k x 128 rows.
// A_MN = true : A is given TRANSPOSED, (Kd, M) row-major (MN-major operand), map_a
//               box = 64 m x 64 k-rows, two boxes per 128 rows.  This is how dtext = G^T . V
//               reads G without anybody materialising G^T.
// MB          : row blocks of 128 per tile.  These GEMMs run at the L2 -> SM bandwidth limit
//               (measured 12.3 TB/s = the chip-wide L2 cap with 128x128 tiles at 41 % tensor-pipe
//               activity), so a 256x128 tile - two accumulators sharing one B tile - moves 25 %
//               fewer operand bytes per FLOP.  MB = 2 needs the whole K loop in one chunk.
// N_FAST      : tile order.  false: m fastest (CTAs in flight share B tiles);  true: n fastest
//               (CTAs in flight share the A rows - for skinny N this keeps a huge A from being
//               streamed from HBM once per n-block).
template <int BLOCK_N, class Epi, bool A_MN = false, int MB = 1, bool N_FAST = false>
__global__ void __launch_bounds__(num_threads<Epi>(), 1)
gemm_tn_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const GemmShape shape, const typename Epi::Params ep) {
    using L = SmemLayout<BLOCK_N, Epi::STAGE_BYTES, MB>;
    constexpr int STAGES = L::STAGES;
    constexpr int HN = BLOCK_N / Epi::PARTS;
    constexpr int NUM_EPI_WARPS = 4 * Epi::PARTS;
    constexpr int TILE_M = MB * BLOCK_M;
    constexpr int ACC_COLS = MB * BLOCK_N;  // TMEM columns per accumulator buffer
    constexpr uint32_t TMEM_COLS = (2 * ACC_COLS <= 256) ? 256 : 512;
    static_assert(BLOCK_N == 128, "epilogues are written for 128-wide tiles");
    static_assert(MB == 1 || MB == 2, "one or two row blocks");
    static_assert(2 * ACC_COLS <= 512, "two accumulator buffers must fit in TMEM");

    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                ~(uintptr_t)1023);
    uint8_t *epi_stage = smem + STAGES * L::STAGE_BYTES;
    uint8_t *tail = epi_stage + Epi::STAGE_BYTES;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tfull_bar = empty_bar + STAGES;  // [2] accumulator ready
    uint64_t *tempty_bar = tfull_bar + 2;      // [2] accumulator drained
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty_bar + 2);
    float *epi_scratch = reinterpret_cast<float *>(tmem_slot + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const int num_m = (shape.M + TILE_M - 1) / TILE_M;
    const int num_n = (shape.N + BLOCK_N - 1) / BLOCK_N;
    const int num_tiles = num_m * num_n;
    const int k_iters = (shape.Kd + BLOCK_K - 1) / BLOCK_K;
    const int chunk_iters = (MB == 1 && shape.chunk_iters > 0 && shape.chunk_iters < k_iters)
                                ? shape.chunk_iters : k_iters;
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
            mbar_init(&tempty_bar[b], NUM_EPI_WARPS);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer (convergent warp, one lane issues) =====================
        {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m_blk = N_FAST ? t / num_n : t % num_m;
                const int n_blk = N_FAST ? t % num_n : t / num_m;
                for (int kb = 0; kb < k_iters; ++kb) {
                    mbar_wait_warp(&empty_bar[stage], phase ^ 1);
                    if (elect_one()) {
                        mbar_expect_tx(&full_bar[stage], L::STAGE_BYTES);
                        uint8_t *sa = smem + stage * L::STAGE_BYTES;
                        if (A_MN) {
                            int krow = kb * BLOCK_K, col_off = 0;
                            if (shape.a_plane_rows > 0) {
                                const int pl = krow / shape.a_plane_rows;
                                krow -= pl * shape.a_plane_rows;
                                col_off = pl * shape.a_plane_cols;
                            }
#pragma unroll
                            for (int sl = 0; sl < 2 * MB; ++sl)  // 64-wide slabs of the M index
                                tma_load_2d(sa + sl * (BLOCK_K * 128), &map_a, &full_bar[stage],
                                            col_off + m_blk * TILE_M + sl * 64, krow);
                        } else {
#pragma unroll
                            for (int rb = 0; rb < MB; ++rb)
                                tma_load_2d(sa + rb * (BLOCK_M * BLOCK_K * 2), &map_a, &full_bar[stage],
                                            kb * BLOCK_K, m_blk * TILE_M + rb * BLOCK_M);
                        }
                        tma_load_2d(sa + L::A_BYTES, &map_b, &full_bar[stage], kb * BLOCK_K,
                                    n_blk * BLOCK_N);
                    }
                    __syncwarp();
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (convergent warp, one lane issues; see mbar_wait_warp) ====
        {
            constexpr uint32_t idesc = make_idesc_f16(BLOCK_M, BLOCK_N, A_MN ? 1 : 0);
            constexpr uint32_t A_STEP = A_MN ? (UMMA_K * 128 / 16) : 2;
            int stage = 0;
            uint32_t phase = 0;
            uint32_t cc = 0;  // accumulator-chunk counter: buffer = cc & 1
            const uint32_t smem_base = smem_u32(smem);
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                for (int ck = 0; ck < n_chunks; ++ck, ++cc) {
                    const uint32_t buf = cc & 1;
                    mbar_wait_warp(&tempty_bar[buf], ((cc >> 1) & 1) ^ 1);
                    tcgen05_fence_after();
                    const uint32_t tmem_d = tmem_base + buf * ACC_COLS;
                    const int kb0 = ck * chunk_iters;
                    const int kb1 = min(k_iters, kb0 + chunk_iters);
                    for (int kb = kb0; kb < kb1; ++kb) {
                        mbar_wait_warp(&full_bar[stage], phase);
                        tcgen05_fence_after();
                        const uint32_t sa = smem_base + (uint32_t)(stage * L::STAGE_BYTES);
                        const uint32_t b_lo = smem_desc_lo(sa + L::A_BYTES);
                        if (elect_one()) {
#pragma unroll
                            for (int rb = 0; rb < MB; ++rb) {
                                const uint32_t sarb = sa + rb * (BLOCK_M * BLOCK_K * 2);
                                const uint32_t td = tmem_d + rb * BLOCK_N;
                                // MN-major: slabs of 64 m are 8 KB apart, a UMMA_K step is 16 rows = 2048 B;
                                // K-major: +32 bytes per UMMA_K step inside the swizzled 128-byte row
                                const uint32_t a_lo = A_MN ? smem_desc_mn_lo(sarb, BLOCK_K * 128) : smem_desc_lo(sarb);
                                umma_f16_lo(td, a_lo, b_lo, idesc, (uint32_t)(kb > kb0));
#pragma unroll
                                for (int k = 1; k < BLOCK_K / UMMA_K; ++k)
                                    umma_f16_lo(td, a_lo + k * A_STEP, b_lo + k * 2, idesc, 1u);
                            }
                            umma_commit(&empty_bar[stage]);  // frees the smem slot when the MMAs retire
                        }
                        __syncwarp();
                        if (++stage == STAGES) {
                            stage = 0;
                            phase ^= 1;
                        }
                    }
                    if (elect_one()) umma_commit(&tfull_bar[buf]);  // this chunk's partial sum is complete
                    __syncwarp();
                }
            }
        }
    } else {
        // ===================== epilogue warps 2..9 =====================
        EpiCtx c;
        c.slice = 0;
        c.q = warp & 3;               // TMEM lane quarter this warp may access
        c.half = (warp - 2) >> 2;     // column half
        c.warp_slot = warp - 2;
        c.lane = lane;
        c.M = shape.M;
        c.N = shape.N;
        c.tid = c.q * 32 + lane;
        c.scratch = epi_scratch;
        c.stage = epi_stage;
        c.sub = 0;
        c.nsub = 1;
        uint32_t cc = 0;  // epilogue invocations so far (parity of the smem scratch buffers)
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            const int m_blk = N_FAST ? t / num_n : t % num_m;
            c.n_blk = N_FAST ? t % num_n : t / num_m;
            c.n0 = c.n_blk * BLOCK_N + c.half * HN;
            if (MB == 1) {
                c.m_blk = m_blk;
                c.m0 = m_blk * BLOCK_M;
                float acc[HN];
                for (int ck = 0; ck < n_chunks; ++ck, ++cc) {
                    const uint32_t buf = cc & 1;
                    const uint32_t taddr = tmem_base + ((uint32_t)(c.q * 32) << 16) + buf * ACC_COLS +
                                           (uint32_t)(c.half * HN);
                    mbar_wait(&tfull_bar[buf], (cc >> 1) & 1);
                    tcgen05_fence_after();
                    if (ck == 0) {
#pragma unroll
                        for (int ch = 0; ch < HN / 32; ++ch) {
                            float v[32];
                            tmem_ld32(taddr + ch * 32, v);
#pragma unroll
                            for (int i = 0; i < 32; ++i) acc[ch * 32 + i] = v[i];
                        }
                    } else {
#pragma unroll
                        for (int ch = 0; ch < HN / 32; ++ch) {
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
                typename Epi::Tile tile_state{};
                    Epi::begin_tile(ep, c, Epi::preload(ep, c, c.m0, c.n_blk, true), Epi::load_consts(ep), tile_state, true);
                Epi::template run<HN>(ep, c, tile_state, acc);
            } else {
                // two row blocks, one chunk: drain and process them one after the other
                const uint32_t buf = cc & 1;
                mbar_wait(&tfull_bar[buf], (cc >> 1) & 1);
                tcgen05_fence_after();
#pragma unroll 1
                for (int rb = 0; rb < MB; ++rb) {
                    c.m_blk = m_blk * MB + rb;
                    c.m0 = c.m_blk * BLOCK_M;
                    const uint32_t taddr = tmem_base + ((uint32_t)(c.q * 32) << 16) + buf * ACC_COLS +
                                           (uint32_t)(rb * BLOCK_N + c.half * HN);
                    float acc[HN];
#pragma unroll
                    for (int ch = 0; ch < HN / 32; ++ch) {
                        float v[32];
                        tmem_ld32(taddr + ch * 32, v);
#pragma unroll
                        for (int i = 0; i < 32; ++i) acc[ch * 32 + i] = v[i];
                    }
                    if (rb == MB - 1) {
                        tcgen05_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&tempty_bar[buf]);  // both row blocks drained
                    }
                    typename Epi::Tile tile_state{};
                    Epi::begin_tile(ep, c, Epi::preload(ep, c, c.m0, c.n_blk, true), Epi::load_consts(ep), tile_state, true);
                    Epi::template run<HN>(ep, c, tile_state, acc);
                }
                ++cc;
            }
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
