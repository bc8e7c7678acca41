// soft-DTW for short sequences (N, M <= 64) on the 5th-generation tensor cores, one CTA per pair.
//
// Replaces the distance functions (soft_dtw_cuda.py:338-363), the numba wavefront kernels (:34-112)
// and the torch glue around them (:121-175) for the shape BASELINE.json quotes (64 x 64 x 512).
//
// forward (fwd_kernel):
//   TMA (cp.async.bulk.tensor, 3-D map over (d, N, B), 128-byte rows, SWIZZLE_128B) streams x_b and y_b
//   once, 32 features at a time, through a 3-stage ring.  Four "split" warps turn every fp32 value
//   into a tf32 pair in place: hi = x rounded to 10 mantissa bits, lo = x - hi (exact), lo stored
//   right behind hi, so that [x_hi ; x_lo] is ONE 128-row K-major operand and [y_hi ; y_lo] another.
//   One tcgen05.mma.kind::tf32 (M = 128, N = 128, K = 8) per 8 features then yields all four
//   products hi.hi, hi.lo, lo.hi, lo.lo in the four quadrants of a 128 x 128 TMEM accumulator; their
//   sum is the fp32-grade dot product.  The tensor core's own accumulation truncates, so the
//   accumulator is drained into fp32 registers (round-to-nearest adds) every FLUSH chunks; two
//   accumulators alternate so that the MMAs never wait for a drain.
//   Row norms come from the same pass over shared memory.  The distance tile, exp(-D/gamma) as a
//   scaled pair and the anti-diagonal wavefront in the linear domain follow sdtw_fused.cuh (same
//   arithmetic), but all tiles are kept TRANSPOSED ([j][i]): the epilogue thread that owns row i of
//   the accumulator then writes 32 consecutive banks, and the wavefront runs one COLUMN per lane
//   (the recurrence is symmetric in i and j).
//   The per-cell soft-min weights go to global memory column-major ([j][i], float4) for the backward.
// backward (bwd_kernel): see below.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "sdtw_fused.cuh"
#include "tc_gemm.cuh"

namespace sdtc {

#ifdef SDTW_PROFILE
#define SDTC_STAMP(k, who)                                                         \
    do {                                                                           \
        if (blockIdx.x == 0 && threadIdx.x == (who)) fused::g_prof[k] = clock64(); \
    } while (0)
#define SDTC_ACC(var, expr)                \
    do {                                   \
        const long long _t0 = clock64();   \
        expr;                              \
        var += clock64() - _t0;            \
    } while (0)
#define SDTC_ISTAMP(k, cond)                                         \
    do {                                                             \
        if (blockIdx.x == 0 && (cond)) fused::g_prof[k] = clock64(); \
    } while (0)
#else
#define SDTC_ISTAMP(k, cond) \
    do {                     \
    } while (0)
#define SDTC_STAMP(k, who) \
    do {                   \
    } while (0)
#define SDTC_ACC(var, expr) \
    do {                    \
        expr;               \
    } while (0)
#endif

using fused::PVal;
using fused::E_ZERO;
using fused::MB_EMPTY;
using fused::pow2_nonpos;
using fused::sdtw_cell;
using fused::spin_pause;
using fused::lg2_approx;

constexpr int T = 64;            // max sequence length handled here
constexpr int KC = 32;           // features per chunk: 32 fp32 = one 128-byte swizzled row
constexpr int STAGES = 3;
constexpr int FLUSH = 4;         // chunks accumulated inside the tensor core between drains
constexpr int OPER_BYTES = T * KC * 4;          // 8 KB: one operand plane (hi or lo) of x or y
constexpr int STAGE_BYTES = 4 * OPER_BYTES;     // x_hi | x_lo | y_hi | y_lo
constexpr int DS = 64;           // row stride of the DP tiles
constexpr float COS_EPS_F = 1e-8f;
constexpr int THREADS = 320;     // warp 0: TMA, warp 1: MMA + TMEM, warps 2..5: split, warps 6..9: drain

struct FwdTail {
    uint64_t full[STAGES], split[STAGES], empty[STAGES], tfull[2], tempty[2];
    uint32_t tmem_slot, pad_[3];
    float nx[T], ny[T];   // 1 / max(||x_i||, eps), 1 / max(||y_j||, eps)
    unsigned long long mb[2 * T];
};
constexpr int FWD_SMEM = STAGES * STAGE_BYTES + (int)sizeof(FwdTail) + 1024;
// after the GEMM phase the ring is dead and holds the DP tiles (all [j][i], row stride DS):
//   Dt | Wf | We | Pf | Pe, 16 KB each; the exchange scratch of the epilogue lives in Pf | Pe
static_assert(5 * T * DS * 4 <= STAGES * STAGE_BYTES, "DP tiles must fit in the ring");

// kind::tf32 instruction descriptor: tf32 x tf32 -> fp32, K-major A and B unless a_mn / b_mn
__host__ __device__ constexpr uint32_t make_idesc_tf32(int m, int n, int a_mn = 0, int b_mn = 0) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32_lo(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "mov.b64 da, {%1, %5};\n"
        "mov.b64 db, {%2, %5};\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate), "r"(tc::SMEM_DESC_HI)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1,
                                            int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(tc::smem_u32(dst)),
        "l"(map), "r"(tc::smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// poll a mailbox entry (see sdtw_fused.cuh) until it leaves the sentinel; a protocol bug traps instead of
// hanging the GPU
// (deliberately NOT inlined: with the polling loop inside the wavefront loop the compiler put a YIELD into
// every step of the follower warp)
__device__ __noinline__ unsigned long long mb_wait(volatile unsigned long long *slot) {
    unsigned long long v = *slot;
    if (v != MB_EMPTY) return v;
    const long long t0 = clock64();
    while ((v = *slot) == MB_EMPTY) {
        spin_pause();
        if (clock64() - t0 > 8000000000ll) __trap();
    }
    return v;
}
// 32 lanes x 16 consecutive fp32 columns -> 16 registers per thread (lane = row)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        "tcgen05.wait::ld.sync.aligned;\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// bring one box of a 3-D tensor into L2 (no shared-memory destination, no barrier)
__device__ __forceinline__ void tma_prefetch_l2_3d(const CUtensorMap *map, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.prefetch.tensor.3d.L2.global.tile [%0, {%1, %2, %3}];" ::"l"(map), "r"(c0), "r"(c1),
                 "r"(c2)
                 : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// KIND: 0 cosine, 1 negative_dot
template <int KIND>
__global__ void __launch_bounds__(THREADS, 2)
fwd_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_y, int N, int M,
           int d, float gamma, float bandwidth, float *__restrict__ Rout, float *__restrict__ out,
           float4 *__restrict__ cell, float *__restrict__ nx_out, float *__restrict__ ny_out,
           int *__restrict__ overflow, const int *__restrict__ ix, const int *__restrict__ iy) {
    extern __shared__ uint8_t smem_raw[];
    // 1 KB alignment as pointer arithmetic ON the shared array: a round trip through an integer makes the
    // compiler treat every later access as generic (LD/ST instead of LDS/STS), which doubled the time
    // of the latency-bound wavefront loop
    uint8_t *smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    FwdTail &tl = *reinterpret_cast<FwdTail *>(smem + STAGES * STAGE_BYTES);
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nchunks = (d + KC - 1) / KC;
    const int ngroups = (nchunks + FLUSH - 1) / FLUSH;
    // pair b compares sequence bx of the x bank with sequence by of the y bank (identity without index lists)
    const int bx = ix ? ix[b] : b, by = iy ? iy[b] : b;
    SDTC_STAMP(0, 0);

    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&map_x);
        tc::tma_prefetch_desc(&map_y);
        for (int s = 0; s < STAGES; ++s) {
            tc::mbar_init(&tl.full[s], 1);
            tc::mbar_init(&tl.split[s], 128);
            tc::mbar_init(&tl.empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            tc::mbar_init(&tl.tfull[a], 1);
            tc::mbar_init(&tl.tempty[a], 128);
        }
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(&tl.tmem_slot, 256);  // two 128-column accumulators
    tc::tcgen05_fence_before();
    __syncthreads();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = tl.tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        int stage = 0;
        uint32_t phase = 0;
        for (int c = 0; c < nchunks; ++c) {
            tc::mbar_wait_warp(&tl.empty[stage], phase ^ 1);
            if (tc::elect_one()) {
                uint8_t *st = smem + stage * STAGE_BYTES;
                tc::mbar_expect_tx(&tl.full[stage], 2 * OPER_BYTES);
                tma_load_3d(st, &map_x, &tl.full[stage], c * KC, 0, bx);
                tma_load_3d(st + 2 * OPER_BYTES, &map_y, &tl.full[stage], c * KC, 0, by);
            }
            __syncwarp();
            if (++stage == STAGES) {
                stage = 0;
                phase ^= 1;
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = make_idesc_tf32(128, 128);
        int stage = 0;
        uint32_t phase = 0;
        const uint32_t smem_base = tc::smem_u32(smem);
        for (int g = 0; g < ngroups; ++g) {
            const uint32_t buf = g & 1;
            tc::mbar_wait_warp(&tl.tempty[buf], ((g >> 1) & 1) ^ 1);
            tc::tcgen05_fence_after();
            const uint32_t tmem_d = tmem_base + buf * 128;
            const int c_end = min(nchunks, (g + 1) * FLUSH);
            for (int c = g * FLUSH; c < c_end; ++c) {
                tc::mbar_wait_warp(&tl.split[stage], phase);
                tc::tcgen05_fence_after();
                const uint32_t sa = smem_base + (uint32_t)(stage * STAGE_BYTES);
                const uint32_t a_lo = tc::smem_desc_lo(sa);                   // [x_hi ; x_lo], 128 rows
                const uint32_t b_lo = tc::smem_desc_lo(sa + 2 * OPER_BYTES);  // [y_hi ; y_lo], 128 rows
                if (tc::elect_one()) {
#pragma unroll
                    for (int k = 0; k < KC / 8; ++k)   // K = 8 tf32 = 32 bytes per instruction
                        umma_tf32_lo(tmem_d, a_lo + k * 2, b_lo + k * 2, idesc, (uint32_t)((c != g * FLUSH) || k != 0));
                    tc::umma_commit(&tl.empty[stage]);
                }
                __syncwarp();
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            if (tc::elect_one()) tc::umma_commit(&tl.tfull[buf]);
            __syncwarp();
        }
    } else if (warp < 6) {
        // ===================== split warps (128 threads): fp32 -> (tf32 hi, tf32 lo) in place =====================
        const int r = tid - 64;                    // 0..63: row r of x, 64..127: row r - 64 of y
        const uint32_t row_off = (uint32_t)((r < 64 ? 0 : 2 * OPER_BYTES) + (r & 63) * 128);
        const uint32_t sw = (uint32_t)(r & 7);
        float nrm = 0.f;
        int stage = 0;
        uint32_t phase = 0;
        long long t_wait = 0;
        (void)t_wait;
        for (int c = 0; c < nchunks; ++c) {
            SDTC_ACC(t_wait, tc::mbar_wait(&tl.full[stage], phase));
            const uint32_t base = tc::smem_u32(smem + stage * STAGE_BYTES) + row_off;
            float4 v[8];
#pragma unroll
            for (int pc = 0; pc < 8; ++pc)   // a different bank group per lane of a phase
                v[pc] = tc::lds128f(base + (((uint32_t)pc ^ sw) << 4));
#pragma unroll
            for (int pc = 0; pc < 8; ++pc) {
                const uint32_t a = base + (((uint32_t)pc ^ sw) << 4);
                uint32_t hi[4], lo[4];
                fused::split_tf32(v[pc].x, hi[0], lo[0]);
                fused::split_tf32(v[pc].y, hi[1], lo[1]);
                fused::split_tf32(v[pc].z, hi[2], lo[2]);
                fused::split_tf32(v[pc].w, hi[3], lo[3]);
                tc::sts128(a, make_uint4(hi[0], hi[1], hi[2], hi[3]));
                tc::sts128(a + OPER_BYTES, make_uint4(lo[0], lo[1], lo[2], lo[3]));
            }
            tc::fence_proxy_async();     // the tensor core reads shared memory through the async proxy
            tc::mbar_arrive(&tl.split[stage]);
            if (KIND == 0) {             // (after the hand-over: off the pipeline's critical path)
                float n0 = 0.f, n1 = 0.f;
#pragma unroll
                for (int pc = 0; pc < 8; pc += 2) {
                    n0 = fmaf(v[pc].x, v[pc].x, fmaf(v[pc].y, v[pc].y, fmaf(v[pc].z, v[pc].z, fmaf(v[pc].w, v[pc].w, n0))));
                    n1 = fmaf(v[pc + 1].x, v[pc + 1].x,
                              fmaf(v[pc + 1].y, v[pc + 1].y, fmaf(v[pc + 1].z, v[pc + 1].z, fmaf(v[pc + 1].w, v[pc + 1].w, n1))));
                }
                nrm += n0 + n1;
            }
            if (++stage == STAGES) {
                stage = 0;
                phase ^= 1;
            }
        }
        SDTC_STAMP(1, 64);
#ifdef SDTW_PROFILE
        if (blockIdx.x == 0 && tid == 64) fused::g_prof[5] = t_wait;
#endif
        if (KIND == 0) {
            const float nv = sqrtf(nrm);
            const float rn = 1.f / fmaxf(nv, COS_EPS_F);   // torch clamps each norm (soft_dtw_cuda.py:347)
            if (r < 64) {
                tl.nx[r] = rn;
                if (r < N) nx_out[(size_t)b * N + r] = nv;
            } else {
                tl.ny[r - 64] = rn;
                if (r - 64 < M) ny_out[(size_t)b * M + r - 64] = nv;
            }
        }
    } else if (warp < 10) {
        // ===================== drain warps (128 threads): TMEM accumulators -> fp32 running sums =====================
        const int q = warp & 3;                    // TMEM lane quarter this warp may read
        const int arow = q * 32 + lane;            // accumulator row: < 64: x_hi row, >= 64: x_lo row (arow - 64)
        float acc[T];                              // running sum over the drains: columns j = 0..63
#pragma unroll
        for (int j = 0; j < T; ++j) acc[j] = 0.f;
        long long t_drain = 0;
        (void)t_drain;
        for (int g = 0; g < ngroups; ++g) {
            const uint32_t buf = g & 1;
            SDTC_ACC(t_drain, tc::mbar_wait(&tl.tfull[buf], (g >> 1) & 1));
            tc::tcgen05_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * 128;
#pragma unroll
            for (int h = 0; h < 4; ++h) {          // columns 0..63: . y_hi, 64..127: . y_lo
                float v[32];
                tc::tmem_ld32(taddr + h * 32, v);
#pragma unroll
                for (int j = 0; j < 32; ++j) acc[(h & 1) * 32 + j] += v[j];
            }
            tc::tcgen05_fence_before();
            tc::mbar_arrive(&tl.tempty[buf]);
        }
#ifdef SDTW_PROFILE
        if (blockIdx.x == 0 && tid == 192) fused::g_prof[6] = t_drain;
#endif
        // rows < 64 hold (hi.hi + hi.lo), rows >= 64 hold (lo.hi + lo.lo) of x row i = arow & 63: two
        // transposed tiles ([j][i]: 32 consecutive banks per store) that the tile pass below adds up.
        // All MMAs have completed (last tfull), so the ring is dead and can hold them.
        float *dst = reinterpret_cast<float *>(smem + 3 * T * DS * 4) + (arow < 64 ? 0 : T * DS) + (arow & 63);
#pragma unroll
        for (int j = 0; j < T; ++j) dst[j * DS] = acc[j];
    }
    for (int e = tid; e < 2 * T; e += THREADS) tl.mb[e] = MB_EMPTY;
    float *Rb = Rout ? Rout + (size_t)b * (N + 2) * (M + 2) : nullptr;
    if (Rb) {  // border of R as the reference initialises it (soft_dtw_cuda.py:135-136)
        for (int e = tid; e < (N + 2) * (M + 2); e += THREADS) {
            const int rr = e / (M + 2), cc = e % (M + 2);
            if (rr == 0 || cc == 0 || rr == N + 1 || cc == M + 1) Rb[e] = (e == 0) ? 0.f : INFINITY;
        }
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    SDTC_STAMP(2, 0);
    if (warp == 1) {
        tc::tcgen05_fence_after();
        tc::tmem_dealloc(tmem_base, 256);
    }
    // ---- all threads: distance tile and exp(-D/gamma) as mantissa / exponent, transposed ([j][i]).
    // -D log2(e) / gamma is formed in float-float arithmetic (its integer part can reach 2^22 and its
    // fraction is needed to ~1e-7): p + e1 + e2 with p = D c_hi rounded, e1 its exact residual (fma),
    // e2 = D c_lo.
    {
        float *Dt = reinterpret_cast<float *>(smem);
        float *Wf = Dt + T * DS;
        int *We = reinterpret_cast<int *>(Wf + T * DS);
        const float *dotH = reinterpret_cast<const float *>(smem + 3 * T * DS * 4);
        const float *dotL = dotH + T * DS;
        const double cd = 1.4426950408889634 / (double)gamma;  // log2(e) / gamma
        const float c_hi = (float)cd, c_lo = (float)(cd - (double)c_hi);
        bool too_big = false;
#pragma unroll 4
        for (int idx = tid; idx < T * T; idx += THREADS) {
            const int j = idx >> 6, i = idx & 63;
            float v = dotH[idx] + dotL[idx];
            if (KIND == 0) {
                v = expf(1.f - v * (tl.nx[i] * tl.ny[j]));   // nx, ny hold 1 / max(norm, eps)
            } else {
                v = -v;
            }
            const float pp = v * c_hi;
            const float e1 = fmaf(v, c_hi, -pp), e2 = v * c_lo;
            float fl = floorf(-pp);
            float fr = (-pp - fl) - (e1 + e2);      // in (-eps, 1 + eps)
            if (fr < 0.f) { fr += 1.f; fl -= 1.f; }
            if (fr >= 1.f) { fr -= 1.f; fl += 1.f; }
            if (i < N && j < M && !(fabsf(fl) <= (float)fused::WE_LIMIT)) too_big = true;
            const bool pruned = (bandwidth > 0.f) && (fabsf((float)(i - j)) > bandwidth);
            Dt[idx] = v;
            We[idx] = (int)fl;
            Wf[idx] = pruned ? 0.f : fused::ex2_approx(fr);
        }
        if (too_big && overflow != nullptr) *overflow = 1;
    }
    __syncthreads();
    SDTC_STAMP(7, 0);
    // ---- wavefront: two warps, one COLUMN per lane (tiles are [j][i]); lane c handles cell
    // (i = p - c, j = c) at step p.  Warp 1 (columns 32..63) consumes column 31 from warp 0 through a
    // per-step mailbox and runs behind it.  Same recurrence as sdtw_fused.cuh.
    float *Dt = reinterpret_cast<float *>(smem);
    float *Wf = Dt + T * DS;
    int *We = reinterpret_cast<int *>(Wf + T * DS);
    float *Pf = reinterpret_cast<float *>(We + T * DS);
    int *Pe = reinterpret_cast<int *>(Pf + T * DS);
    const int steps = N + M - 1;
    if (warp < 2) {
        const int c = tid;  // column
        const PVal zero = {0.f, E_ZERO};
        PVal left = zero, up_prev = zero;   // `left`: this column's previous cell (i-1, c); `up`: column c-1
        if (c == 0) { up_prev.f = 1.f; up_prev.e = 0; }       // R[0,0] = 0 of the reference: diagonal input of cell (0,0)
        volatile unsigned long long *mb = tl.mb;
        const unsigned ncol = (c < M) ? (unsigned)N : 0u;     // (unsigned)i < ncol  <=>  cell (i, c) exists
        const float *wfp = Wf + c * DS - c;                   // running pointers at row i = p - c
        const int *wep = We + c * DS - c;
        float *pfp = Pf + c * DS - c;
        int *pep = Pe + c * DS - c;
        float wfn = 0.f;
        int wen = 0;
        if (c == 0 && ncol > 0) { wfn = wfp[0]; wen = wep[0]; }
        auto wave = [&](auto second_tag) {
            constexpr bool SECOND = decltype(second_tag)::value;
            // the follower reads the mailbox entry of the NEXT step one step ahead (the leader is
            // normally ahead, so the entry is there and the load is off the dependency chain); only an
            // entry that was still empty is polled at the point of use
            unsigned long long mbn = MB_EMPTY;
            for (int p = 0; p < steps; ++p) {
                const float wf = wfn;
                const int we = wen;
                wfn = 0.f;
                if ((unsigned)(p + 1 - c) < ncol) { wfn = wfp[p + 1]; wen = wep[p + 1]; }
                // ALL lanes read the entry (one broadcast load, warp-uniform value): no divergent region on
                // the chain - a lane-0-only branch here cost the follower ~150 cycles per step
                unsigned long long mbv = mbn;
                if (SECOND) {
                    mbn = mb[p];                                      // for step p + 1
                    if (p >= 1 && mbv == MB_EMPTY) mbv = mb_wait(&mb[p - 1]);   // (warp-uniform) not there yet: poll
                }
                PVal up;  // column c-1 at step p-1
                up.f = __shfl_up_sync(0xffffffffu, left.f, 1);
                up.e = __shfl_up_sync(0xffffffffu, left.e, 1);
                if (lane == 0) {
                    const bool have = SECOND && p >= 1;
                    up.f = have ? __int_as_float((int)(unsigned)(mbv & 0xffffffffull)) : 0.f;
                    up.e = have ? (int)(unsigned)(mbv >> 32) : E_ZERO;
                }
                const int emax = max(up_prev.e, max(up.e, left.e));
                const float sum = up_prev.f * pow2_nonpos(up_prev.e - emax) +
                                  (up.f * pow2_nonpos(up.e - emax) + left.f * pow2_nonpos(left.e - emax));
                const float pn = sum * wf;  // in [1, 12), or 0: unreachable / pruned / outside
                const int bits = __float_as_int(pn);
                PVal nv;
                nv.e = (pn > 0.f) ? emax + we + ((bits >> 23) - 127) : E_ZERO;
                nv.f = (pn > 0.f) ? __int_as_float((bits & 0x007fffff) | 0x3f800000) : 0.f;
                if (!SECOND && lane == 31)
                    mb[p] = ((unsigned long long)(unsigned)nv.e << 32) | (unsigned)__float_as_int(nv.f);
                if ((unsigned)(p - c) < ncol) {
                    pfp[p] = nv.f;
                    pep[p] = nv.e;
                }
                up_prev = up;
                left = nv;
            }
        };
        if (warp == 1) wave(std::true_type{}); else wave(std::false_type{});
        SDTC_STAMP(3, 32);
        SDTC_STAMP(12, 0);
    }
    __syncthreads();
    // ---- all threads: soft-min weights (for the backward), R and the result.  Lanes run along i,
    // so both the [j][i] tiles and the column-major weight array are accessed contiguously.
    {
        const float g_ln2 = gamma * 0.6931471805599453f;
        float4 *cb = cell ? cell + (size_t)b * N * M : nullptr;
        const PVal zero = {0.f, E_ZERO};
#pragma unroll 4
        for (int idx = tid; idx < T * T; idx += THREADS) {
            const int j = idx >> 6, i = idx & 63;
            if (i >= N || j >= M) continue;
            PVal a = zero, bb = zero, c = zero;
            if (i > 0 && j > 0) { a.f = Pf[(j - 1) * DS + i - 1]; a.e = Pe[(j - 1) * DS + i - 1]; }
            if (i == 0 && j == 0) { a.f = 1.f; a.e = 0; }
            if (i > 0) { bb.f = Pf[j * DS + i - 1]; bb.e = Pe[j * DS + i - 1]; }
            if (j > 0) { c.f = Pf[(j - 1) * DS + i]; c.e = Pe[(j - 1) * DS + i]; }
            const bool ok = !((bandwidth > 0.f) && (fabsf((float)(i - j)) > bandwidth));
            float4 w = make_float4(0.f, 0.f, 0.f, Dt[j * DS + i]);
            (void)sdtw_cell(a, bb, c, Wf[j * DS + i], We[j * DS + i], ok, w.x, w.y, w.z);
            if (cb) cb[(size_t)j * N + i] = w;
            if (Rb != nullptr || (i == N - 1 && j == M - 1)) {
                const float pf = Pf[j * DS + i];
                const float Rv = (pf > 0.f) ? -g_ln2 * (lg2_approx(pf) + (float)Pe[j * DS + i]) : INFINITY;
                if (Rb) Rb[(size_t)(i + 1) * (M + 2) + (j + 1)] = Rv;
                if (i == N - 1 && j == M - 1) out[b] = Rv;
            }
        }
    }
    SDTC_STAMP(4, 0);
}


// ----------------------------------------------------------------------------------------------
// backward (bwd_kernel): one CTA per pair.
//   1. the saved soft-min weights of the pair ([j][i] float4, N*M*16 bytes) arrive in shared memory
//      by one bulk copy (cp.async.bulk), so the reverse wavefront
//      E[i,j] = E[i+1,j] w_up + E[i,j+1] w_left + E[i+1,j+1] w_diag  (soft_dtw_cuda.py:79-112)
//      runs without a global access on its 127-step chain: two warps, one COLUMN per lane, a
//      mailbox between them (as in the forward).
//   2. all threads turn E into the coefficient tile A~ (d loss / d <x_i, y_j>, cosine: scaled by
//      the clamped norms) and its row / column sums (the "own vector" coefficients of the cosine
//      distance), and store A~ as a tf32 (hi, lo) pair in the swizzled UMMA layout.
//   3. dX = A~ . Y + own_x (.) X, then dY = A~^T . X + own_y (.) Y, 64 features per item: TMA streams
//      y (pass 0) and x (pass 1) through a 2-stage ring in the space the weights occupied, the split
//      warps make tf32 (hi, lo) pairs in place, and ONE tcgen05.mma.kind::tf32 (M = 128, N = 128) per 8
//      reduction indices multiplies [A~_hi ; A~_lo] by [chunk_hi | chunk_lo].  The streamed chunk is
//      the MN-major B operand exactly as TMA wrote it (32-byte-atom swizzle, the only MN-major layout
//      32-bit operands have); the coefficient tile is the K-major A operand and is turned around
//      between the passes.  Drain warps add the four quadrants (hi and lo rows meet through a warp
//      shuffle), add the own-vector term and write full 128-byte rows.
// ----------------------------------------------------------------------------------------------
constexpr int BSTAGES = 2;
constexpr int BFC = 64;                           // features per item: two 32-wide TMA boxes
constexpr int BSTAGE_BYTES = 4 * OPER_BYTES;      // 32 KB: hi block 0 | hi block 1 | lo block 0 | lo block 1
constexpr int W_BYTES = T * T * 16;               // 64 KB: the weights, then the ring
constexpr int A_BYTES = T * T * 4;                // 16 KB: one plane of the coefficient tile
static_assert(BSTAGES * BSTAGE_BYTES <= W_BYTES, "the ring reuses the weights' space");
// High word of the shared-memory descriptor of an MN-major tf32 operand.  32-bit operands read MN-major
// must use the 128-byte swizzle with 32-byte atoms (layout type 1: 32-byte chunk index XOR (row % 4),
// what TMA writes under CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B); its k-blocks are 4 rows = 512 bytes apart.
constexpr uint32_t DESC_HI_MN32 = (512u >> 4) | (1u << 14) | (1u << 29);
// tcgen05.mma.kind::tf32 with the descriptors' high words given separately (A and B use different layouts)
__device__ __forceinline__ void umma_tf32_ab(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo,
                                             uint32_t b_hi, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "mov.b64 da, {%1, %2};\n"
        "mov.b64 db, {%3, %4};\n"
        "setp.ne.b32 p, %6, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %5, p;\n"
        "}\n" ::"r"(tmem_d),
        "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}

struct BwdTail {
    uint64_t wfull, adone, full[BSTAGES], split[BSTAGES], empty[BSTAGES], tfull[2], tempty[2];
    uint32_t tmem_slot, pad_[3];
    float rnx[T], rny[T];      // 1 / max(norm, eps)
    float ownx[T], owny[T];    // own-vector coefficients
    float rsum[T], csum[T];
    unsigned long long mb[2 * T];
};
constexpr int STG_BYTES = 4 * 32 * 20 * 4;         // 10 KB: one 32 x (16 + 4) float staging tile per drain warp
constexpr int BWD_SMEM = W_BYTES + 2 * A_BYTES + STG_BYTES + (int)sizeof(BwdTail) + 1024;

__device__ __forceinline__ void bulk_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     tc::smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(tc::smem_u32(bar))
                 : "memory");
}
// byte offset of element (row, col) of a 64 x 64 fp32 tile stored as two 32-column slabs of 128-byte rows,
// SWIZZLE_128B (the layout TMA produces and the UMMA descriptors expect)
__device__ __forceinline__ uint32_t swz_off(int row, int col) {
    return (uint32_t)((col >> 5) * (T * 128) + row * 128 + ((((col & 31) >> 2) ^ (row & 7)) << 4) + ((col & 3) << 2));
}
// The 128-row A operand holds the hi plane (0) and the lo plane (1) of the 64-row tile INTERLEAVED in blocks of
// 16 rows: operand row (row / 16) * 32 + plane * 16 + row % 16.  Accumulator row r lives in TMEM lane r, so
// the hi and the lo partial sums of one tile row then sit 16 lanes apart in the SAME warp and meet
// through one shuffle (stacked as [hi ; lo] they sit in different warps and need shared memory and
// barriers: measured 5 k cycles per item).
__device__ __forceinline__ uint32_t swz_off2(int plane, int row, int col) {
    const int r = ((row >> 4) << 5) + (plane << 4) + (row & 15);
    return (uint32_t)((col >> 5) * (2 * T * 128) + r * 128 + ((((col & 31) >> 2) ^ (r & 7)) << 4) + ((col & 3) << 2));
}

template <int KIND>
__global__ void __launch_bounds__(THREADS, 2)
bwd_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_y,
           const float *__restrict__ x, const float *__restrict__ y, const float *__restrict__ grad_out, int N,
           int M, int d, float bandwidth, const float4 *__restrict__ cell, const float *__restrict__ nx_in,
           const float *__restrict__ ny_in, float *__restrict__ Eout, float *__restrict__ dx,
           float *__restrict__ dy, const int *__restrict__ ix, const int *__restrict__ iy) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t *a_hi = smem + W_BYTES, *a_lo = a_hi + A_BYTES;
    float *Et = reinterpret_cast<float *>(a_hi);          // raw E, [j][i], until the coefficient pass
    float *stage = reinterpret_cast<float *>(smem + W_BYTES + 2 * A_BYTES);
    BwdTail &tl = *reinterpret_cast<BwdTail *>(smem + W_BYTES + 2 * A_BYTES + STG_BYTES);
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nchunks = (d + KC - 1) / KC;
    const int steps = N + M - 1;
    const uint32_t wbytes = (uint32_t)(N * M * 16);
    // pair b = (sequence bx of the x bank, sequence by of the y bank); with index lists several pairs may share
    // a sequence, so its gradient is accumulated with atomic adds (the host zeroes dx / dy first)
    const int bx = ix ? ix[b] : b, by = iy ? iy[b] : b;
    const bool accumulate = (ix != nullptr) || (iy != nullptr);
    SDTC_STAMP(8, 0);
#ifdef SDTW_PROFILE
    if (blockIdx.x == 0 && tid == 0) fused::g_prof[21] = 0;
#endif

    if (warp == 0 && lane == 0) {
        tc::tma_prefetch_desc(&map_x);
        tc::tma_prefetch_desc(&map_y);
        tc::mbar_init(&tl.wfull, 1);
        tc::mbar_init(&tl.adone, 1);
        for (int s = 0; s < BSTAGES; ++s) {
            tc::mbar_init(&tl.full[s], 1);
            tc::mbar_init(&tl.split[s], 128);
            tc::mbar_init(&tl.empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            tc::mbar_init(&tl.tfull[a], 1);
            tc::mbar_init(&tl.tempty[a], 128);
        }
        tc::fence_barrier_init();
        // the pair's weights: one contiguous block
        tc::mbar_expect_tx(&tl.wfull, wbytes);
        const uint8_t *src = reinterpret_cast<const uint8_t *>(cell + (size_t)b * N * M);
        for (uint32_t o = 0; o < wbytes; o += 16384u)
            bulk_load(smem + o, src + o, min(16384u, wbytes - o), &tl.wfull);
    }
    if (warp == 1) tc::tmem_alloc(&tl.tmem_slot, 256);   // two 128-column accumulators
    for (int e = tid; e < 2 * T; e += THREADS) tl.mb[e] = MB_EMPTY;
    for (int e = tid; e < T * DS; e += THREADS) Et[e] = 0.f;
    if (tid < T) {
        tl.rsum[tid] = 0.f;
        tl.csum[tid] = 0.f;
        const float nxv = (KIND == 0 && tid < N) ? nx_in[(size_t)b * N + tid] : 1.f;
        const float nyv = (KIND == 0 && tid < M) ? ny_in[(size_t)b * M + tid] : 1.f;
        tl.rnx[tid] = 1.f / fmaxf(nxv, COS_EPS_F);
        tl.rny[tid] = 1.f / fmaxf(nyv, COS_EPS_F);
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    tc::tcgen05_fence_after();
    const uint32_t tmem_base = tl.tmem_slot;

#ifdef SDTC_EARLY_PREFETCH
    // Variant builds only (measured: 0.1165 vs 0.1135 ms per fwd+bwd without it).  HBM is idle while warps 0 and 1
    // run the wavefront and the coefficient pass, so an otherwise idle warp asks for both sequences to be brought
    // into L2 for the streaming phase below - but that phase already runs its own-term prefetch two items ahead,
    // and 67 MB of early prefetches compete with the weights and the gradient stores for L2.
    if (warp == 2) {
        if (tc::elect_one()) {
            const int dlim = min(d, 1024);
            for (int c0 = 0; c0 < dlim; c0 += KC) tma_prefetch_l2_3d(&map_y, c0, 0, by);
            for (int c0 = 0; c0 < dlim; c0 += KC) tma_prefetch_l2_3d(&map_x, c0, 0, bx);
        }
        __syncwarp();
    }
#endif

    // ---- 1. reverse wavefront: lane c = column j; at step p the cell (i = p - c, j = c) ----
    if (warp < 2) {
        tc::mbar_wait(&tl.wfull, 0);
        const int c = tid;
        volatile unsigned long long *mb = tl.mb;
        float own = 0.f;              // E[i+1, c] * w_up(i+1, c): this column's previous step
        float pl = 0.f, pd = 0.f;     // this column's E * w_left, E * w_diag of the previous step
        float nd_prev = 0.f;          // column c+1's E * w_diag from two steps ago
        const unsigned ncol = (c < M) ? (unsigned)N : 0u;       // (unsigned)i < ncol  <=>  cell (i, c) exists
        const uint32_t wbase = tc::smem_u32(smem) + (uint32_t)(c * N - c) * 16u;   // + p * 16: weights of (p - c, c)
        float *erow = Et + c * DS - c;                          // erow[p] = Et[c][p - c]
        const bool corner_ok = !((bandwidth > 0.f) && (fabsf((float)(N - M)) > bandwidth));
        if (c == M - 1 && corner_ok) own = 1.f;                 // E[N-1, M-1] = 1 enters here
        // weights of the step ahead (off the dependency chain); cells that do not exist get zeros
        float4 wn = make_float4(0.f, 0.f, 0.f, 0.f);
        if ((unsigned)(steps - 1 - c) < ncol) wn = tc::lds128f(wbase + (uint32_t)(steps - 1) * 16u);
        auto wave = [&](auto lead_tag) {
            constexpr bool LEAD = decltype(lead_tag)::value;    // warp 1 publishes, warp 0 consumes
            unsigned long long mbn = MB_EMPTY;   // mailbox entry for the next step, read one step ahead by all lanes
            for (int p = steps - 1; p >= 0; --p) {
                const float4 wk = wn;
                wn = make_float4(0.f, 0.f, 0.f, 0.f);
                if ((unsigned)(p - 1 - c) < ncol) wn = tc::lds128f(wbase + (uint32_t)(p - 1) * 16u);
                unsigned long long mbv = mbn;
                if (!LEAD) {
                    mbn = mb[p];                                      // for step p - 1
                    if (p + 1 <= steps - 1 && mbv == MB_EMPTY) mbv = mb_wait(&mb[p + 1]);   // (warp-uniform)
                }
                float nl = __shfl_down_sync(0xffffffffu, pl, 1);   // column c+1, published at step p+1
                float nd = __shfl_down_sync(0xffffffffu, pd, 1);
                if (lane == 31) {
                    const bool have = !LEAD && p + 1 <= steps - 1;
                    nl = have ? __int_as_float((int)(unsigned)(mbv & 0xffffffffull)) : 0.f;
                    nd = have ? __int_as_float((int)(unsigned)(mbv >> 32)) : 0.f;
                }
                const float e = nl + own + nd_prev;
                nd_prev = nd;
                own = e * wk.y;     // w_up(i, c): feeds (i-1, c), the same column one step later
                pl = e * wk.z;      // w_left(i, c): feeds (i, c-1)
                pd = e * wk.x;      // w_diag(i, c): feeds (i-1, c-1)
                if (LEAD && lane == 0) {
                    unsigned long long v = ((unsigned long long)(unsigned)__float_as_int(pd) << 32) |
                                           (unsigned)__float_as_int(pl);
                    if (v == MB_EMPTY) v = 0x7fc000007fc00000ull;   // a NaN pair must not alias the sentinel
                    mb[p] = v;
                }
                if ((unsigned)(p - c) < ncol) erow[p] = e;
            }
        };
        if (warp == 1) wave(std::true_type{}); else wave(std::false_type{});
        SDTC_STAMP(9, 0);
        SDTC_STAMP(13, 32);
    }
    __syncthreads();

    // ---- 2. E -> coefficient tile (registers first: the tile replaces Et in place) ----
    constexpr int PER = (T * T + THREADS - 1) / THREADS;      // 13 cells per thread, all with i = tid & 63
    float av[PER];
    {
        const float go = grad_out[b];
        const int i = tid & 63;
        const float rni = tl.rnx[i];
        float rs = 0.f;
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const int idx = tid + k * THREADS;
            const int j = idx >> 6;
            float a = 0.f, sv = 0.f;
            if (idx < T * T && i < N && j < M) {
                const float e = Et[idx];
                if (Eout) Eout[((size_t)b * N + i) * M + j] = e;
                const float gE = go * e;
                if (KIND == 0) {
                    const float dd = tc::lds128f(tc::smem_u32(smem) + (uint32_t)(j * N + i) * 16u).w;
                    const float a0 = -gE * dd;                                  // d loss / d cos  (D = exp(1 - c))
                    sv = a0 * (1.f - 0.6931471805599453f * lg2_approx(dd));    // times c = 1 - ln D
                    a = a0 * (rni * tl.rny[j]);
                } else {
                    a = -gE;
                }
            }
            av[k] = a;
            if (KIND == 0) {
                rs += sv;
                float cs = sv;                      // a warp holds 32 consecutive i of one j
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) cs += __shfl_xor_sync(0xffffffffu, cs, o);
                if (lane == 0 && idx < T * T) atomicAdd(&tl.csum[j], cs);
            }
        }
        if (KIND == 0) atomicAdd(&tl.rsum[i], rs);
    }
    __syncthreads();   // Et and the weights are dead: A~ takes Et's place, the ring the weights'
    // A~ as a tf32 (hi, lo) pair, K-major with the standard 128-byte swizzle: orientation 0 = rows i,
    // columns j (the A operand of dX = A~ . Y), orientation 1 = rows j, columns i (dY = A~^T . X)
    auto write_a = [&](int orient) {
        const uint32_t ab = tc::smem_u32(a_hi);
        const int i = tid & 63;
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const int idx = tid + k * THREADS;
            if (idx < T * T) {
                uint32_t hi, lo;
                fused::split_tf32(av[k], hi, lo);
                const int j = idx >> 6;
                const int row = orient == 0 ? i : j, col = orient == 0 ? j : i;
                tc::sts32f(ab + swz_off2(0, row, col), __uint_as_float(hi));
                tc::sts32f(ab + swz_off2(1, row, col), __uint_as_float(lo));
            }
        }
        tc::fence_proxy_async();   // written by ordinary stores, read by the tensor core
    };
    write_a(0);
    if (tid < 2 * T) {     // own-vector coefficients: - (sum_q gc c) / (||v|| clamp(||v||)); zero vector -> 0
        const int r = tid & 63;
        const bool isx = tid < T;
        float own = 0.f;
        if (KIND == 0) {
            const float nv = isx ? (r < N ? nx_in[(size_t)b * N + r] : 0.f) : (r < M ? ny_in[(size_t)b * M + r] : 0.f);
            const float sr = isx ? tl.rsum[r] : tl.csum[r];
            own = (nv > 0.f) ? -sr * (isx ? tl.rnx[r] : tl.rny[r]) / nv : 0.f;
        }
        if (isx) tl.ownx[r] = own; else tl.owny[r] = own;
    }
    __syncthreads();
    SDTC_STAMP(10, 0);

    // ---- 3. the gradient GEMMs: pass 0 streams y (dX), pass 1 streams x (dY), 64 features per item.
    // One tcgen05.mma (M = 128, N = 128, K = 8) per 8 reduction indices: A = [A~_hi ; A~_lo] (K-major, 128
    // rows), B = [chunk_hi | chunk_lo] (MN-major, four 32-column blocks), so the four quadrants of the
    // accumulator hold hi.hi, hi.lo, lo.hi, lo.lo.  (Small MMAs cost ~90 cycles each whatever their
    // shape - measured with M = 64, N = 32 - so the work is packed into few large ones.)
    // Items t = 0 .. 2 nitems1 - 1 flow through one 2-stage ring; at the pass boundary all threads meet
    // once to turn the coefficient tile around.
    const int nitems1 = (d + BFC - 1) / BFC;
    const int nitems = 2 * nitems1;
    const int prefetch = min(BSTAGES, nitems1);     // pass-1 items loaded before the boundary
    auto boundary = [&]() {
        named_bar_sync(1, THREADS);   // every pass-0 MMA has completed (the MMA warp waited), every accumulator is drained
        // turn the coefficient tile around in place: (i, j) -> (j, i), both planes (read all, meet, write all;
        // keeping the values in registers through pass 0 instead made the drain loop spill)
        const uint32_t ab = tc::smem_u32(a_hi);
        const int i = tid & 63;
        float th[PER], tlo[PER];
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const int idx = tid + k * THREADS;
            if (idx < T * T) {
                th[k] = tc::lds32f(ab + swz_off2(0, i, idx >> 6));
                tlo[k] = tc::lds32f(ab + swz_off2(1, i, idx >> 6));
            }
        }
        named_bar_sync(1, THREADS);
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            const int idx = tid + k * THREADS;
            if (idx < T * T) {
                tc::sts32f(ab + swz_off2(0, idx >> 6, i), th[k]);
                tc::sts32f(ab + swz_off2(1, idx >> 6, i), tlo[k]);
            }
        }
        tc::fence_proxy_async();
        named_bar_sync(1, THREADS);
    };
    if (warp == 0) {
        int stage = 0;
        uint32_t phase = 0;
        // the own-vector term of item t needs the 64 x 64 tile of the OTHER tensor (x in pass 0, y in pass 1):
        // the drain warps read it from global memory, this warp brings it into L2 two items ahead
        auto prefetch_own = [&](int t) {
            if (KIND != 0 || t >= nitems) return;
            const int c = t < nitems1 ? t : t - nitems1;
            if (tc::elect_one()) {
                const CUtensorMap *mp = t < nitems1 ? &map_x : &map_y;
                tma_prefetch_l2_3d(mp, c * BFC, 0, t < nitems1 ? bx : by);
                if (c * BFC + KC < d) tma_prefetch_l2_3d(mp, c * BFC + KC, 0, t < nitems1 ? bx : by);
            }
            __syncwarp();
        };
        prefetch_own(0);
        prefetch_own(1);
        for (int t = 0; t < nitems; ++t) {
            prefetch_own(t + 2);
            if (t == nitems1 + prefetch) boundary();
#ifdef SDTW_PROFILE
            const long long e0 = clock64();
#endif
            tc::mbar_wait_warp(&tl.empty[stage], phase ^ 1);
#ifdef SDTW_PROFILE
            if (blockIdx.x == 0 && lane == 0) fused::g_prof[21] += clock64() - e0;
#endif
            if (tc::elect_one()) {
                const int c = t < nitems1 ? t : t - nitems1;
                const CUtensorMap *mp = t < nitems1 ? &map_y : &map_x;
                uint8_t *st = smem + stage * BSTAGE_BYTES;
                // a last item may have only its first 32-feature block inside the tensor: the second block then
                // keeps stale data, which only reaches output columns that are never stored
                const bool two = c * BFC + KC < d;
                tc::mbar_expect_tx(&tl.full[stage], two ? 2 * OPER_BYTES : OPER_BYTES);
                const int bb = t < nitems1 ? by : bx;          // pass 0 streams y, pass 1 streams x
                tma_load_3d(st, mp, &tl.full[stage], c * BFC, 0, bb);
                if (two) tma_load_3d(st + OPER_BYTES, mp, &tl.full[stage], c * BFC + KC, 0, bb);
                SDTC_ISTAMP(22, t == 5);
            }
            __syncwarp();
            if (++stage == BSTAGES) {
                stage = 0;
                phase ^= 1;
            }
        }
        if (nitems <= nitems1 + prefetch) boundary();
    } else if (warp == 1) {
        constexpr uint32_t idesc = make_idesc_tf32(128, 128, 0, 1);   // A K-major, B (the streamed chunk) MN-major
        const uint32_t ab = tc::smem_u32(a_hi), ring = tc::smem_u32(smem);
        int stage = 0;
        uint32_t phase = 0;
        for (int t = 0; t < nitems; ++t) {
            if (t == nitems1) {
                // all pass-0 MMAs must have completed before the tile they read is rewritten
                if (tc::elect_one()) tc::umma_commit(&tl.adone);
                __syncwarp();
                tc::mbar_wait_warp(&tl.adone, 0);
                boundary();
            }
            const uint32_t buf = t & 1;
            tc::mbar_wait_warp(&tl.tempty[buf], ((t >> 1) & 1) ^ 1);
            SDTC_ISTAMP(31, t == 5 && lane == 0);
            tc::mbar_wait_warp(&tl.split[stage], phase);
            SDTC_ISTAMP(25, t == 5 && lane == 0);
            tc::tcgen05_fence_after();
            const uint32_t sb = ring + (uint32_t)(stage * BSTAGE_BYTES);
            const uint32_t td = tmem_base + buf * 128;
            if (tc::elect_one()) {
#pragma unroll
                for (int k = 0; k < T / 8; ++k) {
                    // A: slab k / 4 (32 reduction indices, 128 rows of 128 bytes), 32 bytes per step
                    // B: the k-th group of 8 rows in each of the four blocks (LBO = 8 KB between blocks; a real
                    //    LBO is required: with a dummy value the layout-type-1 descriptor faults)
                    const uint32_t a_d = tc::smem_desc_lo(ab + (uint32_t)((k >> 2) * (2 * T * 128) + (k & 3) * 32));
                    const uint32_t b_d = tc::smem_desc_mn_lo(sb + k * 1024, OPER_BYTES);
#ifndef SDTC_EXP_SKIP_BWD_MMA   // experiment builds: no MMAs (wrong results, timing only)
                    umma_tf32_ab(td, a_d, tc::SMEM_DESC_HI, b_d, DESC_HI_MN32, idesc, (uint32_t)(k != 0));
#endif
                }
                tc::umma_commit(&tl.empty[stage]);
                tc::umma_commit(&tl.tfull[buf]);
            }
            __syncwarp();
            SDTC_ISTAMP(26, t == 5 && lane == 0);
#ifdef SDTW_PROFILE
            if (t == 5) {   // profiling only: when have the MMAs of this item actually completed?
                tc::mbar_wait_warp(&tl.tfull[buf], (t >> 1) & 1);
                SDTC_ISTAMP(14, lane == 0);
            }
#endif
            if (++stage == BSTAGES) {
                stage = 0;
                phase ^= 1;
            }
        }
    } else if (warp < 6) {
        // split warps: in-place (hi, lo) split of the streamed chunk; thread = (row, 32-feature block)
        const int r = (tid - 64) & 63, blk = (tid - 64) >> 6;
        const uint32_t row_off = (uint32_t)(blk * OPER_BYTES + r * 128);
        const uint32_t sw = (uint32_t)(r & 7);
        int stage = 0;
        uint32_t phase = 0;
        long long s_wait = 0;
        (void)s_wait;
        for (int t = 0; t < nitems; ++t) {
            if (t == nitems1) boundary();
            SDTC_ACC(s_wait, tc::mbar_wait(&tl.full[stage], phase));
            SDTC_ISTAMP(23, t == 5 && tid == 64);
            const uint32_t base = tc::smem_u32(smem + stage * BSTAGE_BYTES) + row_off;
            float4 v[8];
#pragma unroll
            for (int pc = 0; pc < 8; ++pc) v[pc] = tc::lds128f(base + (((uint32_t)pc ^ sw) << 4));
#pragma unroll
            for (int pc = 0; pc < 8; ++pc) {
                const uint32_t a = base + (((uint32_t)pc ^ sw) << 4);
                uint32_t hi[4], lo[4];
                fused::split_tf32(v[pc].x, hi[0], lo[0]);
                fused::split_tf32(v[pc].y, hi[1], lo[1]);
                fused::split_tf32(v[pc].z, hi[2], lo[2]);
                fused::split_tf32(v[pc].w, hi[3], lo[3]);
                tc::sts128(a, make_uint4(hi[0], hi[1], hi[2], hi[3]));
                tc::sts128(a + 2 * OPER_BYTES, make_uint4(lo[0], lo[1], lo[2], lo[3]));
            }
            SDTC_ISTAMP(30, t == 5 && tid == 64);
            tc::fence_proxy_async();
            tc::mbar_arrive(&tl.split[stage]);
            SDTC_ISTAMP(24, t == 5 && tid == 64);
            if (++stage == BSTAGES) {
                stage = 0;
                phase ^= 1;
            }
        }
#ifdef SDTW_PROFILE
        if (blockIdx.x == 0 && tid == 64) fused::g_prof[20] = s_wait;
#endif
    } else {
        // drain warps.  TMEM lane 32 q + l holds, for tile row m = 16 q + (l & 15), the A~_hi products (l < 16)
        // or the A~_lo products (l >= 16); columns 0..63 are . chunk_hi, 64..127 . chunk_lo.  Lane l and lane
        // l ^ 16 swap partial sums by shuffle: the hi lane ends up with features 0..31 of the item for its
        // row, the lo lane with features 32..63.
        // In step g of an item the hi lane ends up with features 32 g .. 32 g + 15 of its row, the lo lane with
        // 32 g + 16 .. 32 g + 31: together one 128-byte piece of the row.
        // Global memory is accessed through a per-warp staging tile, EIGHT lanes per 128-byte row piece, so a
        // warp instruction touches 4 full lines.  With one lane per row it touched 32 different lines (rows
        // are d floats apart) and the load/store unit, not DRAM, bounded the kernel (measured: 37 k cycles
        // for the loads and as much for the stores, per pair).
        const int q = warp & 3;
        const bool hi_lane = lane < 16;
        float *stg = stage + (warp - 6) * (32 * 20);           // [32 lanes][16 floats + 4 pad]
        const int c8 = lane & 7;                               // float4 column within a 32-feature piece
        const int srow_off = (c8 >= 4 ? 16 : 0);               // second half of the piece: the lo lane's staging row
        for (int t = 0; t < nitems; ++t) {
            if (t == nitems1) boundary();
            const bool px = t < nitems1;
            const int c = px ? t : t - nitems1;
            const uint32_t buf = t & 1;
            const int rows = px ? N : M;
            const float *selfb = px ? x + (size_t)bx * N * d : y + (size_t)by * M * d;
            float *dstb = px ? dx + (size_t)bx * N * d : dy + (size_t)by * M * d;
            const float *ownv = px ? tl.ownx : tl.owny;
            // the own-vector operand of the whole item first (mostly L2 hits): issued before the wait for the MMAs
            float4 sv[8];
#pragma unroll
            for (int g = 0; g < 2; ++g)
#pragma unroll
                for (int rr = 0; rr < 4; ++rr) {
                    const int mr = q * 16 + rr * 4 + (lane >> 3);
                    const int f = c * BFC + g * 32 + c8 * 4;
                    sv[g * 4 + rr] = make_float4(0.f, 0.f, 0.f, 0.f);
#if !defined(SDTC_EXP_NO_GLOBAL) && !defined(SDTC_EXP_NO_LOADS)
                    if (KIND == 0 && mr < rows && f < d) sv[g * 4 + rr] = __ldg(reinterpret_cast<const float4 *>(selfb + (size_t)mr * d + f));
#endif
                }
            tc::mbar_wait(&tl.tfull[buf], (t >> 1) & 1);
            SDTC_ISTAMP(27, t == 5 && tid == 192);
            tc::tcgen05_fence_after();
            const uint32_t ta = tmem_base + ((uint32_t)(q * 32) << 16) + buf * 128;
#pragma unroll
            for (int g = 0; g < 2; ++g) {          // 16 features of each half per step
                float p0[16], p1[16];              // this row's partial sums: features 32 g .. + 15 and 32 g + 16 .. + 31
                {
                    float w[16];
                    tmem_ld16(ta + 64 + g * 32, w);            // . chunk_lo (small terms)
                    tmem_ld16(ta + g * 32, p0);
#pragma unroll
                    for (int u = 0; u < 16; ++u) p0[u] += w[u];
                    tmem_ld16(ta + 64 + g * 32 + 16, w);
                    tmem_ld16(ta + g * 32 + 16, p1);
#pragma unroll
                    for (int u = 0; u < 16; ++u) p1[u] += w[u];
                }
                if (g == 1) {                      // the accumulator has been read completely
                    tc::tcgen05_fence_before();
                    tc::mbar_arrive(&tl.tempty[buf]);
                    SDTC_ISTAMP(28, t == 5 && tid == 192);
                }
#pragma unroll
                for (int u = 0; u < 16; u += 4) {
                    float o[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float recv = __shfl_xor_sync(0xffffffffu, hi_lane ? p1[u + e] : p0[u + e], 16);
                        o[e] = (hi_lane ? p0[u + e] : p1[u + e]) + recv;
                    }
                    *reinterpret_cast<float4 *>(stg + lane * 20 + u) = make_float4(o[0], o[1], o[2], o[3]);
                }
                __syncwarp();
#pragma unroll
                for (int rr = 0; rr < 4; ++rr) {
                    const int tr = rr * 4 + (lane >> 3);               // tile row within this warp's 16
                    const int mr = q * 16 + tr;
                    const int f = c * BFC + g * 32 + c8 * 4;
                    float4 r4 = *reinterpret_cast<const float4 *>(stg + (tr + srow_off) * 20 + (c8 & 3) * 4);
                    if (KIND == 0) {
                        const float own = ownv[mr];
                        const float4 s4 = sv[g * 4 + rr];
                        r4.x = fmaf(own, s4.x, r4.x); r4.y = fmaf(own, s4.y, r4.y);
                        r4.z = fmaf(own, s4.z, r4.z); r4.w = fmaf(own, s4.w, r4.w);
                    }
                    if (mr < rows && f < d) {      // d % 4 == 0: a float4 is all in or all out
#if defined(SDTC_EXP_NO_GLOBAL) || defined(SDTC_EXP_NO_STORES)
                        if (r4.x == 12345.678f)
#endif
                        // streaming store: the gradients are not read again here, x / y should stay in L2
                        if (accumulate) atomicAdd(reinterpret_cast<float4 *>(dstb + (size_t)mr * d + f), r4);
                        else __stcs(reinterpret_cast<float4 *>(dstb + (size_t)mr * d + f), r4);
                    }
                }
                __syncwarp();                      // the staging tile is free again
            }
            SDTC_ISTAMP(29, t == 5 && tid == 192);
        }
    }
    tc::tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc::tcgen05_fence_after();
        tc::tmem_dealloc(tmem_base, 256);
    }
    SDTC_STAMP(11, 0);
}

}  // namespace sdtc
