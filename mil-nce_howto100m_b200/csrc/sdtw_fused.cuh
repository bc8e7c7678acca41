// Fused soft-DTW kernels for short sequences (N, M <= 64), one CTA per pair.
//
// forward : stream x_b, y_b once (cp.async, 3-stage ring of 32-feature chunks),
//           accumulate the 64x64 dot-product tile on the tensor cores with a
//           3xTF32 split (hi.hi + lo.hi + hi.lo, ~2^-21, fp32-grade), turn it into
//           the distance tile in shared memory, then run the anti-diagonal
//           wavefront out of shared memory (no global load on the dependency
//           chain).  Replaces soft_dtw_cuda.py:326-363 + :34-76 + :121-145.
// backward: prefetch the first feature chunks, run the reverse wavefront from the
//           saved soft-min weights, keep the 64x64 coefficient tile in shared
//           memory and apply it to y and x chunk by chunk (3xTF32 again):
//           dX = A~ . Y + own . X, dY = A~^T . X + own . Y, written straight to
//           global.  Replaces soft_dtw_cuda.py:79-112 + :147-175 + autograd of the
//           distance function.
// Both are bound by streaming x and y (256 KB per pair each way) - the DP chain of
// 127 steps overlaps with other resident CTAs (two per SM).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

namespace fused {

#ifdef SDTW_PROFILE
// developer-only phase timestamps (block 0): build with -DSDTW_PROFILE, read with sdtw_debug_read
__device__ long long g_prof[32];
#define SDTW_STAMP(k)                                                  \
    do {                                                               \
        if (blockIdx.x == 0 && threadIdx.x == 0) g_prof[k] = clock64(); \
    } while (0)
#else
#define SDTW_STAMP(k) \
    do {              \
    } while (0)
#endif

constexpr int T = 64;          // max sequence length handled here
constexpr int FK = 32;         // forward feature chunk
constexpr int FS = FK + 4;     // smem row stride (floats): conflict-free fragment loads
constexpr int FSTAGES = 3;
constexpr int BK = 64;         // backward feature chunk
constexpr int BS = BK + 8;     // smem row stride (floats)
constexpr float COS_EPS_F = 1e-8f;
constexpr int DS = 64;         // row stride of the DP tiles (see FwdSmem)
#ifndef SDTW_WRING
#define SDTW_WRING 8
#endif
constexpr int WRING = SDTW_WRING;  // backward: weights are fetched this many steps ahead

// x = hi + lo with hi a tf32 value (10 mantissa bits) rounded to nearest, lo = x - hi exact in fp32.
// Integer add-and-mask instead of cvt.rna.tf32.f32: on sm_100a the cvt expands to a ~5-instruction
// select sequence, and the two converts per element were three quarters of the instructions of
// the MMA loops (ncu source page).  The tensor core reads only the top 19 bits of lo (truncation);
// |lo| <= 2^-11 |x|, so that costs <= 2^-21 |x| per product, the same order as fp32 accumulation.
__device__ __forceinline__ void split_tf32(float x, uint32_t &hi, uint32_t &lo) {
    hi = (__float_as_uint(x) + 0x1000u) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi));
}
// D(16x8) += A(16x8, row) . B(8x8, col), tf32 inputs, fp32 accumulate
__device__ __forceinline__ void mma_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                         uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// c += a . b with both operands split: small terms first
__device__ __forceinline__ void mma_3xtf32(float (&c)[4], const uint32_t (&ah)[4],
                                           const uint32_t (&al)[4], uint32_t bh0, uint32_t bh1,
                                           uint32_t bl0, uint32_t bl1) {
    mma_tf32(c, al[0], al[1], al[2], al[3], bh0, bh1);
    mma_tf32(c, ah[0], ah[1], ah[2], ah[3], bl0, bl1);
    mma_tf32(c, ah[0], ah[1], ah[2], ah[3], bh0, bh1);
}
__device__ __forceinline__ void cp_async16(void *dst, const void *src, int src_bytes) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float lg2_approx(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// rows x cols-float chunk of a (rows_total, d) matrix -> smem [64][stride]; zero fill outside
template <int COLS, int STRIDE>
__device__ __forceinline__ void load_chunk(float *dst, const float *src, int rows_total, int d, int k0,
                                           int tid) {
    constexpr int GR = COLS / 4;  // 16-byte granules per row
#pragma unroll
    for (int h = 0; h < (T * GR) / 256; ++h) {
        const int gid = tid + h * 256;
        const int r = gid / GR, gc = (gid % GR) * 4;
        const bool ok = (r < rows_total) && (k0 + gc < d);
        const float *s = src + (size_t)(ok ? r : 0) * d + (ok ? k0 + gc : 0);
        cp_async16(dst + r * STRIDE + gc, s, ok ? 16 : 0);
    }
}

struct FwdSmem {
    float xs[FSTAGES][T][FS];
    float ys[FSTAGES][T][FS];
    // Row stride DS = 64 on purpose: the wavefront touches element [r][p - r], i.e. word
    // r * (DS - 1) + p, and DS - 1 = 63 is odd, so the 32 lanes of a step hit 32 different
    // banks.  (The usual padding to 65 puts a whole anti-diagonal in ONE bank.)
    float Dt[T][DS];      // distance tile
    float Wf[T][DS];      // exp(-D/gamma) = Wf * 2^We, Wf in [1, 2)
    int We[T][DS];
    float nx[T], ny[T];
    unsigned long long mb[2 * T];  // row 31's (f, e) per step, handed from warp 0 to warp 1
};
// A mailbox entry is one 64-bit shared-memory word written by a single store, so the
// consumer needs no fence (a __threadfence_block here stalls on every outstanding global
// store/prefetch of the producer and tripled the step time): it spins until the entry
// differs from the sentinel.
constexpr unsigned long long MB_EMPTY = 0xffffffffffffffffull;
// ~32 cycles off the shared-memory port between two polls of a mailbox entry
__device__ __forceinline__ void spin_pause() {
    const long long t0 = clock64();
    while (clock64() - t0 < 32) {
    }
}

// ---- scaled linear-domain soft-min recurrence ----
// With P = exp(-R/gamma) the recurrence R = D + softmin_gamma(Ra, Rb, Rc) becomes
// P = exp(-D/gamma) * (Pa + Pb + Pc): three adds and a multiply per cell instead of a
// min, three exps and a log on the 127-step dependency chain.  P spans thousands of
// binades, so every value is carried as (f, e) = f * 2^e with f in [1, 2) and an int
// exponent (f = 0 encodes P = 0, i.e. R = +inf).  fp32 significands lose ~1e-7 per
// step relative, i.e. ~1e-6 * gamma absolute on R; the soft-min weights saved for the
// backward are ratios of neighbouring P and need no large-number subtraction at all.
// The exponent of P = 0 is a sentinel far below every real exponent: the kernels keep real
// exponents inside (-2^29, 2^29) (a cell weight's exponent is bounded by WE_LIMIT, a path has at
// most 127 cells; the host and the tile phase enforce it, see fused::eligible and `overflow`), so
// a real exponent never meets the sentinel and `e - emax` never leaves the int range.
constexpr int E_ZERO = -(1 << 30);
constexpr double WE_LIMIT = 4194304.0;  // 2^22: |floor(-D log2(e) / gamma)| allowed per cell
__device__ __forceinline__ float pow2_nonpos(int k) {  // 2^k for k <= 0, flushing to 0 below 2^-126
    k = max(k, -127);
    return __int_as_float((k + 127) << 23);
}
struct PVal {
    float f;
    int e;
};
// one cell: inputs a (diag), b (up), c (left); w = exp(-D/gamma).  Returns the new value and
// the three soft-min weights (0 for an unreachable cell).
__device__ __forceinline__ PVal sdtw_cell(PVal a, PVal b, PVal c, float wf, int we, bool valid,
                                          float &wa, float &wb, float &wc) {
    const int emax = max(a.e, max(b.e, c.e));
    const float ta = a.f * pow2_nonpos(a.e - emax);
    const float tb = b.f * pow2_nonpos(b.e - emax);
    const float tc = c.f * pow2_nonpos(c.e - emax);
    const float sum = ta + (tb + tc);
    PVal r;
    if (!valid || !(sum > 0.f)) {
        wa = wb = wc = 0.f;
        r.f = 0.f;
        r.e = E_ZERO;
        return r;
    }
    const float pn = sum * wf;  // in [1, 12)
    const int bits = __float_as_int(pn);
    r.e = emax + we + ((bits >> 23) - 127);
    r.f = __int_as_float((bits & 0x007fffff) | 0x3f800000);
    const float inv = __fdividef(1.f, sum);
    wa = ta * inv;
    wb = tb * inv;
    wc = tc * inv;
    return r;
}

// KIND: 0 cosine, 1 negative_dot (the euclidean variant needs the direct (x-y)^2 sum and
// stays on the general path)
template <int KIND>
__global__ void __launch_bounds__(256, 2)
fwd_kernel(const float *__restrict__ x, const float *__restrict__ y, int N, int M, int d, float gamma,
           float bandwidth, float *__restrict__ Rout, float *__restrict__ out,
           float4 *__restrict__ cell, float *__restrict__ nx_out, float *__restrict__ ny_out,
           int *__restrict__ overflow) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    FwdSmem &s = *reinterpret_cast<FwdSmem *>(smem_raw);
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float *xb = x + (size_t)b * N * d;
    const float *yb = y + (size_t)b * M * d;
    const int nchunks = (d + FK - 1) / FK;

    SDTW_STAMP(0);
    // ---- streaming dot-product tile ----
#pragma unroll
    for (int st = 0; st < FSTAGES - 1; ++st) {
        if (st < nchunks) {
            load_chunk<FK, FS>(&s.xs[st][0][0], xb, N, d, st * FK, tid);
            load_chunk<FK, FS>(&s.ys[st][0][0], yb, M, d, st * FK, tid);
        }
        cp_async_commit();
    }
    const int wm = warp >> 1, wn = warp & 1;  // 4 x 2 warps: 16 rows x 32 columns each
    const int g = lane >> 2, t4 = lane & 3;
    float acc[4][4] = {};
    float nrm = 0.f;  // threads 0..63: ||x_row||^2, threads 64..127: ||y_row||^2
    for (int c = 0; c < nchunks; ++c) {
        cp_async_wait<FSTAGES - 2>();
        __syncthreads();
        {   // prefetch chunk c + FSTAGES - 1 into the slot freed by chunk c - 1
            const int pc = c + FSTAGES - 1;
            if (pc < nchunks) {
                load_chunk<FK, FS>(&s.xs[pc % FSTAGES][0][0], xb, N, d, pc * FK, tid);
                load_chunk<FK, FS>(&s.ys[pc % FSTAGES][0][0], yb, M, d, pc * FK, tid);
            }
            cp_async_commit();
        }
        const float(*xs)[FS] = s.xs[c % FSTAGES];
        const float(*ys)[FS] = s.ys[c % FSTAGES];
        if (KIND == 0 && tid < 128) {
            const float(*src)[FS] = (tid < 64) ? xs : ys;
            const int r = tid & 63;
#pragma unroll
            for (int kk = 0; kk < FK; ++kk) {
                const float v = src[r][(kk + r) & (FK - 1)];  // rotated: conflict-free
                nrm = fmaf(v, v, nrm);
            }
        }
#pragma unroll
        for (int k8 = 0; k8 < FK / 8; ++k8) {
            const int kc = k8 * 8 + t4;
            uint32_t ah[4], al[4];
            split_tf32(xs[wm * 16 + g][kc], ah[0], al[0]);
            split_tf32(xs[wm * 16 + g + 8][kc], ah[1], al[1]);
            split_tf32(xs[wm * 16 + g][kc + 4], ah[2], al[2]);
            split_tf32(xs[wm * 16 + g + 8][kc + 4], ah[3], al[3]);
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
                uint32_t bh0, bl0, bh1, bl1;
                split_tf32(ys[wn * 32 + nt * 8 + g][kc], bh0, bl0);
                split_tf32(ys[wn * 32 + nt * 8 + g][kc + 4], bh1, bl1);
                // the tensor core's fp32 accumulation truncates: keep its chains short (3 MMAs
                // from zero) and do the long sum over the features in registers
                float part[4] = {0.f, 0.f, 0.f, 0.f};
                mma_3xtf32(part, ah, al, bh0, bh1, bl0, bl1);
#pragma unroll
                for (int e = 0; e < 4; ++e) acc[nt][e] += part[e];
            }
        }
    }
    cp_async_wait<0>();
    SDTW_STAMP(1);
    if (KIND == 0 && tid < 128) {
        const float nv = sqrtf(nrm);
        if (tid < 64) s.nx[tid] = nv; else s.ny[tid - 64] = nv;
    }
    __syncthreads();
    // distance tile and per-cell weight exp(-D/gamma) as mantissa/exponent
    const double inv_g2 = 1.4426950408889634 / (double)gamma;  // log2(e) / gamma
    bool too_big = false;  // |D| / gamma beyond what the scaled-pair exponents can carry
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int i = wm * 16 + g + ((e >> 1) << 3);
            const int j = wn * 32 + nt * 8 + t4 * 2 + (e & 1);
            float v = acc[nt][e];
            if (KIND == 0) {
                v = v / (fmaxf(s.nx[i], COS_EPS_F) * fmaxf(s.ny[j], COS_EPS_F));
                v = expf(1.f - v);
            } else {
                v = -v;
            }
            s.Dt[i][j] = v;
            const double xx = -(double)v * inv_g2;
            const double fl = floor(xx);
            if (i < N && j < M && !(fabs(fl) <= WE_LIMIT)) too_big = true;  // (also catches NaN)
            s.We[i][j] = (int)fl;
            // a Sakoe-Chiba-pruned cell (soft_dtw_cuda.py:60) gets weight 0: the wavefront then needs no
            // per-step bandwidth test, P = sum * 0 = 0 marks the cell unreachable
            const bool pruned = (bandwidth > 0.f) && (fabsf((float)(i - j)) > bandwidth);
            s.Wf[i][j] = pruned ? 0.f : exp2f((float)(xx - fl));
        }
    if (KIND == 0 && tid < 128) {
        if (tid < 64) { if (tid < N) nx_out[(size_t)b * N + tid] = s.nx[tid]; }
        else { if (tid - 64 < M) ny_out[(size_t)b * M + tid - 64] = s.ny[tid - 64]; }
    }
    // the host then re-runs the batch on the float64 general path (sdtw_forward)
    if (too_big && overflow != nullptr) *overflow = 1;
    for (int e = tid; e < 2 * T; e += 256) s.mb[e] = MB_EMPTY;
    // border of R as the reference initialises it (soft_dtw_cuda.py:135-136)
    float *Rb = Rout ? Rout + (size_t)b * (N + 2) * (M + 2) : nullptr;
    if (Rb) {
        for (int e = tid; e < (N + 2) * (M + 2); e += 256) {
            const int r = e / (M + 2), c = e % (M + 2);
            if (r == 0 || c == 0 || r == N + 1 || c == M + 1) Rb[e] = (e == 0) ? 0.f : INFINITY;
        }
    }
    __syncthreads();

    SDTW_STAMP(2);
    // ---- wavefront: two warps, one row per lane, neighbours by shuffle ----
    // Warp 1 (rows 32..63) consumes row 31 from warp 0 through a per-step mailbox in shared
    // memory and simply runs behind it: no block barrier on the 127-step chain.  The loop
    // body is kept to the recurrence itself (a lone warp issues ~0.2 instructions per cycle,
    // so every instruction in it costs ~5 cycles x 127 steps): it only deposits P = (f, e)
    // in shared memory; weights, R and the global stores are done afterwards by all threads.
    float(*Pf)[DS] = reinterpret_cast<float(*)[DS]>(&s.xs[0][0][0]);  // staging ring is free now
    int(*Pe)[DS] = reinterpret_cast<int(*)[DS]>(&s.ys[0][0][0]);
    static_assert(sizeof(float) * T * DS <= sizeof(float) * FSTAGES * T * FS, "P tile alias");
    const int steps = N + M - 1;
    if (warp < 2) {
        const int r = tid;  // row
        const PVal zero = {0.f, E_ZERO};
        PVal left = zero, up_prev = zero;
        if (r == 0) { up_prev.f = 1.f; up_prev.e = 0; }      // R[0,0] = 0 of the reference: the diagonal input of cell (0,0)
        volatile unsigned long long *mb = s.mb;
        const unsigned mrow = (r < N) ? (unsigned)M : 0u;    // (unsigned)j < mrow  <=>  cell (r, j) exists
        // running pointers at column j = p - r of this lane's row (no index arithmetic in the loop)
        const float *wfp = &s.Wf[r][0] - r;
        const int *wep = &s.We[r][0] - r;
        float *pfp = &Pf[r][0] - r;
        int *pep = &Pe[r][0] - r;
        // operands of the next step are fetched one step ahead (off the dependency chain); a cell that
        // does not exist gets weight 0 and so does a pruned one (see the tile phase)
        float wfn = 0.f;
        int wen = 0;
        if (r == 0 && mrow > 0) { wfn = wfp[0]; wen = wep[0]; }
        auto wave = [&](auto second_tag) {
            constexpr bool SECOND = decltype(second_tag)::value;  // warp 1 consumes row 31, warp 0 publishes it
            for (int p = 0; p < steps; ++p) {
                const float wf = wfn;
                const int we = wen;
                wfn = 0.f;
                if ((unsigned)(p + 1 - r) < mrow) { wfn = wfp[p + 1]; wen = wep[p + 1]; }
                PVal up;  // row r-1 at step p-1
                up.f = __shfl_up_sync(0xffffffffu, left.f, 1);
                up.e = __shfl_up_sync(0xffffffffu, left.e, 1);
                if (lane == 0) {
                    up = zero;
                    if (SECOND && p >= 1) {
                        unsigned long long v;
                        // back off between polls: a tight spin on shared memory starves the producer warp
                        // (measured 3.3x slower steps), a __nanosleep costs ~500 cycles to wake from
                        while ((v = mb[p - 1]) == MB_EMPTY) spin_pause();
                        up.f = __int_as_float((int)(unsigned)(v & 0xffffffffull));
                        up.e = (int)(unsigned)(v >> 32);
                    }
                }
                // the cell (sdtw_cell without the weights): aligned sum of the three inputs times w
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
                if ((unsigned)(p - r) < mrow) {
                    pfp[p] = nv.f;
                    pep[p] = nv.e;
                }
                up_prev = up;
                left = nv;
            }
        };
        if (warp == 1) wave(std::true_type{}); else wave(std::false_type{});
        SDTW_STAMP(3);
    }
    __syncthreads();
    // ---- all threads: soft-min weights (for the backward), R and the result, diagonal by diagonal
    {
        const float g_ln2 = gamma * 0.6931471805599453f;
        float4 *cb = cell ? cell + (size_t)b * N * M : nullptr;   // [j][i]
        const PVal zero = {0.f, E_ZERO};
        const int i = tid & 63;
        for (int p = tid >> 6; p < steps; p += 4) {
            const int j = p - i;
            if (i >= N || j < 0 || j >= M) continue;
            PVal a = zero, bb = zero, c = zero;
            if (i > 0 && j > 0) { a.f = Pf[i - 1][j - 1]; a.e = Pe[i - 1][j - 1]; }
            if (i == 0 && j == 0) { a.f = 1.f; a.e = 0; }
            if (i > 0) { bb.f = Pf[i - 1][j]; bb.e = Pe[i - 1][j]; }
            if (j > 0) { c.f = Pf[i][j - 1]; c.e = Pe[i][j - 1]; }
            const bool ok = !((bandwidth > 0.f) && (fabsf((float)(i - j)) > bandwidth));
            float4 w = make_float4(0.f, 0.f, 0.f, s.Dt[i][j]);
            (void)sdtw_cell(a, bb, c, s.Wf[i][j], s.We[i][j], ok, w.x, w.y, w.z);
            if (cb) cb[(size_t)j * N + i] = w;
            if (Rb != nullptr || (i == N - 1 && j == M - 1)) {
                const float pf = Pf[i][j];
                const float Rv = (pf > 0.f) ? -g_ln2 * (lg2_approx(pf) + (float)Pe[i][j]) : INFINITY;
                if (Rb) Rb[(size_t)(i + 1) * (M + 2) + (j + 1)] = Rv;
                if (i == N - 1 && j == M - 1) out[b] = Rv;
            }
        }
    }
    SDTW_STAMP(4);
}

struct BwdSmem {
    float A[T][BS];             // coefficient tile, pre-scaled
    float xs[2][T][BS];
    float ys[2][T][BS];
    float rsum[T], csum[T];
    float nx[T], ny[T];
    float ownx[T], owny[T], rsx[T], rsy[T];
    unsigned long long mb[2 * T];  // row 32's (E*w_up, E*w_diag) per step, warp 1 -> warp 0
    float4 wring[WRING][T];        // per-row ring of saved weights, filled by cp.async
};

template <int KIND>
__global__ void __launch_bounds__(256, 2)
bwd_kernel(const float *__restrict__ x, const float *__restrict__ y, const float *__restrict__ grad_out,
           int N, int M, int d, float bandwidth, const float4 *__restrict__ cell,
           const float *__restrict__ nx_in, const float *__restrict__ ny_in, float *__restrict__ Eout,
           float *__restrict__ dx, float *__restrict__ dy) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    BwdSmem &s = *reinterpret_cast<BwdSmem *>(smem_raw);
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float *xb = x + (size_t)b * N * d;
    const float *yb = y + (size_t)b * M * d;
    const int nchunks = (d + BK - 1) / BK;

    SDTW_STAMP(8);
    // start streaming the first feature chunk; it lands while the wavefront runs
    load_chunk<BK, BS>(&s.xs[0][0][0], xb, N, d, 0, tid);
    load_chunk<BK, BS>(&s.ys[0][0][0], yb, M, d, 0, tid);
    cp_async_commit();

    for (int e = tid; e < T * BS; e += 256) (&s.A[0][0])[e] = 0.f;
    for (int e = tid; e < 2 * T; e += 256) s.mb[e] = MB_EMPTY;
    if (tid < T) {
        s.rsum[tid] = 0.f;
        s.csum[tid] = 0.f;
        s.nx[tid] = (KIND == 0 && tid < N) ? nx_in[(size_t)b * N + tid] : 1.f;
        s.ny[tid] = (KIND == 0 && tid < M) ? ny_in[(size_t)b * M + tid] : 1.f;
    }
    __syncthreads();

    // ---- reverse wavefront: two warps, one row per lane ----
    // E[i,j] = E[i+1,j] w_up(i+1,j) + E[i,j+1] w_left(i,j+1) + E[i+1,j+1] w_diag(i+1,j+1)
    // Warp 1 (rows 32..63) leads; warp 0 reads row 32's products from a per-step mailbox.
    if (warp < 2) {
        const int r = tid;
        const float4 *cb = cell + (size_t)b * N * M;   // [j][i]
        const int steps = N + M - 1;
        volatile unsigned long long *mb = s.mb;
        float left = 0.f;             // E[r, j+1] * w_left(r, j+1)
        float pu = 0.f, pd = 0.f;     // this row's E*w_up, E*w_diag of the previous step
        float nd_prev = 0.f;          // row r+1's E*w_diag from two steps ago
        // The saved weights come from L2/HBM (~1 us away).  Each lane streams its own row's
        // float4 per step through a small shared-memory ring with cp.async, WRING steps ahead;
        // a lane only ever reads slots it filled itself, so cp.async.wait_group is all the
        // synchronisation needed (register double-buffering exposed the load latency on every
        // step: 1800 cycles per step measured).
        // A lone warp pays ~5-6 cycles per instruction here, 127 times, so the loop body is pared
        // down to the recurrence:
        //  - the saved weights (L2/HBM, ~1 us away) stream through a per-lane shared-memory ring
        //    filled by cp.async WRING steps ahead; the loop is unrolled WRING times so every ring
        //    slot is an immediate offset (the rolled loop spent ~12 instructions per step on slot
        //    arithmetic).  Register prefetch is not an option at this depth: a warp has six
        //    load scoreboards, so eight loads in flight make each wait cover a younger load
        //    (measured 600 cycles per step);
        //  - no validity test on the recurrence: the weights of pruned, unreachable and
        //    out-of-range cells are zero (the forward stores zeros, cp.async zero-fills), so their
        //    E comes out as 0.
        const unsigned mrow = (r < N) ? (unsigned)M : 0u;       // (unsigned)j < mrow  <=>  cell (r, j) exists
        const float4 *src = cb + ((long long)(steps - 1) - r) * N + r;   // weights of cell (r, p - r) of the step being fetched
        const uint32_t ring0 = (uint32_t)__cvta_generic_to_shared(&s.wring[0][r]);
        auto fetch_w = [&](int p, int slot) {                   // slot is a compile-time constant at every call
            const bool okw = (unsigned)(p - r) < mrow;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(ring0 + (uint32_t)(slot * (int)sizeof(float4) * T)),
                         "l"(okw ? src : cb), "r"(okw ? 16 : 0)
                         : "memory");
            cp_async_commit();
            src -= N;
        };
#pragma unroll
        for (int k = 0; k < WRING; ++k) fetch_w(steps - 1 - k, k);
        // E[N-1, M-1] = 1 enters through `left` (nothing else feeds the corner)
        const bool corner_ok = !((bandwidth > 0.f) && (fabsf((float)(N - M)) > bandwidth));
        if (r == N - 1 && corner_ok) left = 1.f;
        auto wave = [&](auto lead_tag) {
            constexpr bool LEAD = decltype(lead_tag)::value;    // warp 1 publishes, warp 0 consumes
            for (int p0 = steps - 1; p0 >= 0; p0 -= WRING) {
                float *arow = &s.A[r][0] + (p0 - r);            // &A[r][j], j = p - r
#pragma unroll
                for (int k = 0; k < WRING; ++k) {
                    const int p = p0 - k;
                    if (p < 0) break;
                    cp_async_wait<WRING - 1>();
                    const float4 wk = s.wring[k][r];
                    fetch_w(p - WRING, k);
                    // row r+1, published at step p+1
                    float nu = __shfl_down_sync(0xffffffffu, pu, 1);
                    float nd = __shfl_down_sync(0xffffffffu, pd, 1);
                    if (lane == 31) {
                        nu = 0.f;
                        nd = 0.f;
                        if (!LEAD && p + 1 <= steps - 1) {
                            unsigned long long v;
                            while ((v = mb[p + 1]) == MB_EMPTY) spin_pause();
                            nu = __int_as_float((int)(unsigned)(v & 0xffffffffull));
                            nd = __int_as_float((int)(unsigned)(v >> 32));
                        }
                    }
                    const float e = nu + left + nd_prev;
                    nd_prev = nd;
                    left = e * wk.z;
                    pu = e * wk.y;
                    pd = e * wk.x;
                    if (LEAD && lane == 0) {
                        // a NaN pair with all bits set would alias the sentinel: canonicalise the payload
                        unsigned long long v = ((unsigned long long)(unsigned)__float_as_int(pd) << 32) |
                                               (unsigned)__float_as_int(pu);
                        if (v == MB_EMPTY) v = 0x7fc000007fc00000ull;
                        mb[p] = v;
                    }
                    if ((unsigned)(p - r) < mrow) arow[-k] = e;  // raw E; turned into gradient coefficients below
                }
            }
        };
        if (warp == 1) wave(std::true_type{}); else wave(std::false_type{});
        SDTW_STAMP(9);
    }
    __syncthreads();
    // ---- all threads: E -> coefficient tile A (and, for cosine, the summand tile S = A * c) ----
    {
        const float4 *cb = cell + (size_t)b * N * M;
        const float go = grad_out[b];
        float(*S)[BS] = s.xs[1];  // stage 1 of the ring is not in flight yet
        const int steps = N + M - 1;
        const int i = tid & 63;
        for (int p = tid >> 6; p < steps; p += 4) {
            const int j = p - i;
            if (i >= N || j < 0 || j >= M) continue;
            const float e = s.A[i][j];
            if (Eout) Eout[((size_t)b * N + i) * M + j] = e;
            const float gE = go * e;
            if (KIND == 0) {
                const float dd = cb[(size_t)j * N + i].w;
                const float a = -gE * dd;                                      // d loss / d cos  (D = exp(1 - c))
                s.A[i][j] = a;
                S[i][j] = a * (1.f - 0.6931471805599453f * lg2_approx(dd));   // times c = 1 - ln D
            } else {
                s.A[i][j] = -gE;
            }
        }
    }
    __syncthreads();
    if (KIND == 0 && tid < 128) {
        const float(*S)[BS] = s.xs[1];
        const int t = tid & 63;
        float acc = 0.f;
        if (tid < 64) {
            if (t < N) for (int j = 0; j < M; ++j) acc += S[t][j];
            s.rsum[t] = acc;
        } else {
            if (t < M) for (int i = 0; i < N; ++i) acc += S[i][t];
            s.csum[t] = acc;
        }
    }
    __syncthreads();
    // per-row scale and "own vector" coefficients (cosine; identity otherwise)
    if (tid < 128) {
        const int r = tid & 63;
        const bool isx = tid < 64;
        float rscale = 1.f, own = 0.f;
        if (KIND == 0) {
            const float rn = isx ? s.nx[r] : s.ny[r];
            const float cn = fmaxf(rn, COS_EPS_F);
            const float sr = isx ? s.rsum[r] : s.csum[r];
            rscale = 1.f / cn;
            own = (rn > 0.f) ? -sr / (rn * cn) : 0.f;
        }
        if (isx) { s.rsx[r] = rscale; s.ownx[r] = own; }
        else { s.rsy[r] = rscale; s.owny[r] = own; }
    }
    __syncthreads();
    if (KIND == 0) {
        for (int e = tid; e < T * T; e += 256) {
            const int i = e >> 6, j = e & 63;
            s.A[i][j] *= s.rsx[i] * s.rsy[j];
        }
        __syncthreads();
    }

    SDTW_STAMP(10);
    // ---- input gradients, 64 features at a time ----
    const int wm = warp >> 1, wn = warp & 1, g = lane >> 2, t4 = lane & 3;
    for (int c = 0; c < nchunks; ++c) {
        if (c + 1 < nchunks) {
            load_chunk<BK, BS>(&s.xs[(c + 1) & 1][0][0], xb, N, d, (c + 1) * BK, tid);
            load_chunk<BK, BS>(&s.ys[(c + 1) & 1][0][0], yb, M, d, (c + 1) * BK, tid);
        }
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();
        const float(*xs)[BS] = s.xs[c & 1];
        const float(*ys)[BS] = s.ys[c & 1];
        float ax[4][4] = {}, ay[4][4] = {};
#pragma unroll
        for (int k8 = 0; k8 < T / 8; ++k8) {
            const int kc = k8 * 8 + t4;
            // dX: A operand = A~[i][j] (m = i, k = j); dY: A operand = A~^T (m = j, k = i)
            uint32_t ah[4], al[4], th[4], tl[4];
            split_tf32(s.A[wm * 16 + g][kc], ah[0], al[0]);
            split_tf32(s.A[wm * 16 + g + 8][kc], ah[1], al[1]);
            split_tf32(s.A[wm * 16 + g][kc + 4], ah[2], al[2]);
            split_tf32(s.A[wm * 16 + g + 8][kc + 4], ah[3], al[3]);
            split_tf32(s.A[kc][wm * 16 + g], th[0], tl[0]);
            split_tf32(s.A[kc][wm * 16 + g + 8], th[1], tl[1]);
            split_tf32(s.A[kc + 4][wm * 16 + g], th[2], tl[2]);
            split_tf32(s.A[kc + 4][wm * 16 + g + 8], th[3], tl[3]);
#pragma unroll
            for (int nt = 0; nt < 4; ++nt) {
                const int n = wn * 32 + nt * 8 + g;
                uint32_t bh0, bl0, bh1, bl1;
                split_tf32(ys[kc][n], bh0, bl0);       // B[k = j][n = feature] = y[j][n]
                split_tf32(ys[kc + 4][n], bh1, bl1);
                mma_3xtf32(ax[nt], ah, al, bh0, bh1, bl0, bl1);
                split_tf32(xs[kc][n], bh0, bl0);       // B[k = i][n = feature] = x[i][n]
                split_tf32(xs[kc + 4][n], bh1, bl1);
                mma_3xtf32(ay[nt], th, tl, bh0, bh1, bl0, bl1);
            }
        }
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const int r = wm * 16 + g + hh * 8;
                const int cl = wn * 32 + nt * 8 + t4 * 2;
                const int gc = c * BK + cl;
                if (gc < d) {
                    if (r < N) {
                        float2 o = make_float2(ax[nt][hh * 2], ax[nt][hh * 2 + 1]);
                        if (KIND == 0) {
                            o.x = fmaf(s.ownx[r], xs[r][cl], o.x);
                            o.y = fmaf(s.ownx[r], xs[r][cl + 1], o.y);
                        }
                        float *dst = dx + ((size_t)b * N + r) * d + gc;
                        if (gc + 1 < d) *reinterpret_cast<float2 *>(dst) = o; else dst[0] = o.x;
                    }
                    if (r < M) {
                        float2 o = make_float2(ay[nt][hh * 2], ay[nt][hh * 2 + 1]);
                        if (KIND == 0) {
                            o.x = fmaf(s.owny[r], ys[r][cl], o.x);
                            o.y = fmaf(s.owny[r], ys[r][cl + 1], o.y);
                        }
                        float *dst = dy + ((size_t)b * M + r) * d + gc;
                        if (gc + 1 < d) *reinterpret_cast<float2 *>(dst) = o; else dst[0] = o.x;
                    }
                }
            }
        __syncthreads();
    }
    cp_async_wait<0>();
    SDTW_STAMP(11);
}

// The general path covers everything else.  gamma: the cosine distance exp(1 - cos) lies in
// [1, e^2], so 127 cells of at most e^2 log2(e) / gamma each stay below 2^29 for gamma >= 3e-6
// (decided here); negative_dot distances are unbounded, so its tile phase checks every cell against
// WE_LIMIT and raises the `overflow` flag that makes sdtw_forward redo the batch on the general path.
// x, y: the loaders issue 16-byte cp.async.
__host__ inline bool eligible(int N, int M, int d, int dist_kind, float gamma, const void *x, const void *y) {
    if (!(N <= T && M <= T && (d % 4 == 0) && dist_kind != 2)) return false;
    if ((((uintptr_t)x | (uintptr_t)y) & 15) != 0) return false;
    if (dist_kind == 0 && !(gamma >= 3e-6f)) return false;
    return true;
}

}  // namespace fused
