// soft-DTW forward / backward for sm_100a.
//
// Replaces the reference's distance functions (soft_dtw_cuda.py:326-363, which
// materialise a (B,N,M,d) broadcast), its numba-CUDA wavefront kernels
// (:34-112) and the eager glue around them (:121-175).
//
// Forward
//   dist_kernel     D[b] = f(x_b . y_b^T) per 64x64 tile (fp32 FMA, direct
//                   (x-y)^2 accumulation for the euclidean variant), row norms
//   dp_fwd_kernel   anti-diagonal wavefront, one CTA per pair, one thread per
//                   row; R is carried in float64 (only add/sub/min in fp64, the
//                   exp/log see fp32 DIFFERENCES) and never round-trips through
//                   global memory; the three soft-min weights of every cell are
//                   saved so that the backward needs no exp and no R at all.
// Backward
//   dp_bwd_kernel   reverse wavefront E[i,j] = sum of successors' E x weight
//   grad_kernel     dX = A . Y + coef * X,  dY = A^T . X + coef * Y  (fp32 FMA)
//
// aux layout (caller-allocated, sdtw_aux_bytes): see struct AuxPlan.
#include <algorithm>
#include <mutex>

#include "common.cuh"
#include "sdtw_fused.cuh"
#include "sdtw_tc.cuh"

namespace {

using b200::align_up;

constexpr float COS_EPS = 1e-8f;  // torch cosine_similarity eps (soft_dtw_cuda.py:347)

struct AuxPlan {
    // cell (M, N) float4 per pair, [j][i] (i contiguous): {w_diag, w_up, w_left, D}
    // acoef (B,N,M): GEMM coefficient of the input gradient, scoef (B,N,M): summand of the
    // "own vector" coefficient; nx (B,N), ny (B,M) norms; rsum (B,N), csum (B,M)
    size_t off_cell, off_d, off_acoef, off_nx, off_ny, off_rsum, off_csum, off_flag, total;
};

AuxPlan make_aux(int B, int N, int M) {
    AuxPlan a;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes, 256); return r; };
    a.off_cell = take((size_t)B * N * M * sizeof(float4));
    a.off_d = take((size_t)B * N * M * 4);
    a.off_acoef = take((size_t)B * N * M * 4);
    a.off_nx = take((size_t)B * N * 4);
    a.off_ny = take((size_t)B * M * 4);
    a.off_rsum = take((size_t)B * N * 4);
    a.off_csum = take((size_t)B * M * 4);
    a.off_flag = take(256);  // int: the fused forward could not represent this batch (see fused::WE_LIMIT)
    a.total = o;
    return a;
}

// ----------------------------------------------------------------------------
// distance tile kernel
// ----------------------------------------------------------------------------
constexpr int DT = 64;   // tile edge
constexpr int DK = 32;   // feature chunk

template <int KIND>
__global__ void __launch_bounds__(256)
dist_kernel(const float *__restrict__ x, const float *__restrict__ y, int N, int M, int d,
            float *__restrict__ Dout, float *__restrict__ nx_out, float *__restrict__ ny_out,
            const int *__restrict__ run_if, const int *__restrict__ ix, const int *__restrict__ iy) {
    // run_if: device flag; the grid exits at once when it is 0 (the fused kernel handled the batch)
    if (run_if != nullptr && *run_if == 0) return;
    __shared__ float xs[DK][DT + 4];
    __shared__ float ys[DK][DT + 4];
    __shared__ float snx[DT], sny[DT];
    const int b = blockIdx.z;
    const int i0 = blockIdx.y * DT, j0 = blockIdx.x * DT;
    const int tid = threadIdx.x;
    const int ty = tid >> 4, tx = tid & 15;  // 16 x 16 threads, 4 x 4 cells each
    const float *xb = x + (size_t)(ix ? ix[b] : b) * N * d;   // pair b = (x sequence ix[b], y sequence iy[b])
    const float *yb = y + (size_t)(iy ? iy[b] : b) * M * d;
    float acc[4][4] = {};
    float nxa[4] = {}, nya[4] = {};
    // loader mapping: 64 rows x 32 features = 512 float4; two per thread
    const int lr = tid >> 3;          // 0..31 (+32)
    const int lc = (tid & 7) * 4;     // feature offset within the chunk
    const bool vec_ok = (d % 4 == 0) && ((((uintptr_t)x | (uintptr_t)y) & 15) == 0);
    for (int k0 = 0; k0 < d; k0 += DK) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int r = lr + h * 32;
            float4 vx = make_float4(0.f, 0.f, 0.f, 0.f), vy = vx;
            const int gi = i0 + r, gj = j0 + r, gk = k0 + lc;
            if (gi < N) {
                if (vec_ok && gk + 3 < d) vx = *reinterpret_cast<const float4 *>(xb + (size_t)gi * d + gk);
                else {
                    if (gk < d) vx.x = xb[(size_t)gi * d + gk];
                    if (gk + 1 < d) vx.y = xb[(size_t)gi * d + gk + 1];
                    if (gk + 2 < d) vx.z = xb[(size_t)gi * d + gk + 2];
                    if (gk + 3 < d) vx.w = xb[(size_t)gi * d + gk + 3];
                }
            }
            if (gj < M) {
                if (vec_ok && gk + 3 < d) vy = *reinterpret_cast<const float4 *>(yb + (size_t)gj * d + gk);
                else {
                    if (gk < d) vy.x = yb[(size_t)gj * d + gk];
                    if (gk + 1 < d) vy.y = yb[(size_t)gj * d + gk + 1];
                    if (gk + 2 < d) vy.z = yb[(size_t)gj * d + gk + 2];
                    if (gk + 3 < d) vy.w = yb[(size_t)gj * d + gk + 3];
                }
            }
            xs[lc][r] = vx.x; xs[lc + 1][r] = vx.y; xs[lc + 2][r] = vx.z; xs[lc + 3][r] = vx.w;
            ys[lc][r] = vy.x; ys[lc + 1][r] = vy.y; ys[lc + 2][r] = vy.z; ys[lc + 3][r] = vy.w;
        }
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < DK; ++k) {
            const float4 a = *reinterpret_cast<const float4 *>(&xs[k][ty * 4]);
            const float4 c = *reinterpret_cast<const float4 *>(&ys[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, cv[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
            for (int p = 0; p < 4; ++p)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (KIND == B200HEAD_DIST_EUCLIDEAN) {
                        const float t = av[p] - cv[q];
                        acc[p][q] = fmaf(t, t, acc[p][q]);
                    } else {
                        acc[p][q] = fmaf(av[p], cv[q], acc[p][q]);
                    }
                }
            if (KIND == B200HEAD_DIST_COSINE) {
#pragma unroll
                for (int p = 0; p < 4; ++p) {
                    nxa[p] = fmaf(av[p], av[p], nxa[p]);
                    nya[p] = fmaf(cv[p], cv[p], nya[p]);
                }
            }
        }
        __syncthreads();
    }
    if (KIND == B200HEAD_DIST_COSINE) {
        if (tx == 0)
#pragma unroll
            for (int p = 0; p < 4; ++p) snx[ty * 4 + p] = sqrtf(nxa[p]);
        if (ty == 0)
#pragma unroll
            for (int q = 0; q < 4; ++q) sny[tx * 4 + q] = sqrtf(nya[q]);
        __syncthreads();
        if (tid < DT) {
            if (blockIdx.x == 0 && i0 + tid < N) nx_out[(size_t)b * N + i0 + tid] = snx[tid];
            if (blockIdx.y == 0 && j0 + tid < M) ny_out[(size_t)b * M + j0 + tid] = sny[tid];
        }
    }
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int gi = i0 + ty * 4 + p;
        if (gi >= N) continue;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int gj = j0 + tx * 4 + q;
            if (gj >= M) continue;
            float v;
            if (KIND == B200HEAD_DIST_COSINE) {
                const float c = acc[p][q] / (fmaxf(snx[ty * 4 + p], COS_EPS) * fmaxf(sny[tx * 4 + q], COS_EPS));
                v = expf(1.f - c);
            } else if (KIND == B200HEAD_DIST_NEGATIVE_DOT) {
                v = -acc[p][q];
            } else {
                v = expf(sqrtf(acc[p][q]));
            }
            Dout[((size_t)b * N + gi) * M + gj] = v;
        }
    }
}

// ----------------------------------------------------------------------------
// forward wavefront
// ----------------------------------------------------------------------------
// thread i owns row i (0-based); at step p it handles column j = p - i.
__global__ void dp_fwd_kernel(const float *__restrict__ D, int N, int M, float gamma, float bandwidth,
                              float *__restrict__ Rout, float *__restrict__ out,
                              float4 *__restrict__ cell, const int *__restrict__ run_if) {
    if (run_if != nullptr && *run_if == 0) return;  // see dist_kernel
    extern __shared__ double sh_up[];  // [2][blockDim.x + 1]
    const int b = blockIdx.x, i = threadIdx.x, T = blockDim.x;
    double *buf0 = sh_up, *buf1 = sh_up + (T + 1);
    const double inf = (double)INFINITY;
    const double g = (double)gamma;
    const double inv_g = 1.0 / g;
    const float *Db = D + (size_t)b * N * M;
    float *Rb = Rout ? Rout + (size_t)b * (N + 2) * (M + 2) : nullptr;
    float4 *cb = cell ? cell + (size_t)b * N * M : nullptr;

    if (Rb) {  // border of the (N+2, M+2) array exactly as the reference initialises it
        for (int e = i; e < (N + 2) * (M + 2); e += T) {
            const int r = e / (M + 2), c = e % (M + 2);
            if (r == 0 || c == 0 || r == N + 1 || c == M + 1) Rb[e] = (e == 0) ? 0.f : INFINITY;
        }
    }
    // buf[k][i+1] = R of row i written at the step of parity k; slot 0 is the virtual row -1
    for (int e = i; e < 2 * (T + 1); e += T) sh_up[e] = inf;
    __syncthreads();

    double left = inf;       // R[i, j-1]
    double diag = inf;       // R[i-1, j-1]
    if (i == 0) diag = 0.0;  // R[-1,-1] = 0 in 0-based terms (reference R[0,0] = 0)
    const int steps = N + M - 1;
    for (int p = 0; p < steps; ++p) {
        double *cur = (p & 1) ? buf1 : buf0;        // written this step
        const double *prev = (p & 1) ? buf0 : buf1;  // written last step
        const int j = p - i;
        const bool active = (i < N) && (j >= 0) && (j < M);
        double rnew = inf;
        if (active) {
            const double up = (i == 0) ? inf : prev[i];  // R[i-1, j], slot (i-1)+1
            const bool pruned = (bandwidth > 0.f) && (fabsf((float)(i - j)) > bandwidth);
            float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
            const float dij = Db[(size_t)i * M + j];
            w.w = dij;
            if (!pruned) {
                const double m = fmin(fmin(diag, up), left);
                if (m < inf) {
                    const float e0 = expf((float)((m - diag) * inv_g));
                    const float e1 = expf((float)((m - up) * inv_g));
                    const float e2 = expf((float)((m - left) * inv_g));
                    const float s = e0 + e1 + e2;
                    rnew = (double)dij + m - g * (double)logf(s);
                    const float inv_s = 1.f / s;
                    w.x = e0 * inv_s;
                    w.y = e1 * inv_s;
                    w.z = e2 * inv_s;
                }
            }
            if (Rb) Rb[(size_t)(i + 1) * (M + 2) + (j + 1)] = (float)rnew;
            if (cb) cb[(size_t)j * N + i] = w;
            diag = up;   // becomes R[i-1, j] = diagonal neighbour of the next column
            left = rnew;
            if (i == N - 1 && j == M - 1) out[b] = (float)rnew;
        }
        cur[i + 1] = rnew;  // inactive rows publish +inf
        __syncthreads();
    }
}

// ----------------------------------------------------------------------------
// backward wavefront
// ----------------------------------------------------------------------------
template <int KIND>
__global__ void dp_bwd_kernel(const float4 *__restrict__ cell, const float *__restrict__ grad_out,
                              int N, int M, float bandwidth, float *__restrict__ Eout,
                              float *__restrict__ acoef, float *__restrict__ rsum,
                              float *__restrict__ csum) {
    extern __shared__ float sh_f[];  // [2][T+1] float2 products, then csum[M]
    const int b = blockIdx.x, i = threadIdx.x, T = blockDim.x;
    float2 *pb0 = reinterpret_cast<float2 *>(sh_f);
    float2 *pb1 = pb0 + (T + 1);
    float *scs = reinterpret_cast<float *>(pb1 + (T + 1));
    const float4 *cb = cell + (size_t)b * N * M;
    const float go = grad_out[b];
    for (int e = i; e < 2 * (T + 1); e += T) pb0[e] = make_float2(0.f, 0.f);
    for (int e = i; e < M; e += T) scs[e] = 0.f;
    __syncthreads();
    float own_left = 0.f;  // E[i, j+1] * w_left(i, j+1)
    float diag = 0.f;      // E[i+1, j+1] * w_diag(i+1, j+1)
    float rs = 0.f;
    const int steps = N + M - 1;
    for (int p = steps - 1; p >= 0; --p) {
        float2 *cur = (p & 1) ? pb1 : pb0;
        const float2 *prev = (p & 1) ? pb0 : pb1;  // written at step p+1
        const int j = p - i;
        const bool active = (i < N) && (j >= 0) && (j < M);
        float2 pub = make_float2(0.f, 0.f);
        if (active) {
            const float2 below = prev[i + 1];  // cell (i+1, j): {E*w_up, E*w_diag}
            const float4 w = cb[(size_t)j * N + i];
            const bool pruned = (bandwidth > 0.f) && (fabsf((float)(i - j)) > bandwidth);
            float e = 0.f;
            if (!pruned) {
                e = (i == N - 1 && j == M - 1) ? 1.f : (below.x + own_left + diag);
            }
            diag = below.y;
            own_left = e * w.z;
            pub = make_float2(e * w.y, e * w.x);
            if (Eout) Eout[((size_t)b * N + i) * M + j] = e;
            const float gE = go * e;
            float a, s = 0.f;
            if (KIND == B200HEAD_DIST_COSINE) {
                a = -gE * w.w;                 // d loss / d cos  (D = exp(1 - c))
                s = a * (1.f - logf(w.w));     // times c
            } else if (KIND == B200HEAD_DIST_NEGATIVE_DOT) {
                a = -gE;
            } else {
                a = -gE * w.w / logf(w.w);     // -(g D / r), r = log D; NaN at r = 0 like torch
                s = -a;
            }
            acoef[((size_t)b * N + i) * M + j] = a;
            rs += s;
            if (KIND != B200HEAD_DIST_NEGATIVE_DOT) atomicAdd(&scs[j], s);
        }
        cur[i] = pub;
        __syncthreads();
    }
    if (i < N) rsum[(size_t)b * N + i] = rs;
    for (int e = i; e < M; e += T) csum[(size_t)b * M + e] = scs[e];
}

// ----------------------------------------------------------------------------
// input-gradient kernel:  out[r, :] = sum_q A(r,q) * rscale(r) * qscale(q) * src[q, :]
//                                     + own(r) * self[r, :]
// with A(r,q) = acoef[r*sr + q*sq] so that one kernel serves dX (A) and dY (A^T).
// ----------------------------------------------------------------------------
constexpr int GR = 64, GD = 64, GQ = 16;

template <int KIND>
__global__ void __launch_bounds__(256)
grad_kernel(const float *__restrict__ acoef, long long sr, long long sq, int R, int Q, int d,
            const float *__restrict__ src, const float *__restrict__ self,
            const float *__restrict__ norm_r, const float *__restrict__ norm_q,
            const float *__restrict__ sum_r, float *__restrict__ out, const int *__restrict__ iq,
            const int *__restrict__ ir, int accumulate) {
    // iq / ir: the pair's sequence in the bank `src` is read from / `self` and `out` belong to (identity when
    // null); accumulate: several pairs may share an output sequence (atomic adds, out zeroed by the host)
    __shared__ float as[GQ][GR + 4];
    __shared__ float ss[GQ][GD + 4];
    const int b = blockIdx.z, r0 = blockIdx.y * GR, d0 = blockIdx.x * GD;
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const float *ab = acoef + (size_t)b * R * Q;  // R*Q == N*M either way
    const int bq = iq ? iq[b] : b, br = ir ? ir[b] : b;
    const float *sb = src + (size_t)bq * Q * d;
    const float *nq = norm_q ? norm_q + (size_t)b * Q : nullptr;
    float acc[4][4] = {};
    for (int q0 = 0; q0 < Q; q0 += GQ) {
        // A tile: GQ x GR coefficients (4 per thread)
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            const int e = tid + h * 256;
            int rr, qq;
            if (sq == 1) { qq = e % GQ; rr = e / GQ; }   // q contiguous in memory
            else { rr = e % GR; qq = e / GR; }           // r contiguous in memory
            const int gr = r0 + rr, gq = q0 + qq;
            float v = 0.f;
            if (gr < R && gq < Q) {
                v = ab[(size_t)gr * sr + (size_t)gq * sq];
                if (KIND == B200HEAD_DIST_COSINE) v /= fmaxf(nq[gq], COS_EPS);
            }
            as[qq][rr] = v;
        }
        // source tile: GQ x GD
#pragma unroll
        for (int h = 0; h < 4; ++h) {
            const int e = tid + h * 256;
            const int qq = e / GD, dd = e % GD;
            const int gq = q0 + qq, gd = d0 + dd;
            ss[qq][dd] = (gq < Q && gd < d) ? sb[(size_t)gq * d + gd] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < GQ; ++k) {
            const float4 a = *reinterpret_cast<const float4 *>(&as[k][ty * 4]);
            const float4 c = *reinterpret_cast<const float4 *>(&ss[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, cv[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
            for (int p = 0; p < 4; ++p)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[p][q] = fmaf(av[p], cv[q], acc[p][q]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int gr = r0 + ty * 4 + p;
        if (gr >= R) continue;
        float rscale = 1.f, own = 0.f;
        if (KIND == B200HEAD_DIST_COSINE) {
            const float rn = norm_r[(size_t)b * R + gr];
            const float cn = fmaxf(rn, COS_EPS);
            rscale = 1.f / cn;
            // - (sum_q gc * c) * x / (||x||_true * clamp(||x||)); zero vector -> 0 (norm subgradient)
            own = (rn > 0.f) ? -sum_r[(size_t)b * R + gr] / (rn * cn) : 0.f;
        } else if (KIND == B200HEAD_DIST_EUCLIDEAN) {
            own = sum_r[(size_t)b * R + gr];
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int gd = d0 + tx * 4 + q;
            if (gd >= d) continue;
            float v = acc[p][q] * rscale;
            if (KIND != B200HEAD_DIST_NEGATIVE_DOT) {
                const float sv = self[((size_t)br * R + gr) * d + gd];
                v = fmaf(own, sv, v);
            }
            if (accumulate) atomicAdd(&out[((size_t)br * R + gr) * d + gd], v);
            else out[((size_t)br * R + gr) * d + gd] = v;
        }
    }
}

// dist_kernel + dp_fwd_kernel: the path for any shape and distance.  run_if (device int, may be
// null): when given and zero, both grids exit immediately.
int general_forward(const float *x, const float *y, int B, int N, int M, int d, float gamma, float bandwidth,
                    int dist_kind, float *R, float *out, float *Dm, float *nx, float *ny, float4 *cell,
                    const int *run_if, cudaStream_t st, const char *who, const int *ix = nullptr,
                    const int *iy = nullptr) {
    dim3 grid((M + DT - 1) / DT, (N + DT - 1) / DT, B);
    if (dist_kind == B200HEAD_DIST_COSINE)
        dist_kernel<B200HEAD_DIST_COSINE><<<grid, 256, 0, st>>>(x, y, N, M, d, Dm, nx, ny, run_if, ix, iy);
    else if (dist_kind == B200HEAD_DIST_NEGATIVE_DOT)
        dist_kernel<B200HEAD_DIST_NEGATIVE_DOT><<<grid, 256, 0, st>>>(x, y, N, M, d, Dm, nx, ny, run_if, ix, iy);
    else
        dist_kernel<B200HEAD_DIST_EUCLIDEAN><<<grid, 256, 0, st>>>(x, y, N, M, d, Dm, nx, ny, run_if, ix, iy);
    const int T = std::max(32, (N + 31) / 32 * 32);
    const size_t sh = 2 * (size_t)(T + 1) * sizeof(double);
    dp_fwd_kernel<<<B, T, sh, st>>>(Dm, N, M, gamma, bandwidth, R, out, cell, run_if);
    b200::note_launches(2);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        b200::set_error("%s: launch of the distance / wavefront kernels failed: %s", who, cudaGetErrorString(e));
        return B200HEAD_ERR_CUDA;
    }
    return 0;
}

int check_args(const char *who, const void *x, const void *y, int B, int N, int M, int d, float gamma,
               int dist_kind) {
    if (!x || !y) {
        b200::set_error("%s: null input pointer", who);
        return B200HEAD_ERR_ARG;
    }
    if (B <= 0 || N <= 0 || M <= 0 || d <= 0) {
        b200::set_error("%s: B, N, M, d must be positive (got %d %d %d %d)", who, B, N, M, d);
        return B200HEAD_ERR_ARG;
    }
    if (N > 1024 || M > 1024) {
        b200::set_error("%s: sequence length %d x %d exceeds 1024 (one thread per row; the reference "
                        "falls back to its CPU path here, this library has none)", who, N, M);
        return B200HEAD_ERR_ARG;
    }
    if (B > 65535) {
        b200::set_error("%s: at most 65535 pairs per call (got %d); split the batch", who, B);
        return B200HEAD_ERR_ARG;
    }
    if (!(gamma > 0.f)) {
        b200::set_error("%s: gamma must be > 0", who);
        return B200HEAD_ERR_ARG;
    }
    if (dist_kind < 0 || dist_kind > 2) {
        b200::set_error("%s: unknown dist_kind %d", who, dist_kind);
        return B200HEAD_ERR_ARG;
    }
    return b200::check_device();
}

}  // namespace

extern "C" size_t sdtw_aux_bytes(int B, int N, int M, int d) {
    (void)d;
    if (B <= 0 || N <= 0 || M <= 0) return 0;
    return make_aux(B, N, M).total;
}

namespace {

// P pairs; pair p compares sequence ix[p] of the x bank (Bx sequences) with sequence iy[p] of the y bank
// (By sequences); ix == iy == nullptr is the plain batched case (P == Bx == By, pair p = (p, p)).
int forward_impl(const char *who, const float *x, int Bx, const float *y, int By, const int *ix, const int *iy,
                 int P, int N, int M, int d, float gamma, float bandwidth, int dist_kind, float *R, float *out,
                 void *aux, size_t aux_bytes, cudaStream_t st) {
    int rc = check_args(who, x, y, P, N, M, d, gamma, dist_kind);
    if (rc) return rc;
    if (!out) {
        b200::set_error("%s: out is null", who);
        return B200HEAD_ERR_ARG;
    }
    const AuxPlan a = make_aux(P, N, M);
    if (!aux || aux_bytes < a.total || ((uintptr_t)aux & 255)) {
        b200::set_error("%s: aux must be 256-byte aligned and >= %zu bytes (got %zu)", who, a.total, aux_bytes);
        return B200HEAD_ERR_WS;
    }
    uint8_t *ax = (uint8_t *)aux;
    float *Dm = reinterpret_cast<float *>(ax + a.off_d);
    float *nx = reinterpret_cast<float *>(ax + a.off_nx);
    float *ny = reinterpret_cast<float *>(ax + a.off_ny);
    int *flag = reinterpret_cast<int *>(ax + a.off_flag);
    const int *run_if = nullptr;
    if (fused::eligible(N, M, d, dist_kind, gamma, x, y)) {
        // short sequences: one fused kernel per pair (tcgen05 distance tile + wavefront)
        float4 *cellp = reinterpret_cast<float4 *>(ax + a.off_cell);
        // exp(1 - cos) is bounded and eligible() has checked gamma; negative_dot distances are not:
        // the kernel raises `flag` when |D| / gamma is beyond its scaled-pair exponents, and the general
        // kernels below then redo the batch (they exit at once when the flag is clear)
        int *fl = (dist_kind == B200HEAD_DIST_COSINE) ? nullptr : flag;
        if (fl) B200_CUDA_OK(cudaMemsetAsync(fl, 0, sizeof(int), st));
#ifdef SDTW_FWD_LEGACY   // variant builds: the mma.sync kernel of round 1, for A/B timing (no index lists)
        const int smem = (int)sizeof(fused::FwdSmem);
        auto kern = (dist_kind == B200HEAD_DIST_COSINE) ? fused::fwd_kernel<0> : fused::fwd_kernel<1>;
        B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        kern<<<P, 256, smem, st>>>(x, y, N, M, d, gamma, bandwidth, R, out, cellp, nx, ny, fl);
#else
        CUtensorMap mx, my;
        rc = b200::make_tmap_3d_f32(&mx, x, (uint64_t)d, (uint64_t)N, (uint64_t)Bx, sdtc::T, 0);
        if (rc) return rc;
        rc = b200::make_tmap_3d_f32(&my, y, (uint64_t)d, (uint64_t)M, (uint64_t)By, sdtc::T, 0);
        if (rc) return rc;
        auto kern = (dist_kind == B200HEAD_DIST_COSINE) ? sdtc::fwd_kernel<0> : sdtc::fwd_kernel<1>;
        B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, sdtc::FWD_SMEM));
        kern<<<P, sdtc::THREADS, sdtc::FWD_SMEM, st>>>(mx, my, N, M, d, gamma, bandwidth, R, out, cellp, nx, ny, fl,
                                                       ix, iy);
#endif
        b200::note_launches(1);
        B200_LAUNCH_OK("sdtw fused forward kernel");
        if (fl == nullptr) return 0;
        run_if = flag;
    }
    return general_forward(x, y, P, N, M, d, gamma, bandwidth, dist_kind, R, out, Dm, nx, ny,
                           reinterpret_cast<float4 *>(ax + a.off_cell), run_if, st, who, ix, iy);
}

int backward_impl(const char *who, const float *x, int Bx, const float *y, int By, const int *ix, const int *iy,
                  int P, const float *grad_out, int N, int M, int d, float gamma, float bandwidth, int dist_kind,
                  const void *aux, size_t aux_bytes, float *E, float *dx, float *dy, cudaStream_t st) {
    int rc = check_args(who, x, y, P, N, M, d, gamma, dist_kind);
    if (rc) return rc;
    if (!grad_out || !dx || !dy) {
        b200::set_error("%s: null pointer argument", who);
        return B200HEAD_ERR_ARG;
    }
    const AuxPlan a = make_aux(P, N, M);
    if (!aux || aux_bytes < a.total || ((uintptr_t)aux & 255)) {
        b200::set_error("%s: aux must be the buffer the forward filled (>= %zu bytes)", who, a.total);
        return B200HEAD_ERR_WS;
    }
    uint8_t *ax = (uint8_t *)const_cast<void *>(aux);
    const float4 *cell = reinterpret_cast<const float4 *>(ax + a.off_cell);
    float *acoef = reinterpret_cast<float *>(ax + a.off_acoef);
    float *nx = reinterpret_cast<float *>(ax + a.off_nx);
    float *ny = reinterpret_cast<float *>(ax + a.off_ny);
    float *rsum = reinterpret_cast<float *>(ax + a.off_rsum);
    float *csum = reinterpret_cast<float *>(ax + a.off_csum);
    const int accumulate = (ix || iy) ? 1 : 0;
    if (accumulate) {   // several pairs may share a sequence: gradients are accumulated with atomic adds
        B200_CUDA_OK(cudaMemsetAsync(dx, 0, (size_t)Bx * N * d * sizeof(float), st));
        if (dy != dx) B200_CUDA_OK(cudaMemsetAsync(dy, 0, (size_t)By * M * d * sizeof(float), st));
    }
    if (fused::eligible(N, M, d, dist_kind, gamma, x, y)) {
#ifdef SDTW_BWD_LEGACY   // variant builds: the mma.sync kernel of round 1, for A/B timing (no index lists)
        const int smem = (int)sizeof(fused::BwdSmem);
        auto kern = (dist_kind == B200HEAD_DIST_COSINE) ? fused::bwd_kernel<0> : fused::bwd_kernel<1>;
        B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        kern<<<P, 256, smem, st>>>(x, y, grad_out, N, M, d, bandwidth, cell, nx, ny, E, dx, dy);
#else
        CUtensorMap mx, my;
        rc = b200::make_tmap_3d_f32(&mx, x, (uint64_t)d, (uint64_t)N, (uint64_t)Bx, sdtc::T, 1);
        if (rc) return rc;
        rc = b200::make_tmap_3d_f32(&my, y, (uint64_t)d, (uint64_t)M, (uint64_t)By, sdtc::T, 1);
        if (rc) return rc;
        auto kern = (dist_kind == B200HEAD_DIST_COSINE) ? sdtc::bwd_kernel<0> : sdtc::bwd_kernel<1>;
        B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, sdtc::BWD_SMEM));
        kern<<<P, sdtc::THREADS, sdtc::BWD_SMEM, st>>>(mx, my, x, y, grad_out, N, M, d, bandwidth, cell, nx, ny, E, dx,
                                                       dy, ix, iy);
#endif
        b200::note_launches(1);
        B200_LAUNCH_OK("sdtw fused backward kernel");
        return 0;
    }
    const int T = std::max(32, (N + 31) / 32 * 32);
    const size_t sh = 2 * (size_t)(T + 1) * sizeof(float2) + (size_t)M * sizeof(float);
    dim3 gx((d + GD - 1) / GD, (N + GR - 1) / GR, P), gy((d + GD - 1) / GD, (M + GR - 1) / GR, P);
#define SDTW_BWD(KIND)                                                                                            \
    do {                                                                                                          \
        dp_bwd_kernel<KIND><<<P, T, sh, st>>>(cell, grad_out, N, M, bandwidth, E, acoef, rsum, csum);             \
        grad_kernel<KIND><<<gx, 256, 0, st>>>(acoef, (long long)M, 1ll, N, M, d, y, x, nx, ny, rsum, dx, iy, ix,  \
                                              accumulate);                                                        \
        grad_kernel<KIND><<<gy, 256, 0, st>>>(acoef, 1ll, (long long)M, M, N, d, x, y, ny, nx, csum, dy, ix, iy,  \
                                              accumulate);                                                        \
    } while (0)
    if (dist_kind == B200HEAD_DIST_COSINE) SDTW_BWD(B200HEAD_DIST_COSINE);
    else if (dist_kind == B200HEAD_DIST_NEGATIVE_DOT) SDTW_BWD(B200HEAD_DIST_NEGATIVE_DOT);
    else SDTW_BWD(B200HEAD_DIST_EUCLIDEAN);
#undef SDTW_BWD
    b200::note_launches(3);
    B200_LAUNCH_OK("sdtw backward kernels");
    return 0;
}

int check_pairs(const char *who, int Bx, int By, const int *ix, const int *iy, int P) {
    if (Bx <= 0 || By <= 0 || P <= 0) {
        b200::set_error("%s: Bx, By and P must be positive (got %d %d %d)", who, Bx, By, P);
        return B200HEAD_ERR_ARG;
    }
    if ((!ix && P != Bx) || (!iy && P != By)) {
        b200::set_error("%s: a bank without an index list must hold exactly P = %d sequences", who, P);
        return B200HEAD_ERR_ARG;
    }
    return 0;
}

}  // namespace

extern "C" int sdtw_forward(const float *x, const float *y, int B, int N, int M, int d, float gamma,
                            float bandwidth, int dist_kind, float *R, float *out, void *aux,
                            size_t aux_bytes, void *stream) {
    return forward_impl("sdtw_forward", x, B, y, B, nullptr, nullptr, B, N, M, d, gamma, bandwidth, dist_kind, R, out,
                        aux, aux_bytes, (cudaStream_t)stream);
}

extern "C" int sdtw_backward(const float *x, const float *y, const float *grad_out, int B, int N,
                             int M, int d, float gamma, float bandwidth, int dist_kind,
                             const void *aux, size_t aux_bytes, float *E, float *dx, float *dy,
                             void *stream) {
    return backward_impl("sdtw_backward", x, B, y, B, nullptr, nullptr, B, grad_out, N, M, d, gamma, bandwidth,
                         dist_kind, aux, aux_bytes, E, dx, dy, (cudaStream_t)stream);
}

extern "C" int sdtw_forward_pairs(const float *x, int Bx, const float *y, int By, const int *ix, const int *iy,
                                  int P, int N, int M, int d, float gamma, float bandwidth, int dist_kind,
                                  float *R, float *out, void *aux, size_t aux_bytes, void *stream) {
    int rc = check_pairs("sdtw_forward_pairs", Bx, By, ix, iy, P);
    if (rc) return rc;
    return forward_impl("sdtw_forward_pairs", x, Bx, y, By, ix, iy, P, N, M, d, gamma, bandwidth, dist_kind, R, out,
                        aux, aux_bytes, (cudaStream_t)stream);
}

extern "C" int sdtw_backward_pairs(const float *x, int Bx, const float *y, int By, const int *ix, const int *iy,
                                   int P, const float *grad_out, int N, int M, int d, float gamma,
                                   float bandwidth, int dist_kind, const void *aux, size_t aux_bytes, float *E,
                                   float *dx, float *dy, void *stream) {
    int rc = check_pairs("sdtw_backward_pairs", Bx, By, ix, iy, P);
    if (rc) return rc;
    return backward_impl("sdtw_backward_pairs", x, Bx, y, By, ix, iy, P, grad_out, N, M, d, gamma, bandwidth,
                         dist_kind, aux, aux_bytes, E, dx, dy, (cudaStream_t)stream);
}

// ----------------------------------------------------------------------------
// Forward without any allocation by anybody: sdtw_forward_noalloc
// ----------------------------------------------------------------------------
// Same kernels as sdtw_forward (dist_kernel + dp_fwd_kernel), but the scratch they
// need between them - the distance matrix D (B,N,M) and the two norm vectors -
// lives in the file-scope device array below instead of a caller-provided `aux`.
// It exists for callers that can hand over input and output pointers but have no
// way to obtain more device memory: a binding from another runtime that owns no
// device allocator, a process whose memory pool has already claimed the whole GPU
// (so a further cudaMalloc would fail), or a latency-critical loop that cannot
// afford the device-wide synchronisation cudaMalloc/cudaFree imply.
//
// The array is part of the library image: every GPU that loads libb200head.so
// holds these 8 MiB until the process exits, whether or not the entry point is
// ever called.  All calls on a device share it and it keeps whatever the last
// call left there.  Nothing in it has to start at zero: every word the wavefront
// kernel reads was written by the distance kernel of the same call, and the
// per-cell weights that the backward would need are not produced here at all
// (forward only).  Two calls that overlap on the same GPU (different streams, or
// different host threads) would overwrite each other's D - callers must order them.
namespace {

constexpr size_t NOALLOC_MAX_BYTES = (size_t)B200HEAD_SDTW_NOALLOC_MAX_BYTES;
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_MAX_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;
void *g_noalloc_dev_ptr[NOALLOC_MAX_DEVICES] = {};

// In host code `g_noalloc_scratch` is only a handle for the symbol, not an address
// a kernel may dereference.  The real device address differs per GPU and is looked
// up once per device, with that device current, through cudaGetSymbolAddress.
int noalloc_scratch_address(void **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        b200::set_error("sdtw_forward_noalloc: cudaGetDevice failed: %s", cudaGetErrorString(e));
        return B200HEAD_ERR_CUDA;
    }
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) {
        b200::set_error("sdtw_forward_noalloc: device ordinal %d not supported (max %d)", dev,
                        NOALLOC_MAX_DEVICES - 1);
        return B200HEAD_ERR_ARG;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_dev_ptr[dev] == nullptr) {
        void *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess || p == nullptr) {
            b200::set_error("sdtw_forward_noalloc: cudaGetSymbolAddress failed on device %d: %s",
                            dev, cudaGetErrorString(e));
            return B200HEAD_ERR_CUDA;
        }
        g_noalloc_dev_ptr[dev] = p;
    }
    *out = g_noalloc_dev_ptr[dev];
    return 0;
}

}  // namespace

extern "C" int sdtw_forward_noalloc(const float *x, const float *y, int B, int N, int M, int d,
                                    float gamma, float bandwidth, int dist_kind, float *R,
                                    float *out, void *stream) {
    int rc = check_args("sdtw_forward_noalloc", x, y, B, N, M, d, gamma, dist_kind);
    if (rc) return rc;
    if (!out) {
        b200::set_error("sdtw_forward_noalloc: out is null");
        return B200HEAD_ERR_ARG;
    }
    // scratch = D (B,N,M) | ||x|| (B,N) | ||y|| (B,M), each part 256-byte aligned
    const size_t off_d = 0;
    const size_t off_nx = align_up(off_d + (size_t)B * N * M * 4, 256);
    const size_t off_ny = align_up(off_nx + (size_t)B * N * 4, 256);
    const size_t need = align_up(off_ny + (size_t)B * M * 4, 256);
    if (need > NOALLOC_MAX_BYTES) {
        b200::set_error("sdtw_forward_noalloc: B=%d N=%d M=%d needs %zu bytes of scratch but the "
                        "built-in buffer holds %zu; use sdtw_forward with a caller-allocated aux",
                        B, N, M, need, NOALLOC_MAX_BYTES);
        return B200HEAD_ERR_WS;
    }
    void *base = nullptr;
    rc = noalloc_scratch_address(&base);
    if (rc) return rc;
    uint8_t *sx = (uint8_t *)base;
    float *Dm = reinterpret_cast<float *>(sx + off_d);
    float *nx = reinterpret_cast<float *>(sx + off_nx);
    float *ny = reinterpret_cast<float *>(sx + off_ny);
    return general_forward(x, y, B, N, M, d, gamma, bandwidth, dist_kind, R, out, Dm, nx, ny, nullptr, nullptr,
                           (cudaStream_t)stream, "sdtw_forward_noalloc");
}

#ifdef SDTW_PROFILE
extern "C" int sdtw_debug_read(long long *host16) {
    return (int)cudaMemcpyFromSymbol(host16, fused::g_prof, sizeof(long long) * 32);
}
#endif
