// MIL-NCE loss head for sm_100a: host orchestration + the small SIMT kernels
// around the tcgen05 GEMM core (tc_gemm.cuh).
//
// Replaces reference loss.py:10-18 (MILNCELoss.forward) and its autograd.
//
// Forward, for the samples [i0, i0+n) this call owns (n == B on one GPU):
//   stage   video/text -> fp16 operands (power-of-two scaled; fp32 inputs are
//           split into hi+lo fp16 planes concatenated along K so that three
//           tensor-core products reproduce the fp32 dot product to ~2^-22)
//   stats   S[range, :] = V[range] . T^T on the tensor cores; the epilogue
//           (EpiRowCol) reduces each tile to row AND column log-sum-exp
//           partials, S itself is never stored.  bf16 inputs use the CTA-pair
//           kernel (tc_gemm2.cuh), fp32 inputs the chunked one-CTA kernel.
//           With n < B (row + column sharding) a second GEMM over the column
//           block S[:, range*K] supplies the column statistic.
//   final   merge the partials, nom_i (exact fp32 dots),
//           den_i = logaddexp(row_i, col_i), loss = sum(den - nom) / B
// Backward (closed form):
//   G       recompute the S tiles -> G (n, B*K) fp16 (x 2^13)  [EpiG; never fp32, never S]
//   dvideo  = G . T          (EpiStore, fp32 out)
//   dtext   = G^T . V        bf16: G read as an MN-major UMMA operand, no transpose;
//                            fp32: a transposed pass writes G^T planes first
// The *_shard entry points are the rows-only multi-GPU variant: the column
// statistic and dtext come out as per-rank partials for the caller to combine.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_gemm.cuh"
#include "tc_gemm2.cuh"

namespace {

using b200::align_up;

constexpr int BLOCK_N = 128;

// Device-side scalars shared between kernels of one call.
struct Scalars {
    unsigned int absmax_bits[2];  // |video|max, |text|max as float bits
    float s_video, s_text;        // power-of-two staging scales
    float inv_ss;                 // 1 / (s_video * s_text): un-scales S
    float unused_[2];
    float den_mid_log2;           // (min den + max den) / 2 in the log2 domain
    int g_fast;                   // 1: den range small enough for the one-exp gradient tile
    unsigned int ticket;          // blocks of merge_finalize_kernel that are done (last one reduces the loss)
    int pad[6];
};

struct Plan {
    int B, K, D, dtype, i0, n;
    int P;               // K-planes: 1 (bf16 inputs) or 3 (fp32 inputs: hi|lo|hi x hi|hi|lo)
    long long BK;
    long long ld_vk, ld_tk;  // leading dims of the K-major staged operands (cols = P*D)
    long long ld_vt, ld_tt;  // transposed operands: Vt (D, P*B), Tt (D, P*BK)
    long long ld_g, ld_gt;   // G (n, P*BK), Gt (n*K, P*B)
    long long ld_g2;         // G2 (B, n*K) for the sharded bf16 backward
    int nb_r, nb_c;          // n-blocks of pass R (over BK columns) and pass C (over B columns)
    int nb_cl;               // 32-row blocks of the owned rows
    int cg;                  // columns per column statistic: K when the epilogue pre-sums the K candidates, else 1
    int chunk_s, chunk_g;    // k-blocks per tensor-core accumulation chunk (S passes / grad GEMMs)
    size_t off_scal, off_vk, off_tk, off_vt, off_tt, off_pr, off_pc, off_diff, off_bsum, off_g, off_gt, total;
};

bool make_plan(Plan &p, int B, int K, int D, int dtype, int i0, int n) {
    if (B <= 0 || K <= 0 || D <= 0 || n <= 0 || i0 < 0 || i0 + n > B) return false;
    if (dtype != B200HEAD_F32 && dtype != B200HEAD_BF16) return false;
    p.B = B; p.K = K; p.D = D; p.dtype = dtype; p.i0 = i0; p.n = n;
    p.P = (dtype == B200HEAD_F32) ? 3 : 1;
    // fp32 inputs: 8 MMA steps per chunk (truncation bias ~5e-7); bf16 inputs: the
    // similarity in one chunk (<= 32 steps at D=512), gradient GEMMs in 64-step chunks.
    p.chunk_s = (dtype == B200HEAD_F32) ? 2 : 16;
    p.chunk_g = (dtype == B200HEAD_F32) ? 2 : 16;
    p.BK = (long long)B * K;
    p.ld_vk = p.ld_tk = (long long)align_up((size_t)p.P * D, 8);
    p.ld_vt = (long long)align_up((size_t)p.P * B, 8);
    p.ld_tt = (long long)align_up((size_t)p.P * p.BK, 8);
    p.ld_g = p.ld_tt;
    p.ld_gt = p.ld_vt;
    // one row partial per epilogue warp column part; bf16 mode runs 256-wide pair tiles (all written)
    p.nb_r = tc::EpiRowCol::PARTS * ((p.P == 1) ? (int)((p.BK + 255) / 256) : (int)((p.BK + BLOCK_N - 1) / BLOCK_N));
    p.cg = (p.P == 1 && (K == 2 || K == 4 || K == 8)) ? K : 1;
    // one column partial per 32-row block (one per epilogue warp quarter); bf16 mode runs 256-row
    // tiles, and every 32-row block of a tile is written
    p.nb_c = (p.P == 1) ? 8 * ((B + 255) / 256) : 4 * ((B + 127) / 128);
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes, 1024); return r; };
    p.off_scal = take(sizeof(Scalars));
    p.off_vk = take((size_t)B * p.ld_vk * 2);
    p.off_tk = take((size_t)p.BK * p.ld_tk * 2);
    // transposed operand copies: only the fp32-accurate path (one-CTA kernels) needs them; the CTA-pair
    // kernel reads text and video as MN-major B operands
    p.off_vt = take(p.P == 1 ? 0 : (size_t)D * p.ld_vt * 2);
    p.off_tt = take(p.P == 1 ? 0 : (size_t)D * p.ld_tt * 2);
    p.off_pr = take((size_t)p.nb_r * n * 4);
    p.nb_cl = 8 * ((n + 255) / 256);  // 32-row blocks of the owned rows (rows-only sharding)
    p.off_pc = take(std::max((size_t)p.nb_c * n * K, (size_t)p.nb_cl * (size_t)p.BK) / p.cg * 4);
    p.off_diff = take((size_t)n * 4 * 2);  // den - nom, then the merged den
    p.off_bsum = take((size_t)((n + 31) / 32) * 16);  // per block: sum of den - nom (double), min den, max den
    p.off_g = take((size_t)n * p.ld_g * 2);
    // P == 3: G^T (n*K, P*B) from the transposed pass.  P == 1: G2 = G[all rows, own columns]
    // (B, n*K), read MN-major by the dtext GEMM; not needed at all when n == B.
    p.ld_g2 = (long long)align_up((size_t)n * K, 8);
    p.off_gt = take(p.P == 3 ? (size_t)n * K * p.ld_gt * 2 : (n != B ? (size_t)B * p.ld_g2 * 2 : 0));
    p.total = o;
    return true;
}

// ----------------------------------------------------------------------------
// staging kernels
// ----------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float to_f32(T x);
template <> __device__ __forceinline__ float to_f32<float>(float x) { return x; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 x) {
    return __bfloat162float(x);
}

template <typename T>
__global__ void absmax_kernel(const T *__restrict__ x, size_t count, unsigned int *out_bits) {
    float m = 0.f;
    constexpr int VEC = 16 / sizeof(T);  // elements per 16-byte load
    const size_t nvec = (((uintptr_t)x & 15) == 0) ? count / VEC : 0;
    const uint4 *xv = reinterpret_cast<const uint4 *>(x);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec;
         i += (size_t)gridDim.x * blockDim.x) {
        const uint4 q = xv[i];
        const T *e = reinterpret_cast<const T *>(&q);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const float v = fabsf(to_f32(e[j]));
            if (v == v && v <= 3.0e38f) m = fmaxf(m, v);  // ignore NaN/inf for the scale choice
        }
    }
    for (size_t i = nvec * VEC + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < count;
         i += (size_t)gridDim.x * blockDim.x) {
        const float v = fabsf(to_f32(x[i]));
        if (v == v && v <= 3.0e38f) m = fmaxf(m, v);
    }
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(out_bits, __float_as_uint(m));
}

// scale = 2^(13 - floor(log2(absmax))): staged magnitudes land in [2^13, 2^14) < fp16 max
__device__ __forceinline__ float pow2_scale(unsigned int absmax_bits) {
    if (absmax_bits == 0) return 1.f;
    int e = (int)((absmax_bits >> 23) & 0xff) - 127;  // floor(log2) for normal floats
    if (((absmax_bits >> 23) & 0xff) == 0) e = -126;
    int k = 13 - e;
    k = max(-100, min(100, k));
    return exp2f((float)k);
}

__global__ void scalars_forward_kernel(Scalars *s) {
    float sv = pow2_scale(s->absmax_bits[0]);
    float st = pow2_scale(s->absmax_bits[1]);
    s->s_video = sv;
    s->s_text = st;
    s->inv_ss = (1.f / sv) * (1.f / st);
}

// finite and within 60 nats: 2^(den_i - den_j) and its two factors stay in fp32 range
__device__ __forceinline__ void set_den_range(Scalars *s, float lo, float hi) {
    const bool ok = (hi - lo) < 60.f && isfinite(hi) && isfinite(lo);
    s->g_fast = ok ? 1 : 0;
    s->den_mid_log2 = ok ? 0.5f * (lo + hi) * tc::LOG2E : 0.f;
}

// one block: gradient scales + the range of den (decides the one-exp path of EpiG)
__global__ void scalars_backward_kernel(Scalars *s, const float *grad_loss, const float *den, int B) {
    __shared__ float smin[32], smax[32];
    tc::pdl_wait();
    tc::pdl_trigger();
    float lo = INFINITY, hi = -INFINITY;
    for (int i = threadIdx.x; i < B; i += blockDim.x) {
        const float d = den[i];
        lo = fminf(lo, d);
        hi = fmaxf(hi, d);
    }
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) {
        smin[threadIdx.x >> 5] = lo;
        smax[threadIdx.x >> 5] = hi;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) {
            lo = fminf(lo, smin[w]);
            hi = fmaxf(hi, smax[w]);
        }
        set_den_range(s, lo, hi);
    }
}

// Plane order along K.  A-side [hi | lo | hi], B-side [hi | hi | lo]:
// sum_k A'[k] B'[k] = hi.hi + lo.hi + hi.lo  (the lo.lo term, ~2^-22, is dropped).
__device__ __forceinline__ void split_store(float xs, int planes, bool a_side, __half *dst,
                                            long long plane_stride) {
    const __half h = __float2half_rn(xs);
    dst[0] = h;
    if (planes == 3) {
        const __half l = __float2half_rn(xs - __half2float(h));
        dst[plane_stride] = a_side ? l : h;
        dst[2 * plane_stride] = a_side ? h : l;
    }
}

// in (rows, D) -> outk (rows, planes*D) K-major [a_side order]
//              -> outt (D, planes*rows) transposed [B-side order] (optional)
template <typename T>
__global__ void stage_operand_kernel(const T *__restrict__ in, int rows, int D,
                                     const float *__restrict__ scale_ptr, int planes, bool a_side,
                                     __half *__restrict__ outk, long long ldk,
                                     __half *__restrict__ outt, long long ldt) {
    __shared__ float tile[32][33];
    const float sc = *scale_ptr;
    const int r0 = blockIdx.y * 32, d0 = blockIdx.x * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int r = r0 + j, d = d0 + threadIdx.x;
        float v = 0.f;
        if (r < rows && d < D) {
            v = to_f32(in[(size_t)r * D + d]) * sc;
            split_store(v, planes, a_side, outk + (size_t)r * ldk + d, D);
        }
        tile[j][threadIdx.x] = v;
    }
    if (outt == nullptr) return;
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int d = d0 + j, r = r0 + threadIdx.x;
        if (r < rows && d < D)
            split_store(tile[threadIdx.x][j], planes, false, outt + (size_t)d * ldt + r, rows);
    }
}

// bf16 fast path of the above (planes == 1, D % 8 == 0, 16-byte aligned rows): 64 x 64 tiles,
// 16-byte global loads and stores on both the K-major copy and the transposed copy.  One launch
// stages both tensors (blockIdx.y < tiles_v: video, else text) and derives the power-of-two
// scales from the absmax bits itself; block (0, 0) publishes them for the later kernels.
struct StageSide {
    const __nv_bfloat16 *in;
    int rows;
    __half *outk;
    long long ldk;
    __half *outt;
    long long ldt;
};
__global__ void __launch_bounds__(256)
stage_bf16_kernel(StageSide sv, StageSide stx, int tiles_v, int D, Scalars *scal) {
    __shared__ __half tile[64][64 + 2];  // +2 halves: conflict-free column reads
    tc::pdl_wait();
    tc::pdl_trigger();
    const float s_v = pow2_scale(scal->absmax_bits[0]), s_t = pow2_scale(scal->absmax_bits[1]);
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {  // nobody reads these in this kernel
        scal->s_video = s_v;
        scal->s_text = s_t;
        scal->inv_ss = (1.f / s_v) * (1.f / s_t);
    }
    const bool is_v = (int)blockIdx.x < tiles_v;  // row tiles on x (no 65535 limit), feature tiles on y
    const __nv_bfloat16 *__restrict__ in = is_v ? sv.in : stx.in;
    __half *__restrict__ outk = is_v ? sv.outk : stx.outk;
    __half *__restrict__ outt = is_v ? sv.outt : stx.outt;
    const int rows = is_v ? sv.rows : stx.rows;
    const long long ldk = is_v ? sv.ldk : stx.ldk, ldt = is_v ? sv.ldt : stx.ldt;
    const float sc = is_v ? s_v : s_t;
    const int r0 = ((int)blockIdx.x - (is_v ? 0 : tiles_v)) * 64, d0 = blockIdx.y * 64;
    const int tid = threadIdx.x;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int g = tid + h * 256;          // 512 granules of 8 elements
        const int rr = g >> 3, dc = (g & 7) * 8;
        const int r = r0 + rr, d = d0 + dc;
        __align__(16) __half hv[8];
        if (r < rows && d < D) {              // D % 8 == 0: a granule is all in or all out
            const uint4 q = *reinterpret_cast<const uint4 *>(in + (size_t)r * D + d);
            const __nv_bfloat16 *e = reinterpret_cast<const __nv_bfloat16 *>(&q);
#pragma unroll
            for (int j = 0; j < 8; ++j) hv[j] = __float2half_rn(__bfloat162float(e[j]) * sc);
            *reinterpret_cast<uint4 *>(outk + (size_t)r * ldk + d) = *reinterpret_cast<uint4 *>(hv);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) hv[j] = __float2half_rn(0.f);
        }
        if (outt != nullptr) {
#pragma unroll
            for (int j = 0; j < 8; ++j) tile[rr][dc + j] = hv[j];
        }
    }
    if (outt == nullptr) return;  // (block-uniform) no transposed copy wanted
    __syncthreads();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int g = tid + h * 256;
        const int dd = g >> 3, rc = (g & 7) * 8;   // transposed granule: feature dd, 8 rows from rc
        const int d = d0 + dd, r = r0 + rc;
        if (d >= D || r >= rows) continue;
        __align__(16) __half hv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) hv[j] = tile[rc + j][dd];
        __half *dst = outt + (size_t)d * ldt + r;
        if (r + 8 <= rows && ((((uintptr_t)dst) & 15) == 0)) {
            *reinterpret_cast<uint4 *>(dst) = *reinterpret_cast<uint4 *>(hv);
        } else {
            for (int j = 0; j < 8 && r + j < rows; ++j) dst[j] = hv[j];
        }
    }
}

__device__ __forceinline__ float warp_max(float v) {
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// bf16 fast path, first pass over the inputs (one launch instead of three): one warp per sample i
// reads video[i] and its K candidates text[iK .. iK+K) once and produces
//   - the running |x|max of both tensors (atomicMax on the float bits, like absmax_kernel)
//   - nom_i = log sum_k exp(<video_i, text_{iK+k}>) from fp32 FMAs on the original values
//     (loss.py:13-15), for the samples in [i0, i0+n) when nom != nullptr.
// D % 8 == 0 and 16-byte aligned rows (checked by the caller).
__global__ void __launch_bounds__(256, 8)
prep_bf16_kernel(const __nv_bfloat16 *__restrict__ video, const __nv_bfloat16 *__restrict__ text, int B,
                 int K, int D, int i0, int n, float *__restrict__ nom, unsigned int *absmax_bits) {
    tc::pdl_wait();
    tc::pdl_trigger();
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    float mv = 0.f, mt = 0.f;
    auto amax8 = [](float m, const float (&f)[8]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float a = fabsf(f[j]);
            if (a == a && a <= 3.0e38f) m = fmaxf(m, a);  // ignore NaN/inf for the scale choice
        }
        return m;
    };
    auto unpack8 = [](const uint4 &q, float (&f)[8]) {
        const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&q);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 v = __bfloat1622float2(h[j]);
            f[2 * j] = v.x;
            f[2 * j + 1] = v.y;
        }
    };
    for (int i = blockIdx.x * wpb + (threadIdx.x >> 5); i < B; i += gridDim.x * wpb) {
        const bool want_nom = (nom != nullptr) && i >= i0 && i < i0 + n;
        const uint4 *vrow = reinterpret_cast<const uint4 *>(video + (size_t)i * D);
        float pm = -INFINITY, ps = 0.f;
#pragma unroll 2
        for (int k = 0; k < K; ++k) {
            const uint4 *trow = reinterpret_cast<const uint4 *>(text + ((size_t)i * K + k) * D);
            float acc = 0.f;
#pragma unroll 2
            for (int c = lane; c < D / 8; c += 32) {
                float fv[8], ft[8];
                unpack8(__ldg(vrow + c), fv);  // re-read per candidate: 1 KB per warp, stays in L1
                unpack8(__ldg(trow + c), ft);
                if (k == 0) mv = amax8(mv, fv);
                mt = amax8(mt, ft);
#pragma unroll
                for (int j = 0; j < 8; ++j) acc = fmaf(fv[j], ft[j], acc);
            }
            if (want_nom) {
                acc = warp_sum(acc);
                if (acc > pm) {
                    ps = ps * expf(pm - acc) + 1.f;
                    pm = acc;
                } else {
                    ps += expf(acc - pm);
                }
            }
        }
        if (want_nom && lane == 0) nom[i] = pm + logf(ps);
    }
    mv = warp_max(mv);
    mt = warp_max(mt);
    __shared__ float sv[8], stx[8];
    if (lane == 0) {
        sv[threadIdx.x >> 5] = mv;
        stx[threadIdx.x >> 5] = mt;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < wpb; ++w) {
            mv = fmaxf(mv, sv[w]);
            mt = fmaxf(mt, stx[w]);
        }
        if (mv > 0.f) atomicMax(&absmax_bits[0], __float_as_uint(mv));
        if (mt > 0.f) atomicMax(&absmax_bits[1], __float_as_uint(mt));
    }
}

// ----------------------------------------------------------------------------
// multi-GPU without an embedding collective: every rank PUBLISHES its bf16 shard in a buffer that its
// peers can read (peer-mapped / symmetric memory), and stages the fp16 operands of ALL ranks straight
// from the peers' buffers (plain 16-byte loads over NVLink; no TMA touches another GPU's memory).
// Slot layout (milnce_peer_slot_bytes): 256-byte header {|video|max bits, |text|max bits} | video (b, D) |
// text (b*K, D), bfloat16.
// ----------------------------------------------------------------------------
constexpr int PEER_MAX = 16;
constexpr size_t PEER_HEADER = 256;
struct PeerSlots {
    const uint8_t *slot[PEER_MAX];
};

// one warp per local sample i: copies video[i] and its K candidates into the slot, accumulates the local
// |x|max of both tensors into the slot header (atomicMax on float bits; the header is zeroed by the host
// before the launch) and produces nom_i = log sum_k exp(<video_i, text_{iK+k}>) from fp32 FMAs (loss.py:13-15).
__global__ void __launch_bounds__(256, 4)
peer_publish_kernel(const __nv_bfloat16 *__restrict__ video, const __nv_bfloat16 *__restrict__ text, int b, int K,
                    int D, uint8_t *__restrict__ slot, float *__restrict__ nom_own) {
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    unsigned int *hdr = reinterpret_cast<unsigned int *>(slot);
    uint4 *sv = reinterpret_cast<uint4 *>(slot + PEER_HEADER);
    uint4 *stx = sv + (size_t)b * D / 8;
    float mv = 0.f, mt = 0.f;
    auto amax8 = [](float m, const float (&f)[8]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float a = fabsf(f[j]);
            if (a == a && a <= 3.0e38f) m = fmaxf(m, a);
        }
        return m;
    };
    auto unpack8 = [](const uint4 &q, float (&f)[8]) {
        const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&q);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 v = __bfloat1622float2(h[j]);
            f[2 * j] = v.x;
            f[2 * j + 1] = v.y;
        }
    };
    const int dv = D / 8;
    // candidates are taken KB at a time with all their loads issued before the first use: a warp that walked
    // the K rows one after the other paid one memory round trip per candidate (30 us for 2048 samples)
    constexpr int KB = 4;
    for (int i = blockIdx.x * wpb + (threadIdx.x >> 5); i < b; i += gridDim.x * wpb) {
        const uint4 *vrow = reinterpret_cast<const uint4 *>(video + (size_t)i * D);
        float pm = -INFINITY, ps = 0.f;
        for (int k0 = 0; k0 < K; k0 += KB) {
            float acc[KB];
#pragma unroll
            for (int u = 0; u < KB; ++u) acc[u] = 0.f;
            for (int c = lane; c < dv; c += 32) {
                const uint4 qv = __ldg(vrow + c);
                uint4 qt[KB];
#pragma unroll
                for (int u = 0; u < KB; ++u)
                    if (k0 + u < K) qt[u] = __ldg(reinterpret_cast<const uint4 *>(text + ((size_t)i * K + k0 + u) * D) + c);
                float fv[8];
                unpack8(qv, fv);
                if (k0 == 0) {
                    mv = amax8(mv, fv);
                    sv[(size_t)i * dv + c] = qv;
                }
#pragma unroll
                for (int u = 0; u < KB; ++u) {
                    if (k0 + u >= K) break;
                    float ft[8];
                    unpack8(qt[u], ft);
                    mt = amax8(mt, ft);
                    stx[((size_t)i * K + k0 + u) * dv + c] = qt[u];
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[u] = fmaf(fv[j], ft[j], acc[u]);
                }
            }
#pragma unroll
            for (int u = 0; u < KB; ++u) {
                if (k0 + u >= K) break;
                const float a = warp_sum(acc[u]);
                if (a > pm) {
                    ps = ps * expf(pm - a) + 1.f;
                    pm = a;
                } else {
                    ps += expf(a - pm);
                }
            }
        }
        if (lane == 0) nom_own[i] = pm + logf(ps);
    }
    mv = warp_max(mv);
    mt = warp_max(mt);
    if (lane == 0) {
        if (mv > 0.f) atomicMax(&hdr[0], __float_as_uint(mv));
        if (mt > 0.f) atomicMax(&hdr[1], __float_as_uint(mt));
    }
}

// bf16 shards -> scaled fp16 K-major operands in the local workspace: the text of ALL W ranks (the B operand of
// the similarity pass and of dvideo) and the video of THIS rank only (the rows scheme never touches the other
// ranks' video rows, so they do not cross NVLink).  The power-of-two scales come from the maximum over the W
// slot headers; block 0 publishes them for the later kernels.  D % 8 == 0, so the staged rows are unpadded and
// both tensors are flat arrays: a grid-stride loop over 16-byte granules, eight independent loads in flight
// per thread, 512 contiguous bytes per warp instruction.  (A 64 x 64-tile version with two loads per thread
// read the peers at 95 GB/s.)
// The sources are visited in ROTATED order - rank r reads the text of ranks r, r+1, ..., r+W-1 (mod W) - because
// all ranks run this kernel at the same moment: in natural order every rank pulled from rank 0 first, then from
// rank 1, ..., and one GPU's egress served all readers at once (29 MB in ~220 us at W = 8, half the rate of W = 2).
__global__ void __launch_bounds__(256)
stage_peers_kernel(PeerSlots ps, int W, int rank, int b, int K, int D, __half *__restrict__ vk,
                   __half *__restrict__ tk, Scalars *scal) {
    // The slot headers are read by ONE warp per block (lane r reads rank r's) and handed on through shared
    // memory.  When every thread read them, each GPU's two header words were hit by ~9 000 warp requests per
    // reader over NVLink, all on one L2 sector: the kernel's time grew by ~30 us per remote rank (218 us at W = 8).
    __shared__ unsigned int s_abs[2];
    if (threadIdx.x < 32) {
        unsigned int hv = 0, ht = 0;
        if ((int)threadIdx.x < W) {
            const unsigned int *h = reinterpret_cast<const unsigned int *>(ps.slot[threadIdx.x]);
            hv = __ldcg(h);               // positive floats compare like their bit patterns
            ht = __ldcg(h + 1);
        }
        for (int o = 16; o > 0; o >>= 1) {
            hv = max(hv, __shfl_xor_sync(0xffffffffu, hv, o));
            ht = max(ht, __shfl_xor_sync(0xffffffffu, ht, o));
        }
        if (threadIdx.x == 0) {
            s_abs[0] = hv;
            s_abs[1] = ht;
        }
    }
    __syncthreads();
    const unsigned int av = s_abs[0], at = s_abs[1];
    const float s_v = pow2_scale(av), s_t = pow2_scale(at);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        scal->absmax_bits[0] = av;
        scal->absmax_bits[1] = at;
        scal->s_video = s_v;
        scal->s_text = s_t;
        scal->inv_ss = (1.f / s_v) * (1.f / s_t);
    }
    const size_t gv = (size_t)b * D / 8, gt = gv * K;      // granules (8 elements) per shard: video, text
    const size_t n_text = gt * W, total = n_text + gv;     // [text of ranks rank, rank+1, ... (mod W) | own video]
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    auto source = [&](size_t g, int &r, size_t &o) {       // text granule g -> (source rank, granule within its text)
        const int p = (int)(g / gt);
        o = g - (size_t)p * gt;
        r = rank + p;
        if (r >= W) r -= W;
    };
    auto convert = [&](size_t g, const uint4 &q) {
        const bool is_v = g >= n_text;
        const float sc = is_v ? s_v : s_t;
        const __nv_bfloat16 *e = reinterpret_cast<const __nv_bfloat16 *>(&q);
        __align__(16) __half hv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) hv[j] = __float2half_rn(__bfloat162float(e[j]) * sc);
        uint4 *dst;
        if (is_v) {
            dst = reinterpret_cast<uint4 *>(vk) + (size_t)rank * gv + (g - n_text);
        } else {
            int r;
            size_t o;
            source(g, r, o);
            dst = reinterpret_cast<uint4 *>(tk) + (size_t)r * gt + o;
        }
        *dst = *reinterpret_cast<uint4 *>(hv);
    };
    auto src = [&](size_t g) -> const uint4 * {
        // (L2-only loads: another GPU wrote this buffer since the last time this SM may have read it)
        if (g >= n_text) return reinterpret_cast<const uint4 *>(ps.slot[rank] + PEER_HEADER) + (g - n_text);
        int r;
        size_t o;
        source(g, r, o);
        return reinterpret_cast<const uint4 *>(ps.slot[r] + PEER_HEADER) + gv + o;
    };
    size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; g + 7 * stride < total; g += 8 * stride) {
        uint4 q[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) q[u] = __ldcg(src(g + u * stride));
#pragma unroll
        for (int u = 0; u < 8; ++u) convert(g + u * stride, q[u]);
    }
    for (; g < total; g += stride) convert(g, __ldcg(src(g)));
}

// Reduce-scatter of the partial text gradient by PULL: every rank left its (B*K, D) bf16 partial in a buffer its
// peers can read; this rank sums block `rank` (its own b*K rows) of all W partials, in rank order, in float32
// (so the result does not depend on timing), and writes its dtext.  Plain 16-byte loads over NVLink.
template <bool OUT_BF16>
__global__ void __launch_bounds__(256)
peer_reduce_kernel(PeerSlots parts, int W, size_t first, size_t count, void *__restrict__ out) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < count; g += stride) {
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        for (int r0 = 0; r0 < W; r0 += 4) {
            uint4 q[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (r0 + u < W) q[u] = __ldcg(reinterpret_cast<const uint4 *>(parts.slot[r0 + u]) + first + g);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (r0 + u >= W) break;
                const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&q[u]);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 v = __bfloat1622float2(h[j]);
                    acc[2 * j] += v.x;
                    acc[2 * j + 1] += v.y;
                }
            }
        }
        if (OUT_BF16) {
            __nv_bfloat162 h[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(acc[2 * j], acc[2 * j + 1]);
            reinterpret_cast<uint4 *>(out)[g] = *reinterpret_cast<uint4 *>(h);
        } else {
            float4 *o = reinterpret_cast<float4 *>(out) + 2 * g;
            o[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
            o[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
        }
    }
}

// ----------------------------------------------------------------------------
// finalize: one warp per owned sample
// ----------------------------------------------------------------------------

// merge the per-tile partial log-sum-exps of 32 samples per block: lane = sample (coalesced
// reads), the block's 32 warps stride over the partials, then combine through smem.
//   part_r[t][w], t < nb_r           -> row_i = lse over all text columns
//   part_c[t][w*kc + k], t < nb_c    -> col_i = lse over all video rows and the K candidates
//                                       (kc = K / column-group statistics per sample)
// out[w] = logaddexp(row_w, col_w).  With fin.den != nullptr the kernel also finishes the forward:
// den, den - nom, and the loss (the last block to finish adds the per-block partial sums in a
// fixed order, in double: deterministic).
struct MergeFin {
    const float *nom;   // indexed by global sample (i0 + w)
    float *den;         // idem; nullptr: plain merge
    float *diff;        // (n) den - nom
    double *bsum;       // (gridDim.x) per-block sums of diff
    float *bmin, *bmax; // (gridDim.x) per-block range of den
    unsigned int *ticket;
    float *loss;
    int i0, B;
    Scalars *range_out; // non-null: the last block also publishes the den range (decides the one-exp
                        // gradient tile), so that a backward on the same workspace needs no scalar kernel
};
__global__ void __launch_bounds__(1024)
merge_partials_kernel(const float *__restrict__ part_r, const float *__restrict__ part_c, int kc, int n,
                      int nb_r, int nb_c, float *__restrict__ out, MergeFin fin) {
    constexpr int NW = 32;
    __shared__ float sm[NW][32], ss[NW][32];
    __shared__ bool is_last;
    tc::pdl_wait();
    tc::pdl_trigger();
    const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
    const int w = blockIdx.x * 32 + lane;
    float m = -INFINITY, s = 0.f;
    // eight independent loads at a time; a batch is reduced with its own maximum (eight independent
    // exponentials) and merged into the running (m, s) once - no 8-long chain of dependent pushes
    auto push8 = [&](const float (&v)[8]) {
        float bm = fmaxf(fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3])), fmaxf(fmaxf(v[4], v[5]), fmaxf(v[6], v[7])));
        if (!(bm > -INFINITY)) return;
        float bs = 0.f;
#pragma unroll
        for (int u = 0; u < 8; ++u) bs += __expf(v[u] - bm);  // exp(-inf) = 0 for absent entries
        if (bm > m) {
            s = s * __expf(m - bm) + bs;
            m = bm;
        } else {
            s += bs * __expf(bm - m);
        }
    };
    if (w < n) {
        for (int t = wp; t < nb_r; t += 8 * NW) {
            float v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u)
                v[u] = (t + u * NW < nb_r) ? __ldg(part_r + (size_t)(t + u * NW) * n + w) : -INFINITY;
            push8(v);
        }
        const size_t nk = (size_t)n * kc;
        for (int k = 0; k < kc; ++k) {
            const float *pc = part_c + (size_t)w * kc + k;
            for (int t = wp; t < nb_c; t += 8 * NW) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    v[u] = (t + u * NW < nb_c) ? __ldg(pc + (size_t)(t + u * NW) * nk) : -INFINITY;
                push8(v);
            }
        }
    }
    sm[wp][lane] = m;
    ss[wp][lane] = s;
    __syncthreads();
    if (wp != 0) return;
    float d = 0.f;
    if (w < n) {
        float mm = sm[0][lane];
        for (int j = 1; j < NW; ++j) mm = fmaxf(mm, sm[j][lane]);
        float tot = 0.f;
        for (int j = 0; j < NW; ++j)
            if (sm[j][lane] > -INFINITY) tot += ss[j][lane] * __expf(sm[j][lane] - mm);
        const float dn = mm + logf(tot);  // = logaddexp(row_i, col_i) = den_i  (loss.py:16-17)
        if (out != nullptr) out[w] = dn;
        if (fin.den != nullptr) {
            fin.den[fin.i0 + w] = dn;
            d = dn - fin.nom[fin.i0 + w];
            fin.diff[w] = d;
        }
    }
    if (fin.den == nullptr) return;
    double acc = (double)d;
    float lo = (w < n) ? fin.den[fin.i0 + w] : INFINITY, hi = (w < n) ? fin.den[fin.i0 + w] : -INFINITY;
    for (int o = 16; o > 0; o >>= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0) {
        fin.bsum[blockIdx.x] = acc;
        fin.bmin[blockIdx.x] = lo;
        fin.bmax[blockIdx.x] = hi;
        __threadfence();
        is_last = (atomicAdd(fin.ticket, 1u) == gridDim.x - 1);
    }
    __syncwarp();
    if (!is_last) return;
    __threadfence();
    acc = 0.0;
    lo = INFINITY;
    hi = -INFINITY;
    for (int b = lane; b < (int)gridDim.x; b += 32) {
        acc += __ldcg(fin.bsum + b);
        lo = fminf(lo, __ldcg(fin.bmin + b));
        hi = fmaxf(hi, __ldcg(fin.bmax + b));
    }
    for (int o = 16; o > 0; o >>= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lane == 0) {
        fin.loss[0] = (float)(acc / (double)fin.B);
        *fin.ticket = 0;
        if (fin.range_out != nullptr) set_den_range(fin.range_out, lo, hi);
    }
}

// warp per owned sample: positives from exact fp32 dot products (loss.py:13-15); generic path
// (the bf16 fast path gets them from prep_bf16_kernel)
template <typename T>
__global__ void finalize_kernel(const T *__restrict__ video, const T *__restrict__ text, int K, int D,
                                int i0, int n, float *__restrict__ nom) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n) return;
    const int i = i0 + w;
    float pm = -INFINITY, ps = 0.f;
    const T *v = video + (size_t)i * D;
    for (int k = 0; k < K; ++k) {
        const T *t = text + ((size_t)i * K + k) * D;
        float acc = 0.f;
        for (int d = lane; d < D; d += 32) acc = fmaf(to_f32(v[d]), to_f32(t[d]), acc);
        acc = warp_sum(acc);
        if (acc > pm) {
            ps = ps * expf(pm - acc) + 1.f;
            pm = acc;
        } else {
            ps += expf(acc - pm);
        }
    }
    if (lane == 0) nom[i] = pm + logf(ps);
}

// Rows-only sharding, after the exchange: pack r (2b + B floats) = rank r's [row_lse of its b samples |
// nom of its b samples | its partial column statistic of all B samples]; the packs are either the rows of one
// all-gathered array or W separate buffers in the ranks' peer-readable memory.  One block:
//   den_j = logaddexp(row_j, logsumexp_r col_part[r][j]), nom_all, loss = mean(den - nom)
// (replaces a cat / logsumexp / logaddexp / mean chain of eight small torch launches).
__global__ void __launch_bounds__(512)
shard_combine_kernel(PeerSlots packs, int W, int b, int B, float *__restrict__ den,
                     float *__restrict__ nom_all, float *__restrict__ loss, unsigned int *ticket,
                     double *__restrict__ block_sums) {
    __shared__ double sh[16];
    __shared__ bool last;
    double acc = 0.0;
    // one sample per thread when the caller gave scratch for a multi-block launch (all 2 + W loads of a sample,
    // possibly from other GPUs' memory, in flight at once: one remote round trip for the whole kernel);
    // otherwise a single block strides over the samples
    constexpr int WB = 8;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < B; j += gridDim.x * blockDim.x) {
        const int r = j / b, i = j - r * b;
        const float *own = reinterpret_cast<const float *>(packs.slot[r]);
        const float row = __ldcg(own + i);
        const float nm = __ldcg(own + b + i);
        float m = -INFINITY, sum = 0.f;
        for (int w0 = 0; w0 < W; w0 += WB) {
            float cp[WB];
#pragma unroll
            for (int k = 0; k < WB; ++k)
                cp[k] = (w0 + k < W) ? __ldcg(reinterpret_cast<const float *>(packs.slot[w0 + k]) + 2 * b + j) : -INFINITY;
            float mx = m;
#pragma unroll
            for (int k = 0; k < WB; ++k) mx = fmaxf(mx, cp[k]);
            if (mx > -INFINITY) {
                float sacc = sum * expf(m - mx);
#pragma unroll
                for (int k = 0; k < WB; ++k) sacc += expf(cp[k] - mx);   // exp(-inf) = 0 for the padding
                sum = sacc;
                m = mx;
            }
        }
        // the row statistic joins last: den = logaddexp(row, logsumexp_r col_part[r])
        const float mm = fmaxf(m, row);
        float d = -INFINITY;
        if (mm > -INFINITY) d = mm + logf(sum * expf(m - mm) + expf(row - mm));
        den[j] = d;
        nom_all[j] = nm;
        acc += (double)(d - nm);
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sh[w];     // fixed order
        if (gridDim.x == 1) {
            loss[0] = (float)(t / (double)B);
            last = false;
        } else {
            block_sums[blockIdx.x] = t;
            __threadfence();
            last = atomicAdd(ticket, 1u) == gridDim.x - 1;
        }
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        double t = 0.0;
        for (int k = 0; k < (int)gridDim.x; ++k) t += *reinterpret_cast<volatile double *>(block_sums + k);   // block order
        loss[0] = (float)(t / (double)B);
        *ticket = 0;        // ready for the next call
    }
}

// ----------------------------------------------------------------------------
// host helpers
// ----------------------------------------------------------------------------
int num_sms() {
    static thread_local int dev_cached = -1, sms = 0;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev != dev_cached) {
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        dev_cached = dev;
    }
    return sms > 0 ? sms : 148;
}

// A_MN = false: A is (M, Kd) with leading dimension lda.
// A_MN = true : A is passed transposed, (Kd, M) with leading dimension lda (MN-major operand).
template <class Epi, bool A_MN = false, int MB = 1, bool N_FAST = false>
int launch_gemm(const void *A, long long lda, const void *Bm, long long ldb, int M, int N, int Kd,
                int chunk_iters, const typename Epi::Params &ep, cudaStream_t stream,
                const char *what, int a_plane_rows = 0, int a_plane_cols = 0) {
    using L = tc::SmemLayout<BLOCK_N, Epi::STAGE_BYTES, MB>;
    CUtensorMap ma, mb;
    // (MN-major A in planes: the matrix is a_plane_rows x (planes * a_plane_cols), see GemmShape)
    const uint64_t a_rows = a_plane_rows > 0 ? (uint64_t)a_plane_rows : (uint64_t)Kd;
    const uint64_t a_cols = a_plane_rows > 0 ? (uint64_t)(Kd / a_plane_rows) * a_plane_cols : (uint64_t)M;
    int rc = A_MN ? b200::make_tmap_2d_f16(&ma, A, a_rows, a_cols, (uint64_t)lda, 64)
                  : b200::make_tmap_2d_f16(&ma, A, (uint64_t)M, (uint64_t)Kd, (uint64_t)lda,
                                           tc::BLOCK_M);
    if (rc) return rc;
    rc = b200::make_tmap_2d_f16(&mb, Bm, (uint64_t)N, (uint64_t)Kd, (uint64_t)ldb, BLOCK_N);
    if (rc) return rc;
    auto kern = tc::gemm_tn_kernel<BLOCK_N, Epi, A_MN, MB, N_FAST>;
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    const int tiles = ((M + MB * tc::BLOCK_M - 1) / (MB * tc::BLOCK_M)) * ((N + BLOCK_N - 1) / BLOCK_N);
    const int grid = tiles < num_sms() ? tiles : num_sms();
    tc::GemmShape shape{M, N, Kd, chunk_iters, a_plane_rows, a_plane_cols, 0u};
    kern<<<grid, tc::num_threads<Epi>(), L::TOTAL, stream>>>(ma, mb, shape, ep);
    b200::note_launches(1);
    B200_LAUNCH_OK(what);
    return 0;
}

// Variant builds (profiles/build_variant.sh) can switch the A-stationary statistics kernel off with
// -DB200_ASTAT=0 for A/B timing; the shipped library has no run-time switches.
#ifndef B200_ASTAT
#define B200_ASTAT 1
#endif
constexpr bool astat_enabled() { return B200_ASTAT != 0; }

// bf16 mode: CTA-pair kernel (tc_gemm2.cuh), 256 x 256 tiles.
// fmt: tc::IDESC_A_BF16 / tc::IDESC_B_BF16 when that operand holds bfloat16 instead of float16.
template <class Epi, bool A_MN = false, bool N_FAST = false, bool RANGE = false, bool B_MN = false>
int launch_gemm2(const void *A, long long lda, const void *Bm, long long ldb, int M, int N, int Kd,
                 const typename Epi::Params &ep, cudaStream_t stream, const char *what, uint32_t fmt = 0,
                 int k_splits = 0) {
    CUtensorMap ma, mb;
    int rc = A_MN ? b200::make_tmap_2d_f16(&ma, A, (uint64_t)Kd, (uint64_t)M, (uint64_t)lda, 64)
                  : b200::make_tmap_2d_f16(&ma, A, (uint64_t)M, (uint64_t)Kd, (uint64_t)lda,
                                           tc::BLOCK_M);
    if (rc) return rc;
    // B_MN: Bm is passed transposed, (Kd, N) with leading dimension ldb
    rc = B_MN ? b200::make_tmap_2d_f16(&mb, Bm, (uint64_t)Kd, (uint64_t)N, (uint64_t)ldb, 64)
              : b200::make_tmap_2d_f16(&mb, Bm, (uint64_t)N, (uint64_t)Kd, (uint64_t)ldb, tc::PAIR_N / 2);
    if (rc) return rc;
    const int tiles = ((M + tc::PAIR_M - 1) / tc::PAIR_M) * ((N + tc::PAIR_N - 1) / tc::PAIR_N);
    const int max_pairs = num_sms() / 2;
    const int units = tiles * (k_splits > 1 ? k_splits : 1);
    const int pairs = units < max_pairs ? units : max_pairs;
    tc::GemmShape shape{M, N, Kd, 0, 0, 0, fmt, k_splits};
    (void)what;
    // short reductions with an epilogue-heavy functor: keep A resident, stream only B
    // (not for EpiG: its 32 KB store-staging buffer leaves only three ring stages next to the resident A and
    // no L1 for its spilled per-tile state: 7680 vs 6310 cycles per tile)
    constexpr bool CAN_ASTAT = !A_MN && !B_MN && Epi::STAGE_BYTES == 0 && !std::is_same<Epi, tc::EpiStore>::value;
    if constexpr (CAN_ASTAT) {
        if (Kd <= tc::ASTAT_MAX_K_ITERS * tc::BLOCK_K && astat_enabled()) {
            using LA = tc::SmemLayoutAS<Epi::STAGE_BYTES>;
            auto kern = tc::gemm_tn_2cta_astat_kernel<Epi>;
            B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, LA::TOTAL));
            B200_CUDA_OK(b200::launch_pdl(kern, dim3(2 * pairs), dim3(tc::num_threads<Epi>()), LA::TOTAL, stream, ma,
                                          mb, shape, ep));
            return 0;
        }
    }
    using L = tc::SmemLayout2<Epi::STAGE_BYTES>;
    auto kern = tc::gemm_tn_2cta_kernel<Epi, A_MN, N_FAST, RANGE, B_MN>;
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    B200_CUDA_OK(b200::launch_pdl(kern, dim3(2 * pairs), dim3(tc::num_threads<Epi>()), L::TOTAL, stream, ma, mb,
                                  shape, ep));
    return 0;
}

// Gradient-tile pass: n-fastest contiguous tile ranges per CTA pair (the epilogue keeps its per-row
// state across the tiles of a row block).  -DB200_EPIG_RANGE=0 in a variant build restores the strided
// m-fastest order (A/B timing).
#ifndef B200_EPIG_RANGE
#define B200_EPIG_RANGE 1
#endif
template <class... Args>
int launch_epig2(Args... a) {
    if constexpr (B200_EPIG_RANGE != 0) return launch_gemm2<tc::EpiG, false, true, true>(a...);
    else return launch_gemm2<tc::EpiG, false, false, false>(a...);
}

// What the GEMMs read: the staged fp16 operands in the workspace and the scalars that un-scale them.
// (Feeding the caller's bf16 tensors to the tensor cores directly is not an option for the gradient
// GEMMs: tcgen05.mma.kind::f16 with A = fp16 gradient tiles and B = bf16 embeddings - two different
// operand formats in one instruction - raises an illegal-instruction fault on sm_100a, and bf16
// gradient tiles cost 1.7e-3 of accuracy.)
struct Operands {
    const __half *vk, *tk;      // K-major video (B, ld_vk) and text (BK, ld_tk)
    long long ld_vk, ld_tk;
    const float *inv_ss, *s_video, *s_text;  // device scalars
};
Operands operands(const Plan &p, uint8_t *ws) {
    Scalars *sc = reinterpret_cast<Scalars *>(ws + p.off_scal);
    return Operands{reinterpret_cast<const __half *>(ws + p.off_vk), reinterpret_cast<const __half *>(ws + p.off_tk),
                    p.ld_vk, p.ld_tk, &sc->inv_ss, &sc->s_video, &sc->s_text};
}

// Stages video/text into the fp16 operands.  When `nom` is non-null the positives' log-sum-exp of
// the owned samples is produced on the way (bf16 fast path) or by finalize_kernel (generic path).
template <typename T>
int stage_inputs(const Plan &p, const T *video, const T *text, uint8_t *ws, cudaStream_t st, float *nom) {
    Scalars *sc = reinterpret_cast<Scalars *>(ws + p.off_scal);
    B200_CUDA_OK(cudaMemsetAsync(sc, 0, sizeof(Scalars), st));
    __half *vk_ = reinterpret_cast<__half *>(ws + p.off_vk), *vt_ = reinterpret_cast<__half *>(ws + p.off_vt);
    __half *tk_ = reinterpret_cast<__half *>(ws + p.off_tk), *tt_ = reinterpret_cast<__half *>(ws + p.off_tt);
    bool fast = false;
    if constexpr (std::is_same<T, __nv_bfloat16>::value) {
        fast = (p.D % 8 == 0) && (((uintptr_t)video & 15) == 0) && (((uintptr_t)text & 15) == 0);
        if (fast) {
            const int wpb = 8;
            const int blocks = std::min((p.B + wpb - 1) / wpb, 8 * num_sms());
            B200_CUDA_OK(b200::launch_pdl(prep_bf16_kernel, dim3(blocks), dim3(wpb * 32), 0, st, video, text, p.B,
                                          p.K, p.D, p.i0, p.n, nom, sc->absmax_bits));
            const int tiles_v = (p.B + 63) / 64, tiles_t = (int)((p.BK + 63) / 64);
            StageSide sv{video, p.B, vk_, p.ld_vk, nullptr, 0};
            StageSide stx{text, (int)p.BK, tk_, p.ld_tk, nullptr, 0};
            B200_CUDA_OK(b200::launch_pdl(stage_bf16_kernel, dim3(tiles_v + tiles_t, (p.D + 63) / 64), dim3(256), 0,
                                          st, sv, stx, tiles_v, p.D, sc));
        }
    }
    if (!fast) {
        const size_t cv = (size_t)p.B * p.D, ct = (size_t)p.BK * p.D;
        absmax_kernel<T><<<(unsigned)std::min<size_t>((cv + 1023) / 1024, 1184), 256, 0, st>>>(
            video, cv, &sc->absmax_bits[0]);
        absmax_kernel<T><<<(unsigned)std::min<size_t>((ct + 1023) / 1024, 1184), 256, 0, st>>>(
            text, ct, &sc->absmax_bits[1]);
        scalars_forward_kernel<<<1, 1, 0, st>>>(sc);
        b200::note_launches(5 + (nom != nullptr ? 1 : 0));  // 2 absmax, scalars, 2 staging, positives
        dim3 blk(32, 8);
        dim3 gv((p.D + 31) / 32, (p.B + 31) / 32), gt((p.D + 31) / 32, (unsigned)((p.BK + 31) / 32));
        // the transposed copies exist only in the three-plane layout (make_plan gives them no room when P == 1)
        __half *vt_out = (p.P == 1) ? nullptr : vt_, *tt_out = (p.P == 1) ? nullptr : tt_;
        stage_operand_kernel<T><<<gv, blk, 0, st>>>(video, p.B, p.D, &sc->s_video, p.P, true, vk_, p.ld_vk,
                                                    vt_out, p.ld_vt);
        stage_operand_kernel<T><<<gt, blk, 0, st>>>(text, (int)p.BK, p.D, &sc->s_text, p.P, false, tk_,
                                                    p.ld_tk, tt_out, p.ld_tt);
        if (nom != nullptr) {
            const int warps_per_block = 8;
            finalize_kernel<T><<<(p.n + warps_per_block - 1) / warps_per_block, warps_per_block * 32, 0, st>>>(
                video, text, p.K, p.D, p.i0, p.n, nom);
        }
    }
    B200_LAUNCH_OK("staging kernels");
    return 0;
}

int check_common(const Plan &p, const void *ws, size_t ws_bytes) {
    if (ws == nullptr || ((uintptr_t)ws & 255)) {
        b200::set_error("workspace must be non-null and 256-byte aligned");
        return B200HEAD_ERR_WS;
    }
    if (ws_bytes < p.total) {
        b200::set_error("workspace too small: need %zu bytes, got %zu", p.total, ws_bytes);
        return B200HEAD_ERR_WS;
    }
    if ((long long)p.B * p.K > 2000000000ll || (long long)p.P * p.BK > 2000000000ll) {
        b200::set_error("B*K too large for 32-bit tile indexing");
        return B200HEAD_ERR_ARG;
    }
    return 0;
}

// statistics GEMM with the column-group size chosen at run time (Plan::cg)
template <int G>
int launch_stats2_g(const float *scale, float *pr, long long ldr, float *pc, long long ldc, const void *A,
                    long long lda, const void *Bm, long long ldb, int M, int N, int Kd, cudaStream_t st,
                    const char *what) {
    typename tc::EpiRowColT<G>::Params e{pr, ldr, pc, ldc, scale};
    return launch_gemm2<tc::EpiRowColT<G>>(A, lda, Bm, ldb, M, N, Kd, e, st, what);
}
template <class... Args>
int launch_stats2(int cg, Args... a) {
    switch (cg) {
        case 2: return launch_stats2_g<2>(a...);
        case 4: return launch_stats2_g<4>(a...);
        case 8: return launch_stats2_g<8>(a...);
        default: return launch_stats2_g<1>(a...);
    }
}

template <typename T>
int forward_impl(const Plan &p, const T *video, const T *text, float *den, float *nom, float *loss,
                 uint8_t *ws, cudaStream_t st) {
    int rc = stage_inputs<T>(p, video, text, ws, st, nom);
    if (rc) return rc;
    Scalars *sc = reinterpret_cast<Scalars *>(ws + p.off_scal);
    const Operands o = operands(p, ws);
    const __half *vk = o.vk, *tk = o.tk;
    float *pr = reinterpret_cast<float *>(ws + p.off_pr);
    float *pc = reinterpret_cast<float *>(ws + p.off_pc);
    float *diff = reinterpret_cast<float *>(ws + p.off_diff);
    const int Kd = p.P * p.D;
    const long long ldc = (long long)p.n * p.K / p.cg;  // column statistics of the owned samples
    if (p.n == p.B) {
        // one pass over S: row statistic per thread, column statistic by a cross-lane reduction
        const char *what = "milnce forward (fused row+column statistics)";
        if (p.P == 1) {
            rc = launch_stats2(p.cg, o.inv_ss, pr, (long long)p.n, pc, ldc, vk, o.ld_vk, tk, o.ld_tk, p.B,
                               (int)p.BK, Kd, st, what);
        } else {
            tc::EpiRowCol::Params e{pr, (long long)p.n, pc, ldc, o.inv_ss};
            rc = launch_gemm<tc::EpiRowCol>(vk, o.ld_vk, tk, o.ld_tk, p.B, (int)p.BK, Kd, p.chunk_s, e, st, what);
        }
        if (rc) return rc;
    } else {
        // sharded: rows [i0, i0+n) x all columns -> row statistic of the owned samples
        const __half *va = vk + (size_t)p.i0 * o.ld_vk;
        // all rows x columns of the owned samples -> their column statistic
        const __half *tb = tk + (size_t)p.i0 * p.K * o.ld_tk;
        if (p.P == 1) {
            rc = launch_stats2(p.cg, o.inv_ss, pr, (long long)p.n, nullptr, 0, va, o.ld_vk, tk, o.ld_tk, p.n,
                               (int)p.BK, Kd, st, "milnce forward row block");
            if (rc) return rc;
            rc = launch_stats2(p.cg, o.inv_ss, nullptr, 0, pc, ldc, vk, o.ld_vk, tb, o.ld_tk, p.B,
                               p.n * p.K, Kd, st, "milnce forward column block");
        } else {
            tc::EpiRowCol::Params er{pr, (long long)p.n, nullptr, 0, o.inv_ss};
            rc = launch_gemm<tc::EpiRowCol>(va, o.ld_vk, tk, o.ld_tk, p.n, (int)p.BK, Kd, p.chunk_s, er, st,
                                            "milnce forward row block");
            if (rc) return rc;
            tc::EpiRowCol::Params ec{nullptr, 0, pc, ldc, o.inv_ss};
            rc = launch_gemm<tc::EpiRowCol>(vk, o.ld_vk, tb, o.ld_tk, p.B, p.n * p.K, Kd, p.chunk_s, ec, st,
                                            "milnce forward column block");
        }
        if (rc) return rc;
    }
    const int nblk = (p.n + 31) / 32;
    double *bsum = reinterpret_cast<double *>(ws + p.off_bsum);
    float *bmin = reinterpret_cast<float *>(bsum + nblk);
    MergeFin fin{nom, den, diff, bsum, bmin, bmin + nblk, &sc->ticket, loss, p.i0, p.B,
                 p.n == p.B ? sc : nullptr};
    B200_CUDA_OK(b200::launch_pdl(merge_partials_kernel, dim3((p.n + 31) / 32), dim3(1024), 0, st, pr, pc,
                                  p.K / p.cg, p.n, p.nb_r, p.nb_c, (float *)nullptr, fin));
    B200_LAUNCH_OK("milnce finalize");
    return 0;
}

template <typename T>
int backward_impl(const Plan &p, const T *video, const T *text, const float *den, const float *nom,
                  const float *grad_loss, void *dvideo, void *dtext, int grad_bf16, int reuse, uint8_t *ws,
                  cudaStream_t st) {
    int rc;
    if (!reuse) {
        rc = stage_inputs<T>(p, video, text, ws, st, nullptr);
        if (rc) return rc;
    }
    Scalars *sc = reinterpret_cast<Scalars *>(ws + p.off_scal);
    // The den range was published by the forward that left this workspace (single-GPU case); the
    // gradient scale grad_loss / (B 2^13 s) is formed by the GEMM epilogues from grad_loss itself.
    // Otherwise (fresh workspace, or den gathered from other ranks) one small kernel derives the range.
    if (!(reuse && p.n == p.B))
        B200_CUDA_OK(b200::launch_pdl(scalars_backward_kernel, dim3(1), dim3(1024), 0, st, sc, grad_loss, den, p.B));
    const float alpha0 = 1.f / ((float)p.B * tc::G_SCALE);
    const Operands o = operands(p, ws);
    const __half *vk = o.vk, *tk = o.tk;
    const __half *vt = reinterpret_cast<const __half *>(ws + p.off_vt);
    const __half *tt = reinterpret_cast<const __half *>(ws + p.off_tt);
    __half *g = reinterpret_cast<__half *>(ws + p.off_g);
    __half *gt = reinterpret_cast<__half *>(ws + p.off_gt);
    const int Kd = p.P * p.D;
    const tc::GFast fast{&sc->g_fast, &sc->den_mid_log2};
    // row block: G (n, BK): row r <-> sample i0+r, column c <-> sample c / K
    tc::EpiG::Params gr{den, nom, g, p.ld_g, p.BK, p.P, p.i0, tc::make_fastdiv(1), 0, tc::make_fastdiv(p.K), o.inv_ss, fast};
    const __half *va = vk + (size_t)p.i0 * o.ld_vk;
    rc = (p.P == 1) ? launch_epig2(va, o.ld_vk, tk, o.ld_tk, p.n, (int)p.BK, Kd, gr, st,
                                           "milnce backward row block")
                    : launch_gemm<tc::EpiG>(va, o.ld_vk, tk, o.ld_tk, p.n, (int)p.BK, Kd, p.chunk_s, gr, st,
                                            "milnce backward row block");
    if (rc) return rc;
    // dvideo (n, D) = G . T      : A = G (n, P*BK), Bm = Tt (D, P*BK)
    tc::EpiStore::Params sv{dvideo, (long long)p.D, alpha0, grad_loss, o.s_text, grad_bf16};
    // n-fastest tile order: the CTAs in flight share G rows through L2 (G was re-read from HBM once
    // per 128-column block of D before: 2.2 GB instead of 0.54 GB for the dtext GEMM).
    rc = (p.P == 1) ? launch_gemm2<tc::EpiStore, false, true, false, true>(g, p.ld_g, tk, o.ld_tk, p.n, p.D, (int)p.BK, sv, st,
                                            "milnce dvideo GEMM")
                    : launch_gemm<tc::EpiStore, false, 1, true>(g, p.ld_g, tt, p.ld_tt, p.n, p.D,
                                                                (int)(p.P * p.BK), p.chunk_g, sv, st,
                                                                "milnce dvideo GEMM");
    if (rc) return rc;
    tc::EpiStore::Params sd{dtext, (long long)p.D, alpha0, grad_loss, o.s_video, grad_bf16};
    if (p.P == 1) {
        // dtext (n*K, D) = G[:, own columns]^T . V, reading G as an MN-major A operand:
        // nobody materialises G^T.  On one GPU (n == B) the G of the row block is the G needed.
        const __half *g2 = g;
        long long ld_g2 = p.ld_g;
        if (p.n != p.B) {
            // column block: G2 (B, n*K): row <-> sample r, column c <-> sample (i0*K + c) / K
            tc::EpiG::Params gc{den, nom, gt, p.ld_g2, (long long)p.n * p.K, 1, 0, tc::make_fastdiv(1), p.i0 * p.K, tc::make_fastdiv(p.K),
                                o.inv_ss, fast};
            rc = launch_epig2(vk, o.ld_vk, tk + (size_t)p.i0 * p.K * o.ld_tk, o.ld_tk, p.B,
                                        p.n * p.K, Kd, gc, st, "milnce backward column block");
            if (rc) return rc;
            g2 = gt;
            ld_g2 = p.ld_g2;
        }
        rc = launch_gemm2<tc::EpiStore, true, true, false, true>(g2, ld_g2, vk, o.ld_vk, p.n * p.K, p.D, p.B, sd, st,
                                    "milnce dtext GEMM (MN-major G)");
        return rc;
    }
    // fp32 inputs (hi/lo planes).  Single-GPU case with B a multiple of 64: read the planes of G
    // ([hi | lo | hi] along its columns) as an MN-major A operand, plane by plane along the
    // reduction, against Vt ([hi | hi | lo] along ITS reduction index): no G^T, no second S pass.
    if (p.n == p.B && p.B % tc::BLOCK_K == 0) {
        return launch_gemm<tc::EpiStore, true, 1, true>(g, p.ld_g, vt, p.ld_vt, (int)p.BK, p.D, p.P * p.B,
                                                        p.chunk_g, sd, st, "milnce dtext GEMM (MN-major planes)",
                                                        p.B, (int)p.BK);
    }
    // otherwise the transposed problem S^T = T . V^T yields G^T (n*K, P*B) row-major
    tc::EpiG::Params gc{den, nom, gt, p.ld_gt, (long long)p.B, p.P, p.i0 * p.K, tc::make_fastdiv(p.K), 0, tc::make_fastdiv(1), o.inv_ss,
                        fast};
    rc = launch_gemm<tc::EpiG>(tk + (size_t)p.i0 * p.K * o.ld_tk, o.ld_tk, vk, o.ld_vk, p.n * p.K,
                               p.B, Kd, p.chunk_s, gc, st, "milnce backward transposed block");
    if (rc) return rc;
    // dtext (n*K, D) = G^T . V   : A = Gt (n*K, P*B), Bm = Vt (D, P*B)
    rc = launch_gemm<tc::EpiStore, false, 1, true>(gt, p.ld_gt, vt, p.ld_vt, p.n * p.K, p.D, p.P * p.B,
                                                   p.chunk_g, sd, st, "milnce dtext GEMM");
    return rc;
}

// ---------------------------------------------------------------------------------------------
// Rows-only sharding (bf16 inputs): a rank computes ONLY the rows of S it owns.
//   forward : row_lse of its samples (complete) and, for EVERY sample j, the partial column
//             statistic over its own rows; ranks exchange those (B floats each) and form den.
//   backward: G[own rows, :] -> dvideo (own rows, complete) and a (B*K, D) PARTIAL dtext that the
//             caller reduce-scatters.  4/W GEMM units per rank instead of 6/W for the row+column
//             block scheme of milnce_forward/backward with n < B.
// ---------------------------------------------------------------------------------------------
// statistics GEMM over the owned rows + the two merges (everything of the rows-only forward after staging)
int forward_shard_tail(const Plan &p, float *row_lse, float *col_part, uint8_t *ws, cudaStream_t st) {
    const Operands o = operands(p, ws);
    const __half *vk = o.vk, *tk = o.tk;
    float *pr = reinterpret_cast<float *>(ws + p.off_pr);
    float *pc = reinterpret_cast<float *>(ws + p.off_pc);
    int rc = launch_stats2(p.cg, o.inv_ss, pr, (long long)p.n, pc, p.BK / p.cg, vk + (size_t)p.i0 * o.ld_vk,
                           o.ld_vk, tk, o.ld_tk, p.n, (int)p.BK, p.D, st, "milnce forward (rows-only shard)");
    if (rc) return rc;
    const MergeFin none{nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0, nullptr};
    merge_partials_kernel<<<(p.n + 31) / 32, 1024, 0, st>>>(pr, nullptr, 1, p.n, p.nb_r, 0, row_lse + p.i0,
                                                            none);
    merge_partials_kernel<<<(p.B + 31) / 32, 1024, 0, st>>>(nullptr, pc, p.K / p.cg, p.B, 0, p.nb_cl, col_part,
                                                            none);
    b200::note_launches(2);
    B200_LAUNCH_OK("milnce shard finalize");
    return 0;
}

int forward_shard_impl(const Plan &p, const __nv_bfloat16 *video, const __nv_bfloat16 *text,
                       float *row_lse, float *col_part, float *nom, uint8_t *ws, cudaStream_t st) {
    int rc = stage_inputs<__nv_bfloat16>(p, video, text, ws, st, nom);
    if (rc) return rc;
    return forward_shard_tail(p, row_lse, col_part, ws, st);
}

// Split-K for the sharded dvideo GEMM.  Its output has only (n/256) x (D/256) pair tiles - 32 at two ranks,
// 8 at eight, against 74 CTA pairs - while the reduction runs over all B*K text rows, so without a split its time
// does not go down with the rank count at all.  Slices: as many as keep one wave of (slice, tile) units within the
// number of pairs, at least 8 k-blocks each, every slice non-empty, partial planes inside `cap_bytes` of scratch.
int dvideo_splits(int n, int D, long long Kd, size_t cap_bytes) {
    const int tiles = ((n + tc::PAIR_M - 1) / tc::PAIR_M) * ((D + tc::PAIR_N - 1) / tc::PAIR_N);
    const int pairs = num_sms() / 2;
    const int k_iters = (int)((Kd + tc::BLOCK_K - 1) / tc::BLOCK_K);
    int s = pairs / std::max(tiles, 1);
    s = std::min(s, k_iters / 8);
    s = std::min<long long>(s, (long long)(cap_bytes / ((size_t)n * D * 4)));
    if (s < 2) return 1;
    const int kps = (k_iters + s - 1) / s;
    s = (k_iters + kps - 1) / kps;      // drop slices the rounding left empty
    return s < 2 ? 1 : s;
}

// out (rows x cols, float32 or bfloat16) = sum of `splits` float32 planes, added in slice order
template <bool OUT_BF16>
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float4 *__restrict__ part, int splits, size_t count4, void *__restrict__ out) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < count4; g += stride) {
        float4 a = part[g];
        for (int k = 1; k < splits; ++k) {
            const float4 b = part[(size_t)k * count4 + g];
            a.x += b.x;
            a.y += b.y;
            a.z += b.z;
            a.w += b.w;
        }
        if (OUT_BF16) {
            __nv_bfloat162 h[2] = {__floats2bfloat162_rn(a.x, a.y), __floats2bfloat162_rn(a.z, a.w)};
            reinterpret_cast<uint2 *>(out)[g] = *reinterpret_cast<uint2 *>(h);
        } else {
            reinterpret_cast<float4 *>(out)[g] = a;
        }
    }
}

// dvideo (n, D) = G[own rows, :] . T of the rows-only scheme, with split-K when the rank's share is small
int dvideo_shard(const Plan &p, const Operands &o, const __half *g, const float *grad_loss, void *dvideo, int grad_bf16,
                 uint8_t *ws, cudaStream_t st) {
    const size_t cap = p.n != p.B ? (size_t)p.B * p.ld_g2 * 2 : 0;   // the G2 region: unused by the rows-only scheme
    const int splits = (p.D % 4 == 0) ? dvideo_splits(p.n, p.D, p.BK, cap) : 1;
    const float alpha = 1.f / ((float)p.B * tc::G_SCALE);
    if (splits <= 1) {
        tc::EpiStore::Params sv{dvideo, (long long)p.D, alpha, grad_loss, o.s_text, grad_bf16};
        return launch_gemm2<tc::EpiStore, false, true, false, true>(g, p.ld_g, o.tk, o.ld_tk, p.n, p.D, (int)p.BK, sv, st,
                                                                    "milnce dvideo GEMM (shard)");
    }
    float *part = reinterpret_cast<float *>(ws + p.off_gt);
    tc::EpiStore::Params sv{part, (long long)p.D, alpha, grad_loss, o.s_text, 0, (long long)p.n * p.D};
    int rc = launch_gemm2<tc::EpiStore, false, true, false, true>(g, p.ld_g, o.tk, o.ld_tk, p.n, p.D, (int)p.BK, sv, st,
                                                                  "milnce dvideo GEMM (shard, split-K)", 0, splits);
    if (rc) return rc;
    const size_t count4 = (size_t)p.n * p.D / 4;
    const int blocks = (int)std::min<size_t>((count4 + 255) / 256, (size_t)8 * num_sms());
    if (grad_bf16)
        splitk_reduce_kernel<true><<<blocks, 256, 0, st>>>(reinterpret_cast<const float4 *>(part), splits, count4, dvideo);
    else
        splitk_reduce_kernel<false><<<blocks, 256, 0, st>>>(reinterpret_cast<const float4 *>(part), splits, count4, dvideo);
    b200::note_launches(1);
    B200_LAUNCH_OK("milnce dvideo split-K reduce");
    return 0;
}

int backward_shard_impl(const Plan &p, const __nv_bfloat16 *video, const __nv_bfloat16 *text,
                        const float *den, const float *nom, const float *grad_loss, void *dvideo,
                        int grad_bf16, void *dtext_partial, int dtext_bf16, int parts, int reuse, uint8_t *ws,
                        cudaStream_t st, void *const *owner_dst = nullptr, int owners = 0) {
    // parts: 1 = G and the partial dtext, 2 = dvideo (from the G a parts-1 call left in the
    // workspace), 3 = both.  Split calls let the caller start the reduce-scatter of dtext
    // while dvideo is still being computed.
    int rc;
    const Operands o = operands(p, ws);
    const __half *vk = o.vk, *tk = o.tk;
    __half *g = reinterpret_cast<__half *>(ws + p.off_g);
    Scalars *sc = reinterpret_cast<Scalars *>(ws + p.off_scal);
    if (parts == 2) return dvideo_shard(p, o, g, grad_loss, dvideo, grad_bf16, ws, st);
    if (!reuse) {
        rc = stage_inputs<__nv_bfloat16>(p, video, text, ws, st, nullptr);
        if (rc) return rc;
    }
    B200_CUDA_OK(b200::launch_pdl(scalars_backward_kernel, dim3(1), dim3(1024), 0, st, sc, grad_loss, den, p.B));
    const tc::GFast fast{&sc->g_fast, &sc->den_mid_log2};
    tc::EpiG::Params gr{den, nom, g, p.ld_g, p.BK, 1, p.i0, tc::make_fastdiv(1), 0, tc::make_fastdiv(p.K), o.inv_ss, fast};
    rc = launch_epig2(vk + (size_t)p.i0 * o.ld_vk, o.ld_vk, tk, o.ld_tk, p.n, (int)p.BK, p.D, gr,
                                st, "milnce backward (rows-only shard) G");
    if (rc) return rc;
    // partial dtext over the owned rows: (B*K, D) = G^T . V[own rows];  A = G read MN-major
    tc::EpiStore::Params sd{dtext_partial, (long long)p.D, 1.f / ((float)p.B * tc::G_SCALE), grad_loss, o.s_video, dtext_bf16};
    if (owner_dst) {
        // the rows of the partial that rank q owns go straight to rank q's buffer (push over NVLink from the
        // epilogue's stores): the reduce-scatter's transfer rides under this GEMM
        sd.seg_rows = (int)(p.BK / owners);
        for (int q = 0; q < 16; ++q) sd.seg[q] = owner_dst[q < owners ? q : 0];
    }
    rc = launch_gemm2<tc::EpiStore, true, true, false, true>(g, p.ld_g, vk + (size_t)p.i0 * o.ld_vk, o.ld_vk, (int)p.BK, p.D, p.n, sd, st,
                                                "milnce partial dtext GEMM (shard)");
    if (rc || parts == 1) return rc;
    return dvideo_shard(p, o, g, grad_loss, dvideo, grad_bf16, ws, st);
}

int check_shard(const Plan &p, int dtype) {
    if (dtype != B200HEAD_BF16) {
        b200::set_error("rows-only sharding is implemented for bfloat16 inputs; use "
                        "milnce_forward/milnce_backward with a range for float32");
        return B200HEAD_ERR_ARG;
    }
    if (p.i0 % 8 != 0 || p.D % 8 != 0) {
        b200::set_error("rows-only sharding needs i0 and D to be multiples of 8 (TMA alignment)");
        return B200HEAD_ERR_ARG;
    }
    return 0;
}

}  // namespace

extern "C" size_t milnce_workspace_bytes(int B, int K, int D, int dtype, int n) {
    Plan p;
    if (!make_plan(p, B, K, D, dtype, 0, n)) return 0;
    return p.total;
}

extern "C" int milnce_forward(const void *video, const void *text, int B, int K, int D, int dtype,
                              int i0, int n, float *den, float *nom, float *loss, void *workspace,
                              size_t workspace_bytes, void *stream) {
    Plan p;
    if (!make_plan(p, B, K, D, dtype, i0, n)) {
        b200::set_error("milnce_forward: bad shape/range B=%d K=%d D=%d dtype=%d i0=%d n=%d", B, K, D,
                        dtype, i0, n);
        return B200HEAD_ERR_ARG;
    }
    if (!video || !text || !den || !nom || !loss) {
        b200::set_error("milnce_forward: null pointer argument");
        return B200HEAD_ERR_ARG;
    }
    int rc = b200::check_device();
    if (rc) return rc;
    rc = check_common(p, workspace, workspace_bytes);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t *ws = (uint8_t *)workspace;
    if (dtype == B200HEAD_F32)
        return forward_impl<float>(p, (const float *)video, (const float *)text, den, nom, loss, ws, st);
    return forward_impl<__nv_bfloat16>(p, (const __nv_bfloat16 *)video, (const __nv_bfloat16 *)text,
                                       den, nom, loss, ws, st);
}

extern "C" int milnce_backward(const void *video, const void *text, const float *den,
                               const float *nom, const float *grad_loss, void *dvideo,
                               void *dtext, int grad_dtype, int B, int K, int D, int dtype, int i0, int n,
                               int reuse_forward_ws, void *workspace, size_t workspace_bytes,
                               void *stream) {
    Plan p;
    if (!make_plan(p, B, K, D, dtype, i0, n)) {
        b200::set_error("milnce_backward: bad shape/range B=%d K=%d D=%d dtype=%d i0=%d n=%d", B, K,
                        D, dtype, i0, n);
        return B200HEAD_ERR_ARG;
    }
    if (!video || !text || !den || !nom || !grad_loss || !dvideo || !dtext) {
        b200::set_error("milnce_backward: null pointer argument");
        return B200HEAD_ERR_ARG;
    }
    if (grad_dtype != B200HEAD_F32 && grad_dtype != B200HEAD_BF16) {
        b200::set_error("milnce_backward: grad_dtype must be B200HEAD_F32 or B200HEAD_BF16");
        return B200HEAD_ERR_ARG;
    }
    const int gb = (grad_dtype == B200HEAD_BF16) ? 1 : 0;
    int rc = b200::check_device();
    if (rc) return rc;
    rc = check_common(p, workspace, workspace_bytes);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t *ws = (uint8_t *)workspace;
    if (dtype == B200HEAD_F32)
        return backward_impl<float>(p, (const float *)video, (const float *)text, den, nom, grad_loss,
                                    dvideo, dtext, gb, reuse_forward_ws, ws, st);
    return backward_impl<__nv_bfloat16>(p, (const __nv_bfloat16 *)video, (const __nv_bfloat16 *)text,
                                        den, nom, grad_loss, dvideo, dtext, gb, reuse_forward_ws, ws, st);
}

#ifdef B200_EPI_STAMP
extern "C" int b200head_debug_stamps(long long *host, size_t count) {
    const size_t have = sizeof(tc::g_stamp) / sizeof(long long);
    B200_CUDA_OK(cudaDeviceSynchronize());
    B200_CUDA_OK(cudaMemcpyFromSymbol(host, tc::g_stamp, sizeof(long long) * (count < have ? count : have)));
    return 0;
}
#endif

extern "C" int b200head_gemm_tn_f16(const void *A, const void *Bm, float *C, int M, int N, int Kd,
                                    int lda, int ldb, int ldc, float alpha, void *stream) {
    if (!A || !Bm || !C || M <= 0 || N <= 0 || Kd <= 0) {
        b200::set_error("b200head_gemm_tn_f16: bad argument");
        return B200HEAD_ERR_ARG;
    }
    int rc = b200::check_device();
    if (rc) return rc;
    tc::EpiStore::Params ep{C, (long long)ldc, alpha, nullptr, nullptr, 0};
    return launch_gemm<tc::EpiStore>(A, lda, Bm, ldb, M, N, Kd, 16, ep, (cudaStream_t)stream,
                                     "gemm_tn_f16 test hook");
}

extern "C" int b200head_gemm_at_f16(const void *At, const void *Bm, float *C, int M, int N, int Kd,
                                    int ldat, int ldb, int ldc, float alpha, void *stream) {
    if (!At || !Bm || !C || M <= 0 || N <= 0 || Kd <= 0) {
        b200::set_error("b200head_gemm_at_f16: bad argument");
        return B200HEAD_ERR_ARG;
    }
    int rc = b200::check_device();
    if (rc) return rc;
    tc::EpiStore::Params ep{C, (long long)ldc, alpha, nullptr, nullptr, 0};
    return launch_gemm<tc::EpiStore, true>(At, ldat, Bm, ldb, M, N, Kd, 16, ep, (cudaStream_t)stream,
                                           "gemm_at_f16 test hook");
}

extern "C" int b200head_gemm2_f16(const void *A, const void *Bm, float *C, int M, int N, int Kd, int lda,
                                  int ldb, int ldc, float alpha, int a_transposed, void *stream) {
    if (!A || !Bm || !C || M <= 0 || N <= 0 || Kd <= 0) {
        b200::set_error("b200head_gemm2_f16: bad argument");
        return B200HEAD_ERR_ARG;
    }
    int rc = b200::check_device();
    if (rc) return rc;
    tc::EpiStore::Params ep{C, (long long)ldc, alpha, nullptr, nullptr, 0};
    // a_transposed: bit 0 = A passed as (Kd, M), bit 1 = Bm passed as (Kd, N), bit 2 = BOTH operands hold
    // bfloat16 instead of float16 (one format per instruction: mixing them faults on sm_100a)
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t fmt = (a_transposed & 4) ? (tc::IDESC_A_BF16 | tc::IDESC_B_BF16) : 0u;
    switch (a_transposed & 3) {
        case 0: return launch_gemm2<tc::EpiStore, false, false>(A, lda, Bm, ldb, M, N, Kd, ep, st, "gemm2 test hook", fmt);
        case 1: return launch_gemm2<tc::EpiStore, true, true>(A, lda, Bm, ldb, M, N, Kd, ep, st, "gemm2 test hook", fmt);
        case 2: return launch_gemm2<tc::EpiStore, false, true, false, true>(A, lda, Bm, ldb, M, N, Kd, ep, st, "gemm2 test hook", fmt);
        default: return launch_gemm2<tc::EpiStore, true, true, false, true>(A, lda, Bm, ldb, M, N, Kd, ep, st, "gemm2 test hook", fmt);
    }
}

extern "C" int milnce_forward_shard(const void *video, const void *text, int B, int K, int D, int dtype,
                                    int i0, int n, float *row_lse, float *col_part, float *nom,
                                    void *workspace, size_t workspace_bytes, void *stream) {
    Plan p;
    if (!make_plan(p, B, K, D, dtype == B200HEAD_BF16 ? dtype : B200HEAD_BF16, i0, n)) {
        b200::set_error("milnce_forward_shard: bad shape/range B=%d K=%d D=%d i0=%d n=%d", B, K, D, i0, n);
        return B200HEAD_ERR_ARG;
    }
    if (!video || !text || !row_lse || !col_part || !nom) {
        b200::set_error("milnce_forward_shard: null pointer argument");
        return B200HEAD_ERR_ARG;
    }
    int rc = check_shard(p, dtype);
    if (rc) return rc;
    rc = b200::check_device();
    if (rc) return rc;
    rc = check_common(p, workspace, workspace_bytes);
    if (rc) return rc;
    return forward_shard_impl(p, (const __nv_bfloat16 *)video, (const __nv_bfloat16 *)text, row_lse,
                              col_part, nom, (uint8_t *)workspace, (cudaStream_t)stream);
}

extern "C" size_t milnce_peer_slot_bytes(int b, int K, int D) {
    if (b <= 0 || K <= 0 || D <= 0) return 0;
    return align_up(PEER_HEADER + (size_t)b * (1 + K) * D * 2, 256);
}

extern "C" int milnce_peer_publish(const void *video_local, const void *text_local, int b, int K, int D, void *slot,
                                   float *nom_own, void *stream) {
    if (!video_local || !text_local || !slot || !nom_own || b <= 0 || K <= 0 || D <= 0) {
        b200::set_error("milnce_peer_publish: bad argument");
        return B200HEAD_ERR_ARG;
    }
    if (D % 8 != 0 || ((((uintptr_t)video_local) | ((uintptr_t)text_local) | ((uintptr_t)slot)) & 15)) {
        b200::set_error("milnce_peer_publish: D must be a multiple of 8 and all pointers 16-byte aligned");
        return B200HEAD_ERR_ARG;
    }
    int rc = b200::check_device();
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    B200_CUDA_OK(cudaMemsetAsync(slot, 0, PEER_HEADER, st));
    const int wpb = 8;
    const int blocks = std::min((b + wpb - 1) / wpb, 8 * num_sms());
    peer_publish_kernel<<<blocks, wpb * 32, 0, st>>>((const __nv_bfloat16 *)video_local, (const __nv_bfloat16 *)text_local,
                                                     b, K, D, (uint8_t *)slot, nom_own);
    b200::note_launches(1);
    B200_LAUNCH_OK("milnce_peer_publish");
    return 0;
}

extern "C" int milnce_forward_shard_peers(const void *const *slots, int W, int B, int K, int D, int i0, int n,
                                          float *row_lse, float *col_part, void *workspace, size_t workspace_bytes,
                                          void *stream) {
    Plan p;
    if (!make_plan(p, B, K, D, B200HEAD_BF16, i0, n)) {
        b200::set_error("milnce_forward_shard_peers: bad shape/range B=%d K=%d D=%d i0=%d n=%d", B, K, D, i0, n);
        return B200HEAD_ERR_ARG;
    }
    if (!slots || !row_lse || !col_part || W < 1 || W > PEER_MAX || B % W != 0) {
        b200::set_error("milnce_forward_shard_peers: bad argument (1 <= W <= %d ranks, B a multiple of W)", PEER_MAX);
        return B200HEAD_ERR_ARG;
    }
    const int b = B / W;
    if (b % 64 != 0 || D % 8 != 0 || n != b || i0 % b != 0) {
        b200::set_error("milnce_forward_shard_peers: the local batch must be a multiple of 64, D of 8, and the "
                        "range one rank's shard");
        return B200HEAD_ERR_ARG;
    }
    int rc = b200::check_device();
    if (rc) return rc;
    rc = check_common(p, workspace, workspace_bytes);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t *ws = (uint8_t *)workspace;
    Scalars *sc = reinterpret_cast<Scalars *>(ws + p.off_scal);
    B200_CUDA_OK(cudaMemsetAsync(sc, 0, sizeof(Scalars), st));
    PeerSlots ps;
    for (int r = 0; r < PEER_MAX; ++r) ps.slot[r] = (const uint8_t *)slots[r < W ? r : 0];
    __half *vk = reinterpret_cast<__half *>(ws + p.off_vk), *tk = reinterpret_cast<__half *>(ws + p.off_tk);
    const int rank = i0 / b;
    const int grid = 4 * num_sms();
    // (ld_vk == ld_tk == D since D % 8 == 0)
    // (Tried: staging the own shard first and overlapping the other ranks' copies, on a side stream, with the
    // similarity pass over the own column block.  The copy kernel did not get onto the SMs next to the
    // one-CTA-per-SM GEMM, whichever was enqueued first and with 128-thread copy blocks, so the phases ran back
    // to back with two extra launches: 209-214 us against 173 us for this form at W = 2.)
    stage_peers_kernel<<<grid, 256, 0, st>>>(ps, W, rank, b, K, D, vk, tk, sc);
    b200::note_launches(1);
    B200_LAUNCH_OK("milnce_forward_shard_peers staging");
    return forward_shard_tail(p, row_lse, col_part, ws, st);
}

extern "C" int milnce_backward_shard_push(const float *den, const float *nom, const float *grad_loss,
                                          void *const *owner_dst, int W, int B, int K, int D, int i0, int n,
                                          void *workspace, size_t workspace_bytes, void *stream) {
    Plan p;
    if (!make_plan(p, B, K, D, B200HEAD_BF16, i0, n)) {
        b200::set_error("milnce_backward_shard_push: bad shape/range B=%d K=%d D=%d i0=%d n=%d", B, K, D, i0, n);
        return B200HEAD_ERR_ARG;
    }
    if (!den || !nom || !grad_loss || !owner_dst || W < 1 || W > PEER_MAX || B % W != 0) {
        b200::set_error("milnce_backward_shard_push: bad argument (1 <= W <= %d ranks, B a multiple of W)", PEER_MAX);
        return B200HEAD_ERR_ARG;
    }
    if (((long long)(B / W) * K) % 128 != 0) {
        b200::set_error("milnce_backward_shard_push: a rank's b*K text rows must be a multiple of 128");
        return B200HEAD_ERR_ARG;
    }
    for (int q = 0; q < W; ++q)
        if (!owner_dst[q] || (((uintptr_t)owner_dst[q]) & 15)) {
            b200::set_error("milnce_backward_shard_push: destination %d is null or not 16-byte aligned", q);
            return B200HEAD_ERR_ARG;
        }
    int rc = check_shard(p, B200HEAD_BF16);
    if (rc) return rc;
    rc = b200::check_device();
    if (rc) return rc;
    rc = check_common(p, workspace, workspace_bytes);
    if (rc) return rc;
    return backward_shard_impl(p, nullptr, nullptr, den, nom, grad_loss, nullptr, 0, nullptr, 1, /*parts=*/1,
                               /*reuse=*/1, (uint8_t *)workspace, (cudaStream_t)stream, owner_dst, W);
}

extern "C" int milnce_peer_reduce_dtext(const void *const *partials, int W, int rank, int b, int K, int D,
                                        void *dtext, int dtext_dtype, void *stream) {
    if (!partials || !dtext || W < 1 || W > PEER_MAX || rank < 0 || rank >= W || b <= 0 || K <= 0 || D <= 0 ||
        (dtext_dtype != B200HEAD_F32 && dtext_dtype != B200HEAD_BF16)) {
        b200::set_error("milnce_peer_reduce_dtext: bad argument (1 <= W <= %d ranks, dtext_dtype F32 or BF16)", PEER_MAX);
        return B200HEAD_ERR_ARG;
    }
    uintptr_t bits = (uintptr_t)dtext;
    for (int r = 0; r < W; ++r) bits |= (uintptr_t)partials[r];
    if (((size_t)b * K * D) % 8 != 0 || (bits & 15)) {
        b200::set_error("milnce_peer_reduce_dtext: b*K*D must be a multiple of 8 and all pointers 16-byte aligned");
        return B200HEAD_ERR_ARG;
    }
    int rc = b200::check_device();
    if (rc) return rc;
    PeerSlots ps;
    for (int r = 0; r < PEER_MAX; ++r) ps.slot[r] = (const uint8_t *)partials[r < W ? r : 0];
    const size_t count = (size_t)b * K * D / 8;
    const int blocks = (int)std::min<size_t>((count + 255) / 256, (size_t)16 * num_sms());
    cudaStream_t st = (cudaStream_t)stream;
    if (dtext_dtype == B200HEAD_BF16)
        peer_reduce_kernel<true><<<blocks, 256, 0, st>>>(ps, W, (size_t)rank * count, count, dtext);
    else
        peer_reduce_kernel<false><<<blocks, 256, 0, st>>>(ps, W, (size_t)rank * count, count, dtext);
    b200::note_launches(1);
    B200_LAUNCH_OK("milnce_peer_reduce_dtext");
    return 0;
}

extern "C" int milnce_shard_combine(const float *gathered, int W, int b, float *den, float *nom,
                                    float *loss, void *stream) {
    if (!gathered || !den || !nom || !loss || W <= 0 || b <= 0) {
        b200::set_error("milnce_shard_combine: bad argument");
        return B200HEAD_ERR_ARG;
    }
    int rc = b200::check_device();
    if (rc) return rc;
    if (W > PEER_MAX) {
        b200::set_error("milnce_shard_combine: at most %d ranks", PEER_MAX);
        return B200HEAD_ERR_ARG;
    }
    PeerSlots ps;
    for (int r = 0; r < PEER_MAX; ++r)
        ps.slot[r] = reinterpret_cast<const uint8_t *>(gathered + (size_t)(r < W ? r : 0) * ((size_t)2 * b + (size_t)W * b));
    shard_combine_kernel<<<1, 512, 0, (cudaStream_t)stream>>>(ps, W, b, W * b, den, nom, loss, nullptr, nullptr);
    b200::note_launches(1);
    B200_LAUNCH_OK("milnce_shard_combine");
    return 0;
}

extern "C" size_t milnce_shard_combine_scratch_bytes(int W, int b) {
    if (W <= 0 || b <= 0) return 0;
    return 256 + 8 * (((size_t)W * b + 511) / 512);
}

extern "C" int milnce_shard_combine_peers(const void *const *packs, int W, int b, float *den, float *nom, float *loss,
                                          void *scratch, void *stream) {
    if (!packs || !den || !nom || !loss || W <= 0 || W > PEER_MAX || b <= 0 || (((uintptr_t)scratch) & 255)) {
        b200::set_error("milnce_shard_combine_peers: bad argument (1 <= W <= %d ranks)", PEER_MAX);
        return B200HEAD_ERR_ARG;
    }
    for (int r = 0; r < W; ++r)
        if (!packs[r] || (((uintptr_t)packs[r]) & 3)) {
            b200::set_error("milnce_shard_combine_peers: pack %d is null or misaligned", r);
            return B200HEAD_ERR_ARG;
        }
    int rc = b200::check_device();
    if (rc) return rc;
    PeerSlots ps;
    for (int r = 0; r < PEER_MAX; ++r) ps.slot[r] = (const uint8_t *)packs[r < W ? r : 0];
    const int B = W * b;
    const int blocks = scratch ? (B + 511) / 512 : 1;
    shard_combine_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>(
        ps, W, b, B, den, nom, loss, reinterpret_cast<unsigned int *>(scratch),
        scratch ? reinterpret_cast<double *>(reinterpret_cast<uint8_t *>(scratch) + 256) : nullptr);
    b200::note_launches(1);
    B200_LAUNCH_OK("milnce_shard_combine_peers");
    return 0;
}

extern "C" int milnce_backward_shard(const void *video, const void *text, const float *den,
                                     const float *nom, const float *grad_loss, void *dvideo,
                                     int dvideo_dtype, void *dtext_partial, int dtext_dtype, int parts, int B, int K, int D,
                                     int dtype, int i0, int n, int reuse_forward_ws, void *workspace,
                                     size_t workspace_bytes, void *stream) {
    Plan p;
    if (!make_plan(p, B, K, D, dtype == B200HEAD_BF16 ? dtype : B200HEAD_BF16, i0, n)) {
        b200::set_error("milnce_backward_shard: bad shape/range B=%d K=%d D=%d i0=%d n=%d", B, K, D, i0, n);
        return B200HEAD_ERR_ARG;
    }
    if (parts < 1 || parts > 3 || (dvideo_dtype != B200HEAD_F32 && dvideo_dtype != B200HEAD_BF16) ||
        (dtext_dtype != B200HEAD_F32 && dtext_dtype != B200HEAD_BF16)) {
        b200::set_error("milnce_backward_shard: parts must be 1, 2 or 3 and dvideo_dtype / dtext_dtype F32 or BF16");
        return B200HEAD_ERR_ARG;
    }
    if (!video || !text || !den || !nom || !grad_loss || ((parts & 2) && !dvideo) ||
        ((parts & 1) && !dtext_partial)) {
        b200::set_error("milnce_backward_shard: null pointer argument");
        return B200HEAD_ERR_ARG;
    }
    int rc = check_shard(p, dtype);
    if (rc) return rc;
    rc = b200::check_device();
    if (rc) return rc;
    rc = check_common(p, workspace, workspace_bytes);
    if (rc) return rc;
    return backward_shard_impl(p, (const __nv_bfloat16 *)video, (const __nv_bfloat16 *)text, den, nom,
                               grad_loss, dvideo, dvideo_dtype == B200HEAD_BF16 ? 1 : 0, dtext_partial,
                               dtext_dtype == B200HEAD_BF16 ? 1 : 0, parts,
                               reuse_forward_ws, (uint8_t *)workspace, (cudaStream_t)stream);
}
