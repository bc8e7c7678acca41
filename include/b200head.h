/* b200head.h - C ABI of the B200-native MIL-NCE / soft-DTW loss head.
 *
 * The reference (KoDohwan/MIL-NCE_HowTo100M) is pure Python and has no FFI
 * layer; its boundary for this path is the Python module API.  Each entry
 * point below is what a binding for that API calls, and cites the reference
 * code it replaces.  The Python mirror of the reference modules lives in
 * mil-nce_howto100m_b200/{loss,soft_dtw_cuda,utils}.py and calls these through
 * ctypes; INTEGRATION.md shows the stub a reference maintainer would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name starts with host_;
 *     the caller allocates everything, the library allocates no device memory;
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*), with
 *     no host synchronisation inside any call;
 *   - tensors are dense row-major; return value 0 = ok, negative = error and
 *     b200head_last_error() (thread-local) explains it;
 *   - kernels are compiled for sm_100a only; on any other device the calls
 *     fail with B200HEAD_ERR_DEVICE.  There is no CPU fallback.
 */
#ifndef B200HEAD_H
#define B200HEAD_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200HEAD_OK 0
#define B200HEAD_ERR_ARG (-1)     /* bad argument / unsupported shape            */
#define B200HEAD_ERR_CUDA (-2)    /* a CUDA runtime or driver call failed        */
#define B200HEAD_ERR_DEVICE (-3)  /* current device is not compute capability 10 */
#define B200HEAD_ERR_WS (-4)      /* workspace too small or misaligned           */

#define B200HEAD_F32 0
#define B200HEAD_BF16 1

#define B200HEAD_DIST_COSINE 0        /* exp(1 - cos)   soft_dtw_cuda.py:338-348 */
#define B200HEAD_DIST_NEGATIVE_DOT 1  /* -<x,y>         soft_dtw_cuda.py:351-363 */
#define B200HEAD_DIST_EUCLIDEAN 2     /* exp(||x-y||)   soft_dtw_cuda.py:326-335 */

/* Message for the last failing call on this host thread ("" if none). */
const char *b200head_last_error(void);
/* ABI version (bumped when a signature changes). */
int b200head_version(void);
/* Number of kernels this library has launched in this process so far (all threads, all
 * devices).  No reference counterpart; bench.py reports the difference over its timed region
 * as `gpu_launches`.  Kernels replayed from a captured CUDA graph are not seen here: count them
 * as (launches made during the capture) x (replays).
 * The library reads no environment variables and has no run-time switches: kernel variants used
 * for A/B timing exist only as compile-time macros of variant builds (profiles/build_variant.sh). */
unsigned long long b200head_launch_count(void);

/* ------------------------------------------------------------------------
 * MIL-NCE loss: replaces MILNCELoss.forward (reference loss.py:10-18) and
 * its autograd backward.  video is (B, D), text is (B*K, D), sample-major
 * (row j*K+k = candidate k of sample j; reference train.py:212-213).
 *
 * The range [i0, i0+n) selects the samples this call owns.  n == B is the
 * single-GPU case.  With n < B a rank computes exactly the slice that the
 * reference's AllGather.backward (utils.py:20-24) would hand back to it:
 * the caller all-gathers video/text first (replacing utils.py:11-17), calls
 * forward with its range, all-gathers `den`, then calls backward.
 * ------------------------------------------------------------------------ */

/* Bytes of scratch needed by milnce_forward/milnce_backward for this shape. */
size_t milnce_workspace_bytes(int B, int K, int D, int dtype, int n);

/* Forward.  den[B], nom[B]: entries [i0, i0+n) are written.
 * loss[0] receives sum_{i in range} (den_i - nom_i) / B  (so the sum over
 * ranks is the reference's mean, loss.py:18).  */
int milnce_forward(const void *video, const void *text, int B, int K, int D, int dtype,
                   int i0, int n, float *den, float *nom, float *loss,
                   void *workspace, size_t workspace_bytes, void *stream);

/* Backward.  den must hold all B entries (all-gathered when sharded); nom
 * needs entries [i0, i0+n).  grad_loss[0] is d(objective)/d(loss), read on
 * the device.  Outputs: dvideo (n, D) = rows [i0, i0+n) of d loss/d video,
 * dtext (n*K, D) = rows [i0*K, (i0+n)*K) of d loss/d text, written as
 * grad_dtype (B200HEAD_F32 or B200HEAD_BF16: autograd hands back gradients in
 * the dtype of the inputs, so a bf16 caller gets them without a conversion pass).
 * reuse_forward_ws != 0 promises that `workspace` is untouched since a
 * milnce_forward call with the same inputs (skips re-staging the operands). */
int milnce_backward(const void *video, const void *text, const float *den, const float *nom,
                    const float *grad_loss, void *dvideo, void *dtext, int grad_dtype,
                    int B, int K, int D, int dtype, int i0, int n, int reuse_forward_ws,
                    void *workspace, size_t workspace_bytes, void *stream);

/* Rows-only sharding (bfloat16 inputs; i0 and D multiples of 8): the multi-GPU scheme that
 * GatheredMILNCELoss uses by default.  Same inputs as above (all-gathered video/text, a range),
 * but a rank computes only the rows of the similarity matrix it owns - 4/W GEMM units per rank
 * instead of 6/W - at the price of one reduce-scatter.  Replaces the same reference code as
 * milnce_forward/backward (loss.py:10-18, utils.py:11-24).
 *   forward : row_lse[i0, i0+n)  = log sum over ALL text columns of exp(S[i, :])      (complete)
 *             col_part[0, B)     = log sum over the OWNED rows i and the K candidates of sample j
 *                                  of exp(S[i, jK+k])   (partial: combine over ranks with logsumexp)
 *             nom[i0, i0+n)      as in milnce_forward
 *             The caller all-gathers row_lse / nom, logsumexp-reduces col_part over ranks and forms
 *             den_j = logaddexp(row_lse_j, col_j), loss = mean(den - nom).
 *   backward: dvideo (n, D) complete for the owned rows, written as dvideo_dtype (F32 or BF16);
 *             dtext_partial (B*K, D), written as dtext_dtype (F32 or BF16), holds only the owned
 *             rows' contribution: reduce-scatter (sum) it over ranks to obtain each rank's
 *             (n*K, D) slice (BF16 halves the bytes on the wire; every partial is then rounded once).
 *             parts = 3 computes both.  parts = 1 computes the gradient tile matrix and
 *             dtext_partial only (dvideo may be NULL); a following parts = 2 call with the same
 *             workspace computes dvideo (dtext_partial may be NULL), so the reduce-scatter of
 *             dtext can run while dvideo is computed. */
int milnce_forward_shard(const void *video, const void *text, int B, int K, int D, int dtype,
                         int i0, int n, float *row_lse, float *col_part, float *nom,
                         void *workspace, size_t workspace_bytes, void *stream);
/* After the exchange of the rows-only scheme: `gathered` is (W, 2b + B) float32, rank r's row =
 * [row_lse of its b samples | nom of its b samples | its col_part of all B = W*b samples] (the
 * all-gather of what milnce_forward_shard produced).  Writes den[B], nom[B] and loss[0] =
 * mean(den - nom); identical on every rank.  (The reference has no counterpart: there the full
 * similarity matrix is reduced on every rank, loss.py:16-18.) */
int milnce_shard_combine(const float *gathered, int W, int b, float *den, float *nom, float *loss,
                         void *stream);
/* The same, with the W packs read where they are: packs[r] = rank r's (2b + B) floats as mapped into this
 * process (host array of W device pointers), e.g. in peer-mapped memory after the caller's cross-rank
 * barrier - the all-gather of the statistics is then not needed.  W <= 16 for both variants.
 * scratch: NULL (one block walks all samples), or a 256-byte aligned device buffer of
 * milnce_shard_combine_scratch_bytes(W, b) bytes whose first 4 bytes are zero before the first call (the kernel
 * leaves them zero): the kernel then runs one sample per thread over several blocks, so that all remote loads are
 * one round trip, and the last block to finish adds the blocks' loss sums in block order (deterministic).
 * One scratch buffer per concurrently running call. */
size_t milnce_shard_combine_scratch_bytes(int W, int b);
int milnce_shard_combine_peers(const void *const *packs, int W, int b, float *den, float *nom, float *loss,
                               void *scratch, void *stream);
/* Rows-only sharding WITHOUT an embedding collective.  Every rank publishes its bfloat16 shard in a buffer
 * that its peers can read directly (peer-mapped memory, e.g. torch symmetric memory: slot pointers of all
 * ranks are exchanged once at set-up), and stages the fp16 operands of all W ranks straight from the
 * peers' buffers with plain loads over NVLink - the gather of utils.py:11-17 and the staging pass become one
 * kernel.  Per step and rank:
 *   milnce_peer_publish      copies the local shard into this rank's slot (milnce_peer_slot_bytes), leaves
 *                            |video|max, |text|max in the slot header and nom[0, b) of the local samples;
 *   <the caller's cross-rank barrier: every slot is complete>
 *   milnce_forward_shard_peers  reads the text of all slots and the video of its own (slots[r] = rank r's
 *                            slot as mapped into THIS process), then as milnce_forward_shard:
 *                            row_lse[i0, i0+n), col_part[0, B).
 * Backward: milnce_backward_shard with reuse_forward_ws = 1 (it reads only the workspace).  The caller must
 * not let a rank overwrite its slot while a peer may still be reading it (a cross-rank exchange between the
 * staging of one step and the publish of the next - e.g. the one that distributes the statistics - is enough).  Needs b = B / W a multiple of 64, D a multiple of 8, W <= 16. */
size_t milnce_peer_slot_bytes(int b, int K, int D);
int milnce_peer_publish(const void *video_local, const void *text_local, int b, int K, int D, void *slot,
                        float *nom_own, void *stream);
int milnce_forward_shard_peers(const void *const *host_slots, int W, int B, int K, int D, int i0, int n,
                               float *row_lse, float *col_part, void *workspace, size_t workspace_bytes,
                               void *stream);
/* parts = 1 of milnce_backward_shard (gradient tiles + partial text gradient, from the workspace the forward
 * left), with the partial DELIVERED instead of stored: the b*K rows of the (B*K, D) partial that rank q owns are
 * written as bfloat16 to owner_dst[q] (host array of W device pointers, each a (b*K, D) buffer, normally in
 * rank q's peer-mapped memory: this rank's block of q's receive area) by the GEMM epilogue's own stores, so the
 * transfer of the reduce-scatter overlaps the GEMM.  Afterwards: milnce_backward_shard(parts = 2) for dvideo,
 * the caller's cross-rank barrier, and milnce_peer_reduce_dtext over the W blocks each rank received
 * (partials[q] = block q, rank = 0).  b*K must be a multiple of 128, W <= 16. */
int milnce_backward_shard_push(const float *den, const float *nom, const float *grad_loss, void *const *owner_dst,
                               int W, int B, int K, int D, int i0, int n, void *workspace, size_t workspace_bytes,
                               void *stream);
/* The gradient collective of the rows-only scheme by pull, for callers whose partial text gradients sit in
 * peer-mapped memory (it replaces reduce_scatter(sum) of dtext_partial; reference counterpart: autograd of
 * the concatenated text tensor + the slice of utils.py:20-24).  partials[r] = rank r's (B*K, D) BFLOAT16
 * dtext_partial as mapped into this process (host array of W device pointers; write it with
 * milnce_backward_shard(parts = 1, dtext_dtype = BF16) and pass the caller's cross-rank barrier first).
 * dtext (b*K, D), float32 or bfloat16 = sum over r of rows [rank*b*K, (rank+1)*b*K) of partials[r], added in
 * rank order in float32: bit-identical from run to run.  W <= 16, b*K*D a multiple of 8. */
int milnce_peer_reduce_dtext(const void *const *partials, int W, int rank, int b, int K, int D, void *dtext,
                             int dtext_dtype, void *stream);
int milnce_backward_shard(const void *video, const void *text, const float *den, const float *nom,
                          const float *grad_loss, void *dvideo, int dvideo_dtype, void *dtext_partial,
                          int dtext_dtype, int parts, int B, int K, int D, int dtype, int i0, int n,
                          int reuse_forward_ws,
                          void *workspace, size_t workspace_bytes, void *stream);

/* ------------------------------------------------------------------------
 * soft-DTW: replaces SoftDTW.forward's distance function + _SoftDTWCUDA
 * (reference soft_dtw_cuda.py:326-363, 115-175, kernels :34-112).
 * x is (B, N, d), y is (B, M, d), float32; 1 <= N, M <= 1024.
 * gamma > 0; bandwidth <= 0 disables Sakoe-Chiba pruning (:60, :104).
 * ------------------------------------------------------------------------ */

/* Bytes of `aux` that sdtw_forward fills and sdtw_backward consumes. */
size_t sdtw_aux_bytes(int B, int N, int M, int d);

/* Forward.  out[B] = R[:, N, M].  R (B, N+2, M+2) float32 may be NULL; when
 * given it is filled exactly like the reference's R (+inf border, R[0,0]=0,
 * pruned cells +inf; soft_dtw_cuda.py:135-145).  aux (sdtw_aux_bytes, 256-byte
 * aligned) is always required: it holds the distance matrix and the per-cell
 * soft-min weights that sdtw_backward consumes.  B <= 65535 per call. */
int sdtw_forward(const float *x, const float *y, int B, int N, int M, int d,
                 float gamma, float bandwidth, int dist_kind,
                 float *R, float *out, void *aux, size_t aux_bytes, void *stream);

/* Backward.  grad_out[B]; dx (B, N, d), dy (B, M, d) float32.  E (B, N, M)
 * float32 may be NULL; when given it receives d out/d D, the reference's
 * E[:, 1:N+1, 1:M+1] (soft_dtw_cuda.py:174-175) before the grad_out scale. */
int sdtw_backward(const float *x, const float *y, const float *grad_out,
                  int B, int N, int M, int d, float gamma, float bandwidth, int dist_kind,
                  const void *aux, size_t aux_bytes, float *E, float *dx, float *dy,
                  void *stream);

/* The same on INDEX PAIRS: pair p compares sequence ix[p] of the x bank (Bx, N, d) with sequence iy[p]
 * of the y bank (By, M, d); ix / iy are device int32 arrays of length P, or NULL for the identity (that
 * bank must then hold exactly P sequences).  This is how the reference's composite losses are served
 * without materialising their operand copies: SDTW_3 builds all b*b pairs by
 * expand(b, b, n, d).reshape(-1, n, d) (reference loss.py:101-105, 112-116, 123-127), CDTW repeats one
 * sequence 8 times (loss.py:29).  out[P], R (P, N+2, M+2) and aux (sdtw_aux_bytes(P, N, M, d)) are per
 * pair.  In the backward the gradient of a sequence is the SUM over the pairs it takes part in: dx
 * (Bx, N, d) and dy (By, M, d) are zeroed by the call and accumulated with atomic adds when an index
 * list is given (summation order, hence the last float32 bits, then varies from run to run); dx and dy
 * must not overlap.  E (P, N, M) optional as in sdtw_backward. */
int sdtw_forward_pairs(const float *x, int Bx, const float *y, int By, const int *ix, const int *iy, int P,
                       int N, int M, int d, float gamma, float bandwidth, int dist_kind,
                       float *R, float *out, void *aux, size_t aux_bytes, void *stream);
int sdtw_backward_pairs(const float *x, int Bx, const float *y, int By, const int *ix, const int *iy, int P,
                        const float *grad_out, int N, int M, int d, float gamma, float bandwidth,
                        int dist_kind, const void *aux, size_t aux_bytes, float *E, float *dx, float *dy,
                        void *stream);

/* Forward for callers that cannot allocate device memory at all.  It has no
 * counterpart in the reference (there the scratch is torch tensors,
 * soft_dtw_cuda.py:135-136).  Same arguments and results as sdtw_forward, minus
 * `aux`: the distance matrix and norms that the two forward kernels exchange sit in
 * a buffer compiled into the library (B200HEAD_SDTW_NOALLOC_MAX_BYTES bytes of
 * every GPU that loads it, held until exit).  Meant for a binding from a runtime
 * that owns no device allocator, a process whose pool already holds the whole GPU,
 * or a loop that must avoid the device-wide stall of cudaMalloc/cudaFree.  The call
 * performs no cudaMalloc/cudaFree and uses no allocating library.
 *   - forward only: no aux is produced, so sdtw_backward cannot follow it;
 *   - limit: align256(4*B*N*M) + align256(4*B*N) + align256(4*B*M) must not exceed
 *     B200HEAD_SDTW_NOALLOC_MAX_BYTES, else B200HEAD_ERR_WS with the sizes in
 *     b200head_last_error() (e.g. 256 pairs of 64x64 need 4.3 MiB and fit);
 *   - the first call on each device looks up the buffer's address for that device
 *     (the device must be current); make that first call before starting any
 *     stream capture;
 *   - the buffer is shared by all calls on a device and is not cleared between
 *     them (nothing in it needs to start at zero); calls that overlap in time on
 *     the same GPU race on it, so the caller must serialise them. */
#define B200HEAD_SDTW_NOALLOC_MAX_BYTES (8u << 20)
int sdtw_forward_noalloc(const float *x, const float *y, int B, int N, int M, int d,
                         float gamma, float bandwidth, int dist_kind,
                         float *R, float *out, void *stream);

/* ------------------------------------------------------------------------
 * Test / diagnostic hooks for the tcgen05 GEMM core (C = alpha * A . B^T with
 * A (M, Kd) and Bm (N, Kd) float16 row-major, C float32 (M, N)).  No
 * reference counterpart; used by the GPU tests to validate the TMA / UMMA
 * descriptors independently of the loss epilogues.
 * ------------------------------------------------------------------------ */
int b200head_gemm_tn_f16(const void *A, const void *Bm, float *C, int M, int N, int Kd,
                         int lda, int ldb, int ldc, float alpha, void *stream);
/* Same product with A passed transposed: At is (Kd, M) row-major, i.e. the A operand
 * is read MN-major by the tensor core (the path dtext = G^T . V uses to consume G). */
int b200head_gemm_at_f16(const void *At, const void *Bm, float *C, int M, int N, int Kd,
                         int ldat, int ldb, int ldc, float alpha, void *stream);
/* Same products through the CTA-pair kernel (tcgen05.mma.cta_group::2, 256 x 256 tiles) that
 * the bf16 path uses; a_transposed bit 0: A is passed as (Kd, M) like b200head_gemm_at_f16, bit 1: Bm is
 * passed as (Kd, N) (MN-major B operand, how the gradient GEMMs read text and video); bit 2: both
 * operands hold bfloat16 instead of float16. */
int b200head_gemm2_f16(const void *A, const void *Bm, float *C, int M, int N, int Kd,
                       int lda, int ldb, int ldc, float alpha, int a_transposed, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200HEAD_H */
