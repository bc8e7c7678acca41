// Two-CTA (cta_group::2) variant of the tcgen05 GEMM core.
//
// The one-CTA kernel in tc_gemm.cuh runs at the chip-wide L2 -> SM bandwidth limit
// (12.3 TB/s measured): a 128x128 tile needs 128 B/clk/SM of operands at full tensor
// rate, a 256x128 tile 96.  Here a CTA PAIR (two SMs of one TPC, a 2-CTA cluster)
// computes a 256 x 256 tile with tcgen05.mma.cta_group::2: each CTA stages only its
// 128 rows of A and its 128-row half of B (32 KB per k-block per SM, i.e. 64 B/clk/SM),
// the leader CTA issues one M=256, N=256 MMA per K=16 step for both SMs, and each SM
// keeps its 128 x 256 half of the accumulator in its own TMEM (2 x 256 columns,
// double-buffered).  Epilogues are the same functors as in tc_gemm.cuh.
//
// Synchronisation (barrier objects live at the same smem offsets in both CTAs):
//   full[s]    leader's only.  Both CTAs' TMA loads complete_tx on it; each producer
//              thread arrives on it with its own expect_tx (count 2).
//   empty[s]   per CTA.  The leader's tcgen05.commit multicasts the arrival to both.
//   tfull[b]   per CTA, multicast commit as well: accumulator b is complete.
//   tempty[b]  leader's only.  All 16 epilogue warps of the pair arrive (remote arrive
//              from the follower) once they have drained buffer b into registers.
#pragma once
#include "tc_gemm.cuh"

namespace tc {

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` (a shared::cta address in this CTA) inside CTA `rank`
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_expect_tx_cluster(uint32_t bar_cluster_addr, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;" ::"r"(bar_cluster_addr),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
// TMA load whose completion is signalled on a barrier that may live in the peer CTA
__device__ __forceinline__ void tma_load_2d_2sm(void *dst, const CUtensorMap *map,
                                                uint32_t bar_cluster_addr, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
        "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t *slot_in_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     smem_u32(slot_in_smem)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                 : "memory");
}
__device__ __forceinline__ void umma_f16_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                             uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives on `bar` (same offset) in both CTAs of the pair once the issued MMAs have retired
__device__ __forceinline__ void umma_commit_2sm(uint64_t *bar) {
    const uint16_t mask = 3;
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 "
        "[%0], %1;" ::"r"(smem_u32(bar)),
        "h"(mask)
        : "memory");
}

constexpr int PAIR_M = 256, PAIR_N = 256;

template <int EPI_STAGE_BYTES>
struct SmemLayout2 {
    static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;      // this CTA's 128 rows of A
    static constexpr int B_BYTES = (PAIR_N / 2) * BLOCK_K * 2; // this CTA's half of B
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int TAIL_BYTES = 8192;
    static constexpr int MAX_SMEM = 232448;
    static constexpr int STAGES_FIT = (MAX_SMEM - 1024 - TAIL_BYTES - EPI_STAGE_BYTES) / STAGE_BYTES;
    // epilogues that read per-sample vectors from global memory (EPI_STAGE_BYTES > 0: EpiG) keep
    // the ring at 4 stages so that ~55 KB of the SM stay L1 cache and those reads hit
    static constexpr int STAGES_MAX = EPI_STAGE_BYTES > 0 ? 4 : 6;
    static constexpr int STAGES = STAGES_FIT > STAGES_MAX ? STAGES_MAX : STAGES_FIT;
    static constexpr int TOTAL = STAGES * STAGE_BYTES + EPI_STAGE_BYTES + TAIL_BYTES + 1024;
    static_assert(STAGES >= 3, "pipeline too shallow");
};

template <class Epi, bool A_MN = false, bool N_FAST = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(num_threads<Epi>(), 1)
gemm_tn_2cta_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                    const GemmShape shape, const typename Epi::Params ep) {
    using L = SmemLayout2<Epi::STAGE_BYTES>;
    constexpr int STAGES = L::STAGES;
    constexpr int HN = 128 / Epi::PARTS;   // columns per epilogue thread per pass
    constexpr int NUM_EPI_WARPS = 4 * Epi::PARTS;
    constexpr uint32_t TMEM_COLS = 512;    // 2 buffers x 256 columns

    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                                ~(uintptr_t)1023);
    uint8_t *epi_stage = smem + STAGES * L::STAGE_BYTES;
    uint8_t *tail = epi_stage + Epi::STAGE_BYTES;
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(tail);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tfull_bar = empty_bar + STAGES;
    uint64_t *tempty_bar = tfull_bar + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty_bar + 2);
    float *epi_scratch = reinterpret_cast<float *>(tmem_slot + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();  // 0 = leader
    const int pair = blockIdx.x >> 1;
    const int num_pairs = gridDim.x >> 1;

    const int num_m = (shape.M + PAIR_M - 1) / PAIR_M;
    const int num_n = (shape.N + PAIR_N - 1) / PAIR_N;
    const int num_tiles = num_m * num_n;
    const int k_iters = (shape.Kd + BLOCK_K - 1) / BLOCK_K;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 2);   // one arrive(+expect_tx) per CTA of the pair
            mbar_init(&empty_bar[s], 1);  // the leader's multicast commit
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull_bar[b], 1);
            mbar_init(&tempty_bar[b], 2 * NUM_EPI_WARPS);  // epilogue warps of both CTAs
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_2sm(tmem_slot, TMEM_COLS);
    tcgen05_fence_before();
    __syncthreads();
    cluster_sync_all();  // the peer's barriers are initialised before anything signals them
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = pair; t < num_tiles; t += num_pairs) {
                const int m_blk = N_FAST ? t / num_n : t % num_m;
                const int n_blk = N_FAST ? t % num_n : t / num_m;
                const int m0 = m_blk * PAIR_M + (int)rank * BLOCK_M;        // my rows of A
                const int n0 = n_blk * PAIR_N + (int)rank * (PAIR_N / 2);   // my half of B
                for (int kb = 0; kb < k_iters; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    const uint32_t lead_full = mapa_u32(smem_u32(&full_bar[stage]), 0);
                    mbar_expect_tx_cluster(lead_full, L::STAGE_BYTES);
                    uint8_t *sa = smem + stage * L::STAGE_BYTES;
                    if (A_MN) {
                        tma_load_2d_2sm(sa, &map_a, lead_full, m0, kb * BLOCK_K);
                        tma_load_2d_2sm(sa + L::A_BYTES / 2, &map_a, lead_full, m0 + 64, kb * BLOCK_K);
                    } else {
                        tma_load_2d_2sm(sa, &map_a, lead_full, kb * BLOCK_K, m0);
                    }
                    tma_load_2d_2sm(sa + L::A_BYTES, &map_b, lead_full, kb * BLOCK_K, n0);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA, one thread) =====================
        if (rank == 0 && lane == 0) {
            constexpr uint32_t idesc = make_idesc_f16(PAIR_M, PAIR_N, A_MN ? 1 : 0);
            int stage = 0;
            uint32_t phase = 0;
            uint32_t cc = 0;
            for (int t = pair; t < num_tiles; t += num_pairs, ++cc) {
                const uint32_t buf = cc & 1;
                mbar_wait(&tempty_bar[buf], ((cc >> 1) & 1) ^ 1);
                tcgen05_fence_after();
                const uint32_t tmem_d = tmem_base + buf * PAIR_N;
                for (int kb = 0; kb < k_iters; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tcgen05_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * L::STAGE_BYTES);
                    const uint64_t bdesc = make_smem_desc(sa + L::A_BYTES);
                    if (A_MN) {
                        const uint64_t adesc = make_smem_desc_mn(sa, L::A_BYTES / 2);
#pragma unroll
                        for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                            umma_f16_2sm(tmem_d, adesc + (uint64_t)(k * (UMMA_K * 128 / 16)),
                                         bdesc + (uint64_t)(k * 2), idesc, (uint32_t)((kb | k) != 0));
                    } else {
                        const uint64_t adesc = make_smem_desc(sa);
#pragma unroll
                        for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                            umma_f16_2sm(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2),
                                         idesc, (uint32_t)((kb | k) != 0));
                    }
                    umma_commit_2sm(&empty_bar[stage]);  // frees this slot in BOTH CTAs
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit_2sm(&tfull_bar[buf]);  // accumulator complete, both CTAs
            }
        }
    } else {
        // ===================== epilogue warps 2..9 (both CTAs) =====================
        EpiCtx c;
        c.q = warp & 3;
        c.half = (warp - 2) >> 2;
        c.warp_slot = warp - 2;
        c.lane = lane;
        c.M = shape.M;
        c.N = shape.N;
        c.tid = c.q * 32 + lane;
        c.etid = (warp - 2) * 32 + lane;
        c.scratch = epi_scratch;
        c.stage = epi_stage;
        c.nthreads = NUM_EPI_WARPS * 32;
        c.nsub = 2;
        uint32_t cc = 0;
        int runs = 0;
        for (int t = pair; t < num_tiles; t += num_pairs, ++cc) {
            const int m_blk = N_FAST ? t / num_n : t % num_m;
            const int n_blk = N_FAST ? t % num_n : t / num_m;
            c.m_blk = m_blk * 2 + (int)rank;   // 128-row block index of this CTA's rows
            c.m0 = c.m_blk * BLOCK_M;
            const uint32_t buf = cc & 1;
            typename Epi::Tile tile_state = Epi::begin_tile(ep, c);  // overlaps the wait below
            mbar_wait(&tfull_bar[buf], (cc >> 1) & 1);
            tcgen05_fence_after();
#pragma unroll 1
            for (int nb = 0; nb < 2; ++nb) {   // the 256-wide tile as two 128-wide column blocks
                c.n_blk = n_blk * 2 + nb;
                c.sub = nb;
                c.n0 = c.n_blk * 128 + c.half * HN;
                c.par = runs & 1;
                ++runs;
                const uint32_t taddr = tmem_base + ((uint32_t)(c.q * 32) << 16) + buf * PAIR_N +
                                       (uint32_t)(nb * 128 + c.half * HN);
                float acc[HN];
#pragma unroll
                for (int ch = 0; ch < HN / 32; ++ch) {
                    float v[32];
                    tmem_ld32(taddr + ch * 32, v);
#pragma unroll
                    for (int i = 0; i < 32; ++i) acc[ch * 32 + i] = v[i];
                }
                if (nb == 1) {
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(mapa_u32(smem_u32(&tempty_bar[buf]), 0));
                }
                Epi::template run<HN>(ep, c, tile_state, acc);
            }
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    cluster_sync_all();  // nobody may still be signalling into, or reading from, the peer
    if (warp == 1) {
        tcgen05_fence_after();
        tmem_dealloc_2sm(tmem_base, TMEM_COLS);
    }
}

}  // namespace tc
