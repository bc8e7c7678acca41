// Microbenchmark: issue rate of legacy mma.sync (tf32 m16n8k8, f16 m16n8k16) and FFMA on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_sync_rate mma_sync_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, int iters) {
    float c[4][4] = {};
    uint32_t a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, b0 = a0 * 3, b1 = a0 * 5;
    float f0 = threadIdx.x * 1e-3f, f1 = 1.0001f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (MODE == 0)
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[u][0]), "+f"(c[u][1]), "+f"(c[u][2]), "+f"(c[u][3])
                             : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else if (MODE == 1)
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+f"(c[u][0]), "+f"(c[u][1]), "+f"(c[u][2]), "+f"(c[u][3])
                             : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
            else {
#pragma unroll
                for (int j = 0; j < 4; ++j) c[u][j] = fmaf(c[u][j], f1, f0);
            }
        }
    }
    float s = 0;
    for (int u = 0; u < 4; ++u)
        for (int j = 0; j < 4; ++j) s += c[u][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char *name, double macs_per_warp_instr, int per_iter) {
    float *out;
    cudaMalloc(&out, 148 * 8 * 256 * 4);
    const int iters = 20000;
    k<MODE><<<148 * 8, 256>>>(out, 10);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<148 * 8, 256>>>(out, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double warps = 148.0 * 8 * 8;
    double instr = warps * iters * per_iter;
    printf("%-28s %8.3f ms  %.3e warp-instr/s  %.1f TMAC/s  (%.2f warp-instr/clk/SM @1.9GHz)\n", name, ms,
           instr / (ms * 1e-3), instr * macs_per_warp_instr / (ms * 1e-3) / 1e12,
           instr / (ms * 1e-3) / 148 / 1.9e9);
    cudaFree(out);
}

int main() {
    run<0>("mma.sync m16n8k8 tf32", 16 * 8 * 8, 4);
    run<1>("mma.sync m16n8k16 f16", 16 * 8 * 16, 4);
    run<2>("ffma (x16 per iter)", 32, 16);
    return 0;
}
