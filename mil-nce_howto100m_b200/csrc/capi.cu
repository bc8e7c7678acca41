// Error reporting, device check and TMA descriptor construction for libb200head.
#include <atomic>
#include <mutex>
#include <string.h>

#include <stdlib.h>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

const char *get_error() { return g_err; }

int check_device() {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        set_error("cudaGetDevice failed: %s (no CUDA device? this library has no CPU path)",
                  cudaGetErrorString(e));
        return B200HEAD_ERR_CUDA;
    }
    static thread_local int cached_dev = -1, cached_major = 0;
    if (cached_dev != dev) {
        int major = 0;
        e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
        if (e != cudaSuccess) {
            set_error("cudaDeviceGetAttribute failed: %s", cudaGetErrorString(e));
            return B200HEAD_ERR_CUDA;
        }
        cached_dev = dev;
        cached_major = major;
    }
    if (cached_major != 10) {
        set_error("device %d has compute capability %d.x; kernels are built for sm_100a only",
                  dev, cached_major);
        return B200HEAD_ERR_DEVICE;
    }
    return 0;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                  const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
                cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    });
    return fn;
}

int make_tmap_2d_f16(CUtensorMap *map, const void *base, uint64_t rows, uint64_t cols,
                     uint64_t ld, uint32_t box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return B200HEAD_ERR_CUDA;
    }
    if (((uintptr_t)base & 15) || ((ld * 2) & 15)) {
        set_error("TMA operand must be 16-byte aligned with a 16-byte multiple row pitch "
                  "(base=%p ld=%llu)", base, (unsigned long long)ld);
        return B200HEAD_ERR_ARG;
    }
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstr[1] = {ld * 2};
    cuuint32_t box[2] = {64, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), gdim, gstr,
                    box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%llu cols=%llu ld=%llu)",
                  (int)r, (unsigned long long)rows, (unsigned long long)cols,
                  (unsigned long long)ld);
        return B200HEAD_ERR_CUDA;
    }
    return 0;
}

// fp32 tensor (inner, rows, batch), dense: box = 32 floats (128 bytes) x box_rows x 1, SWIZZLE_128B with
// 16-byte atoms, or with 32-byte atoms when atom32 != 0 (what an MN-major tf32 UMMA operand needs).
// Rows or columns of a box that fall outside the tensor are zero-filled.
int make_tmap_3d_f32(CUtensorMap *map, const void *base, uint64_t inner, uint64_t rows, uint64_t batch,
                     uint32_t box_rows, int atom32) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled not available from the driver");
        return B200HEAD_ERR_CUDA;
    }
    if (((uintptr_t)base & 15) || ((inner * 4) & 15)) {
        set_error("TMA operand must be 16-byte aligned with a 16-byte multiple row pitch (base=%p inner=%llu)",
                  base, (unsigned long long)inner);
        return B200HEAD_ERR_ARG;
    }
    cuuint64_t gdim[3] = {inner, rows, batch};
    cuuint64_t gstr[2] = {inner * 4, inner * rows * 4};
    cuuint32_t box[3] = {32, box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void *>(base), gdim, gstr, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (3-D fp32) failed with CUresult %d (inner=%llu rows=%llu batch=%llu)",
                  (int)r, (unsigned long long)inner, (unsigned long long)rows, (unsigned long long)batch);
        return B200HEAD_ERR_CUDA;
    }
    return 0;
}

}  // namespace b200

extern "C" const char *b200head_last_error(void) { return b200::get_error(); }
extern "C" int b200head_version(void) { return 3; }

namespace b200 {
// Programmatic dependent launch measured neutral on the benchmark shape (0.993 vs 0.997 ms), so it
// is compiled out by default; -DB200_PDL=1 in a variant build turns it on.  No run-time switch.
#ifndef B200_PDL
#define B200_PDL 0
#endif
bool pdl_enabled() { return B200_PDL != 0; }

static std::atomic<unsigned long long> g_launches{0};
void note_launches(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace b200

extern "C" unsigned long long b200head_launch_count(void) {
    return b200::g_launches.load(std::memory_order_relaxed);
}
