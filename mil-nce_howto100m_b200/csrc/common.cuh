// Shared host-side helpers for the b200head C-ABI library.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200head.h"

namespace b200 {

// Thread-local error message behind b200head_last_error().
void set_error(const char *fmt, ...);
const char *get_error();

// 0 if the current device is compute capability 10.x, else an error code.
int check_device();

// cuTensorMapEncodeTiled resolved at run time through cudaGetDriverEntryPoint so
// that the library loads (and exports its symbols) on machines without libcuda.
// Describes a row-major fp16/bf16 matrix `rows x cols` with leading dimension
// `ld` elements; box = 64 columns (128 bytes, SWIZZLE_128B) x box_rows rows.
int make_tmap_2d_f16(CUtensorMap *map, const void *base, uint64_t rows, uint64_t cols,
                     uint64_t ld, uint32_t box_rows);

// fp32 tensor (inner, rows, batch): box = 32 floats (128 bytes) x box_rows x 1, zero fill; 128-byte
// swizzle with 16-byte atoms (atom32 == 0) or 32-byte atoms (atom32 != 0)
int make_tmap_3d_f32(CUtensorMap *map, const void *base, uint64_t inner, uint64_t rows, uint64_t batch,
                     uint32_t box_rows, int atom32);

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Launch with programmatic stream serialization: the grid may be scheduled while the previous
// kernel of the stream is still draining, so its prologue (barrier set-up, TMEM allocation,
// descriptor prefetch) and the launch latency overlap that kernel's tail.  Every kernel launched
// this way executes griddepcontrol.wait (tc::pdl_wait) before it touches global memory, which
// blocks until the previous grid has completed and its writes are visible.  Compiled out unless the
// library is built with -DB200_PDL=1: it measured neutral on the benchmark shape.
bool pdl_enabled();
// Every kernel launch the library makes is counted (b200head_launch_count in the C ABI).
void note_launches(unsigned n);
template <class... KArgs, class... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    note_launches(1);
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

#define B200_CUDA_OK(expr)                                                                 \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            b200::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e),        \
                            __FILE__, __LINE__);                                           \
            return B200HEAD_ERR_CUDA;                                                      \
        }                                                                                  \
    } while (0)

#define B200_LAUNCH_OK(what)                                                               \
    do {                                                                                   \
        cudaError_t _e = cudaGetLastError();                                               \
        if (_e != cudaSuccess) {                                                           \
            b200::set_error("launch of %s failed: %s", what, cudaGetErrorString(_e));      \
            return B200HEAD_ERR_CUDA;                                                      \
        }                                                                                  \
    } while (0)

}  // namespace b200
