// Shared host/device helpers for libfsb200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include <nvtx3/nvToolsExt.h>

#include "../../include/fsb200.h"

namespace fsb {

// thread-local error text + launch counter (defined in api.cu)
void set_error(const char* fmt, ...);
void count_launch(int n = 1);

// NVTX range around every C-ABI entry (the reference has no tracing hooks; the new code brings
// its own).  Header-only NVTX3: without a profiler attached a push/pop is a null function-pointer check.
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

inline int cuda_fail(cudaError_t e, const char* what) {
  set_error("%s: %s", what, cudaGetErrorString(e));
  return FSB_ERR_CUDA;
}

#define FSB_REQUIRE(cond, code, ...)  \
  do {                                \
    if (!(cond)) {                    \
      ::fsb::set_error(__VA_ARGS__);  \
      return (code);                  \
    }                                 \
  } while (0)

#define FSB_CHECK_LAUNCH(name)                                   \
  do {                                                           \
    cudaError_t e__ = cudaGetLastError();                        \
    if (e__ != cudaSuccess) return ::fsb::cuda_fail(e__, name);  \
    ::fsb::count_launch();                                       \
  } while (0)

// Programmatic dependent launch (PDL).  Every kernel of the library calls pdl_wait() before it touches
// memory another kernel may have written and pdl_launch_dependents() as soon as its prologue is done; with
// the launch attribute below the next kernel on the stream (also inside a captured CUDA graph) is scheduled
// while this one drains, so its launch latency and prologue (barrier init, TMEM allocation, descriptor and
// weight prefetch) overlap the predecessor's tail.  Without the attribute both instructions are no-ops.
bool pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                 Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int round_up(int a, int b) { return ceil_div(a, b) * b; }

// storage type <-> float
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// four consecutive outputs of storage type T as one vector store (16 B for fp32, 8 B for 16-bit types)
template <typename T> struct Quad;
template <> struct Quad<float> {
  static __device__ __forceinline__ void store(float* p, float a, float b, float c, float d) {
    *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
  }
};
template <> struct Quad<__half> {
  static __device__ __forceinline__ void store(__half* p, float a, float b, float c, float d) {
    __half2 lo = __floats2half2_rn(a, b), hi = __floats2half2_rn(c, d);
    uint2 v;
    v.x = *reinterpret_cast<uint32_t*>(&lo);
    v.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(p) = v;
  }
};
template <> struct Quad<__nv_bfloat16> {
  static __device__ __forceinline__ void store(__nv_bfloat16* p, float a, float b, float c, float d) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
    uint2 v;
    v.x = *reinterpret_cast<uint32_t*>(&lo);
    v.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(p) = v;
  }
};

struct LevelPtrs {
  float* p[FSB_MAX_LEVELS];
};

// Banded init correlation (opt-in): level l stores, for the pixel in image column w1, the bw0 >> l
// columns [s_l, s_l + (bw0 >> l)) with s_l = s0 >> l and s0 = floor((w1 - off) / align) * align -- the part of the row
// a lookup with disparities in [0, d_band] and radius r can touch (plus alignment slack), instead of all W2 columns.
// bw0 == 0 means full rows.  Columns outside [0, W2 >> l) are stored as zeros; a lookup outside the band reads zeros
// where the reference reads real correlations (core/geometry.py:57 does not clamp the disparity).
struct CorrBand {
  int bw0, off, align;
};
__host__ __device__ inline int band_floor_div(int a, int b) { return a >= 0 ? a / b : -((-a + b - 1) / b); }
__host__ __device__ inline int band_start(const CorrBand& bd, int w1) { return band_floor_div(w1 - bd.off, bd.align) * bd.align; }
inline CorrBand make_corr_band(int d_band, int r, int L) {
  CorrBand bd;
  bd.align = 4 << (L - 1);                    // every level's start stays a multiple of 4 columns (16-byte stores)
  const int left = (r + 2) << (L - 1);        // level-0 columns a coarsest-level tap reaches left of w1 - d_band ...
  const int right = (r + 3) << (L - 1);       // ... and right of w1 (taps floor(x) and floor(x) + 1, one column of rounding slack)
  bd.off = d_band + left;
  bd.bw0 = round_up(d_band + left + right + bd.align - 1, bd.align);
  return bd;
}

// corr_umma.cu: init correlation on tcgen05 (nprod 1 = TF32, 3 = 3xTF32 split, fp32-accurate)
int launch_corr_umma(const float* f1, const float* f2, const LevelPtrs& lv, int B, int C, int H, int W1, int W2, int L,
                     int nprod, cudaStream_t stream);

// corr_tma.cu: the same with the NCHW planes fed by TMA as MN-major operands (needs W % 4 == 0, 16-byte aligned bases)
bool corr_tma_eligible(const float* f1, const float* f2, int W1, int W2);
int launch_corr_tma(const float* f1, const float* f2, const LevelPtrs& lv, int B, int C, int H, int W1, int W2, int L,
                    int nprod, cudaStream_t stream, CorrBand band = CorrBand{0, 0, 0});

}  // namespace fsb
