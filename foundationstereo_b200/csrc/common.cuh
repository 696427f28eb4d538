// Shared host/device helpers for libfsb200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/fsb200.h"

namespace fsb {

// thread-local error text + launch counter (defined in api.cu)
void set_error(const char* fmt, ...);
void count_launch(int n = 1);

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

// corr_umma.cu: init correlation on tcgen05 (nprod 1 = TF32, 3 = 3xTF32 split, fp32-accurate)
int launch_corr_umma(const float* f1, const float* f2, const LevelPtrs& lv, int B, int C, int H, int W1, int W2, int L,
                     int nprod, cudaStream_t stream);

}  // namespace fsb
