// The steps either side of the refinement loop, both pure HBM streams:
//   * disparity_regression (core/submodule.py:431-435), optionally with the classifier softmax of
//     core/foundation_stereo.py:218 fused in, so the (B,D,H,W) probability tensor is never written;
//   * context_upsample (core/submodule.py:456-468), optionally with the softmax over the 9 weights of
//     core/foundation_stereo.py:188 and the `disp*4.` of :189 fused in.
#include "common.cuh"

namespace fsb {
namespace {

// 4 (fp32: 16 B) consecutive elements of storage type T as floats
template <typename T> struct Vec4;
template <> struct Vec4<float> {
  static __device__ __forceinline__ void load(const float* p, float v[4]) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
  }
};
template <> struct Vec4<__half> {
  static __device__ __forceinline__ void load(const __half* p, float v[4]) {
    const uint2 q = __ldg(reinterpret_cast<const uint2*>(p));
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&q.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&q.y));
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
  }
};
template <> struct Vec4<__nv_bfloat16> {
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float v[4]) {
    const uint2 q = __ldg(reinterpret_cast<const uint2*>(p));
    const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&q.x));
    const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&q.y));
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
  }
};

template <typename T, int V> __device__ __forceinline__ void load_vec(const T* p, float v[V]) {
  if constexpr (V == 4) {
    Vec4<T>::load(p, v);
  } else {
    v[0] = to_f32<T>(__ldg(p));
  }
}

constexpr int kRegLanes = 32;   // pixel vectors per CTA
constexpr int kRegGroup = 4;    // disparities per thread and trip: kRegGroup independent 16-byte loads in flight

// x (B,D,HW) -> out (B,HW).  CTA = 32 pixel vectors x SLICES disparity slices; slice j walks the disparity groups
// j, j+SLICES, ... of kRegGroup consecutive d; the partial sums (LOGITS: running softmax triples max / sum exp /
// sum d*exp, rescaled once per group) meet in shared memory.  Every byte of x is read once, 16 B per lane.
template <typename T, typename TO, int V, int SLICES, bool LOGITS>
__global__ void __launch_bounds__(SLICES* kRegLanes)
disp_regression_kernel(const T* __restrict__ x, TO* __restrict__ out, int D, int HW) {
  __shared__ float sm[3][SLICES][kRegLanes * V + 1];
  const int lane = threadIdx.x, slice = threadIdx.y;
  const int b = blockIdx.y;
  const int p0 = (blockIdx.x * kRegLanes + lane) * V;
  pdl_wait();
  pdl_launch_dependents();
  float m[V], s[V], t[V];
#pragma unroll
  for (int i = 0; i < V; ++i) { m[i] = -INFINITY; s[i] = 0.f; t[i] = 0.f; }
  if (p0 < HW) {
    const T* src = x + ((size_t)b * D) * HW + p0;
    for (int d0 = slice * kRegGroup; d0 < D; d0 += SLICES * kRegGroup) {
      float v[kRegGroup][V];
#pragma unroll
      for (int g = 0; g < kRegGroup; ++g) {
        if (d0 + g < D) {
          load_vec<T, V>(src + (size_t)(d0 + g) * HW, v[g]);
        } else {
#pragma unroll
          for (int i = 0; i < V; ++i) v[g][i] = LOGITS ? -INFINITY : 0.f;
        }
      }
#pragma unroll
      for (int i = 0; i < V; ++i) {
        if constexpr (LOGITS) {
          float gm = v[0][i];
#pragma unroll
          for (int g = 1; g < kRegGroup; ++g) gm = fmaxf(gm, v[g][i]);
          const float mn = fmaxf(m[i], gm);
          if (mn > -INFINITY) {  // a group of -inf logits on top of nothing adds nothing
            const float c = __expf(m[i] - mn);  // 0 on the first finite group (m = -inf)
            float gs = 0.f, gt = 0.f;
#pragma unroll
            for (int g = 0; g < kRegGroup; ++g) {
              const float e = __expf(v[g][i] - mn);
              gs += e;
              gt = fmaf((float)(d0 + g), e, gt);
            }
            s[i] = fmaf(s[i], c, gs);
            t[i] = fmaf(t[i], c, gt);
            m[i] = mn;
          }
        } else {
#pragma unroll
          for (int g = 0; g < kRegGroup; ++g) t[i] = fmaf(v[g][i], (float)(d0 + g), t[i]);
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < V; ++i) {
    sm[0][slice][lane * V + i] = t[i];
    if constexpr (LOGITS) { sm[1][slice][lane * V + i] = s[i]; sm[2][slice][lane * V + i] = m[i]; }
  }
  __syncthreads();
  // thread (slice, lane) finishes pixel elements e = slice*32 + lane (+ SLICES*32 ...) of the CTA's 32*V pixels
  for (int e = slice * kRegLanes + lane; e < kRegLanes * V; e += SLICES * kRegLanes) {
    const int p = blockIdx.x * kRegLanes * V + e;
    if (p < HW) {
      float res;
      if constexpr (LOGITS) {
        float mm = -INFINITY;
#pragma unroll
        for (int j = 0; j < SLICES; ++j) mm = fmaxf(mm, sm[2][j][e]);
        float S = 0.f, Tt = 0.f;
#pragma unroll
        for (int j = 0; j < SLICES; ++j) {
          const float mj = sm[2][j][e];
          const float c = (mj == mm) ? 1.0f : __expf(mj - mm);
          S = fmaf(sm[1][j][e], c, S);
          Tt = fmaf(sm[0][j][e], c, Tt);
        }
        res = Tt / S;
      } else {
        res = 0.f;
#pragma unroll
        for (int j = 0; j < SLICES; ++j) res += sm[0][j][e];
      }
      out[(size_t)b * HW + p] = from_f32<TO>(res);
    }
  }
}

// disp (B,h,w) fp32, wgt (B,9,4h,4w) of type T -> out (B,4h,4w) fp32.
// Thread = the four full-resolution columns of one low-resolution pixel in one full-resolution row: nine 16-byte
// (fp32) weight loads, nine neighbour disparities (zero outside the image, F.unfold's padding), one 16-byte store.
template <typename T, bool LOGITS>
__global__ void __launch_bounds__(256)
context_upsample_kernel(const float* __restrict__ disp, const T* __restrict__ wgt, float* __restrict__ out, int h,
                        int w, float scale, long long total) {
  pdl_wait();
  pdl_launch_dependents();
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int xl = (int)(idx % w);
  const long long r = idx / w;
  const int Y = (int)(r % (4 * h));
  const int b = (int)(r / (4 * h));
  const int yl = Y >> 2;
  const size_t plane = (size_t)16 * h * w;
  const T* wp = wgt + (size_t)b * 9 * plane + (size_t)Y * (4 * w) + 4 * xl;
  float wv[9][4];
#pragma unroll
  for (int k = 0; k < 9; ++k) Vec4<T>::load(wp + k * plane, wv[k]);
  float dn[9];
  const float* dp = disp + (size_t)b * h * w;
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const int yy = yl + ky - 1, xx = xl + kx - 1;
      const bool in = (yy >= 0) && (yy < h) && (xx >= 0) && (xx < w);
      dn[ky * 3 + kx] = in ? __ldg(dp + (size_t)yy * w + xx) * scale : 0.f;
    }
  }
  float o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if constexpr (LOGITS) {
      float mm = wv[0][i];
#pragma unroll
      for (int k = 1; k < 9; ++k) mm = fmaxf(mm, wv[k][i]);
      float S = 0.f, acc = 0.f;
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const float ex = __expf(wv[k][i] - mm);
        S += ex;
        acc = fmaf(dn[k], ex, acc);
      }
      o[i] = acc / S;
    } else {
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < 9; ++k) acc = fmaf(dn[k], wv[k][i], acc);
      o[i] = acc;
    }
  }
  *reinterpret_cast<float4*>(out + ((size_t)b * 4 * h + Y) * (4 * w) + 4 * xl) = make_float4(o[0], o[1], o[2], o[3]);
}

template <typename T, typename TO, int V, int SLICES, bool LOGITS>
cudaError_t launch_regression_v(const void* x, void* out, int B, int D, int HW, cudaStream_t stream) {
  const dim3 grid(ceil_div(HW, kRegLanes * V), B), block(kRegLanes, SLICES);
  return launch_kernel(disp_regression_kernel<T, TO, V, SLICES, LOGITS>, grid, block, 0, stream, (const T*)x, (TO*)out,
                       D, HW);
}

template <typename T, typename TO, bool LOGITS>
int launch_regression(const void* x, void* out, int B, int D, int HW, cudaStream_t stream) {
  const bool vec = (HW % 4 == 0) && aligned16(x);
  // 4 slices (128 threads) once the grid alone fills the SMs: all CTAs resident in one wave (16 per SM);
  // 8 slices for small batches, where the disparity axis has to supply the parallelism
  const bool many = (long long)ceil_div(HW, kRegLanes * (vec ? 4 : 1)) * B >= 4 * 148;
  cudaError_t e;
  if (vec) e = many ? launch_regression_v<T, TO, 4, 4, LOGITS>(x, out, B, D, HW, stream)
                    : launch_regression_v<T, TO, 4, 8, LOGITS>(x, out, B, D, HW, stream);
  else     e = many ? launch_regression_v<T, TO, 1, 4, LOGITS>(x, out, B, D, HW, stream)
                    : launch_regression_v<T, TO, 1, 8, LOGITS>(x, out, B, D, HW, stream);
  if (e != cudaSuccess) return cuda_fail(e, "disp_regression_kernel");
  FSB_CHECK_LAUNCH("disp_regression_kernel");
  return FSB_OK;
}

template <typename T>
int launch_upsample(const float* disp, const void* wgt, float* out, int B, int h, int w, int logits, float scale,
                    cudaStream_t stream) {
  const long long total = (long long)B * 4 * h * w;
  const unsigned grid = (unsigned)((total + 255) / 256);
  cudaError_t e;
  if (logits)
    e = launch_kernel(context_upsample_kernel<T, true>, dim3(grid), dim3(256), 0, stream, disp, (const T*)wgt, out, h, w,
                      scale, total);
  else
    e = launch_kernel(context_upsample_kernel<T, false>, dim3(grid), dim3(256), 0, stream, disp, (const T*)wgt, out, h,
                      w, scale, total);
  if (e != cudaSuccess) return cuda_fail(e, "context_upsample_kernel");
  FSB_CHECK_LAUNCH("context_upsample_kernel");
  return FSB_OK;
}

}  // namespace
}  // namespace fsb

extern "C" {

int fsb_disparity_regression(const void* x, void* out, int B, int D, int H, int W, int logits, int dtype,
                             fsb_stream_t stream) {
  using namespace fsb;
  FSB_REQUIRE(x != nullptr && out != nullptr, FSB_ERR_ARG, "fsb_disparity_regression: NULL buffer");
  FSB_REQUIRE(B > 0 && D > 0 && H > 0 && W > 0, FSB_ERR_ARG, "fsb_disparity_regression: non-positive size");
  FSB_REQUIRE((long long)H * W < (1ll << 31) && B <= 65535, FSB_ERR_SHAPE, "fsb_disparity_regression: H*W or B too large");
  cudaStream_t s = (cudaStream_t)stream;
  const int HW = H * W;
  switch (dtype) {
    case FSB_F32:
      return logits ? launch_regression<float, float, true>(x, out, B, D, HW, s)
                    : launch_regression<float, float, false>(x, out, B, D, HW, s);
    case FSB_F16:
      return logits ? launch_regression<__half, float, true>(x, out, B, D, HW, s)
                    : launch_regression<__half, __half, false>(x, out, B, D, HW, s);
    case FSB_BF16:
      return logits ? launch_regression<__nv_bfloat16, float, true>(x, out, B, D, HW, s)
                    : launch_regression<__nv_bfloat16, __nv_bfloat16, false>(x, out, B, D, HW, s);
    default:
      break;
  }
  FSB_REQUIRE(false, FSB_ERR_ARG, "fsb_disparity_regression: unknown dtype %d", dtype);
  return FSB_ERR_ARG;
}

int fsb_context_upsample(const float* disp_low, const void* weights, float* out, int B, int h, int w, int logits,
                         float scale, int dtype, fsb_stream_t stream) {
  using namespace fsb;
  FSB_REQUIRE(disp_low != nullptr && weights != nullptr && out != nullptr, FSB_ERR_ARG,
              "fsb_context_upsample: NULL buffer");
  FSB_REQUIRE(B > 0 && h > 0 && w > 0, FSB_ERR_ARG, "fsb_context_upsample: non-positive size");
  FSB_REQUIRE(aligned16(weights) && aligned16(out), FSB_ERR_ALIGN, "fsb_context_upsample: buffers must be 16-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case FSB_F32: return launch_upsample<float>(disp_low, weights, out, B, h, w, logits, scale, s);
    case FSB_F16: return launch_upsample<__half>(disp_low, weights, out, B, h, w, logits, scale, s);
    case FSB_BF16: return launch_upsample<__nv_bfloat16>(disp_low, weights, out, B, h, w, logits, scale, s);
    default: break;
  }
  FSB_REQUIRE(false, FSB_ERR_ARG, "fsb_context_upsample: unknown dtype %d", dtype);
  return FSB_ERR_ARG;
}

}  // extern "C"
