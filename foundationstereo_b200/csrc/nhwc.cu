// Small HBM-bound helpers of the update block (CUDA cores): layout changes between the
// reference's fp32 NCHW tensors and the 16-bit NHWC activations the tcgen05 convolution
// reads, pool2x / interp (core/update.py:72-80), the two 7x7 convolutions that are not
// GEMM-shaped (1 input channel: BasicMotionEncoder.convd1, core/update.py:58; depthwise:
// EdgeNextConvEncoder.dwconv, core/submodule.py:569) and the disparity update
// (core/foundation_stereo.py:242).
#include "common.cuh"

namespace fsb {

// ------------------------------------------------------------------ NCHW <-> NHWC
// tile = 32 pixels x 32 channels through shared memory; both sides coalesced.
template <typename T>
__global__ void __launch_bounds__(256)
nchw_to_nhwc_kernel(const float* __restrict__ src, T* __restrict__ dst16, float* __restrict__ dst32, int C,
                    int HW, int Cpad) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  __shared__ float t[32][33];
  const int b = blockIdx.z;
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const float* s = src + (size_t)b * C * HW;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = c0 + ty + 8 * i, p = p0 + tx;
    t[ty + 8 * i][tx] = (c < C && p < HW) ? s[(size_t)c * HW + p] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = p0 + ty + 8 * i, c = c0 + tx;
    if (p >= HW) continue;
    const float v = t[tx][ty + 8 * i];
    if (dst16 != nullptr && c < Cpad) dst16[((size_t)b * HW + p) * Cpad + c] = from_f32<T>(v);
    if (dst32 != nullptr && c < C) dst32[((size_t)b * HW + p) * C + c] = v;
  }
}

__global__ void __launch_bounds__(256)
nhwc_to_nchw_kernel(const float* __restrict__ src, float* __restrict__ dst, int C, int HW) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  __shared__ float t[32][33];
  const int b = blockIdx.z;
  const int p0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = p0 + ty + 8 * i, c = c0 + tx;
    t[ty + 8 * i][tx] = (c < C && p < HW) ? src[((size_t)b * HW + p) * C + c] : 0.f;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = c0 + ty + 8 * i, p = p0 + tx;
    if (c < C && p < HW) dst[((size_t)b * C + c) * HW + p] = t[tx][ty + 8 * i];
  }
}

// ------------------------------------------------------------------ pool2x / interp
// one thread = V channels of one output pixel (V = 8 when C allows it: two float4 loads per tap, one 16-byte
// store; V = 4: one float4 load, 8-byte store)
template <typename T, int V>
__device__ __forceinline__ void store_vec(T* p, const float (&a)[V]) {
  if constexpr (V == 8) {
    __align__(16) T t[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) t[i] = from_f32<T>(a[i]);
    *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(t);
  } else {
    Quad<T>::store(p, a[0], a[1], a[2], a[3]);
  }
}
template <typename S> __device__ __forceinline__ float2 unpack2s(uint32_t v);
template <> __device__ __forceinline__ float2 unpack2s<__half>(uint32_t v) { return __half22float2(*reinterpret_cast<const __half2*>(&v)); }
template <> __device__ __forceinline__ float2 unpack2s<__nv_bfloat16>(uint32_t v) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v));
}
// 16-bit source: V = 8 -> one 16-byte load, V = 4 -> one 8-byte load
template <int V, typename S>
__device__ __forceinline__ void load_vec(const S* p, float (&a)[V]) {
  uint32_t w[V / 2];
  if constexpr (V == 8) {
    const uint4 t = __ldg(reinterpret_cast<const uint4*>(p));
    w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w;
  } else {
    const uint2 t = __ldg(reinterpret_cast<const uint2*>(p));
    w[0] = t.x; w[1] = t.y;
  }
#pragma unroll
  for (int q = 0; q < V / 2; ++q) {
    const float2 f = unpack2s<S>(w[q]);
    a[2 * q] = f.x; a[2 * q + 1] = f.y;
  }
}
template <int V>
__device__ __forceinline__ void load_vec(const float* p, float (&a)[V]) {
#pragma unroll
  for (int q = 0; q < V / 4; ++q) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(p) + q);
    a[4 * q] = t.x; a[4 * q + 1] = t.y; a[4 * q + 2] = t.z; a[4 * q + 3] = t.w;
  }
}

template <typename T, int V, typename S = float>
__global__ void __launch_bounds__(256)
pool2x_kernel(const S* __restrict__ src, T* __restrict__ dst, int CV, int H, int W, int Ho, int Wo,
              size_t total) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int cv = (int)(i % CV);
  size_t r = i / CV;
  const int xo = (int)(r % Wo);
  r /= Wo;
  const int yo = (int)(r % Ho);
  const int b = (int)(r / Ho);
  const S* s = src + ((size_t)b * H * W * CV + cv) * V;
  float a[V];
#pragma unroll
  for (int j = 0; j < V; ++j) a[j] = 0.f;
#pragma unroll
  for (int dy = -1; dy <= 1; ++dy) {
    const int y = 2 * yo + dy;
    if (y < 0 || y >= H) continue;
#pragma unroll
    for (int dx = -1; dx <= 1; ++dx) {
      const int x = 2 * xo + dx;
      if (x < 0 || x >= W) continue;
      float v[V];
      load_vec<V>(s + ((size_t)y * W + x) * CV * V, v);
#pragma unroll
      for (int j = 0; j < V; ++j) a[j] += v[j];
    }
  }
  const float k = 1.f / 9.f;  // count_include_pad=True: the divisor is always 9
#pragma unroll
  for (int j = 0; j < V; ++j) a[j] *= k;
  store_vec<T, V>(dst + i * V, a);
}

template <typename T, int V, typename S = float>
__global__ void __launch_bounds__(256)
resize_kernel(const S* __restrict__ src, T* __restrict__ dst, int CV, int Hs, int Ws, int Hd, int Wd,
              float sy, float sx, size_t total) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int cv = (int)(i % CV);
  size_t r = i / CV;
  const int xo = (int)(r % Wd);
  r /= Wd;
  const int yo = (int)(r % Hd);
  const int b = (int)(r / Hd);
  // align_corners=True: src = dst * (in-1)/(out-1)
  const float fy = sy * (float)yo, fx = sx * (float)xo;
  const int y0 = min((int)fy, Hs - 1), x0 = min((int)fx, Ws - 1);
  const int y1 = min(y0 + 1, Hs - 1), x1 = min(x0 + 1, Ws - 1);
  const float ly = fy - (float)y0, lx = fx - (float)x0;
  const S* s = src + ((size_t)b * Hs * Ws * CV + cv) * V;
  float v00[V], v01[V], v10[V], v11[V], o[V];
  load_vec<V>(s + ((size_t)y0 * Ws + x0) * CV * V, v00);
  load_vec<V>(s + ((size_t)y0 * Ws + x1) * CV * V, v01);
  load_vec<V>(s + ((size_t)y1 * Ws + x0) * CV * V, v10);
  load_vec<V>(s + ((size_t)y1 * Ws + x1) * CV * V, v11);
  const float w00 = (1.f - ly) * (1.f - lx), w01 = (1.f - ly) * lx, w10 = ly * (1.f - lx), w11 = ly * lx;
#pragma unroll
  for (int j = 0; j < V; ++j) o[j] = w00 * v00[j] + w01 * v01[j] + w10 * v10[j] + w11 * v11[j];
  store_vec<T, V>(dst + i * V, o);
}

// ------------------------------------------------------------------ 7x7, one input channel
// CTA = 32 (x) x kC1Rows (y) pixels, the 49 x Cout weights (12.5 KB) staged once per 256 pixels.  warp = one row,
// lane = one pair of output channels, walking the row in strips of 8 pixels: per kernel row a lane reads the 14
// inputs of the strip (broadcast 16-byte loads) and 7 weight pairs for 112 FMAs, so the kernel issues FMAs, not
// shared-memory loads (thread = pixel x 8 channels needed 3 loads per 8 FMAs and ran at a third of this).
// A warp stores 128 contiguous bytes per pixel.
constexpr int kC1MaxCout = 64;
constexpr int kC1Rows = 8;
constexpr int kC1Strip = 8;
constexpr int kC1PatchW = 40;  // 32 + 6 columns, padded so that every strip starts 16-byte aligned

template <typename T> __device__ __forceinline__ void store_pair(T* p, float a, float b);
template <> __device__ __forceinline__ void store_pair<__half>(__half* p, float a, float b) {
  *reinterpret_cast<__half2*>(p) = __floats2half2_rn(a, b);
}
template <> __device__ __forceinline__ void store_pair<__nv_bfloat16>(__nv_bfloat16* p, float a, float b) {
  *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b);
}

template <typename T>
__global__ void __launch_bounds__(256, 4)
conv7x7_c1_kernel(const float* __restrict__ disp, const float* __restrict__ weight, const float* __restrict__ bias,
                  T* __restrict__ out, int ld_out, int Cout, T* __restrict__ disp_out, int ld_disp, int disp_ch,
                  int H, int W) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  __shared__ __align__(16) float wsm[49][kC1MaxCout];
  __shared__ __align__(16) float patch[kC1Rows + 6][kC1PatchW];
  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * 32, y0 = blockIdx.y * kC1Rows, b = blockIdx.z;
  for (int i = tid; i < 49 * kC1MaxCout; i += 256) {  // parameters: no kernel writes them
    const int co = i / 49, t = i - co * 49;
    wsm[t][co] = co < Cout ? weight[i] : 0.f;
  }
  pdl_wait();  // inputs written by the previous kernel are complete and visible
  const float* d = disp + (size_t)b * H * W;
  for (int i = tid; i < (kC1Rows + 6) * kC1PatchW; i += 256) {
    const int ry = i / kC1PatchW, rx = i - ry * kC1PatchW;
    const int yy = y0 + ry - 3, xx = x0 + rx - 3;
    patch[ry][rx] = (rx < 38 && yy >= 0 && yy < H && xx >= 0 && xx < W) ? d[(size_t)yy * W + xx] : 0.f;
  }
  __syncthreads();
  const int lane = tid & 31, r = tid >> 5;
  const int y = y0 + r, c = 2 * lane;
  if (y >= H) return;
  const bool cok = c < Cout;  // idle channel lanes still carry the disparity copy below
  const float b0 = cok ? __ldg(bias + c) : 0.f, b1 = cok ? __ldg(bias + c + 1) : 0.f;
#pragma unroll 1
  for (int s0 = 0; s0 < 32; s0 += kC1Strip) {
    if (x0 + s0 >= W) break;
    float acc[kC1Strip][2];
#pragma unroll
    for (int j = 0; j < kC1Strip; ++j) { acc[j][0] = b0; acc[j][1] = b1; }
#pragma unroll 1  // unrolled, ptxas hoists all 49 weight pairs and spills; 32 warps per SM cover the load latency
    for (int ky = 0; ky < 7; ++ky) {
      float in[kC1Strip + 8];  // 14 used
      const float4* prow = reinterpret_cast<const float4*>(&patch[r + ky][s0]);
#pragma unroll
      for (int q = 0; q < (kC1Strip + 8) / 4; ++q) {
        const float4 t = prow[q];
        in[4 * q] = t.x; in[4 * q + 1] = t.y; in[4 * q + 2] = t.z; in[4 * q + 3] = t.w;
      }
#pragma unroll
      for (int kx = 0; kx < 7; ++kx) {
        const float2 w = *reinterpret_cast<const float2*>(&wsm[ky * 7 + kx][c]);
#pragma unroll
        for (int j = 0; j < kC1Strip; ++j) {
          acc[j][0] = fmaf(in[j + kx], w.x, acc[j][0]);
          acc[j][1] = fmaf(in[j + kx], w.y, acc[j][1]);
        }
      }
    }
    const size_t pix0 = ((size_t)b * H + y) * W + x0 + s0;
#pragma unroll
    for (int j = 0; j < kC1Strip; ++j)
      if (cok && x0 + s0 + j < W) store_pair<T>(out + (pix0 + j) * ld_out + c, fmaxf(acc[j][0], 0.f), fmaxf(acc[j][1], 0.f));
    if (disp_out != nullptr && lane < kC1Strip && x0 + s0 + lane < W)
      disp_out[(pix0 + lane) * ld_disp + disp_ch] = from_f32<T>(patch[r + 3][s0 + lane + 3]);
  }
}

// ------------------------------------------------------------------ depthwise 7x7
// CTA = 8 (y) x 16 (x) output pixels x 64 channels.  The 14 x 22 input patch is staged in
// shared memory in its 16-bit storage type (39 KB), the 49 x 64 weights as fp32.
// thread = (4 channels, strip of 8 consecutive x, one row): per kernel row it reads 14
// input vectors (8 B) and 7 weight vectors (16 B) for 224 FMAs, so the kernel is FMA-bound
// instead of shared-memory-bound; a warp = 16 channel quads x 2 strips reads/writes whole
// 128-byte runs and its shared-memory accesses are conflict free.
constexpr int kDwTY = 8, kDwTX = 16, kDwC = 64, kDwPY = kDwTY + 6, kDwPX = kDwTX + 6, kDwR = 8;

template <typename T> __device__ __forceinline__ float4 load4_f32(const T* p);
template <> __device__ __forceinline__ float4 load4_f32<__half>(const __half* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x));
  const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}
template <> __device__ __forceinline__ float4 load4_f32<__nv_bfloat16>(const __nv_bfloat16* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.x));
  const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}

template <typename T> __device__ __forceinline__ float2 unpack2f(uint32_t v);
template <> __device__ __forceinline__ float2 unpack2f<__half>(uint32_t v) {
  return __half22float2(*reinterpret_cast<const __half2*>(&v));
}
template <> __device__ __forceinline__ float2 unpack2f<__nv_bfloat16>(uint32_t v) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v));
}

// PERSISTENT: a CTA keeps the 49 x 64 weights of its channel group in shared memory and walks the pixel tiles
// tile = blockIdx.x, blockIdx.x + gridDim.x, ...; the input patch of the NEXT tile streams in with cp.async
// (zero fill outside the image) while the current one is computed -- as one-tile CTAs the kernel spent 40 % of
// its stall samples waiting for the patch and re-read the weights per tile.
__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gsrc, bool valid) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  const int n = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(n) : "memory");
}

template <typename T>
__global__ void __launch_bounds__(256)
dwconv7x7_kernel(const T* __restrict__ x, const float* __restrict__ weight, const float* __restrict__ bias,
                 T* __restrict__ out, int C, int H, int W, int ntiles) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  extern __shared__ __align__(16) uint8_t dsm_raw[];
  float* wsm = reinterpret_cast<float*>(dsm_raw);                        // [49][kDwC]
  T* patch0 = reinterpret_cast<T*>(dsm_raw + 49 * kDwC * sizeof(float));  // 2 x [kDwPY][kDwPX][kDwC]
  constexpr int kPatchElems = kDwPY * kDwPX * kDwC;
  const int tid = threadIdx.x;
  const int tiles_x = (W + kDwTX - 1) / kDwTX, tiles_y = (H + kDwTY - 1) / kDwTY;
  const int c0 = blockIdx.y * kDwC;
  for (int i = tid; i < 49 * kDwC; i += 256) {  // parameters: no kernel writes them, no need to wait
    const int c = i / 49, t = i - c * 49;
    wsm[t * kDwC + c] = weight[(size_t)(c0 + c) * 49 + t];
  }
  pdl_wait();  // inputs written by the previous kernel are complete and visible

  auto prefetch = [&](int tile, T* patch) {
    const int txi = tile % tiles_x, r1 = tile / tiles_x;
    const int tx0 = txi * kDwTX, ty0 = (r1 % tiles_y) * kDwTY, b = r1 / tiles_y;
    // 8 channels (16 bytes) per copy, zero outside the image
    for (int i = tid; i < kDwPY * kDwPX * (kDwC / 8); i += 256) {
      const int c8 = i & (kDwC / 8 - 1), pp = i >> 3;
      const int py = pp / kDwPX, px = pp - py * kDwPX;
      const int yy = ty0 + py - 3, xx = tx0 + px - 3;
      const bool ok = yy >= 0 && yy < H && xx >= 0 && xx < W;
      const T* src = x + (((size_t)b * H + (ok ? yy : 0)) * W + (ok ? xx : 0)) * C + c0 + c8 * 8;
      cp_async16_zfill(patch + pp * kDwC + c8 * 8, src, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  const int cq = tid & 15, xs = (tid >> 4) & 1, row = tid >> 5;
  const float4 bv = *reinterpret_cast<const float4*>(bias + c0 + cq * 4);
  int buf = 0;
  if ((int)blockIdx.x < ntiles) prefetch(blockIdx.x, patch0);
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, buf ^= 1) {
    const T* patch = patch0 + buf * kPatchElems;
    const int next = tile + gridDim.x;
    // the other buffer was read by the previous iteration's compute, which ended with a __syncthreads
    if (next < ntiles) {
      prefetch(next, patch0 + (buf ^ 1) * kPatchElems);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const int txi = tile % tiles_x, r1 = tile / tiles_x;
    const int tx0 = txi * kDwTX, ty0 = (r1 % tiles_y) * kDwTY, b = r1 / tiles_y;
    const int y = ty0 + row, xb = tx0 + xs * kDwR;
    // fp32 accumulation as PACKED pairs (fma.rn.f32x2 = FFMA2): half the issue slots of scalar FFMAs for the same FMAs.
    // It does not raise the FMA rate (tools/fma_bench.cu: scalar 122.6, packed 113.1 FMA/clk/SM) and this kernel measured
    // 80 us in the scalar, packed and mixed forms alike: it is latency-bound at 13 warps per SM.
    float2 acc[kDwR][2];
#pragma unroll
    for (int j = 0; j < kDwR; ++j) { acc[j][0] = make_float2(bv.x, bv.y); acc[j][1] = make_float2(bv.z, bv.w); }
#pragma unroll
    for (int ky = 0; ky < 7; ++ky) {  // unrolled: the loads of row ky+1 are scheduled under the FMAs of row ky
      float2 in[kDwR + 6][2];
      const T* prow = patch + ((row + ky) * kDwPX + xs * kDwR) * kDwC + cq * 4;
#pragma unroll
      for (int j = 0; j < kDwR + 6; ++j) {
        const uint2 u = *reinterpret_cast<const uint2*>(prow + j * kDwC);
        in[j][0] = unpack2f<T>(u.x);
        in[j][1] = unpack2f<T>(u.y);
      }
#pragma unroll
      for (int kx = 0; kx < 7; ++kx) {
        const float4 w = *reinterpret_cast<const float4*>(wsm + (ky * 7 + kx) * kDwC + cq * 4);
        const float2 w01 = make_float2(w.x, w.y), w23 = make_float2(w.z, w.w);
#pragma unroll
        for (int j = 0; j < kDwR; ++j) {
          acc[j][0] = __ffma2_rn(in[j + kx][0], w01, acc[j][0]);
          acc[j][1] = __ffma2_rn(in[j + kx][1], w23, acc[j][1]);
        }
      }
    }
    if (y < H) {
      T* o = out + (((size_t)b * H + y) * W + xb) * C + c0 + cq * 4;
#pragma unroll
      for (int j = 0; j < kDwR; ++j)
        if (xb + j < W) Quad<T>::store(o + (size_t)j * C, acc[j][0].x, acc[j][0].y, acc[j][1].x, acc[j][1].y);
    }
    __syncthreads();  // everyone is done with `patch` before the next iteration's prefetch overwrites it
  }
}

// ------------------------------------------------------------------ 3x3 convolution C -> 1 (DispHead.conv[4])
// delta[b,y,x] = bias + sum_{t=(ky,kx)} <in[b, y+ky-1, x+kx-1, :], w[t][:]>   (core/update.py:28, zero padding 1).
// As an N = 16 tensor-core tile this layer used 6 % of the tensor pipe for 43 us; it is 1152 MACs per pixel of
// plain dot product.  CTA = 8 x 32 output pixels.  Phase 1: for every pixel of the 10 x 34 input patch the NINE
// dot products with the per-tap weight vectors (each input pixel is read ONCE, 4 lanes x 64 contiguous bytes per
// pixel, packed fp32 FMAs, weights broadcast from shared memory), reduced over the 4 lanes by shuffles into a
// shared [9][10][34] table; phase 2: an output pixel sums its 9 shifted entries.  Optionally also disp += delta
// (core/foundation_stereo.py:242), which saves the refinement loop a launch.
constexpr int kC1TY = 8, kC1TX = 32, kC1PY = kC1TY + 2, kC1PX = kC1TX + 2, kC1PS = kC1PX + 1;

template <typename T>
__global__ void __launch_bounds__(256, 2)
conv3x3_cout1_kernel(const T* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                     float* __restrict__ delta, float* __restrict__ disp, int C, int H, int W) {
  pdl_launch_dependents();
  extern __shared__ __align__(16) float c1_sm[];
  float* ws = c1_sm;                      // [9][C]
  float* part = c1_sm + 9 * C;            // [9][kC1PY][kC1PS]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, sub = lane & 3;
  for (int i = tid; i < 9 * C; i += 256) ws[i] = w[i];  // parameters: no kernel writes them
  const int tiles_x = (W + kC1TX - 1) / kC1TX, tiles_y = (H + kC1TY - 1) / kC1TY;
  const int txi = blockIdx.x % tiles_x, r1 = blockIdx.x / tiles_x;
  const int tx0 = txi * kC1TX, ty0 = (r1 % tiles_y) * kC1TY, b = r1 / tiles_y;
  pdl_wait();
  __syncthreads();
  // ---- phase 1: 9 partial dot products per patch pixel.  A warp takes 32 consecutive patch pixels per trip: lane
  // group g = lane >> 2 owns the FOUR pixels p0 + g + 8k (k = 0..3), so every weight vector read from shared memory
  // (an LDS.128 costs four wavefronts however many lanes share an address) feeds four pixels' FMAs -- with one pixel
  // per lane group the kernel was shared-memory bound (8 M wavefronts, 48 us).
  constexpr int kPix = 4;
  for (int p0 = warp * 32; p0 < kC1PY * kC1PX; p0 += 256) {
    const T* src[kPix];
    bool ok[kPix];
    int pp[kPix];
#pragma unroll
    for (int k = 0; k < kPix; ++k) {
      pp[k] = p0 + (lane >> 2) + 8 * k;
      const int py = pp[k] / kC1PX, px = pp[k] - py * kC1PX;
      const int yy = ty0 + py - 1, xx = tx0 + px - 1;
      ok[k] = pp[k] < kC1PY * kC1PX && yy >= 0 && yy < H && xx >= 0 && xx < W;
      src[k] = x + (((size_t)b * H + (ok[k] ? yy : 0)) * W + (ok[k] ? xx : 0)) * C + sub * 8;
    }
    float2 acc[kPix][9];
#pragma unroll
    for (int k = 0; k < kPix; ++k)
#pragma unroll
      for (int t = 0; t < 9; ++t) acc[k][t] = make_float2(0.f, 0.f);
    // the 4 lanes of a pixel read 64 contiguous bytes per chunk of 32 channels; two CTAs per SM (128 registers) cover
    // the load latency of each other
    for (int c0 = 0; c0 < C; c0 += 32) {
      float2 v[kPix][4];
#pragma unroll
      for (int k = 0; k < kPix; ++k) {
        const uint4 u = ok[k] ? *reinterpret_cast<const uint4*>(src[k] + c0) : make_uint4(0u, 0u, 0u, 0u);
        v[k][0] = unpack2f<T>(u.x); v[k][1] = unpack2f<T>(u.y); v[k][2] = unpack2f<T>(u.z); v[k][3] = unpack2f<T>(u.w);
      }
      const float* wc = ws + c0 + sub * 8;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        const float4 wa = *reinterpret_cast<const float4*>(wc + t * C);
        const float4 wb = *reinterpret_cast<const float4*>(wc + t * C + 4);
        const float2 w0 = make_float2(wa.x, wa.y), w1 = make_float2(wa.z, wa.w);
        const float2 w2 = make_float2(wb.x, wb.y), w3 = make_float2(wb.z, wb.w);
#pragma unroll
        for (int k = 0; k < kPix; ++k) {
          acc[k][t] = __ffma2_rn(v[k][0], w0, acc[k][t]);
          acc[k][t] = __ffma2_rn(v[k][1], w1, acc[k][t]);
          acc[k][t] = __ffma2_rn(v[k][2], w2, acc[k][t]);
          acc[k][t] = __ffma2_rn(v[k][3], w3, acc[k][t]);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < kPix; ++k) {
      const int py = pp[k] / kC1PX, px = pp[k] - py * kC1PX;
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        float r = acc[k][t].x + acc[k][t].y;
        r += __shfl_xor_sync(0xffffffffu, r, 1);
        r += __shfl_xor_sync(0xffffffffu, r, 2);
        if (sub == (t & 3) && pp[k] < kC1PY * kC1PX) part[(t * kC1PY + py) * kC1PS + px] = r;
      }
    }
  }
  __syncthreads();
  // ---- phase 2
  const int oy = tid >> 5, ox = tid & 31;
  const int y = ty0 + oy, xo = tx0 + ox;
  if (y < H && xo < W) {
    float v = __ldg(bias);
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) v += part[((ky * 3 + kx) * kC1PY + oy + ky) * kC1PS + ox + kx];
    const size_t o = ((size_t)b * H + y) * W + xo;
    delta[o] = v;
    if (disp) disp[o] += v;
  }
}

__global__ void __launch_bounds__(256) axpy_kernel(float* __restrict__ y, const float* __restrict__ x, size_t n) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] += x[i];
}

}  // namespace fsb

// -------------------------------------------------------------------------------------
#define FSB_DISPATCH16(dtype, CALL)                                   \
  do {                                                                \
    if ((dtype) == FSB_F16) { using T = __half; CALL; }               \
    else { using T = __nv_bfloat16; CALL; }                           \
  } while (0)

static inline bool is16(int dtype) { return dtype == FSB_F16 || dtype == FSB_BF16; }

extern "C" int fsb_nchw_to_nhwc(const float* src, void* dst16, float* dst32, int B, int C, int H, int W, int Cpad,
                                int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_nchw_to_nhwc");
  using namespace fsb;
  FSB_REQUIRE(src && (dst16 || dst32), FSB_ERR_ARG, "fsb_nchw_to_nhwc: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, FSB_ERR_ARG, "fsb_nchw_to_nhwc: non-positive size");
  FSB_REQUIRE(!dst16 || (is16(dtype) && Cpad >= C), FSB_ERR_ARG, "fsb_nchw_to_nhwc: dtype=%d Cpad=%d C=%d", dtype, Cpad, C);
  if (!dst16) Cpad = C;
  const int HW = H * W;
  FSB_REQUIRE(B < 65536 && ceil_div(Cpad, 32) < 65536, FSB_ERR_SHAPE, "fsb_nchw_to_nhwc: grid too large");
  dim3 grid(ceil_div(HW, 32), ceil_div(Cpad, 32), B);
  FSB_DISPATCH16(dtype, (launch_kernel(nchw_to_nhwc_kernel<T>, dim3(grid), dim3(256), 0, (cudaStream_t)stream, src, (T*)dst16, dst32, C, HW, Cpad)));
  FSB_CHECK_LAUNCH("fsb_nchw_to_nhwc");
  return FSB_OK;
}

extern "C" int fsb_nhwc_to_nchw(const float* src, float* dst, int B, int C, int H, int W, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_nhwc_to_nchw");
  using namespace fsb;
  FSB_REQUIRE(src && dst, FSB_ERR_ARG, "fsb_nhwc_to_nchw: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, FSB_ERR_ARG, "fsb_nhwc_to_nchw: non-positive size");
  FSB_REQUIRE(B < 65536 && ceil_div(C, 32) < 65536, FSB_ERR_SHAPE, "fsb_nhwc_to_nchw: grid too large");
  dim3 grid(ceil_div(H * W, 32), ceil_div(C, 32), B);
  launch_kernel(nhwc_to_nchw_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, src, dst, C, H * W);
  FSB_CHECK_LAUNCH("fsb_nhwc_to_nchw");
  return FSB_OK;
}

extern "C" int fsb_pool2x_nhwc(const float* src, void* dst, int B, int C, int H, int W, int dtype,
                               fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_pool2x_nhwc");
  using namespace fsb;
  FSB_REQUIRE(src && dst, FSB_ERR_ARG, "fsb_pool2x_nhwc: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0 && is16(dtype), FSB_ERR_ARG, "fsb_pool2x_nhwc: bad size or dtype");
  FSB_REQUIRE(C % 4 == 0 && aligned16(src), FSB_ERR_SHAPE, "fsb_pool2x_nhwc: C=%d must be a multiple of 4", C);
  const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
  const int V = (C % 8 == 0 && aligned16(dst)) ? 8 : 4;
  const size_t total = (size_t)B * Ho * Wo * (C / V);
  const unsigned blocks = (unsigned)((total + 255) / 256);
  if (V == 8) {
    FSB_DISPATCH16(dtype, (launch_kernel(pool2x_kernel<T, 8>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, src, (T*)dst, C / 8, H, W, Ho, Wo, total)));
  } else {
    FSB_DISPATCH16(dtype, (launch_kernel(pool2x_kernel<T, 4>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, src, (T*)dst, C / 4, H, W, Ho, Wo, total)));
  }
  FSB_CHECK_LAUNCH("fsb_pool2x_nhwc");
  return FSB_OK;
}

extern "C" int fsb_resize_nhwc(const float* src, void* dst, int B, int C, int Hs, int Ws, int Hd, int Wd,
                               int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_resize_nhwc");
  using namespace fsb;
  FSB_REQUIRE(src && dst, FSB_ERR_ARG, "fsb_resize_nhwc: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && Hs > 0 && Ws > 0 && Hd > 0 && Wd > 0 && is16(dtype), FSB_ERR_ARG,
              "fsb_resize_nhwc: bad size or dtype");
  FSB_REQUIRE(C % 4 == 0 && aligned16(src), FSB_ERR_SHAPE, "fsb_resize_nhwc: C=%d must be a multiple of 4", C);
  const float sy = Hd > 1 ? (float)(Hs - 1) / (float)(Hd - 1) : 0.f;
  const float sx = Wd > 1 ? (float)(Ws - 1) / (float)(Wd - 1) : 0.f;
  const int V = (C % 8 == 0 && aligned16(dst)) ? 8 : 4;
  const size_t total = (size_t)B * Hd * Wd * (C / V);
  const unsigned blocks = (unsigned)((total + 255) / 256);
  if (V == 8) {
    FSB_DISPATCH16(dtype, (launch_kernel(resize_kernel<T, 8>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, src, (T*)dst, C / 8, Hs, Ws, Hd, Wd, sy, sx, total)));
  } else {
    FSB_DISPATCH16(dtype, (launch_kernel(resize_kernel<T, 4>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, src, (T*)dst, C / 4, Hs, Ws, Hd, Wd, sy, sx, total)));
  }
  FSB_CHECK_LAUNCH("fsb_resize_nhwc");
  return FSB_OK;
}

// pool2x / interp with the source in the 16-bit storage type (the hidden state as the reference's autocast region holds
// it): half the bytes of the fp32 forms above
extern "C" int fsb_pool2x_nhwc16(const void* src, void* dst, int B, int C, int H, int W, int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_pool2x_nhwc16");
  using namespace fsb;
  FSB_REQUIRE(src && dst, FSB_ERR_ARG, "fsb_pool2x_nhwc16: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0 && is16(dtype), FSB_ERR_ARG, "fsb_pool2x_nhwc16: bad size or dtype");
  FSB_REQUIRE(C % 8 == 0 && aligned16(src) && aligned16(dst), FSB_ERR_SHAPE,
              "fsb_pool2x_nhwc16: C=%d must be a multiple of 8, pointers 16-byte aligned", C);
  const int Ho = (H + 1) / 2, Wo = (W + 1) / 2;
  const size_t total = (size_t)B * Ho * Wo * (C / 8);
  const unsigned blocks = (unsigned)((total + 255) / 256);
  FSB_DISPATCH16(dtype, (launch_kernel(pool2x_kernel<T, 8, T>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, (const T*)src, (T*)dst, C / 8, H, W, Ho, Wo, total)));
  FSB_CHECK_LAUNCH("fsb_pool2x_nhwc16");
  return FSB_OK;
}

extern "C" int fsb_resize_nhwc16(const void* src, void* dst, int B, int C, int Hs, int Ws, int Hd, int Wd, int dtype,
                                 fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_resize_nhwc16");
  using namespace fsb;
  FSB_REQUIRE(src && dst, FSB_ERR_ARG, "fsb_resize_nhwc16: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && Hs > 0 && Ws > 0 && Hd > 0 && Wd > 0 && is16(dtype), FSB_ERR_ARG,
              "fsb_resize_nhwc16: bad size or dtype");
  FSB_REQUIRE(C % 8 == 0 && aligned16(src) && aligned16(dst), FSB_ERR_SHAPE,
              "fsb_resize_nhwc16: C=%d must be a multiple of 8, pointers 16-byte aligned", C);
  const float sy = Hd > 1 ? (float)(Hs - 1) / (float)(Hd - 1) : 0.f;
  const float sx = Wd > 1 ? (float)(Ws - 1) / (float)(Wd - 1) : 0.f;
  const size_t total = (size_t)B * Hd * Wd * (C / 8);
  const unsigned blocks = (unsigned)((total + 255) / 256);
  FSB_DISPATCH16(dtype, (launch_kernel(resize_kernel<T, 8, T>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, (const T*)src, (T*)dst, C / 8, Hs, Ws, Hd, Wd, sy, sx, total)));
  FSB_CHECK_LAUNCH("fsb_resize_nhwc16");
  return FSB_OK;
}

extern "C" int fsb_conv7x7_c1(const float* disp, const float* weight, const float* bias, void* out, int ld_out,
                              int Cout, void* disp_out, int ld_disp, int disp_ch, int B, int H, int W, int dtype,
                              fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_conv7x7_c1");
  using namespace fsb;
  FSB_REQUIRE(disp && weight && bias && out, FSB_ERR_ARG, "fsb_conv7x7_c1: NULL pointer");
  FSB_REQUIRE(B > 0 && H > 0 && W > 0 && is16(dtype), FSB_ERR_ARG, "fsb_conv7x7_c1: bad size or dtype");
  FSB_REQUIRE(Cout > 0 && Cout <= kC1MaxCout && Cout % 8 == 0 && ld_out % 4 == 0 && ld_out >= Cout, FSB_ERR_SHAPE,
              "fsb_conv7x7_c1: Cout=%d (multiple of 8, <= %d), ld_out=%d", Cout, kC1MaxCout, ld_out);
  FSB_REQUIRE(!disp_out || (disp_ch >= 0 && disp_ch < ld_disp), FSB_ERR_ARG, "fsb_conv7x7_c1: disp_ch=%d ld_disp=%d", disp_ch, ld_disp);
  FSB_REQUIRE(H < 65536 && B < 65536, FSB_ERR_SHAPE, "fsb_conv7x7_c1: grid too large");
  dim3 grid(ceil_div(W, 32), ceil_div(H, kC1Rows), B);
  FSB_DISPATCH16(dtype, (launch_kernel(conv7x7_c1_kernel<T>, dim3(grid), dim3(256), 0, (cudaStream_t)stream, disp, weight, bias, (T*)out, ld_out, Cout,
                                                                                       (T*)disp_out, ld_disp, disp_ch, H, W)));
  FSB_CHECK_LAUNCH("fsb_conv7x7_c1");
  return FSB_OK;
}

extern "C" int fsb_dwconv7x7_nhwc(const void* x, const float* weight, const float* bias, void* out, int B, int C,
                                  int H, int W, int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_dwconv7x7_nhwc");
  using namespace fsb;
  FSB_REQUIRE(x && weight && bias && out, FSB_ERR_ARG, "fsb_dwconv7x7_nhwc: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0 && is16(dtype), FSB_ERR_ARG, "fsb_dwconv7x7_nhwc: bad size or dtype");
  FSB_REQUIRE(C % kDwC == 0 && aligned16(x) && aligned16(out) && aligned16(bias), FSB_ERR_SHAPE, "fsb_dwconv7x7_nhwc: C=%d must be a multiple of %d", C, kDwC);
  const size_t smem = (size_t)49 * kDwC * sizeof(float) + (size_t)2 * kDwPY * kDwPX * kDwC * 2;
  const long long ntiles = (long long)ceil_div(W, kDwTX) * ceil_div(H, kDwTY) * B;
  FSB_REQUIRE(ntiles < (1ll << 31), FSB_ERR_SHAPE, "fsb_dwconv7x7_nhwc: too many tiles");
  // two CTAs per SM (91 KB each); every channel group gets the whole machine
  int sms = 0, dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
    sms = 148;
  const int groups = C / kDwC;
  long long per_group = (2ll * sms + groups - 1) / groups;
  if (per_group > ntiles) per_group = ntiles;
  dim3 grid((unsigned)per_group, (unsigned)groups);
  if (dtype == FSB_F16) {
    cudaError_t e = cudaFuncSetAttribute(dwconv7x7_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "fsb_dwconv7x7_nhwc: cudaFuncSetAttribute");
    launch_kernel(dwconv7x7_kernel<__half>, dim3(grid), dim3(256), smem, (cudaStream_t)stream, (const __half*)x, weight, bias, (__half*)out, C, H, W,
                  (int)ntiles);
  } else {
    cudaError_t e = cudaFuncSetAttribute(dwconv7x7_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "fsb_dwconv7x7_nhwc: cudaFuncSetAttribute");
    launch_kernel(dwconv7x7_kernel<__nv_bfloat16>, dim3(grid), dim3(256), smem, (cudaStream_t)stream, (const __nv_bfloat16*)x, weight, bias,
                  (__nv_bfloat16*)out, C, H, W, (int)ntiles);
  }
  FSB_CHECK_LAUNCH("fsb_dwconv7x7_nhwc");
  return FSB_OK;
}

extern "C" int fsb_conv3x3_cout1(const void* x, const float* weight, const float* bias, float* delta, float* disp_accum,
                                 int B, int C, int H, int W, int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_conv3x3_cout1");
  using namespace fsb;
  FSB_REQUIRE(x && weight && bias && delta, FSB_ERR_ARG, "fsb_conv3x3_cout1: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0 && is16(dtype), FSB_ERR_ARG, "fsb_conv3x3_cout1: bad size or dtype");
  FSB_REQUIRE(C % 32 == 0 && C <= 1024 && aligned16(x) && aligned16(weight), FSB_ERR_SHAPE,
              "fsb_conv3x3_cout1: C=%d must be a multiple of 32 (<= 1024), x and weight 16-byte aligned", C);
  const long long tiles = (long long)B * ceil_div(H, kC1TY) * ceil_div(W, kC1TX);
  FSB_REQUIRE(tiles < (1ll << 31), FSB_ERR_SHAPE, "fsb_conv3x3_cout1: too many tiles");
  const size_t smem = (size_t)(9 * C + 9 * kC1PY * kC1PS) * sizeof(float);
  if (dtype == FSB_F16) {
    cudaError_t e = cudaFuncSetAttribute(conv3x3_cout1_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "fsb_conv3x3_cout1: cudaFuncSetAttribute");
    launch_kernel(conv3x3_cout1_kernel<__half>, dim3((unsigned)tiles), dim3(256), smem, (cudaStream_t)stream, (const __half*)x, weight, bias, delta, disp_accum, C, H, W);
  } else {
    cudaError_t e = cudaFuncSetAttribute(conv3x3_cout1_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "fsb_conv3x3_cout1: cudaFuncSetAttribute");
    launch_kernel(conv3x3_cout1_kernel<__nv_bfloat16>, dim3((unsigned)tiles), dim3(256), smem, (cudaStream_t)stream, (const __nv_bfloat16*)x, weight, bias, delta, disp_accum, C, H, W);
  }
  FSB_CHECK_LAUNCH("fsb_conv3x3_cout1");
  return FSB_OK;
}

extern "C" int fsb_axpy_f32(float* y, const float* x, size_t n, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_axpy_f32");
  using namespace fsb;
  FSB_REQUIRE(y && x, FSB_ERR_ARG, "fsb_axpy_f32: NULL pointer");
  if (n == 0) return FSB_OK;
  launch_kernel(axpy_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, (cudaStream_t)stream, y, x, n);
  FSB_CHECK_LAUNCH("fsb_axpy_f32");
  return FSB_OK;
}
