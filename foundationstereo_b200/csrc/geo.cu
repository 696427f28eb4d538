// Geometry-encoding volume: init correlation + pyramid (kernel 2, fp32 parity path),
// geometry-volume transpose + pyramid (kernel 2b) and the per-iteration lookup (kernel 3).
//
// Replaces Combined_Geo_Encoding_Volume (core/geometry.py:17-77) and the H == 1 use of
// bilinear_sampler (core/utils/utils.py:44-55) of the reference.
#include "common.cuh"
#include "lookup.cuh"

#include <type_traits>

namespace fsb {

// =====================================================================================
// Kernel 2 (FSB_PREC_FP32): corr[b,h,w1,w2] = <f1hat(:,w1), f2hat(:,w2)>, fp32 FFMA.
// One CTA = 64 (w1) x 128 (w2) outputs of one image row; the (C,W) planes are already
// "K-major with W contiguous", so the smem tiles are filled with coalesced loads.  The
// L2 norms are accumulated from the same tiles and applied in the epilogue; the
// epilogue also emits every pooled pyramid level from registers (a thread owns 8
// consecutive w2), so the W x W tensor is never re-read.
// =====================================================================================
constexpr int kCorrKC = 16;

template <bool VEC>
__device__ __forceinline__ void store_run4(float* row, int col, int len, const float* v) {
  if (VEC) {
    if (col < len) *reinterpret_cast<float4*>(row + col) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (col + i < len) row[col + i] = v[i];
  }
}

__global__ void __launch_bounds__(256)
corr_fp32_kernel(const float* __restrict__ f1, const float* __restrict__ f2, LevelPtrs lv, int C, int H,
                 int W1, int W2, int L) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  __shared__ __align__(16) float As[kCorrKC][64];
  __shared__ __align__(16) float Bs[kCorrKC][128];
  __shared__ float invA[64], invB[128];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int bh = blockIdx.z, b = bh / H, h = bh - b * H;
  const int w1_0 = blockIdx.y * 64, w2_0 = blockIdx.x * 128;
  const size_t HW1 = (size_t)H * W1, HW2 = (size_t)H * W2;
  const float* a_base = f1 + (size_t)b * C * HW1 + (size_t)h * W1;
  const float* b_base = f2 + (size_t)b * C * HW2 + (size_t)h * W2;

  float acc[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  float ss = 0.f;  // thread t<64: column t of A; 64<=t<192: column t-64 of B

  for (int c0 = 0; c0 < C; c0 += kCorrKC) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + i * 256, k = idx >> 6, w = idx & 63;
      As[k][w] = (c0 + k < C && w1_0 + w < W1) ? a_base[(size_t)(c0 + k) * HW1 + w1_0 + w] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int idx = tid + i * 256, k = idx >> 7, w = idx & 127;
      Bs[k][w] = (c0 + k < C && w2_0 + w < W2) ? b_base[(size_t)(c0 + k) * HW2 + w2_0 + w] : 0.f;
    }
    __syncthreads();
    if (tid < 64) {
#pragma unroll
      for (int k = 0; k < kCorrKC; ++k) ss = fmaf(As[k][tid], As[k][tid], ss);
    } else if (tid < 192) {
#pragma unroll
      for (int k = 0; k < kCorrKC; ++k) ss = fmaf(Bs[k][tid - 64], Bs[k][tid - 64], ss);
    }
#pragma unroll
    for (int k = 0; k < kCorrKC; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[k][tx * 8]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[k][tx * 8 + 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  // F.normalize denominators: max(||x||, 1e-12)
  if (tid < 64) invA[tid] = fmaxf(sqrtf(ss), 1e-12f);
  else if (tid < 192) invB[tid - 64] = fmaxf(sqrtf(ss), 1e-12f);
  __syncthreads();

  const bool vec0 = (W2 & 3) == 0;
  const int c2 = w2_0 + tx * 8;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int w1 = w1_0 + ty * 4 + i;
    if (w1 >= W1) continue;
    const size_t p = (size_t)bh * W1 + w1;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = __fdiv_rn(__fdiv_rn(acc[i][j], invA[ty * 4 + i]), invB[tx * 8 + j]);
    float* row = lv.p[0] + p * W2;
    if (vec0) { store_run4<true>(row, c2, W2, v); store_run4<true>(row, c2 + 4, W2, v + 4); }
    else      { store_run4<false>(row, c2, W2, v); store_run4<false>(row, c2 + 4, W2, v + 4); }
    // pooled levels: mean of adjacent pairs of the previous level (floor length)
    int n = 8, len = W2, col = c2;
    for (int l = 1; l < L; ++l) {
      n >>= 1; len >>= 1; col >>= 1;
      if (n == 0) break;  // host guarantees L <= 4
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < n) v[j] = (v[2 * j] + v[2 * j + 1]) * 0.5f;
      float* r = lv.p[l] + p * len;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < n && col + j < len) r[col + j] = v[j];
    }
  }
}

// =====================================================================================
// Kernel 2b: (B,Cg,D,H,W) -> level l (B*H*W, D>>l, Cg).  Tile = 32 w x 8 d x all Cg.
// Reads are 128-byte runs along W, writes are runs of (8>>l)*Cg floats per pixel.
// =====================================================================================
constexpr int kGeoTW = 32, kGeoTD = 8;

template <int LVL>
__device__ __forceinline__ float pooled(const float* t, int Cg, int jj, int c, int w) {
  if constexpr (LVL == 0) {
    return t[(jj * Cg + c) * (kGeoTW + 1) + w];
  } else {
    return (pooled<LVL - 1>(t, Cg, 2 * jj, c, w) + pooled<LVL - 1>(t, Cg, 2 * jj + 1, c, w)) * 0.5f;
  }
}

template <int LVL>
__device__ __forceinline__ void write_level(const float* t, float* out, int Cg, int D, int d0, size_t p0,
                                            int wvalid, int tid) {
  const int Dl = D >> LVL;
  const int rows = kGeoTD >> LVL;  // pooled rows in this tile
  const int run = rows * Cg;
  const int dl0 = d0 >> LVL;
  for (int i = tid; i < kGeoTW * run; i += 256) {
    const int w = i / run, j = i - w * run;
    const int jj = j / Cg, c = j - jj * Cg;
    if (w < wvalid && dl0 + jj < Dl) out[((p0 + w) * Dl + dl0 + jj) * Cg + c] = pooled<LVL>(t, Cg, jj, c, w);
  }
}

template <typename TIN>
__global__ void __launch_bounds__(256)
geo_pyramid_kernel(const TIN* __restrict__ vol, LevelPtrs lv, int Cg, int D, int H, int W, int L) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  extern __shared__ float t[];  // [kGeoTD*Cg][kGeoTW+1], row index = dd*Cg + c
  const int tid = threadIdx.x;
  const int w0 = blockIdx.x * kGeoTW, d0 = blockIdx.y * kGeoTD;
  const int bh = blockIdx.z, b = bh / H, h = bh - b * H;
  const size_t HW = (size_t)H * W;
  const TIN* src = vol + (size_t)b * Cg * D * HW + (size_t)h * W + w0;
  const int n = Cg * kGeoTD * kGeoTW;
  for (int i = tid; i < n; i += 256) {
    const int w = i & (kGeoTW - 1);
    const int r = i >> 5;  // = c*kGeoTD + dd
    const int c = r >> 3, dd = r & 7;
    float v = 0.f;
    if (d0 + dd < D && w0 + w < W) v = to_f32<TIN>(src[((size_t)c * D + d0 + dd) * HW + w]);
    t[(dd * Cg + c) * (kGeoTW + 1) + w] = v;
  }
  __syncthreads();
  const size_t p0 = (size_t)bh * W + w0;
  const int wvalid = min(kGeoTW, W - w0);
  write_level<0>(t, lv.p[0], Cg, D, d0, p0, wvalid, tid);
  if (L > 1) write_level<1>(t, lv.p[1], Cg, D, d0, p0, wvalid, tid);
  if (L > 2) write_level<2>(t, lv.p[2], Cg, D, d0, p0, wvalid, tid);
  if (L > 3) write_level<3>(t, lv.p[3], Cg, D, d0, p0, wvalid, tid);
}

// Kernel 2b, fast path for compile-time Cg (28 in FoundationStereo): same tile (32 w x 8 d x Cg)
// but NO per-element index arithmetic -- the generic kernel above spends ~100 instructions per
// element on runtime divisions and is issue-bound at 1/3 of HBM speed.  Here thread j owns
// element j = dd*Cg + c of every pixel's 8*Cg-float output run (level 0) and loops over the 32
// pixels: smem reads t[j][w] are conflict free (row stride 33), global writes are whole
// 128-byte lines of a contiguous run, and the pooled levels reuse the same shared tile.
template <typename TIN> __device__ __forceinline__ float2 geo_unpack2(uint32_t v) { return make_float2(0.f, 0.f); }
template <> __device__ __forceinline__ float2 geo_unpack2<__half>(uint32_t v) {
  return __half22float2(*reinterpret_cast<const __half2*>(&v));
}
template <> __device__ __forceinline__ float2 geo_unpack2<__nv_bfloat16>(uint32_t v) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v));
}

template <int CG, typename TIN>
__global__ void __launch_bounds__(256)
geo_pyramid_fast_kernel(const TIN* __restrict__ vol, LevelPtrs lv, int D, int H, int W, int L) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  constexpr int R0 = kGeoTD * CG;  // rows of the tile = floats of a level-0 run
  static_assert(R0 <= 256, "one thread per run element");
  __shared__ float t[R0][kGeoTW + 1];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int w0 = blockIdx.x * kGeoTW, d0 = blockIdx.y * kGeoTD;
  const int bh = blockIdx.z, b = bh / H, h = bh - b * H;
  const size_t HW = (size_t)H * W;
  const TIN* src = vol + (size_t)b * CG * D * HW + (size_t)h * W + w0;
  const bool wok = w0 + lane < W;
  // loads: warp = one (c, dd) row of 32 consecutive w (one 128-byte line), 8 rows in flight per warp
  // (all CG loads of a warp are issued before the first is used: ~14 KB in flight per CTA)
  static_assert(kGeoTD == 8, "row r = c*8 + dd with r = warp + 8k  =>  dd = warp, c = k");
  bool paired = false;
  if constexpr (sizeof(TIN) == 2 && (CG % 2) == 0) {
    // 16-bit volume: a row of 32 consecutive w is only 64 bytes, and with one row per warp load the kernel ran at half
    // the fp32 form's speed (it is bound by requests in flight, not bytes: 1588 vs 767 us at 8 pairs).  Here a warp load
    // covers TWO (c, d) rows -- lanes 0-15 one, lanes 16-31 the other, two elements (4 bytes) per lane -- so a tile
    // takes half as many loads as the fp32 form.  Needs even W and an even tile start (4-byte aligned pairs).
    paired = (W & 1) == 0 && (reinterpret_cast<uintptr_t>(vol) & 3) == 0;
    if (paired) {
      const int hl = lane >> 4, wl = 2 * (lane & 15);
      const bool wok2 = w0 + wl < W;   // W even: the pair is inside or outside as a whole
      uint32_t u[CG / 2];
#pragma unroll
      for (int k = 0; k < CG / 2; ++k) {
        const int pi = warp + 8 * k;            // 0 .. 4*CG-1: (channel, pair of d rows)
        const int c = pi >> 2, dd = 2 * (pi & 3) + hl;
        u[k] = (wok2 && d0 + dd < D)
                   ? __ldg(reinterpret_cast<const uint32_t*>(src + ((size_t)c * D + d0 + dd) * HW + wl)) : 0u;
      }
#pragma unroll
      for (int k = 0; k < CG / 2; ++k) {
        const int pi = warp + 8 * k;
        const int c = pi >> 2, dd = 2 * (pi & 3) + hl;
        const float2 f = geo_unpack2<TIN>(u[k]);
        t[dd * CG + c][wl] = f.x;
        t[dd * CG + c][wl + 1] = f.y;
      }
    }
  }
  if (!paired) {
    float v[CG];
    const bool dok = wok && d0 + warp < D;
#pragma unroll
    for (int k = 0; k < CG; ++k) v[k] = dok ? to_f32<TIN>(__ldg(src + ((size_t)k * D + d0 + warp) * HW + lane)) : 0.f;
#pragma unroll
    for (int k = 0; k < CG; ++k) t[warp * CG + k][lane] = v[k];
  }
  __syncthreads();
  const size_t p0 = (size_t)bh * W + w0;
  const int wvalid = min(kGeoTW, W - w0);
  // level 0: run element j = dd*CG + c
  if (tid < R0) {
    const int dd = tid / CG;
    if (d0 + dd < D) {
      float* o = lv.p[0] + (p0 * D + d0) * CG + tid;
      for (int w = 0; w < wvalid; ++w) o[(size_t)w * D * CG] = t[tid][w];
    }
  }
  // pooled levels: element j = jj*CG + c of a run of (8 >> l)*CG floats; the warps not needed by a
  // level are the ones that start on the next (levels are issued from the top thread ids down)
  if (L > 1) {
    const int u = 255 - tid;
    if (u < R0 / 2) {
      const int jj = u / CG, c = u - jj * CG, D1 = D >> 1, dl = (d0 >> 1) + jj;
      if (dl < D1) {
        float* o = lv.p[1] + (p0 * D1 + dl) * CG + c;
        const float* a = t[(2 * jj) * CG + c];
        const float* bb = t[(2 * jj + 1) * CG + c];
        for (int w = 0; w < wvalid; ++w) o[(size_t)w * D1 * CG] = (a[w] + bb[w]) * 0.5f;
      }
    }
  }
  if (L > 2) {
    for (int l = 2; l < L; ++l) {
      const int rows = kGeoTD >> l, Dl = D >> l;
      if (tid < rows * CG) {
        const int jj = tid / CG, c = tid - jj * CG, dl = (d0 >> l) + jj;
        if (dl < Dl) {
          float* o = lv.p[l] + (p0 * Dl + dl) * CG + c;
          for (int w = 0; w < wvalid; ++w) {
            // mean of adjacent pairs, level by level (same rounding as pooling level l-1 again)
            float acc;
            if (l == 2) {
              const float x0 = (t[(4 * jj) * CG + c][w] + t[(4 * jj + 1) * CG + c][w]) * 0.5f;
              const float x1 = (t[(4 * jj + 2) * CG + c][w] + t[(4 * jj + 3) * CG + c][w]) * 0.5f;
              acc = (x0 + x1) * 0.5f;
            } else {
              float y[2];
#pragma unroll
              for (int q = 0; q < 2; ++q) {
                const int e = 8 * jj + 4 * q;
                const float x0 = (t[e * CG + c][w] + t[(e + 1) * CG + c][w]) * 0.5f;
                const float x1 = (t[(e + 2) * CG + c][w] + t[(e + 3) * CG + c][w]) * 0.5f;
                y[q] = (x0 + x1) * 0.5f;
              }
              acc = (y[0] + y[1]) * 0.5f;
            }
            o[(size_t)w * Dl * CG] = acc;
          }
        }
      }
    }
  }
}

// =====================================================================================
// Kernel 3: lookup (core/geometry.py:43-65, bilinear_sampler core/utils/utils.py:44-55).
//   positions: exactly as the reference computes them (x -> 2x/(n-1)-1 -> ((.+1)/2)*(n-1), fp32,
//   no FMA contraction); window base = floor of the first tap, per-tap fraction relative to base+k.
//   In the (BHW, D_l, Cg) layout the (K+1) x Cg window of a pixel is one contiguous run.
// =====================================================================================
constexpr int kLkPix = 32;
constexpr int kLkThreads = 256;
constexpr int kLkKMax = 9;  // radius <= 4


// =====================================================================================
// Kernel 3, fused output (out_layout 1/2): the same samples written as the 16-bit NHWC tensor
// (B,H,W,Cpad) that encoder.convc1 (core/update.py:56,63) reads as its GEMM A operand -- the
// "packing of the motion-encoder input".  One WARP per pixel, LANE = CHANNEL:
//   phase A: lanes 0..K-1 / 16..16+K-1 compute the geo / corr sampling positions of every
//            level (same fp32 round trip as above); window bases broadcast by shuffle, the
//            per-tap fractions stay in the lane that computed them;
//   phase B: in the (BHW, D_l, Cg) layout row dd of a pixel's window is Cg contiguous floats:
//            lane c loads element c of the K+1 rows (one coalesced 4*Cg-byte request per row,
//            no shared-memory staging), lane Cg loads the K+1 correlation taps.  The loads of
//            ALL levels are issued before the first use (memory-level parallelism);
//   phase C: lane c lerps its K outputs (fractions by shuffle) into a shared-memory row in
//            the reference's channel order (level, c*K + k | K corr taps), which leaves as
//            16-byte coalesced stores.
// Channel groups beyond the first 32 (Cg + 1 > 32) are handled by repeating B/C per group.
// =====================================================================================
constexpr int kLkWarps = 4;

template <typename T, int KT, int LT, int CGT>
__global__ void __launch_bounds__(kLkWarps * 32, 12)
lookup_nhwc_kernel(LevelPtrs geo, LevelPtrs cor, const float* __restrict__ disp, const float* __restrict__ coords,
                   T* __restrict__ out, int Cgrt, int D, int W2, int Lrt, int r, int Cpad, long long npix, int W1, CorrBand band) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  extern __shared__ __align__(16) uint8_t lsm[];
  constexpr int K = KT;
  constexpr int NL = LT > 0 ? LT : 4;
  const int L = LT > 0 ? LT : Lrt;
  const int Cg = CGT > 0 ? CGT : Cgrt;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long pix = (long long)blockIdx.x * kLkWarps + warp;
  if (pix >= npix) return;
  T* orow = reinterpret_cast<T*>(lsm + (size_t)warp * Cpad * sizeof(T));
  const float dsp = __ldg(disp + pix), xco = __ldg(coords + pix);
  const int nout = K * (Cg + 1);
  const int bs0 = band.bw0 > 0 ? band_start(band, (int)(pix % W1)) : 0;  // the pixel's image column names its band

  for (int c0 = 0; c0 < Cg; c0 += 32) {
    const int c = c0 + lane;
    const bool geo_ch = c < Cg;
    float v[NL][K + 1], cv[NL], frac[NL];
    // ---- phases A + B, all levels before the first use
#pragma unroll
    for (int l = 0; l < NL; ++l) {
      if (l >= L) break;
      const float inv = 1.f / (float)(1 << l);
      const float d = dsp * inv;  // exact: power-of-two scale
      lookup_level<K>(geo.p[l], cor.p[l], pix, Cg, D >> l, W2 >> l, d, __fsub_rn(xco * inv, d), r, lane, c, geo_ch,
                      v[l], cv[l], frac[l], band.bw0 > 0 ? band.bw0 >> l : W2 >> l, bs0 >> l);
    }
    // ---- phase C.  The 16-bit result cannot tell lo*(1-f) + hi*f from lo + f*(hi-lo): the shorter form is used.
#pragma unroll
    for (int l = 0; l < NL; ++l) {
      if (l >= L) break;
      T* o = orow + l * nout + c * K;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const float f = __shfl_sync(0xffffffffu, frac[l], k);
        const float s = fmaf(f, v[l][k + 1] - v[l][k], v[l][k]);
        if (geo_ch) o[k] = from_f32<T>(s);
      }
      if (c0 == 0) {
        const float hi = __shfl_down_sync(0xffffffffu, cv[l], 1);
        const float f = __shfl_sync(0xffffffffu, frac[l], (lane + 16) & 31);
        if (lane < K) orow[l * nout + Cg * K + lane] = from_f32<T>(fmaf(f, hi - cv[l], cv[l]));
      }
    }
  }
  for (int j = L * nout + lane; j < Cpad; j += 32) orow[j] = from_f32<T>(0.f);
  __syncwarp();
  const uint4* src = reinterpret_cast<const uint4*>(orow);
  uint4* dst = reinterpret_cast<uint4*>(out + (size_t)pix * Cpad);
  for (int i = lane; i < Cpad * (int)sizeof(T) / 16; i += 32) dst[i] = src[i];
}

// Kernel 3, reference output (out_layout 0): (B, L*K*(Cg+1), H, W) fp32, the tensor geo_fn returns in the
// reference (core/geometry.py:60-65).  Same lane = channel gather as above; one CTA owns 32 consecutive
// pixels of an image row, each warp gathers four of them per level into a [K*(Cg+1)][32 + 1] shared tile
// (channel-major: bank = 9c + k + pixel, conflict free), and the tile leaves as 128-byte rows along W.
// The lerp keeps the reference's form lo*(1-f) + hi*f.
template <int KT, int CGT>
__global__ void __launch_bounds__(kLkThreads, 4)
lookup_nchw_kernel(LevelPtrs geo, LevelPtrs cor, const float* __restrict__ disp, const float* __restrict__ coords,
                   float* __restrict__ out, int Cgrt, int D, int H, int W, int W2, int L, int r, CorrBand band) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  extern __shared__ float sm[];  // [nout][33]
  constexpr int K = KT;
  constexpr int PPW = kLkPix / (kLkThreads / 32);  // pixels per warp
  const int Cg = CGT > 0 ? CGT : Cgrt;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int w0 = blockIdx.x * kLkPix;
  const int bh = blockIdx.y, b = bh / H, h = bh - b * H;
  const long long p0 = (long long)bh * W + w0;
  const int npix = min(kLkPix, W - w0);
  const int nout = K * (Cg + 1);
  const size_t HW = (size_t)H * W;
  float* obase = out + ((size_t)b * L * nout * H + h) * W + w0;
  float dsp[PPW], xco[PPW];
#pragma unroll
  for (int i = 0; i < PPW; ++i) {
    const int px = warp * PPW + i;
    dsp[i] = px < npix ? __ldg(disp + p0 + px) : 0.f;
    xco[i] = px < npix ? __ldg(coords + p0 + px) : 0.f;
  }
  for (int l = 0; l < L; ++l) {
    const int Dl = D >> l, Wl = W2 >> l;
    const float inv = 1.f / (float)(1 << l);
    for (int c0 = 0; c0 < Cg; c0 += 32) {
      const int c = c0 + lane;
      const bool geo_ch = c < Cg;
#pragma unroll
      for (int i0 = 0; i0 < PPW; i0 += 2) {
        float v[2][K + 1], cv[2], frac[2];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int px = warp * PPW + i0 + i;
          const float d = dsp[i0 + i] * inv;
          // a slot beyond the row end gathers the row's last pixel again (harmless, never stored)
          const int bs0 = band.bw0 > 0 ? band_start(band, w0 + min(px, npix - 1)) : 0;
          lookup_level<K>(geo.p[l], cor.p[l], p0 + min(px, npix - 1), Cg, Dl, Wl, d, __fsub_rn(xco[i0 + i] * inv, d), r, lane, c, geo_ch,
                          v[i], cv[i], frac[i], band.bw0 > 0 ? band.bw0 >> l : Wl, bs0 >> l);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const int px = warp * PPW + i0 + i;
          const bool live = px < npix;  // warp-uniform
          float* t = sm + (c * K) * (kLkPix + 1) + px;
#pragma unroll
          for (int k = 0; k < K; ++k) {
            const float f = __shfl_sync(0xffffffffu, frac[i], k);
            if (live && geo_ch) t[k * (kLkPix + 1)] = v[i][k] * (1.f - f) + v[i][k + 1] * f;
          }
          if (c0 == 0) {
            const float hi = __shfl_down_sync(0xffffffffu, cv[i], 1);
            const float f = __shfl_sync(0xffffffffu, frac[i], (lane + 16) & 31);
            if (live && lane < K) sm[(Cg * K + lane) * (kLkPix + 1) + px] = cv[i] * (1.f - f) + hi * f;
          }
        }
      }
    }
    __syncthreads();
    if (lane < npix) {
      float* ol = obase + (size_t)l * nout * HW + lane;
      for (int j = warp; j < nout; j += kLkThreads / 32) ol[(size_t)j * HW] = sm[j * (kLkPix + 1) + lane];
    }
    __syncthreads();
  }
}

static size_t lookup_smem_bytes(int Cg, int r) {
  const int K = 2 * r + 1;
  return (size_t)K * (Cg + 1) * (kLkPix + 1) * sizeof(float);
}

}  // namespace fsb

// -------------------------------------------------------------------------------------
extern "C" int fsb_corr_pyramid(const float* f1, const float* f2, float* const* corr_levels, int B, int C,
                                int H, int W1, int W2, int L, int precision, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_corr_pyramid");
  using namespace fsb;
  FSB_REQUIRE(f1 && f2 && corr_levels, FSB_ERR_ARG, "fsb_corr_pyramid: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && H > 0 && W1 > 0 && W2 > 0, FSB_ERR_ARG, "fsb_corr_pyramid: non-positive size");
  FSB_REQUIRE(L >= 1 && L <= 4, FSB_ERR_SHAPE, "fsb_corr_pyramid: num_levels=%d outside [1,4]", L);
  FSB_REQUIRE((W2 >> (L - 1)) >= 1, FSB_ERR_SHAPE, "fsb_corr_pyramid: W2=%d too small for %d levels", W2, L);
  FSB_REQUIRE(precision == FSB_PREC_FP32 || precision == FSB_PREC_TF32 || precision == FSB_PREC_TF32X3, FSB_ERR_ARG,
              "fsb_corr_pyramid: precision %d (FSB_PREC_FP32, FSB_PREC_TF32 or FSB_PREC_TF32X3)", precision);
  FSB_REQUIRE((long long)B * H < 65536, FSB_ERR_SHAPE, "fsb_corr_pyramid: B*H=%lld >= 65536", (long long)B * H);
  LevelPtrs lv{};
  for (int l = 0; l < L; ++l) {
    FSB_REQUIRE(corr_levels[l], FSB_ERR_ARG, "fsb_corr_pyramid: level %d pointer is NULL", l);
    FSB_REQUIRE(aligned16(corr_levels[l]), FSB_ERR_ALIGN, "fsb_corr_pyramid: level %d not 16-byte aligned", l);
    lv.p[l] = corr_levels[l];
  }
  if (precision != FSB_PREC_FP32 && corr_tma_eligible(f1, f2, W1, W2))
    return launch_corr_tma(f1, f2, lv, B, C, H, W1, W2, L, precision == FSB_PREC_TF32X3 ? 3 : 1, (cudaStream_t)stream);
  if (precision != FSB_PREC_FP32)
    return launch_corr_umma(f1, f2, lv, B, C, H, W1, W2, L, precision == FSB_PREC_TF32X3 ? 3 : 1, (cudaStream_t)stream);
  dim3 grid(ceil_div(W2, 128), ceil_div(W1, 64), B * H);
  launch_kernel(corr_fp32_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, f1, f2, lv, C, H, W1, W2, L);
  FSB_CHECK_LAUNCH("fsb_corr_pyramid");
  return FSB_OK;
}

namespace fsb {
template <typename TIN>
static int launch_geo_pyramid(const TIN* vol, const LevelPtrs& lv, int B, int Cg, int D, int H, int W, int L,
                              cudaStream_t stream) {
  const size_t smem = (size_t)kGeoTD * Cg * (kGeoTW + 1) * sizeof(float);
  FSB_REQUIRE(smem <= 227 * 1024, FSB_ERR_SHAPE, "fsb_geo_pyramid: Cg=%d needs %zu B shared memory", Cg, smem);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(geo_pyramid_kernel<TIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "fsb_geo_pyramid: cudaFuncSetAttribute");
  }
  dim3 grid(ceil_div(W, kGeoTW), ceil_div(D, kGeoTD), B * H);
  if (Cg == 28)
    launch_kernel(geo_pyramid_fast_kernel<28, TIN>, dim3(grid), dim3(256), 0, stream, vol, lv, D, H, W, L);
  else
    launch_kernel(geo_pyramid_kernel<TIN>, dim3(grid), dim3(256), smem, stream, vol, lv, Cg, D, H, W, L);
  FSB_CHECK_LAUNCH("fsb_geo_pyramid");
  return FSB_OK;
}
}  // namespace fsb

extern "C" int fsb_geo_pyramid_in(const void* vol, int vol_dtype, float* const* geo_levels, int B, int Cg, int D, int H,
                                  int W, int L, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_geo_pyramid_in");
  using namespace fsb;
  FSB_REQUIRE(vol && geo_levels, FSB_ERR_ARG, "fsb_geo_pyramid: NULL pointer");
  FSB_REQUIRE(vol_dtype == FSB_F32 || vol_dtype == FSB_F16 || vol_dtype == FSB_BF16, FSB_ERR_ARG,
              "fsb_geo_pyramid: vol_dtype %d", vol_dtype);
  FSB_REQUIRE(B > 0 && Cg > 0 && D > 0 && H > 0 && W > 0, FSB_ERR_ARG, "fsb_geo_pyramid: non-positive size");
  FSB_REQUIRE(L >= 1 && L <= 4, FSB_ERR_SHAPE, "fsb_geo_pyramid: num_levels=%d outside [1,4]", L);
  FSB_REQUIRE((D >> (L - 1)) >= 1, FSB_ERR_SHAPE, "fsb_geo_pyramid: D=%d too small for %d levels", D, L);
  FSB_REQUIRE((long long)B * H < 65536 && ceil_div(D, kGeoTD) < 65536, FSB_ERR_SHAPE, "fsb_geo_pyramid: grid too large");
  LevelPtrs lv{};
  for (int l = 0; l < L; ++l) {
    FSB_REQUIRE(geo_levels[l], FSB_ERR_ARG, "fsb_geo_pyramid: level %d pointer is NULL", l);
    lv.p[l] = geo_levels[l];
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (vol_dtype == FSB_F16) return launch_geo_pyramid((const __half*)vol, lv, B, Cg, D, H, W, L, st);
  if (vol_dtype == FSB_BF16) return launch_geo_pyramid((const __nv_bfloat16*)vol, lv, B, Cg, D, H, W, L, st);
  return launch_geo_pyramid((const float*)vol, lv, B, Cg, D, H, W, L, st);
}

extern "C" int fsb_geo_pyramid(const float* vol, float* const* geo_levels, int B, int Cg, int D, int H, int W,
                               int L, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_geo_pyramid");
  return fsb_geo_pyramid_in(vol, FSB_F32, geo_levels, B, Cg, D, H, W, L, stream);
}

static int geo_lookup_impl(float* const* geo_levels, float* const* corr_levels, const float* disp,
                           const float* coords, void* out, int B, int Cg, int D, int H, int W, int W2, int L,
                           int r, int out_layout, fsb_stream_t stream, fsb::CorrBand band) {
  using namespace fsb;
  FSB_REQUIRE(geo_levels && corr_levels && disp && coords && out, FSB_ERR_ARG, "fsb_geo_lookup: NULL pointer");
  FSB_REQUIRE(B > 0 && Cg > 0 && D > 0 && H > 0 && W > 0 && W2 > 0, FSB_ERR_ARG, "fsb_geo_lookup: non-positive size");
  FSB_REQUIRE(L >= 1 && L <= 4, FSB_ERR_SHAPE, "fsb_geo_lookup: num_levels=%d outside [1,4]", L);
  FSB_REQUIRE(r >= 1 && 2 * r + 1 <= kLkKMax, FSB_ERR_SHAPE, "fsb_geo_lookup: radius=%d outside [1,4]", r);
  // a level of length 1 divides by zero in the reference's coordinate normalisation
  FSB_REQUIRE((D >> (L - 1)) >= 2 && (W2 >> (L - 1)) >= 2, FSB_ERR_SHAPE,
              "fsb_geo_lookup: coarsest level shorter than 2 (D=%d, W2=%d, L=%d)", D, W2, L);
  FSB_REQUIRE(out_layout >= 0 && out_layout <= 2, FSB_ERR_ARG, "fsb_geo_lookup: out_layout %d (0 = fp32 NCHW, 1 = fp16 NHWC, 2 = bf16 NHWC)", out_layout);
  LevelPtrs g{}, c{};
  for (int l = 0; l < L; ++l) {
    FSB_REQUIRE(geo_levels[l] && corr_levels[l], FSB_ERR_ARG, "fsb_geo_lookup: level %d pointer is NULL", l);
    g.p[l] = geo_levels[l];
    c.p[l] = corr_levels[l];
  }
  if (out_layout != 0) {
    const int K = 2 * r + 1;
    const int Cpad = round_up(L * K * (Cg + 1), 64);
    FSB_REQUIRE(aligned16(out), FSB_ERR_ALIGN, "fsb_geo_lookup: out not 16-byte aligned");
    const size_t smem = (size_t)kLkWarps * Cpad * 2;
    FSB_REQUIRE(smem <= 227 * 1024, FSB_ERR_SHAPE, "fsb_geo_lookup: Cg=%d r=%d L=%d needs %zu B shared memory", Cg, r, L, smem);
    const long long npix = (long long)B * H * W;
    FSB_REQUIRE(npix / kLkWarps < (1ll << 31) - 1, FSB_ERR_SHAPE, "fsb_geo_lookup: too many pixels");
    const unsigned blocks = (unsigned)((npix + kLkWarps - 1) / kLkWarps);
#define FSB_LAUNCH_LK(T, KT, LT)                                                                                   \
  do {                                                                                                             \
    auto kern = Cg == 28 ? lookup_nhwc_kernel<T, KT, LT, 28> : lookup_nhwc_kernel<T, KT, LT, 0>;                                                                   \
    if (smem > 48 * 1024) {                                                                                        \
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);          \
      if (e != cudaSuccess) return cuda_fail(e, "fsb_geo_lookup: cudaFuncSetAttribute");                           \
    }                                                                                                              \
    launch_kernel(kern, dim3(blocks), dim3(kLkWarps * 32), smem, (cudaStream_t)stream, g, c, disp, coords,         \
                  (T*)out, Cg, D, W2, L, r, Cpad, npix, W, band);                                                  \
  } while (0)
#define FSB_LAUNCH_LK_K(T, LT)                                                                                     \
  do {                                                                                                             \
    switch (r) {                                                                                                   \
      case 1: FSB_LAUNCH_LK(T, 3, LT); break;                                                                      \
      case 2: FSB_LAUNCH_LK(T, 5, LT); break;                                                                      \
      case 3: FSB_LAUNCH_LK(T, 7, LT); break;                                                                      \
      default: FSB_LAUNCH_LK(T, 9, LT); break;                                                                     \
    }                                                                                                              \
  } while (0)
    if (out_layout == 1) {
      if (L == 2) FSB_LAUNCH_LK_K(__half, 2); else FSB_LAUNCH_LK_K(__half, 0);
    } else {
      if (L == 2) FSB_LAUNCH_LK_K(__nv_bfloat16, 2); else FSB_LAUNCH_LK_K(__nv_bfloat16, 0);
    }
#undef FSB_LAUNCH_LK_K
#undef FSB_LAUNCH_LK
    FSB_CHECK_LAUNCH("fsb_geo_lookup");
    return FSB_OK;
  }
  FSB_REQUIRE((long long)B * H < 65536, FSB_ERR_SHAPE, "fsb_geo_lookup: B*H too large");
  const size_t smem = lookup_smem_bytes(Cg, r);
  FSB_REQUIRE(smem <= 227 * 1024, FSB_ERR_SHAPE, "fsb_geo_lookup: Cg=%d r=%d needs %zu B shared memory", Cg, r, smem);
  dim3 grid(ceil_div(W, kLkPix), B * H);
#define FSB_LAUNCH_LKN(KT)                                                                                         \
  do {                                                                                                             \
    auto kern = Cg == 28 ? lookup_nchw_kernel<KT, 28> : lookup_nchw_kernel<KT, 0>;                                 \
    if (smem > 48 * 1024) {                                                                                        \
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);          \
      if (e != cudaSuccess) return cuda_fail(e, "fsb_geo_lookup: cudaFuncSetAttribute");                           \
    }                                                                                                              \
    launch_kernel(kern, dim3(grid), dim3(kLkThreads), smem, (cudaStream_t)stream, g, c, disp, coords, (float*)out, \
                  Cg, D, H, W, W2, L, r, band);                                                                    \
  } while (0)
  switch (r) {
    case 1: FSB_LAUNCH_LKN(3); break;
    case 2: FSB_LAUNCH_LKN(5); break;
    case 3: FSB_LAUNCH_LKN(7); break;
    default: FSB_LAUNCH_LKN(9); break;
  }
#undef FSB_LAUNCH_LKN
  FSB_CHECK_LAUNCH("fsb_geo_lookup");
  return FSB_OK;
}

extern "C" int fsb_geo_lookup(float* const* geo_levels, float* const* corr_levels, const float* disp,
                              const float* coords, void* out, int B, int Cg, int D, int H, int W, int W2, int L,
                              int r, int out_layout, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_geo_lookup");
  return geo_lookup_impl(geo_levels, corr_levels, disp, coords, out, B, Cg, D, H, W, W2, L, r, out_layout, stream,
                         fsb::CorrBand{0, 0, 0});
}

// ---- banded init correlation (opt-in) ---------------------------------------------------------------------------
extern "C" int fsb_corr_band_geometry(int d_band, int radius, int L, int* bw0, int* off, int* align) {
  using namespace fsb;
  FSB_REQUIRE(d_band >= 1 && radius >= 1 && L >= 1 && L <= 4, FSB_ERR_ARG, "fsb_corr_band_geometry: d_band=%d radius=%d L=%d",
              d_band, radius, L);
  const CorrBand bd = make_corr_band(d_band, radius, L);
  if (bw0) *bw0 = bd.bw0;
  if (off) *off = bd.off;
  if (align) *align = bd.align;
  return FSB_OK;
}

extern "C" int fsb_corr_pyramid_band(const float* f1, const float* f2, float* const* corr_levels, int B, int C, int H,
                                     int W, int L, int precision, int d_band, int radius, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_corr_pyramid_band");
  using namespace fsb;
  FSB_REQUIRE(f1 && f2 && corr_levels, FSB_ERR_ARG, "fsb_corr_pyramid_band: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, FSB_ERR_ARG, "fsb_corr_pyramid_band: non-positive size");
  FSB_REQUIRE(L >= 1 && L <= 4 && d_band >= 1 && radius >= 1, FSB_ERR_ARG, "fsb_corr_pyramid_band: L=%d d_band=%d radius=%d", L,
              d_band, radius);
  FSB_REQUIRE(precision == FSB_PREC_TF32 || precision == FSB_PREC_TF32X3, FSB_ERR_ARG,
              "fsb_corr_pyramid_band: precision %d (the band form exists on the tensor-core path: FSB_PREC_TF32X3 or FSB_PREC_TF32)",
              precision);
  FSB_REQUIRE((long long)B * H < 65536, FSB_ERR_SHAPE, "fsb_corr_pyramid_band: B*H=%lld >= 65536", (long long)B * H);
  FSB_REQUIRE((W & 3) == 0 && aligned16(f1) && aligned16(f2), FSB_ERR_SHAPE,
              "fsb_corr_pyramid_band: W=%d must be a multiple of 4 and the features 16-byte aligned (TMA-fed kernel)", W);
  LevelPtrs lv{};
  for (int l = 0; l < L; ++l) {
    FSB_REQUIRE(corr_levels[l], FSB_ERR_ARG, "fsb_corr_pyramid_band: level %d pointer is NULL", l);
    FSB_REQUIRE(aligned16(corr_levels[l]), FSB_ERR_ALIGN, "fsb_corr_pyramid_band: level %d not 16-byte aligned", l);
    lv.p[l] = corr_levels[l];
  }
  return launch_corr_tma(f1, f2, lv, B, C, H, W, W, L, precision == FSB_PREC_TF32X3 ? 3 : 1, (cudaStream_t)stream,
                         make_corr_band(d_band, radius, L));
}

extern "C" int fsb_geo_lookup_band(float* const* geo_levels, float* const* corr_levels, const float* disp,
                                   const float* coords, void* out, int B, int Cg, int D, int H, int W, int L, int r,
                                   int out_layout, int d_band, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_geo_lookup_band");
  FSB_REQUIRE(d_band >= 1, FSB_ERR_ARG, "fsb_geo_lookup_band: d_band=%d", d_band);
  FSB_REQUIRE(r >= 1 && L >= 1 && L <= 4, FSB_ERR_ARG, "fsb_geo_lookup_band: radius=%d L=%d", r, L);
  return geo_lookup_impl(geo_levels, corr_levels, disp, coords, out, B, Cg, D, H, W, W, L, r, out_layout, stream,
                         fsb::make_corr_band(d_band, r, L));
}
