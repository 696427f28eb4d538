// The gather of kernel 3 (core/geometry.py:43-65, bilinear_sampler core/utils/utils.py:44-55) as device functions shared by
// the lookup kernels (geo.cu) and the lookup fused into encoder.convc1 (lookconv.cu).
#pragma once
#include "common.cuh"

namespace fsb {

__device__ __forceinline__ float ref_unnormalised(float x, int n) {
  // bilinear_sampler: xgrid = 2*x/(n-1) - 1 ; grid_sample(align_corners=True): ((g+1)/2)*(n-1)
  const float nm1 = (float)(n - 1);
  const float g = __fsub_rn(__fdiv_rn(__fmul_rn(2.f, x), nm1), 1.f);
  return __fmul_rn(__fmul_rn(__fadd_rn(g, 1.f), 0.5f), nm1);
}

// Phases A + B for one level of one pixel (whole warp, see lookup_nhwc_kernel): returns this lane's
// geo window v[0..K] (channel c of rows gbase .. gbase+K, zero outside [0, Dl)), this lane's correlation
// tap cv (lane j <= K holds tap cbase + j), and this lane's per-tap fraction (lane k: geo tap k, lane 16+k:
// corr tap k).  Window bases are warp-uniform, so the "whole window inside" fast path is a uniform branch.
// `crow` / `cshift`: row length and first stored column of the correlation level (Wl and 0 for full rows; bw0 >> l and
// s_l for a banded pyramid, where taps outside the stored band read as zero).
template <int K>
__device__ __forceinline__ void lookup_level(const float* __restrict__ gl, const float* __restrict__ cl, long long pix,
                                             int Cg, int Dl, int Wl, float d, float xcor, int r, int lane, int c,
                                             bool geo_ch, float (&v)[K + 1], float& cv, float& frac, int crow, int cshift) {
  const bool is_cor_lane = lane >= 16;
  const int kk = lane & 15;
  const float xc = is_cor_lane ? xcor : d;
  const int n = is_cor_lane ? Wl : Dl;
  const float ix = ref_unnormalised(__fadd_rn(xc, (float)(kk - r)), n);
  const float ix0 = __shfl_sync(0xffffffffu, ix, lane & 16);  // tap 0 of this lane's half
  const int base = (int)fminf(fmaxf(floorf(ix0), -1.0e6f), 1.0e6f);
  frac = ix - (float)(base + kk);
  const int gbase = __shfl_sync(0xffffffffu, base, 0), cbase = __shfl_sync(0xffffffffu, base, 16);
  const float* g = gl + ((size_t)pix * Dl + gbase) * Cg + c;  // dereferenced only where the row exists
  if (gbase >= 0 && gbase + K < Dl) {
    if (geo_ch) {
#pragma unroll
      for (int j = 0; j <= K; ++j) v[j] = __ldg(g + j * Cg);
    }
  } else {
#pragma unroll
    for (int j = 0; j <= K; ++j) v[j] = (geo_ch && (unsigned)(gbase + j) < (unsigned)Dl) ? __ldg(g + j * Cg) : 0.f;
  }
  const int x = cbase + lane, xb = x - cshift;
  cv = (lane <= K && (unsigned)x < (unsigned)Wl && (unsigned)xb < (unsigned)crow) ? __ldg(cl + (size_t)pix * crow + xb) : 0.f;
}

}  // namespace fsb
