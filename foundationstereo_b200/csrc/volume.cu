// Kernel 1: group-wise correlation volume + concat volume in one launch.
//
// Replaces build_gwc_volume / groupwise_correlation (core/submodule.py:388-412) and
// build_concat_volume (core/submodule.py:416-427) of the reference.
//
// HBM-write bound: per pixel the kernel must emit (G + 2*Cp) * D values and read only
// 2*C + 2*Cp, so everything here is about spending as few instructions per 16-byte store as
// possible.  One CTA owns one image row (b,h) and one "unit" of output channels:
//   unit <  G : correlation group g = unit.  A thread loads one column of the group (C/G
//               channels of one pixel, coalesced along W), L2-normalises it in registers
//               (x / max(||x||, 1e-12), fp32, once per row instead of once per disparity as
//               the reference does) and writes it to shared memory; the right row gets a zero
//               apron of round4(D)+4 columns, so the sliding disparity window needs no bounds
//               test.  Outputs are produced as 4(d) x 4(w) register blocks (three 128-bit
//               shared loads per channel feed 16 FMAs); blocks that lie entirely in the
//               reference's w < d zero triangle (32 % of them at W=160, D=104) are never
//               computed, only zero-filled.
//   unit >= G : a chunk of kCatChunk concat channels, left half then right half.  A thread
//               owns one float4 column and walks down the disparities: the left half stores
//               the same register D/nrow times; the right half keeps four copies of the row in
//               shared memory, copy k shifted by k floats, so the window of disparity d is ONE
//               aligned 128-bit shared load from copy (d & 3).
// Every global store is a full 16-byte (fp32) or 8-byte (fp16/bf16) vector along W.
#include "common.cuh"

namespace fsb {

constexpr int kVolThreads = 256;
constexpr int kCatChunk = 3;  // concat channels per CTA (Cp = 12 -> 4 + 4 units)

// number of non-zero 4x4 blocks in block rows [0, dq): row j keeps the blocks w4 >= j
__host__ __device__ inline int vol_row_start(int dq, int nw4) {
  const int m = dq < nw4 ? dq : nw4;
  return m * nw4 - (m * (m - 1)) / 2;
}

template <typename T, bool VEC>
__device__ __forceinline__ void store_quad(T* o, const float4 v, int w, int W) {
  if (VEC) {
    Quad<T>::store(o, v.x, v.y, v.z, v.w);
  } else {
    const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (w + i < W) o[i] = from_f32<T>(vv[i]);
  }
}

template <typename T, bool VEC>
__global__ void __launch_bounds__(kVolThreads, 4)
volume_kernel(const T* __restrict__ fl, const T* __restrict__ fr, const T* __restrict__ pl,
              const T* __restrict__ pr, T* __restrict__ gwc, T* __restrict__ cat, int C, int Cp, int G,
              int H, int W, int D, size_t gwc_bs, size_t cat_bs) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x;
  // unit is the fastest grid index: the FMA-heavy correlation CTAs and the store-only concat CTAs of a row
  // are co-resident on an SM, so the arithmetic of the former hides under the HBM writes of the latter
  const int b = blockIdx.y / H;
  const int h = blockIdx.y - b * H;
  const int unit = blockIdx.x;
  const int Wp = (W + 3) & ~3;
  const int Dpad = ((D + 3) & ~3) + 4;
  const int RS = Dpad + Wp;  // row stride of an apron-padded row
  const int nw4 = Wp >> 2;
  const size_t HW = (size_t)H * W;

  if (unit < G) {
    // ------------------------------------------------------------ correlation group
    if (gwc == nullptr) return;
    const int cpg = C / G;
    float* Ls = smem;             // [cpg][Wp]
    float* Rs = smem + cpg * Wp;  // [cpg][RS]
    // zero apron (and the <= 3 pad columns on the right of both rows)
    for (int j = tid; j < Dpad; j += kVolThreads)
      for (int c = 0; c < cpg; ++c) Rs[c * RS + j] = 0.f;
    if (Wp != W && tid < cpg * (Wp - W)) {
      const int c = tid / (Wp - W), j = W + tid - c * (Wp - W);
      Ls[c * Wp + j] = 0.f;
      Rs[c * RS + Dpad + j] = 0.f;
    }
    // F.normalize: x / max(sqrt(sum x^2), eps), one column per thread (second pass hits L1)
    const size_t row0 = ((size_t)(b * C + unit * cpg) * H + h) * W;
    for (int it = tid; it < 2 * W; it += kVolThreads) {
      const bool right = it >= W;
      const int w = right ? it - W : it;
      const T* src = (right ? fr : fl) + row0 + w;
      float ss = 0.f;
      for (int c = 0; c < cpg; ++c) {
        const float x = to_f32<T>(src[(size_t)c * HW]);
        ss = fmaf(x, x, ss);
      }
      // x / max(||x||, eps) as x * (1 / max(||x||, eps)): one IEEE division per column instead of one
      // per element (<= 1.5 ulp from the reference's quotient, far inside the 1e-4 tolerance)
      const float inv = __fdiv_rn(1.f, fmaxf(sqrtf(ss), 1e-12f));
      float* dst = right ? (Rs + Dpad + w) : (Ls + w);
      const int st = right ? RS : Wp;
      for (int c = 0; c < cpg; ++c) dst[c * st] = to_f32<T>(src[(size_t)c * HW]) * inv;
    }
    __syncthreads();

    const int nd4 = (D + 3) >> 2;
    T* out0 = gwc + (size_t)b * gwc_bs + ((size_t)unit * D) * HW + (size_t)h * W;
    // zero triangle: out[d][w] = 0 for the float4 columns w4 < d/4 (the rest of w < d falls out of the apron)
    {
      const int lane = tid & 31, warp = tid >> 5;
      for (int d = warp; d < D; d += kVolThreads / 32) {
        const int nz = min(d >> 2, nw4);
        for (int w4 = lane; w4 < nz; w4 += 32) {
          T* o = out0 + (size_t)d * HW + 4 * w4;
          if (VEC) {
            Quad<T>::store(o, 0.f, 0.f, 0.f, 0.f);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (4 * w4 + i < W) o[i] = from_f32<T>(0.f);
          }
        }
      }
    }
    const int total = vol_row_start(nd4, nw4);
    for (int it = tid; it < total; it += kVolThreads) {
      // block row dq = the last one that starts at or before item `it`
      int lo = 0, hi = min(nd4, nw4) - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (vol_row_start(mid, nw4) <= it) lo = mid; else hi = mid - 1;
      }
      const int dq = lo;
      const int w = (dq + it - vol_row_start(dq, nw4)) << 2;
      const int d0 = dq << 2;
      float acc[4][4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[j][i] = 0.f;
      const float* lp = Ls + w;
      const float* rp = Rs + Dpad + w - d0 - 4;
#pragma unroll 4
      for (int c = 0; c < cpg; ++c) {
        const float4 l = *reinterpret_cast<const float4*>(lp + c * Wp);
        const float4 ra = *reinterpret_cast<const float4*>(rp + c * RS);
        const float4 rb = *reinterpret_cast<const float4*>(rp + c * RS + 4);
        const float lv[4] = {l.x, l.y, l.z, l.w};
        const float rr[8] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int i = 0; i < 4; ++i) acc[j][i] = fmaf(lv[i], rr[4 + i - j], acc[j][i]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int d = d0 + j;
        if (d < D) {
          T* o = out0 + (size_t)d * HW + w;
          if (VEC) {
            Quad<T>::store(o, acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
          } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              if (w + i < W) o[i] = from_f32<T>(acc[j][i]);
          }
        }
      }
    }
    return;
  }

  // ---------------------------------------------------------------- concat channels
  if (cat == nullptr) return;
  const int nchunk = (Cp + kCatChunk - 1) / kCatChunk;
  const int u = unit - G;
  const bool right = u >= nchunk;
  const int c0 = (right ? u - nchunk : u) * kCatChunk;
  const int nc = min(kCatChunk, Cp - c0);
  const T* src = right ? pr : pl;
  // four copies of each source row, copy k shifted right by k floats: [k][c][RS + 4]
  const int CS = RS + 4;
  for (int i = tid; i < 4 * nc * CS; i += kVolThreads) smem[i] = 0.f;
  __syncthreads();
  for (int c = 0; c < nc; ++c)
    for (int w = tid; w < W; w += kVolThreads) {
      const float v = to_f32<T>(src[((size_t)(b * Cp + c0 + c) * H + h) * W + w]);
#pragma unroll
      for (int k = 0; k < 4; ++k) smem[(k * nc + c) * CS + Dpad + w + k] = v;
    }
  __syncthreads();
  // thread = (disparity phase dsub, float4 column w4); rows of the output are W floats apart by HW
  const int nrow = nw4 < kVolThreads ? kVolThreads / nw4 : 1;
  const int dsub = nw4 < kVolThreads ? tid / nw4 : 0;
  const int wstep = nw4 < kVolThreads ? nw4 : kVolThreads;
  if (dsub >= nrow) return;
  const size_t dstep = (size_t)nrow * HW;  // elements between two disparities of this thread
  for (int w4 = tid - dsub * nw4; w4 < nw4; w4 += wstep) {
    const int w = w4 << 2;
    for (int c = 0; c < nc; ++c) {
      T* o = cat + (size_t)b * cat_bs + ((size_t)((right ? Cp : 0) + c0 + c) * D + dsub) * HW + (size_t)h * W + w;
      if (!right) {
        const float4 v = *reinterpret_cast<const float4*>(smem + c * CS + Dpad + w);
#pragma unroll 4
        for (int d = dsub; d < D; d += nrow, o += dstep) store_quad<T, VEC>(o, v, w, W);
      } else {
        // R[w - d + i], i = 0..3  ==  copy (d & 3) at the aligned offset Dpad + w - (d & ~3).  The thread's
        // disparities d, d + nrow, ... are walked four at a time: disparity d + 4*nrow has the same copy as d
        // and an offset 4*nrow floats lower, so the loop body is LDS.128 + STG.128 + two adds per store.
        const float* sp[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int d = dsub + q * nrow;
          sp[q] = smem + ((d & 3) * nc + c) * CS + Dpad + w - (d & ~3);
        }
        for (int d = dsub; d < D; d += 4 * nrow) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (d + q * nrow < D) store_quad<T, VEC>(o, *reinterpret_cast<const float4*>(sp[q]), w, W);
            sp[q] -= 4 * nrow;
            o += dstep;
          }
        }
      }
    }
  }
}

template <typename T>
static int launch_volume(const void* fl, const void* fr, const void* pl, const void* pr, void* gwc,
                         void* cat, int B, int C, int Cp, int G, int H, int W, int D, size_t gwc_bs,
                         size_t cat_bs, cudaStream_t stream) {
  const int Wp = round_up(W, 4), Dpad = round_up(D, 4) + 4, RS = Dpad + Wp;
  const int cpg = gwc ? C / G : 0;
  const size_t smem_gwc = (size_t)(cpg * Wp + cpg * RS) * sizeof(float);
  const size_t smem_cat = cat ? (size_t)4 * kCatChunk * (RS + 4) * sizeof(float) : 0;
  const size_t smem = smem_gwc > smem_cat ? smem_gwc : smem_cat;
  FSB_REQUIRE(smem <= 227 * 1024, FSB_ERR_SHAPE,
              "fsb_volume_gwc_concat: row of C/G=%d x W=%d with D=%d needs %zu B shared memory (> 227 KB)",
              cpg, W, D, smem);
  const int nchunk = cat ? ceil_div(Cp, kCatChunk) : 0;
  dim3 grid(G + 2 * nchunk, B * H);
  const size_t esz = sizeof(T);
  const bool vec = (W % 4 == 0) && ((reinterpret_cast<uintptr_t>(gwc) % (4 * esz)) == 0) &&
                   ((reinterpret_cast<uintptr_t>(cat) % (4 * esz)) == 0);
  auto kern = vec ? volume_kernel<T, true> : volume_kernel<T, false>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, "fsb_volume_gwc_concat: cudaFuncSetAttribute");
  }
  launch_kernel(kern, dim3(grid), dim3(kVolThreads), smem, stream, (const T*)fl, (const T*)fr, (const T*)pl, (const T*)pr,
                                            (T*)gwc, (T*)cat, C, Cp, G, H, W, D, gwc_bs, cat_bs);
  FSB_CHECK_LAUNCH("fsb_volume_gwc_concat");
  return FSB_OK;
}

}  // namespace fsb

static int volume_entry(const void* fl, const void* fr, const void* pl, const void* pr, void* gwc, void* cat, int B,
                        int C, int Cp, int G, int H, int W, int D, size_t gwc_bs, size_t cat_bs, int dtype,
                        fsb_stream_t stream, const char* who) {
  using namespace fsb;
  FSB_REQUIRE(gwc || cat, FSB_ERR_ARG, "%s: both outputs are NULL", who);
  FSB_REQUIRE(!gwc || (fl && fr), FSB_ERR_ARG, "%s: gwc requested but fl/fr NULL", who);
  FSB_REQUIRE(!cat || (pl && pr), FSB_ERR_ARG, "%s: cat requested but pl/pr NULL", who);
  FSB_REQUIRE(B > 0 && H > 0 && W > 0 && D > 0, FSB_ERR_ARG, "%s: non-positive size B=%d H=%d W=%d D=%d", who, B, H, W, D);
  FSB_REQUIRE(!gwc || (G > 0 && C > 0 && C % G == 0), FSB_ERR_ARG, "%s: C=%d not divisible by G=%d", who, C, G);
  FSB_REQUIRE(!cat || Cp > 0, FSB_ERR_ARG, "%s: Cp=%d", who, Cp);
  FSB_REQUIRE((long long)B * H < 65536 && G + 2 * ceil_div(Cp > 0 ? Cp : 1, kCatChunk) < 65536, FSB_ERR_SHAPE,
              "%s: grid too large", who);
  if (!gwc) G = 0;
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case FSB_F32: return launch_volume<float>(fl, fr, pl, pr, gwc, cat, B, C, Cp, G, H, W, D, gwc_bs, cat_bs, s);
    case FSB_F16: return launch_volume<__half>(fl, fr, pl, pr, gwc, cat, B, C, Cp, G, H, W, D, gwc_bs, cat_bs, s);
    case FSB_BF16: return launch_volume<__nv_bfloat16>(fl, fr, pl, pr, gwc, cat, B, C, Cp, G, H, W, D, gwc_bs, cat_bs, s);
    default: FSB_REQUIRE(false, FSB_ERR_ARG, "%s: unknown dtype %d", who, dtype);
  }
}

extern "C" int fsb_volume_gwc_concat(const void* fl, const void* fr, const void* pl, const void* pr,
                                     void* gwc, void* cat, int B, int C, int Cp, int G, int H, int W,
                                     int D, int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_volume_gwc_concat");
  const size_t hw = (size_t)(H > 0 ? H : 0) * (W > 0 ? W : 0) * (D > 0 ? D : 0);
  return volume_entry(fl, fr, pl, pr, gwc, cat, B, C, Cp, G, H, W, D, (size_t)(G > 0 ? G : 0) * hw,
                      (size_t)2 * (Cp > 0 ? Cp : 0) * hw, dtype, stream, "fsb_volume_gwc_concat");
}

extern "C" int fsb_volume_combined(const void* fl, const void* fr, const void* pl, const void* pr, void* out, int B,
                                   int C, int Cp, int G, int H, int W, int D, int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_volume_combined");
  using namespace fsb;
  FSB_REQUIRE(fl && fr && pl && pr && out, FSB_ERR_ARG, "fsb_volume_combined: NULL pointer");
  FSB_REQUIRE(B > 0 && H > 0 && W > 0 && D > 0 && G > 0 && Cp > 0, FSB_ERR_ARG, "fsb_volume_combined: non-positive size");
  const size_t hw = (size_t)H * W * D;
  const size_t bs = (size_t)(G + 2 * Cp) * hw;
  const size_t esz = dtype == FSB_F32 ? 4 : 2;
  void* cat = static_cast<char*>(out) + (size_t)G * hw * esz;
  return volume_entry(fl, fr, pl, pr, out, cat, B, C, Cp, G, H, W, D, bs, bs, dtype, stream, "fsb_volume_combined");
}
