// Kernel 1s: the cost volume fused with its consumer -- OPT-IN, it replaces part of corr_stem.
//
//   out = corr_stem[0](cat([build_gwc_volume(fl, fr), build_concat_volume(pl, pr)], 1))
//
// i.e. core/foundation_stereo.py:207-213 up to and including the 1x1x1 Conv3d(32 -> 28) of corr_stem (:164-165), in ONE
// launch: the (B, 8 + 24, D, H, W) volume, its torch.cat copy and the convolution's read of it never exist.  Per voxel
//   out[o, d, w] = PL[o, w] + PR[o, w - d] + sum_g Wg[o, g] * gwc[g, d, w]          (PR and gwc are 0 for w < d)
// with PL = b + Wl . pl and PR = Wr . pr (the left / right halves of the concat volume are constant / shifted along d,
// so their share of the convolution is computed once per pixel, not once per voxel) and gwc the group-wise correlation
// of the L2-normalised features (core/submodule.py:388-412), all in fp32.
//
// Two launches: `stem_normalise_kernel` writes the L2-normalised features (fp32, once per pixel and group) to a
// caller-owned scratch tensor; `volume_stem_kernel`, CTA = one image row (b, h) x a chunk of DC disparities (as many as
// fit shared memory: 24 at W = 160), 512 threads:
//   0. pl / pr rows -> shared; PL [Co][W], PR [Co][apron + W] (a zero apron of D columns stands for w < d);
//   1. per correlation group: the normalised rows of the group are copied into shared memory (128-bit loads) and the
//      DC x W correlations are produced as 4(d) x 4(w) register blocks (as in volume_kernel, volume.cu) into gw [G][DC][W];
//   2. a thread owns one voxel (d, w) -- consecutive lanes = consecutive w, so every shared access is conflict free and
//      every global store coalesced: its G correlations stay in registers while it walks the Co output channels
//      (G FMAs + PL + PR per output).
// (The first form normalised the rows inside every CTA -- 13 times per row at 8 disparities per chunk -- and spent half its
// time there: 2.8 ms at 8 pairs against 2.9 ms for the stock path.)
// HBM: writes Co/(G + 2 Cp) = 28/32 of the volume kernel's bytes and saves the cat copy and the convolution's read.
#include "common.cuh"

namespace fsb {

constexpr int kStemThreads = 512;
constexpr int kStemMaxDC = 32;  // disparities per CTA: the largest multiple of 4 whose working set fits shared memory
constexpr int kStemMaxCpg = 32; // channels per group held in registers while a column is normalised
constexpr int kStemMaxCo = 32;
constexpr int kStemMaxG = 8;

// F.normalize (core/submodule.py:394-395): x / max(||x||_2, 1e-12) over the channels of a group, fp32.
// thread = (side, b, g, h, w); the column stays in registers between the norm and the scaling.
template <typename T>
__global__ void __launch_bounds__(256)
stem_normalise_kernel(const T* __restrict__ fl, const T* __restrict__ fr, float* __restrict__ nl, float* __restrict__ nr,
                      int C, int G, int H, int W, long long ncol) {
  pdl_launch_dependents();
  pdl_wait();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2 * ncol) return;
  const bool right = i >= ncol;
  const long long j = right ? i - ncol : i;          // ((b * G + g) * H + h) * W + w
  const size_t HW = (size_t)H * W;
  const long long bg = j / (long long)HW;
  const size_t hw = (size_t)(j - bg * (long long)HW);
  const int cpg = C / G;
  const size_t off = (size_t)bg * cpg * HW + hw;      // b * C + g * cpg = bg * cpg
  const T* src = (right ? fr : fl) + off;
  float* dst = (right ? nr : nl) + off;
  if (cpg <= kStemMaxCpg) {
    float x[kStemMaxCpg];
    float ss = 0.f;
#pragma unroll
    for (int c = 0; c < kStemMaxCpg; ++c) {
      x[c] = c < cpg ? to_f32<T>(src[(size_t)c * HW]) : 0.f;
      ss = fmaf(x[c], x[c], ss);
    }
    const float inv = __fdiv_rn(1.f, fmaxf(sqrtf(ss), 1e-12f));
#pragma unroll
    for (int c = 0; c < kStemMaxCpg; ++c)
      if (c < cpg) dst[(size_t)c * HW] = x[c] * inv;
  } else {
    float ss = 0.f;
    for (int c = 0; c < cpg; ++c) {
      const float x = to_f32<T>(src[(size_t)c * HW]);
      ss = fmaf(x, x, ss);
    }
    const float inv = __fdiv_rn(1.f, fmaxf(sqrtf(ss), 1e-12f));
    for (int c = 0; c < cpg; ++c) dst[(size_t)c * HW] = to_f32<T>(src[(size_t)c * HW]) * inv;
  }
}

// the correlation columns of the convolution's weight as a kernel parameter: with the output-channel loop unrolled every
// FFMA of phase 2 takes its weight straight from the constant bank (as shared-memory reads they were 8 LDS per output)
struct StemGroupWeights {
  float w[kStemMaxCo][kStemMaxG];
};

template <typename T>
__global__ void __launch_bounds__(kStemThreads, 1)
volume_stem_kernel(const float* __restrict__ nl, const float* __restrict__ nr, const T* __restrict__ pl, const T* __restrict__ pr,
                   const float* __restrict__ wst, const float* __restrict__ bst, const __grid_constant__ StemGroupWeights gwt,
                   T* __restrict__ out, int C, int Cp, int G, int Co, int H, int W, int D, int DC, int stage_floats) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x;
  const int b = blockIdx.y / H, h = blockIdx.y - b * H;
  const int d0 = blockIdx.x * DC;
  const int Wp = (W + 3) & ~3;
  const int Dpad = ((D + 3) & ~3) + 4;
  const int RS = Dpad + Wp;
  const int nw4 = Wp >> 2;
  const int cpg = C / G;
  const int K = G + 2 * Cp;   // input channels of the stem convolution
  const size_t HW = (size_t)H * W;
  // shared: PL [Co][Wp] | PR [Co][RS] | gw [G][DC][Wp] | stage: Ls [cpg][Wp], Rs [cpg][RS] | wsm [Co][K] + bias [Co]
  float* PL = smem;
  float* PR = PL + Co * Wp;
  float* gw = PR + Co * RS;
  float* Ls = gw + G * DC * Wp;
  float* Rs = Ls + cpg * Wp;
  float* wsm = Ls + stage_floats;   // stage = max(cpg (Wp + RS), 2 Cp Wp) floats
  float* bsm = wsm + Co * K;

  for (int i = tid; i < Co * K; i += kStemThreads) wsm[i] = __ldg(wst + i);
  for (int i = tid; i < Co; i += kStemThreads) bsm[i] = __ldg(bst + i);
  // ---- phase 0: pl / pr rows into the (still free) Ls / Rs region, then PL / PR
  {
    float* pls = Ls;             // [Cp][Wp]
    float* prs = Ls + Cp * Wp;   // [Cp][Wp]
    for (int i = tid; i < 2 * Cp * Wp; i += kStemThreads) {
      const int side = i / (Cp * Wp), rem = i - side * Cp * Wp, c = rem / Wp, w = rem - c * Wp;
      const T* src = side ? pr : pl;
      pls[i] = w < W ? to_f32<T>(src[((size_t)(b * Cp + c) * H + h) * W + w]) : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < Co * Wp; i += kStemThreads) {
      const int o = i / Wp, w = i - o * Wp;
      float al = bsm[o], ar = 0.f;
      const float* wl = wsm + o * K + G;
      for (int c = 0; c < Cp; ++c) {
        al = fmaf(wl[c], pls[c * Wp + w], al);
        ar = fmaf(wl[Cp + c], prs[c * Wp + w], ar);
      }
      PL[i] = al;
      PR[o * RS + Dpad + w] = w < W ? ar : 0.f;
    }
    for (int i = tid; i < Co * Dpad; i += kStemThreads) {
      const int o = i / Dpad;
      PR[o * RS + (i - o * Dpad)] = 0.f;   // w < d: the right half of the concat volume is zero
    }
    __syncthreads();
  }
  // ---- phase 1: group-wise correlations of the chunk, one group at a time
  for (int g = 0; g < G; ++g) {
    for (int j = tid; j < Dpad * cpg; j += kStemThreads) {
      const int c = j / Dpad;
      Rs[c * RS + (j - c * Dpad)] = 0.f;
    }
    if (Wp != W)
      for (int j = tid; j < cpg * (Wp - W); j += kStemThreads) {
        const int c = j / (Wp - W), w = W + j - c * (Wp - W);
        Ls[c * Wp + w] = 0.f;
        Rs[c * RS + Dpad + w] = 0.f;
      }
    // the group's normalised rows: (b, g cpg + c, h, :) of the scratch tensors
    const size_t row0 = ((size_t)(b * C + g * cpg) * H + h) * W;
    if ((W & 3) == 0) {
      const int n4 = W >> 2;
      for (int it = tid; it < 2 * cpg * n4; it += kStemThreads) {
        const bool right = it >= cpg * n4;
        const int r2 = right ? it - cpg * n4 : it, c = r2 / n4, w4 = r2 - c * n4;
        const float4 v = __ldg(reinterpret_cast<const float4*>((right ? nr : nl) + row0 + (size_t)c * HW) + w4);
        *reinterpret_cast<float4*>(right ? (Rs + c * RS + Dpad + 4 * w4) : (Ls + c * Wp + 4 * w4)) = v;
      }
    } else {
      for (int it = tid; it < 2 * cpg * W; it += kStemThreads) {
        const bool right = it >= cpg * W;
        const int r2 = right ? it - cpg * W : it, c = r2 / W, w = r2 - c * W;
        const float v = __ldg((right ? nr : nl) + row0 + (size_t)c * HW + w);
        *(right ? (Rs + c * RS + Dpad + w) : (Ls + c * Wp + w)) = v;
      }
    }
    __syncthreads();
    // 4(d) x 4(w) register blocks: (DC / 4) x nw4 of them
    for (int it = tid; it < (DC / 4) * nw4; it += kStemThreads) {
      const int dq = it / nw4, w = (it - dq * nw4) << 2;
      const int dd = d0 + 4 * dq;   // first disparity of the block (blocks past D in the last chunk are not computed)
      float acc[4][4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[j][i] = 0.f;
      if (dd < D) {
        const float* lp = Ls + w;
        const float* rp = Rs + Dpad + w - dd - 4;   // >= Rs: dd <= round4(D) - 4, Dpad = round4(D) + 4
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
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<float4*>(gw + ((size_t)g * DC + 4 * dq + j) * Wp + w) =
            make_float4(acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
    }
    __syncthreads();
  }
  // ---- phase 2: the 1x1x1 convolution, one voxel per thread and step
  for (int it = tid; it < DC * W; it += kStemThreads) {
    const int dl = it / W, w = it - dl * W;
    const int d = d0 + dl;
    if (d >= D) break;
    float cg[kStemMaxG];
#pragma unroll
    for (int g = 0; g < kStemMaxG; ++g) cg[g] = g < G ? gw[((size_t)g * DC + dl) * Wp + w] : 0.f;
    T* op = out + (((size_t)b * Co * D + d) * H + h) * W + w;
    const float* plp = PL + w;
    const float* prp = PR + Dpad + w - d;
    const size_t ostride = (size_t)D * HW;
#pragma unroll
    for (int o = 0; o < kStemMaxCo; ++o) {
      if (o < Co) {
        float a = *plp + *prp;
#pragma unroll
        for (int g = 0; g < kStemMaxG; ++g) a = fmaf(gwt.w[o][g], cg[g], a);   // rows / columns past Co / G are zero
        *op = from_f32<T>(a);
        plp += Wp;
        prp += RS;
        op += ostride;
      }
    }
  }
}

}  // namespace fsb

extern "C" int fsb_volume_stem(const void* fl, const void* fr, const void* pl, const void* pr, const float* weight,
                               const float* bias, const float* wg_host, float* scratch, void* out, int B, int C, int Cp, int G, int Co, int H, int W,
                               int D, int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_volume_stem");
  using namespace fsb;
  FSB_REQUIRE(fl && fr && pl && pr && weight && bias && wg_host && scratch && out, FSB_ERR_ARG, "fsb_volume_stem: NULL pointer");
  FSB_REQUIRE(B > 0 && C > 0 && Cp > 0 && G > 0 && Co > 0 && H > 0 && W > 0 && D > 0, FSB_ERR_ARG, "fsb_volume_stem: non-positive size");
  FSB_REQUIRE(dtype == FSB_F32 || dtype == FSB_F16 || dtype == FSB_BF16, FSB_ERR_ARG, "fsb_volume_stem: dtype %d", dtype);
  FSB_REQUIRE(C % G == 0 && G <= kStemMaxG && Co <= kStemMaxCo, FSB_ERR_SHAPE, "fsb_volume_stem: C=%d G=%d Co=%d (C %% G == 0, G <= %d, Co <= %d)",
              C, G, Co, kStemMaxG, kStemMaxCo);
  FSB_REQUIRE((long long)B * H < 65536, FSB_ERR_SHAPE, "fsb_volume_stem: B*H=%lld >= 65536", (long long)B * H);
  FSB_REQUIRE(aligned16(scratch), FSB_ERR_ALIGN, "fsb_volume_stem: scratch must be 16-byte aligned");
  const int Wp = (W + 3) & ~3, Dpad = ((D + 3) & ~3) + 4, RS = Dpad + Wp, cpg = C / G, K = G + 2 * Cp;
  const int stage = cpg * (Wp + RS) > 2 * Cp * Wp ? cpg * (Wp + RS) : 2 * Cp * Wp;   // floats
  const size_t fixed = sizeof(float) * ((size_t)Co * Wp + (size_t)Co * RS + (size_t)stage + (size_t)Co * K + Co);
  int DC = 0;   // disparities per CTA: as many as fit (the group rows are re-staged per chunk)
  for (int dc = kStemMaxDC; dc >= 4; dc -= 4)
    if (fixed + sizeof(float) * (size_t)G * dc * Wp <= 227 * 1024 - 64) { DC = dc; break; }
  FSB_REQUIRE(DC > 0, FSB_ERR_SHAPE, "fsb_volume_stem: W=%d D=%d C/G=%d: one row's working set (%zu B + the correlations) exceeds shared memory",
              W, D, cpg, fixed);
  if (DC > ((D + 3) & ~3)) DC = (D + 3) & ~3;
  const size_t smem = fixed + sizeof(float) * (size_t)G * DC * Wp;
  // the (Co, G) correlation block of the weight travels to the kernel by value (1 KB of launch parameters)
  StemGroupWeights gwt;
  memset(&gwt, 0, sizeof(gwt));
  for (int o = 0; o < Co; ++o)
    for (int g = 0; g < G; ++g) gwt.w[o][g] = wg_host[o * G + g];
  float* nl = scratch;
  float* nr = scratch + (size_t)B * C * H * W;
  const long long ncol = (long long)B * G * H * W;
  FSB_REQUIRE((2 * ncol + 255) / 256 < (1ll << 31), FSB_ERR_SHAPE, "fsb_volume_stem: too many pixels");
  const unsigned nblk = (unsigned)((2 * ncol + 255) / 256);
  dim3 grid(ceil_div(D, DC), B * H);
#define FSB_LAUNCH_STEM(T)                                                                                          \
  do {                                                                                                             \
    launch_kernel(stem_normalise_kernel<T>, dim3(nblk), dim3(256), 0, (cudaStream_t)stream, (const T*)fl, (const T*)fr, \
                  nl, nr, C, G, H, W, ncol);                                                                       \
    FSB_CHECK_LAUNCH("fsb_volume_stem (normalise)");                                                               \
    auto kern = volume_stem_kernel<T>;                                                                             \
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);            \
    if (e != cudaSuccess) return cuda_fail(e, "fsb_volume_stem: cudaFuncSetAttribute");                            \
    launch_kernel(kern, grid, dim3(kStemThreads), smem, (cudaStream_t)stream, (const float*)nl, (const float*)nr,  \
                  (const T*)pl, (const T*)pr, weight, bias, gwt, (T*)out, C, Cp, G, Co, H, W, D, DC, stage);            \
  } while (0)
  if (dtype == FSB_F32) FSB_LAUNCH_STEM(float);
  else if (dtype == FSB_F16) FSB_LAUNCH_STEM(__half);
  else FSB_LAUNCH_STEM(__nv_bfloat16);
#undef FSB_LAUNCH_STEM
  FSB_CHECK_LAUNCH("fsb_volume_stem");
  return FSB_OK;
}
