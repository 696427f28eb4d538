// Kernel 1: group-wise correlation volume + concat volume in one launch.
//
// Replaces build_gwc_volume / groupwise_correlation (core/submodule.py:388-412) and
// build_concat_volume (core/submodule.py:416-427) of the reference.
//
// HBM-write bound: per pixel the kernel must emit (G + 2*Cp) * D values and read only
// 2*C + 2*Cp.  One CTA owns one image row (b,h) and one "unit" of output channels:
//   unit <  G            : correlation group g = unit        (D*W outputs, cpg FMAs each)
//   unit >= G            : a chunk of kCatChunk concat channels, left half then right half
// The source rows of the unit are staged ONCE in shared memory as fp32 with a zero
// apron of round4(D)+4 columns on the left, so the sliding disparity window needs no
// bounds test and the w < d zero region of the reference falls out of the arithmetic.
// The per-group L2 normalisation (x / max(||x||, 1e-12), fp32) is done once per row in
// shared memory instead of once per disparity as the reference does.
// Each thread produces a 4(d) x 4(w) register block per step: three 128-bit shared
// loads per channel feed 16 FMAs, and every global store is a full 16-byte (fp32) or
// 8-byte (fp16/bf16) vector along W.
#include "common.cuh"

namespace fsb {

constexpr int kVolThreads = 256;
constexpr int kCatChunk = 3;  // concat channels per CTA (Cp = 12 -> 4 + 4 units)

template <typename T, bool VEC>
__global__ void __launch_bounds__(kVolThreads)
volume_kernel(const T* __restrict__ fl, const T* __restrict__ fr, const T* __restrict__ pl,
              const T* __restrict__ pr, T* __restrict__ gwc, T* __restrict__ cat, int C, int Cp, int G,
              int H, int W, int D, size_t gwc_bs, size_t cat_bs) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x;
  const int b = blockIdx.x / H;
  const int h = blockIdx.x - b * H;
  const int unit = blockIdx.y;
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
    for (int i = tid; i < cpg * Wp + cpg * RS; i += kVolThreads) smem[i] = 0.f;
    __syncthreads();
    const size_t row0 = ((size_t)(b * C + unit * cpg) * H + h) * W;
    for (int i = tid; i < cpg * W; i += kVolThreads) {
      const int c = i / W, w = i - c * W;
      const size_t src = row0 + (size_t)c * HW + w;
      Ls[c * Wp + w] = to_f32<T>(fl[src]);
      Rs[c * RS + Dpad + w] = to_f32<T>(fr[src]);
    }
    __syncthreads();
    // F.normalize: x / max(sqrt(sum x^2), eps)
    for (int w = tid; w < 2 * W; w += kVolThreads) {
      float* col = (w < W) ? (Ls + w) : (Rs + Dpad + (w - W));
      const int st = (w < W) ? Wp : RS;
      float ss = 0.f;
      for (int c = 0; c < cpg; ++c) ss = fmaf(col[c * st], col[c * st], ss);
      const float den = fmaxf(sqrtf(ss), 1e-12f);
      for (int c = 0; c < cpg; ++c) col[c * st] = __fdiv_rn(col[c * st], den);
    }
    __syncthreads();

    const int nd4 = (D + 3) >> 2;
    T* out0 = gwc + (size_t)b * gwc_bs + ((size_t)unit * D) * HW + (size_t)h * W;
    for (int it = tid; it < nd4 * nw4; it += kVolThreads) {
      const int dq = it / nw4;
      const int w = (it - dq * nw4) << 2;
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
  for (int i = tid; i < nc * RS; i += kVolThreads) smem[i] = 0.f;
  __syncthreads();
  for (int i = tid; i < nc * W; i += kVolThreads) {
    const int c = i / W, w = i - c * W;
    smem[c * RS + Dpad + w] = to_f32<T>(src[((size_t)(b * Cp + c0 + c) * H + h) * W + w]);
  }
  __syncthreads();
  const int per_c = D * nw4;
  for (int it = tid; it < nc * per_c; it += kVolThreads) {
    const int c = it / per_c;
    const int rem = it - c * per_c;
    const int d = rem / nw4;
    const int w = (rem - d * nw4) << 2;
    const float* s = smem + c * RS + Dpad + w - (right ? d : 0);
    const float v0 = s[0], v1 = s[1], v2 = s[2], v3 = s[3];
    T* o = cat + (size_t)b * cat_bs + ((size_t)((right ? Cp : 0) + c0 + c) * D + d) * HW + (size_t)h * W + w;
    if (VEC) {
      Quad<T>::store(o, v0, v1, v2, v3);
    } else {
      const float v[4] = {v0, v1, v2, v3};
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (w + i < W) o[i] = from_f32<T>(v[i]);
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
  const size_t smem_cat = cat ? (size_t)kCatChunk * RS * sizeof(float) : 0;
  const size_t smem = smem_gwc > smem_cat ? smem_gwc : smem_cat;
  FSB_REQUIRE(smem <= 227 * 1024, FSB_ERR_SHAPE,
              "fsb_volume_gwc_concat: row of C/G=%d x W=%d with D=%d needs %zu B shared memory (> 227 KB)",
              cpg, W, D, smem);
  const int nchunk = cat ? ceil_div(Cp, kCatChunk) : 0;
  dim3 grid(B * H, G + 2 * nchunk);
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
  FSB_REQUIRE((long long)B * H < (1ll << 31) && G + 2 * ceil_div(Cp > 0 ? Cp : 1, kCatChunk) < 65536, FSB_ERR_SHAPE,
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
  const size_t hw = (size_t)(H > 0 ? H : 0) * (W > 0 ? W : 0) * (D > 0 ? D : 0);
  return volume_entry(fl, fr, pl, pr, gwc, cat, B, C, Cp, G, H, W, D, (size_t)(G > 0 ? G : 0) * hw,
                      (size_t)2 * (Cp > 0 ? Cp : 0) * hw, dtype, stream, "fsb_volume_gwc_concat");
}

extern "C" int fsb_volume_combined(const void* fl, const void* fr, const void* pl, const void* pr, void* out, int B,
                                   int C, int Cp, int G, int H, int W, int D, int dtype, fsb_stream_t stream) {
  using namespace fsb;
  FSB_REQUIRE(fl && fr && pl && pr && out, FSB_ERR_ARG, "fsb_volume_combined: NULL pointer");
  FSB_REQUIRE(B > 0 && H > 0 && W > 0 && D > 0 && G > 0 && Cp > 0, FSB_ERR_ARG, "fsb_volume_combined: non-positive size");
  const size_t hw = (size_t)H * W * D;
  const size_t bs = (size_t)(G + 2 * Cp) * hw;
  const size_t esz = dtype == FSB_F32 ? 4 : 2;
  void* cat = static_cast<char*>(out) + (size_t)G * hw * esz;
  return volume_entry(fl, fr, pl, pr, out, cat, B, C, Cp, G, H, W, D, bs, bs, dtype, stream, "fsb_volume_combined");
}
