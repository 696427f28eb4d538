// Kernel 2, TMA-fed form: the init correlation (core/geometry.py:68-77) with the pyramid (core/geometry.py:31,38-40)
// in the epilogue, for feature maps whose rows are 16-byte aligned (W % 4 == 0).
//
// GEMM view for one image row (b,h):  D[w1, w2] = sum_c f1[c, w1] * f2[c, w2],  M = W1, N = W2, K = C.
// The (C, W) planes of NCHW features are MN-major operands (the pixel index is the contiguous one), and
// tcgen05.mma kind::tf32 reads MN-major tiles from shared memory (instruction-descriptor bits 15 / 16).  So nothing is
// transposed: a TMA box of 32 channels x 32 pixels (4 KB, rows of 128 B) lands in shared memory as it lies in HBM and
// is multiplied in place.  32-bit MN-major operands want the 32-byte-atom form of the 128-byte swizzle on both sides:
// CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B for the TMA, layout type SWIZZLE_128B_BASE32B (1) in the UMMA descriptor, whose
// canonical atom is 32 pixels x 4 channels (512 B); plain SWIZZLE_128B multiplies zeros (measured).  corr_umma.cu's register transpose of the planes into K-major tiles --
// what bounded that kernel -- is gone; the threads only run one elementwise pass per chunk:
//     lo = x - tf32_truncate(x)  into a second tile at the same (swizzled) offset, and  sum of squares per pixel.
// fp32 parity on the TF32 pipe as in corr_umma.cu: hi*hi + hi*lo + lo*hi into one fp32 TMEM accumulator, where "hi" is
// the raw tile (the TF32 datapath ignores the 13 low mantissa bits).  NPROD == 1 is single-pass TF32.
//
// CTA = (m tile of rows_m <= 128 w1, n tile of <= 256 w2, one (b,h)), 256 threads, two CTAs per SM at W = 160.
// K streams in chunks of 32 channels: two raw stages (TMA for chunk ch+1 is issued as soon as the MMAs of chunk ch-1
// have released its stage) and one lo stage.  Epilogue as in corr_umma.cu: thread = row w1, 16 columns per tcgen05.ld,
// normalisation by 1 / (max(|f1|,eps) max(|f2|,eps)), level 0 and every pooled level from the same registers.
#include "umma.cuh"

#include <stdlib.h>
#include <string.h>

namespace fsb {
namespace {

constexpr int kCtThreads = 256;
constexpr int kCtBM = 128;
constexpr int kCtKC = 32;      // channels per chunk = rows of a box
constexpr int kCtBoxW = 32;    // pixels per box row (128 bytes)
constexpr uint32_t kCtBox = kCtKC * kCtBoxW * 4;  // 4096 bytes

struct CorrTmaParams {
  CUtensorMap map1, map2;  // (W, H, C, B) fp32, box (32, 1, 32, 1), SWIZZLE_128B
  LevelPtrs lv;
  int C, H, W1, W2, L;
  int rows_m;   // rows of w1 per m tile
  int boxes_a;  // boxes of the A region of a stage = ceil(rows_m / 32)
  int boxes_b;  // boxes of the B region = ceil(min(W2, 256) padded to 16 / 32)
  int stage0, stage1;  // level 0 / level 1 leave through shared memory as fully coalesced 16-byte stores
  CorrBand band;       // bw0 > 0: banded output (one n window per m tile, rows of bw0 >> l columns)
};

// MN-major SWIZZLE_128B_BASE32B operand: atom = 32 pixels (128 B) x 4 channels (rows of 128 B, 512 B per atom);
// consecutive pixel blocks of 32 are `mn_stride` bytes apart (one box), consecutive 4-channel groups 512 B; one
// kind::tf32 instruction (K = 8) reads two of them.
__device__ __forceinline__ uint64_t ct_desc_mn128(uint32_t addr, uint32_t mn_stride) {
  const uint32_t lbo = mn_stride, sbo = 512u;  // leading = between 32-pixel blocks, stride = between 4-row groups
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)1 << 61;  // SWIZZLE_128B_BASE32B
  return d;
}
__device__ __forceinline__ void ct_mma_tf32(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

template <int NPROD>
__global__ void __launch_bounds__(kCtThreads)
corr_tma_kernel(const __grid_constant__ CorrTmaParams p) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ __align__(1024) uint8_t ct_smem[];
  __shared__ __align__(8) uint64_t bars[4];  // full[2], mma_done, acc
  __shared__ uint32_t tmem_slot;
  __shared__ float nA[kCtBM], nB[256];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int m0 = blockIdx.x * p.rows_m;
  const int bh = blockIdx.z, b = bh / p.H, h = bh - b * p.H;
  const int W1 = p.W1, W2 = p.W2;
  const int rowsA = min(p.rows_m, W1 - m0);
  const bool banded = p.band.bw0 > 0;
  // band mode: ONE n window per m tile = the union of the rows' bands, [s0(m0), s0(m0 + rowsA - 1) + bw0) clipped to the
  // image (the host sized rows_m so that it is at most 256 columns wide)
  const int n0 = banded ? max(0, band_start(p.band, m0)) : blockIdx.y * 256;
  const int Nv = banded ? min(W2, band_start(p.band, m0 + rowsA - 1) + p.band.bw0) - n0 : min(256, W2 - n0);
  const int Npad = (Nv + 15) & ~15;
  const int ldA = (rowsA + kCtBoxW - 1) / kCtBoxW, ldB = (Nv + kCtBoxW - 1) / kCtBoxW;  // boxes to fetch
  const int nboxes = p.boxes_a + p.boxes_b;
  const uint32_t stage_bytes = (uint32_t)nboxes * kCtBox;
  const uint32_t base = (smem_u32(ct_smem) + 1023u) & ~1023u;
  uint8_t* gbase = ct_smem + (base - smem_u32(ct_smem));
  const uint32_t lo_base = base + 2 * stage_bytes;
  const uint32_t off_b = (uint32_t)p.boxes_a * kCtBox;
  const uint32_t bar_full = smem_u32(&bars[0]), bar_mma = smem_u32(&bars[2]), bar_acc = smem_u32(&bars[3]);
  const int nchunks = (p.C + kCtKC - 1) / kCtKC;

  int tcols = 32;
  while (tcols < Npad) tcols <<= 1;
  if (tid == 0) {
    mbar_init(bar_full, 1);
    mbar_init(bar_full + 8, 1);
    mbar_init(bar_mma, 1);
    mbar_init(bar_acc, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // the first two chunks are requested before TMEM allocation and the set-up barrier (thread 0 initialised the barriers
  // it posts to; everyone else meets them after the __syncthreads below)
  auto issue_tma = [&](int ch) {  // thread 0: chunk ch -> raw stage ch & 1
    const uint32_t s = ch & 1, dst = base + s * stage_bytes, bar = bar_full + 8 * s;
    mbar_expect_tx(bar, (uint32_t)(ldA + ldB) * kCtBox);
    for (int i = 0; i < ldA; ++i) tma_load_4d(dst + i * kCtBox, &p.map1, bar, m0 + i * kCtBoxW, h, ch * kCtKC, b);
    for (int i = 0; i < ldB; ++i) tma_load_4d(dst + off_b + i * kCtBox, &p.map2, bar, n0 + i * kCtBoxW, h, ch * kCtKC, b);
  };
  if (tid == 0) {
    issue_tma(0);
    if (nchunks > 1) issue_tma(1);
  }

  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "r"((uint32_t)tcols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  // instruction descriptor, kind::tf32: D fp32, A/B tf32, both MN-major (bits 15, 16), M = 128, N = Npad
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(Npad >> 3) << 17) |
                         ((uint32_t)(kCtBM >> 4) << 24);
  // elementwise pass: thread = 16-byte slot j of row kr of every 8-row atom of the boxes tg, tg+4, ...; its four
  // pixels are the same in every atom (the swizzle depends on kr only), so their sums of squares stay here
  const int j = tid & 7, kr = (tid >> 3) & 7, tg = tid >> 6;
  const uint32_t slot = (uint32_t)kr * 128u + (uint32_t)j * 16u;
  float acc[3][4];
#pragma unroll
  for (int u = 0; u < 3; ++u)
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[u][i] = 0.f;

  for (int ch = 0; ch < nchunks; ++ch) {
    const uint32_t s = ch & 1;
    if (ch >= 1) {
      mbar_wait(bar_mma, (ch - 1) & 1);  // the MMAs of chunk ch-1 are done: lo tile and the other raw stage are free
      if (tid == 0 && ch + 1 < nchunks) issue_tma(ch + 1);
    }
    mbar_wait(bar_full + 8 * s, (ch >> 1) & 1);
    if (NPROD == 3 && tid == 0) {
      // hi * hi needs only the raw tiles: it runs on the tensor pipe while the threads write the lo tiles
      tc_fence_after();
      const uint32_t sa = base + s * stage_bytes;
#pragma unroll
      for (int kg = 0; kg < 4; ++kg)
        ct_mma_tf32(tmem, ct_desc_mn128(sa + kg * 1024, kCtBox), ct_desc_mn128(sa + off_b + kg * 1024, kCtBox), idesc,
                    (ch > 0 || kg > 0) ? 1u : 0u);
    }
    uint8_t* raw = gbase + (size_t)s * stage_bytes;
    uint8_t* lo = gbase + 2 * (size_t)stage_bytes;
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int bx = tg + 4 * u;            // box index over [A boxes fetched][B boxes fetched]
      if (bx < ldA + ldB) {
        const uint32_t boff = (bx < ldA ? (uint32_t)bx * kCtBox : off_b + (uint32_t)(bx - ldA) * kCtBox) + slot;
#pragma unroll
        for (int kg = 0; kg < 4; ++kg) {
          const float4 x = *reinterpret_cast<const float4*>(raw + boff + kg * 1024);
          acc[u][0] = fmaf(x.x, x.x, acc[u][0]);
          acc[u][1] = fmaf(x.y, x.y, acc[u][1]);
          acc[u][2] = fmaf(x.z, x.z, acc[u][2]);
          acc[u][3] = fmaf(x.w, x.w, acc[u][3]);
          if (NPROD == 3) {
            float4 l;
            l.x = x.x - __uint_as_float(__float_as_uint(x.x) & 0xffffe000u);
            l.y = x.y - __uint_as_float(__float_as_uint(x.y) & 0xffffe000u);
            l.z = x.z - __uint_as_float(__float_as_uint(x.z) & 0xffffe000u);
            l.w = x.w - __uint_as_float(__float_as_uint(x.w) & 0xffffe000u);
            *reinterpret_cast<float4*>(lo + boff + kg * 1024) = l;
          }
        }
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> visible to the MMA
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint32_t sa = base + s * stage_bytes;
#pragma unroll
      for (int kg = 0; kg < 4; ++kg) {  // 8 channels (one atom row group, 1024 B) per MMA
        const uint64_t ahi = ct_desc_mn128(sa + kg * 1024, kCtBox);
        const uint64_t bhi = ct_desc_mn128(sa + off_b + kg * 1024, kCtBox);
        if (NPROD == 1) ct_mma_tf32(tmem, ahi, bhi, idesc, (ch > 0 || kg > 0) ? 1u : 0u);
        if (NPROD == 3) {
          const uint64_t alo = ct_desc_mn128(lo_base + kg * 1024, kCtBox);
          const uint64_t blo = ct_desc_mn128(lo_base + off_b + kg * 1024, kCtBox);
          ct_mma_tf32(tmem, ahi, blo, idesc, 1u);
          ct_mma_tf32(tmem, alo, bhi, idesc, 1u);
        }
      }
      umma_commit(bar_mma);
      if (ch == nchunks - 1) umma_commit(bar_acc);
    }
  }
  // ||f||^2: the thread's partial sums -> scratch[kr][pixel] (eight threads per pixel quad, one per atom row kr),
  // summed in a fixed order (bit-reproducible results).  Scratch = the raw stage the last chunk did NOT use: its MMAs
  // were waited for above and nothing was fetched into it since.
  {
    float* scratch = reinterpret_cast<float*>(gbase + (size_t)(nchunks & 1) * stage_bytes);  // [8][kCtBM + 256]
    const int quad = (j ^ ((kr & 3) << 1)) * 4;  // SWIZZLE_128B_ATOM_32B: 32-byte units XORed with the row's low bits
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int bx = tg + 4 * u;
      if (bx < ldA + ldB) {
        const int pix = bx < ldA ? bx * kCtBoxW + quad : kCtBM + (bx - ldA) * kCtBoxW + quad;
        *reinterpret_cast<float4*>(scratch + kr * (kCtBM + 256) + pix) = make_float4(acc[u][0], acc[u][1], acc[u][2], acc[u][3]);
      }
    }
    __syncthreads();
    if (tid < ldA * kCtBoxW) {
      float t = 0.f;
#pragma unroll
      for (int k8 = 0; k8 < 8; ++k8) t += scratch[k8 * (kCtBM + 256) + tid];
      nA[tid] = fmaxf(sqrtf(t), 1e-12f);
    }
    if (tid < ldB * kCtBoxW) {
      float t = 0.f;
#pragma unroll
      for (int k8 = 0; k8 < 8; ++k8) t += scratch[k8 * (kCtBM + 256) + kCtBM + tid];
      nB[tid] = 1.f / fmaxf(sqrtf(t), 1e-12f);
    }
  }
  __syncthreads();

  // output staging (all MMAs are done once bar_acc has fired, the operand stages are free): level 0 over the two raw
  // stages, level 1 in the lo stage
  float* const out0s = reinterpret_cast<float*>(gbase);
  float* const out1s = reinterpret_cast<float*>(gbase + 2 * (size_t)stage_bytes);
  const int stride0s = W2 + 4, stride1s = (W2 >> 1) + 4;
  {
    // epilogue on all 8 warps: warp w reads TMEM lane quarter (w & 3); warps 0-3 take the even 16-column chunks,
    // warps 4-7 the odd ones
    mbar_wait(bar_acc, 0);
    tc_fence_after();
    const int q = warp & 3;
    const int r = q * 32 + lane, w1 = m0 + r;
    const bool valid = r < rowsA;
    const float ia = 1.f / nA[r];
    const size_t px = (size_t)bh * W1 + (valid ? w1 : 0);
    const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16);
    float* out0 = out0s;
    float* out1 = out1s;
    const int stride0 = stride0s, stride1 = stride1s;
    for (int c0 = (warp >> 2) * 16; c0 < Npad; c0 += 32) {
      uint32_t raw[16];
      tmem_ld16(taddr + c0, raw);
      tmem_ld_wait();
      if (!valid) continue;
      float v[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(raw[i]) * ia * nB[c0 + i];  // nB holds 1/max(|f2|, eps)
      const int col = n0 + c0;
      if (banded) {
        // row w1 keeps the columns [s0, s0 + bw0) of level 0 and [s0 >> l, (s0 + bw0) >> l) of level l; everything is a
        // multiple of 4 columns (s0, n0, c0 are multiples of align = 4 << (L-1), W2 of 4), so whole quads move
        const int s0 = band_start(p.band, w1);
        float* brow = p.lv.p[0] + px * p.band.bw0;
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          const int a = col + 4 * q4, jb = a - s0;
          if (a < W2 && jb >= 0 && jb < p.band.bw0)
            *reinterpret_cast<float4*>(brow + jb) = make_float4(v[4 * q4], v[4 * q4 + 1], v[4 * q4 + 2], v[4 * q4 + 3]);
        }
        int n = 16, len = W2, cl = col, sl = s0, bw = p.band.bw0;
        for (int l = 1; l < p.L; ++l) {
          n >>= 1; len >>= 1; cl >>= 1; sl >>= 1; bw >>= 1;   // s0 is an exact multiple of 2^l: the shift is a division
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (i < n) v[i] = (v[2 * i] + v[2 * i + 1]) * 0.5f;
          float* rl = p.lv.p[l] + px * bw;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int jb = cl + i - sl;
            if (i < n && cl + i < len && jb >= 0 && jb < bw) rl[jb] = v[i];
          }
        }
        continue;
      }
      if (p.stage0) {  // n0 == 0: row r of the staged tile, stride W2 + 4 floats (conflict-free 16-byte stores)
        float* srow = out0 + r * stride0 + c0;
        if (c0 + 16 <= W2) {
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4)
            *reinterpret_cast<float4*>(srow + 4 * q4) = make_float4(v[4 * q4], v[4 * q4 + 1], v[4 * q4 + 2], v[4 * q4 + 3]);
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (c0 + i < W2) srow[i] = v[i];
        }
      } else {
        float* row = p.lv.p[0] + px * W2 + col;
        if (col + 16 <= W2) {
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4)
            *reinterpret_cast<float4*>(row + 4 * q4) = make_float4(v[4 * q4], v[4 * q4 + 1], v[4 * q4 + 2], v[4 * q4 + 3]);
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i)
            if (col + i < W2) row[i] = v[i];
        }
      }
      // pooled levels: mean of adjacent pairs of the previous level, floor length
      int n = 16, len = W2, cl = col;
      for (int l = 1; l < p.L; ++l) {
        n >>= 1; len >>= 1; cl >>= 1;
        if (n == 0) break;
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (i < n) v[i] = (v[2 * i] + v[2 * i + 1]) * 0.5f;
        if (l == 1 && p.stage1) {
          float* srow = out1 + r * stride1 + cl;
          if (cl + 8 <= len) {
            *reinterpret_cast<float4*>(srow) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(srow + 4) = make_float4(v[4], v[5], v[6], v[7]);
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              if (cl + i < len) srow[i] = v[i];
          }
          continue;
        }
        float* rl = p.lv.p[l] + px * len + cl;
        if (n >= 4 && (len & 3) == 0 && cl + n <= len) {
          *reinterpret_cast<float4*>(rl) = make_float4(v[0], v[1], v[2], v[3]);
          if (n == 8) *reinterpret_cast<float4*>(rl + 4) = make_float4(v[4], v[5], v[6], v[7]);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (i < n && cl + i < len) rl[i] = v[i];
        }
      }
    }
    if (banded && valid && warp < 4) {
      // band positions that fall outside the image (w1 near the left / right border) hold zeros
      const int s0 = band_start(p.band, w1);
      int len = W2, sl = s0, bw = p.band.bw0;
      for (int l = 0; l < p.L; ++l) {
        float* rl = p.lv.p[l] + px * bw;
        for (int jb = 0; jb < min(bw, -sl); ++jb) rl[jb] = 0.f;          // columns left of the image
        for (int jb = max(0, len - sl); jb < bw; ++jb) rl[jb] = 0.f;     // columns right of it
        len >>= 1; sl >>= 1; bw >>= 1;
      }
    }
    tc_fence_before();
  }
  if (p.stage0 || p.stage1) {
    // The staged tile leaves as ONE contiguous run per level (rows of a tile are consecutive in the output): every warp
    // instruction writes 512 contiguous bytes = 16 full sectors, instead of 32 half sectors in 32 different rows.
    __syncthreads();
    const size_t prow0 = (size_t)bh * W1 + m0;
    if (p.stage0) {
      const int per_row = W2 >> 2, total = rowsA * per_row;
      float4* dst = reinterpret_cast<float4*>(p.lv.p[0] + prow0 * W2);
      for (int i = tid; i < total; i += kCtThreads) {
        const int rr = i / per_row, c4 = i - rr * per_row;
        dst[i] = *reinterpret_cast<const float4*>(out0s + rr * stride0s + 4 * c4);
      }
    }
    if (p.stage1) {
      const int per_row = W2 >> 3, total = rowsA * per_row;
      float4* dst = reinterpret_cast<float4*>(p.lv.p[1] + prow0 * (W2 >> 1));
      for (int i = tid; i < total; i += kCtThreads) {
        const int rr = i / per_row, c4 = i - rr * per_row;
        dst[i] = *reinterpret_cast<const float4*>(out1s + rr * stride1s + 4 * c4);
      }
    }
  }
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)tcols) : "memory");
  }
}

int make_map(CUtensorMap* map, const float* f, int B, int C, int H, int W) {
  EncodeTiledFn enc = encode_fn();
  FSB_REQUIRE(enc != nullptr, FSB_ERR_CUDA, "fsb_corr_pyramid: cuTensorMapEncodeTiled not available from the driver");
  const cuuint64_t dims[4] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)C, (cuuint64_t)B};
  const cuuint64_t strides[3] = {(cuuint64_t)W * 4, (cuuint64_t)H * W * 4, (cuuint64_t)C * H * W * 4};
  const cuuint32_t box[4] = {kCtBoxW, 1, kCtKC, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(f), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  FSB_REQUIRE(r == CUDA_SUCCESS, FSB_ERR_CUDA, "fsb_corr_pyramid: cuTensorMapEncodeTiled failed (%d)", (int)r);
  return FSB_OK;
}

}  // namespace

bool corr_tma_eligible(const float* f1, const float* f2, int W1, int W2) {
  static int enabled = -1;
  if (enabled < 0) {
    const char* e = getenv("FSB_CORR_TMA");
    enabled = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return enabled && (W1 & 3) == 0 && (W2 & 3) == 0 && aligned16(f1) && aligned16(f2);
}

int launch_corr_tma(const float* f1, const float* f2, const LevelPtrs& lv, int B, int C, int H, int W1, int W2, int L,
                    int nprod, cudaStream_t stream, CorrBand band) {
  CorrTmaParams p;
  memset(&p, 0, sizeof(p));
  int rc = make_map(&p.map1, f1, B, C, H, W1);
  if (rc != FSB_OK) return rc;
  rc = make_map(&p.map2, f2, B, C, H, W2);
  if (rc != FSB_OK) return rc;
  p.lv = lv;
  p.C = C; p.H = H; p.W1 = W1; p.W2 = W2; p.L = L;
  p.band = band;
  int mtiles = ceil_div(W1, kCtBM);
  p.rows_m = round_up(ceil_div(W1, mtiles), 4);  // e.g. W1 = 160: two CTAs of 80 rows instead of 128 + 32
  if (band.bw0 > 0) {
    // the n window of an m tile is rows_m - 1 + bw0 + align - 1 columns at most and must fit ONE 256-column MMA tile
    FSB_REQUIRE(W1 == W2, FSB_ERR_SHAPE, "fsb_corr_pyramid_band: W1 = %d != W2 = %d", W1, W2);
    const int max_rows = (256 - band.bw0 - band.align) & ~3;
    FSB_REQUIRE(max_rows >= 32, FSB_ERR_SHAPE,
                "fsb_corr_pyramid_band: a band of %d columns (d_band, radius, levels) leaves no room for a pixel tile in a "
                "256-column window; use the full-matrix mode", band.bw0);
    const int cap = max_rows < kCtBM ? max_rows : kCtBM;
    mtiles = ceil_div(W1, cap);
    p.rows_m = round_up(ceil_div(W1, mtiles), 4);
    if (p.rows_m > cap) p.rows_m = cap;
    mtiles = ceil_div(W1, p.rows_m);
  }
  p.boxes_a = ceil_div(p.rows_m, kCtBoxW);
  p.boxes_b = ceil_div(round_up(W2 < 256 ? W2 : 256, 16), kCtBoxW);
  // the MMA reads M = 128 rows = 4 pixel blocks from the A region: blocks past boxes_a fall into the B region (their
  // accumulator rows are never read), so a stage must hold at least 4 boxes
  if (p.boxes_a + p.boxes_b < 4) p.boxes_b = 4 - p.boxes_a;
  const size_t stage = (size_t)(p.boxes_a + p.boxes_b) * kCtBox;
  // staged epilogue: one n tile, rows of 16-byte multiples, the staged tile fits the freed operand stages
  p.stage0 = band.bw0 == 0 && W2 <= 256 && aligned16(lv.p[0]) && (size_t)p.rows_m * (W2 + 4) * 4 <= 2 * stage;
  p.stage1 = band.bw0 == 0 && L >= 2 && W2 <= 256 && (W2 & 7) == 0 && aligned16(lv.p[1]) &&
             (size_t)p.rows_m * ((W2 >> 1) + 4) * 4 <= stage;
  const size_t smem = 3 * stage + 1024;
  FSB_REQUIRE(smem <= 220 * 1024, FSB_ERR_SHAPE, "fsb_corr_pyramid: %zu B shared memory", smem);
  dim3 grid(mtiles, band.bw0 > 0 ? 1 : ceil_div(W2, 256), B * H);
  auto kern = nprod == 3 ? corr_tma_kernel<3> : corr_tma_kernel<1>;
  cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (err != cudaSuccess) return cuda_fail(err, "fsb_corr_pyramid: cudaFuncSetAttribute");
  err = launch_kernel(kern, grid, dim3(kCtThreads), smem, stream, p);
  if (err != cudaSuccess) return cuda_fail(err, "fsb_corr_pyramid: launch");
  FSB_CHECK_LAUNCH("fsb_corr_pyramid");
  return FSB_OK;
}

}  // namespace fsb
