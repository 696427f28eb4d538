// Kernel 2 on the tensor cores: per-row all-pairs init correlation as a tcgen05 / TMEM GEMM with
// the avg-pool pyramid fused into its epilogue.
//
// Replaces Combined_Geo_Encoding_Volume.corr (core/geometry.py:68-77: F.normalize over C, then
// einsum 'aijk,aijh->ajkh') and the init_corr pyramid (core/geometry.py:31,38-40).
//
// GEMM view for one image row (b,h):  D[w1, w2] = sum_c f1[c, w1] * f2[c, w2], M = W1, N = W2, K = C.
// The (C, W) planes of NCHW features are "M-major"; instead of asking the tensor core for that
// layout, the CTA transposes while it stages: coalesced loads along W, 16-byte shared-memory
// stores into the K-major 128-byte-swizzled tile layout (row = pixel, 32 fp32 of K per row), which
// is the layout tcgen05.mma kind::tf32 reads.  The same pass accumulates ||f||^2 per pixel; the L2
// normalisation is applied in the epilogue (D / max(|f1|,eps) / max(|f2|,eps)), so the features
// are read exactly once.
//
// fp32 parity on a TF32 pipe: NPROD == 3 splits every operand into hi = tf32(x) and
// lo = tf32(x - hi) and issues hi*hi + hi*lo + lo*hi into the same fp32 TMEM accumulator; the
// dropped lo*lo term is ~2^-22 relative, so the result matches the fp32 reference to ~1e-6.
// NPROD == 1 is plain single-pass TF32 (reported separately, ~5e-4).
//
// CTA = (m tile of 128 w1, n tile of <= 256 w2, one (b,h)); K streams in chunks of 32 channels
// through two shared-memory buffers; accumulator = 128 lanes x Npad fp32 columns of TMEM.
// Epilogue: thread = row w1, 32 columns per tcgen05.ld, level 0 stored as 128-byte runs and every
// pooled level (mean of adjacent pairs, floor length) formed in registers from the same values.
#include "common.cuh"

namespace fsb {

constexpr int kCuThreads = 256;
constexpr int kCuBM = 128;
constexpr int kCuKC = 32;  // fp32 per 128-byte row

__device__ __forceinline__ uint32_t cu_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool cu_mbar_try(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void cu_mbar_wait(uint32_t bar, uint32_t parity) {
  if (cu_mbar_try(bar, parity)) return;
  const long long t0 = clock64();
  uint32_t spins = 0;
  while (!cu_mbar_try(bar, parity))
    if ((++spins & 255u) == 0 && clock64() - t0 > 4000000000ll) __trap();
}
__device__ __forceinline__ float tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ uint64_t cu_desc_k128(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ void cu_mma_tf32(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void cu_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void cu_tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr) : "memory");
}

// Stage one K chunk (32 channels) of `rows_pad` pixels starting at w0 into K-major SWIZZLE_128B
// tiles: row r = pixel, 16-byte chunk j holds channels c0+4j..c0+4j+3 and lives at chunk slot
// j ^ (r & 7).  A work item is a 4 (pixels) x 4 (channels) block: four 16-byte loads along W (one
// per channel), a register transpose, four 16-byte shared stores (one per pixel row).  Item index
// = rq*8 + j with j fastest, so the 8 lanes of a store phase hit 8 different swizzle slots (no bank
// conflict) and item (rq, j) is the same for every chunk: ||.||^2 partials stay in registers.
// hi = x with the 13 low mantissa bits cleared (what the TF32 datapath reads), lo = x - hi (exact
// in fp32; the datapath truncates it to TF32 again, leaving ~2^-20 relative).
// The chunk is fetched in two halves so that the loads of chunk ch+1 are in flight while the MMAs of chunk ch run:
// cu_load (global -> registers, issued right after the MMAs of the previous chunk) and cu_store (registers ->
// swizzled hi/lo tiles, after the buffer is free).
template <int ITEMS, bool VEC>
__device__ __forceinline__ void cu_load(const float* __restrict__ plane, size_t HW, int w0, int W, int rows_pad, int c0,
                                        int C, float (&v)[ITEMS][4][4]) {
#pragma unroll
  for (int it = 0; it < ITEMS; ++it) {
    const int idx = threadIdx.x + it * kCuThreads;
    if (idx < rows_pad * 2) {  // (rows_pad / 4) pixel quads x 8 channel quads
      const int j = idx & 7, rq = idx >> 3;
      const int w = w0 + 4 * rq;
#pragma unroll
      for (int q = 0; q < 4; ++q) {  // [channel][pixel]
        const int c = c0 + 4 * j + q;
        const float* src = plane + (size_t)c * HW + w;
        if (c < C && VEC && w + 3 < W) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(src));
          v[it][q][0] = t.x; v[it][q][1] = t.y; v[it][q][2] = t.z; v[it][q][3] = t.w;
        } else {
#pragma unroll
          for (int i = 0; i < 4; ++i) v[it][q][i] = (c < C && w + i < W) ? __ldg(src + i) : 0.f;
        }
      }
    }
  }
}
template <int NPROD, int ITEMS>
__device__ __forceinline__ void cu_store(uint8_t* hi, uint8_t* lo, int rows_pad, const float (&v)[ITEMS][4][4],
                                         float (&ss)[ITEMS][4]) {
#pragma unroll
  for (int it = 0; it < ITEMS; ++it) {
    const int idx = threadIdx.x + it * kCuThreads;
    if (idx < rows_pad * 2) {
      const int j = idx & 7, rq = idx >> 3;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = 4 * rq + i;
        const uint32_t off = (uint32_t)r * 128u + (uint32_t)((j ^ (r & 7)) << 4);
        float h[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          ss[it][i] = fmaf(v[it][q][i], v[it][q][i], ss[it][i]);
          h[q] = NPROD == 3 ? __uint_as_float(__float_as_uint(v[it][q][i]) & 0xffffe000u) : v[it][q][i];
        }
        *reinterpret_cast<float4*>(hi + off) = make_float4(h[0], h[1], h[2], h[3]);
        if (NPROD == 3)
          *reinterpret_cast<float4*>(lo + off) =
              make_float4(v[it][0][i] - h[0], v[it][1][i] - h[1], v[it][2][i] - h[2], v[it][3][i] - h[3]);
      }
    }
  }
}

template <int NPROD>
__global__ void __launch_bounds__(kCuThreads)
corr_umma_kernel(const float* __restrict__ f1, const float* __restrict__ f2, LevelPtrs lv, int C, int H, int W1,
                 int W2, int L, int nbuf, int rows_m) {
  pdl_launch_dependents();  // the next kernel may start its prologue
  pdl_wait();               // inputs written by the previous kernel are complete and visible
  extern __shared__ __align__(1024) uint8_t cu_smem[];
  __shared__ __align__(8) uint64_t bars[3];
  __shared__ uint32_t tmem_slot;
  __shared__ float ssA[8][kCuBM], ssB[8][256];  // [channel quad j][pixel] partial sums of squares
  __shared__ float nA[kCuBM], nB[256];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // rows_m <= 128 rows of w1 per CTA (W1 split evenly over the m tiles, a multiple of 4): only those rows of the
  // 128-row A tile are staged and stored; the MMA still runs M = 128, the other accumulator rows are never read
  const int m0 = blockIdx.x * rows_m, n0 = blockIdx.y * 256;
  const int rowsA = min(rows_m, (W1 - m0 + 3) & ~3);
  const int bh = blockIdx.z, b = bh / H, h = bh - b * H;
  const int Nv = min(256, W2 - n0);         // valid columns of this tile
  const int Npad = (Nv + 15) & ~15;         // MMA N
  const size_t HW1 = (size_t)H * W1, HW2 = (size_t)H * W2;
  const float* a_plane = f1 + (size_t)b * C * HW1 + (size_t)h * W1;
  const float* b_plane = f2 + (size_t)b * C * HW2 + (size_t)h * W2;

  const uint32_t base = (cu_smem_u32(cu_smem) + 1023u) & ~1023u;
  uint8_t* gbase = cu_smem + (base - cu_smem_u32(cu_smem));
  const uint32_t a_bytes = kCuBM * 128, b_bytes = (uint32_t)Npad * 128;
  const uint32_t buf_bytes = (NPROD == 3 ? 2u : 1u) * (a_bytes + b_bytes);  // [A_hi][A_lo][B_hi][B_lo]
  const uint32_t off_alo = a_bytes, off_bhi = (NPROD == 3 ? 2u : 1u) * a_bytes, off_blo = off_bhi + b_bytes;

  const uint32_t bar0 = cu_smem_u32(&bars[0]), bar_acc = cu_smem_u32(&bars[2]);
  int tcols = 32;
  while (tcols < Npad) tcols <<= 1;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_acc) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(cu_smem_u32(&tmem_slot)),
                 "r"((uint32_t)tcols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  // instruction descriptor, kind::tf32: D fp32, A/B tf32, both K-major, M = 128, N = Npad
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(Npad >> 3) << 17) | ((uint32_t)(kCuBM >> 4) << 24);

  float pa[1][4] = {{0.f, 0.f, 0.f, 0.f}};
  float pb[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
  // 16-byte loads along W need W % 4 == 0 (every (c,h) row then starts 16-byte aligned) and an aligned base
  const bool vecA = (W1 & 3) == 0 && (reinterpret_cast<uintptr_t>(f1) & 15) == 0;
  const bool vecB = (W2 & 3) == 0 && (reinterpret_cast<uintptr_t>(f2) & 15) == 0;
  const int nchunks = (C + kCuKC - 1) / kCuKC;
  float va[1][4][4], vb[2][4][4];
  auto load_chunk = [&](int ch) {
    if (vecA) cu_load<1, true>(a_plane, HW1, m0, W1, rowsA, ch * kCuKC, C, va);
    else      cu_load<1, false>(a_plane, HW1, m0, W1, rowsA, ch * kCuKC, C, va);
    if (vecB) cu_load<2, true>(b_plane, HW2, n0, W2, Npad, ch * kCuKC, C, vb);
    else      cu_load<2, false>(b_plane, HW2, n0, W2, Npad, ch * kCuKC, C, vb);
  };
  load_chunk(0);
  for (int ch = 0; ch < nchunks; ++ch) {
    // nbuf == 2: double buffered (chunk ch waits for the MMAs of chunk ch-2); nbuf == 1: one buffer, so
    // two CTAs fit an SM and overlap each other instead (chunk ch waits for the MMAs of chunk ch-1)
    const int buf = nbuf == 2 ? (ch & 1) : 0;
    if (nbuf == 2) {
      if (ch >= 2) cu_mbar_wait(bar0 + 8 * buf, ((ch >> 1) - 1) & 1);
    } else if (ch >= 1) {
      cu_mbar_wait(bar0, (ch - 1) & 1);
    }
    uint8_t* g = gbase + (size_t)buf * buf_bytes;
    cu_store<NPROD, 1>(g, g + off_alo, rowsA, va, pa);
    cu_store<NPROD, 2>(g + off_bhi, g + off_blo, Npad, vb, pb);
    if (ch + 1 < nchunks) load_chunk(ch + 1);  // in flight across the barrier, the MMAs and the wait above
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy stores -> visible to the MMA
    __syncthreads();
    if (tid == 0) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t s = base + buf * buf_bytes;
      const uint64_t ahi = cu_desc_k128(s), alo = cu_desc_k128(s + off_alo);
      const uint64_t bhi = cu_desc_k128(s + off_bhi), blo = cu_desc_k128(s + off_blo);
#pragma unroll
      for (int k = 0; k < 4; ++k) {  // 8 tf32 (32 bytes) per MMA: +2 in the descriptor's 16-byte address field
        cu_mma_tf32(tmem, ahi + 2 * k, bhi + 2 * k, idesc, (ch > 0 || k > 0) ? 1u : 0u);
        if (NPROD == 3) {
          cu_mma_tf32(tmem, ahi + 2 * k, blo + 2 * k, idesc, 1u);
          cu_mma_tf32(tmem, alo + 2 * k, bhi + 2 * k, idesc, 1u);
        }
      }
      cu_commit(bar0 + 8 * buf);
      if (ch == nchunks - 1) cu_commit(bar_acc);
    }
  }
  // ||f||: partials -> shared -> per pixel max(sqrt(ss), 1e-12)  (F.normalize's denominator)
  {
    const int j = tid & 7, rq = tid >> 3;  // item 0 of A: 32 pixel quads x 8 channel quads
#pragma unroll
    for (int i = 0; i < 4; ++i) ssA[j][4 * rq + i] = pa[0][i];
  }
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int idx = tid + it * kCuThreads;
    if (idx < Npad * 2) {
      const int j = idx & 7, rq = idx >> 3;
#pragma unroll
      for (int i = 0; i < 4; ++i) ssB[j][4 * rq + i] = pb[it][i];
    }
  }
  __syncthreads();
  if (tid < kCuBM) {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += ssA[j][tid];
    nA[tid] = fmaxf(sqrtf(s), 1e-12f);
  }
  if (tid < Npad) {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += ssB[j][tid];
    nB[tid] = 1.f / fmaxf(sqrtf(s), 1e-12f);
  }
  __syncthreads();

  {
    // epilogue on all 8 warps: warp w reads TMEM lane quarter (w & 3); warps 0-3 take the even
    // 16-column chunks, warps 4-7 the odd ones
    cu_mbar_wait(bar_acc, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int q = warp & 3;
    const int r = q * 32 + lane, w1 = m0 + r;
    const bool valid = r < rowsA && w1 < W1;
    const float ia = 1.f / nA[r];
    const size_t p = (size_t)bh * W1 + (valid ? w1 : 0);
    const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16);
    const bool vec = (W2 & 3) == 0;
    for (int c0 = (warp >> 2) * 16; c0 < Npad; c0 += 32) {
      uint32_t raw[16];
      cu_tmem_ld16(taddr + c0, raw);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (!valid) continue;
      float v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(raw[j]) * ia * nB[c0 + j];  // nB holds 1/max(|f2|, eps)
      const int col = n0 + c0;
      float* row = lv.p[0] + p * W2 + col;
      if (vec && col + 16 <= W2) {
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4)
          *reinterpret_cast<float4*>(row + 4 * q4) = make_float4(v[4 * q4], v[4 * q4 + 1], v[4 * q4 + 2], v[4 * q4 + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (col + j < W2) row[j] = v[j];
      }
      // pooled levels: mean of adjacent pairs of the previous level, floor length
      int n = 16, len = W2, cl = col;
      for (int l = 1; l < L; ++l) {
        n >>= 1; len >>= 1; cl >>= 1;
        if (n == 0) break;
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j < n) v[j] = (v[2 * j] + v[2 * j + 1]) * 0.5f;
        float* rl = lv.p[l] + p * len + cl;
        if (n >= 4 && (len & 3) == 0 && cl + n <= len) {
          *reinterpret_cast<float4*>(rl) = make_float4(v[0], v[1], v[2], v[3]);
          if (n == 8) *reinterpret_cast<float4*>(rl + 4) = make_float4(v[4], v[5], v[6], v[7]);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j < n && cl + j < len) rl[j] = v[j];
        }
      }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)tcols) : "memory");
  }
}

int launch_corr_umma(const float* f1, const float* f2, const LevelPtrs& lv, int B, int C, int H, int W1, int W2, int L,
                     int nprod, cudaStream_t stream) {
  const int Npad = round_up(W2 < 256 ? W2 : 256, 16);
  const size_t buf = (size_t)(nprod == 3 ? 2 : 1) * ((size_t)kCuBM * 128 + (size_t)Npad * 128);
  // two CTAs per SM (<= ~96 KB of dynamic shared memory each) beat one double-buffered CTA: the chunk
  // loop of a CTA is a serial load -> stage -> MMA chain
  const int nbuf = (2 * buf + 1024 <= 96 * 1024) ? 2 : 1;
  const size_t smem = nbuf * buf + 1024;
  FSB_REQUIRE(smem <= 227 * 1024 - 16 * 1024, FSB_ERR_SHAPE, "fsb_corr_pyramid: %zu B shared memory", smem);
  const int mtiles = ceil_div(W1, kCuBM);
  const int rows_m = round_up(ceil_div(W1, mtiles), 4);  // e.g. W1 = 160: two CTAs of 80 rows instead of 128 + 32
  dim3 grid(mtiles, ceil_div(W2, 256), B * H);
  auto kern = nprod == 3 ? corr_umma_kernel<3> : corr_umma_kernel<1>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return cuda_fail(e, "fsb_corr_pyramid: cudaFuncSetAttribute");
  launch_kernel(kern, dim3(grid), dim3(kCuThreads), smem, stream, f1, f2, lv, C, H, W1, W2, L, nbuf, rows_m);
  FSB_CHECK_LAUNCH("fsb_corr_pyramid");
  return FSB_OK;
}

}  // namespace fsb
