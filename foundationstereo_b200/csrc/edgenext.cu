// Kernel 4d: a whole EdgeNextConvEncoder block (core/submodule.py:565-591, norm=None; used twice by DispHead,
// core/update.py:26-27) as ONE tcgen05 kernel:
//     t   = dwconv7x7(x) + bd                              depthwise, pad 3      (core/submodule.py:581)
//     h   = GELU(t @ W1^T + b1)                            pwconv1, 128 -> 512   (:584-585)
//     y   = h @ W2^T + b2                                  pwconv2, 512 -> 128   (:586)
//     out = x + gamma * y                                  layer scale + residual (:587-590)
// csrc/mlp.cu does everything after the depthwise convolution; the depthwise convolution itself was a second launch
// (csrc/nhwc.cu, 80-105 us at 8 pairs against 88-115 us for the MLP) whose 39 MB output made a round trip through
// L2 / HBM.  Here four extra warps of the MLP's persistent CTA compute the depthwise tile straight into the shared
// memory A operand of G1 (K-major SWIZZLE_128B, what the TMA load used to deliver), under the GELU epilogue of the
// previous tile:
//   tile = 8 x 16 pixels of one image (128 GEMM rows); the input patch (14 x 22 pixels) streams in with cp.async, 64
//   channels at a time (39 KB), zero-filled outside the image = the convolution's padding;
//   dw thread = 4 channels x a strip of 4 pixels x 2 rows, packed fp32 FMAs, the same accumulation order as
//   dwconv7x7_kernel (results are bit-identical to the two-launch path);
//   MMA / GELU / second GEMM / final epilogue exactly as in mlp.cu (see there for the pipeline), with ONE X slot: the
//   dw warps refill it as soon as G1(3) of the previous tile has read it.
// Warps: 0 = TMEM + MMA issuer, 1 = weight producer (TMA), 2..17 = epilogue (16), 18..25 = depthwise (8).
#include "common.cuh"
#include "umma.cuh"

namespace fsb {

constexpr int kEbC = 128;             // block channels
constexpr int kEbF = 512;             // hidden features (expansion 4)
constexpr int kEbChunks = kEbF / 128;
constexpr int kEbEpiWarps = 16;
constexpr int kEbDwWarps = 8;
constexpr int kEbThreads = (2 + kEbEpiWarps + kEbDwWarps) * 32;  // 704
constexpr int kEbWStages = 3;
constexpr int kEbTile = 128 * 64 * 2;  // 16 KB: 128 rows x 64 16-bit elements, SWIZZLE_128B
constexpr int kEbTY = 8, kEbTX = 16, kEbPY = kEbTY + 6, kEbPX = kEbTX + 6;
constexpr int kEbPatch = kEbPY * kEbPX * 64 * 2;  // one 64-channel half of the input patch: 39424 B

struct alignas(64) EbParams {
  CUtensorMap w1map, w2map;
  const void* x;        // block input (B,H,W,128) 16-bit NHWC = depthwise input AND residual
  const float* dw_w;    // (128, 49) fp32
  const float* dw_b;    // (128)
  const float* b1;
  const float* b2;
  const float* gamma;
  void* out;            // (B,H,W,128) 16-bit NHWC
  int H, W, tiles_x, tiles_y, ntiles;
  uint32_t idesc;
};

// shared-memory carve-up (offsets from the 1024-byte aligned base)
constexpr uint32_t kEbOffX = 0;                                   // ONE slot x 2 sub-tiles (K halves)
constexpr uint32_t kEbOffW = kEbOffX + 2 * kEbTile;               // ring
constexpr uint32_t kEbOffH = kEbOffW + kEbWStages * kEbTile;      // 2 buffers x 2 sub-tiles
constexpr uint32_t kEbOffP = kEbOffH + 2 * 2 * kEbTile;           // input patch, one 64-channel half
constexpr uint32_t kEbOffDw = kEbOffP + ((kEbPatch + 127) & ~127);  // depthwise weights [49][128] fp32
constexpr uint32_t kEbSmem = kEbOffDw + 49 * kEbC * 4 + 1024;

__device__ __forceinline__ void eb_cp_async16_zfill(uint32_t smem_dst, const void* gsrc, bool valid) {
  const int n = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_dst), "l"(gsrc), "r"(n) : "memory");
}
__device__ __forceinline__ void eb_bar_dw() {  // the four depthwise warps only (named barrier 1)
  asm volatile("bar.sync 1, %0;" ::"n"(kEbDwWarps * 32) : "memory");
}
template <typename T> __device__ __forceinline__ float2 eb_unpack2(uint32_t v);
template <> __device__ __forceinline__ float2 eb_unpack2<__half>(uint32_t v) {
  return __half22float2(*reinterpret_cast<const __half2*>(&v));
}
template <> __device__ __forceinline__ float2 eb_unpack2<__nv_bfloat16>(uint32_t v) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v));
}

template <typename T>
__global__ void __launch_bounds__(kEbThreads, 1)
edgenext_block_kernel(const __grid_constant__ EbParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // barriers: x_full x_empty | w_full[3] w_empty[3] | a1_full[2] a1_empty[2] h_full[2] h_empty[2] a2_full[2] a2_empty[2]
  __shared__ __align__(8) uint64_t bars[2 + 2 * kEbWStages + 12];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float b1_s[kEbF];
  __shared__ __align__(16) float b2_s[kEbC], g_s[kEbC], dwb_s[kEbC];

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t x_full = smem_u32(&bars[0]), x_empty = smem_u32(&bars[1]);
  const uint32_t w_full = smem_u32(&bars[2]), w_empty = smem_u32(&bars[2 + kEbWStages]);
  const uint32_t a1_full = smem_u32(&bars[2 + 2 * kEbWStages]), a1_empty = a1_full + 16;
  const uint32_t h_full = a1_full + 32, h_empty = a1_full + 48;
  const uint32_t a2_full = a1_full + 64, a2_empty = a1_full + 80;

  if (warp == 0 && lane == 0) {
    mbar_init(x_full, kEbDwWarps);
    mbar_init(x_empty, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(a1_full + 8 * i, 1);
      mbar_init(a1_empty + 8 * i, kEbEpiWarps);
      mbar_init(h_full + 8 * i, kEbEpiWarps);
      mbar_init(h_empty + 8 * i, 1);
      mbar_init(a2_full + 8 * i, 1);
      mbar_init(a2_empty + 8 * i, kEbEpiWarps);
    }
    for (int i = 0; i < kEbWStages; ++i) {
      mbar_init(w_full + 8 * i, 1);
      mbar_init(w_empty + 8 * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1 && lane == 0) {
    tma_prefetch_desc(&p.w1map);
    tma_prefetch_desc(&p.w2map);
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp >= 2 && warp < 2 + kEbEpiWarps) {
    const int i = threadIdx.x - 64;  // 0..511
    b1_s[i] = __ldg(p.b1 + i);
    if (i < kEbC) {
      b2_s[i] = __ldg(p.b2 + i);
      g_s[i] = __ldg(p.gamma + i);
      dwb_s[i] = __ldg(p.dw_b + i);
    }
  }
  if (warp >= 2 + kEbEpiWarps) {
    // depthwise weights (C, 49) -> [tap][channel] fp32: parameters, no kernel writes them
    float* wsm = reinterpret_cast<float*>(gbase + kEbOffDw);
    for (int i = threadIdx.x - (2 + kEbEpiWarps) * 32; i < 49 * kEbC; i += kEbDwWarps * 32) {
      const int c = i / 49, t = i - c * 49;
      wsm[t * kEbC + c] = __ldg(p.dw_w + i);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;  // acc1: cols [0,256), acc2: cols [256,512)
  pdl_launch_dependents();
  if (warp != 1) pdl_wait();  // the weight producer only reads parameters no kernel writes

  const int tiles_per_img = p.tiles_x * p.tiles_y;

  if (warp == 1) {
    // ================================================================= W producer (consumption order)
    int s = 0;
    uint32_t ph = 0;
    bool first_round = true;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
      // sequence of 8 GEMM slices: G1(0) G1(1) G2(0) G1(2) G2(1) G1(3) G2(2) G2(3), two K halves each
#pragma unroll 1
      for (int q = 0; q < 8; ++q) {
        const int is2 = (0xD4 >> q) & 1;                         // bits 2,4,6,7 -> GEMM 2
        const int j = is2 ? (q == 2 ? 0 : q == 4 ? 1 : q == 6 ? 2 : 3) : (q == 0 ? 0 : q == 1 ? 1 : q == 3 ? 2 : 3);
#pragma unroll 1
        for (int k = 0; k < 2; ++k) {
          if (!first_round) mbar_wait(w_empty + 8 * s, ph ^ 1);
          __syncwarp();
          if (elect_one()) {
            mbar_expect_tx(w_full + 8 * s, kEbTile);
            const uint32_t dst = base + kEbOffW + s * kEbTile;
            if (is2) tma_load_2d(dst, &p.w2map, w_full + 8 * s, j * 128 + k * 64, 0);
            else     tma_load_2d(dst, &p.w1map, w_full + 8 * s, k * 64, j * 128);
          }
          if (++s == kEbWStages) { s = 0; ph ^= 1; first_round = false; }
        }
      }
    }
  } else if (warp == 0) {
    // ================================================================= MMA issuer
    int ws = 0;
    uint32_t wph = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
      const uint32_t acc2 = tmem_base + 256u + (uint32_t)(it & 1) * 128u;
      mbar_wait(x_full, (uint32_t)(it & 1));                                  // the dw warps wrote X for this tile
      mbar_wait(a2_empty + 8 * (it & 1), ((uint32_t)(it >> 1) & 1) ^ 1);      // final epilogue of tile it-2 drained this buffer
      tc_fence_after();
#pragma unroll 1
      for (int q = 0; q < 8; ++q) {
        const int is2 = (0xD4 >> q) & 1;
        const int j = is2 ? (q == 2 ? 0 : q == 4 ? 1 : q == 6 ? 2 : 3) : (q == 0 ? 0 : q == 1 ? 1 : q == 3 ? 2 : 3);
        const int b = j & 1;
        const uint32_t use = (uint32_t)(it * 2 + (j >> 1));  // how often acc1[b] / H[b] were used before
        uint32_t a_addr, d;
        if (!is2) {
          mbar_wait(a1_empty + 8 * b, (use & 1) ^ 1);  // E1 of the chunk two back has read acc1[b]
          a_addr = base + kEbOffX;
          d = tmem_base + (uint32_t)b * 128u;
        } else {
          mbar_wait(h_full + 8 * b, use & 1);  // E1(j) wrote H[b]
          a_addr = base + kEbOffH + b * 2 * kEbTile;
          d = acc2;
        }
        tc_fence_after();
#pragma unroll 1
        for (int k = 0; k < 2; ++k) {
          mbar_wait(w_full + 8 * ws, wph);
          tc_fence_after();
          __syncwarp();
          const uint64_t adesc = umma_desc_k128(a_addr + k * kEbTile);
          const uint64_t bdesc = umma_desc_k128(base + kEbOffW + ws * kEbTile);
          const uint32_t accum = (!is2) ? (uint32_t)(k > 0) : (uint32_t)(j > 0 || k > 0);
          if (elect_one()) {
            umma_f16(d, adesc, bdesc, p.idesc, accum);
            umma_f16(d, adesc + 2, bdesc + 2, p.idesc, 1);
            umma_f16(d, adesc + 4, bdesc + 4, p.idesc, 1);
            umma_f16(d, adesc + 6, bdesc + 6, p.idesc, 1);
            umma_commit(w_empty + 8 * ws);
          }
          if (++ws == kEbWStages) { ws = 0; wph ^= 1; }
        }
        if (elect_one()) {
          if (!is2) {
            umma_commit(a1_full + 8 * b);
            if (j == 3) umma_commit(x_empty);  // X free once the last G1 has read it: the dw warps refill it
          } else {
            umma_commit(h_empty + 8 * b);
            if (j == 3) umma_commit(a2_full + 8 * (it & 1));
          }
        }
      }
    }
  } else if (warp < 2 + kEbEpiWarps) {
    // ================================================================= epilogue warps
    // warp e = warp - 2: TMEM lane quarter (warp & 3) -- the hardware rule -- and 32-column group (e >> 2)
    const int e = warp - 2;
    const int q = warp & 3, g = e >> 2;
    const int r = q * 32 + lane;  // GEMM row = pixel (r >> 4, r & 15) of the 8 x 16 tile
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    // H sub-tile (g >> 1), 16-byte units (g & 1) * 4 .. + 3 of row r, SWIZZLE_128B: unit ^ (r & 7)
    uint8_t* hrow = gbase + kEbOffH + (g >> 1) * kEbTile + r * 128;
    const int u0 = (g & 1) * 4, sw = r & 7;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
#pragma unroll 1
      for (int j = 0; j < kEbChunks; ++j) {
        const int b = j & 1;
        const uint32_t use = (uint32_t)(it * 2 + (j >> 1));
        mbar_wait(a1_full + 8 * b, use & 1);
        tc_fence_after();
        uint32_t v[32];
        tmem_ld32(lane_addr + (uint32_t)(b * 128 + g * 32), v);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(a1_empty + 8 * b);  // acc1[b] may be overwritten by G1(j+2)
        const float* bias = b1_s + j * 128 + g * 32;
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float2 bb = *reinterpret_cast<const float2*>(bias + 2 * i);
          const float2 y = gelu2_f(__fadd2_rn(make_float2(__uint_as_float(v[2 * i]), __uint_as_float(v[2 * i + 1])), bb));
          pk[i] = pack2<T>(y.x, y.y);
        }
        mbar_wait(h_empty + 8 * b, (use & 1) ^ 1);  // G2 of the chunk two back has read H[b]
        uint8_t* hb = hrow + b * 2 * kEbTile;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          *reinterpret_cast<uint4*>(hb + (((u0 + i) ^ sw) << 4)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to tcgen05.mma
        __syncwarp();
        if (lane == 0) mbar_arrive(h_full + 8 * b);
      }
      // ---- final epilogue of the tile: out = x + gamma * (acc2 + b2)
      const int ab = it & 1;
      const uint32_t use2 = (uint32_t)(it >> 1);
      mbar_wait(a2_full + 8 * ab, use2 & 1);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld32(lane_addr + 256u + (uint32_t)(ab * 128 + g * 32), v);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(a2_empty + 8 * ab);
      const int bimg = tile / tiles_per_img, trem = tile - bimg * tiles_per_img;
      const int tyi = trem / p.tiles_x, txi = trem - tyi * p.tiles_x;
      const int y = tyi * kEbTY + (r >> 4), x = txi * kEbTX + (r & 15);
      if (y < p.H && x < p.W) {
        const size_t pix = ((size_t)bimg * p.H + y) * p.W + x;
        float f[32], rr[32];
        load32<T>(reinterpret_cast<const T*>(p.x) + pix * kEbC + g * 32, rr);
        const float4* g4 = reinterpret_cast<const float4*>(g_s + g * 32);
        const float4* b4 = reinterpret_cast<const float4*>(b2_s + g * 32);
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          const float4 gg = g4[i4], bb = b4[i4];
          f[4 * i4 + 0] = fmaf(gg.x, __uint_as_float(v[4 * i4 + 0]) + bb.x, rr[4 * i4 + 0]);
          f[4 * i4 + 1] = fmaf(gg.y, __uint_as_float(v[4 * i4 + 1]) + bb.y, rr[4 * i4 + 1]);
          f[4 * i4 + 2] = fmaf(gg.z, __uint_as_float(v[4 * i4 + 2]) + bb.z, rr[4 * i4 + 2]);
          f[4 * i4 + 3] = fmaf(gg.w, __uint_as_float(v[4 * i4 + 3]) + bb.w, rr[4 * i4 + 3]);
        }
        store32<T>(reinterpret_cast<T*>(p.out) + pix * kEbC + g * 32, f);
      }
    }
  } else {
    // ================================================================= depthwise warps
    constexpr int kRows = kEbTY / (kEbDwWarps / 2);             // rows of the tile per thread (2 warps cover a row group)
    const int d = (int)threadIdx.x - (2 + kEbEpiWarps) * 32;  // 0 .. 32 * kEbDwWarps - 1
    const int cq = d & 15, xs = (d >> 4) & 3, rg = d >> 6;     // 4 channels, x strip of 4 pixels, rows rg*kRows .. + kRows - 1
    const float* wsm = reinterpret_cast<const float*>(gbase + kEbOffDw);
    const T* xin = reinterpret_cast<const T*>(p.x);
    auto prefetch = [&](int tile, int half) {   // 14 x 22 pixels x 64 channels, 16 bytes per copy, zero outside the image
      const int bimg = tile / tiles_per_img, trem = tile - bimg * tiles_per_img;
      const int tyi = trem / p.tiles_x, txi = trem - tyi * p.tiles_x;
      const int ty0 = tyi * kEbTY - 3, tx0 = txi * kEbTX - 3;
      for (int i = d; i < kEbPY * kEbPX * 8; i += kEbDwWarps * 32) {
        const int pp = i >> 3, c8 = i & 7;
        const int py = pp / kEbPX, px = pp - py * kEbPX;
        const int yy = ty0 + py, xx = tx0 + px;
        const bool ok = yy >= 0 && yy < p.H && xx >= 0 && xx < p.W;
        const T* src = xin + (((size_t)bimg * p.H + (ok ? yy : 0)) * p.W + (ok ? xx : 0)) * kEbC + half * 64 + c8 * 8;
        eb_cp_async16_zfill(base + kEbOffP + (uint32_t)pp * 128u + (uint32_t)c8 * 16u, src, ok);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int it = 0;
    if ((int)blockIdx.x < p.ntiles) prefetch(blockIdx.x, 0);
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
#pragma unroll 1
      for (int half = 0; half < 2; ++half) {
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        eb_bar_dw();  // the patch of (tile, half) is complete and visible to the four warps
        if (half == 0) mbar_wait(x_empty, (uint32_t)((it & 1) ^ 1));  // G1(3) of the previous tile has read X (passes at it = 0)
        const float4 bv = *reinterpret_cast<const float4*>(dwb_s + half * 64 + cq * 4);
        const uint8_t* patch = gbase + kEbOffP;
#pragma unroll 1
        for (int i = 0; i < kRows; ++i) {
          const int row = rg * kRows + i;
          float2 acc[4][2];
#pragma unroll
          for (int j = 0; j < 4; ++j) { acc[j][0] = make_float2(bv.x, bv.y); acc[j][1] = make_float2(bv.z, bv.w); }
#pragma unroll
          for (int ky = 0; ky < 7; ++ky) {
            float2 in[10][2];
            const uint8_t* prow = patch + ((row + ky) * kEbPX + xs * 4) * 128 + cq * 8;
#pragma unroll
            for (int j = 0; j < 10; ++j) {
              const uint2 u = *reinterpret_cast<const uint2*>(prow + j * 128);
              in[j][0] = eb_unpack2<T>(u.x);
              in[j][1] = eb_unpack2<T>(u.y);
            }
#pragma unroll
            for (int kx = 0; kx < 7; ++kx) {
              const float4 w = *reinterpret_cast<const float4*>(wsm + (ky * 7 + kx) * kEbC + half * 64 + cq * 4);
              const float2 w01 = make_float2(w.x, w.y), w23 = make_float2(w.z, w.w);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                acc[j][0] = __ffma2_rn(in[j + kx][0], w01, acc[j][0]);
                acc[j][1] = __ffma2_rn(in[j + kx][1], w23, acc[j][1]);
              }
            }
          }
          // 8 bytes (4 channels) of GEMM row r into the K-major SWIZZLE_128B operand tile of K half `half`
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int r = row * 16 + xs * 4 + j;
            uint8_t* dst = gbase + kEbOffX + half * kEbTile + r * 128 + ((((cq >> 1) ^ (r & 7))) << 4) + (cq & 1) * 8;
            *reinterpret_cast<uint2*>(dst) = make_uint2(pack2<T>(acc[j][0].x, acc[j][0].y), pack2<T>(acc[j][1].x, acc[j][1].y));
          }
        }
        eb_bar_dw();  // everyone is done with the patch before the next one lands in it
        if (half == 0) prefetch(tile, 1);
        else if (tile + (int)gridDim.x < p.ntiles) prefetch(tile + gridDim.x, 0);
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes of X -> visible to tcgen05.mma
      __syncwarp();
      if (lane == 0) mbar_arrive(x_full);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  __syncwarp();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

}  // namespace fsb

// ------------------------------------------------------------------------------------------------------
extern "C" int fsb_edgenext_block(const void* x, const float* dw_w, const float* dw_b, const void* w1, const float* b1,
                                  const void* w2, const float* b2, const float* gamma, void* out, int B, int H, int W,
                                  int C, int F, int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_edgenext_block");
  using namespace fsb;
  FSB_REQUIRE(x && dw_w && dw_b && w1 && b1 && w2 && b2 && gamma && out, FSB_ERR_ARG, "fsb_edgenext_block: NULL pointer");
  FSB_REQUIRE(x != out, FSB_ERR_ARG, "fsb_edgenext_block: out must not alias x (a tile reads its neighbours' inputs)");
  FSB_REQUIRE(dtype == FSB_F16 || dtype == FSB_BF16, FSB_ERR_ARG, "fsb_edgenext_block: dtype %d (need f16/bf16)", dtype);
  FSB_REQUIRE(C == kEbC && F == kEbF, FSB_ERR_SHAPE, "fsb_edgenext_block: C=%d F=%d (built for %d -> %d -> %d)", C, F, kEbC,
              kEbF, kEbC);
  FSB_REQUIRE(B > 0 && H > 0 && W > 0, FSB_ERR_ARG, "fsb_edgenext_block: non-positive size");
  FSB_REQUIRE(aligned16(x) && aligned16(w1) && aligned16(w2) && aligned16(out) && aligned16(dw_b) && aligned16(b1) &&
                  aligned16(b2) && aligned16(gamma),
              FSB_ERR_ALIGN, "fsb_edgenext_block: pointers must be 16-byte aligned");
  EncodeTiledFn enc = encode_fn();
  FSB_REQUIRE(enc != nullptr, FSB_ERR_CUDA, "fsb_edgenext_block: cuTensorMapEncodeTiled not available from the driver");
  EbParams p;
  memset(&p, 0, sizeof(p));
  const CUtensorMapDataType dt = dtype == FSB_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  auto map2d = [&](CUtensorMap* m, const void* ptr, cuuint64_t inner, cuuint64_t rows, const char* what) -> int {
    cuuint64_t dims[2] = {inner, rows};
    cuuint64_t strides[1] = {inner * 2};
    cuuint32_t box[2] = {64, 128};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FSB_REQUIRE(r == CUDA_SUCCESS, FSB_ERR_CUDA, "fsb_edgenext_block: cuTensorMapEncodeTiled(%s) -> %d", what, (int)r);
    return FSB_OK;
  };
  int rc;
  if ((rc = map2d(&p.w1map, w1, kEbC, kEbF, "w1")) != FSB_OK) return rc;
  if ((rc = map2d(&p.w2map, w2, kEbF, kEbC, "w2")) != FSB_OK) return rc;
  p.x = x;
  p.dw_w = dw_w;
  p.dw_b = dw_b;
  p.b1 = b1;
  p.b2 = b2;
  p.gamma = gamma;
  p.out = out;
  p.H = H;
  p.W = W;
  p.tiles_x = ceil_div(W, kEbTX);
  p.tiles_y = ceil_div(H, kEbTY);
  const long long ntiles = (long long)B * p.tiles_x * p.tiles_y;
  FSB_REQUIRE(ntiles < (1ll << 31), FSB_ERR_SHAPE, "fsb_edgenext_block: too many tiles");
  p.ntiles = (int)ntiles;
  const uint32_t fmt = dtype == FSB_F16 ? 0u : 1u;
  p.idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  auto kern = dtype == FSB_F16 ? edgenext_block_kernel<__half> : edgenext_block_kernel<__nv_bfloat16>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kEbSmem);
  if (e != cudaSuccess) return cuda_fail(e, "fsb_edgenext_block: cudaFuncSetAttribute");
  const int grid = p.ntiles < sm_count() ? p.ntiles : sm_count();
  launch_kernel(kern, dim3(grid), dim3(kEbThreads), kEbSmem, (cudaStream_t)stream, p);
  FSB_CHECK_LAUNCH("fsb_edgenext_block");
  return FSB_OK;
}
