// Kernel 4: 3x3 / 1x1 convolution of the update block as a tcgen05 implicit GEMM.
//
// Replaces the nn.Conv2d / nn.Linear calls of BasicSelectiveMultiUpdateBlock
// (core/update.py:20-159, core/submodule.py:565-591) together with the elementwise work
// that follows them (ReLU, GELU, layer scale + residual, and the sigmoid / tanh / blend of
// RaftConvGRU + SelectiveConvGRU, core/update.py:90-95,112-119).
//
// GEMM view:  D[pixel, cout] = sum_{tap, cin} X[pixel + tap, cin] * W[cout, tap, cin]
//   M tile  = 128 pixels = a TH x TW patch of one image (TH*TW = 128)
//   K step  = one tap x 64 input channels: ONE 4-D TMA box (64 ch, TW, TH, 1) of the NHWC
//             activation tensor at the shifted coordinate (x0+dx, y0+dy).  The box lands in
//             shared memory as 128 rows x 128 B with the 128-byte swizzle, which is exactly
//             the K-major operand layout tcgen05.mma wants; out-of-image rows/columns are
//             zero-filled by the TMA unit, which is the convolution's zero padding.  No
//             im2col buffer exists anywhere.
//   N tile  = block_n output channels; weights are pre-packed [rows][Ktot] K-major, one 2-D
//             TMA box (64, block_n) per K step.
//   D       = fp32 accumulators in TMEM (up to two: the 3x3 "large" and 1x1 "small" GRU).
//
// Warp roles (352 threads): warp 0 = TMA producer of the activation tiles, warp 10 = TMA
// producer of the weight tiles (one elected lane each), warp 1 = TMEM allocator + MMA issuer
// (one elected lane), warps 2..9 = epilogue (tcgen05.ld, one pixel per thread, 32 output
// channels per load, two warps per TMEM lane quarter).  A pipeline stage holds `kper` K steps (1, 2 or 4).  Stages are handed
// over with mbarriers: full[s] (both producers -> MMA, transaction bytes), empty[s]
// (tcgen05.commit -> producers), accum (last commit -> epilogue).
//
// Why two producer warps and multi-step stages: measured on B200 (tools/tma_bench2.cu), ONE
// issuing thread sustains one mbarrier round (try_wait + expect_tx) per ~370 cycles plus ~117
// cycles per TMA instruction regardless of the bytes moved, while a 128 x 128 x 64 MMA step
// lasts 256 cycles -- a single producer issuing two loads per 64-wide K step (~600 cycles)
// starves the tensor pipe.  Independent issuers scale linearly up to the L2 limit (~95
// B/cycle/SM).  The single-thread loops below are therefore kept free of integer division
// and parameter-space indexing beyond one word per step.
#include "common.cuh"
#include "umma.cuh"

#include <cuda.h>

namespace fsb {

constexpr int kConvThreads = 352;  // 11 warps
constexpr int kWarpB = 10;           // weight producer warp
constexpr int kMaxSteps = 144;
constexpr int kMaxRows = 512;                // output rows (padded Cout) whose bias a persistent CTA keeps in shared memory
constexpr int kBM = 128;
constexpr int kBK = 64;                      // 16-bit elements per K step (128 B)
constexpr int kATileBytes = kBM * kBK * 2;   // 16 KB

// K-step word: bits 0-1 source, 2-3 dy+1, 4-5 dx+1, 6 accumulator, 8-15 channel chunk of the source, 16-23 the
// 64-wide column block of the weight matrix (= the step's position in tap order), 24 empty slot (row-reuse mode)
constexpr uint32_t kStepNull = 1u << 24;
__host__ __device__ inline uint32_t make_step(int src, int dy, int dx, int acc, int chunk, int wblock) {
  return (uint32_t)src | ((uint32_t)(dy + 1) << 2) | ((uint32_t)(dx + 1) << 4) | ((uint32_t)acc << 6) |
         ((uint32_t)chunk << 8) | ((uint32_t)wblock << 16);
}

struct alignas(64) ConvParams {
  CUtensorMap amap[FSB_MAX_SRC];
  CUtensorMap wmap;
  CUtensorMap wmap_half;  // pair mode: box (64, block_n / 2), each CTA of a pair holds half of the weight tile
  CUtensorMap omap;  // 16-bit NHWC output, box (32 channels, bw, 32/bw, 1): the TMA-store epilogue of the tap kernel
  int ostore;        // 1: the epilogue leaves through shared memory + cp.async.bulk.tensor stores
  CUtensorMap gmap[4];  // gate outputs: GATE_ZR z_l, z_s, rh_l, rh_s (32-channel boxes); GATE_Q [0] h16_out (16 ch), [1] h32_out (16 fp32)
  uint32_t steps[kMaxSteps];
  int nsteps;
  int H, W, tiles_x, tiles_y, tw_log2;
  int BN, nacc, tmem_cols, stages, kper;
  int vr;  // row-reuse mode: a stage = ONE activation box of TH+2 rows serving the three vertical taps (+ the 1x1 segment)
  int nbuf, ntiles, ntiles_n, nbatch;  // persistent tap kernel: TMEM accumulator buffers, work items, N tiles per pixel tile
  int SA, SB;  // halo kernel: activation / weight ring depths
  uint32_t idesc;
  int rows, Cout;
  const float* bias;
  void* out;
  int ld_out, ch_off;
  float* out_f32;
  float out_scale;
  const void* res;
  int ld_res;
  const float* gamma;
  const float* h32;
  const float* att;
  void* z_l;
  void* z_s;
  void* rh_l;
  void* rh_s;
  float* h32_out;
  void* h16_out;
  int hidden;
};

// ------------------------------------------------------------------------------------ epilogue
// One warp-level tensor store of a staged tile: rows = the warp's 32 pixels (lane = row), NU 16-byte units per row
// (4: SWIZZLE_64B, unit u of row r at u ^ ((r >> 1) & 3); 2: no swizzle).  Rows outside the image are clipped by
// the TMA unit.
template <int NU>
__device__ __forceinline__ void stage_and_store(uint8_t* stage, const uint4 (&units)[NU], const CUtensorMap* map, int c0,
                                                int wx0, int wy0, int b) {
  const int lane = threadIdx.x & 31;
  uint8_t* row = stage + lane * (NU * 16);
  const int sw = NU == 4 ? ((lane >> 1) & 3) : 0;
#pragma unroll
  for (int u = 0; u < NU; ++u) *reinterpret_cast<uint4*>(row + ((u ^ sw) << 4)) = units[u];
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  if (lane == 0) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 ::"l"(map), "r"(smem_u32(stage)), "r"(c0), "r"(wx0), "r"(wy0), "r"(b)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
}
// before a warp rewrites its store tiles: the tensor stores issued so far have read them
__device__ __forceinline__ void store_tiles_free() {
  if ((threadIdx.x & 31) == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  __syncwarp();
}

// One thread = one pixel (TMEM lane); `taddr` = lane quarter + first column of the pixel tile's
// accumulators; the thread handles the 32-column (16 for GATE_Q) chunks c_first, c_first + c_stride, ...
template <typename T, int EPI>
__device__ __forceinline__ void conv_epilogue(const ConvParams& p, const float* bias_s, uint32_t taddr, bool valid,
                                              size_t pix, int b, int y, int x, int n0, int nt, int c_first,
                                              int c_stride, int bstride, uint32_t wait_bar = 0,
                                              uint32_t wait_parity = 0, uint8_t* stage = nullptr) {
  // wait_bar != 0: the caller has NOT yet waited for the accumulators; the epilogue does it itself, after it has
  // issued whatever global loads do not depend on them
  if constexpr (EPI != FSB_EPI_GATE_Q) {
    if (wait_bar) {
      mbar_wait(wait_bar, wait_parity);
      tc_fence_after();
    }
  }
  if constexpr (EPI == FSB_EPI_GATE_ZR) {
    const int half = p.BN >> 1;
    const int cpa = p.BN >> 5;  // chunks per accumulator
    for (int ci = c_first; ci < p.nacc * cpa; ci += c_stride) {
      const int a = ci / cpa, c = (ci - a * cpa) << 5;
      T* zb = reinterpret_cast<T*>(a == 0 ? p.z_l : p.z_s);
      T* rb = reinterpret_cast<T*>(a == 0 ? p.rh_l : p.rh_s);
      const float* bias = bias_s + a * bstride;
      {
        uint32_t v[32];
        tmem_ld32(taddr + a * p.BN + c, v);
        tmem_ld_wait();
        const bool gstore = stage != nullptr && p.ostore;
        if (!valid && !gstore) continue;
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = sigmoid_f(__uint_as_float(v[j]) + bias[c + j]);
        const bool is_z = c < half;
        const int ch = nt * half + (is_z ? c : c - half);
        if (!is_z) {
          float h[32];
          load32f(p.h32 + pix * p.hidden + ch, h);  // pix = 0 for rows outside the image: a valid address, result clipped
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] *= h[j];
        }
        if (gstore) {
          store_tiles_free();
          uint4 units[4];
#pragma unroll
          for (int u = 0; u < 4; ++u)
            units[u] = make_uint4(pack2<T>(f[8 * u], f[8 * u + 1]), pack2<T>(f[8 * u + 2], f[8 * u + 3]),
                                  pack2<T>(f[8 * u + 4], f[8 * u + 5]), pack2<T>(f[8 * u + 6], f[8 * u + 7]));
          const int wx0 = __shfl_sync(0xffffffffu, x, 0), wy0 = __shfl_sync(0xffffffffu, y, 0);
          stage_and_store<4>(stage, units, &p.gmap[(is_z ? 0 : 2) + a], ch, wx0, wy0, b);
        } else {
          store32<T>((is_z ? zb : rb) + pix * p.hidden + ch, f);
        }
      }
    }
  } else if constexpr (EPI == FSB_EPI_GATE_Q) {
    // 16 columns per step.  The hidden state and the z gates of a step do not depend on the accumulators: they
    // are loaded one step ahead (the first step before the accumulator barrier), so their L2 latency -- long
    // while the TMA producers saturate L2 -- hides under the previous step's arithmetic instead of adding up
    // eight times per tile (measured: 138 -> 97 us for gru04.q at 4 pairs with the loads removed).
    const float att = valid ? __ldg(p.att + pix) : 0.f;
    const float* hbase = p.h32 + pix * p.hidden + n0;
    const T* zlbase = reinterpret_cast<const T*>(p.z_l) + pix * p.hidden + n0;
    const T* zsbase = reinterpret_cast<const T*>(p.z_s) + pix * p.hidden + n0;
    float4 hq[4];
    uint4 zlq[2], zsq[2];
    auto load_step = [&](int c) {
      if (valid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) hq[i] = *reinterpret_cast<const float4*>(hbase + c + 4 * i);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          zlq[i] = *reinterpret_cast<const uint4*>(zlbase + c + 8 * i);
          zsq[i] = *reinterpret_cast<const uint4*>(zsbase + c + 8 * i);
        }
      }
    };
    load_step(c_first * 16);
    if (wait_bar) {
      mbar_wait(wait_bar, wait_parity);
      tc_fence_after();
    }
    for (int c = c_first * 16; c < p.BN; c += 16 * c_stride) {
      uint32_t v0[16], v1[16];
      tmem_ld16(taddr + c, v0);
      tmem_ld16(taddr + p.BN + c, v1);
      float h[16];
      uint32_t zlw[8], zsw[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) { h[4 * i] = hq[i].x; h[4 * i + 1] = hq[i].y; h[4 * i + 2] = hq[i].z; h[4 * i + 3] = hq[i].w; }
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        zlw[4 * i] = zlq[i].x; zlw[4 * i + 1] = zlq[i].y; zlw[4 * i + 2] = zlq[i].z; zlw[4 * i + 3] = zlq[i].w;
        zsw[4 * i] = zsq[i].x; zsw[4 * i + 1] = zsq[i].y; zsw[4 * i + 2] = zsq[i].z; zsw[4 * i + 3] = zsq[i].w;
      }
      if (c + 16 * c_stride < p.BN) load_step(c + 16 * c_stride);  // next step's operands, in flight during this one
      tmem_ld_wait();
      const bool gstore = stage != nullptr && p.ostore;
      if (!valid && !gstore) continue;
      const int ch = n0 + c;
      float* ho = p.h32_out + pix * p.hidden + ch;
      uint4* h16o = reinterpret_cast<uint4*>(reinterpret_cast<T*>(p.h16_out) + pix * p.hidden + ch);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const float2 zl2 = unpack2<T>(zlw[j >> 1]), zs2 = unpack2<T>(zsw[j >> 1]);
        const float zl = (j & 1) ? zl2.y : zl2.x, zs = (j & 1) ? zs2.y : zs2.x;
        const float ql = tanh_f(__uint_as_float(v0[j]) + bias_s[c + j]);
        const float qs = tanh_f(__uint_as_float(v1[j]) + bias_s[bstride + c + j]);
        const float hl = (1.f - zl) * h[j] + zl * ql;
        const float hs = (1.f - zs) * h[j] + zs * qs;
        h[j] = hs * att + hl * (1.f - att);
      }
      if (gstore) {
        store_tiles_free();
        uint4 u16[2], u32[4];
#pragma unroll
        for (int o = 0; o < 2; ++o)
          u16[o] = make_uint4(pack2<T>(h[8 * o], h[8 * o + 1]), pack2<T>(h[8 * o + 2], h[8 * o + 3]),
                              pack2<T>(h[8 * o + 4], h[8 * o + 5]), pack2<T>(h[8 * o + 6], h[8 * o + 7]));
#pragma unroll
        for (int o = 0; o < 4; ++o)
          u32[o] = make_uint4(__float_as_uint(h[4 * o]), __float_as_uint(h[4 * o + 1]), __float_as_uint(h[4 * o + 2]),
                              __float_as_uint(h[4 * o + 3]));
        const int wx0 = __shfl_sync(0xffffffffu, x, 0), wy0 = __shfl_sync(0xffffffffu, y, 0);
        stage_and_store<4>(stage, u32, &p.gmap[1], ch, wx0, wy0, b);
        stage_and_store<2>(stage + 2048, u16, &p.gmap[0], ch, wx0, wy0, b);
        continue;
      }
#pragma unroll
      for (int o = 0; o < 2; ++o) {
        *reinterpret_cast<float4*>(ho + 8 * o) = make_float4(h[8 * o], h[8 * o + 1], h[8 * o + 2], h[8 * o + 3]);
        *reinterpret_cast<float4*>(ho + 8 * o + 4) = make_float4(h[8 * o + 4], h[8 * o + 5], h[8 * o + 6], h[8 * o + 7]);
        uint4 u;
        u.x = pack2<T>(h[8 * o], h[8 * o + 1]); u.y = pack2<T>(h[8 * o + 2], h[8 * o + 3]);
        u.z = pack2<T>(h[8 * o + 4], h[8 * o + 5]); u.w = pack2<T>(h[8 * o + 6], h[8 * o + 7]);
        h16o[o] = u;
      }
    }
  } else {
    T* out = reinterpret_cast<T*>(p.out);
    for (int c = c_first * 32; c < p.BN; c += 32 * c_stride) {
      uint32_t v[32];
      tmem_ld32(taddr + c, v);
      tmem_ld_wait();
      const int n = n0 + c;
      const bool full = (n + 32 <= p.Cout);
      if (stage != nullptr && full) {
        // Coalesced path: the warp's 32 pixels x 32 channels go to a 2 KB shared tile (SWIZZLE_64B: 16-byte unit
        // u of row r at u ^ ((r >> 1) & 3), conflict free) and leave as ONE tensor store; rows outside the image
        // are clipped by the TMA unit.  Per-lane stores of 64 B at pixel stride touched 32 lines per instruction
        // and sustained ~1.6 TB/s (tools/conv_sweep.py: time of the 1x1 layers ~ 0.6 us per MB of output).
        store_tiles_free();
        uint4 units[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          uint32_t w[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int e = 8 * u + 2 * j;
            float a0 = __uint_as_float(v[e]) + bias_s[c + e], a1 = __uint_as_float(v[e + 1]) + bias_s[c + e + 1];
            if constexpr (EPI == FSB_EPI_RELU) { a0 = fmaxf(a0, 0.f); a1 = fmaxf(a1, 0.f); }
            if constexpr (EPI == FSB_EPI_GELU) { a0 = gelu_f(a0); a1 = gelu_f(a1); }
            w[j] = pack2<T>(a0, a1);
          }
          units[u] = make_uint4(w[0], w[1], w[2], w[3]);
        }
        const int wx0 = __shfl_sync(0xffffffffu, x, 0), wy0 = __shfl_sync(0xffffffffu, y, 0);  // lane 0 = the warp's first pixel
        stage_and_store<4>(stage, units, &p.omap, p.ch_off + n, wx0, wy0, b);
        continue;
      }
      if (!valid || n >= p.Cout) continue;
      float f[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float a = __uint_as_float(v[j]) + bias_s[c + j];  // rows beyond Cout carry zero bias
        if constexpr (EPI == FSB_EPI_RELU) a = fmaxf(a, 0.f);
        if constexpr (EPI == FSB_EPI_GELU) a = gelu_f(a);
        f[j] = a;
      }
      if constexpr (EPI == FSB_EPI_SCALE_RES) {
        float rres[32];
        load32<T>(reinterpret_cast<const T*>(p.res) + pix * p.ld_res + n, rres);  // Cout % 32 == 0 (host check)
#pragma unroll
        for (int j = 0; j < 32; ++j) f[j] = rres[j] + __ldg(p.gamma + n + j) * f[j];
      }
      if (out != nullptr) {
        T* o = out + pix * p.ld_out + p.ch_off + n;
        if (full) {
          store32<T>(o, f);
        } else {
          // ragged last chunk (e.g. the 127 motion channels): whole 16-byte vectors first, scalars for the rest
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (n + 8 * q + 8 <= p.Cout) {
              uint4 u;
              u.x = pack2<T>(f[8 * q], f[8 * q + 1]); u.y = pack2<T>(f[8 * q + 2], f[8 * q + 3]);
              u.z = pack2<T>(f[8 * q + 4], f[8 * q + 5]); u.w = pack2<T>(f[8 * q + 6], f[8 * q + 7]);
              reinterpret_cast<uint4*>(o)[q] = u;
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (n + 8 * q + j < p.Cout) o[8 * q + j] = from_f32<T>(f[8 * q + j]);
            }
          }
        }
      }
      if (p.out_f32 != nullptr) {
        const size_t hw = (size_t)p.H * p.W;
        float* o = p.out_f32 + ((size_t)b * p.Cout + n) * hw + (size_t)y * p.W + x;
#pragma unroll 4
        for (int j = 0; j < 32; ++j)
          if (full || n + j < p.Cout) o[(size_t)j * hw] = f[j] * p.out_scale;
      }
    }
  }
}

// ------------------------------------------------------------------------------------ kernel
// PERSISTENT: one CTA per SM walks the work list  tile = blockIdx.x, blockIdx.x + gridDim.x, ...  with
// tile = m_tile * ntiles_n + n_tile (N fastest: the CTAs that run side by side read the same activation
// boxes, which then come from L2 instead of HBM).  The shared-memory ring, its phases and the two
// producers run straight through tile boundaries, and the accumulators are double-buffered in TMEM
// (2 x nacc x block_n columns): while the eight epilogue warps drain buffer i & 1 the MMA warp is
// already accumulating tile i + 1 into the other one.  Measured before this change (tools/conv_sweep.py):
// a CTA that lives for one small-K tile spends ~15k cycles on launch, TMEM allocation, pipeline fill and
// a serial epilogue around 512 cycles of MMA, and the GRU q gate lost 45 % of its time to an epilogue
// nothing overlapped.
// CL == 2 ("pair mode"): a cluster of two CTAs on two adjacent pixel tiles and the SAME N tile runs ONE
// tcgen05.mma.cta_group::2 stream (M = 256): each CTA stages its own 128-pixel activation tile and HALF of the
// weight tile (rows [r*bn/2, (r+1)*bn/2)), the leader CTA (rank 0) issues the MMAs for both, each CTA's
// accumulators live in its own TMEM and each CTA runs its own epilogue.  What this buys is bytes: a tap-layout CTA
// is bound by what an SM can ingest from L2 (~67 B/cycle, section 6 of DESIGN.md) and in pair mode an SM takes in
// 16 KB + bn*64 B per K step instead of 16 KB + bn*128 B.  (Multicasting the weight tile to two independent
// CTAs does not help: each SM still ingests the whole tile.)
// Barriers: full[s] lives in the leader (its two producers arrive with the bytes of BOTH CTAs; every TMA load of
// either CTA completes on it: .cta_group::2 loads with the peer bit of the barrier address cleared); empty[s],
// tfull[b] are signalled in both CTAs by multicast commits of the leader; tempty[b] of the leader collects the
// epilogue warps of both CTAs (the peer arrives remotely).
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;  // shared::cluster address of the same object in the even CTA of a pair
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_4d_2sm(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar & kPeerBitMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_f16_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint32_t bar) {  // arrives on `bar` in BOTH CTAs of the pair
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {  // arrive on the leader CTA's copy of `bar`
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, 0;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}"
      ::"r"(bar) : "memory");
}

template <typename T, int EPI, int CL>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_umma_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * 8 + 4];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float bias_s[2 * kMaxRows];  // [nacc][rows]

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);  // warp-uniform for the compiler
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int kper = p.kper;
  const uint32_t b_tile_bytes = (uint32_t)(p.BN / CL) * (kBK * 2);  // this CTA's part of a weight tile
  const uint32_t a_row_bytes = (uint32_t)(kBK * 2) << p.tw_log2;  // one image row of the pixel tile
  const uint32_t a_stage_bytes = p.vr ? a_row_bytes * (uint32_t)((kBM >> p.tw_log2) + 2) : (uint32_t)kper * kATileBytes;
  const uint32_t stage_bytes = a_stage_bytes + (uint32_t)kper * b_tile_bytes;
  const int S = p.stages;
  const uint32_t bar_full = smem_u32(&bars[0]);
  const uint32_t bar_empty = smem_u32(&bars[8]);
  const uint32_t bar_tfull = smem_u32(&bars[16]);   // [2] accumulator buffer complete (MMA -> epilogue)
  const uint32_t bar_tempty = smem_u32(&bars[18]);  // [2] accumulator buffer drained (epilogue -> MMA)
  const int nbuf = p.nbuf;
  const uint32_t buf_cols = (uint32_t)(p.nacc * p.BN);
  const int TW = 1 << p.tw_log2, TH = kBM >> p.tw_log2;
  const int ntn = p.ntiles_n;
  // work list: CL == 1: (pixel tile, N tile); CL == 2: (PAIR of pixel tiles, N tile), CTA r of the cluster takes pixel
  // tile 2*pair + r (one past the end for an odd count: all-zero operands, nothing stored)
  const int crank = CL == 2 ? (int)cluster_ctarank() : 0;
  const bool leader = crank == 0;
  const int wfirst = CL == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int wstep = CL == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int nwork = p.ntiles;
  const int mtiles = p.tiles_x * p.tiles_y * p.nbatch;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(bar_full + 8 * s, 2);   // one expect_tx arrival from each producer
      mbar_init(bar_empty + 8 * s, 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_tfull + 8 * i, 1);
      mbar_init(bar_tempty + 8 * i, 8 * CL);  // one arrival per epilogue warp (of both CTAs in pair mode)
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&p.amap[0]);
  }
  if (warp == kWarpB && lane == 0) tma_prefetch_desc(&p.wmap);
  if (warp == 1) {
    if (CL == 2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                   "r"((uint32_t)p.tmem_cols)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                   "r"((uint32_t)p.tmem_cols)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  if (warp >= 2 && warp < kWarpB) {
    for (int i = threadIdx.x - 64; i < p.nacc * p.rows; i += 256) bias_s[i] = __ldg(p.bias + i);
  }
  tc_fence_before();
  __syncthreads();
  if (CL == 2) cluster_sync_all();  // both CTAs' barriers and TMEM exist before either touches the other
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  // PDL: the prologue above (and the weight producer below, which only reads parameters that no kernel
  // writes) overlaps the tail of the previous kernel; everything that touches activations waits here.
  pdl_launch_dependents();
  if (warp != kWarpB) pdl_wait();

  if (warp == 0) {
    // ================================================== TMA producer: activation tiles
    int s = 0;
    uint32_t ph = 0;
    bool first_round = true;
    for (int wt = wfirst; wt < nwork; wt += wstep) {
      const int mtile = CL == 2 ? 2 * (wt / ntn) + crank : wt / ntn;
      const int txi = mtile % p.tiles_x, r1 = mtile / p.tiles_x;
      const int tyi = r1 % p.tiles_y, b = r1 / p.tiles_y;  // b == nbatch for the pad tile of an odd count: zero fill
      const int x0 = txi * TW, y0 = tyi * TH;
      for (int i = 0; i < p.nsteps; i += kper) {
        const int n = min(kper, p.nsteps - i);
        if (!first_round) mbar_wait(bar_empty + 8 * s, ph ^ 1);
        __syncwarp();
        if (elect_one()) {
          const uint32_t full = bar_full + 8 * s;
          uint32_t dst = smem_base + s * stage_bytes;
          if (p.vr) {
            // one box of TH + 2 image rows: the K steps of this stage are its three vertical taps
            if (leader) mbar_expect_tx(full, (uint32_t)CL * a_stage_bytes);
            const uint32_t st = p.steps[i];
            const int src = st & 3, dx = (int)((st >> 4) & 3) - 1, chunk = (st >> 8) & 0xff;
            if (CL == 2) tma_load_4d_2sm(dst, &p.amap[src], full, chunk * kBK, x0 + dx, y0 - 1, b);
            else         tma_load_4d(dst, &p.amap[src], full, chunk * kBK, x0 + dx, y0 - 1, b);
          } else {
          if (leader) mbar_expect_tx(full, (uint32_t)(CL * n) * kATileBytes);  // pair mode: the peer's tiles complete here too
          for (int j = 0; j < n; ++j) {
            const uint32_t st = p.steps[i + j];
            const int src = st & 3, dy = (int)((st >> 2) & 3) - 1, dx = (int)((st >> 4) & 3) - 1, chunk = (st >> 8) & 0xff;
            if (CL == 2) tma_load_4d_2sm(dst, &p.amap[src], full, chunk * kBK, x0 + dx, y0 + dy, b);
            else         tma_load_4d(dst, &p.amap[src], full, chunk * kBK, x0 + dx, y0 + dy, b);
            dst += kATileBytes;
          }
          }
        }
        if (++s == S) { s = 0; ph ^= 1; first_round = false; }
      }
    }
  } else if (warp == kWarpB) {
    // ====================================================== TMA producer: weight tiles
    int s = 0;
    uint32_t ph = 0;
    bool first_round = true;
    for (int wt = wfirst; wt < nwork; wt += wstep) {
      const int n0 = (wt % ntn) * p.BN;
      for (int i = 0; i < p.nsteps; i += kper) {
        const int n = min(kper, p.nsteps - i);
        if (!first_round) mbar_wait(bar_empty + 8 * s, ph ^ 1);
        __syncwarp();
        if (elect_one()) {
          const uint32_t full = bar_full + 8 * s;
          uint32_t dst = smem_base + s * stage_bytes + a_stage_bytes;
          int nload = 0;
          for (int j = 0; j < n; ++j) nload += (p.steps[i + j] & kStepNull) ? 0 : 1;
          if (leader) mbar_expect_tx(full, (uint32_t)(CL * nload) * b_tile_bytes);
          for (int j = 0; j < n; ++j) {
            const uint32_t st = p.steps[i + j];
            if (!(st & kStepNull)) {
              const int k0 = (int)((st >> 16) & 0xff) * kBK;
              if (CL == 2) tma_load_2d_2sm(dst, &p.wmap_half, full, k0, n0 + crank * (p.BN >> 1));
              else         tma_load_2d(dst, &p.wmap, full, k0, n0);
            }
            dst += b_tile_bytes;
          }
        }
        if (++s == S) { s = 0; ph ^= 1; first_round = false; }
      }
    }
  } else if (warp == 1 && leader) {
    // ====================================================================== MMA issuer (pair mode: the leader CTA only)
    int s = 0;
    uint32_t ph = 0;
    const uint32_t idesc = p.idesc;
    int it = 0;
    for (int wt = wfirst; wt < nwork; wt += wstep, ++it) {
      const int buf = nbuf == 2 ? (it & 1) : 0;
      const uint32_t use = (uint32_t)(nbuf == 2 ? (it >> 1) : it);  // how often this buffer was used before
      mbar_wait(bar_tempty + 8 * buf, (use & 1) ^ 1);                // drained by the epilogue (passes at first use)
      tc_fence_after();
      const uint32_t d_base = tmem_base + (uint32_t)buf * buf_cols;
      uint32_t started = 0;  // bit a: accumulator a already holds a partial sum
      for (int i = 0; i < p.nsteps; i += kper) {
        const int n = min(kper, p.nsteps - i);
        mbar_wait(bar_full + 8 * s, ph);
        tc_fence_after();
        __syncwarp();
        const uint32_t a_addr = smem_base + s * stage_bytes;
        uint64_t adesc = umma_desc_k128(a_addr);
        uint64_t bdesc = umma_desc_k128(a_addr + a_stage_bytes);
        const uint64_t adesc0 = adesc;
        for (int j = 0; j < n; ++j) {
          const uint32_t st = p.steps[i + j];
          if (st & kStepNull) {  // row-reuse mode: slot without a K step
            bdesc += b_tile_bytes >> 4;
            continue;
          }
          if (p.vr) adesc = adesc0 + (uint64_t)((((st >> 2) & 3) * a_row_bytes) >> 4);  // the tap's window of the box
          const int acc = (st >> 6) & 1;
          const uint32_t d = d_base + (uint32_t)(acc * p.BN);
          const uint32_t first = (started >> acc) & 1;
          if (elect_one()) {
            // +32 B along K inside the swizzle atom = +2 in the (addr >> 4) field
            if (CL == 2) {
              umma_f16_2sm(d, adesc, bdesc, idesc, first);
              umma_f16_2sm(d, adesc + 2, bdesc + 2, idesc, 1);
              umma_f16_2sm(d, adesc + 4, bdesc + 4, idesc, 1);
              umma_f16_2sm(d, adesc + 6, bdesc + 6, idesc, 1);
            } else {
              umma_f16(d, adesc, bdesc, idesc, first);
              umma_f16(d, adesc + 2, bdesc + 2, idesc, 1);
              umma_f16(d, adesc + 4, bdesc + 4, idesc, 1);
              umma_f16(d, adesc + 6, bdesc + 6, idesc, 1);
            }
          }
          started |= 1u << acc;
          adesc += kATileBytes >> 4;
          bdesc += b_tile_bytes >> 4;
        }
        if (elect_one()) {  // frees the stage (of both CTAs) when these MMAs have read it
          if (CL == 2) umma_commit_2sm(bar_empty + 8 * s);
          else         umma_commit(bar_empty + 8 * s);
        }
        if (++s == S) { s = 0; ph ^= 1; }
      }
      if (elect_one()) {  // all MMAs of the tile done -> accumulators complete (in both CTAs)
        if (CL == 2) umma_commit_2sm(bar_tfull + 8 * buf);
        else         umma_commit(bar_tfull + 8 * buf);
      }
    }
  } else if (warp >= 2 && warp < kWarpB) {
    // ======================================================================== epilogue
    // 8 warps: warp w reads TMEM lane quarter (w & 3); the two warps of a quarter split the
    // 32-column chunks between them (grp 0: even chunks, grp 1: odd chunks), which halves the
    // serial instruction chain of the epilogue -- the whole cost of the small-K layers.
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int grp = (warp - 2) >> 2;
    const int r = q * 32 + lane;
    const int ty = r >> p.tw_log2, tx = r & (TW - 1);
    // 2 KB store tile of this warp, behind the operand ring
    uint8_t* stage = p.ostore ? smem_raw + (smem_base - smem_u32(smem_raw)) + (size_t)S * stage_bytes + (size_t)(warp - 2) * 3072
                              : nullptr;
    int it = 0;
    for (int wt = wfirst; wt < nwork; wt += wstep, ++it) {
      const int nt = wt % ntn;
      const int mtile = CL == 2 ? 2 * (wt / ntn) + crank : wt / ntn;
      const int txi = mtile % p.tiles_x, r1 = mtile / p.tiles_x;
      const int tyi = r1 % p.tiles_y, b = r1 / p.tiles_y;
      const int y = tyi * TH + ty, x = txi * TW + tx;
      const bool valid = (y < p.H) && (x < p.W) && (mtile < mtiles);
      const size_t pix = valid ? ((size_t)b * p.H + y) * p.W + x : 0;
      const int buf = nbuf == 2 ? (it & 1) : 0;
      const uint32_t use = (uint32_t)(nbuf == 2 ? (it >> 1) : it);
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)buf * buf_cols;
      const int n0 = nt * p.BN;

      conv_epilogue<T, EPI>(p, bias_s + n0, taddr, valid, pix, b, y, x, n0, nt, grp, 2, p.rows, bar_tfull + 8 * buf, use & 1,
                            stage);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CL == 2) mbar_arrive_leader(bar_tempty + 8 * buf);
        else         mbar_arrive(bar_tempty + 8 * buf);
      }
    }
    if (p.ostore && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // stores done before smem goes away
  }
  __syncwarp();
  __syncthreads();
  if (CL == 2) cluster_sync_all();  // no CTA leaves (or frees TMEM) while its peer may still use it
  if (warp == 1) {
    tc_fence_after();
    if (CL == 2)
      asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)p.tmem_cols) : "memory");
    else
      asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)p.tmem_cols) : "memory");
  }
}

// ------------------------------------------------------------------------ kernel, halo variant
// Same GEMM, different operand feed (FSB_CONV_LAYOUT_HALO).  A CTA owns TWO vertically stacked
// 128-pixel tiles (32 rows x 8 columns of the image) and, per 64-channel chunk and per horizontal
// tap dx, loads ONE activation box of (32 + 2) rows x 8 pixels x 64 channels (34 KB) instead of
// one 16 KB box per tap and tile: a tile row of 8 pixels is exactly one 1024-byte swizzle atom,
// so the vertical taps dy = -1, 0, +1 of tile mt are the SAME box read through UMMA descriptors
// that start (mt*16 + 1 + dy) atoms into it -- all 1024-byte aligned, no re-load.  Activation
// traffic per chunk drops from 2 x 9 x 16 KB to 3 x 34 KB (2.8x), and every weight tile feeds 256
// pixels instead of 128 (weight traffic per FLOP halves).  K order of the packed weights:
// segment -> source -> 64-channel chunk -> dx -> dy.  Activation and weight tiles travel through
// separate rings (an activation slot lives for three weight steps).
// Step word: bits 0-1 source, 2-3 row offset 1+dy, 4-5 dx+1, 6 accumulator, 7 first step of an
// activation slot, 8-15 channel chunk, 16 last step of the slot.
constexpr int kHaloRows = 34;
constexpr int kASlotBytes = kHaloRows * 8 * 128;  // 34816

__host__ __device__ inline uint32_t make_halo_step(int src, int dyrow, int dx, int acc, int chunk, int first, int last) {
  return (uint32_t)src | ((uint32_t)dyrow << 2) | ((uint32_t)(dx + 1) << 4) | ((uint32_t)acc << 6) |
         ((uint32_t)first << 7) | ((uint32_t)chunk << 8) | ((uint32_t)last << 16);
}

template <typename T, int EPI>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_halo_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[4 + 4 + 8 + 8 + 1];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float bias_s[2 * 256];

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);  // warp-uniform for the compiler
  const int lane = threadIdx.x & 31;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int SA = p.SA, SB = p.SB;
  const uint32_t b_bytes = (uint32_t)p.BN * (kBK * 2);
  const uint32_t ring_b = smem_base + (uint32_t)SA * kASlotBytes;
  const uint32_t fullA = smem_u32(&bars[0]), emptyA = smem_u32(&bars[4]);
  const uint32_t fullB = smem_u32(&bars[8]), emptyB = smem_u32(&bars[16]);
  const uint32_t bar_accum = smem_u32(&bars[24]);

  int t = blockIdx.x;
  const int txi = t % p.tiles_x;
  t /= p.tiles_x;
  const int tyi = t % p.tiles_y;
  const int b = t / p.tiles_y;
  const int x0 = txi * 8, y0 = tyi * 32;
  const int nt = blockIdx.y;
  const int n0 = nt * p.BN;
  const int nmt = (y0 + 16 < p.H) ? 2 : 1;  // the lower tile may lie entirely below the image

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < SA; ++i) { mbar_init(fullA + 8 * i, 1); mbar_init(emptyA + 8 * i, 1); }
    for (int i = 0; i < SB; ++i) { mbar_init(fullB + 8 * i, 1); mbar_init(emptyB + 8 * i, 1); }
    mbar_init(bar_accum, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&p.amap[0]);
  }
  if (warp == kWarpB && lane == 0) tma_prefetch_desc(&p.wmap);
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "r"((uint32_t)p.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp >= 2 && warp < kWarpB) {
    for (int i = threadIdx.x - 64; i < p.nacc * p.BN; i += 256) {
      const int a = i / p.BN, c = i - a * p.BN;
      bias_s[i] = __ldg(p.bias + (size_t)a * p.rows + n0 + c);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  // PDL: the prologue above (and the weight producer below, which only reads parameters that no kernel
  // writes) overlaps the tail of the previous kernel; everything that touches activations waits here.
  pdl_launch_dependents();
  if (warp != kWarpB) pdl_wait();

  if (warp == 0) {
    // ================================================== TMA producer: activation boxes
    int sa = 0, nslot = 0;
    uint32_t ph = 0;
    for (int i = 0; i < p.nsteps; ++i) {
      const uint32_t st = p.steps[i];
      if (!((st >> 7) & 1)) continue;
      if (nslot >= SA) mbar_wait(emptyA + 8 * sa, ph ^ 1);
      __syncwarp();
      if (elect_one()) {
        const int src = st & 3, dx = (int)((st >> 4) & 3) - 1, chunk = (st >> 8) & 0xff;
        mbar_expect_tx(fullA + 8 * sa, kASlotBytes);
        tma_load_4d(smem_base + sa * kASlotBytes, &p.amap[src], fullA + 8 * sa, chunk * kBK, x0 + dx, y0 - 1, b);
      }
      ++nslot;
      if (++sa == SA) { sa = 0; ph ^= 1; }
    }
  } else if (warp == kWarpB) {
    // ====================================================== TMA producer: weight tiles
    int sb = 0;
    uint32_t ph = 0;
    for (int i = 0; i < p.nsteps; ++i) {
      if (i >= SB) mbar_wait(emptyB + 8 * sb, ph ^ 1);
      __syncwarp();
      if (elect_one()) {
        mbar_expect_tx(fullB + 8 * sb, b_bytes);
        tma_load_2d(ring_b + sb * b_bytes, &p.wmap, fullB + 8 * sb, i * kBK, n0);
      }
      if (++sb == SB) { sb = 0; ph ^= 1; }
    }
  } else if (warp == 1) {
    // ====================================================================== MMA issuer
    uint32_t started = 0;  // bit (mt*2 + acc)
    int sa = 0, sb = 0;
    uint32_t pha = 0, phb = 0, a_base = smem_base;
    const uint32_t idesc = p.idesc;
    const uint32_t mt_cols = (uint32_t)(p.nacc * p.BN);
    for (int i = 0; i < p.nsteps; ++i) {
      const uint32_t st = p.steps[i];
      if ((st >> 7) & 1) {
        mbar_wait(fullA + 8 * sa, pha);
        a_base = smem_base + sa * kASlotBytes;
      }
      mbar_wait(fullB + 8 * sb, phb);
      tc_fence_after();
      __syncwarp();
      const int acc = (st >> 6) & 1, dyrow = (st >> 2) & 3;
      const uint64_t bdesc = umma_desc_k128(ring_b + sb * b_bytes);
      for (int mt = 0; mt < nmt; ++mt) {
        const uint64_t adesc = umma_desc_k128(a_base + (uint32_t)(mt * 16 + dyrow) * 1024u);
        const uint32_t d = tmem_base + mt * mt_cols + (uint32_t)(acc * p.BN);
        const uint32_t bit = 1u << (mt * 2 + acc);
        const uint32_t first = (started & bit) ? 1u : 0u;
        if (elect_one()) {
          umma_f16(d, adesc, bdesc, idesc, first);
          umma_f16(d, adesc + 2, bdesc + 2, idesc, 1);
          umma_f16(d, adesc + 4, bdesc + 4, idesc, 1);
          umma_f16(d, adesc + 6, bdesc + 6, idesc, 1);
        }
        started |= bit;
      }
      if (elect_one()) umma_commit(emptyB + 8 * sb);
      if (++sb == SB) { sb = 0; phb ^= 1; }
      if ((st >> 16) & 1) {
        if (elect_one()) umma_commit(emptyA + 8 * sa);
        if (++sa == SA) { sa = 0; pha ^= 1; }
      }
    }
    if (elect_one()) umma_commit(bar_accum);
  } else if (warp < kWarpB) {
    // ======================================================================== epilogue
    // warps 2-5: upper tile, warps 6-9: lower tile; warp w reads TMEM lane quarter (w & 3)
    const int q = warp & 3, mt = (warp - 2) >> 2;
    const int r = q * 32 + lane;
    const int y = y0 + mt * 16 + (r >> 3), x = x0 + (r & 7);
    const bool valid = (y < p.H) && (x < p.W);
    const size_t pix = valid ? ((size_t)b * p.H + y) * p.W + x : 0;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(mt * p.nacc * p.BN);
    mbar_wait(bar_accum, 0);
    tc_fence_after();
    if (mt < nmt) conv_epilogue<T, EPI>(p, bias_s, taddr, valid, pix, b, y, x, n0, nt, 0, 1, p.BN);
    tc_fence_before();
  }
  __syncwarp();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)p.tmem_cols)
                 : "memory");
  }
}

// ------------------------------------------------------------------------------------- host
static int pick_tile(int H, int W, int* tw_log2) {
  // TH x TW = 128, both powers of two: fewest tiles, ties -> wider rows (longer contiguous runs)
  long best = -1;
  int best_l = 4;
  for (int l = 7; l >= 2; --l) {
    const int tw = 1 << l, th = kBM >> l;
    const long n = (long)ceil_div(W, tw) * ceil_div(H, th);
    if (best < 0 || n < best) {
      best = n;
      best_l = l;
    }
  }
  *tw_log2 = best_l;
  return (int)best;
}

template <typename T, int CL>
static auto kernel_for_cl(int epi) -> void (*)(const ConvParams) {
  switch (epi) {
    case FSB_EPI_BIAS: return conv_umma_kernel<T, FSB_EPI_BIAS, CL>;
    case FSB_EPI_RELU: return conv_umma_kernel<T, FSB_EPI_RELU, CL>;
    case FSB_EPI_GATE_ZR: return conv_umma_kernel<T, FSB_EPI_GATE_ZR, CL>;
    case FSB_EPI_GATE_Q: return conv_umma_kernel<T, FSB_EPI_GATE_Q, CL>;
    case FSB_EPI_GELU: return conv_umma_kernel<T, FSB_EPI_GELU, CL>;
    case FSB_EPI_SCALE_RES: return conv_umma_kernel<T, FSB_EPI_SCALE_RES, CL>;
    default: return nullptr;
  }
}
template <typename T>
static auto kernel_for(int epi, int cl) -> void (*)(const ConvParams) {
  return cl == 2 ? kernel_for_cl<T, 2>(epi) : kernel_for_cl<T, 1>(epi);
}

template <typename T>
static auto halo_kernel_for(int epi) -> void (*)(const ConvParams) {
  switch (epi) {
    case FSB_EPI_BIAS: return conv_halo_kernel<T, FSB_EPI_BIAS>;
    case FSB_EPI_RELU: return conv_halo_kernel<T, FSB_EPI_RELU>;
    case FSB_EPI_GATE_ZR: return conv_halo_kernel<T, FSB_EPI_GATE_ZR>;
    case FSB_EPI_GATE_Q: return conv_halo_kernel<T, FSB_EPI_GATE_Q>;
    case FSB_EPI_GELU: return conv_halo_kernel<T, FSB_EPI_GELU>;
    case FSB_EPI_SCALE_RES: return conv_halo_kernel<T, FSB_EPI_SCALE_RES>;
    default: return nullptr;
  }
}

}  // namespace fsb

extern "C" int fsb_conv2d_nhwc(const fsb_conv_args* a, int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_conv2d_nhwc");
  using namespace fsb;
  FSB_REQUIRE(a != nullptr, FSB_ERR_ARG, "fsb_conv2d_nhwc: NULL args");
  FSB_REQUIRE(dtype == FSB_F16 || dtype == FSB_BF16, FSB_ERR_ARG, "fsb_conv2d_nhwc: dtype %d (need f16/bf16)", dtype);
  FSB_REQUIRE(a->nsrc >= 1 && a->nsrc <= FSB_MAX_SRC, FSB_ERR_ARG, "fsb_conv2d_nhwc: nsrc=%d", a->nsrc);
  FSB_REQUIRE(a->B > 0 && a->H > 0 && a->W > 0, FSB_ERR_ARG, "fsb_conv2d_nhwc: non-positive size");
  FSB_REQUIRE(a->nseg == 1 || a->nseg == 2, FSB_ERR_ARG, "fsb_conv2d_nhwc: nseg=%d", a->nseg);
  FSB_REQUIRE(a->weight && a->bias, FSB_ERR_ARG, "fsb_conv2d_nhwc: NULL weight/bias");
  const int BN = a->block_n;
  FSB_REQUIRE(BN == 16 || (BN >= 32 && BN <= 256 && BN % 32 == 0), FSB_ERR_SHAPE,
              "fsb_conv2d_nhwc: block_n=%d (16 or a multiple of 32 up to 256)", BN);
  FSB_REQUIRE(a->nseg * BN <= 512, FSB_ERR_SHAPE, "fsb_conv2d_nhwc: %d accumulators of %d columns exceed TMEM", a->nseg, BN);
  FSB_REQUIRE(a->rows > 0 && a->rows % BN == 0 && a->Cout > 0 && a->Cout <= a->rows, FSB_ERR_SHAPE,
              "fsb_conv2d_nhwc: rows=%d block_n=%d Cout=%d", a->rows, BN, a->Cout);
  for (int i = 0; i < a->nsrc; ++i) {
    FSB_REQUIRE(a->src[i].ptr && a->src[i].C > 0 && a->src[i].C % kBK == 0, FSB_ERR_SHAPE,
                "fsb_conv2d_nhwc: source %d needs a non-NULL pointer and C %% 64 == 0 (C=%d)", i, a->src[i].C);
    FSB_REQUIRE(aligned16(a->src[i].ptr), FSB_ERR_ALIGN, "fsb_conv2d_nhwc: source %d not 16-byte aligned", i);
  }
  FSB_REQUIRE(aligned16(a->weight), FSB_ERR_ALIGN, "fsb_conv2d_nhwc: weight not 16-byte aligned");
  const int epi = a->epilogue;
  if (epi == FSB_EPI_GATE_ZR) {
    FSB_REQUIRE(a->h32 && a->z_l && a->rh_l && (a->nseg == 1 || (a->z_s && a->rh_s)), FSB_ERR_ARG,
                "fsb_conv2d_nhwc: GATE_ZR needs h32, z_*, rh_*");
    FSB_REQUIRE(BN % 64 == 0 && a->hidden > 0 && a->hidden % 32 == 0 && a->rows == 2 * a->hidden, FSB_ERR_SHAPE,
                "fsb_conv2d_nhwc: GATE_ZR needs block_n %% 64 == 0 and rows == 2*hidden (rows=%d hidden=%d)", a->rows, a->hidden);
  } else if (epi == FSB_EPI_GATE_Q) {
    FSB_REQUIRE(a->nseg == 2 && a->h32 && a->att && a->z_l && a->z_s && a->h32_out && a->h16_out, FSB_ERR_ARG,
                "fsb_conv2d_nhwc: GATE_Q needs two segments, h32, att, z_l, z_s, h32_out, h16_out");
    FSB_REQUIRE(BN % 32 == 0 && a->rows == a->hidden && a->Cout == a->hidden, FSB_ERR_SHAPE,
                "fsb_conv2d_nhwc: GATE_Q needs rows == Cout == hidden");
  } else {
    FSB_REQUIRE(epi == FSB_EPI_BIAS || epi == FSB_EPI_RELU || epi == FSB_EPI_GELU || epi == FSB_EPI_SCALE_RES,
                FSB_ERR_ARG, "fsb_conv2d_nhwc: unknown epilogue %d", epi);
    FSB_REQUIRE(a->nseg == 1, FSB_ERR_ARG, "fsb_conv2d_nhwc: plain epilogues take one segment");
    FSB_REQUIRE(a->out || a->out_f32, FSB_ERR_ARG, "fsb_conv2d_nhwc: no output");
    if (a->out) {
      FSB_REQUIRE(a->ld_out >= a->ch_off + a->Cout && a->ld_out % 8 == 0 && a->ch_off % 8 == 0 && aligned16(a->out),
                  FSB_ERR_SHAPE, "fsb_conv2d_nhwc: out needs ld_out %% 8 == 0, ch_off %% 8 == 0 (ld_out=%d ch_off=%d Cout=%d)",
                  a->ld_out, a->ch_off, a->Cout);
    }
    if (epi == FSB_EPI_SCALE_RES)
      FSB_REQUIRE(a->res && a->gamma && a->Cout % 32 == 0 && a->ld_res % 8 == 0, FSB_ERR_ARG,
                  "fsb_conv2d_nhwc: SCALE_RES needs res, gamma, Cout %% 32 == 0");
  }

  const bool halo = a->layout == FSB_CONV_LAYOUT_HALO;
  FSB_REQUIRE(halo || a->layout == FSB_CONV_LAYOUT_TAP, FSB_ERR_ARG, "fsb_conv2d_nhwc: unknown weight layout %d", a->layout);
  FSB_REQUIRE(!halo || 2 * a->nseg * BN <= 512, FSB_ERR_SHAPE,
              "fsb_conv2d_nhwc: halo layout keeps 2 pixel tiles x %d accumulators of %d columns in TMEM (> 512)", a->nseg, BN);

  EncodeTiledFn enc = encode_fn();
  FSB_REQUIRE(enc != nullptr, FSB_ERR_CUDA, "fsb_conv2d_nhwc: cuTensorMapEncodeTiled not available from the driver");

  ConvParams p;
  memset(&p, 0, sizeof(p));
  int tw_log2 = 4;
  pick_tile(a->H, a->W, &tw_log2);
  if (halo) tw_log2 = 3;                                  // 8-pixel tile rows = one swizzle atom
  const int TW = 1 << tw_log2, TH = halo ? 32 : (kBM >> tw_log2);  // halo: two stacked 16 x 8 tiles per CTA
  p.H = a->H;
  p.W = a->W;
  p.tiles_x = ceil_div(a->W, TW);
  p.tiles_y = ceil_div(a->H, TH);
  p.tw_log2 = tw_log2;
  p.BN = BN;
  p.nacc = a->nseg;
  // tap kernel: two accumulator buffers when they fit the 512 TMEM columns (epilogue of tile i overlaps the MMAs of
  // tile i + 1); halo kernel: two stacked pixel tiles
  p.nbuf = (!halo && 2 * a->nseg * BN <= 512) ? 2 : 1;
  int cols = 32;
  while (cols < (halo ? 2 : p.nbuf) * a->nseg * BN) cols <<= 1;
  p.tmem_cols = cols;
  FSB_REQUIRE(halo || a->nseg * a->rows <= 2 * kMaxRows, FSB_ERR_SHAPE, "fsb_conv2d_nhwc: %d x %d bias values exceed %d",
              a->nseg, a->rows, 2 * kMaxRows);

  // K-step program + Ktot
  int n = 0;
  for (int s = 0; s < a->nseg; ++s) {
    const fsb_conv_seg& sg = a->seg[s];
    FSB_REQUIRE(sg.ksize == 1 || sg.ksize == 3, FSB_ERR_SHAPE, "fsb_conv2d_nhwc: segment %d ksize=%d", s, sg.ksize);
    FSB_REQUIRE(sg.nsrc >= 1 && sg.nsrc <= FSB_MAX_SRC, FSB_ERR_ARG, "fsb_conv2d_nhwc: segment %d nsrc=%d", s, sg.nsrc);
    const int rad = sg.ksize / 2;
    if (halo) {
      // K order: source -> 64-channel chunk -> dx -> dy (one activation box per (source, chunk, dx))
      for (int k = 0; k < sg.nsrc; ++k) {
        const int si = sg.src[k];
        FSB_REQUIRE(si >= 0 && si < a->nsrc, FSB_ERR_ARG, "fsb_conv2d_nhwc: segment %d source index %d", s, si);
        for (int c = 0; c < a->src[si].C / kBK; ++c)
          for (int dx = -rad; dx <= rad; ++dx)
            for (int dy = -rad; dy <= rad; ++dy) {
              FSB_REQUIRE(n < kMaxSteps, FSB_ERR_SHAPE, "fsb_conv2d_nhwc: more than %d K steps", kMaxSteps);
              p.steps[n++] = make_halo_step(si, dy + 1, dx, s, c, dy == -rad, dy == rad);
            }
      }
      continue;
    }
    for (int dy = -rad; dy <= rad; ++dy)
      for (int dx = -rad; dx <= rad; ++dx)
        for (int k = 0; k < sg.nsrc; ++k) {
          const int si = sg.src[k];
          FSB_REQUIRE(si >= 0 && si < a->nsrc, FSB_ERR_ARG, "fsb_conv2d_nhwc: segment %d source index %d", s, si);
          for (int c = 0; c < a->src[si].C / kBK; ++c) {
            FSB_REQUIRE(n < kMaxSteps, FSB_ERR_SHAPE, "fsb_conv2d_nhwc: more than %d K steps", kMaxSteps);
            p.steps[n] = make_step(si, dy, dx, s, c, n);
            ++n;
          }
        }
  }
  p.nsteps = n;
  const int Ktot = n * kBK;

  // Pipeline shape of the persistent tap kernel: one CTA per SM owns ~200 KB of shared memory; a stage holds
  // two K steps for N tiles up to 128 (each producer pays one mbarrier round per two loads), one above.
  // pair mode (tcgen05.mma.cta_group::2 over two pixel tiles, half a weight tile per CTA): at least two pixel tiles
  // and halves that are whole 8-row swizzle atoms; FSB_CONV_PAIR=0 switches it off (A/B measurements)
  const long long mtiles0 = (long long)p.tiles_x * p.tiles_y * a->B;
  static const int pair_env = [] { const char* e = getenv("FSB_CONV_PAIR"); return e ? atoi(e) : 1; }();
  // (short K loops lose: 1x1 128->512 30 -> 43 us, while 3x3 512->256 gains 177 -> 135 us; tools/conv_sweep.py)
  int nsteps_all = 0;
  for (int sgi = 0; sgi < a->nseg; ++sgi) {
    int ch = 0;
    for (int k = 0; k < a->seg[sgi].nsrc; ++k) {
      const int si = a->seg[sgi].src[k];
      if (si >= 0 && si < a->nsrc) ch += a->src[si].C / kBK;
    }
    nsteps_all += a->seg[sgi].ksize * a->seg[sgi].ksize * ch;
  }
  const int cl = (!halo && pair_env != 0 && mtiles0 >= 2 && BN % 16 == 0 && BN >= 32 && nsteps_all >= 24) ? 2 : 1;
  const size_t step_bytes = (size_t)kATileBytes + (size_t)(BN / cl) * kBK * 2;
  int kper = (BN / cl) <= 128 ? 2 : 1, stages = a->stages;
  if (kper > n) kper = n;
  // the TMA-store epilogue (plain epilogues with a 16-bit output, decided below with the same condition) keeps
  // 16 KB of store tiles behind the ring
  const bool ostore_plain = !halo && a->out != nullptr && a->out_f32 == nullptr &&
                            (epi == FSB_EPI_BIAS || epi == FSB_EPI_RELU || epi == FSB_EPI_GELU) && a->Cout >= 32;
  // gate epilogues: only in pair mode, where a stage is small enough that the 24 KB of tiles do not cost a ring stage
  static const int gate_store_env = [] { const char* e = getenv("FSB_GATE_STORE"); return e ? atoi(e) : 1; }();
  const bool ostore_gate = !halo && cl == 2 && gate_store_env != 0 && (epi == FSB_EPI_GATE_ZR || epi == FSB_EPI_GATE_Q) &&
                           a->hidden % 32 == 0;
  const bool ostore = ostore_plain || ostore_gate;
  // Bytes in flight are what the operand feed runs on (a 144 KB ring was 10-25 % slower than a 192 KB one on the
  // same layer): 227 KB per SM - ~5 KB static - 1 KB alignment slack - 16 KB of store tiles leave 204 KB
  const size_t ring_budget = (ostore ? 196 : 220) * 1024;  // 24 KB of store tiles
  if (kper == 2 && ring_budget / (2 * step_bytes) < 3) kper = 1;  // at least three stages in flight
  if (stages <= 0) stages = (int)(ring_budget / (kper * step_bytes));
  if (stages > 8) stages = 8;
  while (stages > 1 && stages * kper * step_bytes > ring_budget) --stages;
  if (stages < 1) stages = 1;
  size_t smem = (size_t)stages * kper * step_bytes + 1024 + (ostore ? 8 * 3072 : 0);
  // Row-reuse mode (pair mode, 3x3 segment [+ a 1x1 segment over the same sources]): a stage holds ONE activation
  // box of TH + 2 image rows per (dx, source, chunk) -- its three vertical taps are windows of that box, the 1x1
  // segment reads the centre window of the dx = 0 box -- and the 3 (4) weight tiles that go with it.  An SM then
  // ingests (TH+2)/(3 TH) of the activation bytes (20 KB instead of 48 KB per three K steps for 8 x 16 tiles), which
  // takes the block_n = 128 layers from the L2-ingest bound to the MMA bound.
  static const int vr_env = [] { const char* e = getenv("FSB_CONV_VR"); return e ? atoi(e) : 1; }();
  bool vr = !halo && cl == 2 && vr_env != 0 && tw_log2 >= 3 && a->seg[0].ksize == 3 && (a->nseg == 1 || a->seg[1].ksize == 1);
  if (vr) {
    const int nslots = a->nseg == 2 ? 4 : 3;
    const size_t vr_stage = (size_t)(TH + 2) * TW * kBK * 2 + (size_t)nslots * (BN / cl) * kBK * 2;
    int chs = 0;
    for (int k = 0; k < a->seg[0].nsrc; ++k) chs += a->src[a->seg[0].src[k]].C / kBK;
    // 1x1 segment: a (source, chunk) that the 3x3 segment reads too rides in the fourth slot of that dx = 0 group;
    // one it does not read (the GRU q gate: r_small * h) gets a group of its own
    int ride[FSB_MAX_SRC], own_chunks = 0;  // ride[i] = weight column block of source i in the 1x1 segment, -1: none
    for (int i = 0; i < FSB_MAX_SRC; ++i) ride[i] = -1;
    if (a->nseg == 2) {
      int off = 9 * chs;
      for (int k = 0; k < a->seg[1].nsrc; ++k) {
        const int si = a->seg[1].src[k];
        ride[si] = off;
        off += a->src[si].C / kBK;
        bool shared = false;
        for (int k0 = 0; k0 < a->seg[0].nsrc; ++k0) shared = shared || a->seg[0].src[k0] == si;
        if (!shared) own_chunks += a->src[si].C / kBK;
      }
    }
    const int vr_steps = (3 * chs + own_chunks) * nslots;
    if (ring_budget / vr_stage < 3 || vr_steps > kMaxSteps) {
      vr = false;
    } else {
      int m = 0;
      bool in_seg0[FSB_MAX_SRC] = {false, false, false, false};
      for (int dx = -1; dx <= 1; ++dx) {
        int choff = 0;
        for (int k = 0; k < a->seg[0].nsrc; ++k) {
          const int si = a->seg[0].src[k], nch = a->src[si].C / kBK;
          in_seg0[si] = true;
          for (int c = 0; c < nch; ++c) {
            for (int dy = -1; dy <= 1; ++dy)  // weight column block = the step's position in tap order
              p.steps[m++] = make_step(si, dy, dx, 0, c, ((dy + 1) * 3 + (dx + 1)) * chs + choff + c);
            if (nslots == 4)
              p.steps[m++] = (dx == 0 && ride[si] >= 0) ? make_step(si, 0, 0, 1, c, ride[si] + c)
                                                        : (make_step(si, 0, dx, 0, c, 0) | kStepNull);
          }
          choff += nch;
        }
      }
      if (a->nseg == 2)
        for (int k = 0; k < a->seg[1].nsrc; ++k) {
          const int si = a->seg[1].src[k];
          if (in_seg0[si]) continue;
          for (int c = 0; c < a->src[si].C / kBK; ++c) {
            p.steps[m++] = make_step(si, 0, 0, 1, c, ride[si] + c);  // the centre window of its own box
            for (int j = 1; j < nslots; ++j) p.steps[m++] = make_step(si, 0, 0, 1, c, 0) | kStepNull;
          }
        }
      p.nsteps = m;
      kper = nslots;
      stages = a->stages > 0 ? a->stages : (int)(ring_budget / vr_stage);
      if (stages > 8) stages = 8;
      while (stages > 1 && stages * vr_stage > ring_budget) --stages;
      smem = (size_t)stages * vr_stage + 1024 + (ostore ? 8 * 3072 : 0);
    }
  }
  p.vr = vr ? 1 : 0;
  p.stages = stages;
  p.kper = kper;
  if (halo) {
    // activation ring: 3 boxes of 34 KB; weight ring: what is left of 220 KB (2..8 tiles)
    const size_t b_bytes = (size_t)BN * kBK * 2;
    p.SA = 3;
    int sb = (int)((220 * 1024 - 1024 - (size_t)p.SA * kASlotBytes) / b_bytes);
    if (sb > 8) sb = 8;
    if (sb > n) sb = n;
    FSB_REQUIRE(sb >= 2 || n < 2, FSB_ERR_SHAPE, "fsb_conv2d_nhwc: block_n=%d leaves no room for a weight ring", BN);
    p.SB = sb < 1 ? 1 : sb;
    smem = (size_t)p.SA * kASlotBytes + (size_t)p.SB * b_bytes + 1024;
  }

  const CUtensorMapDataType dt = dtype == FSB_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  for (int i = 0; i < a->nsrc; ++i) {
    const cuuint64_t C = (cuuint64_t)a->src[i].C;
    cuuint64_t dims[4] = {C, (cuuint64_t)a->W, (cuuint64_t)a->H, (cuuint64_t)a->B};
    cuuint64_t strides[3] = {C * 2, C * 2 * a->W, C * 2 * a->W * a->H};
    cuuint32_t box[4] = {(cuuint32_t)kBK, (cuuint32_t)TW, (cuuint32_t)(halo ? kHaloRows : (vr ? TH + 2 : TH)), 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&p.amap[i], dt, 4, const_cast<void*>(a->src[i].ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FSB_REQUIRE(r == CUDA_SUCCESS, FSB_ERR_CUDA, "fsb_conv2d_nhwc: cuTensorMapEncodeTiled(source %d) -> %d", i, (int)r);
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)Ktot, (cuuint64_t)a->rows};
    cuuint64_t strides[1] = {(cuuint64_t)Ktot * 2};
    cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)BN};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&p.wmap, dt, 2, const_cast<void*>(a->weight), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FSB_REQUIRE(r == CUDA_SUCCESS, FSB_ERR_CUDA, "fsb_conv2d_nhwc: cuTensorMapEncodeTiled(weight) -> %d", (int)r);
  }
  // TMA-store epilogue of the tap kernel: plain epilogues with a 16-bit NHWC output
  p.ostore = 0;
  if (ostore) {
    const int bw = TW < 32 ? TW : 32, bh = 32 / bw;
    // (C, W, H, B) map of an NHWC output with a (cbox channels, bw, bh, 1) box = the 32 pixels of one epilogue warp
    auto out_map = [&](CUtensorMap* m, void* ptr, int ld, int cbox, CUtensorMapDataType type, int esz, CUtensorMapSwizzle sw,
                       const char* what) -> int {
      cuuint64_t dims[4] = {(cuuint64_t)ld, (cuuint64_t)a->W, (cuuint64_t)a->H, (cuuint64_t)a->B};
      cuuint64_t strides[3] = {(cuuint64_t)ld * esz, (cuuint64_t)ld * esz * a->W, (cuuint64_t)ld * esz * a->W * a->H};
      cuuint32_t box[4] = {(cuuint32_t)cbox, (cuuint32_t)bw, (cuuint32_t)bh, 1};
      cuuint32_t estr[4] = {1, 1, 1, 1};
      CUresult r = enc(m, type, 4, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                       CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      FSB_REQUIRE(r == CUDA_SUCCESS, FSB_ERR_CUDA, "fsb_conv2d_nhwc: cuTensorMapEncodeTiled(%s) -> %d", what, (int)r);
      return FSB_OK;
    };
    int rc = FSB_OK;
    if (ostore_plain) {
      rc = out_map(&p.omap, a->out, a->ld_out, 32, dt, 2, CU_TENSOR_MAP_SWIZZLE_64B, "out");
    } else if (epi == FSB_EPI_GATE_ZR) {
      void* ptrs[4] = {a->z_l, a->z_s, a->rh_l, a->rh_s};
      for (int i = 0; i < 4 && rc == FSB_OK; ++i)
        if (ptrs[i] != nullptr) rc = out_map(&p.gmap[i], ptrs[i], a->hidden, 32, dt, 2, CU_TENSOR_MAP_SWIZZLE_64B, "gate output");
    } else {
      rc = out_map(&p.gmap[0], a->h16_out, a->hidden, 16, dt, 2, CU_TENSOR_MAP_SWIZZLE_NONE, "h16_out");
      if (rc == FSB_OK)
        rc = out_map(&p.gmap[1], a->h32_out, a->hidden, 16, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, CU_TENSOR_MAP_SWIZZLE_64B, "h32_out");
    }
    if (rc != FSB_OK) return rc;
    p.ostore = 1;
  }
  if (cl == 2) {
    cuuint64_t dims[2] = {(cuuint64_t)Ktot, (cuuint64_t)a->rows};
    cuuint64_t strides[1] = {(cuuint64_t)Ktot * 2};
    cuuint32_t box[2] = {(cuuint32_t)kBK, (cuuint32_t)(BN / 2)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&p.wmap_half, dt, 2, const_cast<void*>(a->weight), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FSB_REQUIRE(r == CUDA_SUCCESS, FSB_ERR_CUDA, "fsb_conv2d_nhwc: cuTensorMapEncodeTiled(weight half) -> %d", (int)r);
  }
  // instruction descriptor, kind::f16: D fp32, A/B f16|bf16, both K-major, M = 128 (256 over a CTA pair), N = block_n
  const uint32_t fmt = dtype == FSB_F16 ? 0u : 1u;
  p.idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((cl * kBM) >> 4) << 24);
  p.rows = a->rows;
  p.Cout = a->Cout;
  p.bias = a->bias;
  p.out = a->out;
  p.ld_out = a->ld_out;
  p.ch_off = a->ch_off;
  p.out_f32 = a->out_f32;
  p.out_scale = a->out_scale;
  p.res = a->res;
  p.ld_res = a->ld_res;
  p.gamma = a->gamma;
  p.h32 = a->h32;
  p.att = a->att;
  p.z_l = a->z_l;
  p.z_s = a->z_s;
  p.rh_l = a->rh_l;
  p.rh_s = a->rh_s;
  p.h32_out = a->h32_out;
  p.h16_out = a->h16_out;
  p.hidden = a->hidden;

  auto kern = halo ? (dtype == FSB_F16 ? halo_kernel_for<__half>(epi) : halo_kernel_for<__nv_bfloat16>(epi))
                   : (dtype == FSB_F16 ? kernel_for<__half>(epi, cl) : kernel_for<__nv_bfloat16>(epi, cl));
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return cuda_fail(e, "fsb_conv2d_nhwc: cudaFuncSetAttribute");
  const long long mtiles = (long long)p.tiles_x * p.tiles_y * a->B;
  FSB_REQUIRE(mtiles * (a->rows / BN) < (1ll << 31) && a->rows / BN < 65536, FSB_ERR_SHAPE, "fsb_conv2d_nhwc: grid too large");
  p.ntiles_n = a->rows / BN;
  p.nbatch = a->B;
  p.ntiles = (int)((cl == 2 ? (mtiles + 1) / 2 : mtiles) * p.ntiles_n);  // work items: (pixel tile | pair of them, N tile)
  dim3 grid((unsigned)mtiles, (unsigned)(a->rows / BN));
  if (halo) {
    launch_kernel(kern, dim3(grid), dim3(kConvThreads), smem, (cudaStream_t)stream, p);
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kConvThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (pdl_enabled()) {
      attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[na].val.programmaticStreamSerializationAllowed = 1;
      ++na;
    }
    int nctas = p.ntiles < sm_count() ? p.ntiles : sm_count();
    if (cl == 2) {
      attr[na].id = cudaLaunchAttributeClusterDimension;
      attr[na].val.clusterDim.x = 2;
      attr[na].val.clusterDim.y = 1;
      attr[na].val.clusterDim.z = 1;
      ++na;
      // how many pairs the machine holds at one CTA per SM (queried once per shared-memory size)
      static int max_clusters[64] = {0};
      const int slot = (int)((smem >> 12) & 63);
      int nc = max_clusters[slot];
      if (nc == 0) {
        cfg.gridDim = dim3(2 * sm_count());
        cfg.attrs = attr;
        cfg.numAttrs = na;
        if (cudaOccupancyMaxActiveClusters(&nc, kern, &cfg) != cudaSuccess || nc <= 0) {
          cudaGetLastError();
          nc = sm_count() / 2;
        }
        max_clusters[slot] = nc;
      }
      nctas = 2 * (p.ntiles < nc ? p.ntiles : nc);
    }
    cfg.gridDim = dim3((unsigned)nctas);
    cfg.attrs = attr;
    cfg.numAttrs = na;
    cudaError_t le = cudaLaunchKernelEx(&cfg, kern, p);
    if (le != cudaSuccess) return cuda_fail(le, "fsb_conv2d_nhwc: cudaLaunchKernelEx");
  }
  FSB_CHECK_LAUNCH("fsb_conv2d_nhwc");
  return FSB_OK;
}
