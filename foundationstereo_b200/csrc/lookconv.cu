// Kernel 3+4: the per-iteration geometry lookup (core/geometry.py:43-65, core/utils/utils.py:44-55) FUSED into the first
// layer of the motion encoder, encoder.convc1 1x1 + ReLU (core/update.py:56,63) -- SURVEY.md 8(f) row 1: the lookup
// tensor (L*K*(Cg+1) = 522 channels per pixel) never exists in HBM.
//
//   c1[p, :] = ReLU(W1 . lookup(p) + b1)          p = pixel, lookup(p) rounded to the 16-bit operand type
//
// One persistent CTA per SM walks 128-pixel tiles.  The K dimension of the GEMM is split BY PYRAMID LEVEL: level l owns
// the operand half A[l] (Kl = K*(Cg+1) rounded up to 64 channels = 5 chunks of 64 for Cg = 28) and the columns
// [l*Kl, (l+1)*Kl) of the packed weight matrix, so the two halves ping-pong:
//   * 16 GATHER warps produce A[l] straight into shared memory.  A pass = 4 consecutive pixels: positions with the
//     reference's fp32 round trip (lookup.cuh), computed co-operatively -- one division per tap, shuffled to the lanes
//     that use them --, lane (pixel slot, channel quad) loads the K+1 window rows of its four channels as float4 (the
//     window of a pixel is one contiguous run in the (BHW, D_l, Cg) layout), lerps 4 x K samples and writes them -- 72
//     contiguous bytes of row p of the K-major SWIZZLE_128B tile [chunk][128 rows][64 channels], the layout TMA would
//     have delivered from an NHWC lookup tensor -- as four 16-byte and one 8-byte store; lanes 28..31 do the
//     correlation taps;
//   * a TMA warp streams W1 (Cout = 256 rows, 16-bit K-major) from L2 through a two-stage ring of 64-channel chunks;
//   * one elected thread issues tcgen05.mma (M = 128 pixels, N = 256, K = 16 per instruction) over the chunks of A[l]
//     while the gather warps fill A[1-l], into one of TWO 256-column TMEM accumulators, so that
//   * 4 EPILOGUE warps (bias, ReLU, 16-bit, 64 contiguous bytes per lane and 32 columns) drain tile i under tile i+1.
// The first form of this kernel (one 144 KB operand, all levels, then the MMAs) serialised gather and MMA -- the weight
// stream is latency-bound at 64 KB in flight, 7 us per tile -- and ran 255 us against 93 + 54 us for the two launches.
// Same operand values as fsb_geo_lookup(out_layout 1/2) followed by the convc1 launch of fsb_conv2d_nhwc -- rounded to 16
// bits exactly as the lookup kernel rounds its output --, fp32 accumulation in TMEM in another K order (by level): the two
// paths agree to accumulation-order rounding (4e-4 of the 16-bit result's range in the tests).
#include "common.cuh"
#include "lookup.cuh"
#include "umma.cuh"

namespace fsb {

constexpr int kLcK = 9;                 // taps (radius 4, core/foundation_stereo.py:176)
constexpr int kLcMaxLC = 5;             // 64-channel chunks per level: K*(Cg+1) <= 320
constexpr int kLcN = 256;               // convc1 output channels (core/update.py:56)
constexpr int kLcGatherWarps = 16;
constexpr int kLcDepth = 1;             // passes in flight per gather warp (2: software-pipelined, needs 8 warps of 128 registers)
constexpr int kLcEpiWarps = 4;
constexpr int kLcThreads = (2 + kLcEpiWarps + kLcGatherWarps) * 32;  // 704
constexpr int kLcATile = 128 * 128;     // one 64-channel chunk of the A operand: 128 rows x 128 B
constexpr int kLcAHalf = kLcMaxLC * kLcATile;  // one level's operand: 80 KB
constexpr int kLcWTile = kLcN * 128;    // one 64-channel chunk of W1: 256 rows x 128 B
constexpr int kLcWStages = 2;
constexpr uint32_t kLcOffA = 0;
constexpr uint32_t kLcOffW = kLcOffA + 2 * kLcAHalf;
constexpr uint32_t kLcSmem = kLcOffW + kLcWStages * kLcWTile + 1024;  // 230400 B

struct alignas(64) LcParams {
  CUtensorMap wmap;
  LevelPtrs geo, cor;
  const float* disp;
  const float* coords;
  const float* bias;
  void* out;          // (npix, 256) 16-bit
  long long npix;
  int Cg, D, W2, L, r, nlc, ntiles;   // nlc = chunks per level
  uint32_t idesc;
};

template <typename T, int LT, int CGT>
__global__ void __launch_bounds__(kLcThreads, 1)
lookup_convc1_kernel(const __grid_constant__ LcParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // barriers: a_full[2] a_empty[2] | w_full[2] w_empty[2] | acc_full[2] acc_empty[2]
  __shared__ __align__(8) uint64_t bars[4 + 2 * kLcWStages + 4];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float bias_s[kLcN];

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t a_full = smem_u32(&bars[0]), a_empty = smem_u32(&bars[2]);
  const uint32_t w_full = smem_u32(&bars[4]), w_empty = smem_u32(&bars[4 + kLcWStages]);
  const uint32_t acc_full = smem_u32(&bars[4 + 2 * kLcWStages]), acc_empty = acc_full + 16;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(a_full + 8 * i, kLcGatherWarps);
      mbar_init(a_empty + 8 * i, 1);
      mbar_init(acc_full + 8 * i, 1);
      mbar_init(acc_empty + 8 * i, kLcEpiWarps);
    }
    for (int i = 0; i < kLcWStages; ++i) {
      mbar_init(w_full + 8 * i, 1);
      mbar_init(w_empty + 8 * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1 && lane == 0) tma_prefetch_desc(&p.wmap);
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp >= 2 && warp < 2 + kLcEpiWarps) {
    for (int i = threadIdx.x - 64; i < kLcN; i += kLcEpiWarps * 32) bias_s[i] = __ldg(p.bias + i);  // parameters
  }
  if (warp >= 2 + kLcEpiWarps) {
    // the operand starts as zeros: the pad channels [K*(Cg+1), Kl) of a level are never written again (W1 is zero
    // there, but 0 x NaN would not be), rows past npix of the last tile keep finite values
    uint4* a4 = reinterpret_cast<uint4*>(gbase + kLcOffA);
    for (int i = threadIdx.x - (2 + kLcEpiWarps) * 32; i < 2 * kLcAHalf / 16; i += kLcGatherWarps * 32)
      a4[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;  // accumulator b: columns [256 b, 256 b + 256)
  pdl_launch_dependents();
  if (warp != 1) pdl_wait();  // the weight producer only reads parameters no kernel writes

  if (warp == 1) {
    // ================================================================= W1 producer (consumption order: level, chunk)
    int s = 0;
    uint32_t ph = 0;
    bool first_round = true;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
#pragma unroll 1
      for (int k = 0; k < LT * p.nlc; ++k) {
        if (!first_round) mbar_wait(w_empty + 8 * s, ph ^ 1);
        __syncwarp();
        if (elect_one()) {
          mbar_expect_tx(w_full + 8 * s, kLcWTile);
          tma_load_2d(base + kLcOffW + s * kLcWTile, &p.wmap, w_full + 8 * s, k * 64, 0);
        }
        if (++s == kLcWStages) { s = 0; ph ^= 1; first_round = false; }
      }
    }
  } else if (warp == 0) {
    // ================================================================= MMA issuer
    int ws = 0;
    uint32_t wph = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
      const int b = it & 1;
      const uint32_t d = tmem_base + (uint32_t)b * 256u;
      mbar_wait(acc_empty + 8 * b, ((uint32_t)(it >> 1) & 1) ^ 1);  // the epilogue of tile it-2 drained this accumulator
#pragma unroll 1
      for (int h = 0; h < LT; ++h) {
        mbar_wait(a_full + 8 * h, (uint32_t)(it & 1));              // the gather warps wrote level h of this tile
        tc_fence_after();
#pragma unroll 1
        for (int k = 0; k < p.nlc; ++k) {
          mbar_wait(w_full + 8 * ws, wph);
          tc_fence_after();
          __syncwarp();
          const uint64_t adesc = umma_desc_k128(base + kLcOffA + h * kLcAHalf + k * kLcATile);
          const uint64_t bdesc = umma_desc_k128(base + kLcOffW + ws * kLcWTile);
          if (elect_one()) {
            umma_f16(d, adesc, bdesc, p.idesc, (uint32_t)(h > 0 || k > 0));
            umma_f16(d, adesc + 2, bdesc + 2, p.idesc, 1);
            umma_f16(d, adesc + 4, bdesc + 4, p.idesc, 1);
            umma_f16(d, adesc + 6, bdesc + 6, p.idesc, 1);
            umma_commit(w_empty + 8 * ws);
          }
          if (++ws == kLcWStages) { ws = 0; wph ^= 1; }
        }
        if (elect_one()) {
          umma_commit(a_empty + 8 * h);                     // this level's operand may be refilled
          if (h == LT - 1) umma_commit(acc_full + 8 * b);   // the accumulator is complete
        }
      }
    }
  } else if (warp < 2 + kLcEpiWarps) {
    // ================================================================= epilogue warps (TMEM lane quarter = warp & 3)
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    T* out = reinterpret_cast<T*>(p.out);
    int it = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
      const int b = it & 1;
      mbar_wait(acc_full + 8 * b, (uint32_t)(it >> 1) & 1);
      tc_fence_after();
      const long long pix = (long long)tile * 128 + r;
#pragma unroll 1
      for (int g = 0; g < kLcN / 32; ++g) {
        uint32_t v[32];
        tmem_ld32(lane_addr + (uint32_t)(b * 256 + g * 32), v);
        tmem_ld_wait();
        float f[32];
        const float4* b4 = reinterpret_cast<const float4*>(bias_s + g * 32);
#pragma unroll
        for (int i4 = 0; i4 < 8; ++i4) {
          const float4 bb = b4[i4];
          f[4 * i4 + 0] = fmaxf(__uint_as_float(v[4 * i4 + 0]) + bb.x, 0.f);
          f[4 * i4 + 1] = fmaxf(__uint_as_float(v[4 * i4 + 1]) + bb.y, 0.f);
          f[4 * i4 + 2] = fmaxf(__uint_as_float(v[4 * i4 + 2]) + bb.z, 0.f);
          f[4 * i4 + 3] = fmaxf(__uint_as_float(v[4 * i4 + 3]) + bb.w, 0.f);
        }
        if (pix < p.npix) store32<T>(out + (size_t)pix * kLcN + g * 32, f);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(acc_empty + 8 * b);
    }
  } else {
    // ================================================================= gather warps
    // A pass = 4 consecutive pixels of the tile for one level.
    //   positions: lane (kind, slot, sub) = (lane >> 4, (lane >> 2) & 3, lane & 3) computes taps sub, sub + 4, sub + 8 of
    //     pixel `slot` for the geometry (kind 0) or correlation (kind 1) coordinate -- the same fp32 round trip as
    //     lookup_level (lookup.cuh), one division per tap instead of nine per thread; window base and per-tap fractions
    //     reach their consumers by shuffle;
    //   gather: lane (slot, q) = pixel slot 0..3 x channel quad q < Cg/4 loads the K+1 window rows of its four channels
    //     as float4 (the window of a pixel is one contiguous run in the (BHW, D_l, Cg) layout), lerps 4 x K samples and
    //     writes them -- channels (4q..4q+3)*K + k are 72 contiguous bytes of the operand row -- as four 16-byte and
    //     one 8-byte store; lanes 28..31 do the K correlation taps of the four pixels.
    // The gather is bound by LATENCY (window loads and the dependent position arithmetic), i.e. by the number of warps:
    // 16 warps with one pass in flight each (80 registers) ran 152 us at 8 pairs, 8 warps software-pipelined two deep
    // (kLcDepth = 2: the loads of pass s+1 issued before pass s is lerped and stored; 128 registers) 212 us.
    constexpr int K = kLcK;
    constexpr int kPasses = 32 / kLcGatherWarps;       // passes of a warp per tile and level
    constexpr int kSteps = LT * kPasses;               // per tile; even, so a step's buffer parity is the same in every tile
    static_assert(kLcDepth == 1 || kSteps % 2 == 0, "two-deep pipeline over an odd number of steps per tile");
    const int gw = warp - (2 + kLcEpiWarps);
    const int Cg = CGT > 0 ? CGT : p.Cg, r = p.r, nq = Cg >> 2;
    const bool is_cor = lane >= 28;
    const int slot = is_cor ? lane - 28 : lane / nq;
    const int q = is_cor ? 0 : lane - slot * nq;
    const bool active = is_cor || slot < 4;
    const int pkind = lane >> 4, pslot = (lane >> 2) & 3, psub = lane & 3;
    const int src = (is_cor ? 16 : 0) + (slot & 3) * 4;   // first producer lane of this consumer's (kind, pixel)
    const int cor_byte = 72 * nq;   // = 2 K Cg: byte offset of the first correlation channel inside a level's operand row
    // disparity / x coordinate of this lane's producer pixel in the warp's passes of a tile, fetched one tile ahead
    float nd[kPasses], nx[kPasses];
    auto fetch_dx = [&](int tile) {
#pragma unroll
      for (int i = 0; i < kPasses; ++i) {
        const long long ppix = (long long)tile * 128 + 4 * (gw + kLcGatherWarps * i) + pslot;
        const bool ok = tile < p.ntiles && ppix < p.npix;
        nd[i] = ok ? __ldg(p.disp + ppix) : 0.f;
        nx[i] = ok ? __ldg(p.coords + ppix) : 0.f;
      }
    };
    // level pointers in registers (the steps below are unrolled over the level)
    const float* const geo0 = p.geo.p[0];
    const float* const geo1 = p.geo.p[LT > 1 ? 1 : 0];
    const float* const cor0 = p.cor.p[0];
    const float* const cor1 = p.cor.p[LT > 1 ? 1 : 0];
    // pipeline state of the two passes in flight
    float4 v[kLcDepth][K + 1];
    float pfr[kLcDepth][3];   // the producer role's fractions (taps sub, sub + 4, sub + 8): shuffled to the consumers at use
    bool live[kLcDepth];

    // positions + window loads of pass `pi` of level `h` of `tile` into buffer `b` (h, pi, b: constants after unrolling)
    auto issue = [&](int tile, int h, int pi, float dsp, float xco, int b) {
      const int Dl = p.D >> h, Wl = p.W2 >> h;
      const float inv = 1.f / (float)(1 << h);
      const int pass = gw + kLcGatherWarps * pi;
      float pf[3];
      const float d = dsp * inv;   // exact: power-of-two scale
      const float xc = pkind ? __fsub_rn(xco * inv, d) : d;
      const int n = pkind ? Wl : Dl;
#pragma unroll
      for (int rd = 0; rd < 3; ++rd) pf[rd] = ref_unnormalised(__fadd_rn(xc, (float)(psub + 4 * rd - r)), n);
      const float ix0 = __shfl_sync(0xffffffffu, pf[0], lane & ~3);
      const int pbase = (int)fminf(fmaxf(floorf(ix0), -1.0e6f), 1.0e6f);
#pragma unroll
      for (int rd = 0; rd < 3; ++rd) pfr[b][rd] = pf[rd] - (float)(pbase + psub + 4 * rd);
      const int base_ = __shfl_sync(0xffffffffu, pbase, src);
      const long long pix = (long long)tile * 128 + 4 * pass + (slot & 3);
      live[b] = active && tile < p.ntiles && pix < p.npix;
      const bool inside = base_ >= 0 && base_ + K < (is_cor ? Wl : Dl);
      const bool fast = __all_sync(0xffffffffu, !live[b] || inside);   // whole windows inside: no per-row tests
      if (live[b]) {
        if (!is_cor) {
          const float4* g4 = reinterpret_cast<const float4*>((h == 0 ? geo0 : geo1) + ((long long)pix * Dl + base_) * Cg) + q;
          if (fast) {
#pragma unroll
            for (int j = 0; j <= K; ++j) v[b][j] = __ldg(g4 + j * nq);
          } else {
#pragma unroll
            for (int j = 0; j <= K; ++j)
              v[b][j] = (unsigned)(base_ + j) < (unsigned)Dl ? __ldg(g4 + j * nq) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
        } else {
          const float* c1 = (h == 0 ? cor0 : cor1) + (long long)pix * Wl + base_;
          if (fast) {
#pragma unroll
            for (int j = 0; j <= K; ++j) v[b][j].x = __ldg(c1 + j);
          } else {
#pragma unroll
            for (int j = 0; j <= K; ++j) v[b][j].x = (unsigned)(base_ + j) < (unsigned)Wl ? __ldg(c1 + j) : 0.f;
          }
        }
      }
    };
    // lerp + store of the pass in buffer `b` into row 4 pass + slot of level h's operand
    auto consume = [&](int h, int pi, int b) {
      float frac[K];
#pragma unroll
      for (int k = 0; k < K; ++k) frac[k] = __shfl_sync(0xffffffffu, pfr[b][k >> 2], src + (k & 3));
      if (!live[b]) return;
      const int row = 4 * (gw + kLcGatherWarps * pi) + (slot & 3);
      uint8_t* arow = gbase + kLcOffA + h * kLcAHalf + row * 128;
      const int sw = row & 7;
      auto unit_ptr = [&](int U) -> uint8_t* {   // 16-byte unit U of the level's operand row
        return arow + (U >> 3) * kLcATile + (((U & 7) ^ sw) << 4);
      };
      if (!is_cor) {
        uint32_t w[2 * K];   // 36 samples in channel order (4q + j) * K + k, packed in pairs
#pragma unroll
        for (int i = 0; i < 2 * K; ++i) {
          float s2[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int idx = 2 * i + e, j = idx / K, k = idx - j * K;
            const float lo = j == 0 ? v[b][k].x : j == 1 ? v[b][k].y : j == 2 ? v[b][k].z : v[b][k].w;
            const float hi = j == 0 ? v[b][k + 1].x : j == 1 ? v[b][k + 1].y : j == 2 ? v[b][k + 1].z : v[b][k + 1].w;
            s2[e] = fmaf(frac[k], hi - lo, lo);   // the lookup kernel's form (geo.cu)
          }
          w[i] = pack2<T>(s2[0], s2[1]);
        }
        const int U0 = (72 * q) >> 4;
        if ((q & 1) == 0) {
#pragma unroll
          for (int u = 0; u < 4; ++u)
            *reinterpret_cast<uint4*>(unit_ptr(U0 + u)) = make_uint4(w[4 * u], w[4 * u + 1], w[4 * u + 2], w[4 * u + 3]);
          *reinterpret_cast<uint2*>(unit_ptr(U0 + 4)) = make_uint2(w[16], w[17]);
        } else {
          *reinterpret_cast<uint2*>(unit_ptr(U0) + 8) = make_uint2(w[0], w[1]);
#pragma unroll
          for (int u = 0; u < 4; ++u)
            *reinterpret_cast<uint4*>(unit_ptr(U0 + 1 + u)) = make_uint4(w[2 + 4 * u], w[3 + 4 * u], w[4 + 4 * u], w[5 + 4 * u]);
        }
      } else {
        float s[K];
#pragma unroll
        for (int k = 0; k < K; ++k) s[k] = fmaf(frac[k], v[b][k + 1].x - v[b][k].x, v[b][k].x);
        const uint32_t c0 = pack2<T>(s[0], s[1]), c1 = pack2<T>(s[2], s[3]), c2 = pack2<T>(s[4], s[5]), c3 = pack2<T>(s[6], s[7]);
        const int Uc = cor_byte >> 4;
        uint8_t* u1 = unit_ptr(Uc + 1);
        if (nq & 1) {   // the K taps start in the second half of unit Uc
          *reinterpret_cast<uint2*>(unit_ptr(Uc) + 8) = make_uint2(c0, c1);
          *reinterpret_cast<uint2*>(u1) = make_uint2(c2, c3);
          *reinterpret_cast<T*>(u1 + 8) = from_f32<T>(s[8]);
        } else {
          *reinterpret_cast<uint4*>(unit_ptr(Uc)) = make_uint4(c0, c1, c2, c3);
          *reinterpret_cast<T*>(u1) = from_f32<T>(s[8]);
        }
      }
    };

    fetch_dx(blockIdx.x);
    float cd[kPasses], cx[kPasses];
#pragma unroll
    for (int i = 0; i < kPasses; ++i) { cd[i] = nd[i]; cx[i] = nx[i]; }
    fetch_dx(blockIdx.x + gridDim.x);
    if (kLcDepth == 2) issue(blockIdx.x, 0, 0, cd[0], cx[0], 0);   // prologue: step 0 of the first tile
    int it = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
#pragma unroll
      for (int st = 0; st < kSteps; ++st) {
        const int h = st / kPasses, pi = st - h * kPasses, b = kLcDepth == 2 ? (st & 1) : 0;
        if (kLcDepth == 1) {
          issue(tile, h, pi, cd[pi], cx[pi], 0);
          if (st + 1 == kSteps) {
#pragma unroll
            for (int i = 0; i < kPasses; ++i) { cd[i] = nd[i]; cx[i] = nx[i]; }   // the next tile's coordinates
            fetch_dx(tile + 2 * gridDim.x);
          }
        } else if (st + 1 < kSteps) {   // the loads of the NEXT step go out first
          const int h1 = (st + 1) / kPasses, pi1 = (st + 1) - h1 * kPasses;
          issue(tile, h1, pi1, cd[pi1], cx[pi1], b ^ 1);
        } else {
#pragma unroll
          for (int i = 0; i < kPasses; ++i) { cd[i] = nd[i]; cx[i] = nx[i]; }   // the next tile's coordinates
          fetch_dx(tile + 2 * gridDim.x);
          issue(tile + gridDim.x, 0, 0, cd[0], cx[0], b ^ 1);
        }
        if (pi == 0) mbar_wait(a_empty + 8 * h, (uint32_t)((it & 1) ^ 1));   // the previous tile's MMAs have read level h
        consume(h, pi, b);
        if (pi == kPasses - 1) {
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes of A -> visible to tcgen05.mma
          __syncwarp();
          if (lane == 0) mbar_arrive(a_full + 8 * h);
        }
      }
    }
  }
  __syncwarp();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

}  // namespace fsb

extern "C" int fsb_lookup_convc1(float* const* geo_levels, float* const* corr_levels, const float* disp, const float* coords,
                                 const void* w, const float* bias, void* out, int B, int Cg, int D, int H, int W, int W2, int L,
                                 int r, int Cout, int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_lookup_convc1");
  using namespace fsb;
  FSB_REQUIRE(geo_levels && corr_levels && disp && coords && w && bias && out, FSB_ERR_ARG, "fsb_lookup_convc1: NULL pointer");
  FSB_REQUIRE(B > 0 && Cg > 0 && D > 0 && H > 0 && W > 0 && W2 > 0, FSB_ERR_ARG, "fsb_lookup_convc1: non-positive size");
  FSB_REQUIRE(dtype == FSB_F16 || dtype == FSB_BF16, FSB_ERR_ARG, "fsb_lookup_convc1: dtype %d (need f16/bf16)", dtype);
  FSB_REQUIRE(L == 1 || L == 2, FSB_ERR_SHAPE, "fsb_lookup_convc1: num_levels=%d (the fused form keeps two levels' operands of "
              "%d channels in shared memory: L <= 2; use fsb_geo_lookup + fsb_conv2d_nhwc)", L, kLcMaxLC * 64);
  FSB_REQUIRE(r == 4, FSB_ERR_SHAPE, "fsb_lookup_convc1: radius=%d (built for 4; use fsb_geo_lookup + fsb_conv2d_nhwc)", r);
  FSB_REQUIRE(Cg % 4 == 0 && Cg <= 28, FSB_ERR_SHAPE, "fsb_lookup_convc1: Cg=%d (four pixels x Cg/4 channel quads per warp: Cg %% 4 == 0, "
              "Cg <= 28; use fsb_geo_lookup + fsb_conv2d_nhwc)", Cg);
  FSB_REQUIRE(Cout == kLcN, FSB_ERR_SHAPE, "fsb_lookup_convc1: Cout=%d (built for %d, core/update.py:56)", Cout, kLcN);
  FSB_REQUIRE((D >> (L - 1)) >= 2 && (W2 >> (L - 1)) >= 2, FSB_ERR_SHAPE,
              "fsb_lookup_convc1: coarsest level shorter than 2 (D=%d, W2=%d, L=%d)", D, W2, L);
  FSB_REQUIRE(aligned16(w) && aligned16(out), FSB_ERR_ALIGN, "fsb_lookup_convc1: w / out must be 16-byte aligned");
  const int Kl = round_up(kLcK * (Cg + 1), 64);   // operand channels per level
  const int Cpad = L * Kl;
  const long long npix = (long long)B * H * W;
  FSB_REQUIRE(npix < (1ll << 31) - 128, FSB_ERR_SHAPE, "fsb_lookup_convc1: too many pixels");
  EncodeTiledFn enc = encode_fn();
  FSB_REQUIRE(enc != nullptr, FSB_ERR_CUDA, "fsb_lookup_convc1: cuTensorMapEncodeTiled not available from the driver");
  LcParams p;
  memset(&p, 0, sizeof(p));
  for (int l = 0; l < L; ++l) {
    FSB_REQUIRE(geo_levels[l] && corr_levels[l], FSB_ERR_ARG, "fsb_lookup_convc1: level %d pointer is NULL", l);
    p.geo.p[l] = geo_levels[l];
    p.cor.p[l] = corr_levels[l];
  }
  {
    const CUtensorMapDataType dt = dtype == FSB_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    cuuint64_t dims[2] = {(cuuint64_t)Cpad, (cuuint64_t)kLcN};
    cuuint64_t strides[1] = {(cuuint64_t)Cpad * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)kLcN};
    cuuint32_t estr[2] = {1, 1};
    CUresult rc = enc(&p.wmap, dt, 2, const_cast<void*>(w), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FSB_REQUIRE(rc == CUDA_SUCCESS, FSB_ERR_CUDA, "fsb_lookup_convc1: cuTensorMapEncodeTiled(w) -> %d", (int)rc);
  }
  p.disp = disp;
  p.coords = coords;
  p.bias = bias;
  p.out = out;
  p.npix = npix;
  p.Cg = Cg;
  p.D = D;
  p.W2 = W2;
  p.L = L;
  p.r = r;
  p.nlc = Kl / 64;
  p.ntiles = (int)((npix + 127) / 128);
  const uint32_t fmt = dtype == FSB_F16 ? 0u : 1u;
  p.idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(kLcN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  const int grid = p.ntiles < sm_count() ? p.ntiles : sm_count();
#define FSB_LAUNCH_LC(T, LT)                                                                                          \
  do {                                                                                                                \
    auto kern = Cg == 28 ? lookup_convc1_kernel<T, LT, 28> : lookup_convc1_kernel<T, LT, 0>;                                                                       \
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kLcSmem);            \
    if (e != cudaSuccess) return cuda_fail(e, "fsb_lookup_convc1: cudaFuncSetAttribute");                             \
    launch_kernel(kern, dim3(grid), dim3(kLcThreads), kLcSmem, (cudaStream_t)stream, p);                              \
  } while (0)
  if (dtype == FSB_F16) {
    if (L == 2) FSB_LAUNCH_LC(__half, 2); else FSB_LAUNCH_LC(__half, 1);
  } else {
    if (L == 2) FSB_LAUNCH_LC(__nv_bfloat16, 2); else FSB_LAUNCH_LC(__nv_bfloat16, 1);
  }
#undef FSB_LAUNCH_LC
  FSB_CHECK_LAUNCH("fsb_lookup_convc1");
  return FSB_OK;
}
