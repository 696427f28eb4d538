// Kernel 4b: the pointwise MLP of an EdgeNeXt block as ONE tcgen05 kernel.
//
// Replaces, for one EdgeNextConvEncoder block of the disparity head (core/submodule.py:565-591 with
// norm=None, called from DispHead, core/update.py:20-32), everything after the depthwise convolution:
//     h   = GELU(t @ W1^T + b1)          pwconv1, 128 -> 512   (core/submodule.py:584-585)
//     y   = h @ W2^T + b2                pwconv2, 512 -> 128   (:586)
//     out = res + gamma * y              layer scale + residual (:587-590)
// As two convolution launches the 512-wide hidden tensor (79 MB per 4 pairs) went to HBM and back and
// the GELU ran as the serial epilogue of a 512-cycle GEMM.  Here a persistent CTA (one per SM) keeps
// the hidden activations on chip:
//   tile = 128 pixels; the hidden features are processed in four chunks of 128:
//     G1(j): acc1[j&1] (TMEM, 128 cols) = X (smem, TMA) x W1[128j..128j+127,:]^T
//     E1(j): 16 epilogue warps: tcgen05.ld -> +b1 -> GELU -> 16-bit -> H[j&1] in shared memory, written
//            directly in the K-major SWIZZLE_128B layout tcgen05.mma reads (fence.proxy.async)
//     G2(j): acc2[t&1] (TMEM, 128 cols) += H[j&1] x W2[:,128j..128j+127]^T
//   MMA issue order G1(0) G1(1) G2(0) G1(2) G2(1) G1(3) G2(2) G2(3): the tensor pipe works on chunk
//   j+1 while the epilogue warps run the GELU of chunk j; both accumulators are double-buffered, so the
//   final epilogue (+b2, *gamma, +res, 16-bit store) of tile t overlaps the GEMMs of tile t+1.
//   W1 / W2 (256 KB) stream from L2 through a 5-stage ring of 16 KB tiles in consumption order.
// The kernel is bound by the GELU arithmetic (512 per pixel), which is what should bound this op; the GELU runs on
// the packed fp32 datapath with one MUFU per element (gelu2_f, umma.cuh).
#include "common.cuh"
#include "umma.cuh"

namespace fsb {

constexpr int kMlpC = 128;             // block channels
constexpr int kMlpF = 512;             // hidden features (expansion 4)
constexpr int kMlpChunks = kMlpF / 128;
constexpr int kMlpEpiWarps = 16;
constexpr int kMlpThreads = (3 + kMlpEpiWarps) * 32;  // warp 0: X producer, 1: MMA, 2: W producer, 3..18: epilogue
constexpr int kMlpWStages = 5;
constexpr int kMlpTile = 128 * 64 * 2;  // 16 KB: 128 rows x 64 16-bit elements, SWIZZLE_128B

struct alignas(64) MlpParams {
  CUtensorMap xmap, w1map, w2map;
  const float* b1;
  const float* b2;
  const float* gamma;
  const void* res;
  void* out;
  long long npix;
  int ntiles;
  uint32_t idesc;
};

// shared-memory carve-up (offsets from the 1024-byte aligned base)
constexpr uint32_t kMlpOffX = 0;                              // 2 slots x 2 sub-tiles
constexpr uint32_t kMlpOffW = kMlpOffX + 2 * 2 * kMlpTile;    // ring
constexpr uint32_t kMlpOffH = kMlpOffW + kMlpWStages * kMlpTile;  // 2 buffers x 2 sub-tiles
constexpr uint32_t kMlpSmem = kMlpOffH + 2 * 2 * kMlpTile + 1024;

template <typename T>
__global__ void __launch_bounds__(kMlpThreads, 1)
edgenext_mlp_kernel(const __grid_constant__ MlpParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // barriers: x_full[2] x_empty[2] w_full[5] w_empty[5] a1_full[2] a1_empty[2] h_full[2] h_empty[2] a2_full[2] a2_empty[2]
  __shared__ __align__(8) uint64_t bars[4 + 2 * kMlpWStages + 12];
  __shared__ uint32_t tmem_slot;
  __shared__ __align__(16) float b1_s[kMlpF];
  __shared__ __align__(16) float b2_s[kMlpC], g_s[kMlpC];

  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t x_full = smem_u32(&bars[0]), x_empty = smem_u32(&bars[2]);
  const uint32_t w_full = smem_u32(&bars[4]), w_empty = smem_u32(&bars[4 + kMlpWStages]);
  const uint32_t a1_full = smem_u32(&bars[4 + 2 * kMlpWStages]), a1_empty = a1_full + 16;
  const uint32_t h_full = a1_full + 32, h_empty = a1_full + 48;
  const uint32_t a2_full = a1_full + 64, a2_empty = a1_full + 80;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(x_full + 8 * i, 1);
      mbar_init(x_empty + 8 * i, 1);
      mbar_init(a1_full + 8 * i, 1);
      mbar_init(a1_empty + 8 * i, kMlpEpiWarps);
      mbar_init(h_full + 8 * i, kMlpEpiWarps);
      mbar_init(h_empty + 8 * i, 1);
      mbar_init(a2_full + 8 * i, 1);
      mbar_init(a2_empty + 8 * i, kMlpEpiWarps);
    }
    for (int i = 0; i < kMlpWStages; ++i) {
      mbar_init(w_full + 8 * i, 1);
      mbar_init(w_empty + 8 * i, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    tma_prefetch_desc(&p.xmap);
  }
  if (warp == 2 && lane == 0) {
    tma_prefetch_desc(&p.w1map);
    tma_prefetch_desc(&p.w2map);
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (warp >= 3) {
    const int i = threadIdx.x - 96;  // 0..511
    b1_s[i] = __ldg(p.b1 + i);
    if (i < kMlpC) {
      b2_s[i] = __ldg(p.b2 + i);
      g_s[i] = __ldg(p.gamma + i);
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;  // acc1: cols [0,256), acc2: cols [256,512)
  pdl_launch_dependents();
  if (warp != 2) pdl_wait();  // the weight producer only reads parameters no kernel writes

  if (warp == 0) {
    // ================================================================= X producer
    int it = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
      const int s = it & 1;
      const uint32_t use = (uint32_t)(it >> 1);
      mbar_wait(x_empty + 8 * s, (use & 1) ^ 1);
      __syncwarp();
      if (elect_one()) {
        mbar_expect_tx(x_full + 8 * s, 2 * kMlpTile);
        const uint32_t dst = base + kMlpOffX + s * 2 * kMlpTile;
        tma_load_2d(dst, &p.xmap, x_full + 8 * s, 0, tile * 128);
        tma_load_2d(dst + kMlpTile, &p.xmap, x_full + 8 * s, 64, tile * 128);
      }
    }
  } else if (warp == 2) {
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
            mbar_expect_tx(w_full + 8 * s, kMlpTile);
            const uint32_t dst = base + kMlpOffW + s * kMlpTile;
            if (is2) tma_load_2d(dst, &p.w2map, w_full + 8 * s, j * 128 + k * 64, 0);
            else     tma_load_2d(dst, &p.w1map, w_full + 8 * s, k * 64, j * 128);
          }
          if (++s == kMlpWStages) { s = 0; ph ^= 1; first_round = false; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================================================= MMA issuer
    int ws = 0;
    uint32_t wph = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
      const int xs = it & 1;
      const uint32_t xuse = (uint32_t)(it >> 1);
      const uint32_t acc2 = tmem_base + 256u + (uint32_t)(it & 1) * 128u;
      mbar_wait(x_full + 8 * xs, xuse & 1);
      mbar_wait(a2_empty + 8 * (it & 1), (xuse & 1) ^ 1);  // final epilogue of tile it-2 drained this buffer
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
          a_addr = base + kMlpOffX + xs * 2 * kMlpTile;
          d = tmem_base + (uint32_t)b * 128u;
        } else {
          mbar_wait(h_full + 8 * b, use & 1);  // E1(j) wrote H[b]
          a_addr = base + kMlpOffH + b * 2 * kMlpTile;
          d = acc2;
        }
        tc_fence_after();
#pragma unroll 1
        for (int k = 0; k < 2; ++k) {
          mbar_wait(w_full + 8 * ws, wph);
          tc_fence_after();
          __syncwarp();
          const uint64_t adesc = umma_desc_k128(a_addr + k * kMlpTile);
          const uint64_t bdesc = umma_desc_k128(base + kMlpOffW + ws * kMlpTile);
          const uint32_t accum = (!is2) ? (uint32_t)(k > 0) : (uint32_t)(j > 0 || k > 0);
          if (elect_one()) {
            umma_f16(d, adesc, bdesc, p.idesc, accum);
            umma_f16(d, adesc + 2, bdesc + 2, p.idesc, 1);
            umma_f16(d, adesc + 4, bdesc + 4, p.idesc, 1);
            umma_f16(d, adesc + 6, bdesc + 6, p.idesc, 1);
            umma_commit(w_empty + 8 * ws);
          }
          if (++ws == kMlpWStages) { ws = 0; wph ^= 1; }
        }
        if (elect_one()) {
          if (!is2) {
            umma_commit(a1_full + 8 * b);
            if (j == 3) umma_commit(x_empty + 8 * xs);  // X slot free once the last G1 has read it
          } else {
            umma_commit(h_empty + 8 * b);
            if (j == 3) umma_commit(a2_full + 8 * (it & 1));
          }
        }
      }
    }
  } else {
    // ================================================================= epilogue warps
    // warp e = warp - 3: TMEM lane quarter (warp & 3) -- the hardware rule -- and 32-column group (e >> 2)
    const int e = warp - 3;
    const int q = warp & 3, g = e >> 2;
    const int r = q * 32 + lane;  // pixel row of the tile
    const uint32_t lane_addr = tmem_base + ((uint32_t)(q * 32) << 16);
    // H sub-tile (g >> 1), 16-byte units (g & 1) * 4 .. + 3 of row r, SWIZZLE_128B: unit ^ (r & 7)
    uint8_t* hrow = gbase + kMlpOffH + (g >> 1) * kMlpTile + r * 128;
    const int u0 = (g & 1) * 4, sw = r & 7;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x, ++it) {
#pragma unroll 1
      for (int j = 0; j < kMlpChunks; ++j) {
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
        uint8_t* hb = hrow + b * 2 * kMlpTile;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          *reinterpret_cast<uint4*>(hb + (((u0 + i) ^ sw) << 4)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to tcgen05.mma
        __syncwarp();
        if (lane == 0) mbar_arrive(h_full + 8 * b);
      }
      // ---- final epilogue of the tile: out = res + gamma * (acc2 + b2)
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
      const long long pix = (long long)tile * 128 + r;
      if (pix < p.npix) {
        float f[32], rr[32];
        load32<T>(reinterpret_cast<const T*>(p.res) + pix * kMlpC + g * 32, rr);
        // gamma / b2 of the warp's 32 columns as 16-byte broadcast loads (8 + 8 LDS.128 instead of 64 LDS.32)
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
        store32<T>(reinterpret_cast<T*>(p.out) + pix * kMlpC + g * 32, f);
      }
    }
  }
  __syncwarp();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

}  // namespace fsb

// ------------------------------------------------------------------------------------------------------
extern "C" int fsb_edgenext_mlp(const void* x, const void* w1, const float* b1, const void* w2, const float* b2,
                                const float* gamma, const void* res, void* out, long long npix, int C, int F, int dtype,
                                fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_edgenext_mlp");
  using namespace fsb;
  FSB_REQUIRE(x && w1 && b1 && w2 && b2 && gamma && res && out, FSB_ERR_ARG, "fsb_edgenext_mlp: NULL pointer");
  FSB_REQUIRE(dtype == FSB_F16 || dtype == FSB_BF16, FSB_ERR_ARG, "fsb_edgenext_mlp: dtype %d (need f16/bf16)", dtype);
  FSB_REQUIRE(C == kMlpC && F == kMlpF, FSB_ERR_SHAPE, "fsb_edgenext_mlp: C=%d F=%d (built for %d -> %d -> %d)", C, F, kMlpC,
              kMlpF, kMlpC);
  FSB_REQUIRE(npix > 0 && npix < (1ll << 31) - 128, FSB_ERR_SHAPE, "fsb_edgenext_mlp: npix=%lld", npix);
  FSB_REQUIRE(aligned16(x) && aligned16(w1) && aligned16(w2) && aligned16(res) && aligned16(out), FSB_ERR_ALIGN,
              "fsb_edgenext_mlp: pointers must be 16-byte aligned");
  EncodeTiledFn enc = encode_fn();
  FSB_REQUIRE(enc != nullptr, FSB_ERR_CUDA, "fsb_edgenext_mlp: cuTensorMapEncodeTiled not available from the driver");
  MlpParams p;
  memset(&p, 0, sizeof(p));
  const CUtensorMapDataType dt = dtype == FSB_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  auto map2d = [&](CUtensorMap* m, const void* ptr, cuuint64_t inner, cuuint64_t rows, const char* what) -> int {
    cuuint64_t dims[2] = {inner, rows};
    cuuint64_t strides[1] = {inner * 2};
    cuuint32_t box[2] = {64, 128};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FSB_REQUIRE(r == CUDA_SUCCESS, FSB_ERR_CUDA, "fsb_edgenext_mlp: cuTensorMapEncodeTiled(%s) -> %d", what, (int)r);
    return FSB_OK;
  };
  int rc;
  if ((rc = map2d(&p.xmap, x, kMlpC, (cuuint64_t)npix, "x")) != FSB_OK) return rc;
  if ((rc = map2d(&p.w1map, w1, kMlpC, kMlpF, "w1")) != FSB_OK) return rc;
  if ((rc = map2d(&p.w2map, w2, kMlpF, kMlpC, "w2")) != FSB_OK) return rc;
  p.b1 = b1;
  p.b2 = b2;
  p.gamma = gamma;
  p.res = res;
  p.out = out;
  p.npix = npix;
  p.ntiles = (int)((npix + 127) / 128);
  const uint32_t fmt = dtype == FSB_F16 ? 0u : 1u;
  p.idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(128 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  auto kern = dtype == FSB_F16 ? edgenext_mlp_kernel<__half> : edgenext_mlp_kernel<__nv_bfloat16>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMlpSmem);
  if (e != cudaSuccess) return cuda_fail(e, "fsb_edgenext_mlp: cudaFuncSetAttribute");
  const int grid = p.ntiles < sm_count() ? p.ntiles : sm_count();
  launch_kernel(kern, dim3(grid), dim3(kMlpThreads), kMlpSmem, (cudaStream_t)stream, p);
  FSB_CHECK_LAUNCH("fsb_edgenext_mlp");
  return FSB_OK;
}
