// TMA feed-rate micro-benchmark 2: software-pipelined producer(s), `ops` TMA operations per mbarrier phase.
//   mode 0: 2-D tensor map box 64 x rows (128-byte rows, SWIZZLE_128B); mode 1: cp.async.bulk contiguous rows*128 B
//   nwarps issuing warps per CTA, each with its own ring of S barriers / buffers.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_bench2.bin tools/tma_bench2.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t b) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(b) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t ph) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(ph) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t ph) {
  long long t0 = clock64();
  while (!mbar_try(bar, ph)) { if (clock64() - t0 > 2000000000ll) __trap(); }
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}
struct P {
  CUtensorMap map2;
  const uint8_t* flat;
  int mode, rows, ops, steps, stages, kper;
  long long* cycles;
};
__global__ void __launch_bounds__(128) pipe_kernel(const __grid_constant__ P p) {
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ __align__(8) uint64_t bars[4][16];
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const uint32_t opbytes = (uint32_t)p.rows * 128u;
  const uint32_t stbytes = opbytes * p.ops;
  const int S = p.stages;
  const uint32_t base = ((smem_u32(sm) + 1023u) & ~1023u) + warp * S * stbytes;
  {
    if ((threadIdx.x & 31) == 0) {
      for (int s = 0; s < S; ++s) mbar_init(smem_u32(&bars[warp][s]), 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const long long t0 = clock64();
    // lean loop: no integer division (running stage index / phase bit, incrementing coordinates)
    int s = 0;
    uint32_t ph = 0;
    int c0 = 0, c1 = 0;
    const uint8_t* src = p.flat + (size_t)((blockIdx.x * 4 + warp) % 64) * 65536;
    const uint8_t* src_end = p.flat + (size_t)1024 * 65536 - 65536;
    const int nops = p.ops, steps = p.steps, mode = p.mode, rows = p.rows;
    for (int i = 0; i < steps + S; ++i) {
      const uint32_t full = smem_u32(&bars[warp][s]);
      if (i >= S) {
        while (!mbar_try(full, ph ^ 1)) {}
      }
      __syncwarp();
      if (i < steps && elect_one()) {
        mbar_expect_tx(full, stbytes);
        uint32_t dst = base + s * stbytes;
        for (int j = 0; j < nops; ++j) {
          if (mode == 0) {
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         ::"r"(dst), "l"(&p.map2), "r"(full), "r"(c0), "r"(c1) : "memory");
            c0 += 64;
            if (c0 >= 4608) { c0 = 0; c1 += rows; if (c1 >= 512) c1 = 0; }
          } else {
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst), "l"(src), "r"(opbytes), "r"(full) : "memory");
            src += opbytes;
            if (src >= src_end) src = p.flat;
          }
          dst += opbytes;
        }
      }
      if (++s == S) { s = 0; ph ^= 1; }
    }
    if ((threadIdx.x & 31) == 0) p.cycles[blockIdx.x * 4 + warp] = clock64() - t0;
  }
}
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main() {
  void* f = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q));
  EncodeFn enc = (EncodeFn)f;
  const int K = 4608, R = 512;
  uint8_t *w, *flat;
  CK(cudaMalloc(&w, (size_t)R * K * 2));
  CK(cudaMalloc(&flat, (size_t)1024 * 65536));
  CK(cudaMemset(w, 0, (size_t)R * K * 2));
  CK(cudaMemset(flat, 0, (size_t)1024 * 65536));
  long long* cyc;
  CK(cudaMalloc(&cyc, 148 * 4 * 4 * 8));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  CK(cudaFuncSetAttribute(pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
  struct Cfg { int mode, rows, ops, stages, nwarps, cps; };
  Cfg cfgs[] = {
    {0, 128, 1, 4, 1, 1}, {0, 128, 2, 4, 1, 1}, {0, 128, 4, 2, 1, 1}, {0, 64, 2, 4, 1, 1}, {0, 256, 1, 4, 1, 1},
    {0, 128, 1, 4, 2, 1}, {0, 128, 1, 2, 4, 1}, {0, 128, 2, 2, 2, 1},
    {0, 128, 1, 4, 1, 2}, {0, 128, 1, 2, 1, 4}, {0, 128, 2, 2, 1, 2},
    {1, 128, 1, 4, 1, 1}, {1, 128, 2, 4, 1, 1}, {1, 256, 1, 4, 1, 1}, {1, 384, 1, 3, 1, 1}, {1, 512, 1, 3, 1, 1},
    {1, 128, 1, 4, 2, 1}, {1, 128, 1, 2, 4, 1}, {1, 128, 1, 4, 1, 2}, {1, 256, 1, 2, 1, 2}, {1, 64, 1, 4, 1, 1}, {1, 64, 4, 4, 1, 1},
  };
  for (auto& c : cfgs) {
    P p{};
    p.mode = c.mode; p.rows = c.rows; p.ops = c.ops; p.steps = 256 / c.ops; p.stages = c.stages; p.kper = K / 64; p.flat = flat; p.cycles = cyc;
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)R};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)(c.rows > 256 ? 256 : c.rows)};
    cuuint32_t es[2] = {1, 1};
    enc(&p.map2, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, w, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    size_t smem = (size_t)c.nwarps * c.stages * c.ops * c.rows * 128 + 1024;
    if (c.cps == 1 && smem < 120 * 1024) smem = 120 * 1024;        // force one CTA per SM
    if (c.cps == 2 && smem < 80 * 1024) smem = 80 * 1024;
    if (smem > 220 * 1024) { printf("skip (smem)\n"); continue; }
    const int grid = 148 * c.cps;
    for (int rep = 0; rep < 2; ++rep) pipe_kernel<<<grid, 32 * c.nwarps, smem>>>(p);
    CK(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    pipe_kernel<<<grid, 32 * c.nwarps, smem>>>(p);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    long long h[148 * 16];
    CK(cudaMemcpy(h, cyc, sizeof(long long) * grid * 4, cudaMemcpyDeviceToHost));
    long long mx = 0;
    for (int i = 0; i < grid; ++i) for (int wv = 0; wv < c.nwarps; ++wv) mx = h[i * 4 + wv] > mx ? h[i * 4 + wv] : mx;
    const double bytes_sm = (double)p.steps * c.ops * c.rows * 128 * c.nwarps * c.cps;
    printf("mode %d rows %3d ops/bar %d stages %d warps %d ctas/sm %d : %7.1f cyc/op (per issuer)  %6.1f B/cyc/SM  %6.1f us  %6.0f GB/s\n",
           c.mode, c.rows, c.ops, c.stages, c.nwarps, c.cps, (double)mx / (p.steps * c.ops), bytes_sm / mx, ms * 1e3,
           bytes_sm * 148 / (ms * 1e-3) / 1e9);
  }
  return 0;
}
