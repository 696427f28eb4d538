// Micro-benchmark: how fast can ONE CTA per SM stream operand tiles into shared memory?
//   mode 0: 2-D tensor map, box 64 x ROWS (rows of 128 B, row pitch `pitch` bytes), SWIZZLE_128B
//   mode 1: cp.async.bulk (1-D, contiguous ROWS*128 bytes), no tensor map
//   mode 2: 4-D NHWC activation map, box (64 ch, 16, 8, 1) at shifted coordinates
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/tma_bench tools/tma_bench.cu
// Run:   gpurun_out/tma_bench
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t b) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(b) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t ph) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(ph) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t ph) {
  long long t0 = clock64();
  while (!mbar_try(bar, ph)) { if (clock64() - t0 > 2000000000ll) __trap(); }
}

struct P {
  CUtensorMap map2, map4;
  const uint8_t* flat;
  int mode, rows, steps, stages, ksteps_per_row, H, W, Cchunks;
  long long* cycles;
};

__global__ void __launch_bounds__(64) stream_kernel(const __grid_constant__ P p) {
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ __align__(8) uint64_t bars[32];
  const uint32_t base = (smem_u32(sm) + 1023u) & ~1023u;
  const uint32_t bytes = (uint32_t)p.rows * 128u;
  const int S = p.stages;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&bars[s]), 1); mbar_init(smem_u32(&bars[16 + s]), 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const long long t0 = clock64();
  if (threadIdx.x == 0) {
    for (int i = 0; i < p.steps; ++i) {
      const int s = i % S;
      if (i >= S) mbar_wait(smem_u32(&bars[16 + s]), ((i / S) - 1) & 1);
      const uint32_t full = smem_u32(&bars[s]);
      const uint32_t dst = base + s * bytes;
      mbar_expect_tx(full, bytes);
      if (p.mode == 0) {
        const int c0 = (i % p.ksteps_per_row) * 64, c1 = ((blockIdx.x * 7 + i / p.ksteps_per_row) % (512 / p.rows)) * p.rows;
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(dst), "l"(&p.map2), "r"(full), "r"(c0), "r"(c1) : "memory");
      } else if (p.mode == 1) {
        const uint8_t* src = p.flat + ((size_t)((blockIdx.x * 131 + i) % 2048)) * bytes;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(src), "r"(bytes), "r"(full) : "memory");
      } else {
        const int tx = blockIdx.x % (p.W / 16), ty = (blockIdx.x / (p.W / 16)) % (p.H / 8);
        const int tap = i % 9, ch = (i / 9) % p.Cchunks;
        asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                     ::"r"(dst), "l"(&p.map4), "r"(full), "r"(ch * 64), "r"(tx * 16 + tap % 3 - 1), "r"(ty * 8 + tap / 3 - 1), "r"(0) : "memory");
      }
    }
  } else if (threadIdx.x == 32) {
    for (int i = 0; i < p.steps; ++i) {
      const int s = i % S;
      mbar_wait(smem_u32(&bars[s]), (i / S) & 1);
      mbar_arrive(smem_u32(&bars[16 + s]));
    }
    p.cycles[blockIdx.x] = clock64() - t0;
  }
}


// burst test: `nops` TMA ops on ONE mbarrier per round, one wait per round (no producer/consumer handshake).
// waitmode 0: try_wait loop, 1: test_wait spin
__global__ void __launch_bounds__(32) burst_kernel(const __grid_constant__ P p, int nops, int rounds, int waitmode) {
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = (smem_u32(sm) + 1023u) & ~1023u;
  const uint32_t bytes = (uint32_t)p.rows * 128u;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const uint32_t full = smem_u32(&bar);
    const long long t0 = clock64();
    for (int r = 0; r < rounds; ++r) {
      mbar_expect_tx(full, bytes * nops);
      for (int j = 0; j < nops; ++j) {
        const int i = r * nops + j;
        const uint32_t dst = base + j * bytes;
        if (p.mode == 0) {
          const int c0 = (i % p.ksteps_per_row) * 64, c1 = ((blockIdx.x * 7 + i / p.ksteps_per_row) % (512 / p.rows)) * p.rows;
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                       ::"r"(dst), "l"(&p.map2), "r"(full), "r"(c0), "r"(c1) : "memory");
        } else {
          const uint8_t* src = p.flat + ((size_t)((blockIdx.x * 131 + i) % 2048)) * bytes;
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"(dst), "l"(src), "r"(bytes), "r"(full) : "memory");
        }
      }
      if (waitmode == 0) {
        mbar_wait(full, r & 1);
      } else {
        uint32_t ok = 0;
        while (!ok) {
          asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(full), "r"((uint32_t)(r & 1)) : "memory");
        }
      }
    }
    p.cycles[blockIdx.x] = clock64() - t0;
  }
}

// single-thread software pipeline over S barriers (no consumer thread): does issuing to different mbarriers pipeline?
__global__ void __launch_bounds__(32) selfpipe_kernel(const __grid_constant__ P p) {
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ __align__(8) uint64_t bars[16];
  const uint32_t base = (smem_u32(sm) + 1023u) & ~1023u;
  const uint32_t bytes = (uint32_t)p.rows * 128u;
  const int S = p.stages;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) mbar_init(smem_u32(&bars[s]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const long long t0 = clock64();
    for (int i = 0; i < p.steps + S; ++i) {
      const int s = i % S;
      const uint32_t full = smem_u32(&bars[s]);
      if (i >= S) mbar_wait(full, ((i / S) - 1) & 1);
      if (i >= p.steps) continue;
      mbar_expect_tx(full, bytes);
      const uint32_t dst = base + s * bytes;
      if (p.mode == 0) {
        const int c0 = (i % p.ksteps_per_row) * 64, c1 = ((blockIdx.x * 7 + i / p.ksteps_per_row) % (512 / p.rows)) * p.rows;
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(dst), "l"(&p.map2), "r"(full), "r"(c0), "r"(c1) : "memory");
      } else {
        const uint8_t* src = p.flat + ((size_t)((blockIdx.x * 131 + i) % 2048)) * bytes;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(src), "r"(bytes), "r"(full) : "memory");
      }
    }
    p.cycles[blockIdx.x] = clock64() - t0;
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
  const int K = 4608, R = 512;                 // weight matrix [R][K] fp16
  const int H = 120, W = 160, C = 512;         // activation NHWC fp16
  uint8_t *w, *act, *flat;
  CK(cudaMalloc(&w, (size_t)R * K * 2));
  CK(cudaMalloc(&act, (size_t)H * W * C * 2));
  CK(cudaMalloc(&flat, (size_t)2048 * 32768));
  CK(cudaMemset(w, 0, (size_t)R * K * 2));
  CK(cudaMemset(act, 0, (size_t)H * W * C * 2));
  CK(cudaMemset(flat, 0, (size_t)2048 * 32768));
  long long* cyc;
  CK(cudaMalloc(&cyc, 148 * 8));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));

  for (int mode = 0; mode < 3; ++mode)
    for (int rows : {128})
      for (int stages : {2, 8})
        for (int grid : {1, 148}) {
          if (mode == 2 && rows != 128) continue;
          if ((size_t)stages * rows * 128 > 190 * 1024) continue;
          P p{};
          p.mode = mode; p.rows = rows; p.steps = 288; p.stages = stages; p.ksteps_per_row = K / 64;
          p.H = H; p.W = W; p.Cchunks = C / 64; p.flat = flat; p.cycles = cyc;
          {
            cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)R};
            cuuint64_t strides[1] = {(cuuint64_t)K * 2};
            cuuint32_t box[2] = {64, (cuuint32_t)(rows > 256 ? 256 : rows)};
            cuuint32_t es[2] = {1, 1};
            CUresult r = enc(&p.map2, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, w, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { printf("encode2 %d\n", (int)r); return 1; }
          }
          {
            cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, 1};
            cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)C * 2 * W, (cuuint64_t)C * 2 * W * H};
            cuuint32_t box[4] = {64, 16, 8, 1};
            cuuint32_t es[4] = {1, 1, 1, 1};
            CUresult r = enc(&p.map4, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, act, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { printf("encode4 %d\n", (int)r); return 1; }
          }
          const size_t smem = (size_t)stages * rows * 128 + 1024;
          for (int rep = 0; rep < 2; ++rep) stream_kernel<<<grid, 64, smem>>>(p);
          CK(cudaDeviceSynchronize());
          cudaEventRecord(e0);
          stream_kernel<<<grid, 64, smem>>>(p);
          cudaEventRecord(e1);
          CK(cudaDeviceSynchronize());
          float ms;
          cudaEventElapsedTime(&ms, e0, e1);
          long long h[148];
          CK(cudaMemcpy(h, cyc, grid * 8, cudaMemcpyDeviceToHost));
          long long mx = 0;
          for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
          const double bytes = (double)p.steps * rows * 128;
          printf("mode %d rows %3d stages %d grid %3d : %7.1f cyc/step  %6.1f B/cyc/SM  %7.2f us  %7.1f GB/s total\n", mode, rows,
                 stages, grid, (double)mx / p.steps, bytes / mx, ms * 1e3, bytes * grid / (ms * 1e-3) / 1e9);
        }
  CK(cudaFuncSetAttribute(selfpipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  for (int mode = 0; mode < 2; ++mode)
    for (int stages : {1, 2, 4, 8})
      for (int grid : {1, 148}) {
        P p{};
        const int rows = 128;
        p.mode = mode; p.rows = rows; p.steps = 256; p.stages = stages; p.ksteps_per_row = K / 64; p.flat = flat; p.cycles = cyc;
        cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)R};
        cuuint64_t strides[1] = {(cuuint64_t)K * 2};
        cuuint32_t box[2] = {64, (cuuint32_t)rows};
        cuuint32_t es[2] = {1, 1};
        enc(&p.map2, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, w, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        const size_t smem = (size_t)stages * rows * 128 + 1024;
        selfpipe_kernel<<<grid, 32, smem>>>(p);
        CK(cudaDeviceSynchronize());
        selfpipe_kernel<<<grid, 32, smem>>>(p);
        CK(cudaDeviceSynchronize());
        long long h[148];
        CK(cudaMemcpy(h, cyc, grid * 8, cudaMemcpyDeviceToHost));
        long long mx = 0;
        for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
        printf("selfpipe mode %d stages %d grid %3d : %7.1f cyc/step  %6.1f B/cyc/SM\n", mode, stages, grid, (double)mx / p.steps,
               (double)p.steps * rows * 128 / mx);
      }
  CK(cudaFuncSetAttribute(burst_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  for (int mode = 0; mode < 2; ++mode)
    for (int waitmode = 0; waitmode < 2; ++waitmode)
      for (int nops : {1, 2, 4, 8})
        for (int grid : {1, 148}) {
          P p{};
          const int rows = 128;
          p.mode = mode; p.rows = rows; p.ksteps_per_row = K / 64; p.flat = flat; p.cycles = cyc;
          cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)R};
          cuuint64_t strides[1] = {(cuuint64_t)K * 2};
          cuuint32_t box[2] = {64, (cuuint32_t)rows};
          cuuint32_t es[2] = {1, 1};
          enc(&p.map2, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, w, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
          const int rounds = 64;
          const size_t smem = (size_t)nops * rows * 128 + 1024;
          burst_kernel<<<grid, 32, smem>>>(p, nops, rounds, waitmode);
          CK(cudaDeviceSynchronize());
          burst_kernel<<<grid, 32, smem>>>(p, nops, rounds, waitmode);
          CK(cudaDeviceSynchronize());
          long long h[148];
          CK(cudaMemcpy(h, cyc, grid * 8, cudaMemcpyDeviceToHost));
          long long mx = 0;
          for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
          printf("burst mode %d wait %d nops %d grid %3d : %7.1f cyc/round  %6.1f B/cyc/SM\n", mode, waitmode, nops, grid,
                 (double)mx / rounds, (double)rounds * nops * rows * 128 / mx);
        }
  return 0;
}
