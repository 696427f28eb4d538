// FP32 FMA issue rates on sm_100a: scalar FFMA, packed FFMA2 (fma.rn.f32x2) and a 1:2 mix, 16 independent chains per
// thread, 8 warps x 4 CTAs per SM.  Prints FMA / clk / SM from the kernel's own clock64 span.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fma_bench.bin tools/fma_bench.cu && tools/fma_bench.bin
#include <cuda_runtime.h>
#include <cstdio>
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, long long* cyc) {
  float2 a[8];
  for (int i = 0; i < 8; ++i) a[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
  const float2 w = make_float2(1.0001f, 0.9999f), c = make_float2(1e-3f, -1e-3f);
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) { a[i].x = fmaf(a[i].x, w.x, c.x); a[i].y = fmaf(a[i].y, w.y, c.y); }
      if (MODE == 1) { a[i] = __ffma2_rn(a[i], w, c); }
      if (MODE == 2) { if (i & 1) a[i] = __ffma2_rn(a[i], w, c); else { a[i].x = fmaf(a[i].x, w.x, c.x); a[i].y = fmaf(a[i].y, w.y, c.y); } }
    }
  }
  const long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
int main() {
  int dev = 0, sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int ctas = sms * 4, iters = 20000;
  float* out; long long* cyc;
  cudaMalloc(&out, ctas * 256 * sizeof(float));
  cudaMallocManaged(&cyc, ctas * sizeof(long long));
  const char* names[3] = {"scalar FFMA", "packed FFMA2", "mix 1 FFMA2 : 2 FFMA"};
  for (int m = 0; m < 3; ++m) {
    for (int rep = 0; rep < 2; ++rep) {
      if (m == 0) k<0><<<ctas, 256>>>(out, iters, cyc);
      if (m == 1) k<1><<<ctas, 256>>>(out, iters, cyc);
      if (m == 2) k<2><<<ctas, 256>>>(out, iters, cyc);
      cudaDeviceSynchronize();
    }
    long long mx = 0;
    for (int i = 0; i < ctas; ++i) mx = cyc[i] > mx ? cyc[i] : mx;
    const double fma_per_sm = 4.0 * 256 * 16.0 * iters;  // 4 CTAs x 256 threads x 16 FMAs per iteration
    printf("%-22s %8.1f FMA/clk/SM  (%lld cycles)\n", names[m], fma_per_sm / (double)mx, mx);
  }
  return 0;
}
