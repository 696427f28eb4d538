// tcgen05 / TMA / mbarrier PTX wrappers, the UMMA shared-memory descriptor and the epilogue math shared by the
// tensor-core kernels of libfsb200 (conv.cu: update-block convolutions, mlp.cu: fused EdgeNeXt MLP).
#pragma once
#include "common.cuh"

#include <cuda.h>

namespace fsb {

// ------------------------------------------------------------------------------------ PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// A wait that cannot hang the GPU: after ~2 s of SM clocks the kernel traps (a pipeline bug
// then shows up as a launch failure instead of a dead box).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 255u) == 0 && clock64() - t0 > 4000000000ll) __trap();
  }
}

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}

// One lane of a converged warp (elect.sync).  The role loops below are executed by the WHOLE warp with
// warp-uniform state (the warp index comes from a shuffle, so the compiler knows the role branches are
// uniform) and only the TMA / MMA / commit instructions sit under this predicate.  Written the other
// way round -- `if (lane == 0) { loop }` -- the compiler cannot keep the operands of the uniform-datapath
// instructions (UTMALDG, UTCHMMA) in uniform registers and wraps every one of them in an
// ELECT / R2UR / BRA.U.ANY "waterfall" loop, which made instruction issue the bottleneck of the kernel.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile, 128-byte swizzle: rows of 128 B, 8-row atoms of 1024 B (SBO), descriptor
// version 1 (Blackwell), layout type 2 (SWIZZLE_128B).  `addr` must be 1024-byte aligned; stepping
// 16 elements along K inside the atom adds 32 B to the start address.
__device__ __forceinline__ uint64_t umma_desc_k128(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFFu) >> 4);  // start address, bits [0,14)
  d |= (uint64_t)1 << 16;                   // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;         // stride byte offset between 8-row atoms
  d |= (uint64_t)1 << 46;                   // descriptor version
  d |= (uint64_t)2 << 61;                   // SWIZZLE_128B
  return d;
}

// ------------------------------------------------------------------------------- epilogue math
// GELU(x) = 0.5 x (1 + erf(x / sqrt 2)) with erf by Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7, far below the
// 16-bit output rounding), written for the instruction count -- the 4 x 128 x 512 GELUs per pixel tile of the
// EdgeNeXt MLP are pure epilogue work: bare MUFU.RCP / MUFU.EX2 (no range fix-ups: 1 + 0.33|x| >= 1 and the
// exponent is <= 0), constants folded (|x|/sqrt2, -log2(e)/2), sign by a bit operation: 16 instructions.
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// sigmoid / tanh on the bare MUFU units: exp(-x) = inf gives rcp(inf) = 0, exp(2x) = inf gives 1 - 0
__device__ __forceinline__ float sigmoid_f(float x) { return rcp_approx(1.f + ex2_approx(-1.4426950408889634f * x)); }
__device__ __forceinline__ float tanh_f(float x) {
  return fmaf(-2.f, rcp_approx(ex2_approx(2.8853900817779268f * x) + 1.f), 1.f);
}
__device__ __forceinline__ float gelu_f(float x) {
  const float ax = fabsf(x);
  const float t = rcp_approx(fmaf(0.3275911f * 0.70710678118654752f, ax, 1.f));  // 1 / (1 + p |x|/sqrt2)
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  const float e = ex2_approx(ax * ax * (-0.5f * 1.4426950408889634f));  // exp(-(x/sqrt2)^2)
  const float erf_abs = fmaf(-p * t, e, 1.f);                           // erf(|x|/sqrt2)
  const float hx = 0.5f * x;
  return fmaf(hx, copysignf(erf_abs, x), hx);
}

// Two GELUs at once on the packed fp32 datapath (fma.rn.f32x2 = FFMA2, sm_100+: a scalar FFMA issues every other
// cycle per scheduler, FFMA2 does two lanes' worth per instruction) with ONE MUFU per element instead of two:
//     GELU(x) = max(x, 0) - 0.5 |x| erfc(|x| / sqrt 2),     erfc(a / sqrt 2) = 2^-g(a),
// g = a degree-6 polynomial fitted to -log2 erfc(a / sqrt 2) on [0, 7] weighted by the GELU error it causes
// (tools/fit_gelu.py): |GELU error| <= 3e-7 absolute in fp32 Horner arithmetic, the same class as the
// Abramowitz-Stegun form above (1.5e-7 in erf) and far below the 16-bit output rounding.  |x| is clamped at 7,
// where 0.5 |x| erfc < 1e-11.  Per PAIR: 9 fma-pipe + 5 alu + 2 MUFU instructions (the scalar form: 24 + 3 + 4).
__device__ __forceinline__ float2 gelu2_f(float2 x) {
  const float2 a = make_float2(fminf(fabsf(x.x), 7.f), fminf(fabsf(x.y), 7.f));
  const float2 c6 = make_float2(-3.12827798e-05f, -3.12827798e-05f), c5 = make_float2(7.52653030e-04f, 7.52653030e-04f);
  const float2 c4 = make_float2(-8.02420138e-03f, -8.02420138e-03f), c3 = make_float2(5.33217035e-02f, 5.33217035e-02f);
  const float2 c2 = make_float2(4.58840702e-01f, 4.58840702e-01f), c1 = make_float2(1.15117804e+00f, 1.15117804e+00f);
  const float2 c0 = make_float2(-4.34961722e-06f, -4.34961722e-06f);
  float2 g = __ffma2_rn(c6, a, c5);
  g = __ffma2_rn(g, a, c4);
  g = __ffma2_rn(g, a, c3);
  g = __ffma2_rn(g, a, c2);
  g = __ffma2_rn(g, a, c1);
  g = __ffma2_rn(g, a, c0);
  const float2 e = make_float2(ex2_approx(-g.x), ex2_approx(-g.y));
  const float2 r = make_float2(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f));
  return __ffma2_rn(__fmul2_rn(a, e), make_float2(-0.5f, -0.5f), r);
}

template <typename T> __device__ __forceinline__ uint32_t pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
template <typename T> __device__ __forceinline__ float2 unpack2(uint32_t v);
template <> __device__ __forceinline__ float2 unpack2<__half>(uint32_t v) {
  return __half22float2(*reinterpret_cast<__half2*>(&v));
}
template <> __device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t v) {
  return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&v));
}

// 32 consecutive channels of one pixel, 16-bit storage: 64 B = 4 x 16 B
template <typename T> __device__ __forceinline__ void store32(T* dst, const float (&f)[32]) {
  uint4* d = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    uint4 u;
    u.x = pack2<T>(f[8 * q + 0], f[8 * q + 1]);
    u.y = pack2<T>(f[8 * q + 2], f[8 * q + 3]);
    u.z = pack2<T>(f[8 * q + 4], f[8 * q + 5]);
    u.w = pack2<T>(f[8 * q + 6], f[8 * q + 7]);
    d[q] = u;
  }
}
template <typename T> __device__ __forceinline__ void load32(const T* src, float (&f)[32]) {
  const uint4* s = reinterpret_cast<const uint4*>(src);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint4 u = s[q];
    float2 a = unpack2<T>(u.x), b = unpack2<T>(u.y), c = unpack2<T>(u.z), d = unpack2<T>(u.w);
    f[8 * q + 0] = a.x; f[8 * q + 1] = a.y; f[8 * q + 2] = b.x; f[8 * q + 3] = b.y;
    f[8 * q + 4] = c.x; f[8 * q + 5] = c.y; f[8 * q + 6] = d.x; f[8 * q + 7] = d.y;
  }
}
__device__ __forceinline__ void load32f(const float* src, float (&f)[32]) {
  const float4* s = reinterpret_cast<const float4*>(src);
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float4 u = s[q];
    f[4 * q + 0] = u.x; f[4 * q + 1] = u.y; f[4 * q + 2] = u.z; f[4 * q + 3] = u.w;
  }
}
__device__ __forceinline__ void store32f(float* dst, const float (&f)[32]) {
  float4* d = reinterpret_cast<float4*>(dst);
#pragma unroll
  for (int q = 0; q < 8; ++q) d[q] = make_float4(f[4 * q + 0], f[4 * q + 1], f[4 * q + 2], f[4 * q + 3]);
}

// ------------------------------------------------------------------------------------- host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}

static int sm_count() {
  static int n = [] {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
      v = 148;
    return v;
  }();
  return n;
}

}  // namespace fsb
