// fsb_edgenext_block_noalloc: one EdgeNextConvEncoder block (depthwise 7x7 -> pointwise MLP) for callers that cannot
// allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in
// the reference.
//
// The two-launch form of the block (fsb_dwconv7x7_nhwc -> fsb_edgenext_mlp, csrc/nhwc.cu + csrc/mlp.cu) needs a scratch
// tensor for the depthwise output, which every other caller of the library allocates itself.  Here that scratch is the
// file-scope __device__ array below: every GPU that loads libfsb200.so holds its 8 MiB until the process exits, and the
// entry point calls no cudaMalloc / cudaFree / allocating routine.
//
//   * g_dw_scratch is a HOST-SIDE STAND-IN for the array in host code, not a device address.  The device address is
//     taken once per device, with that device current, by cudaGetSymbolAddress (scratch_for_current_device) and only
//     that pointer is handed to the kernels.
//   * Every call shares the one array and it keeps the last call's contents.  Nothing has to start at zero: the
//     depthwise launch overwrites all npix * C elements the MLP launch then reads (rows past npix of the last
//     128-pixel tile are zero-filled by the TMA unit from the tensor map's bounds, not read from the array).
//   * Calls that overlap on the GPU -- two streams, or two threads of one process on the same device -- RACE on the
//     array.  Calls on one stream are ordered and safe.
#include "common.cuh"

#include <mutex>

namespace fsb {

constexpr long long kNoAllocMaxPix = 32768;  // pixels (B*H*W) per call: one 96x312 KITTI or 120x160 pair at 1/4 res
constexpr int kNoAllocC = 128;               // channels of the block (DispHead, core/update.py:27)
constexpr size_t kNoAllocBytes = (size_t)kNoAllocMaxPix * kNoAllocC * 2;  // 16-bit elements: 8 MiB

static __device__ __align__(256) unsigned char g_dw_scratch[kNoAllocBytes];

constexpr int kMaxDevices = 64;
static void* g_scratch_ptr[kMaxDevices];  // device address of g_dw_scratch per device ordinal (nullptr: not yet taken)
static std::mutex g_scratch_mutex;

// Device address of the scratch array on the CURRENT device (taken on first use per device; no allocation: the array
// is part of the module image the runtime loaded onto that device).
static int scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "fsb_edgenext_block_noalloc: cudaGetDevice");
  FSB_REQUIRE(dev >= 0 && dev < kMaxDevices, FSB_ERR_CUDA, "fsb_edgenext_block_noalloc: device ordinal %d (max %d)", dev,
              kMaxDevices - 1);
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (g_scratch_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_dw_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "fsb_edgenext_block_noalloc: cudaGetSymbolAddress(g_dw_scratch)");
    FSB_REQUIRE(ptr != nullptr && (reinterpret_cast<uintptr_t>(ptr) & 255u) == 0, FSB_ERR_CUDA,
                "fsb_edgenext_block_noalloc: scratch address %p is not 256-byte aligned", ptr);
    g_scratch_ptr[dev] = ptr;
  }
  *out = g_scratch_ptr[dev];
  return FSB_OK;
}

}  // namespace fsb

extern "C" size_t fsb_noalloc_max_pixels(void) { return (size_t)fsb::kNoAllocMaxPix; }

extern "C" int fsb_edgenext_block_noalloc(const void* x, const float* dw_w, const float* dw_b, const void* w1,
                                          const float* b1, const void* w2, const float* b2, const float* gamma, void* out,
                                          int B, int H, int W, int C, int F, int dtype, fsb_stream_t stream) {
  fsb::NvtxRange nvtx_range__("fsb_edgenext_block_noalloc");
  using namespace fsb;
  FSB_REQUIRE(x && dw_w && dw_b && w1 && b1 && w2 && b2 && gamma && out, FSB_ERR_ARG, "fsb_edgenext_block_noalloc: NULL pointer");
  FSB_REQUIRE(x != out, FSB_ERR_ARG, "fsb_edgenext_block_noalloc: out must not alias x (x is the residual)");
  FSB_REQUIRE(dtype == FSB_F16 || dtype == FSB_BF16, FSB_ERR_ARG, "fsb_edgenext_block_noalloc: dtype %d (need f16/bf16)", dtype);
  FSB_REQUIRE(B > 0 && H > 0 && W > 0, FSB_ERR_ARG, "fsb_edgenext_block_noalloc: B=%d H=%d W=%d", B, H, W);
  FSB_REQUIRE(C == kNoAllocC, FSB_ERR_SHAPE, "fsb_edgenext_block_noalloc: C=%d (the static scratch is laid out for %d channels)",
              C, kNoAllocC);
  const long long npix = (long long)B * H * W;
  FSB_REQUIRE(npix <= kNoAllocMaxPix, FSB_ERR_SHAPE,
              "fsb_edgenext_block_noalloc: B*H*W = %lld pixels exceed the static scratch of %lld pixels (%zu bytes); "
              "use fsb_edgenext_block or fsb_dwconv7x7_nhwc + fsb_edgenext_mlp with a caller-owned buffer",
              npix, kNoAllocMaxPix, kNoAllocBytes);
  void* scratch = nullptr;
  int rc = scratch_for_current_device(&scratch);
  if (rc != FSB_OK) return rc;
  // depthwise 7x7 + bias: x -> scratch (every element of scratch[0, npix*C) is written)
  rc = fsb_dwconv7x7_nhwc(x, dw_w, dw_b, scratch, B, C, H, W, dtype, stream);
  if (rc != FSB_OK) return rc;
  // out = x + gamma * (GELU(scratch @ w1^T + b1) @ w2^T + b2)
  return fsb_edgenext_mlp(scratch, w1, b1, w2, b2, gamma, x, out, npix, C, F, dtype, stream);
}
