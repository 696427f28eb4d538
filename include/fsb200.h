/* fsb200.h -- C ABI of libfsb200.so: B200 (sm_100a) kernels for the FoundationStereo
 * cost-volume + iterative-refinement hot path.
 *
 * The reference (TongZhe2016/FoundationStereo) has no FFI: the seam is Python name
 * binding.  Each entry point below names the reference function it
 * replaces (paths relative to the reference checkout).  foundationstereo_b200/ binds
 * these through ctypes and registers them as torch.library ops; INTEGRATION.md shows
 * the stub a reference maintainer would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to a dense, contiguous tensor unless stated;
 *     the caller owns all buffers (PyTorch allocates them), nothing is allocated here;
 *   - every call is asynchronous on `stream` (a cudaStream_t), never synchronises,
 *     keeps no global state and is safe to capture in a CUDA graph;
 *   - return value: FSB_OK (0) or a negative fsb_status; the message for the calling
 *     thread is available from fsb_last_error();
 *   - there is no CPU path: without a CUDA device every compute entry returns
 *     FSB_ERR_CUDA.
 */
#ifndef FSB200_H_
#define FSB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FSB_VERSION 100 /* major*100 + minor */
#define FSB_MAX_LEVELS 6

typedef void* fsb_stream_t; /* cudaStream_t */

enum fsb_status {
  FSB_OK = 0,
  FSB_ERR_ARG = -1,   /* null pointer / non-positive size / unknown enum */
  FSB_ERR_SHAPE = -2, /* shape outside what the kernels support */
  FSB_ERR_ALIGN = -3, /* pointer not aligned as documented */
  FSB_ERR_CUDA = -4   /* CUDA runtime/driver error, message has the detail */
};

enum fsb_dtype { FSB_F32 = 0, FSB_F16 = 1, FSB_BF16 = 2 };

/* precision of the tensor-core kernels */
enum fsb_precision {
  FSB_PREC_FP32 = 0, /* fp32 FFMA, the parity mode (<= 1e-4 rel of the fp32 reference) */
  FSB_PREC_TF32 = 1, /* tcgen05 kind::tf32, fp32 accumulate in TMEM */
  FSB_PREC_F16 = 2,  /* tcgen05 kind::f16, fp16 operands, fp32 accumulate in TMEM */
  FSB_PREC_BF16 = 3  /* tcgen05 kind::f16, bf16 operands, fp32 accumulate in TMEM */
};

int fsb_version(void);
const char* fsb_last_error(void);
/* number of kernels launched by this thread through the library since load
 * (bench.py's gpu_launches) */
uint64_t fsb_launch_count(void);

/* ---- kernel 1: build_gwc_volume + build_concat_volume ----------------------------
 * replaces core/submodule.py:399-412 (with groupwise_correlation :388-397) and
 * core/submodule.py:416-427.
 *   fl, fr : (B,C,H,W)   left / right 1/4-res features
 *   pl, pr : (B,Cp,H,W)  projected features (proj_cmb output); may be NULL with cat NULL
 *   gwc    : (B,G,D,H,W) out;  gwc[b,g,d,h,w] = <Lhat_g(w), Rhat_g(w-d)> for w>=d else 0
 *   cat    : (B,2Cp,D,H,W) out; [0,Cp) = pl at every d, [Cp,2Cp) = pr shifted by d, 0 for w<d
 * `dtype` is the storage type of all six tensors; the math is fp32 (autocast-off in
 * the reference).  Either output may be NULL to skip it.                              */
int fsb_volume_gwc_concat(const void* fl, const void* fr, const void* pl, const void* pr,
                          void* gwc, void* cat, int B, int C, int Cp, int G, int H, int W,
                          int D, int dtype, fsb_stream_t stream);

/* ---- kernel 2: init correlation + its avg-pool pyramid ---------------------------
 * replaces Combined_Geo_Encoding_Volume.corr (core/geometry.py:68-77) and the
 * init_corr pyramid (core/geometry.py:31,38-40).
 *   f1, f2      : (B,C,H,W1), (B,C,H,W2) fp32
 *   corr_levels : HOST array of L device pointers; level l is (B*H*W1, W2 >> l) fp32,
 *                 level l+1 = mean of adjacent pairs of level l (floor length)
 * precision: FSB_PREC_FP32 (parity) or FSB_PREC_TF32.                                */
int fsb_corr_pyramid(const float* f1, const float* f2, float* const* corr_levels, int B,
                     int C, int H, int W1, int W2, int L, int precision, fsb_stream_t stream);

/* ---- kernel 2b: geometry-volume transpose + pyramid -------------------------------
 * replaces the permute/reshape/contiguous and avg_pool2d pyramid of
 * Combined_Geo_Encoding_Volume.__init__ (core/geometry.py:28-36).
 *   vol        : (B,Cg,D,H,W) fp32
 *   geo_levels : HOST array of L device pointers; level l is (B*H*W, D >> l, Cg) fp32
 * LAYOUT NOTE: the reference keeps (B*H*W, Cg, 1, D_l).  Here the disparity axis is
 * outside the channel axis so that the 2r+2 taps x Cg channels a lookup needs are ONE
 * contiguous run per pixel and level.                                                */
int fsb_geo_pyramid(const float* vol, float* const* geo_levels, int B, int Cg, int D, int H,
                    int W, int L, fsb_stream_t stream);

/* ---- kernel 3: per-iteration lookup ------------------------------------------------
 * replaces Combined_Geo_Encoding_Volume.__call__ (core/geometry.py:43-65) with
 * bilinear_sampler (core/utils/utils.py:44-55) folded in.
 *   geo_levels, corr_levels : as produced by the two calls above (HOST arrays)
 *   disp   : (B,1,H,W) fp32;  coords : (B,H,W,1) fp32 (x index of the pixel)
 *   out    : out_layout 0: (B, L*K*(Cg+1), H, W) fp32   -- the reference's tensor
 *            out_layout 1: (B, H, W, Cpad) fp16, Cpad = L*K*(Cg+1) rounded up to 64,
 *                          pad channels zero -- the A operand of encoder.convc1
 *   K = 2r+1, channel order per level: geo c*K+k (c<Cg), then K correlation taps.    */
int fsb_geo_lookup(float* const* geo_levels, float* const* corr_levels, const float* disp,
                   const float* coords, void* out, int B, int Cg, int D, int H, int W, int W2,
                   int L, int r, int out_layout, fsb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* FSB200_H_ */
