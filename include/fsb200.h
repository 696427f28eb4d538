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

#define FSB_VERSION 130 /* major*100 + minor */
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
  FSB_PREC_BF16 = 3, /* tcgen05 kind::f16, bf16 operands, fp32 accumulate in TMEM */
  FSB_PREC_TF32X3 = 4 /* tcgen05 kind::tf32 on hi/lo split operands (hi*hi + hi*lo + lo*hi): fp32-accurate
                         (~1e-6 relative), the tensor-core parity mode */
};

int fsb_version(void);
const char* fsb_last_error(void);
/* number of kernels launched by this thread through the library since load
 * (bench.py's gpu_launches) */
uint64_t fsb_launch_count(void);
/* Programmatic dependent launch between consecutive kernels of a stream / captured graph (default on, or
 * FSB_PDL=0 in the environment).  Returns the previous setting. */
int fsb_set_pdl(int enable);

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

/* The same launch writing ONE tensor out (B, G+2Cp, D, H, W) = torch.cat([gwc, cat], 1), which is what
 * core/foundation_stereo.py:207-212 assembles with two Python loops and a torch.cat. */
int fsb_volume_combined(const void* fl, const void* fr, const void* pl, const void* pr, void* out, int B, int C,
                        int Cp, int G, int H, int W, int D, int dtype, fsb_stream_t stream);

/* ---- kernel 1 fused with its consumer -- OPT-IN: it computes part of corr_stem ----------
 * replaces core/foundation_stereo.py:207-213 up to and including corr_stem[0], the 1x1x1 Conv3d(G + 2 Cp -> Co) of
 * :164-165:    out = corr_stem[0](cat([build_gwc_volume(fl, fr), build_concat_volume(pl, pr)], 1))
 * in one launch; the (B, G + 2 Cp, D, H, W) volume, its torch.cat copy and the convolution's read never exist.
 *   fl, fr, pl, pr : as fsb_volume_gwc_concat (all of `dtype`)
 *   weight : (Co, G + 2 Cp) fp32 row-major = corr_stem[0].weight (Co, 32, 1, 1, 1): columns [gwc groups | left proj |
 *            right proj], the channel order of the reference's cat;   bias : (Co) fp32
 *   wg_host: HOST pointer to the (Co, G) correlation block weight[:, :G], row-major fp32 (the one host pointer of the
 *            library: these 224 values reach the kernel as launch parameters, so that every FFMA of the 1x1x1 convolution
 *            takes its weight from the constant bank); read during the call, not retained
 *   scratch: 2 * B * C * H * W floats, caller-owned, 16-byte aligned (the L2-normalised features, written by a first
 *            launch and read by the second)
 *   out    : (B, Co, D, H, W) of `dtype`, contiguous.  fp32 arithmetic throughout.
 * Limits: G <= 8, Co <= 32, and one image row's working set must fit shared memory (W = 312, D = 104 does, the
 * W = 512 panorama does not: FSB_ERR_SHAPE). */
int fsb_volume_stem(const void* fl, const void* fr, const void* pl, const void* pr, const float* weight,
                    const float* bias, const float* wg_host, float* scratch, void* out, int B, int C, int Cp, int G, int Co, int H, int W, int D,
                    int dtype, fsb_stream_t stream);

/* ---- kernel 2: init correlation + its avg-pool pyramid ---------------------------
 * replaces Combined_Geo_Encoding_Volume.corr (core/geometry.py:68-77) and the
 * init_corr pyramid (core/geometry.py:31,38-40).
 *   f1, f2      : (B,C,H,W1), (B,C,H,W2) fp32
 *   corr_levels : HOST array of L device pointers; level l is (B*H*W1, W2 >> l) fp32,
 *                 level l+1 = mean of adjacent pairs of level l (floor length)
 * precision: FSB_PREC_FP32 (CUDA-core FFMA), FSB_PREC_TF32X3 (tcgen05, fp32-accurate: both meet the
 * 1e-4 parity bar) or FSB_PREC_TF32 (tcgen05 single pass, ~5e-4, reported separately).   */
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
/* The same with `vol` in its storage type (fsb_dtype): the filtered volume is fp16 when it leaves the reference's
 * autocast region and only becomes fp32 through the caller's `.float()` (core/foundation_stereo.py:229); reading it
 * as it is halves this kernel's input bytes.  The levels are fp32 and bit-identical to converting first. */
int fsb_geo_pyramid_in(const void* vol, int vol_dtype, float* const* geo_levels, int B, int Cg, int D, int H,
                       int W, int L, fsb_stream_t stream);

/* ---- kernel 3: per-iteration lookup ------------------------------------------------
 * replaces Combined_Geo_Encoding_Volume.__call__ (core/geometry.py:43-65) with
 * bilinear_sampler (core/utils/utils.py:44-55) folded in.
 *   geo_levels, corr_levels : as produced by the two calls above (HOST arrays)
 *   disp   : (B,1,H,W) fp32;  coords : (B,H,W,1) fp32 (x index of the pixel)
 *   out    : out_layout 0: (B, L*K*(Cg+1), H, W) fp32   -- the reference's tensor
 *            out_layout 1: (B, H, W, Cpad) fp16, Cpad = L*K*(Cg+1) rounded up to 64,
 *                          pad channels zero -- the A operand of encoder.convc1
 *            out_layout 2: the same in bf16
 *   K = 2r+1, channel order per level: geo c*K+k (c<Cg), then K correlation taps.    */
int fsb_geo_lookup(float* const* geo_levels, float* const* corr_levels, const float* disp,
                   const float* coords, void* out, int B, int Cg, int D, int H, int W, int W2,
                   int L, int r, int out_layout, fsb_stream_t stream);

/* ---- kernels 2 / 3, BANDED form (opt-in; north_star kernel 2 "a banded tcgen05/TMEM GEMM") ---------
 * The lookup of core/geometry.py:57 samples the init correlation of the pixel in image column w1 at
 * x = (w1 - disp) / 2^l + dx, |dx| <= r.  For disparities in [0, d_band] that is a band of the W x W matrix; the band
 * form computes and stores only that band: level l is (B*H*W, bw0 >> l) fp32 and the row of a pixel in column w1
 * holds the level-l columns [s0 >> l, (s0 + bw0) >> l), s0 = floor((w1 - off) / align) * align, zeros where that leaves
 * the image.  fsb_corr_band_geometry gives (bw0, off, align) for (d_band, radius, L): align = 4 << (L-1),
 * off = d_band + ((radius + 2) << (L-1)), bw0 = d_band + ((2 radius + 5) << (L-1)) + align - 1 rounded up to align.
 * PARITY STATEMENT: inside the band (0 <= disp <= d_band) lookups are identical to the full-matrix form; outside it
 * they read ZEROS where the reference reads real correlations (it never clamps the disparity).  The full-matrix
 * entry points above stay the drop-in default.  Needs W1 == W2 = W, W % 4 == 0 and bw0 + align <= 224. */
int fsb_corr_band_geometry(int d_band, int radius, int L, int* bw0, int* off, int* align);
int fsb_corr_pyramid_band(const float* f1, const float* f2, float* const* corr_levels, int B, int C, int H, int W, int L,
                          int precision /* FSB_PREC_TF32X3 | FSB_PREC_TF32 */, int d_band, int radius, fsb_stream_t stream);
/* fsb_geo_lookup over a banded correlation pyramid built with the same (d_band, r, L) */
int fsb_geo_lookup_band(float* const* geo_levels, float* const* corr_levels, const float* disp, const float* coords,
                        void* out, int B, int Cg, int D, int H, int W, int L, int r, int out_layout, int d_band,
                        fsb_stream_t stream);

/* ---- kernels 3+4 fused: lookup -> encoder.convc1 --------------------------------------
 * replaces Combined_Geo_Encoding_Volume.__call__ (core/geometry.py:43-65, bilinear_sampler core/utils/utils.py:44-55)
 * TOGETHER WITH the first layer of BasicMotionEncoder, cor = ReLU(convc1(corr)) (core/update.py:56,63): the
 * L*K*(Cg+1)-channel lookup tensor is produced in shared memory as the GEMM operand and never exists in HBM.
 *   geo_levels / corr_levels / disp / coords : as fsb_geo_lookup
 *   w    : (Cout, L*Kl) 16-bit row-major, Kl = round_up(K*(Cg+1), 64): the columns of convc1.weight
 *          (Cout, L*K*(Cg+1), 1, 1) that belong to level l (the reference's channel order inside a level: c*K + k for
 *          the geometry samples, then the K correlation taps) at [l*Kl, l*Kl + K*(Cg+1)), zeros in between -- the GEMM's
 *          K dimension is split by level so that one level's MMAs run under the other level's gather
 *   bias : (Cout) fp32;   out : (B,H,W,Cout) 16-bit NHWC
 * Same values as fsb_geo_lookup(out_layout 1 / 2) followed by the 1x1 convolution with FSB_EPI_RELU.
 * Limits: radius 4, L <= 2, Cg a multiple of 4 and <= 28, Cout = 256 (the reference's inference configuration); anything else returns
 * FSB_ERR_SHAPE and the two separate calls are the path. */
int fsb_lookup_convc1(float* const* geo_levels, float* const* corr_levels, const float* disp, const float* coords,
                      const void* w, const float* bias, void* out, int B, int Cg, int D, int H, int W, int W2, int L,
                      int r, int Cout, int dtype, fsb_stream_t stream);

/* ---- kernel 4: 3x3 / 1x1 convolution as a tcgen05 implicit GEMM with fused epilogue ----
 * The convolutions of the update block (core/update.py:20-159): SelectiveConvGRU conv0/conv1,
 * the RaftConvGRU gate convolutions convz/convr/convq with their sigmoid / tanh / blend
 * (core/update.py:83-119), BasicMotionEncoder (:51-70), DispHead (:20-32) with the
 * pointwise layers of EdgeNextConvEncoder (core/submodule.py:565-591) and the mask head
 * (core/update.py:134-139).
 *
 * Activations are NHWC with 16-bit elements (fp16 or bf16), channel count a multiple of 64;
 * one GEMM row is one pixel.  The input of a convolution is the channel-concatenation of up
 * to FSB_MAX_SRC tensors ("sources"), which replaces the torch.cat calls of the reference.
 * A call may hold two "segments": segment 0 (ksize 1 or 3) accumulates into TMEM accumulator
 * 0, segment 1 (ksize 1) into accumulator 1 -- the large (3x3) and small (1x1) GRU of
 * SelectiveConvGRU run as ONE launch that shares the activation tiles.
 *
 * weight: packed [rows][Ktot] 16-bit, K-major.  The K axis is, in order: for each segment,
 *   for each tap (ky,kx row-major), for each source of the segment, for each channel of the
 *   source.  rows = n_tiles * block_n (rows >= Cout; rows beyond Cout are zero).
 * bias  : fp32 [nseg][rows].
 *
 * epilogue (acc_s = accumulator s + bias):
 *   FSB_EPI_BIAS/RELU/GELU : out[pix, ch_off + n] = act(acc0)            (16-bit NHWC, ld_out)
 *                            out_f32[b, n, y, x]  = out_scale * act(acc0) (fp32 NCHW, optional)
 *   FSB_EPI_SCALE_RES      : out = res[pix, n] + gamma[n] * acc0          (EdgeNeXt layer scale)
 *   FSB_EPI_GATE_ZR        : rows of a tile are [z (block_n/2) | r (block_n/2)] of hidden channels
 *                            [t*block_n/2, (t+1)*block_n/2); z_l/z_s = sigmoid(acc0/acc1) z part,
 *                            rh_l/rh_s = sigmoid(acc0/acc1 r part) * h32      (core/update.py:91-93)
 *   FSB_EPI_GATE_Q         : q_l = tanh(acc0), q_s = tanh(acc1);
 *                            h' = att*((1-z_s)h + z_s q_s) + (1-att)*((1-z_l)h + z_l q_l)
 *                            -> h32_out (fp32 NHWC) and h16_out (16-bit NHWC) (core/update.py:94-95,119)
 */
#define FSB_MAX_SRC 4

enum fsb_epilogue {
  FSB_EPI_BIAS = 0,
  FSB_EPI_RELU = 1,
  FSB_EPI_GATE_ZR = 2,
  FSB_EPI_GATE_Q = 3,
  FSB_EPI_GELU = 4,
  FSB_EPI_SCALE_RES = 5
};

/* K order of the packed weights = operand-feed strategy of the kernel:
 *   FSB_CONV_LAYOUT_TAP  : segment -> tap (ky,kx row-major) -> source -> channel.  One 128-pixel tile per CTA,
 *                          one 16 KB activation box per tap and 64-channel chunk.
 *   FSB_CONV_LAYOUT_HALO : segment -> source -> 64-channel chunk -> kx -> ky -> channel.  Two stacked 16x8-pixel
 *                          tiles per CTA; per chunk and kx ONE (32+2)-row activation box serves the three ky taps
 *                          of both tiles through shifted UMMA descriptors, and every weight tile feeds 256 pixels.
 *                          Needs 2 * nseg * block_n <= 512 TMEM columns. */
enum fsb_conv_layout { FSB_CONV_LAYOUT_TAP = 0, FSB_CONV_LAYOUT_HALO = 1 };

typedef struct fsb_conv_src {
  const void* ptr; /* (B,H,W,C) 16-bit */
  int C;           /* channels = pixel stride, multiple of 64 */
} fsb_conv_src;

typedef struct fsb_conv_seg {
  int ksize;             /* 1 or 3; stride 1, zero padding ksize/2 */
  int nsrc;              /* sources concatenated along channels, in this order */
  int src[FSB_MAX_SRC];  /* indices into fsb_conv_args.src */
} fsb_conv_seg;

typedef struct fsb_conv_args {
  fsb_conv_src src[FSB_MAX_SRC];
  int nsrc;
  int B, H, W;
  fsb_conv_seg seg[2];
  int nseg;
  const void* weight; /* packed, see above */
  const float* bias;  /* [nseg][rows] */
  int rows;           /* n_tiles * block_n */
  int Cout;           /* valid output channels per accumulator */
  int block_n;        /* 16, 32, 64, 96, 128, 160, 192, 224 or 256; nseg*block_n <= 512 */
  int stages;         /* smem pipeline depth, 0 = choose */
  int layout;         /* fsb_conv_layout: K order of `weight` and operand feed of the kernel */
  int epilogue;       /* fsb_epilogue */
  void* out;          /* 16-bit NHWC or NULL */
  int ld_out;         /* channels per pixel of out */
  int ch_off;         /* first channel written */
  float* out_f32;     /* fp32 NCHW (B,Cout,H,W) or NULL */
  float out_scale;
  const void* res;    /* SCALE_RES: 16-bit NHWC residual */
  int ld_res;
  const float* gamma; /* SCALE_RES: [Cout] */
  const float* h32;   /* gates: fp32 NHWC (B,H,W,hidden) */
  const float* att;   /* GATE_Q: fp32 (B,H,W) */
  void* z_l;          /* gates: 16-bit NHWC (B,H,W,hidden); written by ZR, read by Q */
  void* z_s;
  void* rh_l;         /* ZR out: r*h, the first source of the q convolution */
  void* rh_s;
  float* h32_out;     /* GATE_Q out (may alias h32) */
  void* h16_out;      /* GATE_Q out */
  int hidden;         /* channels of h (128) */
} fsb_conv_args;

int fsb_conv2d_nhwc(const fsb_conv_args* args, int dtype /* FSB_F16 | FSB_BF16 */, fsb_stream_t stream);
/* sizeof(fsb_conv_args) as compiled into the library: FFI bindings compare it with their own
 * mirror of the struct before the first call */
size_t fsb_conv_args_size(void);

/* ---- small NHWC helpers of the update block (HBM-bound, CUDA cores) ------------------- */
/* (B,C,H,W) fp32 -> (B,H,W,Cpad) 16-bit (pad channels zero) and, if dst32 != NULL, (B,H,W,C) fp32 */
int fsb_nchw_to_nhwc(const float* src, void* dst16, float* dst32, int B, int C, int H, int W, int Cpad,
                     int dtype, fsb_stream_t stream);
/* (B,H,W,C) fp32 -> (B,C,H,W) fp32 */
int fsb_nhwc_to_nchw(const float* src, float* dst, int B, int C, int H, int W, fsb_stream_t stream);
/* pool2x (core/update.py:72-73): 3x3 average, stride 2, pad 1, padded zeros counted.
 * src fp32 NHWC (B,H,W,C) -> dst 16-bit NHWC (B,Ho,Wo,C), Ho=(H+1)/2 */
int fsb_pool2x_nhwc(const float* src, void* dst, int B, int C, int H, int W, int dtype, fsb_stream_t stream);
/* interp (core/update.py:78-80): bilinear, align_corners=True, fp32 NHWC (B,Hs,Ws,C) -> 16-bit (B,Hd,Wd,C) */
int fsb_resize_nhwc(const float* src, void* dst, int B, int C, int Hs, int Ws, int Hd, int Wd, int dtype,
                    fsb_stream_t stream);
/* The same two with the source in the 16-bit storage type `dtype` (src, dst both (..,C) 16-bit, C % 8 == 0): under the
 * reference's autocast the hidden state these read IS 16-bit (core/foundation_stereo.py:239-240), and reading it so
 * halves the bytes.  Results are rounded to `dtype` like the fp32-source forms. */
int fsb_pool2x_nhwc16(const void* src, void* dst, int B, int C, int H, int W, int dtype, fsb_stream_t stream);
int fsb_resize_nhwc16(const void* src, void* dst, int B, int C, int Hs, int Ws, int Hd, int Wd, int dtype,
                      fsb_stream_t stream);
/* BasicMotionEncoder.convd1 (core/update.py:58,66): 7x7, 1 -> Cout channels, pad 3, + bias, ReLU.
 * disp fp32 (B,1,H,W); weight fp32 (Cout,49); out 16-bit NHWC (B,H,W,ld_out) channels [0,Cout).
 * If disp_out != NULL also writes disp itself as channel disp_ch of disp_out (ld_disp channels/pixel):
 * the last channel of the motion features (core/update.py:70).                                    */
int fsb_conv7x7_c1(const float* disp, const float* weight, const float* bias, void* out, int ld_out, int Cout,
                   void* disp_out, int ld_disp, int disp_ch, int B, int H, int W, int dtype, fsb_stream_t stream);
/* depthwise 7x7, pad 3, + bias (EdgeNextConvEncoder.dwconv, core/submodule.py:569,581).
 * x, out 16-bit NHWC (B,H,W,C); weight fp32 (C,49) */
int fsb_dwconv7x7_nhwc(const void* x, const float* weight, const float* bias, void* out, int B, int C, int H, int W,
                       int dtype, fsb_stream_t stream);
/* The pointwise MLP of one EdgeNextConvEncoder block (core/submodule.py:584-590, norm=None) in one launch:
 *     out = res + gamma * (GELU(x @ w1^T + b1) @ w2^T + b2)
 * x (the dwconv output), res (the block input), out: 16-bit NHWC flattened to (npix, C); w1 (F,C), w2 (C,F) 16-bit
 * row-major (the nn.Linear weights as stored); b1 (F), b2 (C), gamma (C) fp32.  C = 128, F = 512 (dim 128,
 * expan_ratio 4, core/update.py:27).  The 4C-wide hidden activations stay in shared memory / TMEM.        */
int fsb_edgenext_mlp(const void* x, const void* w1, const float* b1, const void* w2, const float* b2,
                     const float* gamma, const void* res, void* out, long long npix, int C, int F, int dtype,
                     fsb_stream_t stream);
/* DispHead's last layer (core/update.py:28): 3x3 convolution C -> 1, pad 1, + bias, on CUDA cores (1152 MACs per pixel
 * of plain dot product; as a tensor-core tile it used 6 % of the pipe).
 *   x 16-bit NHWC (B,H,W,C), C a multiple of 32; weight fp32 [9][C] (tap (ky,kx) row-major, then channel); bias fp32 [1];
 *   delta fp32 (B,1,H,W) out.  If disp_accum != NULL also disp_accum += delta (the disparity update of
 *   core/foundation_stereo.py:242 folded into the same launch). */
int fsb_conv3x3_cout1(const void* x, const float* weight, const float* bias, float* delta, float* disp_accum, int B,
                      int C, int H, int W, int dtype, fsb_stream_t stream);
/* A WHOLE EdgeNextConvEncoder block (core/submodule.py:565-591, norm=None) in one launch:
 *     out = x + gamma * (GELU(dwconv7x7(x; dw_w, dw_b) @ w1^T + b1) @ w2^T + b2)
 * = fsb_dwconv7x7_nhwc followed by fsb_edgenext_mlp (bit-identical to that pair), with the depthwise output tile produced
 * in shared memory as the A operand of the first GEMM instead of making a round trip through memory.
 * x, out 16-bit NHWC (B,H,W,C), out != x; dw_w fp32 (C,49); w1 (F,C), w2 (C,F) 16-bit row-major; C = 128, F = 512. */
int fsb_edgenext_block(const void* x, const float* dw_w, const float* dw_b, const void* w1, const float* b1,
                       const void* w2, const float* b2, const float* gamma, void* out, int B, int H, int W, int C,
                       int F, int dtype, fsb_stream_t stream);
/* fsb_edgenext_block for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and result as fsb_edgenext_block, computed as fsb_dwconv7x7_nhwc -> fsb_edgenext_mlp with the
 * depthwise output held in a file-scope `static __device__` array of the library (csrc/noalloc.cu) instead of a
 * caller-owned buffer; nothing is allocated or freed by the call.
 *   - limit: C = 128 and B*H*W <= fsb_noalloc_max_pixels() (32768 pixels = 8 MiB of 16-bit scratch, held by every GPU
 *     that loads the library until the process exits); larger calls return FSB_ERR_SHAPE with a message naming the limit;
 *   - all calls share the one array (it keeps the last call's contents; nothing in it has to start at zero): calls on
 *     one stream are ordered, calls that OVERLAP on the GPU (two streams / threads on one device) race on it.     */
int fsb_edgenext_block_noalloc(const void* x, const float* dw_w, const float* dw_b, const void* w1, const float* b1,
                               const void* w2, const float* b2, const float* gamma, void* out, int B, int H, int W,
                               int C, int F, int dtype, fsb_stream_t stream);
size_t fsb_noalloc_max_pixels(void);
/* y += x (fp32, n elements): the disparity update of core/foundation_stereo.py:242 */
int fsb_axpy_f32(float* y, const float* x, size_t n, fsb_stream_t stream);


/* ---- the steps either side of the refinement loop (HBM-bound streams) ------------ */
/* disparity_regression (core/submodule.py:431-435): out[b,0,h,w] = sum_d x[b,d,h,w] * d.
 *   x (B,D,H,W) of `dtype`; logits = 0: x are probabilities, out has the dtype of x (the reference's result dtype);
 *   logits = 1: x are the classifier's logits and the softmax over D of core/foundation_stereo.py:218 is fused in
 *   (one pass, running max / sum), out is fp32 (softmax runs in fp32 under autocast).                           */
int fsb_disparity_regression(const void* x, void* out, int B, int D, int H, int W, int logits, int dtype,
                             fsb_stream_t stream);
/* context_upsample (core/submodule.py:456-468): convex 4x upsampling,
 *   out[b,Y,X] = sum_{k=ky*3+kx} scale * disp_low[b, Y/4+ky-1, X/4+kx-1] * weights[b,k,Y,X]   (zero outside the image).
 *   disp_low (B,1,h,w) fp32; weights (B,9,4h,4w) of `dtype`; out (B,4h,4w) fp32.
 *   logits = 1: weights are the logits of spx_gru and the softmax over the 9 of core/foundation_stereo.py:188 is
 *   fused in; scale = 4 folds the `disp*4.` of :189 (pass 1 for the reference's call).  weights/out 16-byte aligned. */
int fsb_context_upsample(const float* disp_low, const void* weights, float* out, int B, int h, int w, int logits,
                         float scale, int dtype, fsb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* FSB200_H_ */
