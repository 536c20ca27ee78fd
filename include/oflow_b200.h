/*
 * oflow_b200.h - C ABI of liboflow_b200.so: the B200 (sm_100a) hot path of the
 * unsupervised optical-flow training step of xianlopez/optical_flow.
 *
 * Boundary rules:
 *   - every pointer is a DEVICE pointer unless the name says host; the caller
 *     owns all memory (inputs, outputs, gradients, workspaces);
 *   - tensors are contiguous NHWC; "dtype" selects the storage type of
 *     activations (OF_F32 = reference layout, OF_BF16 = conv fast path);
 *     gradients produced by the HBM-bound ops are always float32;
 *   - functions are asynchronous on `stream` (a cudaStream_t passed as void*),
 *     re-entrant, and allocate nothing;
 *   - return value: 0 = ok, < 0 = argument/shape error (the reference's
 *     trace-time asserts), > 0 = cudaError_t.  of_last_error() describes it.
 *
 * Each entry point cites the reference lines it replaces (paths relative to
 * the reference tree).
 */
#ifndef OFLOW_B200_H
#define OFLOW_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { OF_F32 = 0, OF_BF16 = 1 };
/* OF_WARP_REF_IJ reproduces model.py:65-71 literally (sampling point =
 * [row, col] + flow, read as (x, y) by the sampler: SURVEY F4);
 * OF_WARP_XY is the conventional order (x = col + flow[0], y = row + flow[1]). */
enum { OF_WARP_REF_IJ = 0, OF_WARP_XY = 1 };
/* conv epilogue activations */
enum { OF_ACT_NONE = 0, OF_ACT_RELU = 1, OF_ACT_LEAKY = 2 };

#define OF_OK 0
#define OF_ERR_ARG (-1)
#define OF_ERR_SHAPE (-2)
#define OF_ERR_UNSUPPORTED (-3)
#define OF_ERR_STATE (-4)

int of_version(void);
const char* of_last_error(void);
/* number of CUDA kernels this library has launched in this process (bench evidence) */
unsigned long long of_launch_count(void);
/* Select a kernel variant by name (e.g. "OF_CV_NO_TC" = "1": SIMT cost volume instead of the tcgen05 one;
 * "OF_LOSS_GENERIC": direct-gather loss kernel).  value NULL unsets.  Process-wide, not thread-safe against
 * concurrent launches; meant for tests and tuning scripts.  The shipped library reads NO environment variables. */
int of_set_option(const char* name, const char* value);
/* Per-launch timeline, for bench.py's in-step kernel shares (no ncu needed): after of_prof_enable(1, stream) every
 * launch this library makes records a CUDA event on `stream` (all launches must go to that one stream; not during
 * graph capture); of_prof_collect synchronises and returns up to max_recs records (launch order): kernel name, the
 * device time since the previous launch's end, and the work (FLOPs) the host code noted for it (0 = not noted).
 * of_prof_enable(0, NULL) switches it off. */
typedef struct { char name[40]; double work; float ms; } of_prof_rec;
int of_prof_enable(int on, void* stream);
int of_prof_collect(of_prof_rec* out, int max_recs);
/* Adds a boundary named `name` to the timeline without launching anything: the interval since the previous record
 * (e.g. a delay kernel the caller queued so that the host runs ahead of the device) is booked under that name. */
int of_prof_marker(const char* name);

/* ---- sampler ------------------------------------------------------------
 * transformations.py:85-129 bilinear_interpolation (+ :70-81 gather).
 * in (B,H,W,C) dtype; pts (B,H,W,2) f32 absolute (x,y); out (B,H,W,C) dtype.
 * idx (nullable): int32 (B,H,W,4) = clipped x0,x1,y0,y1 (bit-exact check). */
int of_bilinear_sample_fwd(const void* in, const float* pts, void* out, int32_t* idx,
                           int B, int H, int W, int C, int dtype, void* stream);
/* TF autodiff of the above.  d_out f32 (B,H,W,C).  d_in f32 (B,H,W,C),
 * nullable, zeroed by the callee then scatter-added.  d_pts f32 nullable. */
int of_bilinear_sample_bwd(const void* in, const float* pts, const float* d_out,
                           float* d_in, float* d_pts,
                           int B, int H, int W, int C, int dtype, void* stream);

/* model.py:55-73 warp_features (grid built in-kernel). flow (B,H,W,2) f32. */
int of_warp_features_fwd(const void* feats, const float* flow, void* out, int32_t* idx,
                         int B, int H, int W, int C, int mode, int dtype, void* stream);
int of_warp_features_bwd(const void* feats, const float* flow, const float* d_out,
                         float* d_feats, float* d_flow,
                         int B, int H, int W, int C, int mode, int dtype, void* stream);

/* same, with the incoming gradient d_out in dtype d_out_dtype (bf16 needs bf16 feats) */
int of_warp_features_bwd_g(const void* feats, const float* flow, const void* d_out, int d_out_dtype,
                           float* d_feats, float* d_flow,
                           int B, int H, int W, int C, int mode, int dtype, void* stream);

/* old/image_warp.py:7-78 image_warp (== image_warp_tf :81-153); f32 only. */
int of_image_warp_fwd(const float* im, const float* flow, float* out, int32_t* idx,
                      int B, int H, int W, int C, void* stream);
int of_image_warp_bwd(const float* im, const float* flow, const float* d_out,
                      float* d_im, float* d_flow,
                      int B, int H, int W, int C, void* stream);

/* ---- resize -------------------------------------------------------------
 * tf.image.resize bilinear, half-pixel centres, no antialias
 * (loss.py:18, old/loss.py:15-22, model.py:77).  f32. */
int of_resize_bilinear(const float* in, float* out, int B, int H, int W, int C,
                       int out_h, int out_w, void* stream);
/* model.py:76-77 upscale_flow: x2 bilinear, values * 2.0.  flow (B,h,w,2). */
int of_upscale_flow_fwd(const float* flow, float* out, int B, int h, int w, void* stream);
/* adjoint: d_out (B,2h,2w,2) -> d_flow (B,h,w,2) (overwritten, deterministic). */
int of_upscale_flow_bwd(const float* d_out, float* d_flow, int B, int h, int w, void* stream);

/* ---- cost volume ----------------------------------------------------------
 * model.py:29-42 create_cost_volume.  f1,f2 (B,H,W,C) dtype; cv has D=(2m+1)^2
 * channels written at channel offset 0 of a tensor whose pixel stride is
 * cv_stride elements (cv_stride >= D; lets the caller place it inside a padded
 * concat buffer); cv dtype = dtype. */
int of_cost_volume_fwd(const void* f1, const void* f2, void* cv,
                       int B, int H, int W, int C, int max_disp, int cv_stride,
                       int dtype, void* stream);
/* d_cv (B,H,W,cv_stride) dtype -> d_f1, d_f2 f32 (B,H,W,C), overwritten. */
int of_cost_volume_bwd(const void* f1, const void* f2, const void* d_cv,
                       float* d_f1, float* d_f2,
                       int B, int H, int W, int C, int max_disp, int cv_stride,
                       int dtype, void* stream);

/* Same gradient with bf16 tensors end to end (the tensor-core conv path's activation dtype):
 * d_f1 = dcv/df1 (+ d_f1_add, nullable bf16 (B,H,W,C): the gradient reaching f1 through the decoder's
 * first conv, model.py:100-104), d_f2 = dcv/df2; both bf16 (B,H,W,C), overwritten.  max_disp == 3 only. */
int of_cost_volume_bwd_bf16(const void* f1, const void* f2, const void* d_cv, void* d_f1, void* d_f2,
                            const void* d_f1_add, int B, int H, int W, int C, int max_disp, int cv_stride,
                            void* stream);

/* ---- fused pyramid losses -------------------------------------------------
 * loss.py:5-32 LossLayer: for each scale s, resize the full-res 6-channel
 * batch to flow s's size, warp frame 2 by the flow (REF_IJ), mean |diff|,
 * average over scales.  ONE launch computes the scalar loss and (if d_flows
 * != NULL) d loss / d flow_s for every scale.
 *   imgs (B,H,W,6) f32; flows/d_flows: HOST arrays of num_scales device
 *   pointers, flow s is (B,hs[s],ws[s],2) f32; loss: device float[1];
 *   workspace: of_loss_workspace_bytes(B, H, W, num_scales) device bytes, 16-byte aligned,
 *   zeroed ONCE by the caller after allocation and then reusable by every call of that size
 *   or smaller (header + an L2-resident float4 pyramid of the resized pair). */
size_t of_loss_workspace_bytes(int B, int H, int W, int num_scales);
int of_loss_l1_pyramid_fwdbwd(const float* imgs, const float* const* flows, float* const* d_flows,
                              const int* hs, const int* ws, int num_scales,
                              float* loss, int B, int H, int W, int mode,
                              void* workspace, void* stream);
/* old/loss.py:7-105 MyLoss photometric (Charbonnier, image_warp of frame 1)
 * + 0.2 * first-order Charbonnier flow smoothness, / batch.  Same calling
 * convention; flows ordered coarse -> fine like y_pred[0] (old/loss.py:24-25). */
int of_loss_charbonnier_smooth_fwdbwd(const float* imgs, const float* const* flows, float* const* d_flows,
                                      const int* hs, const int* ws, int num_scales,
                                      float* loss, int B, int H, int W,
                                      void* workspace, void* stream);

/* ---- pooling / elementwise ------------------------------------------------
 * model.py:17 MaxPool2D(): 2x2, stride 2, valid.  in (B,H,W,C) -> (B,H/2,W/2,C) */
int of_maxpool2_fwd(const void* in, void* out, int B, int H, int W, int C, int dtype, void* stream);
/* gradient goes to the first maximum in window scan order; d_in overwritten. */
int of_maxpool2_bwd(const void* in, const void* d_out, void* d_in,
                    int B, int H, int W, int C, int dtype, void* stream);
/* one pass for the stem's backward entry (model.py:12-17): dz = (maxpool2_bwd(in, d_out) + add) * act'(in);
 * in = the pool input (stem output, post-activation), add = a gradient that reaches it directly. */
int of_maxpool2_bwd_add_act(const void* in, const void* d_out, const void* add, void* dz,
                            int B, int H, int W, int C, int act, int dtype, void* stream);
/* resnet/models.py:31-34: out = relu(a + b) */
int of_add_relu_fwd(const void* a, const void* b, void* out, size_t n, int dtype, void* stream);
/* d = d_out * (out > 0) */
int of_relu_bwd(const void* out, const void* d_out, void* d, size_t n, int dtype, void* stream);

/* ---- convolution ------------------------------------------------------------
 * Keras Conv2D(padding='same') on NHWC / HWIO (model.py:12-15,104-114;
 * resnet/models.py:17-30).  TF asymmetric SAME padding is computed inside.
 *
 * fp32 SIMT path (exact-parity mode; also used for Cin=3 / Cout=2 layers):
 *   y = act(conv(x, w) + bias);  act per OF_ACT_*, LeakyReLU alpha = 0.3.
 *   x (N,H,W,Cin), w (KH,KW,Cin,Cout) f32, bias (Cout) f32 nullable.
 *   in_dtype/out_dtype select the storage type of x / y. */
int of_conv2d_fprop(const void* x, const float* w, const float* bias, void* y,
                    int N, int H, int W, int Cin, int KH, int KW, int Cout, int stride,
                    int act, int in_dtype, int out_dtype, void* stream);
/* dx = conv_transpose(dy, w); dy (N,Ho,Wo,Cout), dx (N,H,W,Cin) overwritten. */
int of_conv2d_dgrad(const void* dy, const float* w, void* dx,
                    int N, int H, int W, int Cin, int KH, int KW, int Cout, int stride,
                    int dy_dtype, int dx_dtype, void* stream);
/* dw (KH,KW,Cin,Cout) f32 and db (Cout) f32 (nullable), both overwritten. */
int of_conv2d_wgrad(const void* x, const void* dy, float* dw, float* db,
                    int N, int H, int W, int Cin, int KH, int KW, int Cout, int stride,
                    int x_dtype, int dy_dtype, void* stream);
/* bf16 tensor-core path (tcgen05 / TMEM implicit GEMM, fp32 accumulate) - the fast path for
 * every conv with >= 32 in/out channels.  One descriptor drives fprop and dgrad:
 *   out[p, n] = epi( sum_{tap, k} src[p*stride + sign*(tap - pad), k] * wpacked[tap][n][k] )
 *   epi(v) = act(v + bias[n] + add[p, n]) * act'(mask[p, n])
 * fprop: sign=+1, wpacked = [tap][Cout][K over the (64-padded) sources];
 * dgrad (stride 1): sign=-1, src = dz, wpacked = [tap][Cin segment][Cout], mask = the producing
 * layer's stored activation output so the next dz is written directly.
 * Sources are bf16 NHWC (channel counts and pixel strides multiples of 8); up to two sources
 * replace tf.concat (model.py:100-102). */
typedef struct {
  const void* src[2]; int src_c[2]; int src_stride[2]; int num_src;
  const void* wpacked; int n_rows; int k_pad;
  int N, H, W;              /* source spatial dims */
  int Ho, Wo;               /* output spatial dims */
  int KH, KW, stride, pad_t, pad_l, sign;
  /* optional strided VIEW of source 0 (0 = dense NHWC): element strides between consecutive W / H / N
   * coordinates, and separate W / H convolution strides.  Lets the Cin = 3 stem run as a 7x1 conv over
   * overlapping 8-pixel windows of the packed image (of_pack_image_pair8): W pitch 2 pixels = 16 elements. */
  long long view_w_stride, view_h_stride, view_n_stride; int stride_w, stride_h;
  int n_mma, n_valid;       /* MMA N (multiple of 16, <= 256) and real output channels */
  void* out; int out_dtype; int out_stride;
  const float* bias; int act;
  const void* add; int add_dtype; int add_stride;
  const void* mask; int mask_dtype; int mask_stride; int mask_act;
  /* Sign-bit activation masks (3x3 / stride-1 layers with Cout %% 16 == 0, Cout <= 128): sign_out (nullable)
   * receives 16 bytes per output pixel, bit c = (activated output channel c > 0); a later dgrad passes that
   * buffer as `mask` with mask_dtype = OF_MASK_BITS instead of re-reading the bf16 activation. */
  void* sign_out;
  /* Split output (3x3 / stride-1, bf16): output columns >= out_split are written to out2 (pixel stride out2_stride,
   * column - out_split) - the dgrad of a conv whose input is a concat writes each segment's gradient tensor from
   * ONE launch (model.py:100-104). */
  void* out2; int out2_stride; int out_split;
  /* add_mode 0 (default): y = act(conv + bias + add).  add_mode 1 (3x3 / stride-1 layers, bf16 out, Cout %% 16 == 0):
   * y = relu(act(conv + bias) + add) - resnet/models.py:24-35 applies ReLU to conv2 + BN BEFORE the residual add and
   * again after it; sign_out (if given) records the INNER activation's sign, which is what conv2's backward needs. */
  int add_mode;
} of_tc_conv;
#define OF_MASK_BITS 2
int of_conv_tc_run(const of_tc_conv* c, void* stream);
/* fp32 HWIO weights -> bf16 packed [taps][rows][cols] (cols %% 64 == 0).  transpose=0: rows = co,
 * cols = packed ci; transpose=1: rows = packed ci, cols = co.  Segment s maps original ci in
 * [seg_off, seg_off+seg_cnt) to packed index seg_base + (ci - seg_off); everything else is 0. */
int of_conv_tc_pack_weights(const float* w_hwio, void* packed, int taps, int Cin, int Cout, int rows, int cols,
                            int transpose, int nseg, const int* seg_off, const int* seg_cnt, const int* seg_base,
                            void* stream);
/* The same packing for many tensors in one launch.  jobs_device: DEVICE array of njobs entries sorted by
 * block_begin; job j owns blocks [block_begin_j, block_begin_{j+1}) of 2048 output elements each
 * (ceil(taps*rows*cols / 2048) blocks); total_blocks = their sum. */
typedef struct {
  const float* w; void* out;
  int taps, Cin, Cout, rows, cols, transpose, nseg;
  int seg_off[3], seg_cnt[3], seg_base[3];
  int block_begin, reserved;
} of_pack_job;
int of_conv_tc_pack_weights_batch(const of_pack_job* jobs_device, int njobs, int total_blocks, void* stream);
/* dw[tap][ci][co] (f32 HWIO, overwritten) = sum_pix x[pix (+) tap][ci] * dy[pix][co]; x = up to two
 * bf16 NHWC sources (x_valid real channels each, the first mapping to original ci = x_ci0). */
typedef struct {
  const void* x[2]; int x_c[2]; int x_stride[2]; int x_valid[2]; int x_ci0[2]; int num_src;
  const void* dy; int dy_c; int dy_stride;
  float* dw;
  int N, H, W, Ho, Wo, KH, KW, stride, pad_t, pad_l, Cin, Cout;
  long long view_w_stride, view_h_stride, view_n_stride; int stride_w, stride_h;   /* as in of_tc_conv */
  float* db;   /* nullable: bias gradient db[co] = sum_pix dy[pix][co] (f32, Cout), fused into the same launch */
} of_tc_wgrad;
int of_conv_tc_wgrad(const of_tc_wgrad* c, void* stream);
/* Stem (model.py:12, Cin = 3, 7x7 / 2) on tensor cores.  of_pack_image_pair8: imgs (B,H,W,6) f32 -> bf16
 * (2B, (H+6)/2, W+6, 2, 4): frame 1s then frame 2s, TF-SAME zero border (2 rows / cols before, 4 after); one 16-byte
 * texel = the pixels of two consecutive rows (2 pr - 2, 2 pr - 1) x 4 channels (3 used).  A window of 8 texels is one
 * 64-element K chunk ((kw, p, c) -> kw*8 + p*4 + c, kh = 2 tap + p) and the 7x7/2 conv is a 4-tap stride-1 conv
 * over the view {W pitch 16 elements, H pitch (W+6)*8}.  of_conv1_pack_weights: folded f32 (7,7,3,Cout) -> bf16
 * [4][Cout][64]; of_conv1_unpack_wgrad: dw_view f32 [4][64][Cout] -> (7,7,3,Cout). */
int of_pack_image_pair8(const float* imgs, void* out, int B, int H, int W, void* stream);
int of_conv1_pack_weights(const float* w_hwio, void* packed, int Cout, void* stream);
int of_conv1_unpack_wgrad(const float* dw_view, float* dw_hwio, int Cout, void* stream);
/* Backward of the flow head, model.py:114 Conv2D(2, 3, padding='same'), in one pass: x (N,H,W,Cin<=32) dtype =
 * the head's input (= the previous conv's post-activation output), dy (N,H,W,2) f32, w (3,3,Cin,2) f32.
 * dw (3,3,Cin,2) f32 and db (2) f32 (nullable) are overwritten; dz (nullable, dtype, (N,H,W,Cin)) =
 * dgrad * act'(x), i.e. already the previous layer's pre-activation gradient (act = its activation). */
int of_flow_head_bwd(const void* x, const float* dy, const float* w, float* dw, float* db, void* dz,
                     int N, int H, int W, int Cin, int act, int dtype, void* stream);
/* out (N,H,W,C) = dy (N,Ho,Wo,C) with zeros inserted (out[2i,2j] = dy[i,j]): a stride-2 conv's
 * dgrad is the stride-1 dgrad of this tensor. */
int of_zero_upsample2(const void* dy, void* out, int N, int Ho, int Wo, int H, int W, int C, int dtype,
                      void* stream);
/* Same tensor, but only the non-zero quarter is written (out[2i,2j] = dy[i,j]): `out` is a buffer the caller
 * zero-filled once and keeps for this layer, so the zeros of earlier steps are still in place. */
int of_scatter_upsample2(const void* dy, void* out, int N, int Ho, int Wo, int H, int W, int C, int dtype,
                         void* stream);
/* db[c] = sum over pixels of dy[p, c] (dy has pixel stride dy_stride) */
int of_colsum(const void* dy, float* db, size_t npix, int C, int dy_stride, int dtype, void* stream);

/* conv_mode "bf16x3" (tensor-core parity mode; stands where SURVEY 7 step 7 planned a tf32 mode for the fp32 convolutions of
 * model.py:12-15,104-114): v = hi + lo with hi = bf16(v), lo = bf16(v - hi); pixels of C channels written with pixel stride
 * C_pad (padding channels zero); out_dtype OF_BF16 (operands) or OF_F32 (weights, bf16-representable values). */
int of_split_bf16(const float* src, void* hi, void* lo, size_t npix, int C, int C_pad, int out_dtype, void* stream);

/* dz = dy * act'(y)  (y = stored activation output) */
int of_act_bwd(const void* y, const void* dy, void* dz, size_t n, int act, int dtype, void* stream);
/* the same from a sign-bit mask (of_tc_conv.sign_out: 16 bytes per pixel, bit c = activated channel c > 0) instead
 * of the activation tensor: dz[p][c] = dy[p][c] * (bit ? 1 : slope(act)); C %% 8 == 0, C <= 128; dy / dz dtype. */
int of_act_bwd_bits(const void* bits, const void* dy, void* dz, size_t npix, int C, int act, int dtype, void* stream);

/* ---- glue -------------------------------------------------------------------
 * model.py:122-123,131-132: split (B,H,W,6) into the two frames stacked on the
 * batch axis, (2B,H,W,3), so the shared-weight encoder runs once. */
int of_split_image_pair(const float* imgs, void* out, int B, int H, int W, int out_dtype, void* stream);
/* strided channel-slice copy (tf.concat of model.py:100-102 and its gradient
 * split): dst[p*dst_stride + c] (=|+=) src[p*src_stride + c], c < nch. */
int of_copy_channels(const void* src, int src_stride, void* dst, int dst_stride, int nch, size_t npix,
                     int src_dtype, int dst_dtype, int accumulate, void* stream);
/* Inference-mode BatchNormalization (model.py:14, resnet/models.py:19,24)
 * folded into the preceding conv: w' = w*s, b' = (b-mean)*s + beta,
 * s = gamma/sqrt(var+eps); w is (K, Cout) = HWIO flattened; b nullable. */
int of_bn_fold_fwd(const float* w, const float* b, const float* gamma, const float* beta, const float* mean,
                   const float* var, float eps, float* w_folded, float* b_folded, int K, int Cout, void* stream);
/* exact chain rule from (dw', db') back to (dw, db, dgamma, dbeta); db nullable. */
int of_bn_fold_bwd(const float* w, const float* b, const float* gamma, const float* mean, const float* var,
                   float eps, const float* dw_folded, const float* db_folded, float* dw, float* db, float* dgamma,
                   float* dbeta, int K, int Cout, void* stream);

/* All BN folds of a model in ONE launch each way.  jobs_device: device array of of_bn_job (one per conv + BN pair;
 * pointers as in of_bn_fold_fwd / of_bn_fold_bwd, b / db nullable).  Forward: a job owns the blocks
 * [block_begin, next job's block_begin), OF_BN_FWD_BLOCK elements of K*Cout each; backward: one block per output
 * channel, block_begin = the running sum of Cout. */
#define OF_BN_FWD_BLOCK 2048
typedef struct {
  const float* w; const float* b; const float* gamma; const float* beta; const float* mean; const float* var;
  float* wf; float* bf;                                    /* forward outputs (folded weight / bias) */
  const float* dwf; const float* dbf;                      /* backward inputs (gradients of the folded pair) */
  float* dw; float* db; float* dgamma; float* dbeta;       /* backward outputs */
  int K, Cout, fwd_block_begin, bwd_block_begin;
} of_bn_job;
int of_bn_fold_fwd_batch(const of_bn_job* jobs_device, int njobs, int total_blocks, float eps, void* stream);
int of_bn_fold_bwd_batch(const of_bn_job* jobs_device, int njobs, int total_blocks, float eps, void* stream);

/* ---- data path ------------------------------------------
 * data_reader.py:45-64 read_item + :36-41 batch assembly on the device: img1 / img2 = decoded uint8 BGR frames
 * (src_h, src_w, 3) in device memory; out = the item's (H, W, 6) f32 slice of the batch:
 * cv2.resize(INTER_LINEAR, bit-exact fixed point) -> / 255 -> minus [123,117,104]/255; flip != 0 swaps the pair
 * (the np.random.rand() < 0.5 branch of :46-51 is drawn by the caller). */
int of_preprocess_pair(const uint8_t* img1, const uint8_t* img2, int src_h, int src_w, float* out,
                       int H, int W, int flip, void* stream);
/* The whole batch in one launch: frames = (B, 2, src_h, src_w, 3) uint8 in device memory (one H2D copy of the
 * decoded pairs), out = (B, H, W, 6) f32; flips = B HOST bytes (nullable = no swaps). */
int of_preprocess_batch(const uint8_t* frames, int B, int src_h, int src_w, float* out, int H, int W,
                        const uint8_t* flips, void* stream);

/* ---- depth / pose re-projection ---------------------------
 * transformations.py:12-67,133-191.  points are (B,H,W,4) f32 homogeneous, depth (B,H,W,1) f32, pixel
 * coordinates (B,H,W,2) f32 (x, y) - what of_bilinear_sample_fwd takes; T (B,4,4) f32 row-major on the
 * device; intrinsics K / Kinv are 9 floats row-major in HOST memory (constructor constants of the layers).
 * Nullable gradient outputs are skipped; d_T is zeroed by the callee, then reduced over the pixels. */
/* BackprojectLayer.__call__ :24-30: points = [depth * (Kinv @ [x, y, 1]), 1] */
int of_backproject_fwd(const float* depth, const float* kinv_host, float* points, int B, int H, int W, void* stream);
int of_backproject_bwd(const float* d_points, const float* kinv_host, float* d_depth, int B, int H, int W,
                       void* stream);
/* ProjectLayer.__call__ :39-44: hom = [K | 0] @ p; pix = hom[:2] / (hom[2] + 1e-7) */
int of_project_fwd(const float* k_host, const float* points, float* pix, int B, int H, int W, void* stream);
int of_project_bwd(const float* k_host, const float* points, const float* d_pix, float* d_points, int B, int H,
                   int W, void* stream);
/* transform3d :61-67: out[b,i,j] = T[b] @ points[b,i,j] */
int of_transform3d_fwd(const float* T, const float* points, float* out, int B, int H, int W, void* stream);
int of_transform3d_bwd(const float* T, const float* points, const float* d_out, float* d_points, float* d_T, int B,
                       int H, int W, void* stream);
/* WarpLayer.__call__ :52-58 coordinate chain in one pass: pix = project(transform3d(T, points)); the image
 * lookup that follows is of_bilinear_sample_fwd/bwd on pix */
int of_view_coords_fwd(const float* k_host, const float* T, const float* points, float* pix, int B, int H, int W,
                       void* stream);
int of_view_coords_bwd(const float* k_host, const float* T, const float* points, const float* d_pix,
                       float* d_points, float* d_T, int B, int H, int W, void* stream);
/* make_transformation_matrix :176-191 (+ rotation_from_axisangle :133-173): raw (B,6) = axis-angle, translation
 * -> (B,4,4); invert != 0 gives the inverse rigid transform */
int of_make_transformation_matrix_fwd(const float* raw, int invert, float* out, int B, void* stream);
int of_make_transformation_matrix_bwd(const float* raw, int invert, const float* d_out, float* d_raw, int B,
                                      void* stream);

/* ---- optimiser ------------------------------------------------------------
 * train.py:34,56 tf.keras.optimizers.Adam: p -= lr_t * m / (sqrt(v) + eps),
 * lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t); grad is multiplied by grad_scale
 * first (1/world_size after the sum-allreduce). */
int of_adam_step(float* param, const float* grad, float* m, float* v, size_t n,
                 int step, float lr, float beta1, float beta2, float eps,
                 float grad_scale, void* stream);

/* The same update with the optimiser state in device memory: *state holds the number of COMPLETED steps and the
 * learning rate; the call applies step (state->step + 1) and then increments state->step on the stream.  No host
 * scalar that changes from step to step is a kernel argument, so a captured CUDA graph of the train step replays. */
typedef struct { int step; float lr; } of_adam_state;
int of_adam_step_dev(float* param, const float* grad, float* m, float* v, size_t n, of_adam_state* state,
                     float beta1, float beta2, float eps, float grad_scale, void* stream);

/* ---- training-mode BatchNormalization (OPT-IN) --------------
 * The reference's step runs layers.BatchNormalization (model.py:14; resnet/models.py:19,24) in inference mode
 * (SURVEY F5), which is what of_bn_fold_* implement; these four passes are the batch-statistics variant.
 * x / y / dy / dx: (npix, C) NHWC rows, dtype f32 or bf16, C % 4 == 0 and C / 4 dividing 256; all statistics f32.
 *   Every function takes `shift` (nullable = 0), a per-channel offset the sums are taken about; the layer passes a
 *   snapshot of its moving mean (identical on every replica), which removes the E[x^2] - E[x]^2 cancellation.
 *   stats:       sum[c] = sum_p (x - shift), sumsq[c] = sum_p (x - shift)^2.  Deterministic: per-CTA partials in
 *                `workspace` (of_bn_train_workspace_bytes(C), 16-byte aligned) are added in a fixed order by the last
 *                CTA - no floating-point atomics.  For cross-replica (sync) BN the caller all-reduces both buffers and
 *                passes the GLOBAL count below.
 *   apply:       mean = shift + sum / count, var = sumsq / count - (sum / count)^2 (biased); y = act(gamma (x - mean) /
 *                sqrt(var + eps) + beta); moving_* (nullable pair) <- moving * momentum + batch * (1 - momentum), the
 *                variance unbiased (count / (count - 1)) like TF's fused kernel.  Keras defaults: eps 1e-3, momentum 0.99.
 *                `shift` must not alias moving_mean.
 *   bwd_reduce:  dy' = dy * act'(y); dsum[c] = sum_p dy' (= dbeta), dsum_xhat[c] = sum_p dy' * xhat (= dgamma)
 *                (same deterministic two-stage sum; all-reduce them as well for sync BN).
 *   bwd_apply:   dx = gamma / sqrt(var + eps) * (dy' - dsum / count - xhat * dsum_xhat / count). */
size_t of_bn_train_workspace_bytes(int C);
int of_bn_train_stats(const void* x, const float* shift, float* sum, float* sumsq, size_t npix, int C, int dtype,
                      void* workspace, void* stream);
/* Allocation-free variant of of_bn_train_stats.  It has no counterpart in the reference and was added:
 * some callers cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same kernel, same deterministic result bit for bit, but the per-CTA partials live in a
 * file-scope `static __device__` array inside the library (2 424 896 bytes, held by every GPU that loads the library
 * until exit), so the caller passes only inputs and outputs.  Limit: C <= 256 (larger C returns OF_ERR_SHAPE with a
 * message; use of_bn_train_stats with a caller workspace).  The call performs no cudaMalloc / cudaFree and calls no
 * allocating library routine; the array's device address is looked up once per device (cudaGetSymbolAddress, with
 * that device current) on the first call.  Every call on a device shares the array: calls that overlap on the GPU
 * (different streams without an ordering between them) race; calls in stream order are safe. */
int of_bn_train_stats_static(const void* x, const float* shift, float* sum, float* sumsq, size_t npix, int C, int dtype,
                             void* stream);
int of_bn_train_apply(const void* x, const float* sum, const float* sumsq, const float* shift, float count,
                      const float* gamma, const float* beta, float eps, int act, void* y, float* moving_mean,
                      float* moving_var, float momentum, size_t npix, int C, int dtype, void* stream);
int of_bn_train_bwd_reduce(const void* x, const void* y, const void* dy, const float* sum, const float* sumsq,
                           const float* shift, float count, float eps, int act, float* dsum, float* dsum_xhat,
                           size_t npix, int C, int dtype, void* workspace, void* stream);
int of_bn_train_bwd_apply(const void* x, const void* y, const void* dy, const float* sum, const float* sumsq,
                          const float* shift, float count, const float* gamma, float eps, int act, const float* dsum,
                          const float* dsum_xhat, void* dx, size_t npix, int C, int dtype, void* stream);

/* Opt-in L2 regulariser: the reference declares kernel_regularizer / bias_regularizer = l2(1e-4)
 * on the encoder convs (model.py:7,13; resnet/models.py:5,18,23) but train.py never adds model.losses, so the default
 * step does not call this.  Keras l2(l): grad[i] += 2 l w[i]; *loss (nullable, f32 scalar) += l * sum(w^2). */
int of_l2_regularizer(const float* w, float* grad, size_t n, float l2, float* loss, void* stream);

/* ---- collectives (NCCL over NVLink; nothing in the reference) --------------
 * One communicator per process (rank == GPU).  unique_id is the 128-byte
 * ncclUniqueId produced by of_comm_unique_id() on rank 0 and shipped to the
 * other ranks by the caller (torch.distributed broadcast). */
int of_comm_unique_id(void* host_id128);
int of_comm_init(const void* host_id128, int world_size, int rank, void** comm_out);
int of_comm_allreduce_sum_f32(void* comm, float* buf, size_t n, void* stream);
int of_comm_destroy(void* comm);

#ifdef __cplusplus
}
#endif
#endif /* OFLOW_B200_H */
