// On-device replacement of the reference's per-item preprocessing, data_reader.py:45-64 (read_item) and
// :36-41 (batch assembly): two decoded uint8 BGR frames -> cv2.resize(INTER_LINEAR) -> float32 / 255 -> minus
// the [123,117,104]/255 means -> the (H, W, 6) float32 slice of the batch, with the pair-order swap of :46-51
// as a flag.  One launch per item; HBM-bound (reads 2 x h x w x 3 bytes, writes H x W x 24 bytes).
//
// cv2.resize on uint8 is OpenCV's 11-bit fixed-point bilinear; it is reproduced bit for bit (see
// oracle/ref_data.py for the restated algorithm and tests/golden/data.npz for vectors from the real cv2):
// every rounding step is an explicit _rn / _rd intrinsic so no FMA contraction can change a result.
#include "common.cuh"

struct AxisTap { int s0, s1, c0, c1; };

// resize.cpp, INTER_LINEAR: f = (float)((d + 0.5) * scale - 0.5); s = floor(f); f -= s.
// x axis: outside the image (s, f) is reset; y axis: weights kept, row indices clipped on fetch.
__device__ __forceinline__ AxisTap axis_tap(int d, double scale, int src, bool vertical) {
  const double fd = __dadd_rn(__dmul_rn(__dadd_rn((double)d, 0.5), scale), -0.5);
  float f = __double2float_rn(fd);
  int s = __float2int_rd(f);
  f = __fsub_rn(f, (float)s);
  AxisTap t;
  if (vertical) {
    t.s0 = min(max(s, 0), src - 1);
    t.s1 = min(max(s + 1, 0), src - 1);
  } else {
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= src - 1) { s = src - 1; f = 0.f; }
    t.s0 = s;
    t.s1 = min(s + 1, src - 1);
  }
  t.c0 = __float2int_rn(__fmul_rn(__fsub_rn(1.f, f), 2048.f));     // cvRound = round half to even
  t.c1 = __float2int_rn(__fmul_rn(f, 2048.f));
  return t;
}

// blockIdx.y = item of the batch: frames (items, 2, sh, sw, 3) uint8, out (items, H, W, 6); bit i of flip_mask swaps
// pair i.  With item_stride == 0 (single-item entry point) img1 / img2 are the two frames themselves.
__global__ void __launch_bounds__(256)
preprocess_pair_kernel(const uint8_t* __restrict__ img1, const uint8_t* __restrict__ img2, size_t item_stride,
                       unsigned long long flip_mask, int sh, int sw, float* __restrict__ out, int H, int W,
                       double scale_y, double scale_x, double m0, double m1, double m2) {
  const size_t total = (size_t)H * W * 2;
  if (item_stride != 0) {
    const uint8_t* a = img1 + (size_t)blockIdx.y * item_stride;
    const uint8_t* b = a + (size_t)sh * sw * 3;
    const bool flip = (flip_mask >> blockIdx.y) & 1ull;
    img1 = flip ? b : a;
    img2 = flip ? a : b;
    out += (size_t)blockIdx.y * H * W * 6;
  }
  const bool same = sh == H && sw == W;                              // cv2.resize returns a copy
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const int which = (int)(e & 1);
    const size_t p = e >> 1;
    const int x = (int)(p % W), y = (int)(p / W);
    const uint8_t* src = which ? img2 : img1;
    int v[3];
    if (same) {
      const uint8_t* q = src + ((size_t)y * sw + x) * 3;
      v[0] = q[0]; v[1] = q[1]; v[2] = q[2];
    } else {
      const AxisTap tx = axis_tap(x, scale_x, sw, false), ty = axis_tap(y, scale_y, sh, true);
      const uint8_t* r0 = src + (size_t)ty.s0 * sw * 3;
      const uint8_t* r1 = src + (size_t)ty.s1 * sw * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int h0 = r0[tx.s0 * 3 + c] * tx.c0 + r0[tx.s1 * 3 + c] * tx.c1;        // HResizeLinear (int32)
        const int h1 = r1[tx.s0 * 3 + c] * tx.c0 + r1[tx.s1 * 3 + c] * tx.c1;
        const int r = (((ty.c0 * (h0 >> 4)) >> 16) + ((ty.c1 * (h1 >> 4)) >> 16) + 2) >> 2;   // VResizeLinear
        v[c] = min(max(r, 0), 255);
      }
    }
    const double mean[3] = {m0, m1, m2};
    float* o = out + p * 6 + which * 3;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float f = __fdiv_rn((float)v[c], 255.0f);                 // img.astype(float32) / 255.0
      o[c] = __double2float_rn(__dsub_rn((double)f, mean[c]));       // float32 - float64 means, stored as float32
    }
  }
}

extern "C" {

int of_preprocess_pair(const uint8_t* img1, const uint8_t* img2, int src_h, int src_w, float* out, int H, int W,
                       int flip, void* stream) {
  OF_REQUIRE(img1 && img2 && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(src_h > 0 && src_w > 0 && H > 0 && W > 0, OF_ERR_SHAPE, "bad image dims %dx%d -> %dx%d", src_h, src_w, H, W);
  // resize.cpp: inv_scale = dsize / ssize; scale = 1. / inv_scale
  const double scale_x = 1.0 / ((double)W / (double)src_w), scale_y = 1.0 / ((double)H / (double)src_h);
  size_t total = (size_t)H * W * 2;
  size_t blocks = (total + 255) / 256;
  if (blocks > (size_t)OF_NUM_SMS * 8) blocks = (size_t)OF_NUM_SMS * 8;
  preprocess_pair_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(flip ? img2 : img1, flip ? img1 : img2, 0, 0ull,
                                                                        src_h, src_w, out, H, W, scale_y, scale_x,
                                                                        123.0 / 255.0, 117.0 / 255.0, 104.0 / 255.0);
  return of_launch_check("preprocess_pair_kernel");
}

int of_preprocess_batch(const uint8_t* frames, int B, int src_h, int src_w, float* out, int H, int W,
                        const uint8_t* flips, void* stream) {
  OF_REQUIRE(frames && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(B > 0 && src_h > 0 && src_w > 0 && H > 0 && W > 0, OF_ERR_SHAPE, "bad batch dims");
  const double scale_x = 1.0 / ((double)W / (double)src_w), scale_y = 1.0 / ((double)H / (double)src_h);
  const size_t item = (size_t)2 * src_h * src_w * 3;
  size_t total = (size_t)H * W * 2;
  size_t blocks = (total + 255) / 256;
  const size_t cap = ((size_t)OF_NUM_SMS * 16 + B - 1) / B;           // ~16 CTAs per SM over the whole batch
  if (blocks > cap) blocks = cap;
  for (int b0 = 0; b0 < B; b0 += 64) {                                // 64 flip bits per launch
    const int nb = B - b0 < 64 ? B - b0 : 64;
    unsigned long long mask = 0ull;
    if (flips)
      for (int i = 0; i < nb; ++i) mask |= (unsigned long long)(flips[b0 + i] != 0) << i;
    preprocess_pair_kernel<<<dim3((unsigned)blocks, (unsigned)nb), 256, 0, (cudaStream_t)stream>>>(
        frames + (size_t)b0 * item, nullptr, item, mask, src_h, src_w, out + (size_t)b0 * H * W * 6, H, W, scale_y, scale_x,
        123.0 / 255.0, 117.0 / 255.0, 104.0 / 255.0);
    if (int r = of_launch_check("preprocess_pair_kernel")) return r;
  }
  return 0;
}

}  // extern "C"
