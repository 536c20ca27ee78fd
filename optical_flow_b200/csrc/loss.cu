// K6/K7: fused multi-scale warp + photometric (+ smoothness) loss, forward and
// backward in ONE launch.
//
//   K6  loss.py:5-32   LossLayer: resize full-res pair -> warp frame 2 (REF_IJ
//       quirk, model.py:55-73) -> mean |frame1 - warped| per scale, / num_scales.
//   K7  old/loss.py:7-105 MyLoss: image_warp(frame 1) vs frame 2, Charbonnier
//       photometric + 0.2 * Charbonnier first-order flow smoothness, / batch.
//
// No pyramid is materialised: every resized pixel is rebuilt on the fly from
// the 2x2 full-resolution pixels tf.image.resize would read (half-pixel
// centres, no antialias), so HBM sees the image pair once (the coarser scales
// hit L2), the flows once and the flow gradients once.  The scalar is reduced
// deterministically (per-CTA double partials, last CTA sums in index order).
#include <stdlib.h>
#include "common.cuh"

#define LOSS_MAX_SCALES 8
#define LTILE 16

struct ScaleDesc {
  const float* flow;
  float* d_flow;
  int h, w;
  int tiles_x, tiles_y;
  long long tile_begin;   // first global tile index of this scale
  float sy, sx;           // H / h, W / w  (tf.image.resize scale)
  float inv_n;            // K6: 1 / (B*h*w*3*num_scales);  K7: 1 / B
};

struct LossParams {
  ScaleDesc sc[LOSS_MAX_SCALES];
  int num_scales;
  long long total_tiles;
  const float* imgs;
  float* loss;
  void* ws;
  int B, H, W;
  int mode;
};

struct Lerp2 { int lo, hi; float t; };
__device__ __forceinline__ Lerp2 tf_lerp2(int o, float scale, int in_size) {
  float src = __fsub_rn(__fmul_rn(__fadd_rn((float)o, 0.5f), scale), 0.5f);
  float fl = floorf(src);
  Lerp2 l;
  l.lo = max((int)fl, 0);
  l.hi = min((int)ceilf(src), in_size - 1);
  l.t = __fsub_rn(src, fl);
  return l;
}

// one pixel (3 channels starting at channel offset `co`) of the virtual resized image
__device__ __forceinline__ float3 fetch_resized(const float* __restrict__ img_b, int r, int c, int co,
                                                int H, int W, float sy, float sx) {
  Lerp2 ly = tf_lerp2(r, sy, H), lx = tf_lerp2(c, sx, W);
  const float* r0 = img_b + (size_t)ly.lo * W * 6 + co;
  const float* r1 = img_b + (size_t)ly.hi * W * 6 + co;
  const float* tl = r0 + (size_t)lx.lo * 6;
  const float* tr = r0 + (size_t)lx.hi * 6;
  const float* bl = r1 + (size_t)lx.lo * 6;
  const float* br = r1 + (size_t)lx.hi * 6;
  float3 o;
  {
    float a = __ldg(tl), b = __ldg(tr), c2 = __ldg(bl), d = __ldg(br);
    float top = a + (b - a) * lx.t, bot = c2 + (d - c2) * lx.t;
    o.x = top + (bot - top) * ly.t;
  }
  {
    float a = __ldg(tl + 1), b = __ldg(tr + 1), c2 = __ldg(bl + 1), d = __ldg(br + 1);
    float top = a + (b - a) * lx.t, bot = c2 + (d - c2) * lx.t;
    o.y = top + (bot - top) * ly.t;
  }
  {
    float a = __ldg(tl + 2), b = __ldg(tr + 2), c2 = __ldg(bl + 2), d = __ldg(br + 2);
    float top = a + (b - a) * lx.t, bot = c2 + (d - c2) * lx.t;
    o.z = top + (bot - top) * ly.t;
  }
  return o;
}

__device__ __forceinline__ int clampi2(int v, int lo, int hi) { return min(max(v, lo), hi); }
__device__ __forceinline__ float sgn(float v) { return (v > 0.f) ? 1.f : ((v < 0.f) ? -1.f : 0.f); }

// KIND 0: L1 pyramid (loss.py), KIND 1: Charbonnier + smoothness (old/loss.py)
template <int KIND>
__global__ void __launch_bounds__(256) fused_loss_kernel(const __grid_constant__ LossParams P) {
  const int B = P.B, H = P.H, W = P.W;
  float acc = 0.f;
  for (long long tile = blockIdx.x; tile < P.total_tiles; tile += gridDim.x) {
    int s = 0;
#pragma unroll
    for (int k = 1; k < LOSS_MAX_SCALES; ++k)
      if (k < P.num_scales && tile >= P.sc[k].tile_begin) s = k;
    const ScaleDesc& S = P.sc[s];
    long long lt = tile - S.tile_begin;
    int tx = (int)(lt % S.tiles_x);
    long long r = lt / S.tiles_x;
    int ty = (int)(r % S.tiles_y);
    int b = (int)(r / S.tiles_y);
    int i = ty * LTILE + (threadIdx.x >> 4);
    int j = tx * LTILE + (threadIdx.x & 15);
    if (i >= S.h || j >= S.w) continue;
    const int h = S.h, w = S.w;
    size_t pix = ((size_t)b * h + i) * w + j;
    float2 f = __ldg(reinterpret_cast<const float2*>(S.flow) + pix);
    const float* img_b = P.imgs + (size_t)b * H * W * 6;

    int x0, x1, y0, y1;
    float ax, bx, ay, by;
    if (KIND == 0) {
      // model.py:65-71 + transformations.py:93-125
      float x, y;
      if (P.mode == OF_WARP_REF_IJ) { x = __fadd_rn((float)i, f.x); y = __fadd_rn((float)j, f.y); }
      else { x = __fadd_rn((float)j, f.x); y = __fadd_rn((float)i, f.y); }
      int xf = __float2int_rd(x), yf = __float2int_rd(y);
      x0 = clampi2(xf, 0, w - 1);
      x1 = clampi2(xf == 0x7fffffff ? xf : xf + 1, 0, w - 1);
      y0 = clampi2(yf, 0, h - 1);
      y1 = clampi2(yf == 0x7fffffff ? yf : yf + 1, 0, h - 1);
      ax = __fsub_rn((float)x1, x);
      ay = __fsub_rn((float)y1, y);
      bx = __fsub_rn(1.f, ax);
      by = __fsub_rn(1.f, ay);
    } else {
      // old/image_warp.py:28-56
      float fx = floorf(f.x), fy = floorf(f.y);
      long long xl = (long long)j + (long long)__float2int_rd(f.x);
      long long yl = (long long)i + (long long)__float2int_rd(f.y);
      int xi = (int)max(min(xl, 0x7ffffffell), -0x7fffffffll);
      int yi = (int)max(min(yl, 0x7ffffffell), -0x7fffffffll);
      x0 = clampi2(xi, 0, w - 1);
      x1 = clampi2(xi + 1, 0, w - 1);
      y0 = clampi2(yi, 0, h - 1);
      y1 = clampi2(yi + 1, 0, h - 1);
      bx = __fsub_rn(f.x, fx);
      by = __fsub_rn(f.y, fy);
      ax = __fsub_rn(1.f, bx);
      ay = __fsub_rn(1.f, by);
    }
    // K6 warps frame 2 (channels 3..5) and compares with frame 1; K7 the reverse
    const int co_warp = (KIND == 0) ? 3 : 0;
    const int co_ref = (KIND == 0) ? 0 : 3;
    float3 v00 = fetch_resized(img_b, y0, x0, co_warp, H, W, S.sy, S.sx);
    float3 v01 = fetch_resized(img_b, y1, x0, co_warp, H, W, S.sy, S.sx);
    float3 v10 = fetch_resized(img_b, y0, x1, co_warp, H, W, S.sy, S.sx);
    float3 v11 = fetch_resized(img_b, y1, x1, co_warp, H, W, S.sy, S.sx);
    float3 ref = fetch_resized(img_b, i, j, co_ref, H, W, S.sy, S.sx);
    float w00 = ax * ay, w01 = ax * by, w10 = bx * ay, w11 = bx * by;
    float3 wp;
    wp.x = w00 * v00.x + w01 * v01.x + w10 * v10.x + w11 * v11.x;
    wp.y = w00 * v00.y + w01 * v01.y + w10 * v10.y + w11 * v11.y;
    wp.z = w00 * v00.z + w01 * v01.z + w10 * v10.z + w11 * v11.z;
    // d warped / d coord0, coord1
    float3 gx, gy;
    gx.x = ay * (v10.x - v00.x) + by * (v11.x - v01.x);
    gx.y = ay * (v10.y - v00.y) + by * (v11.y - v01.y);
    gx.z = ay * (v10.z - v00.z) + by * (v11.z - v01.z);
    gy.x = ax * (v01.x - v00.x) + bx * (v11.x - v10.x);
    gy.y = ax * (v01.y - v00.y) + bx * (v11.y - v10.y);
    gy.z = ax * (v01.z - v00.z) + bx * (v11.z - v10.z);

    float d0, d1;
    if (KIND == 0) {
      float ex = ref.x - wp.x, ey = ref.y - wp.y, ez = ref.z - wp.z;   // loss.py:28
      acc += (fabsf(ex) + fabsf(ey) + fabsf(ez)) * S.inv_n;
      float sx_ = -sgn(ex) * S.inv_n, sy_ = -sgn(ey) * S.inv_n, sz_ = -sgn(ez) * S.inv_n;
      d0 = sx_ * gx.x + sy_ * gx.y + sz_ * gx.z;
      d1 = sx_ * gy.x + sy_ * gy.y + sz_ * gy.z;
    } else {
      const float eps = 1e-4f;
      float ex = wp.x - ref.x, ey = wp.y - ref.y, ez = wp.z - ref.z;   // old/loss.py:41
      float err = sqrtf(ex * ex + ey * ey + ez * ez + eps);              // :42-43
      float lsum = err;
      float k = S.inv_n / err;
      d0 = k * (ex * gx.x + ey * gx.y + ez * gx.z);
      d1 = k * (ex * gy.x + ey * gy.y + ez * gy.z);
      // smoothness (old/loss.py:70-94): each forward difference is counted at its
      // first pixel; the gradient is gathered from the <= 4 differences touching (i, j)
      const float2* fl = reinterpret_cast<const float2*>(S.flow) + (size_t)b * h * w;
      const float lam = 0.2f * S.inv_n;
      if (i + 1 < h) {
        float2 n = __ldg(fl + (size_t)(i + 1) * w + j);
        float dx = f.x - n.x, dy = f.y - n.y;
        float sx2 = sqrtf(dx * dx + eps), sy2 = sqrtf(dy * dy + eps);
        lsum += 0.2f * (sx2 + sy2);
        d0 += lam * dx / sx2;
        d1 += lam * dy / sy2;
      }
      if (i > 0) {
        float2 n = __ldg(fl + (size_t)(i - 1) * w + j);
        float dx = n.x - f.x, dy = n.y - f.y;
        d0 -= lam * dx / sqrtf(dx * dx + eps);
        d1 -= lam * dy / sqrtf(dy * dy + eps);
      }
      if (j + 1 < w) {
        float2 n = __ldg(fl + (size_t)i * w + j + 1);
        float dx = f.x - n.x, dy = f.y - n.y;
        float sx2 = sqrtf(dx * dx + eps), sy2 = sqrtf(dy * dy + eps);
        lsum += 0.2f * (sx2 + sy2);
        d0 += lam * dx / sx2;
        d1 += lam * dy / sy2;
      }
      if (j > 0) {
        float2 n = __ldg(fl + (size_t)i * w + j - 1);
        float dx = n.x - f.x, dy = n.y - f.y;
        d0 -= lam * dx / sqrtf(dx * dx + eps);
        d1 -= lam * dy / sqrtf(dy * dy + eps);
      }
      acc += lsum * S.inv_n;
    }
    if (S.d_flow != nullptr) reinterpret_cast<float2*>(S.d_flow)[pix] = make_float2(d0, d1);
  }

  // CTA reduction in double, then the deterministic grid sum
  __shared__ double s_part[8];
  double d = (double)acc;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = d;
  __syncthreads();
  double cta = 0.0;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < 8; ++k) cta += s_part[k];
  }
  double total;
  if (grid_sum_last(cta, P.ws, &total)) *P.loss = (float)total;
}

#define LOSS_MAX_GRID (OF_NUM_SMS * 8)

// fast path (loss_fused.cu): H, W multiples of 16, power-of-two pyramid
size_t of_lf_workspace_bytes(int B, int H, int W, int num_scales);
bool of_lf_supported(int B, int H, int W, int num_scales, const int* hs, const int* ws);
int of_lf2_run(const float* imgs, const float* const* flows, float* const* d_flows, const int* hs, const int* ws,
               int num_scales, float* loss, int B, int H, int W, int mode, void* workspace, void* stream);
int of_lf_run(const float* imgs, const float* const* flows, float* const* d_flows, const int* hs, const int* ws,
              int num_scales, float* loss, int B, int H, int W, int mode, void* workspace, void* stream);

static int run_loss(int kind, const float* imgs, const float* const* flows, float* const* d_flows, const int* hs,
                    const int* ws, int num_scales, float* loss, int B, int H, int W, int mode, void* workspace,
                    void* stream) {
  OF_REQUIRE(imgs && flows && hs && ws && loss && workspace, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(num_scales >= 1 && num_scales <= LOSS_MAX_SCALES, OF_ERR_ARG, "num_scales %d not in [1,%d]",
             num_scales, LOSS_MAX_SCALES);
  OF_REQUIRE(B > 0 && H > 0 && W > 0, OF_ERR_SHAPE, "bad dims");
  LossParams P;
  memset(&P, 0, sizeof(P));
  long long tiles = 0;
  for (int s = 0; s < num_scales; ++s) {
    OF_REQUIRE(flows[s] != nullptr, OF_ERR_ARG, "flow %d is null", s);
    OF_REQUIRE(hs[s] > 0 && ws[s] > 0, OF_ERR_SHAPE, "bad scale dims");
    if (kind == 0) {
      // loss.py:11-15 shape asserts
      int eh = (int)((double)H / (double)(1 << (s + 1))), ew = (int)((double)W / (double)(1 << (s + 1)));
      OF_REQUIRE(hs[s] == eh && ws[s] == ew, OF_ERR_SHAPE, "flow %d is %dx%d, LossLayer expects %dx%d", s, hs[s],
                 ws[s], eh, ew);
    } else {
      // old/loss.py:28-32 shape asserts (coarse -> fine, last one full resolution)
      int div = 1 << (num_scales - 1 - s);
      OF_REQUIRE(hs[s] == H / div && ws[s] == W / div, OF_ERR_SHAPE, "flow %d is %dx%d, MyLoss expects %dx%d", s,
                 hs[s], ws[s], H / div, W / div);
    }
    ScaleDesc& S = P.sc[s];
    S.flow = flows[s];
    S.d_flow = d_flows ? d_flows[s] : nullptr;
    S.h = hs[s];
    S.w = ws[s];
    S.tiles_x = (ws[s] + LTILE - 1) / LTILE;
    S.tiles_y = (hs[s] + LTILE - 1) / LTILE;
    S.tile_begin = tiles;
    tiles += (long long)B * S.tiles_x * S.tiles_y;
    S.sy = (float)H / (float)hs[s];
    S.sx = (float)W / (float)ws[s];
    S.inv_n = kind == 0 ? (float)(1.0 / ((double)B * hs[s] * ws[s] * 3.0 * num_scales)) : (float)(1.0 / (double)B);
  }
  P.num_scales = num_scales;
  P.total_tiles = tiles;
  P.imgs = imgs;
  P.loss = loss;
  P.ws = reinterpret_cast<char*>(workspace) + 16384;   // [0,16K) belongs to the fast path's counters
  P.B = B; P.H = H; P.W = W;
  P.mode = mode;
  long long g = tiles < LOSS_MAX_GRID ? tiles : LOSS_MAX_GRID;
  if (kind == 0)
    fused_loss_kernel<0><<<(int)g, 256, 0, (cudaStream_t)stream>>>(P);
  else
    fused_loss_kernel<1><<<(int)g, 256, 0, (cudaStream_t)stream>>>(P);
  return of_launch_check("fused_loss_kernel");
}

extern "C" {

// workspace: header (ticket, per-image counters, one double per CTA) + the L2-resident float4
// pyramid of the fast path.  Must be zeroed once by the caller after allocation (the kernels
// re-arm their counters themselves).  Sized for (B, H, W, num_scales); 16-byte aligned.
size_t of_loss_workspace_bytes(int B, int H, int W, int num_scales) {
  size_t generic = 32768;   // two 16 KB headers (fast path, generic path: ticket + one double per CTA)
  size_t fast = (B > 0 && H > 0 && W > 0 && num_scales > 0) ? of_lf_workspace_bytes(B, H, W, num_scales) : 0;
  return generic > fast ? generic : fast;
}

int of_loss_l1_pyramid_fwdbwd(const float* imgs, const float* const* flows, float* const* d_flows, const int* hs,
                              const int* ws, int num_scales, float* loss, int B, int H, int W, int mode,
                              void* workspace, void* stream) {
  OF_REQUIRE(mode == OF_WARP_REF_IJ || mode == OF_WARP_XY, OF_ERR_ARG, "bad warp mode %d", mode);
  OF_REQUIRE(imgs && flows && hs && ws && loss && workspace, OF_ERR_ARG, "null pointer argument");
  if (of_lf_supported(B, H, W, num_scales, hs, ws) && !of_tune_env("OF_LOSS_GENERIC")) {
    for (int s = 0; s < num_scales; ++s) OF_REQUIRE(flows[s] != nullptr, OF_ERR_ARG, "flow %d is null", s);
    // warp-specialised kernel (loss_fused2.cu) by default; option OF_LF_V1 selects the first-generation kernel
    if (of_tune_env("OF_LF_V1"))
      return of_lf_run(imgs, flows, d_flows, hs, ws, num_scales, loss, B, H, W, mode, workspace, stream);
    return of_lf2_run(imgs, flows, d_flows, hs, ws, num_scales, loss, B, H, W, mode, workspace, stream);
  }
  return run_loss(0, imgs, flows, d_flows, hs, ws, num_scales, loss, B, H, W, mode, workspace, stream);
}

int of_loss_charbonnier_smooth_fwdbwd(const float* imgs, const float* const* flows, float* const* d_flows,
                                      const int* hs, const int* ws, int num_scales, float* loss, int B, int H,
                                      int W, void* workspace, void* stream) {
  return run_loss(1, imgs, flows, d_flows, hs, ws, num_scales, loss, B, H, W, 0, workspace, stream);
}

}  // extern "C"
