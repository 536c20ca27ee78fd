// K5/K9/K11/K13 + elementwise: bilinear resize, flow upscale (fwd/bwd),
// 2x2 max-pool (fwd/bwd), add+ReLU, activation backward, Adam.
// All HBM-bound: coalesced NHWC access, float4 vectors where the shape allows,
// grids sized in multiples of the SM count.
#include "common.cuh"

static inline int ew_grid(size_t work_items, int threads = 256) {
  size_t blocks = (work_items + threads - 1) / threads;
  size_t cap = (size_t)OF_NUM_SMS * 16;
  if (blocks > cap) blocks = cap;
  return (int)(blocks < 1 ? 1 : blocks);
}

// ---- tf.image.resize (bilinear, half-pixel centres, no antialias) -----------
// loss.py:18, old/loss.py:15-22, model.py:77.  TF's
// compute_interpolation_weights: in = (o + 0.5) * scale - 0.5,
// lower = max(floor(in), 0), upper = min(ceil(in), size - 1), lerp = in - floor(in).
struct Lerp { int lo, hi; float t; };
__device__ __forceinline__ Lerp tf_lerp(int o, float scale, int in_size) {
  float src = __fsub_rn(__fmul_rn(__fadd_rn((float)o, 0.5f), scale), 0.5f);
  float fl = floorf(src);
  Lerp l;
  l.lo = max((int)fl, 0);
  l.hi = min((int)ceilf(src), in_size - 1);
  l.t = __fsub_rn(src, fl);
  return l;
}

__global__ void __launch_bounds__(256) resize_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                     int B, int H, int W, int C, int oh, int ow, float mul) {
  const float sy = (float)H / (float)oh, sx = (float)W / (float)ow;
  size_t total = (size_t)B * oh * ow * C;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(e % C);
    size_t r = e / C;
    int x = (int)(r % ow); r /= ow;
    int y = (int)(r % oh);
    int b = (int)(r / oh);
    Lerp ly = tf_lerp(y, sy, H), lx = tf_lerp(x, sx, W);
    const float* base = in + (size_t)b * H * W * C + c;
    float tl = __ldg(base + ((size_t)ly.lo * W + lx.lo) * C), tr = __ldg(base + ((size_t)ly.lo * W + lx.hi) * C);
    float bl = __ldg(base + ((size_t)ly.hi * W + lx.lo) * C), br = __ldg(base + ((size_t)ly.hi * W + lx.hi) * C);
    float top = tl + (tr - tl) * lx.t;
    float bot = bl + (br - bl) * lx.t;
    out[e] = (top + (bot - top) * ly.t) * mul;
  }
}

// adjoint of the x2 upscale: every low-res pixel gathers from the <=5x5 high-res
// pixels whose stencil can touch it (deterministic, no atomics).
__global__ void __launch_bounds__(256) upscale2_bwd_kernel(const float* __restrict__ d_out, float* __restrict__ d_in,
                                                           int B, int h, int w, int C, float mul) {
  const int oh = 2 * h, ow = 2 * w;
  size_t total = (size_t)B * h * w * C;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(e % C);
    size_t r = e / C;
    int x = (int)(r % w); r /= w;
    int y = (int)(r % h);
    int b = (int)(r / h);
    float acc = 0.f;
    for (int oy = max(2 * y - 2, 0); oy <= min(2 * y + 2, oh - 1); ++oy) {
      Lerp ly = tf_lerp(oy, 0.5f, h);
      float wy = (ly.lo == y ? 1.f - ly.t : 0.f) + (ly.hi == y ? ly.t : 0.f);
      if (wy == 0.f) continue;
      for (int ox = max(2 * x - 2, 0); ox <= min(2 * x + 2, ow - 1); ++ox) {
        Lerp lx = tf_lerp(ox, 0.5f, w);
        float wx = (lx.lo == x ? 1.f - lx.t : 0.f) + (lx.hi == x ? lx.t : 0.f);
        if (wx == 0.f) continue;
        acc += wy * wx * __ldg(d_out + (((size_t)b * oh + oy) * ow + ox) * C + c);
      }
    }
    d_in[e] = acc * mul;
  }
}

// ---- MaxPool2D() 2x2 / 2 / valid  (model.py:17) ------------------------------
template <typename T>
__global__ void __launch_bounds__(256) maxpool2_fwd_kernel(const T* __restrict__ in, T* __restrict__ out,
                                                           int B, int H, int W, int C) {
  const int oh = H / 2, ow = W / 2, CV = C / 4;
  size_t total = (size_t)B * oh * ow * CV;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int v = (int)(e % CV);
    size_t r = e / CV;
    int x = (int)(r % ow); r /= ow;
    int y = (int)(r % oh);
    int b = (int)(r / oh);
    const T* p = in + (((size_t)b * H + 2 * y) * W + 2 * x) * C + v * 4;
    float4 a = ld4<T>(p), bq = ld4<T>(p + C), c = ld4<T>(p + (size_t)W * C), d = ld4<T>(p + (size_t)W * C + C);
    float4 m;
    m.x = fmaxf(fmaxf(a.x, bq.x), fmaxf(c.x, d.x));
    m.y = fmaxf(fmaxf(a.y, bq.y), fmaxf(c.y, d.y));
    m.z = fmaxf(fmaxf(a.z, bq.z), fmaxf(c.z, d.z));
    m.w = fmaxf(fmaxf(a.w, bq.w), fmaxf(c.w, d.w));
    st4<T>(out + (((size_t)b * oh + y) * ow + x) * C + v * 4, m);
  }
}

// bf16, 8 channels (16 bytes) per access (the 8-byte version ran at 75 % of the HBM peak)
__global__ void __launch_bounds__(256) maxpool2_fwd8_kernel(const __nv_bfloat16* __restrict__ in,
                                                            __nv_bfloat16* __restrict__ out, int B, int H, int W, int C) {
  const int oh = H / 2, ow = W / 2, CV = C / 8;
  size_t total = (size_t)B * oh * ow * CV;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int v = (int)(e % CV);
    size_t r = e / CV;
    int x = (int)(r % ow); r /= ow;
    int y = (int)(r % oh);
    int b = (int)(r / oh);
    const __nv_bfloat16* p = in + (((size_t)b * H + 2 * y) * W + 2 * x) * C + v * 8;
    float a[8], bq[8], c[8], d[8], m[8];
    ld8_bf16(p, a); ld8_bf16(p + C, bq); ld8_bf16(p + (size_t)W * C, c); ld8_bf16(p + (size_t)W * C + C, d);
#pragma unroll
    for (int k = 0; k < 8; ++k) m[k] = fmaxf(fmaxf(a[k], bq[k]), fmaxf(c[k], d[k]));
    st8_bf16(out + (((size_t)b * oh + y) * ow + x) * C + v * 8, m);
  }
}

__device__ __forceinline__ void route4(float a, float b, float c, float d, float g, float& ga, float& gb,
                                       float& gc, float& gd) {
  // gradient to the first maximum in window scan order (row-major)
  float m = fmaxf(fmaxf(a, b), fmaxf(c, d));
  ga = gb = gc = gd = 0.f;
  if (a == m) ga = g;
  else if (b == m) gb = g;
  else if (c == m) gc = g;
  else gd = g;
}

template <typename T>
__global__ void __launch_bounds__(256) maxpool2_bwd_kernel(const T* __restrict__ in, const T* __restrict__ d_out,
                                                           T* __restrict__ d_in, int B, int H, int W, int C) {
  const int oh = H / 2, ow = W / 2, CV = C / 4;
  size_t total = (size_t)B * oh * ow * CV;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int v = (int)(e % CV);
    size_t r = e / CV;
    int x = (int)(r % ow); r /= ow;
    int y = (int)(r % oh);
    int b = (int)(r / oh);
    size_t o = (((size_t)b * H + 2 * y) * W + 2 * x) * C + v * 4;
    const T* p = in + o;
    float4 a = ld4<T>(p), bq = ld4<T>(p + C), c = ld4<T>(p + (size_t)W * C), d = ld4<T>(p + (size_t)W * C + C);
    float4 g = ld4<T>(d_out + (((size_t)b * oh + y) * ow + x) * C + v * 4);
    float4 ga, gb, gc, gd;
    route4(a.x, bq.x, c.x, d.x, g.x, ga.x, gb.x, gc.x, gd.x);
    route4(a.y, bq.y, c.y, d.y, g.y, ga.y, gb.y, gc.y, gd.y);
    route4(a.z, bq.z, c.z, d.z, g.z, ga.z, gb.z, gc.z, gd.z);
    route4(a.w, bq.w, c.w, d.w, g.w, ga.w, gb.w, gc.w, gd.w);
    st4<T>(d_in + o, ga);
    st4<T>(d_in + o + C, gb);
    st4<T>(d_in + o + (size_t)W * C, gc);
    st4<T>(d_in + o + (size_t)W * C + C, gd);
  }
}

__device__ __forceinline__ float act_grad(float y, int act);
__device__ __forceinline__ float4 add_mask4(float4 g, float4 a, float4 y, int act) {
  return make_float4((g.x + a.x) * act_grad(y.x, act), (g.y + a.y) * act_grad(y.y, act), (g.z + a.z) * act_grad(y.z, act),
                     (g.w + a.w) * act_grad(y.w, act));
}
// model.py:12-17 backward in one pass: dz = (maxpool2_bwd(in, d_out) + add) * act'(in), where `in` is the
// stem's post-activation output (pool input), `add` the gradient reaching it directly (the finest decoder
// level reads it too) - replaces three full-tensor passes (pool bwd, accumulate, activation bwd).
template <typename T>
__global__ void __launch_bounds__(256) maxpool2_bwd_add_act_kernel(const T* __restrict__ in, const T* __restrict__ d_out,
                                                                   const T* __restrict__ add, T* __restrict__ dz, int B,
                                                                   int H, int W, int C, int act) {
  const int oh = H / 2, ow = W / 2, CV = C / 4;
  size_t total = (size_t)B * oh * ow * CV;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int v = (int)(e % CV);
    size_t r = e / CV;
    int x = (int)(r % ow); r /= ow;
    int y = (int)(r % oh);
    int b = (int)(r / oh);
    size_t o = (((size_t)b * H + 2 * y) * W + 2 * x) * C + v * 4;
    const size_t o1 = o + C, o2 = o + (size_t)W * C, o3 = o2 + C;
    float4 a = ld4<T>(in + o), bq = ld4<T>(in + o1), c = ld4<T>(in + o2), d = ld4<T>(in + o3);
    float4 g = ld4<T>(d_out + (((size_t)b * oh + y) * ow + x) * C + v * 4);
    float4 ga, gb, gc, gd;
    route4(a.x, bq.x, c.x, d.x, g.x, ga.x, gb.x, gc.x, gd.x);
    route4(a.y, bq.y, c.y, d.y, g.y, ga.y, gb.y, gc.y, gd.y);
    route4(a.z, bq.z, c.z, d.z, g.z, ga.z, gb.z, gc.z, gd.z);
    route4(a.w, bq.w, c.w, d.w, g.w, ga.w, gb.w, gc.w, gd.w);
    st4<T>(dz + o, add_mask4(ga, ld4<T>(add + o), a, act));
    st4<T>(dz + o1, add_mask4(gb, ld4<T>(add + o1), bq, act));
    st4<T>(dz + o2, add_mask4(gc, ld4<T>(add + o2), c, act));
    st4<T>(dz + o3, add_mask4(gd, ld4<T>(add + o3), d, act));
  }
}

// bf16, 8 channels (16 bytes) per access: twice the bytes in flight per thread (the 8-byte version ran at 48 % of
// the HBM peak).  In the step (16 x 192 x 640 x 64: 819 MB of reads + writes) it takes 125 us = 6.5 TB/s, the measured HBM
// peak; a word-by-word form at 63 registers / 4 CTAs per SM ran at the same speed (round 2, call 75) and was not kept.
__global__ void __launch_bounds__(256) maxpool2_bwd_add_act8_kernel(const __nv_bfloat16* __restrict__ in,
                                                                    const __nv_bfloat16* __restrict__ d_out,
                                                                    const __nv_bfloat16* __restrict__ add,
                                                                    __nv_bfloat16* __restrict__ dz, int B, int H, int W,
                                                                    int C, int act) {
  const int oh = H / 2, ow = W / 2, CV = C / 8;
  size_t total = (size_t)B * oh * ow * CV;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int v = (int)(e % CV);
    size_t r = e / CV;
    int x = (int)(r % ow); r /= ow;
    int y = (int)(r % oh);
    int b = (int)(r / oh);
    const size_t o0 = (((size_t)b * H + 2 * y) * W + 2 * x) * C + v * 8;
    const size_t o1 = o0 + C, o2 = o0 + (size_t)W * C, o3 = o2 + C;
    float a[8], bq[8], c[8], d[8], g[8], s0[8], s1[8], s2[8], s3[8];
    ld8_bf16(in + o0, a); ld8_bf16(in + o1, bq); ld8_bf16(in + o2, c); ld8_bf16(in + o3, d);
    ld8_bf16(d_out + (((size_t)b * oh + y) * ow + x) * C + v * 8, g);
    ld8_bf16(add + o0, s0); ld8_bf16(add + o1, s1); ld8_bf16(add + o2, s2); ld8_bf16(add + o3, s3);
    float r0[8], r1[8], r2[8], r3[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float ga, gb, gc, gd;
      route4(a[k], bq[k], c[k], d[k], g[k], ga, gb, gc, gd);
      r0[k] = (ga + s0[k]) * act_grad(a[k], act);
      r1[k] = (gb + s1[k]) * act_grad(bq[k], act);
      r2[k] = (gc + s2[k]) * act_grad(c[k], act);
      r3[k] = (gd + s3[k]) * act_grad(d[k], act);
    }
    st8_bf16(dz + o0, r0); st8_bf16(dz + o1, r1); st8_bf16(dz + o2, r2); st8_bf16(dz + o3, r3);
  }
}

// ---- bf16 hi / lo split (conv_mode "bf16x3") ----------------------------------------------------------------------
// v = hi + lo + O(2^-16 |v|): hi = bf16(v), lo = bf16(v - hi).  Pixels of C channels are written with pixel stride C_pad
// (>= C, padding channels zero).  OUT = __nv_bfloat16 (activations / gradients, the tensor-core operands) or float
// (weights: values already bf16-representable, packed by of_conv_tc_pack_weights without further rounding).
template <typename OUT>
__global__ void __launch_bounds__(256) split_bf16_kernel(const float* __restrict__ src, OUT* __restrict__ hi,
                                                         OUT* __restrict__ lo, size_t npix, int C, int C_pad) {
  const size_t total = npix * (size_t)C_pad;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    const size_t pix = e / C_pad;
    const int c = (int)(e - pix * C_pad);
    float h = 0.f, l = 0.f;
    if (c < C) {
      const float v = src[pix * C + c];
      h = __bfloat162float(__float2bfloat16_rn(v));
      l = __bfloat162float(__float2bfloat16_rn(v - h));
    }
    if constexpr (sizeof(OUT) == 2) { hi[e] = __float2bfloat16_rn(h); lo[e] = __float2bfloat16_rn(l); }
    else { hi[e] = h; lo[e] = l; }
  }
}

// unpadded tensors (C_pad == C), 4 elements per thread: 16-byte loads, 8-byte bf16 / 16-byte fp32 stores
template <typename OUT>
__global__ void __launch_bounds__(256) split_bf16_flat4_kernel(const float* __restrict__ src, OUT* __restrict__ hi,
                                                               OUT* __restrict__ lo, size_t n4) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n4; e += (size_t)gridDim.x * blockDim.x) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(src) + e);
    const float vv[4] = {v.x, v.y, v.z, v.w};
    float h[4], l[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      h[k] = __bfloat162float(__float2bfloat16_rn(vv[k]));
      l[k] = __bfloat162float(__float2bfloat16_rn(vv[k] - h[k]));
    }
    st4<OUT>(hi + e * 4, make_float4(h[0], h[1], h[2], h[3]));
    st4<OUT>(lo + e * 4, make_float4(l[0], l[1], l[2], l[3]));
  }
}

// ---- elementwise -------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) add_relu_kernel(const T* __restrict__ a, const T* __restrict__ b,
                                                       T* __restrict__ out, size_t n4) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n4; e += (size_t)gridDim.x * blockDim.x) {
    float4 x = ld4<T>(a + e * 4), y = ld4<T>(b + e * 4);
    st4<T>(out + e * 4, make_float4(fmaxf(x.x + y.x, 0.f), fmaxf(x.y + y.y, 0.f), fmaxf(x.z + y.z, 0.f),
                                    fmaxf(x.w + y.w, 0.f)));
  }
}

// bf16, 8 elements (16 bytes) per access
__global__ void __launch_bounds__(256) add_relu8_kernel(const __nv_bfloat16* __restrict__ a, const __nv_bfloat16* __restrict__ b,
                                                        __nv_bfloat16* __restrict__ out, size_t n8) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n8; e += (size_t)gridDim.x * blockDim.x) {
    float x[8], y[8], r[8];
    ld8_bf16(a + e * 8, x); ld8_bf16(b + e * 8, y);
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = fmaxf(x[k] + y[k], 0.f);
    st8_bf16(out + e * 8, r);
  }
}

__device__ __forceinline__ float act_grad(float y, int act) {
  // ReLU: tf grad = (y > 0); LeakyReLU(0.3): y > 0 ? 1 : 0.3 (sign(y) == sign(z))
  if (act == OF_ACT_RELU) return y > 0.f ? 1.f : 0.f;
  if (act == OF_ACT_LEAKY) return y > 0.f ? 1.f : 0.3f;
  return 1.f;
}

template <typename T>
__global__ void __launch_bounds__(256) act_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy,
                                                      T* __restrict__ dz, size_t n4, int act) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n4; e += (size_t)gridDim.x * blockDim.x) {
    float4 a = ld4<T>(y + e * 4), g = ld4<T>(dy + e * 4);
    st4<T>(dz + e * 4, make_float4(g.x * act_grad(a.x, act), g.y * act_grad(a.y, act), g.z * act_grad(a.z, act),
                                   g.w * act_grad(a.w, act)));
  }
}

__global__ void __launch_bounds__(256) act_bwd8_kernel(const __nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ dy,
                                                       __nv_bfloat16* __restrict__ dz, size_t n8, int act) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n8; e += (size_t)gridDim.x * blockDim.x) {
    float a[8], g[8], r[8];
    ld8_bf16(y + e * 8, a); ld8_bf16(dy + e * 8, g);
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = g[k] * act_grad(a[k], act);
    st8_bf16(dz + e * 8, r);
  }
}

// activation derivative from a sign-bit mask (16 B per pixel, bit c = activated channel c > 0): one thread per 8
// channels of a pixel (16 bytes of bf16 / 32 bytes of fp32), the mask word is a broadcast within the pixel's threads
template <typename T>
__global__ void __launch_bounds__(256) act_bwd_bits_kernel(const uint32_t* __restrict__ bits, const T* __restrict__ dy,
                                                           T* __restrict__ dz, size_t n8, int c8, float slope) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n8; e += (size_t)gridDim.x * blockDim.x) {
    const size_t pix = e / c8;
    const int g = (int)(e - pix * c8);                       // group of 8 channels within the pixel
    const uint32_t m = (__ldg(bits + pix * 4 + (g >> 2)) >> ((g & 3) * 8)) & 0xffu;
    float v[8];
    if constexpr (sizeof(T) == 2) ld8_bf16(reinterpret_cast<const __nv_bfloat16*>(dy) + e * 8, v);
    else {
      const float4 a = ld4<T>(dy + e * 8), b = ld4<T>(dy + e * 8 + 4);
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] *= ((m >> k) & 1u) ? 1.f : slope;
    if constexpr (sizeof(T) == 2) st8_bf16(reinterpret_cast<__nv_bfloat16*>(dz) + e * 8, v);
    else {
      st4<T>(dz + e * 8, make_float4(v[0], v[1], v[2], v[3]));
      st4<T>(dz + e * 8 + 4, make_float4(v[4], v[5], v[6], v[7]));
    }
  }
}

// ---- Adam (train.py:34,56; Keras epsilon 1e-7, TF update form) ----------------
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                   float* __restrict__ m, float* __restrict__ v, size_t n,
                                                   float lr_t, float omb1, float omb2, float eps, float gscale) {
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    float gr = g[e] * gscale;
    float mm = m[e], vv = v[e];
    mm = mm + (gr - mm) * omb1;
    vv = vv + (gr * gr - vv) * omb2;
    m[e] = mm;
    v[e] = vv;
    p[e] = p[e] - lr_t * mm / (sqrtf(vv) + eps);
  }
}

// Adam with the step count and learning rate in DEVICE memory (of_adam_state), so a CUDA graph of the whole train
// step can be replayed: nothing about the optimiser is baked into the captured kernel arguments.
__global__ void __launch_bounds__(256) adam_dev_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                       float* __restrict__ m, float* __restrict__ v, size_t n,
                                                       const of_adam_state* __restrict__ st, float beta1, float beta2,
                                                       float eps, float gscale) {
  const int step = st->step + 1;                                   // 1-based count of THIS update
  const float lr_t = (float)((double)st->lr * sqrt(1.0 - pow((double)beta2, (double)step)) /
                             (1.0 - pow((double)beta1, (double)step)));
  const float omb1 = 1.0f - beta1, omb2 = 1.0f - beta2;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    float gr = g[e] * gscale;
    float mm = m[e], vv = v[e];
    mm = mm + (gr - mm) * omb1;
    vv = vv + (gr * gr - vv) * omb2;
    m[e] = mm;
    v[e] = vv;
    p[e] = p[e] - lr_t * mm / (sqrtf(vv) + eps);
  }
}
__global__ void adam_advance_kernel(of_adam_state* st) { st->step += 1; }

// ---- opt-in L2 regulariser (model.py:7,13 / resnet/models.py:5,18,23: declared with l2(1e-4) on the encoder convs'
// kernels and biases, but train.py never adds model.losses - SURVEY F6; off by default) ----------------------------
// Keras l2(l): penalty = l * sum(w^2)  ->  grad += 2 l w, loss += l * sum(w^2)
__global__ void __launch_bounds__(256) l2_regularizer_kernel(const float* __restrict__ w, float* __restrict__ grad,
                                                             size_t n, float l2, float* __restrict__ loss) {
  float acc = 0.f;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    const float v = w[e];
    grad[e] += 2.0f * l2 * v;
    acc += v * v;
  }
  if (loss == nullptr) return;
  __shared__ float sm[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += sm[i];
    atomicAdd(loss, l2 * t);
  }
}

extern "C" {

int of_l2_regularizer(const float* w, float* grad, size_t n, float l2, float* loss, void* stream) {
  OF_REQUIRE(w && grad, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(n > 0, OF_ERR_SHAPE, "empty tensor");
  size_t blocks = (n + 255) / 256;
  if (blocks > (size_t)OF_NUM_SMS * 4) blocks = (size_t)OF_NUM_SMS * 4;
  l2_regularizer_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(w, grad, n, l2, loss);
  return of_launch_check("l2_regularizer_kernel");
}

int of_resize_bilinear(const float* in, float* out, int B, int H, int W, int C, int out_h, int out_w,
                       void* stream) {
  OF_REQUIRE(in && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(B > 0 && H > 0 && W > 0 && C > 0 && out_h > 0 && out_w > 0, OF_ERR_SHAPE, "bad dims");
  size_t total = (size_t)B * out_h * out_w * C;
  resize_kernel<<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(in, out, B, H, W, C, out_h, out_w, 1.0f);
  return of_launch_check("resize_kernel");
}

int of_upscale_flow_fwd(const float* flow, float* out, int B, int h, int w, void* stream) {
  OF_REQUIRE(flow && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(B > 0 && h > 0 && w > 0, OF_ERR_SHAPE, "bad dims");
  size_t total = (size_t)B * 4 * h * w * 2;
  resize_kernel<<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(flow, out, B, h, w, 2, 2 * h, 2 * w, 2.0f);
  return of_launch_check("upscale_flow_fwd");
}

int of_upscale_flow_bwd(const float* d_out, float* d_flow, int B, int h, int w, void* stream) {
  OF_REQUIRE(d_out && d_flow, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(B > 0 && h > 0 && w > 0, OF_ERR_SHAPE, "bad dims");
  size_t total = (size_t)B * h * w * 2;
  upscale2_bwd_kernel<<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(d_out, d_flow, B, h, w, 2, 2.0f);
  return of_launch_check("upscale_flow_bwd");
}

int of_maxpool2_fwd(const void* in, void* out, int B, int H, int W, int C, int dtype, void* stream) {
  OF_REQUIRE(in && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(dtype), OF_ERR_ARG, "bad dtype");
  OF_REQUIRE(B > 0 && H >= 2 && W >= 2 && C > 0 && C % 4 == 0, OF_ERR_SHAPE, "maxpool2 needs C %% 4 == 0, H,W >= 2");
  size_t total = (size_t)B * (H / 2) * (W / 2) * (C / 4);
  if (dtype == OF_F32)
    maxpool2_fwd_kernel<float><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>((const float*)in, (float*)out, B, H, W, C);
  else if (C % 8 == 0 && (((uintptr_t)in | (uintptr_t)out) & 15) == 0)
    maxpool2_fwd8_kernel<<<ew_grid(total / 2), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, B, H, W, C);
  else
    maxpool2_fwd_kernel<__nv_bfloat16><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)in, (__nv_bfloat16*)out, B, H, W, C);
  return of_launch_check("maxpool2_fwd");
}

int of_maxpool2_bwd(const void* in, const void* d_out, void* d_in, int B, int H, int W, int C, int dtype,
                    void* stream) {
  OF_REQUIRE(in && d_out && d_in, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(dtype), OF_ERR_ARG, "bad dtype");
  OF_REQUIRE(B > 0 && H >= 2 && W >= 2 && C > 0 && C % 4 == 0 && H % 2 == 0 && W % 2 == 0, OF_ERR_SHAPE,
             "maxpool2_bwd needs even H,W and C %% 4 == 0");
  size_t total = (size_t)B * (H / 2) * (W / 2) * (C / 4);
  if (dtype == OF_F32)
    maxpool2_bwd_kernel<float><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
        (const float*)in, (const float*)d_out, (float*)d_in, B, H, W, C);
  else
    maxpool2_bwd_kernel<__nv_bfloat16><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)in, (const __nv_bfloat16*)d_out, (__nv_bfloat16*)d_in, B, H, W, C);
  return of_launch_check("maxpool2_bwd");
}

int of_maxpool2_bwd_add_act(const void* in, const void* d_out, const void* add, void* dz, int B, int H, int W, int C,
                            int act, int dtype, void* stream) {
  OF_REQUIRE(in && d_out && add && dz, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(dtype) && H % 2 == 0 && W % 2 == 0 && C % 4 == 0, OF_ERR_SHAPE,
             "maxpool2 backward needs even H, W and C %% 4 == 0");
  OF_REQUIRE(act >= OF_ACT_NONE && act <= OF_ACT_LEAKY, OF_ERR_ARG, "bad activation %d", act);
  size_t total = (size_t)B * (H / 2) * (W / 2) * (C / 4);
  if (dtype == OF_F32)
    maxpool2_bwd_add_act_kernel<float><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
        (const float*)in, (const float*)d_out, (const float*)add, (float*)dz, B, H, W, C, act);
  else if (C % 8 == 0 && (((uintptr_t)in | (uintptr_t)d_out | (uintptr_t)add | (uintptr_t)dz) & 15) == 0)
    maxpool2_bwd_add_act8_kernel<<<ew_grid(total / 2), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)in, (const __nv_bfloat16*)d_out, (const __nv_bfloat16*)add, (__nv_bfloat16*)dz, B, H, W, C, act);
  else
    maxpool2_bwd_add_act_kernel<__nv_bfloat16><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)in, (const __nv_bfloat16*)d_out, (const __nv_bfloat16*)add, (__nv_bfloat16*)dz, B, H, W, C, act);
  return of_launch_check("maxpool2_bwd_add_act");
}

int of_split_bf16(const float* src, void* hi, void* lo, size_t npix, int C, int C_pad, int out_dtype, void* stream) {
  OF_REQUIRE(src && hi && lo, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(out_dtype) && npix > 0 && C > 0 && C_pad >= C, OF_ERR_ARG, "bad split arguments");
  const size_t total = npix * (size_t)C_pad;
  if (C_pad == C && total % 4 == 0 && (((uintptr_t)src | (uintptr_t)hi | (uintptr_t)lo) & 15) == 0) {
    if (out_dtype == OF_F32)
      split_bf16_flat4_kernel<float><<<ew_grid(total / 4), 256, 0, (cudaStream_t)stream>>>(src, (float*)hi, (float*)lo, total / 4);
    else
      split_bf16_flat4_kernel<__nv_bfloat16><<<ew_grid(total / 4), 256, 0, (cudaStream_t)stream>>>(
          src, (__nv_bfloat16*)hi, (__nv_bfloat16*)lo, total / 4);
    return of_launch_check("split_bf16_kernel");
  }
  if (out_dtype == OF_F32)
    split_bf16_kernel<float><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(src, (float*)hi, (float*)lo, npix, C, C_pad);
  else
    split_bf16_kernel<__nv_bfloat16><<<ew_grid(total), 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)hi,
                                                                                      (__nv_bfloat16*)lo, npix, C, C_pad);
  return of_launch_check("split_bf16_kernel");
}

int of_add_relu_fwd(const void* a, const void* b, void* out, size_t n, int dtype, void* stream) {
  OF_REQUIRE(a && b && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(dtype) && n % 4 == 0, OF_ERR_ARG, "bad dtype or n %% 4 != 0");
  if (dtype == OF_F32)
    add_relu_kernel<float><<<ew_grid(n / 4), 256, 0, (cudaStream_t)stream>>>((const float*)a, (const float*)b, (float*)out, n / 4);
  else if (n % 8 == 0 && (((uintptr_t)a | (uintptr_t)b | (uintptr_t)out) & 15) == 0)
    add_relu8_kernel<<<ew_grid(n / 8), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b,
                                                                       (__nv_bfloat16*)out, n / 8);
  else
    add_relu_kernel<__nv_bfloat16><<<ew_grid(n / 4), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (__nv_bfloat16*)out, n / 4);
  return of_launch_check("add_relu");
}

int of_act_bwd(const void* y, const void* dy, void* dz, size_t n, int act, int dtype, void* stream) {
  OF_REQUIRE(y && dy && dz, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(dtype) && n % 4 == 0, OF_ERR_ARG, "bad dtype or n %% 4 != 0");
  OF_REQUIRE(act >= OF_ACT_NONE && act <= OF_ACT_LEAKY, OF_ERR_ARG, "bad activation %d", act);
  if (dtype == OF_F32)
    act_bwd_kernel<float><<<ew_grid(n / 4), 256, 0, (cudaStream_t)stream>>>((const float*)y, (const float*)dy, (float*)dz, n / 4, act);
  else if (n % 8 == 0 && (((uintptr_t)y | (uintptr_t)dy | (uintptr_t)dz) & 15) == 0)
    act_bwd8_kernel<<<ew_grid(n / 8), 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)dy,
                                                                      (__nv_bfloat16*)dz, n / 8, act);
  else
    act_bwd_kernel<__nv_bfloat16><<<ew_grid(n / 4), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dz, n / 4, act);
  return of_launch_check("act_bwd");
}

int of_act_bwd_bits(const void* bits, const void* dy, void* dz, size_t npix, int C, int act, int dtype, void* stream) {
  OF_REQUIRE(bits && dy && dz, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(dtype) && C >= 8 && C <= 128 && C % 8 == 0 && npix > 0, OF_ERR_ARG, "bad dtype / C (C %% 8 == 0, <= 128)");
  OF_REQUIRE(act >= OF_ACT_NONE && act <= OF_ACT_LEAKY, OF_ERR_ARG, "bad activation %d", act);
  OF_REQUIRE((((uintptr_t)bits | (uintptr_t)dy | (uintptr_t)dz) & 15) == 0, OF_ERR_ARG, "16-byte alignment");
  const float slope = act == OF_ACT_RELU ? 0.f : (act == OF_ACT_LEAKY ? 0.3f : 1.f);
  const size_t n8 = npix * (size_t)(C / 8);
  if (dtype == OF_F32)
    act_bwd_bits_kernel<float><<<ew_grid(n8), 256, 0, (cudaStream_t)stream>>>((const uint32_t*)bits, (const float*)dy, (float*)dz, n8, C / 8, slope);
  else
    act_bwd_bits_kernel<__nv_bfloat16><<<ew_grid(n8), 256, 0, (cudaStream_t)stream>>>(
        (const uint32_t*)bits, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dz, n8, C / 8, slope);
  return of_launch_check("act_bwd_bits");
}

int of_relu_bwd(const void* out, const void* d_out, void* d, size_t n, int dtype, void* stream) {
  return of_act_bwd(out, d_out, d, n, OF_ACT_RELU, dtype, stream);
}

int of_adam_step(float* param, const float* grad, float* m, float* v, size_t n, int step, float lr, float beta1,
                 float beta2, float eps, float grad_scale, void* stream) {
  OF_REQUIRE(param && grad && m && v, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(step >= 1, OF_ERR_ARG, "Adam step count is 1-based");
  double lr_t = (double)lr * sqrt(1.0 - pow((double)beta2, step)) / (1.0 - pow((double)beta1, step));
  adam_kernel<<<ew_grid(n), 256, 0, (cudaStream_t)stream>>>(param, grad, m, v, n, (float)lr_t, 1.0f - beta1,
                                                            1.0f - beta2, eps, grad_scale);
  return of_launch_check("adam_kernel");
}

int of_adam_step_dev(float* param, const float* grad, float* m, float* v, size_t n, of_adam_state* state, float beta1,
                     float beta2, float eps, float grad_scale, void* stream) {
  OF_REQUIRE(param && grad && m && v && state, OF_ERR_ARG, "null pointer argument");
  adam_dev_kernel<<<ew_grid(n), 256, 0, (cudaStream_t)stream>>>(param, grad, m, v, n, state, beta1, beta2, eps, grad_scale);
  if (int r = of_launch_check("adam_kernel")) return r;
  adam_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(state);          // after every block has read the old count
  return of_launch_check("adam_advance_kernel");
}

}  // extern "C"
