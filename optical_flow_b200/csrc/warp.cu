// K1-K4: bilinear sampler / feature warp / UnFlow image warp, forward + backward.
//
// Replaces transformations.py:70-129 (bilinear_interpolation + gather_nd),
// model.py:55-73 (warp_features) and old/image_warp.py:7-78 (image_warp).
// One kernel family, three coordinate sources:
//   SRC_PTS    absolute (x, y) sampling points              (transformations.py)
//   SRC_REF_IJ [row, col] + flow read as (x, y)              (model.py:65-71)
//   SRC_XY     [col, row] + flow                             (corrected order)
//   SRC_UNFLOW int(floor(flow)) + pos, frac weights          (old/image_warp.py)
// Index arithmetic follows the reference's op order literally so the int32 tap
// indices are bit-exact; products/sums use __fmul_rn/__fadd_rn (no FMA
// contraction) in the reference's accumulate order so fp32 outputs are too.
#include <stdlib.h>
#include "common.cuh"

enum { SRC_PTS = 0, SRC_REF_IJ = 1, SRC_XY = 2, SRC_UNFLOW = 3 };

struct Taps {
  int x0, x1, y0, y1;      // clipped tap coordinates
  float ax, bx, ay, by;    // weights of x0, x1, y0, y1
};

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }
// x0 + 1 without signed overflow for saturated casts
__device__ __forceinline__ int inc_sat(int v) { return v == 0x7fffffff ? v : v + 1; }

template <int SRC>
__device__ __forceinline__ Taps make_taps(float c0, float c1, int i, int j, int H, int W) {
  Taps t;
  if (SRC == SRC_UNFLOW) {
    // old/image_warp.py:28-29,48-56
    float f0 = floorf(c0), f1 = floorf(c1);
    int fx = __float2int_rd(c0), fy = __float2int_rd(c1);
    float xw = __fsub_rn(c0, f0), yw = __fsub_rn(c1, f1);
    // pos + floor(flow) in int32 (saturating inputs keep it inside int range for sane flows)
    long long x0l = (long long)j + fx, y0l = (long long)i + fy;
    int x0 = (int)max(min(x0l, 0x7ffffffell), -0x7fffffffll);
    int y0 = (int)max(min(y0l, 0x7ffffffell), -0x7fffffffll);
    t.x0 = clampi(x0, 0, W - 1);
    t.x1 = clampi(x0 + 1, 0, W - 1);
    t.y0 = clampi(y0, 0, H - 1);
    t.y1 = clampi(y0 + 1, 0, H - 1);
    t.ax = __fsub_rn(1.0f, xw);
    t.bx = xw;
    t.ay = __fsub_rn(1.0f, yw);
    t.by = yw;
  } else {
    float x, y;
    if (SRC == SRC_PTS) {
      x = c0; y = c1;                                   // transformations.py:93-94
    } else if (SRC == SRC_REF_IJ) {
      x = __fadd_rn((float)i, c0);                      // model.py:65-71: [row, col] + flow
      y = __fadd_rn((float)j, c1);
    } else {
      x = __fadd_rn((float)j, c0);
      y = __fadd_rn((float)i, c1);
    }
    int x0 = __float2int_rd(x), y0 = __float2int_rd(y); // :98,:100 floor + cast
    int x1 = inc_sat(x0), y1 = inc_sat(y0);             // :99,:101
    t.x0 = clampi(x0, 0, W - 1);                        // :104-107
    t.x1 = clampi(x1, 0, W - 1);
    t.y0 = clampi(y0, 0, H - 1);
    t.y1 = clampi(y1, 0, H - 1);
    float a = __fsub_rn((float)t.x1, x);                // :116,:120 weights from CLIPPED x1
    float b = __fsub_rn((float)t.y1, y);                // :117,:121
    t.ax = a;
    t.bx = __fsub_rn(1.0f, a);
    t.ay = b;
    t.by = __fsub_rn(1.0f, b);
  }
  return t;
}

__device__ __forceinline__ float blend(float w00, float w01, float w10, float w11,
                                       float v00, float v01, float v10, float v11) {
  // accumulate_n / add_n order: x0y0, x0y1, x1y0, x1y1 (transformations.py:128-129)
  float r = __fmul_rn(w00, v00);
  r = __fadd_rn(r, __fmul_rn(w01, v01));
  r = __fadd_rn(r, __fmul_rn(w10, v10));
  r = __fadd_rn(r, __fmul_rn(w11, v11));
  return r;
}

#define TILE 16  // 16x16 output pixels per CTA pass: keeps the (possibly transposed) source footprint compact

__device__ __forceinline__ bool tile_pixel(long long tile, int p, int H, int W, int tiles_x, int tiles_y,
                                           int& b, int& i, int& j) {
  int tx = (int)(tile % tiles_x);
  long long r = tile / tiles_x;
  int ty = (int)(r % tiles_y);
  b = (int)(r / tiles_y);
  i = ty * TILE + p / TILE;
  j = tx * TILE + p % TILE;
  return i < H && j < W;
}
// column-major walk of a TILE x TJ tile (RUN kernels): pixel p -> (row p % TILE, column p / TILE)
template <int TJ>
__device__ __forceinline__ bool tile_pixel_run(long long tile, int p, int H, int W, int tiles_x, int tiles_y,
                                               int& b, int& i, int& j) {
  int tx = (int)(tile % tiles_x);
  long long r = tile / tiles_x;
  int ty = (int)(r % tiles_y);
  b = (int)(r / tiles_y);
  i = ty * TILE + p % TILE;
  j = tx * TJ + p / TILE;
  return i < H && j < W;
}

// ------------------------------------------------------------------ forward --
template <typename T, int VEC, int SRC>
__global__ void __launch_bounds__(256) sample_fwd_kernel(const T* __restrict__ in,
                                                         const float* __restrict__ coords,
                                                         T* __restrict__ out, int32_t* __restrict__ idx,
                                                         int B, int H, int W, int C) {
  const int CV = C / VEC;
  const int tiles_x = (W + TILE - 1) / TILE, tiles_y = (H + TILE - 1) / TILE;
  const long long ntiles = (long long)B * tiles_x * tiles_y;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    for (int it = threadIdx.x; it < TILE * TILE * CV; it += blockDim.x) {
      int p = it / CV, v = it - p * CV;
      int b, i, j;
      if (!tile_pixel(tile, p, H, W, tiles_x, tiles_y, b, i, j)) continue;
      size_t pix = ((size_t)b * H + i) * W + j;
      float2 c = __ldg(reinterpret_cast<const float2*>(coords) + pix);
      Taps t = make_taps<SRC>(c.x, c.y, i, j, H, W);
      if (idx != nullptr && v == 0) {
        reinterpret_cast<int4*>(idx)[pix] = make_int4(t.x0, t.x1, t.y0, t.y1);
      }
      float w00 = __fmul_rn(t.ax, t.ay), w01 = __fmul_rn(t.ax, t.by);
      float w10 = __fmul_rn(t.bx, t.ay), w11 = __fmul_rn(t.bx, t.by);
      const T* base = in + (size_t)b * H * W * C;
      const T* p00 = base + ((size_t)t.y0 * W + t.x0) * C + v * VEC;
      const T* p01 = base + ((size_t)t.y1 * W + t.x0) * C + v * VEC;
      const T* p10 = base + ((size_t)t.y0 * W + t.x1) * C + v * VEC;
      const T* p11 = base + ((size_t)t.y1 * W + t.x1) * C + v * VEC;
      T* po = out + pix * C + v * VEC;
      if constexpr (VEC == 8) {
        // bf16 only: 16-byte taps - half the threads (and half the redundant tap arithmetic) per pixel
        float a[8], bq[8], cq[8], d[8], r[8];
        ld8_bf16(reinterpret_cast<const __nv_bfloat16*>(p00), a);
        ld8_bf16(reinterpret_cast<const __nv_bfloat16*>(p01), bq);
        ld8_bf16(reinterpret_cast<const __nv_bfloat16*>(p10), cq);
        ld8_bf16(reinterpret_cast<const __nv_bfloat16*>(p11), d);
#pragma unroll
        for (int k = 0; k < 8; ++k) r[k] = blend(w00, w01, w10, w11, a[k], bq[k], cq[k], d[k]);
        st8_bf16(reinterpret_cast<__nv_bfloat16*>(po), r);
      } else if (VEC == 4) {
        float4 a = ld4<T>(p00), bq = ld4<T>(p01), cq = ld4<T>(p10), d = ld4<T>(p11);
        float4 r;
        r.x = blend(w00, w01, w10, w11, a.x, bq.x, cq.x, d.x);
        r.y = blend(w00, w01, w10, w11, a.y, bq.y, cq.y, d.y);
        r.z = blend(w00, w01, w10, w11, a.z, bq.z, cq.z, d.z);
        r.w = blend(w00, w01, w10, w11, a.w, bq.w, cq.w, d.w);
        st4<T>(po, r);
      } else {
        st_f<T>(po, blend(w00, w01, w10, w11, ld_f<T>(p00), ld_f<T>(p01), ld_f<T>(p10), ld_f<T>(p11)));
      }
    }
  }
}

// ----------------------------------------------------------------- backward --
// Warp-shuffle-combined atomics: lanes of a warp that scatter to the same
// address (adjacent pixels share taps under smooth flow) are merged with a
// rank-ordered shuffle tree; only the first lane of each group issues the
// reduction to global memory.
__device__ __forceinline__ float combine_peers(unsigned peers, int rank, int lane, float val) {
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    unsigned src = __fns(peers, 0, rank + off + 1);
    bool ok = (src != 0xffffffffu);
    float o = __shfl_sync(0xffffffffu, val, ok ? (int)src : lane);
    if (ok && (rank & (2 * off - 1)) == 0) val += o;
  }
  return val;
}

// G = dtype of the incoming gradient d_out (float, or bf16 on the tensor-core conv path)
// RUN (VEC == 4, C / 4 == TPP): a thread group keeps row i fixed and walks j through the tile; each of its four
// taps sits in a register slot that accumulates while consecutive pixels hit the SAME destination texel and is
// flushed (one red.global.add.v4) when the address changes.  Under model.py:65-71's row/column swap every output
// column j >= H - 1 - flow reads the clamped last row, so ~70 % of a 192 x 640 level's pixels scatter into one row:
// the run merge turns 4 atomics per pixel there into 4 per 16-pixel run.
// RUN tiles are TILE rows x RUN_TJ columns: the longer the walk along j, the longer the runs that merge (in the clamped
// region one flush per tile instead of one per 16 pixels: 4x fewer atomics on the hot row at RUN_TJ = 64)
#define RUN_TJ 64
#ifndef WARP_BWD_MINB
#define WARP_BWD_MINB 2          // resident CTAs per SM asked of the compiler: 2 = 128 registers (1 lets it take 164 -> ONE CTA per SM: 427 -> 530 us;
                                 // 3 / 4 spill: 629 / 729 us at level 3 bf16)
#endif
template <typename T, int VEC, int TPP, int SRC, typename G = float, bool RUN = false>
__global__ void __launch_bounds__(256, WARP_BWD_MINB) sample_bwd_kernel(const T* __restrict__ in,
                                                         const float* __restrict__ coords,
                                                         const G* __restrict__ d_out,
                                                         float* __restrict__ d_in,
                                                         float* __restrict__ d_coords,
                                                         int B, int H, int W, int C) {
  const int CV = C / VEC;
  constexpr int TJ = RUN ? RUN_TJ : TILE;
  const int tiles_x = (W + TJ - 1) / TJ, tiles_y = (H + TILE - 1) / TILE;
  const long long ntiles = (long long)B * tiles_x * tiles_y;
  const int lane = threadIdx.x & 31;
  const int sub = threadIdx.x % TPP;       // lane within the pixel's thread group
  const int grp = threadIdx.x / TPP;       // pixel slot within the pass
  constexpr int PIX_PER_PASS = 256 / TPP;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    float4 racc[4][VEC == 8 ? 2 : 1];                 // RUN: pending contribution per tap slot
    size_t raddr[4];
    bool rvalid[4] = {false, false, false, false};
    for (int p0 = 0; p0 < TILE * TJ; p0 += PIX_PER_PASS) {
      int p = p0 + grp;
      int b, i, j;
      // RUN walks the tile column by column: pixel p -> (row p % TILE, column p / TILE)
      bool valid = RUN ? tile_pixel_run<TJ>(tile, p, H, W, tiles_x, tiles_y, b, i, j)
                       : tile_pixel(tile, p, H, W, tiles_x, tiles_y, b, i, j);
      size_t pix = valid ? ((size_t)b * H + i) * W + j : 0;
      // (loading the coordinates one pass ahead was tried in round 2: 427 -> 530 us at level 3 bf16, more live registers)
      float2 c = valid ? __ldg(reinterpret_cast<const float2*>(coords) + pix) : make_float2(0.f, 0.f);
      Taps t = make_taps<SRC>(c.x, c.y, valid ? i : 0, valid ? j : 0, H, W);
      float w00 = t.ax * t.ay, w01 = t.ax * t.by, w10 = t.bx * t.ay, w11 = t.bx * t.by;
      size_t boff = (size_t)(valid ? b : 0) * H * W;
      size_t o00 = (boff + (size_t)t.y0 * W + t.x0) * C, o01 = (boff + (size_t)t.y1 * W + t.x0) * C;
      size_t o10 = (boff + (size_t)t.y0 * W + t.x1) * C, o11 = (boff + (size_t)t.y1 * W + t.x1) * C;
      float d0 = 0.f, d1 = 0.f;  // d/d coord0, d/d coord1

      if (TPP == 1 && VEC == 1) {
        // small-C path (images): one thread per pixel, warp-combined scatter
        unsigned pe[4] = {0, 0, 0, 0};
        int rk[4] = {0, 0, 0, 0};
        bool any_dup = false;
        if (d_in != nullptr) {
          size_t keys[4] = {o00, o01, o10, o11};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            unsigned long long key = valid ? (unsigned long long)keys[k] : (0x8000000000000000ull | lane);
            pe[k] = __match_any_sync(0xffffffffu, key);
            rk[k] = __popc(pe[k] & ((1u << lane) - 1u));
          }
          any_dup = __any_sync(0xffffffffu, (__popc(pe[0]) | __popc(pe[1]) | __popc(pe[2]) | __popc(pe[3])) > 1);
        }
        for (int ch = 0; ch < C; ++ch) {
          float g = valid ? ld_f<G>(d_out + pix * C + ch) : 0.f;
          float v00 = ld_f<T>(in + o00 + ch), v01 = ld_f<T>(in + o01 + ch);
          float v10 = ld_f<T>(in + o10 + ch), v11 = ld_f<T>(in + o11 + ch);
          d0 += g * (t.ay * (v10 - v00) + t.by * (v11 - v01));
          d1 += g * (t.ax * (v01 - v00) + t.bx * (v11 - v10));
          if (d_in != nullptr) {
            float vals[4] = {w00 * g, w01 * g, w10 * g, w11 * g};
            size_t offs[4] = {o00, o01, o10, o11};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              float s = vals[k];
              if (any_dup) s = combine_peers(pe[k], rk[k], lane, s);
              if (valid && rk[k] == 0) red_add(d_in + offs[k] + ch, s);
            }
          }
        }
      } else {
        for (int v = sub; v < CV; v += TPP) {
          if (!valid) continue;
          int ch = v * VEC;
          if constexpr (VEC == 8) {
            // bf16 features and bf16 incoming gradient, 16-byte accesses: 8 threads per 64-channel pixel
            float g[8], v00[8], v01[8], v10[8], v11[8];
            ld8_bf16(reinterpret_cast<const __nv_bfloat16*>(d_out) + pix * C + ch, g);
            // both taps of a column on ONE source row (y0 == y1: the clamped region, ~70 % of a 192 x 640 level under
            // model.py:65-71's row / column swap): two loads and two scatter slots instead of four
            const bool same_row = t.y0 == t.y1;
            ld8_bf16(reinterpret_cast<const __nv_bfloat16*>(in) + o00 + ch, v00);
            ld8_bf16(reinterpret_cast<const __nv_bfloat16*>(in) + o10 + ch, v10);
            if (same_row) {
#pragma unroll
              for (int k = 0; k < 8; ++k) { v01[k] = v00[k]; v11[k] = v10[k]; }
            } else {
              ld8_bf16(reinterpret_cast<const __nv_bfloat16*>(in) + o01 + ch, v01);
              ld8_bf16(reinterpret_cast<const __nv_bfloat16*>(in) + o11 + ch, v11);
            }
            // same summation order as the VEC == 4 path over the two 4-channel halves
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int k = 4 * h;
              d0 += g[k] * (t.ay * (v10[k] - v00[k]) + t.by * (v11[k] - v01[k])) +
                    g[k + 1] * (t.ay * (v10[k + 1] - v00[k + 1]) + t.by * (v11[k + 1] - v01[k + 1])) +
                    g[k + 2] * (t.ay * (v10[k + 2] - v00[k + 2]) + t.by * (v11[k + 2] - v01[k + 2])) +
                    g[k + 3] * (t.ay * (v10[k + 3] - v00[k + 3]) + t.by * (v11[k + 3] - v01[k + 3]));
              d1 += g[k] * (t.ax * (v01[k] - v00[k]) + t.bx * (v11[k] - v10[k])) +
                    g[k + 1] * (t.ax * (v01[k + 1] - v00[k + 1]) + t.bx * (v11[k + 1] - v10[k + 1])) +
                    g[k + 2] * (t.ax * (v01[k + 2] - v00[k + 2]) + t.bx * (v11[k + 2] - v10[k + 2])) +
                    g[k + 3] * (t.ax * (v01[k + 3] - v00[k + 3]) + t.bx * (v11[k + 3] - v10[k + 3]));
            }
            if (d_in != nullptr) {
              const size_t offs[4] = {o00 + ch, o01 + ch, o10 + ch, o11 + ch};
              const float ws[4] = {same_row ? w00 + w01 : w00, w01, same_row ? w10 + w11 : w10, w11};
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                if (same_row && (k & 1)) continue;           // folded into slot k - 1 (same address)
                const float4 lo = make_float4(ws[k] * g[0], ws[k] * g[1], ws[k] * g[2], ws[k] * g[3]);
                const float4 hi = make_float4(ws[k] * g[4], ws[k] * g[5], ws[k] * g[6], ws[k] * g[7]);
                if (RUN) {
                  if (rvalid[k] && raddr[k] == offs[k]) {
                    racc[k][0].x += lo.x; racc[k][0].y += lo.y; racc[k][0].z += lo.z; racc[k][0].w += lo.w;
                    racc[k][1].x += hi.x; racc[k][1].y += hi.y; racc[k][1].z += hi.z; racc[k][1].w += hi.w;
                  } else {
                    if (rvalid[k]) { red_add_v4(d_in + raddr[k], racc[k][0]); red_add_v4(d_in + raddr[k] + 4, racc[k][1]); }
                    racc[k][0] = lo; racc[k][1] = hi; raddr[k] = offs[k]; rvalid[k] = true;
                  }
                } else {
                  red_add_v4(d_in + offs[k], lo);
                  red_add_v4(d_in + offs[k] + 4, hi);
                }
              }
            }
          } else if (VEC == 4) {
            float4 g = ld4<G>(d_out + pix * C + ch);
            float4 v00 = ld4<T>(in + o00 + ch), v01 = ld4<T>(in + o01 + ch);
            float4 v10 = ld4<T>(in + o10 + ch), v11 = ld4<T>(in + o11 + ch);
            d0 += g.x * (t.ay * (v10.x - v00.x) + t.by * (v11.x - v01.x)) +
                  g.y * (t.ay * (v10.y - v00.y) + t.by * (v11.y - v01.y)) +
                  g.z * (t.ay * (v10.z - v00.z) + t.by * (v11.z - v01.z)) +
                  g.w * (t.ay * (v10.w - v00.w) + t.by * (v11.w - v01.w));
            d1 += g.x * (t.ax * (v01.x - v00.x) + t.bx * (v11.x - v10.x)) +
                  g.y * (t.ax * (v01.y - v00.y) + t.bx * (v11.y - v10.y)) +
                  g.z * (t.ax * (v01.z - v00.z) + t.bx * (v11.z - v10.z)) +
                  g.w * (t.ax * (v01.w - v00.w) + t.bx * (v11.w - v10.w));
            if (d_in != nullptr) {
              if (RUN) {
                const size_t offs[4] = {o00 + ch, o01 + ch, o10 + ch, o11 + ch};
                const float ws[4] = {w00, w01, w10, w11};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const float4 val = make_float4(ws[k] * g.x, ws[k] * g.y, ws[k] * g.z, ws[k] * g.w);
                  if (rvalid[k] && raddr[k] == offs[k]) {
                    racc[k][0].x += val.x; racc[k][0].y += val.y; racc[k][0].z += val.z; racc[k][0].w += val.w;
                  } else {
                    if (rvalid[k]) red_add_v4(d_in + raddr[k], racc[k][0]);
                    racc[k][0] = val; raddr[k] = offs[k]; rvalid[k] = true;
                  }
                }
              } else {
                red_add_v4(d_in + o00 + ch, make_float4(w00 * g.x, w00 * g.y, w00 * g.z, w00 * g.w));
                red_add_v4(d_in + o01 + ch, make_float4(w01 * g.x, w01 * g.y, w01 * g.z, w01 * g.w));
                red_add_v4(d_in + o10 + ch, make_float4(w10 * g.x, w10 * g.y, w10 * g.z, w10 * g.w));
                red_add_v4(d_in + o11 + ch, make_float4(w11 * g.x, w11 * g.y, w11 * g.z, w11 * g.w));
              }
            }
          } else {
            float g = ld_f<G>(d_out + pix * C + ch);
            float v00 = ld_f<T>(in + o00 + ch), v01 = ld_f<T>(in + o01 + ch);
            float v10 = ld_f<T>(in + o10 + ch), v11 = ld_f<T>(in + o11 + ch);
            d0 += g * (t.ay * (v10 - v00) + t.by * (v11 - v01));
            d1 += g * (t.ax * (v01 - v00) + t.bx * (v11 - v10));
            if (d_in != nullptr) {
              red_add(d_in + o00 + ch, w00 * g);
              red_add(d_in + o01 + ch, w01 * g);
              red_add(d_in + o10 + ch, w10 * g);
              red_add(d_in + o11 + ch, w11 * g);
            }
          }
        }
#pragma unroll
        for (int o = TPP / 2; o > 0; o >>= 1) {
          d0 += __shfl_xor_sync(0xffffffffu, d0, o);
          d1 += __shfl_xor_sync(0xffffffffu, d1, o);
        }
      }
      if (valid && sub == 0 && d_coords != nullptr) {
        reinterpret_cast<float2*>(d_coords)[pix] = make_float2(d0, d1);
      }
    }
    if (RUN) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (rvalid[k]) {
          red_add_v4(d_in + raddr[k], racc[k][0]);
          if constexpr (VEC == 8) red_add_v4(d_in + raddr[k] + 4, racc[k][1]);
        }
    }
  }
}

// ------------------------------------------------------------------- host ----
static int grid_for(long long ntiles) {
  long long g = OF_NUM_SMS * 8;  // 8 resident 256-thread CTAs per SM
  if (ntiles < g) g = ntiles;
  return (int)(g < 1 ? 1 : g);
}

template <typename T, int SRC>
static int launch_fwd(const void* in, const float* coords, void* out, int32_t* idx, int B, int H, int W,
                      int C, cudaStream_t st) {
  long long ntiles = (long long)B * ((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
  int grid = grid_for(ntiles);
  if (sizeof(T) == 2 && C % 8 == 0 && (((uintptr_t)in | (uintptr_t)out) & 15) == 0 && !of_tune_env("OF_WARP_NO_VEC8"))
    sample_fwd_kernel<T, 8, SRC><<<grid, 256, 0, st>>>((const T*)in, coords, (T*)out, idx, B, H, W, C);
  else if (C % 4 == 0)
    sample_fwd_kernel<T, 4, SRC><<<grid, 256, 0, st>>>((const T*)in, coords, (T*)out, idx, B, H, W, C);
  else
    sample_fwd_kernel<T, 1, SRC><<<grid, 256, 0, st>>>((const T*)in, coords, (T*)out, idx, B, H, W, C);
  return of_launch_check("sample_fwd_kernel");
}

template <typename T, int SRC, typename G = float>
static int launch_bwd(const void* in, const float* coords, const void* d_out, float* d_in, float* d_coords,
                      int B, int H, int W, int C, cudaStream_t st) {
  long long ntiles = (long long)B * ((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
  int grid = grid_for(ntiles);
  const int grid_run = grid_for((long long)B * ((W + RUN_TJ - 1) / RUN_TJ) * ((H + TILE - 1) / TILE));
  if (d_in != nullptr) {
    cudaError_t e = cudaMemsetAsync(d_in, 0, (size_t)B * H * W * C * sizeof(float), st);
    if (e != cudaSuccess) { of_set_error("memset d_in: %s", cudaGetErrorString(e)); return (int)e; }
  }
#define L(VEC, TPP)                                                                          \
  sample_bwd_kernel<T, VEC, TPP, SRC, G><<<grid, 256, 0, st>>>((const T*)in, coords, (const G*)d_out, d_in, \
                                                               d_coords, B, H, W, C)
#define LRUN(TPP)                                                                                    \
  sample_bwd_kernel<T, 4, TPP, SRC, G, true><<<grid_run, 256, 0, st>>>((const T*)in, coords, (const G*)d_out, d_in, \
                                                                       d_coords, B, H, W, C)
  const bool run = SRC == SRC_REF_IJ && d_in != nullptr && !of_tune_env("OF_WARP_NO_RUN");
  // bf16 features + bf16 gradient, C = 64 / 128: 16-byte accesses, 8 / 16 threads per pixel
  if (sizeof(T) == 2 && sizeof(G) == 2 && run && (C == 64 || C == 128) && !of_tune_env("OF_WARP_NO_VEC8") &&
      ((((uintptr_t)in | (uintptr_t)d_out) & 15) == 0)) {
    if (C == 64)
      sample_bwd_kernel<T, 8, 8, SRC, G, true><<<grid_run, 256, 0, st>>>((const T*)in, coords, (const G*)d_out, d_in, d_coords, B, H, W, C);
    else
      sample_bwd_kernel<T, 8, 16, SRC, G, true><<<grid_run, 256, 0, st>>>((const T*)in, coords, (const G*)d_out, d_in, d_coords, B, H, W, C);
  } else
  if (C % 4 == 0) {
    int cv = C / 4;
    if (run && cv == 32) LRUN(32);
    else if (run && cv == 16) LRUN(16);
    else if (cv >= 32) L(4, 32);
    else if (cv >= 16) L(4, 16);
    else if (cv >= 8) L(4, 8);
    else if (cv >= 4) L(4, 4);
    else if (cv >= 2) L(4, 2);
    else L(4, 1);
  } else if (C <= 8) {
    L(1, 1);
  } else {
    L(1, 8);
  }
#undef L
#undef LRUN
  return of_launch_check("sample_bwd_kernel");
}

static int check_dims(int B, int H, int W, int C) {
  OF_REQUIRE(B > 0 && H > 0 && W > 0 && C > 0, OF_ERR_SHAPE, "bad dims B=%d H=%d W=%d C=%d", B, H, W, C);
  OF_REQUIRE((long long)B * H * W * C < (1ll << 40), OF_ERR_SHAPE, "tensor too large");
  return 0;
}

template <int SRC>
static int fwd_dispatch(const void* in, const float* coords, void* out, int32_t* idx, int B, int H, int W,
                        int C, int dtype, void* stream) {
  OF_REQUIRE(in && coords && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(dtype), OF_ERR_ARG, "bad dtype %d", dtype);
  if (int r = check_dims(B, H, W, C)) return r;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == OF_F32) return launch_fwd<float, SRC>(in, coords, out, idx, B, H, W, C, st);
  return launch_fwd<__nv_bfloat16, SRC>(in, coords, out, idx, B, H, W, C, st);
}

template <int SRC>
static int bwd_dispatch(const void* in, const float* coords, const void* d_out, float* d_in, float* d_coords,
                        int B, int H, int W, int C, int dtype, void* stream, int g_dtype = OF_F32) {
  OF_REQUIRE(in && coords && d_out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(d_in || d_coords, OF_ERR_ARG, "both gradient outputs are null");
  OF_REQUIRE(of_valid_dtype(dtype), OF_ERR_ARG, "bad dtype %d", dtype);
  if (int r = check_dims(B, H, W, C)) return r;
  cudaStream_t st = (cudaStream_t)stream;
  if (g_dtype == OF_BF16) {
    OF_REQUIRE(dtype == OF_BF16, OF_ERR_UNSUPPORTED, "a bf16 incoming gradient needs bf16 features");
    return launch_bwd<__nv_bfloat16, SRC, __nv_bfloat16>(in, coords, d_out, d_in, d_coords, B, H, W, C, st);
  }
  if (dtype == OF_F32) return launch_bwd<float, SRC>(in, coords, d_out, d_in, d_coords, B, H, W, C, st);
  return launch_bwd<__nv_bfloat16, SRC>(in, coords, d_out, d_in, d_coords, B, H, W, C, st);
}

extern "C" {

int of_bilinear_sample_fwd(const void* in, const float* pts, void* out, int32_t* idx, int B, int H, int W,
                           int C, int dtype, void* stream) {
  return fwd_dispatch<SRC_PTS>(in, pts, out, idx, B, H, W, C, dtype, stream);
}

int of_bilinear_sample_bwd(const void* in, const float* pts, const float* d_out, float* d_in, float* d_pts,
                           int B, int H, int W, int C, int dtype, void* stream) {
  return bwd_dispatch<SRC_PTS>(in, pts, d_out, d_in, d_pts, B, H, W, C, dtype, stream);
}

int of_warp_features_fwd(const void* feats, const float* flow, void* out, int32_t* idx, int B, int H, int W,
                         int C, int mode, int dtype, void* stream) {
  OF_REQUIRE(mode == OF_WARP_REF_IJ || mode == OF_WARP_XY, OF_ERR_ARG, "bad warp mode %d", mode);
  if (mode == OF_WARP_REF_IJ) return fwd_dispatch<SRC_REF_IJ>(feats, flow, out, idx, B, H, W, C, dtype, stream);
  return fwd_dispatch<SRC_XY>(feats, flow, out, idx, B, H, W, C, dtype, stream);
}

int of_warp_features_bwd(const void* feats, const float* flow, const float* d_out, float* d_feats,
                         float* d_flow, int B, int H, int W, int C, int mode, int dtype, void* stream) {
  OF_REQUIRE(mode == OF_WARP_REF_IJ || mode == OF_WARP_XY, OF_ERR_ARG, "bad warp mode %d", mode);
  if (mode == OF_WARP_REF_IJ)
    return bwd_dispatch<SRC_REF_IJ>(feats, flow, d_out, d_feats, d_flow, B, H, W, C, dtype, stream);
  return bwd_dispatch<SRC_XY>(feats, flow, d_out, d_feats, d_flow, B, H, W, C, dtype, stream);
}

int of_warp_features_bwd_g(const void* feats, const float* flow, const void* d_out, int d_out_dtype, float* d_feats,
                           float* d_flow, int B, int H, int W, int C, int mode, int dtype, void* stream) {
  OF_REQUIRE(mode == OF_WARP_REF_IJ || mode == OF_WARP_XY, OF_ERR_ARG, "bad warp mode %d", mode);
  OF_REQUIRE(of_valid_dtype(d_out_dtype), OF_ERR_ARG, "bad d_out dtype %d", d_out_dtype);
  if (mode == OF_WARP_REF_IJ)
    return bwd_dispatch<SRC_REF_IJ>(feats, flow, d_out, d_feats, d_flow, B, H, W, C, dtype, stream, d_out_dtype);
  return bwd_dispatch<SRC_XY>(feats, flow, d_out, d_feats, d_flow, B, H, W, C, dtype, stream, d_out_dtype);
}

int of_image_warp_fwd(const float* im, const float* flow, float* out, int32_t* idx, int B, int H, int W,
                      int C, void* stream) {
  return fwd_dispatch<SRC_UNFLOW>(im, flow, out, idx, B, H, W, C, OF_F32, stream);
}

int of_image_warp_bwd(const float* im, const float* flow, const float* d_out, float* d_im, float* d_flow,
                      int B, int H, int W, int C, void* stream) {
  return bwd_dispatch<SRC_UNFLOW>(im, flow, d_out, d_im, d_flow, B, H, W, C, OF_F32, stream);
}

}  // extern "C"
