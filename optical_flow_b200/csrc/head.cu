// Backward of the flow head - model.py:114 `Conv2D(2, 3, padding='same')` on the 32-channel output of the
// previous LeakyReLU conv - as ONE pass over the activation:
//     dw[kh,kw,ci,co] = sum_p x[p + (kh-1,kw-1), ci] * dy[p, co]            (weight gradient)
//     db[co]          = sum_p dy[p, co]
//     dz[q, ci]       = act'(x[q,ci]) * sum_{kh,kw,co} w[kh,kw,ci,co] * dy[q - (kh-1,kw-1), co]
// (x is the previous layer's post-activation output, so it also is the mask of the chained
// activation derivative: the result is that layer's dz directly.)
// Both sums run over the same 3x3 window of dy around q, so a warp walks a row segment with lane = ci,
// keeps the window (18 floats, warp-uniform loads) in registers, reads x[q, ci] once and issues 36 FMAs
// per pixel.  HBM-bound: x read once, dz written once, dy (2 channels) negligible.
// The generic kernels this replaces (split-K SIMT GEMM with Cout = 2, plus a separate activation pass)
// took 1.1 ms per step at C2; this takes ~0.1 ms.
#include "common.cuh"

#define HD_SEG 64
#define HD_WARPS 8
#define HD_CTAS_PER_SM 6
// Occupancy / prefetch-depth A/B (round 2, call 64; level 3 = 8 x 192 x 640 x 32 bf16, us): default (122 registers,
// 2 CTAs per SM, 4 pixels ahead) 195 | HD_MINB 3 (80 registers + spills) 266 | HD_MINB 4 477 | HD_PF 8 (162 registers) 254
// | HD_MINB 3 + HD_PF 8 406 | HD_PF 2 + HD_MINB 4 342.  The defaults stay.
#ifndef HD_MINB
#define HD_MINB 1          // __launch_bounds__ minimum CTAs per SM
#endif
#ifndef HD_PF
#define HD_PF 4            // pixels whose loads are issued up front per iteration
#endif

template <typename TX>
__global__ void __launch_bounds__(HD_WARPS * 32, HD_MINB)
flow_head_bwd_kernel(const TX* __restrict__ x, const float* __restrict__ dy, const float* __restrict__ w,
                     float* __restrict__ dw, float* __restrict__ db, TX* __restrict__ dz, int N, int H, int W, int Cin,
                     int act) {
  __shared__ float red[HD_WARPS][19][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool ch_ok = lane < Cin;
  const int ci = ch_ok ? lane : 0;
  // the two output channels ride in one packed fp32x2 FMA (sm_100 FFMA2): 18 instead of 36 FMAs per pixel
  float2 wr[9], acc[9];
  float bsum[2] = {0.f, 0.f};
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    wr[t] = ch_ok ? __ldg(reinterpret_cast<const float2*>(w) + (size_t)t * Cin + ci) : make_float2(0.f, 0.f);
    acc[t] = make_float2(0.f, 0.f);
  }
  const int segs_x = (W + HD_SEG - 1) / HD_SEG;
  const long long nseg = (long long)N * H * segs_x;
  const float2* dy2 = reinterpret_cast<const float2*>(dy);
  for (long long s = (long long)blockIdx.x * HD_WARPS + warp; s < nseg; s += (long long)gridDim.x * HD_WARPS) {
    const int sx = (int)(s % segs_x); const long long r = s / segs_x;
    const int y = (int)(r % H), n = (int)(r / H);
    const int x0 = sx * HD_SEG, x1 = min(W, x0 + HD_SEG);
    const size_t rowbase[3] = {((size_t)n * H + (y - 1)) * W, ((size_t)n * H + y) * W, ((size_t)n * H + (y + 1)) * W};
    const bool row_ok[3] = {y - 1 >= 0, true, y + 1 < H};
    float2 c0[3], c1[3], c2[3];                       // window columns q-1, q, q+1; rows y-1, y, y+1
#pragma unroll
    for (int rr = 0; rr < 3; ++rr) {
      c0[rr] = (row_ok[rr] && x0 - 1 >= 0) ? __ldg(dy2 + rowbase[rr] + x0 - 1) : make_float2(0.f, 0.f);
      c1[rr] = row_ok[rr] ? __ldg(dy2 + rowbase[rr] + x0) : make_float2(0.f, 0.f);
    }
    // 4 pixels per iteration: their 12 window columns and 4 activations are loaded up front (16 independent
    // loads in flight per warp), then consumed in order - the walk is latency-bound otherwise
    for (int q0 = x0; q0 < x1; q0 += HD_PF) {
      float2 cn[HD_PF][3];
      float xin[HD_PF];
#pragma unroll
      for (int k = 0; k < HD_PF; ++k) {
#pragma unroll
        for (int rr = 0; rr < 3; ++rr)
          cn[k][rr] = (row_ok[rr] && q0 + k + 1 < W) ? __ldg(dy2 + rowbase[rr] + q0 + k + 1) : make_float2(0.f, 0.f);
        xin[k] = (ch_ok && q0 + k < x1) ? ld_f<TX>(x + (rowbase[1] + q0 + k) * Cin + ci) : 0.f;
      }
#pragma unroll
      for (int k = 0; k < HD_PF; ++k) {
        const int q = q0 + k;
        if (q < x1) {                                   // warp-uniform
#pragma unroll
          for (int rr = 0; rr < 3; ++rr) c2[rr] = cn[k][rr];
          const float xv = xin[k];
          const float2 xv2 = make_float2(xv, xv);
          // three independent partial sums (one per kh): a single 9-deep FFMA2 chain per pixel was the walk's critical path
          float2 d2k[3] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
          // tap (kh,kw) pairs with dy[q - (kh-1, kw-1)] = window[row 2-kh][col 2-kw]
#pragma unroll
          for (int kh = 0; kh < 3; ++kh) {
            const float2 wc[3] = {c0[2 - kh], c1[2 - kh], c2[2 - kh]};
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
              const float2 g = wc[2 - kw];
              const int t = kh * 3 + kw;
              acc[t] = __ffma2_rn(xv2, g, acc[t]);
              d2k[kh] = __ffma2_rn(wr[t], g, d2k[kh]);
            }
          }
          const float d = ((d2k[0].x + d2k[1].x) + d2k[2].x) + ((d2k[0].y + d2k[1].y) + d2k[2].y);
          bsum[0] += c1[1].x; bsum[1] += c1[1].y;
          if (dz != nullptr && ch_ok) {
            float slope = 1.f;
            if (act == OF_ACT_RELU) slope = xv > 0.f ? 1.f : 0.f;
            else if (act == OF_ACT_LEAKY) slope = xv > 0.f ? 1.f : 0.3f;
            st_f<TX>(dz + (rowbase[1] + q) * Cin + ci, d * slope);
          }
#pragma unroll
          for (int rr = 0; rr < 3; ++rr) { c0[rr] = c1[rr]; c1[rr] = c2[rr]; }
        }
      }
    }
  }
  // CTA reduction, then one atomic per value per CTA
  float* mine = &red[warp][0][0];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    mine[(2 * t) * 32 + lane] = acc[t].x;
    mine[(2 * t + 1) * 32 + lane] = acc[t].y;
  }
  if (lane < 2) mine[18 * 32 + lane] = bsum[lane];
  __syncthreads();
  for (int i = threadIdx.x; i < 19 * 32; i += blockDim.x) {
    const int k = i >> 5, l = i & 31;
    float sum = 0.f;
#pragma unroll
    for (int wv = 0; wv < HD_WARPS; ++wv) sum += red[wv][k][l];
    if (k < 18) {
      if (l < Cin) red_add(dw + ((size_t)(k >> 1) * Cin + l) * 2 + (k & 1), sum);
    } else if (l < 2 && db != nullptr) {
      red_add(db + l, sum);
    }
  }
}

extern "C" {

int of_flow_head_bwd(const void* x, const float* dy, const float* w, float* dw, float* db, void* dz, int N, int H,
                     int W, int Cin, int act, int dtype, void* stream) {
  OF_REQUIRE(x && dy && w && dw, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(dtype), OF_ERR_ARG, "bad dtype %d", dtype);
  OF_REQUIRE(N > 0 && H > 0 && W > 0 && Cin > 0 && Cin <= 32, OF_ERR_SHAPE, "flow head backward needs 1 <= Cin <= 32 (got %d)", Cin);
  OF_REQUIRE(act == OF_ACT_NONE || act == OF_ACT_RELU || act == OF_ACT_LEAKY, OF_ERR_ARG, "bad activation %d", act);
  cudaStream_t st = (cudaStream_t)stream;
  OF_CUDA(cudaMemsetAsync(dw, 0, (size_t)9 * Cin * 2 * sizeof(float), st));
  if (db) OF_CUDA(cudaMemsetAsync(db, 0, 2 * sizeof(float), st));
  const long long nseg = (long long)N * H * ((W + HD_SEG - 1) / HD_SEG);
  long long blocks = (nseg + HD_WARPS - 1) / HD_WARPS;
  if (blocks > OF_NUM_SMS * HD_CTAS_PER_SM) blocks = OF_NUM_SMS * HD_CTAS_PER_SM;
  if (dtype == OF_F32)
    flow_head_bwd_kernel<float><<<(int)blocks, HD_WARPS * 32, 0, st>>>((const float*)x, dy, w, dw, db, (float*)dz, N, H, W, Cin, act);
  else
    flow_head_bwd_kernel<__nv_bfloat16><<<(int)blocks, HD_WARPS * 32, 0, st>>>((const __nv_bfloat16*)x, dy, w, dw, db,
                                                                               (__nv_bfloat16*)dz, N, H, W, Cin, act);
  return of_launch_check("flow_head_bwd_kernel");
}

}  // extern "C"
