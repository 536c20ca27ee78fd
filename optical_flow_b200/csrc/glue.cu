// Small data-movement / parameter kernels that sit between the heavy ops:
//   of_split_image_pair  model.py:122-123,131-132: (B,H,W,6) -> (2B,H,W,3) so the shared
//                        encoder runs once on 2B images
//   of_copy_channels     strided channel-slice copy / accumulate with dtype conversion
//                        (tf.concat of model.py:100-102 and its gradient split)
//   of_bn_fold_fwd/bwd   inference-mode BatchNormalization (model.py:14,
//                        resnet/models.py:19,24) folded into the preceding conv:
//                        w' = w * s, b' = (b - mean) * s + beta, s = gamma / sqrt(var + eps),
//                        and the exact chain rule back to (w, b, gamma, beta).
#include "common.cuh"

static inline int g_grid(size_t work, int threads = 256) {
  size_t blocks = (work + threads - 1) / threads;
  size_t cap = (size_t)OF_NUM_SMS * 16;
  if (blocks > cap) blocks = cap;
  return (int)(blocks < 1 ? 1 : blocks);
}

template <typename TO>
__global__ void __launch_bounds__(256) split_pair_kernel(const float* __restrict__ imgs, TO* __restrict__ out,
                                                         size_t npix_b /* B*H*W */) {
  // out[n = half*B + b] pixel p, channel c  <-  imgs[b, p, half*3 + c]
  size_t total = npix_b * 6;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    size_t pix = e / 6;
    int c6 = (int)(e - pix * 6);
    int half = c6 / 3, c = c6 - half * 3;
    st_f<TO>(out + ((size_t)half * npix_b + pix) * 3 + c, __ldg(imgs + e));
  }
}

template <typename TS, typename TD, bool ACC>
__global__ void __launch_bounds__(256) copy_channels_kernel(const TS* __restrict__ src, int src_stride,
                                                            TD* __restrict__ dst, int dst_stride, int nch,
                                                            size_t npix) {
  size_t total = npix * nch;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    size_t pix = e / nch;
    int c = (int)(e - pix * nch);
    float v = ld_f<TS>(src + pix * src_stride + c);
    TD* d = dst + pix * dst_stride + c;
    if (ACC) v += ld_f<TD>(d);
    st_f<TD>(d, v);
  }
}

// 4 channels per thread (float4 / 4 x bf16): every stride, count and base pointer a multiple of 4 elements
template <typename TS, typename TD, bool ACC>
__global__ void __launch_bounds__(256) copy_channels4_kernel(const TS* __restrict__ src, int src_stride,
                                                             TD* __restrict__ dst, int dst_stride, int nch4,
                                                             size_t npix) {
  size_t total = npix * nch4;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    size_t pix = e / nch4;
    int c = (int)(e - pix * nch4) * 4;
    float4 v = ld4<TS>(src + pix * src_stride + c);
    TD* d = dst + pix * dst_stride + c;
    if (ACC) {
      float4 o = ld4<TD>(d);
      v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
    }
    st4<TD>(d, v);
  }
}

__global__ void __launch_bounds__(256) bn_fold_fwd_kernel(const float* __restrict__ w, const float* __restrict__ b,
                                                          const float* __restrict__ gamma,
                                                          const float* __restrict__ beta,
                                                          const float* __restrict__ mean,
                                                          const float* __restrict__ var, float eps,
                                                          float* __restrict__ wf, float* __restrict__ bf, int K,
                                                          int Cout) {
  size_t total = (size_t)K * Cout;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int co = (int)(e % Cout);
    float s = gamma[co] / sqrtf(var[co] + eps);
    wf[e] = w[e] * s;
    if (e < (size_t)Cout) bf[co] = ((b ? b[co] : 0.f) - mean[co]) * s + beta[co];
  }
}

// one CTA per output channel: dgamma needs sum_k dwf[k,co] * w[k,co]
__global__ void __launch_bounds__(256) bn_fold_bwd_kernel(const float* __restrict__ w, const float* __restrict__ b,
                                                          const float* __restrict__ gamma,
                                                          const float* __restrict__ mean,
                                                          const float* __restrict__ var, float eps,
                                                          const float* __restrict__ dwf,
                                                          const float* __restrict__ dbf, float* __restrict__ dw,
                                                          float* __restrict__ db, float* __restrict__ dgamma,
                                                          float* __restrict__ dbeta, int K, int Cout) {
  int co = blockIdx.x;
  float inv_std = 1.0f / sqrtf(var[co] + eps);
  float s = gamma[co] * inv_std;
  float acc = 0.f;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    float g = dwf[(size_t)k * Cout + co];
    acc += g * w[(size_t)k * Cout + co];
    dw[(size_t)k * Cout + co] = g * s;
  }
  __shared__ float sm[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += sm[i];
    float g_b = dbf[co];
    dgamma[co] = (t + g_b * ((b ? b[co] : 0.f) - mean[co])) * inv_std;
    dbeta[co] = g_b;
    if (db) db[co] = g_b * s;
  }
}

// ---- the same two folds for every conv + BN pair of the model, one launch each (26 tiny launches -> 2) ----
__global__ void __launch_bounds__(256) bn_fold_fwd_batch_kernel(const of_bn_job* __restrict__ jobs, int njobs,
                                                                float eps) {
  int j = 0;
  while (j + 1 < njobs && (int)blockIdx.x >= jobs[j + 1].fwd_block_begin) ++j;
  const of_bn_job J = jobs[j];
  const size_t total = (size_t)J.K * J.Cout;
  const size_t e0 = (size_t)((int)blockIdx.x - J.fwd_block_begin) * OF_BN_FWD_BLOCK;
  for (size_t e = e0 + threadIdx.x; e < e0 + OF_BN_FWD_BLOCK && e < total; e += 256) {
    int co = (int)(e % J.Cout);
    float s = J.gamma[co] / sqrtf(J.var[co] + eps);
    J.wf[e] = J.w[e] * s;
    if (e < (size_t)J.Cout) J.bf[co] = ((J.b ? J.b[co] : 0.f) - J.mean[co]) * s + J.beta[co];
  }
}

__global__ void __launch_bounds__(256) bn_fold_bwd_batch_kernel(const of_bn_job* __restrict__ jobs, int njobs,
                                                                float eps) {
  int j = 0;
  while (j + 1 < njobs && (int)blockIdx.x >= jobs[j + 1].bwd_block_begin) ++j;
  const of_bn_job J = jobs[j];
  const int co = (int)blockIdx.x - J.bwd_block_begin, Cout = J.Cout;
  float inv_std = 1.0f / sqrtf(J.var[co] + eps);
  float s = J.gamma[co] * inv_std;
  float acc = 0.f;
  for (int k = threadIdx.x; k < J.K; k += blockDim.x) {
    float g = J.dwf[(size_t)k * Cout + co];
    acc += g * J.w[(size_t)k * Cout + co];
    J.dw[(size_t)k * Cout + co] = g * s;
  }
  __shared__ float sm[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += sm[i];
    float g_b = J.dbf[co];
    J.dgamma[co] = (t + g_b * ((J.b ? J.b[co] : 0.f) - J.mean[co])) * inv_std;
    J.dbeta[co] = g_b;
    if (J.db) J.db[co] = g_b * s;
  }
}

// zero-insertion x2 (stride-2 dgrad = stride-1 dgrad of the zero-upsampled dy)
template <typename T>
__global__ void __launch_bounds__(256) zero_upsample2_kernel(const T* __restrict__ dy, T* __restrict__ out, int N,
                                                             int Ho, int Wo, int H, int W, int C4) {
  size_t total = (size_t)N * H * W * C4;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(e % C4);
    size_t r = e / C4;
    int w = (int)(r % W); r /= W;
    int h = (int)(r % H);
    int n = (int)(r / H);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!(h & 1) && !(w & 1) && (h >> 1) < Ho && (w >> 1) < Wo)
      v = ld4<T>(dy + ((((size_t)n * Ho + (h >> 1)) * Wo + (w >> 1)) * C4 + c) * 4);
    st4<T>(out + e * 4, v);
  }
}

// the non-zero quarter only: out[2i, 2j] = dy[i, j]; the caller keeps the other positions of `out` zero (a buffer it
// zero-filled once and dedicates to this layer), so a step writes 1/4 of the tensor instead of all of it
template <typename T>
__global__ void __launch_bounds__(256) scatter_upsample2_kernel(const T* __restrict__ dy, T* __restrict__ out, int N,
                                                                int Ho, int Wo, int H, int W, int C4) {
  size_t total = (size_t)N * Ho * Wo * C4;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(e % C4);
    size_t r = e / C4;
    int w = (int)(r % Wo); r /= Wo;
    int h = (int)(r % Ho);
    int n = (int)(r / Ho);
    if (2 * h < H && 2 * w < W)
      st4<T>(out + ((((size_t)n * H + 2 * h) * W + 2 * w) * C4 + c) * 4, ld4<T>(dy + e * 4));
  }
}

// the same in 16-byte units (bf16: 8 channels, fp32: 4): a pure copy, twice the bytes in flight per thread for bf16
__global__ void __launch_bounds__(256) scatter_upsample2_u16_kernel(const uint4* __restrict__ dy, uint4* __restrict__ out,
                                                                    int N, int Ho, int Wo, int H, int W, int CU) {
  size_t total = (size_t)N * Ho * Wo * CU;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(e % CU);
    size_t r = e / CU;
    int w = (int)(r % Wo); r /= Wo;
    int h = (int)(r % Ho);
    int n = (int)(r / Ho);
    if (2 * h < H && 2 * w < W) out[(((size_t)n * H + 2 * h) * W + 2 * w) * CU + c] = __ldg(dy + e);
  }
}

// ---- stem (Cin = 3, 7x7 / 2) on tensor cores: packed image + weight (un)packing ------------------
// out (2B, (H+6)/2, W+6, 2, 4) bf16: frame 1s then frame 2s; TF-SAME zero border (2 rows / 2 cols before, 4 after);
// one 16-byte texel = the pixel of TWO consecutive image rows (row pair pr = rows 2 pr - 2, 2 pr - 1) x 4 channels
// (3 used).  A window of 8 texels is then one 64-element K chunk holding two kh taps x 7 kw x 3 channels, and the
// 7x7 / 2 stem is a 4-tap, stride-1 conv over row pairs (K = 256 instead of 7 x 64 = 448).
__global__ void __launch_bounds__(256) pack_pair8_kernel(const float* __restrict__ imgs, uint4* __restrict__ out,
                                                         int B, int H, int W) {
  const int Hp = (H + 6) / 2, Wp = W + 6;
  size_t total = (size_t)2 * B * Hp * Wp;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(e % Wp) - 2;
    size_t r2 = e / Wp;
    int r0 = 2 * (int)(r2 % Hp) - 2;
    int n = (int)(r2 / Hp);
    int half = n >= B, b = n - half * B;
    uint32_t wv[4] = {0u, 0u, 0u, 0u};
    if (c >= 0 && c < W) {
#pragma unroll
      for (int p = 0; p < 2; ++p) {
        const int r = r0 + p;
        if (r >= 0 && r < H) {
          const float* q = imgs + (((size_t)b * H + r) * W + c) * 6 + half * 3;
          __nv_bfloat162 a = __floats2bfloat162_rn(__ldg(q), __ldg(q + 1));
          __nv_bfloat162 d = __floats2bfloat162_rn(__ldg(q + 2), 0.f);
          wv[2 * p] = *reinterpret_cast<uint32_t*>(&a);
          wv[2 * p + 1] = *reinterpret_cast<uint32_t*>(&d);
        }
      }
    }
    out[e] = make_uint4(wv[0], wv[1], wv[2], wv[3]);
  }
}
// w (7,7,3,Cout) f32 -> bf16 [tap][co][kw*8 + p*4 + c],  kh = 2 tap + p  (kh = 7, kw = 7 and c = 3 are zero)
__global__ void __launch_bounds__(256) conv1_pack_w_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out,
                                                           int Cout) {
  int total = 4 * Cout * 64;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    int col = e & 63, co = (e >> 6) % Cout, tap = (e >> 6) / Cout;
    int kw = col >> 3, p = (col >> 2) & 1, c = col & 3;
    int kh = 2 * tap + p;
    float v = (kh < 7 && kw < 7 && c < 3) ? w[((size_t)(kh * 7 + kw) * 3 + c) * Cout + co] : 0.f;
    out[e] = __float2bfloat16_rn(v);
  }
}
// dw_view [tap][kw*8 + p*4 + c][co] f32 -> dw (7,7,3,Cout)
__global__ void __launch_bounds__(256) conv1_unpack_dw_kernel(const float* __restrict__ dwv, float* __restrict__ dw,
                                                              int Cout) {
  int total = 7 * 7 * 3 * Cout;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    int co = e % Cout, t = e / Cout;
    int c = t % 3; t /= 3;
    int kw = t % 7, kh = t / 7;
    dw[e] = dwv[((size_t)(kh >> 1) * 64 + kw * 8 + (kh & 1) * 4 + c) * Cout + co];
  }
}

extern "C" {

int of_zero_upsample2(const void* dy, void* out, int N, int Ho, int Wo, int H, int W, int C, int dtype,
                      void* stream) {
  OF_REQUIRE(dy && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(N > 0 && Ho > 0 && Wo > 0 && H > 0 && W > 0 && C > 0 && C % 4 == 0, OF_ERR_SHAPE, "bad dims");
  OF_REQUIRE(of_valid_dtype(dtype), OF_ERR_ARG, "bad dtype");
  size_t total = (size_t)N * H * W * (C / 4);
  if (dtype == OF_F32)
    zero_upsample2_kernel<float><<<g_grid(total), 256, 0, (cudaStream_t)stream>>>((const float*)dy, (float*)out, N, Ho, Wo, H, W, C / 4);
  else
    zero_upsample2_kernel<__nv_bfloat16><<<g_grid(total), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)dy, (__nv_bfloat16*)out, N, Ho, Wo, H, W, C / 4);
  return of_launch_check("zero_upsample2_kernel");
}

int of_scatter_upsample2(const void* dy, void* out, int N, int Ho, int Wo, int H, int W, int C, int dtype,
                         void* stream) {
  OF_REQUIRE(dy && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(N > 0 && Ho > 0 && Wo > 0 && H > 0 && W > 0 && C > 0 && C % 4 == 0, OF_ERR_SHAPE, "bad dims");
  OF_REQUIRE(of_valid_dtype(dtype), OF_ERR_ARG, "bad dtype");
  size_t total = (size_t)N * Ho * Wo * (C / 4);
  if (dtype == OF_F32)
    scatter_upsample2_kernel<float><<<g_grid(total), 256, 0, (cudaStream_t)stream>>>((const float*)dy, (float*)out, N, Ho, Wo, H, W, C / 4);
  else if (C % 8 == 0 && (((uintptr_t)dy | (uintptr_t)out) & 15) == 0)
    scatter_upsample2_u16_kernel<<<g_grid(total / 2), 256, 0, (cudaStream_t)stream>>>((const uint4*)dy, (uint4*)out, N, Ho, Wo,
                                                                                      H, W, C / 8);
  else
    scatter_upsample2_kernel<__nv_bfloat16><<<g_grid(total), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)dy, (__nv_bfloat16*)out, N, Ho, Wo, H, W, C / 4);
  return of_launch_check("scatter_upsample2_kernel");
}

int of_split_image_pair(const float* imgs, void* out, int B, int H, int W, int out_dtype, void* stream) {
  OF_REQUIRE(imgs && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(B > 0 && H > 0 && W > 0, OF_ERR_SHAPE, "bad dims");
  OF_REQUIRE(of_valid_dtype(out_dtype), OF_ERR_ARG, "bad dtype");
  size_t npix = (size_t)B * H * W;
  if (out_dtype == OF_F32)
    split_pair_kernel<float><<<g_grid(npix * 6), 256, 0, (cudaStream_t)stream>>>(imgs, (float*)out, npix);
  else
    split_pair_kernel<__nv_bfloat16><<<g_grid(npix * 6), 256, 0, (cudaStream_t)stream>>>(imgs, (__nv_bfloat16*)out, npix);
  return of_launch_check("split_pair_kernel");
}

int of_copy_channels(const void* src, int src_stride, void* dst, int dst_stride, int nch, size_t npix,
                     int src_dtype, int dst_dtype, int accumulate, void* stream) {
  OF_REQUIRE(src && dst, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(nch > 0 && src_stride >= nch && dst_stride >= nch, OF_ERR_SHAPE, "bad channel strides");
  OF_REQUIRE(of_valid_dtype(src_dtype) && of_valid_dtype(dst_dtype), OF_ERR_ARG, "bad dtype");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t ssz = src_dtype == OF_F32 ? 4 : 2, dsz = dst_dtype == OF_F32 ? 4 : 2;
  if (nch % 4 == 0 && src_stride % 4 == 0 && dst_stride % 4 == 0 && (uintptr_t)src % (4 * ssz) == 0 &&
      (uintptr_t)dst % (4 * dsz) == 0) {
    int grid4 = g_grid(npix * (nch / 4));
#define GO4(TS, TD)                                                                                         \
  do {                                                                                                      \
    if (accumulate)                                                                                         \
      copy_channels4_kernel<TS, TD, true><<<grid4, 256, 0, st>>>((const TS*)src, src_stride, (TD*)dst, dst_stride, nch / 4, npix); \
    else                                                                                                    \
      copy_channels4_kernel<TS, TD, false><<<grid4, 256, 0, st>>>((const TS*)src, src_stride, (TD*)dst, dst_stride, nch / 4, npix); \
  } while (0)
    if (src_dtype == OF_F32 && dst_dtype == OF_F32) GO4(float, float);
    else if (src_dtype == OF_F32 && dst_dtype == OF_BF16) GO4(float, __nv_bfloat16);
    else if (src_dtype == OF_BF16 && dst_dtype == OF_F32) GO4(__nv_bfloat16, float);
    else GO4(__nv_bfloat16, __nv_bfloat16);
#undef GO4
    return of_launch_check("copy_channels4_kernel");
  }
  int grid = g_grid(npix * nch);
#define GO(TS, TD)                                                                                          \
  do {                                                                                                      \
    if (accumulate)                                                                                         \
      copy_channels_kernel<TS, TD, true><<<grid, 256, 0, st>>>((const TS*)src, src_stride, (TD*)dst, dst_stride, nch, npix); \
    else                                                                                                    \
      copy_channels_kernel<TS, TD, false><<<grid, 256, 0, st>>>((const TS*)src, src_stride, (TD*)dst, dst_stride, nch, npix); \
  } while (0)
  if (src_dtype == OF_F32 && dst_dtype == OF_F32) GO(float, float);
  else if (src_dtype == OF_F32 && dst_dtype == OF_BF16) GO(float, __nv_bfloat16);
  else if (src_dtype == OF_BF16 && dst_dtype == OF_F32) GO(__nv_bfloat16, float);
  else GO(__nv_bfloat16, __nv_bfloat16);
#undef GO
  return of_launch_check("copy_channels_kernel");
}

int of_bn_fold_fwd(const float* w, const float* b, const float* gamma, const float* beta, const float* mean,
                   const float* var, float eps, float* w_folded, float* b_folded, int K, int Cout, void* stream) {
  OF_REQUIRE(w && gamma && beta && mean && var && w_folded && b_folded, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(K > 0 && Cout > 0, OF_ERR_SHAPE, "bad dims");
  bn_fold_fwd_kernel<<<g_grid((size_t)K * Cout), 256, 0, (cudaStream_t)stream>>>(w, b, gamma, beta, mean, var, eps,
                                                                                 w_folded, b_folded, K, Cout);
  return of_launch_check("bn_fold_fwd_kernel");
}

int of_bn_fold_bwd(const float* w, const float* b, const float* gamma, const float* mean, const float* var,
                   float eps, const float* dw_folded, const float* db_folded, float* dw, float* db, float* dgamma,
                   float* dbeta, int K, int Cout, void* stream) {
  OF_REQUIRE(w && gamma && mean && var && dw_folded && db_folded && dw && dgamma && dbeta, OF_ERR_ARG,
             "null pointer argument");
  OF_REQUIRE(K > 0 && Cout > 0, OF_ERR_SHAPE, "bad dims");
  bn_fold_bwd_kernel<<<Cout, 256, 0, (cudaStream_t)stream>>>(w, b, gamma, mean, var, eps, dw_folded, db_folded, dw,
                                                             db, dgamma, dbeta, K, Cout);
  return of_launch_check("bn_fold_bwd_kernel");
}

int of_bn_fold_fwd_batch(const of_bn_job* jobs_device, int njobs, int total_blocks, float eps, void* stream) {
  OF_REQUIRE(jobs_device && njobs > 0 && total_blocks > 0, OF_ERR_ARG, "bad BN job table");
  bn_fold_fwd_batch_kernel<<<total_blocks, 256, 0, (cudaStream_t)stream>>>(jobs_device, njobs, eps);
  return of_launch_check("bn_fold_fwd_batch_kernel");
}

int of_bn_fold_bwd_batch(const of_bn_job* jobs_device, int njobs, int total_blocks, float eps, void* stream) {
  OF_REQUIRE(jobs_device && njobs > 0 && total_blocks > 0, OF_ERR_ARG, "bad BN job table");
  bn_fold_bwd_batch_kernel<<<total_blocks, 256, 0, (cudaStream_t)stream>>>(jobs_device, njobs, eps);
  return of_launch_check("bn_fold_bwd_batch_kernel");
}

int of_pack_image_pair8(const float* imgs, void* out, int B, int H, int W, void* stream) {
  OF_REQUIRE(imgs && out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(B > 0 && H > 0 && W > 0 && H % 2 == 0 && W % 2 == 0, OF_ERR_SHAPE, "bad dims (even H, W)");
  size_t total = (size_t)2 * B * ((H + 6) / 2) * (W + 6);
  pack_pair8_kernel<<<g_grid(total), 256, 0, (cudaStream_t)stream>>>(imgs, (uint4*)out, B, H, W);
  return of_launch_check("pack_pair8_kernel");
}

int of_conv1_pack_weights(const float* w_hwio, void* packed, int Cout, void* stream) {
  OF_REQUIRE(w_hwio && packed && Cout > 0, OF_ERR_ARG, "bad argument");
  conv1_pack_w_kernel<<<g_grid((size_t)4 * Cout * 64), 256, 0, (cudaStream_t)stream>>>(w_hwio, (__nv_bfloat16*)packed, Cout);
  return of_launch_check("conv1_pack_w_kernel");
}

int of_conv1_unpack_wgrad(const float* dw_view, float* dw_hwio, int Cout, void* stream) {
  OF_REQUIRE(dw_view && dw_hwio && Cout > 0, OF_ERR_ARG, "bad argument");
  conv1_unpack_dw_kernel<<<g_grid((size_t)147 * Cout), 256, 0, (cudaStream_t)stream>>>(dw_view, dw_hwio, Cout);
  return of_launch_check("conv1_unpack_dw_kernel");
}

}  // extern "C"
