// Training-mode BatchNormalization (OPT-IN: the reference's step runs its
// layers.BatchNormalization in inference mode - SURVEY F5 - and that is what the fused train step does; these
// kernels are the building blocks of the batch-statistics variant a user may want instead).
//   y = act(gamma * (x - mean_B) / sqrt(var_B + eps) + beta),  mean_B / var_B over (N, H, W), var_B biased
//   moving = moving * momentum + batch * (1 - momentum)        (Keras momentum 0.99, eps 1e-3; the moving variance
//                                                               takes the unbiased batch variance like TF's fused op)
// Four HBM-bound passes over NHWC (fp32 or bf16, C % 4 == 0), per-channel sums in fp32:
//   stats        sum[c] += x, sumsq[c] += x^2                        (buffers the caller can all-reduce: sync BN)
//   apply        y = act(a[c] x + b[c]) (+ moving-statistics update by the first block)
//   bwd_reduce   dsum[c] += dy', dsum_xhat[c] += dy' * xhat,  dy' = dy * act'(y)     (= dbeta, dgamma)
//   bwd_apply    dx = (gamma / sigma) * (dy' - dsum / n - xhat * dsum_xhat / n)
// A CTA walks pixel rows with a fixed channel quad per thread, so the per-channel partials stay in registers and
// leave as one red.global.add per channel per CTA.
#include "common.cuh"

#define BN_THREADS 256

__device__ __forceinline__ float bn_act(float v, int act) {
  return act == OF_ACT_RELU ? fmaxf(v, 0.f) : (act == OF_ACT_LEAKY ? (v > 0.f ? v : 0.3f * v) : v);
}
__device__ __forceinline__ float bn_act_grad(float y, int act) {
  return act == OF_ACT_RELU ? (y > 0.f ? 1.f : 0.f) : (act == OF_ACT_LEAKY ? (y > 0.f ? 1.f : 0.3f) : 1.f);
}

// thread -> (channel quad cq = tid % CV, pixel lane pl = tid / CV); requires BN_THREADS % CV == 0
template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_stats_kernel(const T* __restrict__ x, float* __restrict__ sum,
                                                              float* __restrict__ sumsq, size_t npix, int C) {
  const int CV = C / 4, cq = threadIdx.x % CV, pl = threadIdx.x / CV, PL = BN_THREADS / CV;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = s;
  for (size_t p = (size_t)blockIdx.x * PL + pl; p < npix; p += (size_t)gridDim.x * PL) {
    const float4 v = ld4<T>(x + p * C + cq * 4);
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    q.x += v.x * v.x; q.y += v.y * v.y; q.z += v.z * v.z; q.w += v.w * v.w;
  }
  __shared__ float4 sh[2][BN_THREADS];
  sh[0][threadIdx.x] = s; sh[1][threadIdx.x] = q;
  __syncthreads();
  if (pl == 0) {
    for (int k = 1; k < PL; ++k) {
      const float4 a = sh[0][k * CV + cq], b = sh[1][k * CV + cq];
      s.x += a.x; s.y += a.y; s.z += a.z; s.w += a.w;
      q.x += b.x; q.y += b.y; q.z += b.z; q.w += b.w;
    }
    red_add_v4(sum + cq * 4, s);
    red_add_v4(sumsq + cq * 4, q);
  }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_apply_kernel(const T* __restrict__ x, const float* __restrict__ sum,
                                                              const float* __restrict__ sumsq, float count,
                                                              const float* __restrict__ gamma,
                                                              const float* __restrict__ beta, float eps, int act,
                                                              T* __restrict__ y, float* __restrict__ moving_mean,
                                                              float* __restrict__ moving_var, float momentum,
                                                              size_t npix, int C) {
  const int CV = C / 4, cq = threadIdx.x % CV, pl = threadIdx.x / CV, PL = BN_THREADS / CV;
  float a[4], b[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = cq * 4 + k;
    const float mean = sum[c] / count;
    const float var = fmaxf(sumsq[c] / count - mean * mean, 0.f);
    a[k] = gamma[c] / sqrtf(var + eps);
    b[k] = beta[c] - mean * a[k];
    if (moving_mean != nullptr && blockIdx.x == 0 && pl == 0) {
      const float unbiased = count > 1.f ? var * count / (count - 1.f) : var;
      moving_mean[c] = moving_mean[c] * momentum + mean * (1.f - momentum);
      moving_var[c] = moving_var[c] * momentum + unbiased * (1.f - momentum);
    }
  }
  for (size_t p = (size_t)blockIdx.x * PL + pl; p < npix; p += (size_t)gridDim.x * PL) {
    const float4 v = ld4<T>(x + p * C + cq * 4);
    st4<T>(y + p * C + cq * 4, make_float4(bn_act(a[0] * v.x + b[0], act), bn_act(a[1] * v.y + b[1], act),
                                           bn_act(a[2] * v.z + b[2], act), bn_act(a[3] * v.w + b[3], act)));
  }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_bwd_reduce_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                                   const T* __restrict__ dy,
                                                                   const float* __restrict__ sum,
                                                                   const float* __restrict__ sumsq, float count,
                                                                   float eps, int act, float* __restrict__ dsum,
                                                                   float* __restrict__ dsum_xhat, size_t npix, int C) {
  const int CV = C / 4, cq = threadIdx.x % CV, pl = threadIdx.x / CV, PL = BN_THREADS / CV;
  float mean[4], inv[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = cq * 4 + k;
    mean[k] = sum[c] / count;
    inv[k] = 1.0f / sqrtf(fmaxf(sumsq[c] / count - mean[k] * mean[k], 0.f) + eps);
  }
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = s;
  for (size_t p = (size_t)blockIdx.x * PL + pl; p < npix; p += (size_t)gridDim.x * PL) {
    const size_t o = p * C + cq * 4;
    const float4 xv = ld4<T>(x + o), yv = ld4<T>(y + o), g = ld4<T>(dy + o);
    const float g0 = g.x * bn_act_grad(yv.x, act), g1 = g.y * bn_act_grad(yv.y, act);
    const float g2 = g.z * bn_act_grad(yv.z, act), g3 = g.w * bn_act_grad(yv.w, act);
    s.x += g0; s.y += g1; s.z += g2; s.w += g3;
    q.x += g0 * (xv.x - mean[0]) * inv[0]; q.y += g1 * (xv.y - mean[1]) * inv[1];
    q.z += g2 * (xv.z - mean[2]) * inv[2]; q.w += g3 * (xv.w - mean[3]) * inv[3];
  }
  __shared__ float4 sh[2][BN_THREADS];
  sh[0][threadIdx.x] = s; sh[1][threadIdx.x] = q;
  __syncthreads();
  if (pl == 0) {
    for (int k = 1; k < PL; ++k) {
      const float4 a = sh[0][k * CV + cq], b = sh[1][k * CV + cq];
      s.x += a.x; s.y += a.y; s.z += a.z; s.w += a.w;
      q.x += b.x; q.y += b.y; q.z += b.z; q.w += b.w;
    }
    red_add_v4(dsum + cq * 4, s);
    red_add_v4(dsum_xhat + cq * 4, q);
  }
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_bwd_apply_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                                  const T* __restrict__ dy,
                                                                  const float* __restrict__ sum,
                                                                  const float* __restrict__ sumsq, float count,
                                                                  const float* __restrict__ gamma, float eps, int act,
                                                                  const float* __restrict__ dsum,
                                                                  const float* __restrict__ dsum_xhat,
                                                                  T* __restrict__ dx, size_t npix, int C) {
  const int CV = C / 4, cq = threadIdx.x % CV, pl = threadIdx.x / CV, PL = BN_THREADS / CV;
  float mean[4], inv[4], sc[4], m1[4], m2[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = cq * 4 + k;
    mean[k] = sum[c] / count;
    inv[k] = 1.0f / sqrtf(fmaxf(sumsq[c] / count - mean[k] * mean[k], 0.f) + eps);
    sc[k] = gamma[c] * inv[k];
    m1[k] = dsum[c] / count;
    m2[k] = dsum_xhat[c] / count;
  }
  for (size_t p = (size_t)blockIdx.x * PL + pl; p < npix; p += (size_t)gridDim.x * PL) {
    const size_t o = p * C + cq * 4;
    const float4 xv = ld4<T>(x + o), yv = ld4<T>(y + o), g = ld4<T>(dy + o);
    const float xs[4] = {xv.x, xv.y, xv.z, xv.w}, ys[4] = {yv.x, yv.y, yv.z, yv.w}, gs[4] = {g.x, g.y, g.z, g.w};
    float r[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float gp = gs[k] * bn_act_grad(ys[k], act);
      r[k] = sc[k] * (gp - m1[k] - (xs[k] - mean[k]) * inv[k] * m2[k]);
    }
    st4<T>(dx + o, make_float4(r[0], r[1], r[2], r[3]));
  }
}

// ------------------------------------------------------------------ C ABI --
static inline int bn_grid(size_t npix, int C) {
  const int PL = BN_THREADS / (C / 4);
  size_t g = (npix + (size_t)PL * 8 - 1) / ((size_t)PL * 8);      // >= 8 pixels per thread: amortise the atomics
  const size_t cap = (size_t)OF_NUM_SMS * 8;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}
#define BN_CHECK_DIMS()                                                                                         \
  OF_REQUIRE(npix > 0 && C >= 4 && C % 4 == 0 && C <= 4 * BN_THREADS && BN_THREADS % (C / 4) == 0, OF_ERR_SHAPE, \
             "%s: needs C %% 4 == 0 and C / 4 dividing %d (C = %d)", __func__, BN_THREADS, C);                  \
  OF_REQUIRE(of_valid_dtype(dtype), OF_ERR_ARG, "%s: bad dtype %d", __func__, dtype)

extern "C" int of_bn_train_stats(const void* x, float* sum, float* sumsq, size_t npix, int C, int dtype, void* stream) {
  OF_REQUIRE(x && sum && sumsq, OF_ERR_ARG, "of_bn_train_stats: null pointer");
  BN_CHECK_DIMS();
  cudaStream_t st = (cudaStream_t)stream;
  OF_CUDA(cudaMemsetAsync(sum, 0, (size_t)C * sizeof(float), st));
  OF_CUDA(cudaMemsetAsync(sumsq, 0, (size_t)C * sizeof(float), st));
  if (dtype == OF_F32) bn_stats_kernel<float><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const float*)x, sum, sumsq, npix, C);
  else bn_stats_kernel<__nv_bfloat16><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const __nv_bfloat16*)x, sum, sumsq, npix, C);
  return of_launch_check("bn_stats_kernel");
}

extern "C" int of_bn_train_apply(const void* x, const float* sum, const float* sumsq, float count, const float* gamma,
                                 const float* beta, float eps, int act, void* y, float* moving_mean, float* moving_var,
                                 float momentum, size_t npix, int C, int dtype, void* stream) {
  OF_REQUIRE(x && sum && sumsq && gamma && beta && y, OF_ERR_ARG, "of_bn_train_apply: null pointer");
  OF_REQUIRE((moving_mean == nullptr) == (moving_var == nullptr), OF_ERR_ARG, "of_bn_train_apply: pass both moving statistics or neither");
  OF_REQUIRE(count >= 1.f, OF_ERR_ARG, "of_bn_train_apply: count %f", (double)count);
  BN_CHECK_DIMS();
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == OF_F32)
    bn_apply_kernel<float><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const float*)x, sum, sumsq, count, gamma, beta, eps, act,
                                                                    (float*)y, moving_mean, moving_var, momentum, npix, C);
  else
    bn_apply_kernel<__nv_bfloat16><<<bn_grid(npix, C), BN_THREADS, 0, st>>>(
        (const __nv_bfloat16*)x, sum, sumsq, count, gamma, beta, eps, act, (__nv_bfloat16*)y, moving_mean, moving_var,
        momentum, npix, C);
  return of_launch_check("bn_apply_kernel");
}

extern "C" int of_bn_train_bwd_reduce(const void* x, const void* y, const void* dy, const float* sum, const float* sumsq,
                                      float count, float eps, int act, float* dsum, float* dsum_xhat, size_t npix, int C,
                                      int dtype, void* stream) {
  OF_REQUIRE(x && y && dy && sum && sumsq && dsum && dsum_xhat, OF_ERR_ARG, "of_bn_train_bwd_reduce: null pointer");
  BN_CHECK_DIMS();
  cudaStream_t st = (cudaStream_t)stream;
  OF_CUDA(cudaMemsetAsync(dsum, 0, (size_t)C * sizeof(float), st));
  OF_CUDA(cudaMemsetAsync(dsum_xhat, 0, (size_t)C * sizeof(float), st));
  if (dtype == OF_F32)
    bn_bwd_reduce_kernel<float><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const float*)x, (const float*)y, (const float*)dy, sum,
                                                                         sumsq, count, eps, act, dsum, dsum_xhat, npix, C);
  else
    bn_bwd_reduce_kernel<__nv_bfloat16><<<bn_grid(npix, C), BN_THREADS, 0, st>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy, sum, sumsq, count, eps, act, dsum,
        dsum_xhat, npix, C);
  return of_launch_check("bn_bwd_reduce_kernel");
}

extern "C" int of_bn_train_bwd_apply(const void* x, const void* y, const void* dy, const float* sum, const float* sumsq,
                                     float count, const float* gamma, float eps, int act, const float* dsum,
                                     const float* dsum_xhat, void* dx, size_t npix, int C, int dtype, void* stream) {
  OF_REQUIRE(x && y && dy && sum && sumsq && gamma && dsum && dsum_xhat && dx, OF_ERR_ARG, "of_bn_train_bwd_apply: null pointer");
  BN_CHECK_DIMS();
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == OF_F32)
    bn_bwd_apply_kernel<float><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const float*)x, (const float*)y, (const float*)dy, sum,
                                                                        sumsq, count, gamma, eps, act, dsum, dsum_xhat,
                                                                        (float*)dx, npix, C);
  else
    bn_bwd_apply_kernel<__nv_bfloat16><<<bn_grid(npix, C), BN_THREADS, 0, st>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy, sum, sumsq, count, gamma, eps, act, dsum,
        dsum_xhat, (__nv_bfloat16*)dx, npix, C);
  return of_launch_check("bn_bwd_apply_kernel");
}
