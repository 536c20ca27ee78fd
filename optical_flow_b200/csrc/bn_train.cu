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
// A CTA walks pixel rows with a fixed channel quad per thread, so the per-channel partials stay in registers.
// Reductions are DETERMINISTIC: every CTA stores its per-channel partials to a workspace and the last CTA to
// arrive (ticket) adds them in CTA order in double precision - no floating-point atomics, so two runs (and two
// boxes) give bit-identical statistics.  Sums are taken about a per-channel `shift` (nullable = 0; the layer passes
// its moving mean, which every replica holds identically, so the buffers stay all-reducible):
//   sum = sum_p (x - shift), sumsq = sum_p (x - shift)^2, mean = shift + sum / n, var = sumsq / n - (sum / n)^2,
// which removes the E[x^2] - E[x]^2 cancellation once the moving mean tracks the batch mean.
#include "common.cuh"

#define BN_THREADS 256
#define BN_MAX_GRID (OF_NUM_SMS * 8)
#define BN_WS_HDR 64          // ticket counters of the two reducing kernels

__device__ __forceinline__ float bn_act(float v, int act) {
  return act == OF_ACT_RELU ? fmaxf(v, 0.f) : (act == OF_ACT_LEAKY ? (v > 0.f ? v : 0.3f * v) : v);
}
__device__ __forceinline__ float bn_act_grad(float y, int act) {
  return act == OF_ACT_RELU ? (y > 0.f ? 1.f : 0.f) : (act == OF_ACT_LEAKY ? (y > 0.f ? 1.f : 0.3f) : 1.f);
}

// per-channel batch statistics from the (shifted, possibly all-reduced) sums
struct BnMoments { float mean, var; };
__device__ __forceinline__ BnMoments bn_moments(const float* sum, const float* sumsq, const float* shift, float count,
                                                int c) {
  const float m = sum[c] / count;
  BnMoments r;
  r.var = fmaxf(sumsq[c] / count - m * m, 0.f);
  r.mean = (shift ? shift[c] : 0.f) + m;
  return r;
}

// Deterministic two-stage per-channel reduction of two float4 accumulators per thread (thread -> channel quad
// cq = tid % CV, pixel lane pl = tid / CV).  Stage 1: fixed-order sum over the CTA's pixel lanes, partial to
// part[blockIdx][2][C].  Stage 2 (last CTA by ticket): out0[c] / out1[c] = sum over CTAs in index order, double.
__device__ __forceinline__ void bn_two_stage(float4 s, float4 q, int C, float* __restrict__ part, unsigned* ticket,
                                             float* __restrict__ out0, float* __restrict__ out1) {
  const int CV = C / 4, cq = threadIdx.x % CV, pl = threadIdx.x / CV, PL = BN_THREADS / CV;
  __shared__ float4 sh[2][BN_THREADS];
  __shared__ bool s_last;
  sh[0][threadIdx.x] = s; sh[1][threadIdx.x] = q;
  __syncthreads();
  if (pl == 0) {
    for (int k = 1; k < PL; ++k) {
      const float4 a = sh[0][k * CV + cq], b = sh[1][k * CV + cq];
      s.x += a.x; s.y += a.y; s.z += a.z; s.w += a.w;
      q.x += b.x; q.y += b.y; q.z += b.z; q.w += b.w;
    }
    float* dst = part + (size_t)blockIdx.x * 2 * C;
    *reinterpret_cast<float4*>(dst + cq * 4) = s;
    *reinterpret_cast<float4*>(dst + C + cq * 4) = q;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int v = threadIdx.x; v < 2 * C; v += BN_THREADS) {
    const volatile float* src = part + v;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;     // four independent chains, combined in a fixed order
    unsigned i = 0;
    for (; i + 4 <= gridDim.x; i += 4) {
      a0 += (double)src[(size_t)(i + 0) * 2 * C]; a1 += (double)src[(size_t)(i + 1) * 2 * C];
      a2 += (double)src[(size_t)(i + 2) * 2 * C]; a3 += (double)src[(size_t)(i + 3) * 2 * C];
    }
    for (; i < gridDim.x; ++i) a0 += (double)src[(size_t)i * 2 * C];
    const float r = (float)((a0 + a1) + (a2 + a3));
    if (v < C) out0[v] = r; else out1[v - C] = r;
  }
  if (threadIdx.x == 0) *ticket = 0u;                   // re-arm
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_stats_kernel(const T* __restrict__ x, const float* __restrict__ shift,
                                                              float* __restrict__ sum, float* __restrict__ sumsq,
                                                              float* __restrict__ part, unsigned* ticket, size_t npix,
                                                              int C) {
  const int CV = C / 4, cq = threadIdx.x % CV, pl = threadIdx.x / CV, PL = BN_THREADS / CV;
  float4 sf = make_float4(0.f, 0.f, 0.f, 0.f);
  if (shift != nullptr) sf = *reinterpret_cast<const float4*>(shift + cq * 4);
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = s;
  for (size_t p = (size_t)blockIdx.x * PL + pl; p < npix; p += (size_t)gridDim.x * PL) {
    float4 v = ld4<T>(x + p * C + cq * 4);
    v.x -= sf.x; v.y -= sf.y; v.z -= sf.z; v.w -= sf.w;
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    q.x += v.x * v.x; q.y += v.y * v.y; q.z += v.z * v.z; q.w += v.w * v.w;
  }
  bn_two_stage(s, q, C, part, ticket, sum, sumsq);
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_apply_kernel(const T* __restrict__ x, const float* __restrict__ sum,
                                                              const float* __restrict__ sumsq,
                                                              const float* __restrict__ shift, float count,
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
    const BnMoments mo = bn_moments(sum, sumsq, shift, count, c);
    a[k] = gamma[c] / sqrtf(mo.var + eps);
    b[k] = beta[c] - mo.mean * a[k];
  }
  for (size_t p = (size_t)blockIdx.x * PL + pl; p < npix; p += (size_t)gridDim.x * PL) {
    const float4 v = ld4<T>(x + p * C + cq * 4);
    st4<T>(y + p * C + cq * 4, make_float4(bn_act(a[0] * v.x + b[0], act), bn_act(a[1] * v.y + b[1], act),
                                           bn_act(a[2] * v.z + b[2], act), bn_act(a[3] * v.w + b[3], act)));
  }
}

// moving statistics: a launch of its own AFTER the apply pass, because `shift` may alias moving_mean
__global__ void bn_moving_kernel(const float* __restrict__ sum, const float* __restrict__ sumsq,
                                 const float* __restrict__ shift, float count, float* __restrict__ moving_mean,
                                 float* __restrict__ moving_var, float momentum, int C) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const BnMoments mo = bn_moments(sum, sumsq, shift, count, c);
  const float unbiased = count > 1.f ? mo.var * count / (count - 1.f) : mo.var;
  moving_mean[c] = moving_mean[c] * momentum + mo.mean * (1.f - momentum);
  moving_var[c] = moving_var[c] * momentum + unbiased * (1.f - momentum);
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_bwd_reduce_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                                   const T* __restrict__ dy,
                                                                   const float* __restrict__ sum,
                                                                   const float* __restrict__ sumsq,
                                                                   const float* __restrict__ shift, float count,
                                                                   float eps, int act, float* __restrict__ dsum,
                                                                   float* __restrict__ dsum_xhat,
                                                                   float* __restrict__ part, unsigned* ticket,
                                                                   size_t npix, int C) {
  const int CV = C / 4, cq = threadIdx.x % CV, pl = threadIdx.x / CV, PL = BN_THREADS / CV;
  float mean[4], inv[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const BnMoments mo = bn_moments(sum, sumsq, shift, count, cq * 4 + k);
    mean[k] = mo.mean;
    inv[k] = 1.0f / sqrtf(mo.var + eps);
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
  bn_two_stage(s, q, C, part, ticket, dsum, dsum_xhat);
}

template <typename T>
__global__ void __launch_bounds__(BN_THREADS) bn_bwd_apply_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                                                  const T* __restrict__ dy,
                                                                  const float* __restrict__ sum,
                                                                  const float* __restrict__ sumsq,
                                                                  const float* __restrict__ shift, float count,
                                                                  const float* __restrict__ gamma, float eps, int act,
                                                                  const float* __restrict__ dsum,
                                                                  const float* __restrict__ dsum_xhat,
                                                                  T* __restrict__ dx, size_t npix, int C) {
  const int CV = C / 4, cq = threadIdx.x % CV, pl = threadIdx.x / CV, PL = BN_THREADS / CV;
  float mean[4], inv[4], sc[4], m1[4], m2[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = cq * 4 + k;
    const BnMoments mo = bn_moments(sum, sumsq, shift, count, c);
    mean[k] = mo.mean;
    inv[k] = 1.0f / sqrtf(mo.var + eps);
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
  size_t g = (npix + (size_t)PL * 8 - 1) / ((size_t)PL * 8);      // >= 8 pixels per thread
  const size_t cap = (size_t)BN_MAX_GRID;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}
#define BN_CHECK_DIMS()                                                                                         \
  OF_REQUIRE(npix > 0 && C >= 4 && C % 4 == 0 && C <= 4 * BN_THREADS && BN_THREADS % (C / 4) == 0, OF_ERR_SHAPE, \
             "%s: needs C %% 4 == 0 and C / 4 dividing %d (C = %d)", __func__, BN_THREADS, C);                  \
  OF_REQUIRE(of_valid_dtype(dtype), OF_ERR_ARG, "%s: bad dtype %d", __func__, dtype)

extern "C" size_t of_bn_train_workspace_bytes(int C) { return BN_WS_HDR + (size_t)BN_MAX_GRID * 2 * C * sizeof(float); }

extern "C" int of_bn_train_stats(const void* x, const float* shift, float* sum, float* sumsq, size_t npix, int C,
                                 int dtype, void* workspace, void* stream) {
  OF_REQUIRE(x && sum && sumsq && workspace, OF_ERR_ARG, "of_bn_train_stats: null pointer");
  OF_REQUIRE(((uintptr_t)workspace & 15) == 0, OF_ERR_ARG, "of_bn_train_stats: workspace must be 16-byte aligned");
  BN_CHECK_DIMS();
  cudaStream_t st = (cudaStream_t)stream;
  unsigned* ticket = reinterpret_cast<unsigned*>(workspace);
  float* part = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + BN_WS_HDR);
  OF_CUDA(cudaMemsetAsync(ticket, 0, BN_WS_HDR, st));
  if (dtype == OF_F32)
    bn_stats_kernel<float><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const float*)x, shift, sum, sumsq, part, ticket, npix, C);
  else
    bn_stats_kernel<__nv_bfloat16><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const __nv_bfloat16*)x, shift, sum, sumsq, part,
                                                                            ticket, npix, C);
  return of_launch_check("bn_stats_kernel");
}

// ---- allocation-free variant (no counterpart in the reference) ----
// Same kernel as of_bn_train_stats, but the per-CTA partials and the ticket live in a file-scope device array, so a
// caller that cannot call cudaMalloc (CUDA graph capture, tight serving loops) passes only inputs and outputs.  The
// array is BN_STATIC_BYTES (2.31 MiB, sized for C <= BN_STATIC_MAX_C) and every GPU that loads the library holds it
// until exit.  In host code `g_bn_static_scratch` is only a stand-in for the symbol: its device address is looked up
// with cudaGetSymbolAddress once per device (with that device current) and cached.  All calls on a device share the
// array - calls that overlap on the GPU (two streams) race; the ticket header is cleared on the stream by every call.
#define BN_STATIC_MAX_C 256
#define BN_STATIC_BYTES (BN_WS_HDR + (size_t)BN_MAX_GRID * 2 * BN_STATIC_MAX_C * sizeof(float))
static __device__ __align__(256) unsigned char g_bn_static_scratch[BN_STATIC_BYTES];

#define BN_STATIC_MAX_DEVICES 64
static void* g_bn_static_addr[BN_STATIC_MAX_DEVICES];   // device address of the array, per device ordinal (0 = not looked up)

extern "C" int of_bn_train_stats_static(const void* x, const float* shift, float* sum, float* sumsq, size_t npix, int C,
                                        int dtype, void* stream) {
  OF_REQUIRE(x && sum && sumsq, OF_ERR_ARG, "of_bn_train_stats_static: null pointer");
  BN_CHECK_DIMS();
  OF_REQUIRE(C <= BN_STATIC_MAX_C, OF_ERR_SHAPE,
             "of_bn_train_stats_static: C = %d exceeds the static scratch (C <= %d, %zu bytes); use of_bn_train_stats "
             "with a caller workspace", C, BN_STATIC_MAX_C, (size_t)BN_STATIC_BYTES);
  int dev = -1;
  OF_CUDA(cudaGetDevice(&dev));
  OF_REQUIRE(dev >= 0 && dev < BN_STATIC_MAX_DEVICES, OF_ERR_ARG, "of_bn_train_stats_static: device ordinal %d", dev);
  void* base = __atomic_load_n(&g_bn_static_addr[dev], __ATOMIC_ACQUIRE);
  if (base == nullptr) {                                 // set-up, once per device, with that device current
    OF_CUDA(cudaGetSymbolAddress(&base, g_bn_static_scratch));
    OF_REQUIRE(base != nullptr && ((uintptr_t)base & 255) == 0, OF_ERR_STATE,
               "of_bn_train_stats_static: bad symbol address %p", base);
    __atomic_store_n(&g_bn_static_addr[dev], base, __ATOMIC_RELEASE);   // same value from any thread
  }
  cudaStream_t st = (cudaStream_t)stream;
  unsigned* ticket = reinterpret_cast<unsigned*>(base);
  float* part = reinterpret_cast<float*>(reinterpret_cast<char*>(base) + BN_WS_HDR);
  OF_CUDA(cudaMemsetAsync(ticket, 0, BN_WS_HDR, st));   // the array keeps the last call's contents: the ticket must start at 0
  if (dtype == OF_F32)
    bn_stats_kernel<float><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const float*)x, shift, sum, sumsq, part, ticket, npix, C);
  else
    bn_stats_kernel<__nv_bfloat16><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const __nv_bfloat16*)x, shift, sum, sumsq, part,
                                                                            ticket, npix, C);
  return of_launch_check("bn_stats_kernel");
}

extern "C" int of_bn_train_apply(const void* x, const float* sum, const float* sumsq, const float* shift, float count,
                                 const float* gamma, const float* beta, float eps, int act, void* y, float* moving_mean,
                                 float* moving_var, float momentum, size_t npix, int C, int dtype, void* stream) {
  OF_REQUIRE(x && sum && sumsq && gamma && beta && y, OF_ERR_ARG, "of_bn_train_apply: null pointer");
  OF_REQUIRE((moving_mean == nullptr) == (moving_var == nullptr), OF_ERR_ARG, "of_bn_train_apply: pass both moving statistics or neither");
  OF_REQUIRE(count >= 1.f, OF_ERR_ARG, "of_bn_train_apply: count %f", (double)count);
  BN_CHECK_DIMS();
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == OF_F32)
    bn_apply_kernel<float><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const float*)x, sum, sumsq, shift, count, gamma, beta, eps,
                                                                    act, (float*)y, moving_mean, moving_var, momentum, npix, C);
  else
    bn_apply_kernel<__nv_bfloat16><<<bn_grid(npix, C), BN_THREADS, 0, st>>>(
        (const __nv_bfloat16*)x, sum, sumsq, shift, count, gamma, beta, eps, act, (__nv_bfloat16*)y, moving_mean, moving_var,
        momentum, npix, C);
  if (int r = of_launch_check("bn_apply_kernel")) return r;
  if (moving_mean != nullptr) {
    bn_moving_kernel<<<(C + 127) / 128, 128, 0, st>>>(sum, sumsq, shift, count, moving_mean, moving_var, momentum, C);
    return of_launch_check("bn_moving_kernel");
  }
  return 0;
}

extern "C" int of_bn_train_bwd_reduce(const void* x, const void* y, const void* dy, const float* sum, const float* sumsq,
                                      const float* shift, float count, float eps, int act, float* dsum,
                                      float* dsum_xhat, size_t npix, int C, int dtype, void* workspace, void* stream) {
  OF_REQUIRE(x && y && dy && sum && sumsq && dsum && dsum_xhat && workspace, OF_ERR_ARG, "of_bn_train_bwd_reduce: null pointer");
  OF_REQUIRE(((uintptr_t)workspace & 15) == 0, OF_ERR_ARG, "of_bn_train_bwd_reduce: workspace must be 16-byte aligned");
  BN_CHECK_DIMS();
  cudaStream_t st = (cudaStream_t)stream;
  unsigned* ticket = reinterpret_cast<unsigned*>(workspace);
  float* part = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + BN_WS_HDR);
  OF_CUDA(cudaMemsetAsync(ticket, 0, BN_WS_HDR, st));
  if (dtype == OF_F32)
    bn_bwd_reduce_kernel<float><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const float*)x, (const float*)y, (const float*)dy, sum,
                                                                         sumsq, shift, count, eps, act, dsum, dsum_xhat, part,
                                                                         ticket, npix, C);
  else
    bn_bwd_reduce_kernel<__nv_bfloat16><<<bn_grid(npix, C), BN_THREADS, 0, st>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy, sum, sumsq, shift, count, eps, act, dsum,
        dsum_xhat, part, ticket, npix, C);
  return of_launch_check("bn_bwd_reduce_kernel");
}

extern "C" int of_bn_train_bwd_apply(const void* x, const void* y, const void* dy, const float* sum, const float* sumsq,
                                     const float* shift, float count, const float* gamma, float eps, int act,
                                     const float* dsum, const float* dsum_xhat, void* dx, size_t npix, int C, int dtype,
                                     void* stream) {
  OF_REQUIRE(x && y && dy && sum && sumsq && gamma && dsum && dsum_xhat && dx, OF_ERR_ARG, "of_bn_train_bwd_apply: null pointer");
  BN_CHECK_DIMS();
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == OF_F32)
    bn_bwd_apply_kernel<float><<<bn_grid(npix, C), BN_THREADS, 0, st>>>((const float*)x, (const float*)y, (const float*)dy, sum,
                                                                        sumsq, shift, count, gamma, eps, act, dsum, dsum_xhat,
                                                                        (float*)dx, npix, C);
  else
    bn_bwd_apply_kernel<__nv_bfloat16><<<bn_grid(npix, C), BN_THREADS, 0, st>>>(
        (const __nv_bfloat16*)x, (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy, sum, sumsq, shift, count, gamma, eps, act,
        dsum, dsum_xhat, (__nv_bfloat16*)dx, npix, C);
  return of_launch_check("bn_bwd_apply_kernel");
}
