// fp32 SIMT implicit-GEMM convolution: fprop / dgrad / wgrad.
//
// Keras Conv2D(padding='same') on NHWC / HWIO with TF's asymmetric SAME padding
// (model.py:12-15,104-114; resnet/models.py:17-30).  This is the exact-parity
// path (fp32 FMA accumulate, <= 1e-5 vs the oracle) and the kernel used for the
// layers that are not tensor-core shaped (Cin = 3 stem, Cout = 2 flow heads).
// The bf16 tcgen05 implicit-GEMM path lives in conv_tc.cu.
//
// One generic 64x64x16 register-tiled GEMM; the three passes differ only in how
// A(m,k) and B(k,n) are fetched:
//   fprop  M = pixels(out)  N = Cout  K = KH*KW*Cin   A = im2col(x)     B = w
//   dgrad  M = pixels(in)   N = Cin   K = KH*KW*Cout  A = col2im(dy)    B = w^T
//   wgrad  M = KH*KW*Cin    N = Cout  K = pixels(out) A = im2col(x)^T   B = dy   (split-K, atomics)
#include "common.cuh"

struct ConvGeom {
  int N, H, W, Cin, KH, KW, Cout, stride, Ho, Wo, pt, pl;
};

static inline void same_pad(int in, int k, int s, int* out, int* lo) {
  int o = (in + s - 1) / s;
  int total = (o - 1) * s + k - in;
  if (total < 0) total = 0;
  *out = o;
  *lo = total / 2;
}

template <typename T> __device__ __forceinline__ float ldg_as_f(const T* p);
template <> __device__ __forceinline__ float ldg_as_f<float>(const float* p) { return __ldg(p); }
template <> __device__ __forceinline__ float ldg_as_f<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __bfloat162float(*p);
}

// ---- operand fetchers -------------------------------------------------------
template <typename TX>
struct FpropA {  // A(m = out pixel, k = (kh,kw,ci))
  const TX* x; ConvGeom g;
  __device__ __forceinline__ float operator()(int m, int k) const {
    int wo = m % g.Wo; int r = m / g.Wo; int ho = r % g.Ho; int n = r / g.Ho;
    int ci = k % g.Cin; int tap = k / g.Cin; int kw = tap % g.KW; int kh = tap / g.KW;
    int h = ho * g.stride + kh - g.pt, w = wo * g.stride + kw - g.pl;
    if (h < 0 || h >= g.H || w < 0 || w >= g.W) return 0.f;
    return ldg_as_f<TX>(x + (((size_t)n * g.H + h) * g.W + w) * g.Cin + ci);
  }
};
struct FpropB {  // B(k, n = co) = w[k*Cout + co]
  const float* w; ConvGeom g;
  __device__ __forceinline__ float operator()(int k, int n) const { return __ldg(w + (size_t)k * g.Cout + n); }
};
template <typename TY>
struct DgradA {  // A(m = in pixel, k = (kh,kw,co))
  const TY* dy; ConvGeom g;
  __device__ __forceinline__ float operator()(int m, int k) const {
    int w = m % g.W; int r = m / g.W; int h = r % g.H; int n = r / g.H;
    int co = k % g.Cout; int tap = k / g.Cout; int kw = tap % g.KW; int kh = tap / g.KW;
    int hh = h + g.pt - kh, ww = w + g.pl - kw;
    if (hh < 0 || ww < 0 || hh % g.stride != 0 || ww % g.stride != 0) return 0.f;
    int ho = hh / g.stride, wo = ww / g.stride;
    if (ho >= g.Ho || wo >= g.Wo) return 0.f;
    return ldg_as_f<TY>(dy + (((size_t)n * g.Ho + ho) * g.Wo + wo) * g.Cout + co);
  }
};
struct DgradB {  // B(k = (kh,kw,co), n = ci) = w[(tap*Cin + ci)*Cout + co]
  const float* w; ConvGeom g;
  __device__ __forceinline__ float operator()(int k, int n) const {
    int co = k % g.Cout; int tap = k / g.Cout;
    return __ldg(w + ((size_t)tap * g.Cin + n) * g.Cout + co);
  }
};
template <typename TX>
struct WgradA {  // A(m = (kh,kw,ci), k = out pixel)
  const TX* x; ConvGeom g;
  __device__ __forceinline__ float operator()(int m, int k) const {
    int wo = k % g.Wo; int r = k / g.Wo; int ho = r % g.Ho; int n = r / g.Ho;
    int ci = m % g.Cin; int tap = m / g.Cin; int kw = tap % g.KW; int kh = tap / g.KW;
    int h = ho * g.stride + kh - g.pt, w = wo * g.stride + kw - g.pl;
    if (h < 0 || h >= g.H || w < 0 || w >= g.W) return 0.f;
    return ldg_as_f<TX>(x + (((size_t)n * g.H + h) * g.W + w) * g.Cin + ci);
  }
};
template <typename TY>
struct WgradB {  // B(k = out pixel, n = co)
  const TY* dy; ConvGeom g;
  __device__ __forceinline__ float operator()(int k, int n) const {
    return ldg_as_f<TY>(dy + (size_t)k * g.Cout + n);
  }
};

// ---- epilogues --------------------------------------------------------------
template <typename TO>
struct StoreEpi {  // out[m*ldn + n] = act(acc + bias[n])
  TO* out; const float* bias; int act;
  __device__ __forceinline__ void operator()(int m, int n, int ldn, float v) const {
    if (bias) v += __ldg(bias + n);
    if (act == OF_ACT_RELU) v = fmaxf(v, 0.f);
    else if (act == OF_ACT_LEAKY) v = v > 0.f ? v : 0.3f * v;   // Keras LeakyReLU() alpha = 0.3
    st_f<TO>(out + (size_t)m * ldn + n, v);
  }
};
struct AtomicEpi {  // split-K accumulation
  float* out;
  __device__ __forceinline__ void operator()(int m, int n, int ldn, float v) const {
    red_add(out + (size_t)m * ldn + n, v);
  }
};

#define BM 64
#define BN 64
#define BK 16
#define APAD 4

// A_ALONG_K: consecutive lanes fetch consecutive k (true) or consecutive m (false)
template <typename FA, typename FB, typename EPI, bool A_ALONG_K>
__global__ void __launch_bounds__(256) gemm_simt_kernel(FA fa, FB fb, EPI epi, int M, int N, int K, int k_chunk) {
  __shared__ __align__(16) float As[BK][BM + APAD];
  __shared__ __align__(16) float Bs[BK][BN + APAD];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int k_begin = blockIdx.z * k_chunk;
  const int k_end = min(K, k_begin + k_chunk);
  const int tx = tid & 15, ty = tid >> 4;
  float acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;

  for (int k0 = k_begin; k0 < k_end; k0 += BK) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      int kk, mm;
      if (A_ALONG_K) { kk = tid & 15; mm = (tid >> 4) + 16 * r; }
      else { mm = tid & 63; kk = (tid >> 6) + 4 * r; }
      int m = m0 + mm, k = k0 + kk;
      As[kk][mm] = (m < M && k < k_end) ? fa(m, k) : 0.f;
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      int nn = tid & 63, kk = (tid >> 6) + 4 * r;
      int n = n0 + nn, k = k0 + kk;
      Bs[kk][nn] = (n < N && k < k_end) ? fb(k, n) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n < N) epi(m, n, N, acc[i][j]);
    }
  }
}

// per-channel sum of dy over all pixels: db[co] += partial (db zeroed by the caller;
// gridDim.y CTAs share the pixel range)
template <typename TY>
__global__ void __launch_bounds__(256) colsum_kernel(const TY* __restrict__ dy, float* __restrict__ db,
                                                     size_t P, int C, int ld = 0) {
  if (ld == 0) ld = C;
  __shared__ float s[8][33];
  int c = blockIdx.x * 32 + (threadIdx.x & 31);
  int row = threadIdx.x >> 5;
  float acc = 0.f;
  if (c < C)
    for (size_t p = (size_t)blockIdx.y * 8 + row; p < P; p += (size_t)gridDim.y * 8) acc += ldg_as_f<TY>(dy + p * ld + c);
  s[row][threadIdx.x & 31] = acc;
  __syncthreads();
  if (row == 0 && c < C) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) t += s[r][threadIdx.x & 31];
    red_add(db + c, t);
  }
}

// coalesced variant for C % 4 == 0, C <= 1024: each thread owns 4 channels of every ppp-th pixel
// (a warp reads whole pixel rows, 8/16 bytes per lane), CTA-level reduction, one atomic per channel
template <typename TY>
__global__ void __launch_bounds__(256) colsum4_kernel(const TY* __restrict__ dy, float* __restrict__ db, size_t P,
                                                      int C, int ld) {
  const int cg = C >> 2, ppp = 256 / cg;
  const int g = threadIdx.x % cg, pr = threadIdx.x / cg;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (pr < ppp) {
    // four rows in flight per thread (independent loads, then the adds)
    const size_t stride = (size_t)gridDim.x * ppp;
    size_t p = (size_t)blockIdx.x * ppp + pr;
    for (; p + 3 * stride < P; p += 4 * stride) {
      const float4 v0 = ld4<TY>(dy + p * ld + g * 4), v1 = ld4<TY>(dy + (p + stride) * ld + g * 4);
      const float4 v2 = ld4<TY>(dy + (p + 2 * stride) * ld + g * 4), v3 = ld4<TY>(dy + (p + 3 * stride) * ld + g * 4);
      acc.x += (v0.x + v1.x) + (v2.x + v3.x); acc.y += (v0.y + v1.y) + (v2.y + v3.y);
      acc.z += (v0.z + v1.z) + (v2.z + v3.z); acc.w += (v0.w + v1.w) + (v2.w + v3.w);
    }
    for (; p < P; p += stride) {
      float4 v = ld4<TY>(dy + p * ld + g * 4);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  __shared__ float4 s[256];
  s[threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.x < cg) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < ppp; ++r) {
      float4 v = s[r * cg + threadIdx.x];
      t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
    }
    red_add(db + threadIdx.x * 4 + 0, t.x);
    red_add(db + threadIdx.x * 4 + 1, t.y);
    red_add(db + threadIdx.x * 4 + 2, t.z);
    red_add(db + threadIdx.x * 4 + 3, t.w);
  }
}

// wgrad for tiny Cout (the flow heads, Cout = 2): dw is only KH*KW*Cin x Cout, so a GEMM tile is
// wasted.  One thread per weight row (tap, ci) (+1 thread for the bias), the CTA stages an input halo
// tile and the dy tile in shared memory and every thread walks the tile's pixels; lanes = consecutive
// ci -> conflict-free, dy is a broadcast.  Stride 1 only.
#define SC_TH 8
#define SC_TW 32
template <typename TX, int COUT>
__global__ void __launch_bounds__(1024) smallcout_wgrad_kernel(const TX* __restrict__ x, const float* __restrict__ dy,
                                                               float* __restrict__ dw, float* __restrict__ db,
                                                               ConvGeom g) {
  extern __shared__ float sm[];
  const int HH = SC_TH + g.KH - 1, HW = SC_TW + g.KW - 1;
  float* xs = sm;                                   // [HH][HW][Cin]
  float* dys = sm + (size_t)HH * HW * g.Cin;        // [SC_TH*SC_TW][COUT]
  const int R = g.KH * g.KW * g.Cin;
  const int r = threadIdx.x;
  const int ci = r % g.Cin, tap = r / g.Cin;
  const int kh = tap / g.KW, kw = tap % g.KW;
  const int tiles_x = (g.W + SC_TW - 1) / SC_TW, tiles_y = (g.H + SC_TH - 1) / SC_TH;
  const long long ntiles = (long long)g.N * tiles_x * tiles_y;
  float acc[COUT];
#pragma unroll
  for (int c = 0; c < COUT; ++c) acc[c] = 0.f;
  for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
    int tx = (int)(t % tiles_x); long long q = t / tiles_x;
    int ty = (int)(q % tiles_y); int n = (int)(q / tiles_y);
    int y0 = ty * SC_TH, x0 = tx * SC_TW;
    __syncthreads();
    for (int e = threadIdx.x; e < HH * HW * g.Cin; e += blockDim.x) {
      int c = e % g.Cin; int p = e / g.Cin;
      int hx = p % HW, hy = p / HW;
      int yy = y0 + hy - g.pt, xx = x0 + hx - g.pl;
      float v = 0.f;
      if (yy >= 0 && yy < g.H && xx >= 0 && xx < g.W)
        v = ldg_as_f<TX>(x + (((size_t)n * g.H + yy) * g.W + xx) * g.Cin + c);
      xs[e] = v;
    }
    for (int e = threadIdx.x; e < SC_TH * SC_TW * COUT; e += blockDim.x) {
      int c = e % COUT; int p = e / COUT;
      int yy = y0 + p / SC_TW, xx = x0 + p % SC_TW;
      dys[e] = (yy < g.H && xx < g.W) ? __ldg(dy + (((size_t)n * g.H + yy) * g.W + xx) * COUT + c) : 0.f;
    }
    __syncthreads();
    if (r < R) {
      for (int py = 0; py < SC_TH; ++py) {
        const float* xrow = xs + ((size_t)(py + kh) * HW + kw) * g.Cin + ci;
        const float* drow = dys + (size_t)py * SC_TW * COUT;
#pragma unroll 8
        for (int px = 0; px < SC_TW; ++px) {
          float xv = xrow[(size_t)px * g.Cin];
#pragma unroll
          for (int c = 0; c < COUT; ++c) acc[c] = fmaf(xv, drow[px * COUT + c], acc[c]);
        }
      }
    } else if (r == R && db != nullptr) {
      for (int p = 0; p < SC_TH * SC_TW; ++p)
#pragma unroll
        for (int c = 0; c < COUT; ++c) acc[c] += dys[p * COUT + c];
    }
  }
  if (r < R) {
#pragma unroll
    for (int c = 0; c < COUT; ++c) red_add(dw + (size_t)r * COUT + c, acc[c]);
  } else if (r == R && db != nullptr) {
#pragma unroll
    for (int c = 0; c < COUT; ++c) red_add(db + c, acc[c]);
  }
}

static int make_geom(ConvGeom* g, int N, int H, int W, int Cin, int KH, int KW, int Cout, int stride) {
  OF_REQUIRE(N > 0 && H > 0 && W > 0 && Cin > 0 && Cout > 0 && KH > 0 && KW > 0, OF_ERR_SHAPE, "bad conv dims");
  OF_REQUIRE(stride == 1 || stride == 2, OF_ERR_UNSUPPORTED, "stride %d unsupported", stride);
  g->N = N; g->H = H; g->W = W; g->Cin = Cin; g->KH = KH; g->KW = KW; g->Cout = Cout; g->stride = stride;
  same_pad(H, KH, stride, &g->Ho, &g->pt);
  same_pad(W, KW, stride, &g->Wo, &g->pl);
  OF_REQUIRE((long long)N * H * W < (1ll << 31) && (long long)KH * KW * Cin * Cout < (1ll << 31), OF_ERR_SHAPE,
             "conv too large for 32-bit GEMM indices");
  return 0;
}

template <typename FA, typename FB, typename EPI, bool AK>
static int launch_gemm(FA fa, FB fb, EPI epi, int M, int N, int K, int splits, cudaStream_t st) {
  int k_chunk = ((K + splits - 1) / splits + BK - 1) / BK * BK;
  int z = (K + k_chunk - 1) / k_chunk;
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, z);
  gemm_simt_kernel<FA, FB, EPI, AK><<<grid, 256, 0, st>>>(fa, fb, epi, M, N, K, k_chunk);
  return of_launch_check("gemm_simt_kernel");
}

extern "C" {

int of_conv2d_fprop(const void* x, const float* w, const float* bias, void* y, int N, int H, int W, int Cin, int KH,
                    int KW, int Cout, int stride, int act, int in_dtype, int out_dtype, void* stream) {
  OF_REQUIRE(x && w && y, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(in_dtype) && of_valid_dtype(out_dtype), OF_ERR_ARG, "bad dtype");
  OF_REQUIRE(act >= OF_ACT_NONE && act <= OF_ACT_LEAKY, OF_ERR_ARG, "bad activation %d", act);
  ConvGeom g;
  if (int r = make_geom(&g, N, H, W, Cin, KH, KW, Cout, stride)) return r;
  int M = N * g.Ho * g.Wo, K = KH * KW * Cin;
  cudaStream_t st = (cudaStream_t)stream;
  FpropB fb{w, g};
#define GO(TX, TO)                                                                                  \
  return launch_gemm<FpropA<TX>, FpropB, StoreEpi<TO>, true>(FpropA<TX>{(const TX*)x, g}, fb,       \
                                                             StoreEpi<TO>{(TO*)y, bias, act}, M, Cout, K, 1, st)
  if (in_dtype == OF_F32 && out_dtype == OF_F32) GO(float, float);
  if (in_dtype == OF_F32 && out_dtype == OF_BF16) GO(float, __nv_bfloat16);
  if (in_dtype == OF_BF16 && out_dtype == OF_F32) GO(__nv_bfloat16, float);
  GO(__nv_bfloat16, __nv_bfloat16);
#undef GO
}

int of_conv2d_dgrad(const void* dy, const float* w, void* dx, int N, int H, int W, int Cin, int KH, int KW,
                    int Cout, int stride, int dy_dtype, int dx_dtype, void* stream) {
  OF_REQUIRE(dy && w && dx, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(dy_dtype) && of_valid_dtype(dx_dtype), OF_ERR_ARG, "bad dtype");
  ConvGeom g;
  if (int r = make_geom(&g, N, H, W, Cin, KH, KW, Cout, stride)) return r;
  int M = N * H * W, K = KH * KW * Cout;
  cudaStream_t st = (cudaStream_t)stream;
  DgradB fb{w, g};
#define GO(TY, TO)                                                                                 \
  return launch_gemm<DgradA<TY>, DgradB, StoreEpi<TO>, true>(DgradA<TY>{(const TY*)dy, g}, fb,     \
                                                             StoreEpi<TO>{(TO*)dx, nullptr, OF_ACT_NONE}, M, Cin, K, 1, st)
  if (dy_dtype == OF_F32 && dx_dtype == OF_F32) GO(float, float);
  if (dy_dtype == OF_F32 && dx_dtype == OF_BF16) GO(float, __nv_bfloat16);
  if (dy_dtype == OF_BF16 && dx_dtype == OF_F32) GO(__nv_bfloat16, float);
  GO(__nv_bfloat16, __nv_bfloat16);
#undef GO
}

int of_conv2d_wgrad(const void* x, const void* dy, float* dw, float* db, int N, int H, int W, int Cin, int KH,
                    int KW, int Cout, int stride, int x_dtype, int dy_dtype, void* stream) {
  OF_REQUIRE(x && dy && dw, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(x_dtype) && of_valid_dtype(dy_dtype), OF_ERR_ARG, "bad dtype");
  ConvGeom g;
  if (int r = make_geom(&g, N, H, W, Cin, KH, KW, Cout, stride)) return r;
  int M = KH * KW * Cin, K = N * g.Ho * g.Wo;
  cudaStream_t st = (cudaStream_t)stream;
  OF_CUDA(cudaMemsetAsync(dw, 0, (size_t)M * Cout * sizeof(float), st));
  if (Cout == 2 && stride == 1 && dy_dtype == OF_F32 && M + 1 <= 1024) {
    // flow heads: dedicated small-Cout kernel (also produces db)
    if (db) OF_CUDA(cudaMemsetAsync(db, 0, (size_t)Cout * sizeof(float), st));
    int threads = (M + 1 + 31) / 32 * 32;
    size_t smem = ((size_t)(SC_TH + KH - 1) * (SC_TW + KW - 1) * Cin + (size_t)SC_TH * SC_TW * 2) * sizeof(float);
    if (smem <= 200 * 1024) {
      long long ntiles = (long long)N * ((W + SC_TW - 1) / SC_TW) * ((H + SC_TH - 1) / SC_TH);
      int grid = (int)(ntiles < 2 * OF_NUM_SMS ? ntiles : 2 * OF_NUM_SMS);
      if (x_dtype == OF_F32) {
        OF_CUDA(cudaFuncSetAttribute(smallcout_wgrad_kernel<float, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smallcout_wgrad_kernel<float, 2><<<grid, threads, smem, st>>>((const float*)x, (const float*)dy, dw, db, g);
      } else {
        OF_CUDA(cudaFuncSetAttribute(smallcout_wgrad_kernel<__nv_bfloat16, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smallcout_wgrad_kernel<__nv_bfloat16, 2><<<grid, threads, smem, st>>>((const __nv_bfloat16*)x, (const float*)dy, dw, db, g);
      }
      return of_launch_check("smallcout_wgrad_kernel");
    }
  }
  // split the pixel reduction so the grid covers the machine (~4 CTAs per SM)
  int tiles = ((M + BM - 1) / BM) * ((Cout + BN - 1) / BN);
  int splits = (OF_NUM_SMS * 4 + tiles - 1) / tiles;
  int max_splits = (K + 4 * BK - 1) / (4 * BK);
  if (splits > max_splits) splits = max_splits;
  if (splits < 1) splits = 1;
  int rc;
#define GO(TX, TY)                                                                                  \
  rc = launch_gemm<WgradA<TX>, WgradB<TY>, AtomicEpi, false>(WgradA<TX>{(const TX*)x, g},           \
                                                             WgradB<TY>{(const TY*)dy, g}, AtomicEpi{dw}, M, Cout, K, splits, st)
  if (x_dtype == OF_F32 && dy_dtype == OF_F32) { GO(float, float); }
  else if (x_dtype == OF_F32 && dy_dtype == OF_BF16) { GO(float, __nv_bfloat16); }
  else if (x_dtype == OF_BF16 && dy_dtype == OF_F32) { GO(__nv_bfloat16, float); }
  else { GO(__nv_bfloat16, __nv_bfloat16); }
#undef GO
  if (rc) return rc;
  if (db != nullptr) return of_colsum(dy, db, (size_t)K, Cout, Cout, dy_dtype, stream);
  return 0;
}

int of_colsum(const void* dy, float* db, size_t npix, int C, int dy_stride, int dtype, void* stream) {
  OF_REQUIRE(dy && db, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(of_valid_dtype(dtype) && C > 0 && dy_stride >= C && npix > 0, OF_ERR_ARG, "bad colsum arguments");
  cudaStream_t st = (cudaStream_t)stream;
  OF_CUDA(cudaMemsetAsync(db, 0, (size_t)C * sizeof(float), st));
  if (C % 4 == 0 && C <= 1024 && dy_stride % 4 == 0) {
    int ppp = 256 / (C / 4);
    size_t blocks = (npix + ppp - 1) / ppp;
    if (blocks > (size_t)OF_NUM_SMS * 8) blocks = (size_t)OF_NUM_SMS * 8;
    // every CTA ends with C same-address atomics: give each thread >= 16 pixels so they stay a small share
    size_t cap = npix / ((size_t)ppp * 16);
    if (cap < 1) cap = 1;
    if (blocks > cap) blocks = cap;
    if (dtype == OF_F32)
      colsum4_kernel<float><<<(int)blocks, 256, 0, st>>>((const float*)dy, db, npix, C, dy_stride);
    else
      colsum4_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, st>>>((const __nv_bfloat16*)dy, db, npix, C, dy_stride);
    return of_launch_check("colsum4_kernel");
  }
  int gx = (C + 31) / 32;
  int gy = (OF_NUM_SMS * 4 + gx - 1) / gx;
  if ((size_t)gy * 8 > npix) gy = (int)((npix + 7) / 8);
  dim3 cg(gx, gy);
  if (dtype == OF_F32)
    colsum_kernel<float><<<cg, 256, 0, st>>>((const float*)dy, db, npix, C, dy_stride);
  else
    colsum_kernel<__nv_bfloat16><<<cg, 256, 0, st>>>((const __nv_bfloat16*)dy, db, npix, C, dy_stride);
  return of_launch_check("colsum_kernel");
}

}  // extern "C"
