// K8: cost volume forward/backward (model.py:29-42 create_cost_volume).
//   cv[b,y,x,i*D+j] = sum_c f1[b,y,x,c] * f2[b,y+i-m,x+j-m,c],  D = 2m+1, zero padding,
//   no normalisation.  Not a GEMM: 2*D*D*C flops per pixel against (2C + D*D)
//   elements of traffic -> L1/L2-resident stencil; all gradients in gather
//   form (no atomics, deterministic).
#include "common.cuh"

#define CV_MAX_D 9   // max_disp <= 4

// forward: one thread = one pixel x one displacement row i (D outputs).
template <typename T, int D>
__global__ void __launch_bounds__(256) costvol_fwd_kernel(const T* __restrict__ f1, const T* __restrict__ f2,
                                                          T* __restrict__ cv, int B, int H, int W, int C,
                                                          int cv_stride) {
  constexpr int M = D / 2;
  // 2D tiling: 8 rows x 32 cols of pixels per CTA pass, all D displacement rows
  const int tiles_x = (W + 31) / 32, tiles_y = (H + 7) / 8;
  const long long ntiles = (long long)B * tiles_x * tiles_y * D;
  for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
    int i = (int)(t % D);
    long long r = t / D;
    int tx = (int)(r % tiles_x); r /= tiles_x;
    int ty = (int)(r % tiles_y);
    int b = (int)(r / tiles_y);
    int y = ty * 8 + (threadIdx.x >> 5);
    int x = tx * 32 + (threadIdx.x & 31);
    if (y >= H || x >= W) continue;
    int y2 = y + i - M;
    float acc[D];
#pragma unroll
    for (int j = 0; j < D; ++j) acc[j] = 0.f;
    size_t pix = ((size_t)b * H + y) * W + x;
    if (y2 >= 0 && y2 < H) {
      const T* p1 = f1 + pix * C;
      const T* row2 = f2 + ((size_t)b * H + y2) * W * C;
      for (int c = 0; c < C; c += 4) {
        float4 a = ld4<T>(p1 + c);
#pragma unroll
        for (int j = 0; j < D; ++j) {
          int x2 = x + j - M;
          if (x2 >= 0 && x2 < W) {
            float4 q = ld4<T>(row2 + (size_t)x2 * C + c);
            acc[j] += a.x * q.x + a.y * q.y + a.z * q.z + a.w * q.w;
          }
        }
      }
    }
    T* o = cv + pix * cv_stride + i * D;
#pragma unroll
    for (int j = 0; j < D; ++j) st_f<T>(o + j, acc[j]);
  }
}

// backward: one thread = one pixel x 4 channels; gathers both d_f1 and d_f2.
//   d_f1[p,c] = sum_ij d_cv[p,ij] * f2[p + (i,j) - m, c]
//   d_f2[q,c] = sum_ij d_cv[q - (i,j) + m, ij] * f1[q - (i,j) + m, c]
template <typename T, int D>
__global__ void __launch_bounds__(256) costvol_bwd_kernel(const T* __restrict__ f1, const T* __restrict__ f2,
                                                          const T* __restrict__ d_cv, float* __restrict__ d_f1,
                                                          float* __restrict__ d_f2, int B, int H, int W, int C,
                                                          int cv_stride) {
  constexpr int M = D / 2;
  const int CV4 = C / 4;
  size_t total = (size_t)B * H * W * CV4;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(e % CV4) * 4;
    size_t pix = e / CV4;
    int x = (int)(pix % W);
    size_t r = pix / W;
    int y = (int)(r % H);
    int b = (int)(r / H);
    const size_t img = (size_t)b * H * W;
    float4 a1 = make_float4(0.f, 0.f, 0.f, 0.f), a2 = a1;
    const T* g_here = d_cv + pix * cv_stride;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      int yf = y + i - M;   // f2 row for d_f1
      int yb = y - i + M;   // f1 / d_cv row for d_f2
#pragma unroll
      for (int j = 0; j < D; ++j) {
        int xf = x + j - M, xb = x - j + M;
        if (yf >= 0 && yf < H && xf >= 0 && xf < W) {
          float g = ld_f<T>(g_here + i * D + j);
          float4 q = ld4<T>(f2 + (img + (size_t)yf * W + xf) * C + c);
          a1.x += g * q.x; a1.y += g * q.y; a1.z += g * q.z; a1.w += g * q.w;
        }
        if (yb >= 0 && yb < H && xb >= 0 && xb < W) {
          size_t pb = img + (size_t)yb * W + xb;
          float g = ld_f<T>(d_cv + pb * cv_stride + i * D + j);
          float4 q = ld4<T>(f1 + pb * C + c);
          a2.x += g * q.x; a2.y += g * q.y; a2.z += g * q.z; a2.w += g * q.w;
        }
      }
    }
    *reinterpret_cast<float4*>(d_f1 + pix * C + c) = a1;
    *reinterpret_cast<float4*>(d_f2 + pix * C + c) = a2;
  }
}

// tuned max_disp = 3 path (costvol7.cu)
template <typename T>
int cv7_fwd(const void* f1, const void* f2, void* cv, int B, int H, int W, int C, int cv_stride, cudaStream_t st);
template <typename T>
int cv7_bwd(const void* f1, const void* f2, const void* d_cv, float* d_f1, float* d_f2, int B, int H, int W, int C,
            int cv_stride, cudaStream_t st);
int cv7_bwd_bf16(const void* f1, const void* f2, const void* d_cv, void* d_f1, void* d_f2, const void* d_f1_add, int B,
                 int H, int W, int C, int cv_stride, cudaStream_t st);
// tensor-core path (costvol_tc.cu): 1 = handled, 0 = not applicable
int cvtc_fwd_try(const void* f1, const void* f2, void* cv, int B, int H, int W, int C, int cv_stride, cudaStream_t st);
int cvtc_bwd_try(const void* f1, const void* f2, const void* d_cv, void* d_f1, void* d_f2, const void* d_f1_add, int B,
                 int H, int W, int C, int cv_stride, cudaStream_t st);
template <typename T> static bool cv7_ok(int C, int D) { return D == 7 && (C * (int)sizeof(T)) % 16 == 0; }

template <typename T>
static int cv_fwd(const void* f1, const void* f2, void* cv, int B, int H, int W, int C, int m, int cv_stride,
                  cudaStream_t st) {
  int D = 2 * m + 1;
  if (cv7_ok<T>(C, D)) return cv7_fwd<T>(f1, f2, cv, B, H, W, C, cv_stride, st);
  long long ntiles = (long long)B * ((W + 31) / 32) * ((H + 7) / 8) * D;
  int grid = (int)(ntiles < OF_NUM_SMS * 8 ? ntiles : OF_NUM_SMS * 8);
#define L(DD) costvol_fwd_kernel<T, DD><<<grid, 256, 0, st>>>((const T*)f1, (const T*)f2, (T*)cv, B, H, W, C, cv_stride)
  switch (D) {
    case 1: L(1); break;
    case 3: L(3); break;
    case 5: L(5); break;
    case 7: L(7); break;
    case 9: L(9); break;
  }
#undef L
  return of_launch_check("costvol_fwd_kernel");
}

template <typename T>
static int cv_bwd(const void* f1, const void* f2, const void* d_cv, float* d_f1, float* d_f2, int B, int H, int W,
                  int C, int m, int cv_stride, cudaStream_t st) {
  int D = 2 * m + 1;
  if (cv7_ok<T>(C, D)) return cv7_bwd<T>(f1, f2, d_cv, d_f1, d_f2, B, H, W, C, cv_stride, st);
  size_t total = (size_t)B * H * W * (C / 4);
  size_t blocks = (total + 255) / 256;
  int grid = (int)(blocks < (size_t)OF_NUM_SMS * 16 ? blocks : (size_t)OF_NUM_SMS * 16);
#define L(DD)                                                                                             \
  costvol_bwd_kernel<T, DD><<<grid, 256, 0, st>>>((const T*)f1, (const T*)f2, (const T*)d_cv, d_f1, d_f2, \
                                                  B, H, W, C, cv_stride)
  switch (D) {
    case 1: L(1); break;
    case 3: L(3); break;
    case 5: L(5); break;
    case 7: L(7); break;
    case 9: L(9); break;
  }
#undef L
  return of_launch_check("costvol_bwd_kernel");
}

static int cv_check(int B, int H, int W, int C, int m, int cv_stride, int dtype) {
  OF_REQUIRE(of_valid_dtype(dtype), OF_ERR_ARG, "bad dtype %d", dtype);
  OF_REQUIRE(B > 0 && H > 0 && W > 0 && C > 0, OF_ERR_SHAPE, "bad dims");
  OF_REQUIRE(C % 4 == 0, OF_ERR_SHAPE, "cost volume needs C %% 4 == 0 (got %d)", C);
  OF_REQUIRE(m >= 0 && 2 * m + 1 <= CV_MAX_D, OF_ERR_ARG, "max_disp %d unsupported", m);
  OF_REQUIRE(cv_stride >= (2 * m + 1) * (2 * m + 1), OF_ERR_SHAPE, "cv_stride %d < D*D", cv_stride);
  return 0;
}

extern "C" {

int of_cost_volume_fwd(const void* f1, const void* f2, void* cv, int B, int H, int W, int C, int max_disp,
                       int cv_stride, int dtype, void* stream) {
  OF_REQUIRE(f1 && f2 && cv, OF_ERR_ARG, "null pointer argument");
  if (int r = cv_check(B, H, W, C, max_disp, cv_stride, dtype)) return r;
  if (dtype == OF_F32) return cv_fwd<float>(f1, f2, cv, B, H, W, C, max_disp, cv_stride, (cudaStream_t)stream);
  if (max_disp == 3)
    if (int r = cvtc_fwd_try(f1, f2, cv, B, H, W, C, cv_stride, (cudaStream_t)stream)) return r == 1 ? 0 : r;
  return cv_fwd<__nv_bfloat16>(f1, f2, cv, B, H, W, C, max_disp, cv_stride, (cudaStream_t)stream);
}

int of_cost_volume_bwd(const void* f1, const void* f2, const void* d_cv, float* d_f1, float* d_f2, int B, int H,
                       int W, int C, int max_disp, int cv_stride, int dtype, void* stream) {
  OF_REQUIRE(f1 && f2 && d_cv && d_f1 && d_f2, OF_ERR_ARG, "null pointer argument");
  if (int r = cv_check(B, H, W, C, max_disp, cv_stride, dtype)) return r;
  if (dtype == OF_F32)
    return cv_bwd<float>(f1, f2, d_cv, d_f1, d_f2, B, H, W, C, max_disp, cv_stride, (cudaStream_t)stream);
  return cv_bwd<__nv_bfloat16>(f1, f2, d_cv, d_f1, d_f2, B, H, W, C, max_disp, cv_stride, (cudaStream_t)stream);
}

int of_cost_volume_bwd_bf16(const void* f1, const void* f2, const void* d_cv, void* d_f1, void* d_f2,
                            const void* d_f1_add, int B, int H, int W, int C, int max_disp, int cv_stride,
                            void* stream) {
  OF_REQUIRE(f1 && f2 && d_cv && d_f1 && d_f2, OF_ERR_ARG, "null pointer argument");
  if (int r = cv_check(B, H, W, C, max_disp, cv_stride, OF_BF16)) return r;
  OF_REQUIRE(cv7_ok<__nv_bfloat16>(C, 2 * max_disp + 1), OF_ERR_UNSUPPORTED,
             "bf16-output cost-volume backward needs max_disp == 3 and C %% 8 == 0 (got %d, %d)", max_disp, C);
  if (int r = cvtc_bwd_try(f1, f2, d_cv, d_f1, d_f2, d_f1_add, B, H, W, C, cv_stride, (cudaStream_t)stream))
    return r == 1 ? 0 : r;
  return cv7_bwd_bf16(f1, f2, d_cv, d_f1, d_f2, d_f1_add, B, H, W, C, cv_stride, (cudaStream_t)stream);
}

}  // extern "C"
