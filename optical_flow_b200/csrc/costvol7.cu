// K8, tuned path for max_disp = 3 (the reference's value, model.py:119): shared-memory tiled,
// register-blocked cost volume forward / backward.
//
// A 4 x 64 pixel tile plus its 3-pixel halo is staged in shared memory once per channel chunk
// and every value is reused from registers:
//   fwd  : a thread owns 4 adjacent pixels x 7 column displacements of one displacement row
//          (28 accumulators); per channel word it issues 4 + 10 conflict-free LDS for 56 FMAs.
//          The tile is stored channel-plane major with the columns de-interleaved by 4 so that
//          the 32 lanes (consecutive pixel quads) read consecutive words.
//   bwd  : a warp owns a pixel quad, its lanes the channels; d_cv rows are 16-byte broadcasts,
//          the feature halo is read once per (row displacement, column) and reused by the 4
//          pixels: 8 LDS.128 + 10 LDS.32 for 56 FMAs.  Gather form, no atomics, deterministic.
// Storage type T = float (exact-parity mode: one channel per 32-bit word) or bf16 (two channels
// per word, unpacked with shifts); accumulation is fp32.
#include "common.cuh"

#define CV_TR 4
#define CV_TC 64
#define CV_D 7
#define CV_HR (CV_TR + 6)          // halo rows
#define CV_HC (CV_TC + 6)          // halo cols (70)

template <typename T> struct WordTraits;
template <> struct WordTraits<float> {
  static constexpr int PER_WORD = 1;
  __device__ static __forceinline__ void unpack(uint32_t w, float& lo, float& hi) { lo = __uint_as_float(w); hi = 0.f; }
};
template <> struct WordTraits<__nv_bfloat16> {
  static constexpr int PER_WORD = 2;
  __device__ static __forceinline__ void unpack(uint32_t w, float& lo, float& hi) {
    lo = __uint_as_float(w << 16);
    hi = __uint_as_float(w & 0xffff0000u);
  }
};

// ------------------------------------------------------------------------------ forward --
#define F2_RS 80                     // row stride (words) of a de-interleaved halo row (72 used)
#define F2_PS (CV_HR * F2_RS + 1)    // plane stride, == 1 mod 32
#define F1_RS 80
#define F1_PS (CV_TR * F1_RS + 1)
#define CVF_THREADS 448              // 64 quads x 7 displacement rows
#define CVF_WORDS 32                 // channel words per chunk

__device__ __forceinline__ int deint_pos(int col, int quarter) { return (col & 3) * quarter + (col >> 2); }

template <typename T>
__global__ void __launch_bounds__(CVF_THREADS, 1)
cv7_fwd_kernel(const T* __restrict__ f1, const T* __restrict__ f2, T* __restrict__ cv, int B, int H, int W, int C,
               int cv_stride) {
  using WT = WordTraits<T>;
  constexpr int PW = WT::PER_WORD;
  extern __shared__ uint32_t smw[];
  uint32_t* f2s = smw;                              // [32][CV_HR][F2_RS]
  uint32_t* f1s = smw + (size_t)CVF_WORDS * F2_PS;  // [32][CV_TR][F1_RS]
  const int tiles_x = (W + CV_TC - 1) / CV_TC, tiles_y = (H + CV_TR - 1) / CV_TR;
  const long long ntiles = (long long)B * tiles_x * tiles_y;
  const int t = threadIdx.x;
  const int i = t >> 6, quad = t & 63, r = quad >> 4, qx = quad & 15;
  const int words_total = (C + PW - 1) / PW;
  const uint32_t* f1w = reinterpret_cast<const uint32_t*>(f1);
  const uint32_t* f2w = reinterpret_cast<const uint32_t*>(f2);

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int tx = (int)(tile % tiles_x); long long q = tile / tiles_x;
    int ty = (int)(q % tiles_y); int b = (int)(q / tiles_y);
    const int y0 = ty * CV_TR, x0 = tx * CV_TC;
    float acc[4][CV_D];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
      for (int j = 0; j < CV_D; ++j) acc[k][j] = 0.f;

    for (int w0 = 0; w0 < words_total; w0 += CVF_WORDS) {
      __syncthreads();
      // ---- stage the f2 halo and the f1 tile: 16-byte global loads (4 words), transposing stores
      for (int e = t; e < CV_HR * CV_HC * (CVF_WORDS / 4); e += CVF_THREADS) {
        int wq = e & 7; int p = e >> 3;
        int hx = p % CV_HC, hy = p / CV_HC;
        int yy = y0 + hy - 3, xx = x0 + hx - 3;
        uint4 v = make_uint4(0, 0, 0, 0);
        int wbase = w0 + wq * 4;
        if (yy >= 0 && yy < H && xx >= 0 && xx < W && wbase < words_total)
          v = __ldg(reinterpret_cast<const uint4*>(f2w + (((size_t)b * H + yy) * W + xx) * words_total + wbase));
        uint32_t* dst = f2s + (size_t)(wq * 4) * F2_PS + hy * F2_RS + deint_pos(hx, 18);
        dst[0] = v.x; dst[F2_PS] = v.y; dst[2 * F2_PS] = v.z; dst[3 * F2_PS] = v.w;
      }
      for (int e = t; e < CV_TR * CV_TC * (CVF_WORDS / 4); e += CVF_THREADS) {
        int wq = e & 7; int p = e >> 3;
        int cx = p % CV_TC, cy = p / CV_TC;
        int yy = y0 + cy, xx = x0 + cx;
        uint4 v = make_uint4(0, 0, 0, 0);
        int wbase = w0 + wq * 4;
        if (yy < H && xx < W && wbase < words_total)
          v = __ldg(reinterpret_cast<const uint4*>(f1w + (((size_t)b * H + yy) * W + xx) * words_total + wbase));
        uint32_t* dst = f1s + (size_t)(wq * 4) * F1_PS + cy * F1_RS + deint_pos(cx, 16);
        dst[0] = v.x; dst[F1_PS] = v.y; dst[2 * F1_PS] = v.z; dst[3 * F1_PS] = v.w;
      }
      __syncthreads();
      // ---- 28 accumulators per thread, 14 LDS per channel word
      const uint32_t* a_row = f1s + r * F1_RS + qx;
      const uint32_t* b_row = f2s + (r + i) * F2_RS + qx;
#pragma unroll 2
      for (int cw = 0; cw < CVF_WORDS; ++cw) {
        float alo[4], ahi[4], blo[10], bhi[10];
#pragma unroll
        for (int k = 0; k < 4; ++k) WT::unpack(a_row[(size_t)cw * F1_PS + k * 16], alo[k], ahi[k]);
#pragma unroll
        for (int kj = 0; kj < 10; ++kj)
          WT::unpack(b_row[(size_t)cw * F2_PS + (kj & 3) * 18 + (kj >> 2)], blo[kj], bhi[kj]);
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
          for (int j = 0; j < CV_D; ++j) {
            acc[k][j] = fmaf(alo[k], blo[k + j], acc[k][j]);
            if (PW == 2) acc[k][j] = fmaf(ahi[k], bhi[k + j], acc[k][j]);
          }
      }
    }
    // ---- write cv[p][i*7 + j]
    int yy = y0 + r;
    if (yy < H) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        int xx = x0 + qx * 4 + k;
        if (xx < W) {
          T* o = cv + (((size_t)b * H + yy) * W + xx) * cv_stride + i * CV_D;
#pragma unroll
          for (int j = 0; j < CV_D; ++j) st_f<T>(o + j, acc[k][j]);
        }
      }
    }
  }
}

// ----------------------------------------------------------------------------- backward --
#define CVB_THREADS 512
#define CVB_WORDS 32
// FROM_F2 = true : d_f1[p,c] = sum_ij g[p, ij]              * f2[p + (i-3, j-3), c]
// FROM_F2 = false: d_f2[q,c] = sum_ij g[q - (i-3, j-3), ij] * f1[q - (i-3, j-3), c]
// OUT_BF16: the result (+ the optional `add` operand, a gradient that reaches the same tensor by another
// path) is stored as bf16 - lets the caller write the encoder's incoming gradient directly.
template <typename T, bool FROM_F2, bool OUT_BF16 = false>
__global__ void __launch_bounds__(CVB_THREADS, 1)
cv7_bwd_kernel(const T* __restrict__ feat, const T* __restrict__ d_cv, void* __restrict__ d_out_v, int B, int H,
               int W, int C, int cv_stride, const __nv_bfloat16* __restrict__ add = nullptr) {
  float* d_out = reinterpret_cast<float*>(d_out_v);
  using WT = WordTraits<T>;
  constexpr int PW = WT::PER_WORD;
  extern __shared__ uint32_t smw[];
  uint32_t* fs = smw;                                                        // [CV_HR][CV_HC][32] words
  float* gs = reinterpret_cast<float*>(smw + CV_HR * CV_HC * CVB_WORDS);     // [7][CV_TR][CV_TC][8]
  const int tiles_x = (W + CV_TC - 1) / CV_TC, tiles_y = (H + CV_TR - 1) / CV_TR;
  const long long ntiles = (long long)B * tiles_x * tiles_y;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int words_total = (C + PW - 1) / PW;
  const uint32_t* fw = reinterpret_cast<const uint32_t*>(feat);

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    int tx = (int)(tile % tiles_x); long long q = tile / tiles_x;
    int ty = (int)(q % tiles_y); int b = (int)(q / tiles_y);
    const int y0 = ty * CV_TR, x0 = tx * CV_TC;
    __syncthreads();
    // ---- d_cv tile -> gs[i][r][x][j] (slot 7 unused).  FROM_F2: the pixel's own 49 values;
    //      otherwise gathered from the source pixel q - (i-3, j-3).
    for (int e = t; e < CV_D * CV_TR * CV_TC * CV_D; e += CVB_THREADS) {
      int j = e % CV_D; int rest = e / CV_D;
      int xl = rest % CV_TC; rest /= CV_TC;
      int rr = rest % CV_TR; int i = rest / CV_TR;
      int py, px;
      if (FROM_F2) { py = y0 + rr; px = x0 + xl; }
      else { py = y0 + rr - i + 3; px = x0 + xl - j + 3; }
      float v = 0.f;
      if (py >= 0 && py < H && px >= 0 && px < W && (FROM_F2 || (y0 + rr < H && x0 + xl < W)))
        v = ld_f<T>(d_cv + (((size_t)b * H + py) * W + px) * cv_stride + i * CV_D + j);
      gs[(((size_t)i * CV_TR + rr) * CV_TC + xl) * 8 + j] = v;
    }
    for (int w0 = 0; w0 < words_total; w0 += CVB_WORDS) {
      if (w0 > 0) __syncthreads();
      for (int e = t; e < CV_HR * CV_HC * (CVB_WORDS / 4); e += CVB_THREADS) {
        int wq = e & 7; int p = e >> 3;
        int hx = p % CV_HC, hy = p / CV_HC;
        int yy = y0 + hy - 3, xx = x0 + hx - 3;
        uint4 v = make_uint4(0, 0, 0, 0);
        int wbase = w0 + wq * 4;
        if (yy >= 0 && yy < H && xx >= 0 && xx < W && wbase < words_total)
          v = __ldg(reinterpret_cast<const uint4*>(fw + (((size_t)b * H + yy) * W + xx) * words_total + wbase));
        *reinterpret_cast<uint4*>(fs + (size_t)p * CVB_WORDS + wq * 4) = v;
      }
      __syncthreads();
      // ---- each warp: 4 pixel quads; lane = channel word
      for (int quad = warp; quad < CV_TR * CV_TC / 4; quad += CVB_THREADS / 32) {
        const int r = quad >> 4, qx = quad & 15;
        float alo[4] = {0.f, 0.f, 0.f, 0.f}, ahi[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < CV_D; ++i) {
          float g[4][8];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float4* gp = reinterpret_cast<const float4*>(gs + (((size_t)i * CV_TR + r) * CV_TC + qx * 4 + k) * 8);
            float4 a = gp[0], c4 = gp[1];
            g[k][0] = a.x; g[k][1] = a.y; g[k][2] = a.z; g[k][3] = a.w;
            g[k][4] = c4.x; g[k][5] = c4.y; g[k][6] = c4.z; g[k][7] = c4.w;
          }
          const int hr = FROM_F2 ? (r + i) : (r - i + 6);
          const uint32_t* frow = fs + ((size_t)hr * CV_HC + qx * 4) * CVB_WORDS + lane;
#pragma unroll
          for (int kj = 0; kj < 10; ++kj) {
            float lo, hi;
            WT::unpack(frow[(size_t)kj * CVB_WORDS], lo, hi);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int j = FROM_F2 ? (kj - k) : (k + 6 - kj);
              if (j >= 0 && j < CV_D) {
                alo[k] = fmaf(g[k][j], lo, alo[k]);
                if (PW == 2) ahi[k] = fmaf(g[k][j], hi, ahi[k]);
              }
            }
          }
        }
        int yy = y0 + r;
        int wi = w0 + lane;
        if (yy < H && wi < words_total) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            int xx = x0 + qx * 4 + k;
            if (xx < W) {
              const size_t e = (((size_t)b * H + yy) * W + xx) * C + (size_t)wi * PW;
              if (OUT_BF16 && PW == 2) {
                float lo = alo[k], hi = ahi[k];
                if (add != nullptr) {
                  uint32_t a = __ldg(reinterpret_cast<const uint32_t*>(add + e));
                  lo += __uint_as_float(a << 16);
                  hi += __uint_as_float(a & 0xffff0000u);
                }
                __nv_bfloat162 pk = __floats2bfloat162_rn(lo, hi);
                *reinterpret_cast<__nv_bfloat162*>(reinterpret_cast<__nv_bfloat16*>(d_out_v) + e) = pk;
              } else {
                float* o = d_out + e;
                if (PW == 2) *reinterpret_cast<float2*>(o) = make_float2(alo[k], ahi[k]);
                else *o = alo[k];
              }
            }
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------- host --
template <typename T>
int cv7_fwd(const void* f1, const void* f2, void* cv, int B, int H, int W, int C, int cv_stride, cudaStream_t st) {
  size_t smem = ((size_t)CVF_WORDS * F2_PS + (size_t)CVF_WORDS * F1_PS) * sizeof(uint32_t);
  static bool set = false;
  if (!set) {
    OF_CUDA(cudaFuncSetAttribute(cv7_fwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    OF_CUDA(cudaFuncSetAttribute(cv7_fwd_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    set = true;
  }
  long long ntiles = (long long)B * ((W + CV_TC - 1) / CV_TC) * ((H + CV_TR - 1) / CV_TR);
  int grid = (int)(ntiles < OF_NUM_SMS ? ntiles : OF_NUM_SMS);
  cv7_fwd_kernel<T><<<grid, CVF_THREADS, smem, st>>>((const T*)f1, (const T*)f2, (T*)cv, B, H, W, C, cv_stride);
  return of_launch_check("cv7_fwd_kernel");
}

template <typename T>
int cv7_bwd(const void* f1, const void* f2, const void* d_cv, float* d_f1, float* d_f2, int B, int H, int W, int C,
            int cv_stride, cudaStream_t st) {
  size_t smem = (size_t)CV_HR * CV_HC * CVB_WORDS * sizeof(uint32_t) + (size_t)CV_D * CV_TR * CV_TC * 8 * sizeof(float);
  static bool set = false;
  if (!set) {
    OF_CUDA(cudaFuncSetAttribute(cv7_bwd_kernel<float, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    OF_CUDA(cudaFuncSetAttribute(cv7_bwd_kernel<float, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    OF_CUDA(cudaFuncSetAttribute(cv7_bwd_kernel<__nv_bfloat16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    OF_CUDA(cudaFuncSetAttribute(cv7_bwd_kernel<__nv_bfloat16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    set = true;
  }
  long long ntiles = (long long)B * ((W + CV_TC - 1) / CV_TC) * ((H + CV_TR - 1) / CV_TR);
  int grid = (int)(ntiles < OF_NUM_SMS ? ntiles : OF_NUM_SMS);
  cv7_bwd_kernel<T, true><<<grid, CVB_THREADS, smem, st>>>((const T*)f2, (const T*)d_cv, d_f1, B, H, W, C, cv_stride);
  if (int r = of_launch_check("cv7_bwd_kernel<d_f1>")) return r;
  cv7_bwd_kernel<T, false><<<grid, CVB_THREADS, smem, st>>>((const T*)f1, (const T*)d_cv, d_f2, B, H, W, C, cv_stride);
  return of_launch_check("cv7_bwd_kernel<d_f2>");
}

// bf16 in, bf16 out: d_f1 = cv-gradient (+ d_f1_add), d_f2 = cv-gradient
int cv7_bwd_bf16(const void* f1, const void* f2, const void* d_cv, void* d_f1, void* d_f2, const void* d_f1_add, int B,
                 int H, int W, int C, int cv_stride, cudaStream_t st) {
  typedef __nv_bfloat16 T;
  size_t smem = (size_t)CV_HR * CV_HC * CVB_WORDS * sizeof(uint32_t) + (size_t)CV_D * CV_TR * CV_TC * 8 * sizeof(float);
  static bool set = false;
  if (!set) {
    OF_CUDA(cudaFuncSetAttribute(cv7_bwd_kernel<T, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    OF_CUDA(cudaFuncSetAttribute(cv7_bwd_kernel<T, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    set = true;
  }
  long long ntiles = (long long)B * ((W + CV_TC - 1) / CV_TC) * ((H + CV_TR - 1) / CV_TR);
  int grid = (int)(ntiles < OF_NUM_SMS ? ntiles : OF_NUM_SMS);
  cv7_bwd_kernel<T, true, true><<<grid, CVB_THREADS, smem, st>>>((const T*)f2, (const T*)d_cv, d_f1, B, H, W, C, cv_stride,
                                                                 (const T*)d_f1_add);
  if (int r = of_launch_check("cv7_bwd_kernel<d_f1,bf16>")) return r;
  cv7_bwd_kernel<T, false, true><<<grid, CVB_THREADS, smem, st>>>((const T*)f1, (const T*)d_cv, d_f2, B, H, W, C, cv_stride,
                                                                  nullptr);
  return of_launch_check("cv7_bwd_kernel<d_f2,bf16>");
}

template int cv7_fwd<float>(const void*, const void*, void*, int, int, int, int, int, cudaStream_t);
template int cv7_fwd<__nv_bfloat16>(const void*, const void*, void*, int, int, int, int, int, cudaStream_t);
template int cv7_bwd<float>(const void*, const void*, const void*, float*, float*, int, int, int, int, int, cudaStream_t);
template int cv7_bwd<__nv_bfloat16>(const void*, const void*, const void*, float*, float*, int, int, int, int, int,
                                    cudaStream_t);
