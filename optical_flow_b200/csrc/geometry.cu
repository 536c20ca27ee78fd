// Depth / pose re-projection chain of transformations.py:
//   BackprojectLayer  :12-30    depth -> homogeneous 3-D points        (backproject_*_kernel)
//   transform3d       :61-67    per-sample 4x4 rigid transform         (view_*_kernel, PROJECT = false)
//   ProjectLayer      :33-44    pinhole projection, /(z + 1e-7)        (project_*_kernel)
//   WarpLayer         :47-58    transform3d + project fused, the coords feed K1 (view_*_kernel, PROJECT = true)
//   make_transformation_matrix / rotation_from_axisangle :133-191      (pose_*_kernel)
// All of it is HBM-bound elementwise work on (B,h,w,{1,2,4}) fp32 tensors: one thread per pixel, float4 /
// float2 accesses, intrinsics by value in the parameter block, the pose matrix of the sample in registers.
// The only reduction is d_T (16 sums per sample over h*w pixels): registers -> warp shuffle -> shared ->
// one red.global.add per CTA and entry.
#include "common.cuh"

struct Mat3 { float m[9]; };   // row-major, passed by value (host intrinsics)

// ---------------------------------------------------------------- backproject --
__global__ void __launch_bounds__(256) backproject_fwd_kernel(const float* __restrict__ depth, Mat3 kinv,
                                                              float4* __restrict__ points, int B, int H, int W) {
  const size_t n = (size_t)B * H * W;
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (size_t)gridDim.x * blockDim.x) {
    int j = (int)(p % W), i = (int)((p / W) % H);
    float x = (float)j, y = (float)i;
    // rays = Kinv @ [x, y, 1]  (:19-20)
    float r0 = kinv.m[0] * x + kinv.m[1] * y + kinv.m[2];
    float r1 = kinv.m[3] * x + kinv.m[4] * y + kinv.m[5];
    float r2 = kinv.m[6] * x + kinv.m[7] * y + kinv.m[8];
    float d = __ldg(depth + p);
    points[p] = make_float4(d * r0, d * r1, d * r2, 1.0f);   // :27-28
  }
}

__global__ void __launch_bounds__(256) backproject_bwd_kernel(const float4* __restrict__ d_points, Mat3 kinv,
                                                              float* __restrict__ d_depth, int B, int H, int W) {
  const size_t n = (size_t)B * H * W;
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (size_t)gridDim.x * blockDim.x) {
    int j = (int)(p % W), i = (int)((p / W) % H);
    float x = (float)j, y = (float)i;
    float r0 = kinv.m[0] * x + kinv.m[1] * y + kinv.m[2];
    float r1 = kinv.m[3] * x + kinv.m[4] * y + kinv.m[5];
    float r2 = kinv.m[6] * x + kinv.m[7] * y + kinv.m[8];
    float4 g = __ldg(d_points + p);
    d_depth[p] = g.x * r0 + g.y * r1 + g.z * r2;
  }
}

// -------------------------------------------------------------------- project --
__device__ __forceinline__ float2 project_point(const Mat3& k, float4 q) {
  // K_ext = [K | 0] (:36); pix = hom[:2] / (hom[2] + 1e-7) (:42-43)
  float h0 = k.m[0] * q.x + k.m[1] * q.y + k.m[2] * q.z;
  float h1 = k.m[3] * q.x + k.m[4] * q.y + k.m[5] * q.z;
  float h2 = k.m[6] * q.x + k.m[7] * q.y + k.m[8] * q.z;
  float den = h2 + 1e-7f;
  return make_float2(h0 / den, h1 / den);
}

__device__ __forceinline__ float4 project_point_bwd(const Mat3& k, float4 q, float2 g) {
  float h0 = k.m[0] * q.x + k.m[1] * q.y + k.m[2] * q.z;
  float h1 = k.m[3] * q.x + k.m[4] * q.y + k.m[5] * q.z;
  float h2 = k.m[6] * q.x + k.m[7] * q.y + k.m[8] * q.z;
  float den = h2 + 1e-7f;
  float inv = 1.0f / den;
  float g0 = g.x * inv, g1 = g.y * inv;
  float g2 = -(g.x * h0 + g.y * h1) * inv * inv;
  return make_float4(k.m[0] * g0 + k.m[3] * g1 + k.m[6] * g2,
                     k.m[1] * g0 + k.m[4] * g1 + k.m[7] * g2,
                     k.m[2] * g0 + k.m[5] * g1 + k.m[8] * g2, 0.0f);
}

__global__ void __launch_bounds__(256) project_fwd_kernel(const float4* __restrict__ points, Mat3 k,
                                                          float2* __restrict__ pix, size_t n) {
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (size_t)gridDim.x * blockDim.x)
    pix[p] = project_point(k, __ldg(points + p));
}

__global__ void __launch_bounds__(256) project_bwd_kernel(const float4* __restrict__ points, Mat3 k,
                                                          const float2* __restrict__ d_pix,
                                                          float4* __restrict__ d_points, size_t n) {
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (size_t)gridDim.x * blockDim.x)
    d_points[p] = project_point_bwd(k, __ldg(points + p), __ldg(d_pix + p));
}

// ------------------------------------------------- transform3d (+ projection) --
__device__ __forceinline__ float4 mat4_apply(const float* t, float4 p) {
  return make_float4(t[0] * p.x + t[1] * p.y + t[2] * p.z + t[3] * p.w,
                     t[4] * p.x + t[5] * p.y + t[6] * p.z + t[7] * p.w,
                     t[8] * p.x + t[9] * p.y + t[10] * p.z + t[11] * p.w,
                     t[12] * p.x + t[13] * p.y + t[14] * p.z + t[15] * p.w);
}

// grid = (ctas per sample, B).  PROJECT: out = pixel coordinates (float2) else transformed points (float4).
template <bool PROJECT>
__global__ void __launch_bounds__(256) view_fwd_kernel(const float* __restrict__ T, const float4* __restrict__ points,
                                                       Mat3 k, void* __restrict__ out, int HW) {
  const int b = blockIdx.y;
  float t[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) t[e] = __ldg(T + b * 16 + e);
  const float4* pin = points + (size_t)b * HW;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
    float4 q = mat4_apply(t, __ldg(pin + p));
    if (PROJECT) reinterpret_cast<float2*>(out)[(size_t)b * HW + p] = project_point(k, q);
    else reinterpret_cast<float4*>(out)[(size_t)b * HW + p] = q;
  }
}

template <bool PROJECT>
__global__ void __launch_bounds__(256) view_bwd_kernel(const float* __restrict__ T, const float4* __restrict__ points,
                                                       Mat3 k, const void* __restrict__ d_out,
                                                       float4* __restrict__ d_points, float* __restrict__ d_T, int HW) {
  const int b = blockIdx.y;
  float t[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) t[e] = __ldg(T + b * 16 + e);
  float acc[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) acc[e] = 0.0f;
  const float4* pin = points + (size_t)b * HW;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += gridDim.x * blockDim.x) {
    float4 p1 = __ldg(pin + p);
    float4 g;
    if (PROJECT) {
      float4 q = mat4_apply(t, p1);
      g = project_point_bwd(k, q, __ldg(reinterpret_cast<const float2*>(d_out) + (size_t)b * HW + p));
    } else {
      g = __ldg(reinterpret_cast<const float4*>(d_out) + (size_t)b * HW + p);
    }
    if (d_points != nullptr) {
      d_points[(size_t)b * HW + p] = make_float4(t[0] * g.x + t[4] * g.y + t[8] * g.z + t[12] * g.w,
                                                 t[1] * g.x + t[5] * g.y + t[9] * g.z + t[13] * g.w,
                                                 t[2] * g.x + t[6] * g.y + t[10] * g.z + t[14] * g.w,
                                                 t[3] * g.x + t[7] * g.y + t[11] * g.z + t[15] * g.w);
    }
    const float gv[4] = {g.x, g.y, g.z, g.w}, pv[4] = {p1.x, p1.y, p1.z, p1.w};
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[r * 4 + c] += gv[r] * pv[c];
  }
  if (d_T == nullptr) return;
  __shared__ float s_part[8][16];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int e = 0; e < 16; ++e) {
    float v = warp_sum(acc[e]);
    if (lane == 0) s_part[warp][e] = v;
  }
  __syncthreads();
  if (threadIdx.x < 16) {
    float v = 0.0f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) v += s_part[w][threadIdx.x];
    red_add(d_T + b * 16 + threadIdx.x, v);
  }
}

// ------------------------------------------------------- pose (axis-angle) --
struct Pose {
  float x, y, z, c, s, C, theta, den;
  float R[9];
};

__device__ __forceinline__ Pose pose_rotation(float a0, float a1, float a2) {
  Pose q;
  q.theta = sqrtf(a0 * a0 + a1 * a1 + a2 * a2);          // :137
  q.den = q.theta + 1e-7f;
  q.x = a0 / q.den; q.y = a1 / q.den; q.z = a2 / q.den;  // :138
  q.c = cosf(q.theta); q.s = sinf(q.theta); q.C = 1.0f - q.c;
  float xs = q.x * q.s, ys = q.y * q.s, zs = q.z * q.s;
  float xC = q.x * q.C, yC = q.y * q.C, zC = q.z * q.C;
  float xyC = q.x * yC, yzC = q.y * zC, zxC = q.z * xC;
  q.R[0] = q.x * xC + q.c; q.R[1] = xyC - zs;        q.R[2] = zxC + ys;   // :160-168
  q.R[3] = xyC + zs;       q.R[4] = q.y * yC + q.c;  q.R[5] = yzC - xs;
  q.R[6] = zxC - ys;       q.R[7] = yzC + xs;        q.R[8] = q.z * zC + q.c;
  return q;
}

__global__ void pose_fwd_kernel(const float* __restrict__ raw, int invert, float* __restrict__ out, int B) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const float* r = raw + b * 6;
  Pose q = pose_rotation(r[0], r[1], r[2]);
  float t0 = r[3], t1 = r[4], t2 = r[5];
  float R[9];
#pragma unroll
  for (int e = 0; e < 9; ++e) R[e] = q.R[e];
  if (invert) {                                            // :185-187
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) R[i * 3 + j] = q.R[j * 3 + i];
    float u0 = -(R[0] * t0 + R[1] * t1 + R[2] * t2);
    float u1 = -(R[3] * t0 + R[4] * t1 + R[5] * t2);
    float u2 = -(R[6] * t0 + R[7] * t1 + R[8] * t2);
    t0 = u0; t1 = u1; t2 = u2;
  }
  float* m = out + b * 16;                                 // :188-190
  m[0] = R[0]; m[1] = R[1]; m[2] = R[2];  m[3] = t0;
  m[4] = R[3]; m[5] = R[4]; m[6] = R[5];  m[7] = t1;
  m[8] = R[6]; m[9] = R[7]; m[10] = R[8]; m[11] = t2;
  m[12] = 0.f; m[13] = 0.f; m[14] = 0.f;  m[15] = 1.f;
}

__global__ void pose_bwd_kernel(const float* __restrict__ raw, int invert, const float* __restrict__ d_out,
                                float* __restrict__ d_raw, int B) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const float* r = raw + b * 6;
  const float* g = d_out + b * 16;
  Pose q = pose_rotation(r[0], r[1], r[2]);
  float dR[9] = {g[0], g[1], g[2], g[4], g[5], g[6], g[8], g[9], g[10]};
  float dt[3] = {g[3], g[7], g[11]};
  float dtr[3];
  if (invert) {
    // Q = R^T, t' = -Q t:  dQ = dR - dt' (x) t,  dt = -Q^T dt' = -R dt';  dR_orig = dQ^T
    float t[3] = {r[3], r[4], r[5]};
    float dQ[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) dQ[i * 3 + j] = dR[i * 3 + j] - dt[i] * t[j];
#pragma unroll
    for (int i = 0; i < 3; ++i) dtr[i] = -(q.R[i * 3] * dt[0] + q.R[i * 3 + 1] * dt[1] + q.R[i * 3 + 2] * dt[2]);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) dR[i * 3 + j] = dQ[j * 3 + i];
  } else {
    dtr[0] = dt[0]; dtr[1] = dt[1]; dtr[2] = dt[2];
  }
  const float x = q.x, y = q.y, z = q.z, s = q.s, c = q.c, Cc = q.C;
  float s01 = dR[1] + dR[3], s02 = dR[2] + dR[6], s12 = dR[5] + dR[7];
  float dC = dR[0] * x * x + dR[4] * y * y + dR[8] * z * z + s01 * x * y + s02 * z * x + s12 * y * z;
  float dc = dR[0] + dR[4] + dR[8] - dC;
  float ds = z * (dR[3] - dR[1]) + y * (dR[2] - dR[6]) + x * (dR[7] - dR[5]);
  float dx = 2.f * x * Cc * dR[0] + s01 * y * Cc + s02 * z * Cc + s * (dR[7] - dR[5]);
  float dy = 2.f * y * Cc * dR[4] + s01 * x * Cc + s12 * z * Cc + s * (dR[2] - dR[6]);
  float dz = 2.f * z * Cc * dR[8] + s02 * x * Cc + s12 * y * Cc + s * (dR[3] - dR[1]);
  float dtheta = c * ds - s * dc;
  float a[3] = {r[0], r[1], r[2]};
  float inv = 1.0f / q.den;
  dtheta -= (dx * a[0] + dy * a[1] + dz * a[2]) * inv * inv;
  float nrm = q.theta > 0.f ? 1.0f / q.theta : 0.f;      // d|a|/da at 0 := 0 (autodiff convention)
  float* o = d_raw + b * 6;
  o[0] = dx * inv + dtheta * a[0] * nrm;
  o[1] = dy * inv + dtheta * a[1] * nrm;
  o[2] = dz * inv + dtheta * a[2] * nrm;
  o[3] = dtr[0]; o[4] = dtr[1]; o[5] = dtr[2];
}

// ------------------------------------------------------------------ C ABI --
static inline Mat3 load_mat3(const float* host9) {
  Mat3 m;
  for (int e = 0; e < 9; ++e) m.m[e] = host9[e];
  return m;
}
static inline int grid_1d(size_t n) {
  size_t g = (n + 255) / 256;
  size_t cap = (size_t)OF_NUM_SMS * 8;
  return (int)(g < cap ? (g ? g : 1) : cap);
}
#define GEOM_DIMS_OK(B, H, W) OF_REQUIRE((B) > 0 && (H) > 0 && (W) > 0, OF_ERR_SHAPE, "%s: bad dims B=%d H=%d W=%d", __func__, B, H, W)

extern "C" int of_backproject_fwd(const float* depth, const float* kinv_host, float* points, int B, int H, int W,
                                  void* stream) {
  OF_REQUIRE(depth && kinv_host && points, OF_ERR_ARG, "of_backproject_fwd: null pointer");
  GEOM_DIMS_OK(B, H, W);
  backproject_fwd_kernel<<<grid_1d((size_t)B * H * W), 256, 0, (cudaStream_t)stream>>>(
      depth, load_mat3(kinv_host), reinterpret_cast<float4*>(points), B, H, W);
  return of_launch_check("backproject_fwd_kernel");
}

extern "C" int of_backproject_bwd(const float* d_points, const float* kinv_host, float* d_depth, int B, int H, int W,
                                  void* stream) {
  OF_REQUIRE(d_points && kinv_host && d_depth, OF_ERR_ARG, "of_backproject_bwd: null pointer");
  GEOM_DIMS_OK(B, H, W);
  backproject_bwd_kernel<<<grid_1d((size_t)B * H * W), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(d_points), load_mat3(kinv_host), d_depth, B, H, W);
  return of_launch_check("backproject_bwd_kernel");
}

extern "C" int of_project_fwd(const float* k_host, const float* points, float* pix, int B, int H, int W,
                              void* stream) {
  OF_REQUIRE(k_host && points && pix, OF_ERR_ARG, "of_project_fwd: null pointer");
  GEOM_DIMS_OK(B, H, W);
  size_t n = (size_t)B * H * W;
  project_fwd_kernel<<<grid_1d(n), 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(points),
                                                                    load_mat3(k_host),
                                                                    reinterpret_cast<float2*>(pix), n);
  return of_launch_check("project_fwd_kernel");
}

extern "C" int of_project_bwd(const float* k_host, const float* points, const float* d_pix, float* d_points, int B,
                              int H, int W, void* stream) {
  OF_REQUIRE(k_host && points && d_pix && d_points, OF_ERR_ARG, "of_project_bwd: null pointer");
  GEOM_DIMS_OK(B, H, W);
  size_t n = (size_t)B * H * W;
  project_bwd_kernel<<<grid_1d(n), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(points), load_mat3(k_host), reinterpret_cast<const float2*>(d_pix),
      reinterpret_cast<float4*>(d_points), n);
  return of_launch_check("project_bwd_kernel");
}

static inline dim3 view_grid(int B, int HW) {
  int per = (HW + 256 * 4 - 1) / (256 * 4);   // ~4 pixels per thread so the d_T partials amortise
  int cap = (OF_NUM_SMS * 8 + B - 1) / B;
  if (per > cap) per = cap;
  if (per < 1) per = 1;
  return dim3(per, B);
}

extern "C" int of_transform3d_fwd(const float* T, const float* points, float* out, int B, int H, int W,
                                  void* stream) {
  OF_REQUIRE(T && points && out, OF_ERR_ARG, "of_transform3d_fwd: null pointer");
  GEOM_DIMS_OK(B, H, W);
  Mat3 none = {};
  view_fwd_kernel<false><<<view_grid(B, H * W), 256, 0, (cudaStream_t)stream>>>(
      T, reinterpret_cast<const float4*>(points), none, out, H * W);
  return of_launch_check("view_fwd_kernel<false>");
}

extern "C" int of_transform3d_bwd(const float* T, const float* points, const float* d_out, float* d_points,
                                  float* d_T, int B, int H, int W, void* stream) {
  OF_REQUIRE(T && points && d_out, OF_ERR_ARG, "of_transform3d_bwd: null pointer");
  GEOM_DIMS_OK(B, H, W);
  if (d_T) OF_CUDA(cudaMemsetAsync(d_T, 0, (size_t)B * 16 * sizeof(float), (cudaStream_t)stream));
  Mat3 none = {};
  view_bwd_kernel<false><<<view_grid(B, H * W), 256, 0, (cudaStream_t)stream>>>(
      T, reinterpret_cast<const float4*>(points), none, d_out, reinterpret_cast<float4*>(d_points), d_T, H * W);
  return of_launch_check("view_bwd_kernel<false>");
}

extern "C" int of_view_coords_fwd(const float* k_host, const float* T, const float* points, float* pix, int B, int H,
                                  int W, void* stream) {
  OF_REQUIRE(k_host && T && points && pix, OF_ERR_ARG, "of_view_coords_fwd: null pointer");
  GEOM_DIMS_OK(B, H, W);
  view_fwd_kernel<true><<<view_grid(B, H * W), 256, 0, (cudaStream_t)stream>>>(
      T, reinterpret_cast<const float4*>(points), load_mat3(k_host), pix, H * W);
  return of_launch_check("view_fwd_kernel<true>");
}

extern "C" int of_view_coords_bwd(const float* k_host, const float* T, const float* points, const float* d_pix,
                                  float* d_points, float* d_T, int B, int H, int W, void* stream) {
  OF_REQUIRE(k_host && T && points && d_pix, OF_ERR_ARG, "of_view_coords_bwd: null pointer");
  GEOM_DIMS_OK(B, H, W);
  if (d_T) OF_CUDA(cudaMemsetAsync(d_T, 0, (size_t)B * 16 * sizeof(float), (cudaStream_t)stream));
  view_bwd_kernel<true><<<view_grid(B, H * W), 256, 0, (cudaStream_t)stream>>>(
      T, reinterpret_cast<const float4*>(points), load_mat3(k_host), d_pix, reinterpret_cast<float4*>(d_points), d_T,
      H * W);
  return of_launch_check("view_bwd_kernel<true>");
}

extern "C" int of_make_transformation_matrix_fwd(const float* raw, int invert, float* out, int B, void* stream) {
  OF_REQUIRE(raw && out, OF_ERR_ARG, "of_make_transformation_matrix_fwd: null pointer");
  OF_REQUIRE(B > 0, OF_ERR_SHAPE, "of_make_transformation_matrix_fwd: B=%d", B);
  pose_fwd_kernel<<<(B + 63) / 64, 64, 0, (cudaStream_t)stream>>>(raw, invert, out, B);
  return of_launch_check("pose_fwd_kernel");
}

extern "C" int of_make_transformation_matrix_bwd(const float* raw, int invert, const float* d_out, float* d_raw,
                                                 int B, void* stream) {
  OF_REQUIRE(raw && d_out && d_raw, OF_ERR_ARG, "of_make_transformation_matrix_bwd: null pointer");
  OF_REQUIRE(B > 0, OF_ERR_SHAPE, "of_make_transformation_matrix_bwd: B=%d", B);
  pose_bwd_kernel<<<(B + 63) / 64, 64, 0, (cudaStream_t)stream>>>(raw, invert, d_out, d_raw, B);
  return of_launch_check("pose_bwd_kernel");
}
