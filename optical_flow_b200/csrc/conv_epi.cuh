// Epilogue helpers shared by the tcgen05 convolution kernels (conv_tc.cu, conv_tc3.cu).
#pragma once
#include "common.cuh"

__device__ __forceinline__ float load_any(const void* p, int dtype, size_t idx) {
  return dtype == OF_F32 ? __ldg(reinterpret_cast<const float*>(p) + idx)
                         : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[idx]);
}
__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == OF_ACT_RELU) return fmaxf(v, 0.f);
  if (act == OF_ACT_LEAKY) return v > 0.f ? v : 0.3f * v;
  return v;
}
__device__ __forceinline__ float act_slope(float y, int act) {
  if (act == OF_ACT_RELU) return y > 0.f ? 1.f : 0.f;
  if (act == OF_ACT_LEAKY) return y > 0.f ? 1.f : 0.3f;
  return 1.f;
}

// 16 consecutive channels as floats: two 16-byte (bf16) / four 16-byte (f32) vector loads when the
// run is complete and aligned, scalar (zero-filled) otherwise
__device__ __forceinline__ void load16(const void* base, int dtype, size_t idx, int nv, float (&o)[16]) {
  if (dtype == OF_BF16) {
    const __nv_bfloat16* p = reinterpret_cast<const __nv_bfloat16*>(base) + idx;
    if (nv == 16 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
      uint4 q[2];
      q[0] = __ldg(reinterpret_cast<const uint4*>(p));
      q[1] = __ldg(reinterpret_cast<const uint4*>(p) + 1);
      const uint32_t* w = reinterpret_cast<const uint32_t*>(q);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        o[2 * j] = __uint_as_float(w[j] << 16);
        o[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) o[j] = j < nv ? __bfloat162float(p[j]) : 0.f;
    }
  } else {
    const float* p = reinterpret_cast<const float*>(base) + idx;
    if (nv == 16 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float4 t = __ldg(reinterpret_cast<const float4*>(p) + j);
        o[4 * j] = t.x; o[4 * j + 1] = t.y; o[4 * j + 2] = t.z; o[4 * j + 3] = t.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) o[j] = j < nv ? __ldg(p + j) : 0.f;
    }
  }
}


// 16 values from raw 16-byte vectors (bf16: q[0..1], f32: q[0..3])
__device__ __forceinline__ void unpack16(const uint4 (&q)[4], int dtype, float (&o)[16]) {
  const uint32_t* w = reinterpret_cast<const uint32_t*>(q);
  if (dtype == OF_BF16) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      o[2 * j] = __uint_as_float(w[j] << 16);
      o[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
    }
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j) o[j] = __uint_as_float(w[j]);
  }
}

struct ConvEpi {
  void* out; int out_dtype; long long out_stride;
  const float* bias; int act;
  const void* add; int add_dtype; long long add_stride;
  const void* mask; int mask_dtype; long long mask_stride; int mask_act;
  int n_valid;
  void* out2; long long out2_stride; int split;   // split > 0: output columns >= split go to out2 (column - split)
  void* sign_out;      // nullable: 16 bytes per pixel, bit c = (output channel c > 0) - a later dgrad's activation mask
  int add_post;        // 1: y = relu(act(acc + bias) + add) instead of act(acc + bias + add) (conv_tc3 only)
};
#define OF_MASK_BITS 2   // mask_dtype: `mask` is such a sign-bit buffer (conv_tc3 only)

// 32 bytes in one instruction = one full L2 sector (p 32-byte aligned)
__device__ __forceinline__ void st_global_256(void* p, const uint32_t (&pk)[8]) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]),
               "r"(pk[3]), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7])
               : "memory");
}

// 32 bytes in one load (read-only path): two uint4 halves of one L2 sector (p 32-byte aligned)
__device__ __forceinline__ void ld_global_256(const void* p, uint4& a, uint4& b) {
  asm volatile("ld.global.nc.v8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)
               : "l"(p));
}

struct ConvEpi;
__device__ __forceinline__ void store16(const ConvEpi& E, size_t pix, int c0, const float (&v)[16]);

// 16 consecutive output channels [c0, c0+16) of pixel `pix` from 16 fp32 accumulators:
// (+bias) (+add tensor) -> activation -> (* act'(mask tensor)) -> bf16 / f32 store
__device__ __forceinline__ void epilogue_store16(const ConvEpi& E, size_t pix, int c0, const uint32_t (&rr)[16]) {
  float v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rr[j]);
  int nv = min(16, E.n_valid - c0);
  if (E.bias) {
    float t[16];
    load16(E.bias, OF_F32, (size_t)c0, nv, t);
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] += t[j];
  }
  if (E.add) {
    float t[16];
    load16(E.add, E.add_dtype, pix * E.add_stride + c0, nv, t);
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] += t[j];
  }
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = apply_act(v[j], E.act);
  if (E.mask) {
    float t[16];
    load16(E.mask, E.mask_dtype, pix * E.mask_stride + c0, nv, t);
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] *= act_slope(t[j], E.mask_act);
  }
  if (E.out_dtype == OF_BF16) {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(E.out) + pix * E.out_stride + c0;
    if (nv == 16 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
      uint32_t pk[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        __nv_bfloat162 t = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
        pk[j] = *reinterpret_cast<uint32_t*>(&t);
      }
      if ((reinterpret_cast<uintptr_t>(o) & 31) == 0) {
        st_global_256(o, pk);
      } else {
        reinterpret_cast<uint4*>(o)[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        reinterpret_cast<uint4*>(o)[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) if (j < nv) o[j] = __float2bfloat16_rn(v[j]);
    }
  } else {
    float* o = reinterpret_cast<float*>(E.out) + pix * E.out_stride + c0;
    if (nv == 16 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        reinterpret_cast<float4*>(o)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) if (j < nv) o[j] = v[j];
    }
  }
}

// aligned full chunk: 16 finished values -> bf16 (2 x 16 B) or f32 (4 x 16 B)
__device__ __forceinline__ void store16(const ConvEpi& E, size_t pix, int c0, const float (&v)[16]) {
  if (E.out_dtype == OF_BF16) {
    __nv_bfloat16* o = (E.split > 0 && c0 >= E.split)
                           ? reinterpret_cast<__nv_bfloat16*>(E.out2) + pix * E.out2_stride + (c0 - E.split)
                           : reinterpret_cast<__nv_bfloat16*>(E.out) + pix * E.out_stride + c0;
    uint32_t pk[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      __nv_bfloat162 t = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
      pk[j] = *reinterpret_cast<uint32_t*>(&t);
    }
    if ((reinterpret_cast<uintptr_t>(o) & 31) == 0) {
      // one 256-bit store = one full 32-byte L2 sector (two 16-byte stores are two partial-sector writes: the conv
      // kernels are L2-slice-throughput bound, and the output stores were 17 % of a 128->128 layer's time)
      st_global_256(o, pk);
    } else {
      reinterpret_cast<uint4*>(o)[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      reinterpret_cast<uint4*>(o)[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    }
  } else {
    float* o = reinterpret_cast<float*>(E.out) + pix * E.out_stride + c0;
    if ((reinterpret_cast<uintptr_t>(o) & 31) == 0) {
      uint32_t lo[8], hi[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) { lo[j] = __float_as_uint(v[j]); hi[j] = __float_as_uint(v[8 + j]); }
      st_global_256(o, lo);
      st_global_256(o + 8, hi);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        reinterpret_cast<float4*>(o)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
  }
}
