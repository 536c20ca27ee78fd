// K10: convolution as a tcgen05 / TMEM implicit GEMM (bf16 operands, fp32 accumulate).
//
// Keras Conv2D(padding='same') on NHWC (model.py:12-15,104-114; resnet/models.py:17-30),
// forward and data-gradient.  No im2col buffer: for every filter tap the A tile is a TMA box
// {64 channels, TW, TH, 1 image} of the NHWC activation tensor, shifted by the tap offset -
// out-of-bounds rows/columns are zero-filled by TMA, which IS the SAME padding - landing in
// shared memory as the canonical K-major SWIZZLE_128B operand (128 pixel rows x 128 bytes).
// Stride-2 convs use TMA elementStrides.  The decoder's tf.concat (model.py:100-102) is never
// materialised: the K loop walks up to two source tensors.  B tiles are rows of the packed
// weights [tap][Cout][K].  D (128 pixels x Cout, fp32) lives in TMEM, double-buffered so the
// epilogue of tile i overlaps the MMAs of tile i+1.
//
// Warp roles (192 threads, persistent over output tiles): warps 0-3 epilogue (TMEM lane
// quarter = warp id), warp 4 TMA producer, warp 5 MMA issuer + TMEM allocator.
// Epilogue: (+bias) (+residual/add tensor) -> activation -> (* act'(mask tensor)) -> bf16/f32 store.
//   dgrad = the same kernel on dy with weights packed [tap][Cin][Cout] and negated tap offsets;
//   its epilogue multiplies by act'(y_prev) so the next layer's dz comes out directly.
#include "common.cuh"
#include "tc_common.cuh"
#include "conv_epi.cuh"

#define TC_THREADS 192
#define A_TILE_BYTES 16384   // 128 pixels x 64 bf16
#define MAX_STAGES 8

struct ConvTcParams {
  int num_src;
  int kchunks[2];     // 64-channel chunks per source
  int kbase[2];       // first K column of each source in the packed weights
  int taps, KW, pad_t, pad_l, stride_w, stride_h, sign;
  int NB, Ho, Wo;     // output (GEMM M) space
  int TW, TH, tiles_w, tiles_h, m_tiles;
  int n_mma, n_valid; // MMA N (multiple of 16) and real output channels
  int stages, b_tile_bytes;
  int tmem_cols;
  // epilogue
  void* out; int out_dtype; long long out_stride;
  const float* bias; int act;
  const void* add; int add_dtype; long long add_stride;
  const void* mask; int mask_dtype; long long mask_stride; int mask_act;
};

// SPEC >= 0: compile-time epilogue (same reason as conv_tc3_kernel's): bits 0-1 activation, bit 2 add (bf16), bit 3 mask
// (bf16 tensor), bits 4-5 mask activation; always bf16 output, n_valid == n_mma, 32-byte aligned rows, bias from shared memory
#define TC_SPEC(act, add, mask, mact) ((act) | ((add) << 2) | ((mask) << 3) | ((mact) << 4))
template <int SPEC = -1>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmB, const __grid_constant__ ConvTcParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte aligned tile area (SWIZZLE_128B atoms), then barriers
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = A_TILE_BYTES + P.b_tile_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + (size_t)P.stages * stage_bytes);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tmem_full = empty_bar + MAX_STAGES;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;           // [2]
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P.stages; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { tc::mbar_init(&tmem_full[b], 1); tc::mbar_init(&tmem_empty[b], 4); }
    tc::fence_barrier_init();
  }
  if (warp == 4 && lane == 0) {
    tc::prefetch_tmap(&tmA0);
    if (P.num_src > 1) tc::prefetch_tmap(&tmA1);
    tc::prefetch_tmap(&tmB);
  }
  if (warp == 5) {
    tc::tmem_alloc(tmem_base_slot, (uint32_t)P.tmem_cols);
    tc::tmem_relinquish();
  }
  if (threadIdx.x == 0) OF_PDL_TRIGGER();
  OF_PDL_WAIT();                                     // nothing above touches global memory
  __shared__ float s_bias[256];
  if (SPEC >= 0)
    for (int i = threadIdx.x; i < 256; i += TC_THREADS) s_bias[i] = (P.bias && i < P.n_valid) ? P.bias[i] : 0.f;
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_base_slot;

  const int kblocks_per_tile = P.taps * (P.kchunks[0] + (P.num_src > 1 ? P.kchunks[1] : 0));

  if (warp == 4) {
    // ===================== TMA producer =====================
    if (OF_ONE_LANE()) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < P.m_tiles; tile += gridDim.x) {
        int tw_i = tile % P.tiles_w; int r = tile / P.tiles_w;
        int th_i = r % P.tiles_h; int n = r / P.tiles_h;
        int w0 = tw_i * P.TW * P.stride_w, h0 = th_i * P.TH * P.stride_h;
        for (int s = 0; s < P.num_src; ++s) {
          const CUtensorMap* tmA = s == 0 ? &tmA0 : &tmA1;
          for (int tap = 0; tap < P.taps; ++tap) {
            int kh = tap / P.KW, kw = tap - kh * P.KW;
            int cw = w0 + P.sign * (kw - P.pad_l), ch = h0 + P.sign * (kh - P.pad_t);
            for (int kc = 0; kc < P.kchunks[s]; ++kc) {
              tc::mbar_wait(&empty_bar[stage], phase ^ 1);
              uint8_t* a_dst = tiles + (size_t)stage * stage_bytes;
              uint8_t* b_dst = a_dst + A_TILE_BYTES;
              tc::mbar_expect_tx(&full_bar[stage], (uint32_t)(A_TILE_BYTES + P.n_mma * 128));
              tc::tma_load_4d(a_dst, tmA, &full_bar[stage], kc * 64, cw, ch, n);
              tc::tma_load_3d(b_dst, &tmB, &full_bar[stage], P.kbase[s] + kc * 64, 0, tap);
              if (++stage == P.stages) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 5) {
    // ===================== MMA issuer =====================
    if (OF_ONE_LANE()) {
      const uint32_t idesc = tc::idesc_bf16_f32(128, P.n_mma, 0, 0);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < P.m_tiles; tile += gridDim.x, ++it) {
        int buf = it & 1;
        uint32_t use = (uint32_t)(it >> 1);
        tc::mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);   // epilogue has drained this accumulator
        tc::fence_after_sync();
        uint32_t d_tmem = tmem_base + (uint32_t)(buf * P.n_mma);
        for (int kb = 0; kb < kblocks_per_tile; ++kb) {
          tc::mbar_wait(&full_bar[stage], phase);
          tc::fence_after_sync();
          uint32_t a_addr = tc::smem_u32(tiles + (size_t)stage * stage_bytes);
          uint32_t b_addr = a_addr + A_TILE_BYTES;
          uint64_t adesc = tc::desc_sw128_kmajor(a_addr);
          uint64_t bdesc = tc::desc_sw128_kmajor(b_addr);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the 128-byte swizzle row: +2 in the >>4 address field
            tc::mma_bf16(d_tmem, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) ? 1u : 0u);
          }
          tc::mma_commit(&empty_bar[stage]);            // frees the smem slot when these MMAs finish
          if (kb == kblocks_per_tile - 1) tc::mma_commit(&tmem_full[buf]);
          if (++stage == P.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===================== epilogue warps 0..3 =====================
    int it = 0;
    for (int tile = blockIdx.x; tile < P.m_tiles; tile += gridDim.x, ++it) {
      int buf = it & 1;
      uint32_t use = (uint32_t)(it >> 1);
      int tw_i = tile % P.tiles_w; int r = tile / P.tiles_w;
      int th_i = r % P.tiles_h; int n = r / P.tiles_h;
      int row = warp * 32 + lane;                       // accumulator row == TMEM lane == pixel in tile
      int ph = th_i * P.TH + row / P.TW, pw = tw_i * P.TW + row % P.TW;
      bool valid = ph < P.Ho && pw < P.Wo;
      size_t pix = ((size_t)n * P.Ho + ph) * P.Wo + pw;
      tc::mbar_wait(&tmem_full[buf], use & 1);
      tc::fence_after_sync();
      uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(buf * P.n_mma);
      if constexpr (SPEC >= 0) {
        constexpr int ACT = SPEC & 3, MACT = (SPEC >> 4) & 3;
        constexpr bool ADD = (SPEC >> 2) & 1, MASK = (SPEC >> 3) & 1;
        const int nchunks = P.n_mma >> 4;
        uint32_t rr[2][16];
        uint4 araw[2][2], mraw[2][2];
        const __nv_bfloat16* addp = reinterpret_cast<const __nv_bfloat16*>(P.add) + pix * P.add_stride;
        const __nv_bfloat16* maskp = reinterpret_cast<const __nv_bfloat16*>(P.mask) + pix * P.mask_stride;
        __nv_bfloat16* outp = reinterpret_cast<__nv_bfloat16*>(P.out) + pix * P.out_stride;
        auto issue = [&](int c, int slot) {
          __syncwarp();
          tc::tmem_ld16(taddr + (uint32_t)(c * 16), rr[slot]);
          if (ADD && valid) ld_global_256(addp + c * 16, araw[slot][0], araw[slot][1]);
          if (MASK && valid) ld_global_256(maskp + c * 16, mraw[slot][0], mraw[slot][1]);
        };
        auto finish = [&](int c, int slot) {
          __syncwarp();
          tc::tmem_ld_wait();
          if (c + 1 < nchunks) issue(c + 1, slot ^ 1);
          if (!valid) return;
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rr[slot][j]) + s_bias[c * 16 + j];
          if (ADD) {
            float t[16];
            const uint4 a4[4] = {araw[slot][0], araw[slot][1], make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
            unpack16(a4, OF_BF16, t);
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] += t[j];
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = apply_act(v[j], ACT);
          if (MASK) {
            float t[16];
            const uint4 m4[4] = {mraw[slot][0], mraw[slot][1], make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
            unpack16(m4, OF_BF16, t);
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] *= act_slope(t[j], MACT);
          }
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            __nv_bfloat162 t2 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
            pk[j] = *reinterpret_cast<uint32_t*>(&t2);
          }
          st_global_256(outp + c * 16, pk);
        };
        issue(0, 0);
#pragma unroll 1
        for (int c = 0; c < nchunks; c += 2) {
          finish(c, 0);
          if (c + 1 < nchunks) finish(c + 1, 1);
        }
      } else
      for (int c0 = 0; c0 < P.n_mma; c0 += 16) {
        __syncwarp();                                   // tcgen05.ld is .sync.aligned: reconverge first
        uint32_t rr[16];
        tc::tmem_ld16(taddr + (uint32_t)c0, rr);
        tc::tmem_ld_wait();
        if (valid && c0 < P.n_valid) {
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rr[j]);
        int nv = min(16, P.n_valid - c0);
        if (P.bias) {
          float t[16];
          load16(P.bias, OF_F32, (size_t)c0, nv, t);
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] += t[j];
        }
        if (P.add) {
          float t[16];
          load16(P.add, P.add_dtype, pix * P.add_stride + c0, nv, t);
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] += t[j];
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = apply_act(v[j], P.act);
        if (P.mask) {
          float t[16];
          load16(P.mask, P.mask_dtype, pix * P.mask_stride + c0, nv, t);
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] *= act_slope(t[j], P.mask_act);
        }
        if (P.out_dtype == OF_BF16) {
          __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(P.out) + pix * P.out_stride + c0;
          if (nv == 16 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              __nv_bfloat162 t = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
              pk[j] = *reinterpret_cast<uint32_t*>(&t);
            }
            reinterpret_cast<uint4*>(o)[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            reinterpret_cast<uint4*>(o)[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) if (j < nv) o[j] = __float2bfloat16_rn(v[j]);
          }
        } else {
          float* o = reinterpret_cast<float*>(P.out) + pix * P.out_stride + c0;
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
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&tmem_empty[buf]);
    }
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 5) tc::tmem_dealloc(tmem_base, (uint32_t)P.tmem_cols);
}

// ---- weight packing: fp32 HWIO master weights -> bf16 [tap][rows][cols] -----------------------
// transpose == 0 (fprop): rows = co, cols = K over the (padded) sources;
// transpose == 1 (dgrad): rows = ci within ONE destination segment, cols = co.
// A segment maps original ci in [off, off+cnt) to packed index base + (ci - off).
struct PackSeg { int off[3], cnt[3], base[3], nseg; };

__global__ void __launch_bounds__(256) pack_weights_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ out,
                                                           int taps, int Cin, int Cout, int rows, int cols,
                                                           int transpose, PackSeg seg) {
  size_t total = (size_t)taps * rows * cols;
  for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
    int c = (int)(e % cols);
    size_t r2 = e / cols;
    int r = (int)(r2 % rows);
    int tap = (int)(r2 / rows);
    int packed_ci = transpose ? r : c;
    int co = transpose ? c : r;
    int ci = -1;
#pragma unroll
    for (int s = 0; s < 3; ++s)
      if (s < seg.nseg && packed_ci >= seg.base[s] && packed_ci < seg.base[s] + seg.cnt[s])
        ci = seg.off[s] + packed_ci - seg.base[s];
    float v = 0.f;
    if (ci >= 0 && co < Cout) v = w[((size_t)tap * Cin + ci) * Cout + co];
    out[e] = __float2bfloat16_rn(v);
  }
}

// All of a model's weight packings in ONE launch (the ~76 per-layer launches of 4 us each were 0.3 ms of a 13 ms
// step): a device table of jobs, each owning a contiguous range of 2048-element blocks.
#define PACK_BLOCK_ELEMS 2048
__global__ void __launch_bounds__(256) pack_weights_batch_kernel(const of_pack_job* __restrict__ jobs, int njobs) {
  int j = 0;
  while (j + 1 < njobs && (int)blockIdx.x >= jobs[j + 1].block_begin) ++j;
  const of_pack_job J = jobs[j];
  // 32-bit index arithmetic (a packing has < 2^31 elements; the host checks): the 64-bit divisions were most of this kernel
  const unsigned total = (unsigned)J.taps * (unsigned)J.rows * (unsigned)J.cols;
  const unsigned e0 = (unsigned)((int)blockIdx.x - J.block_begin) * PACK_BLOCK_ELEMS;
  __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(J.out);
  for (unsigned e = e0 + threadIdx.x; e < e0 + PACK_BLOCK_ELEMS && e < total; e += 256) {
    const unsigned r2 = e / (unsigned)J.cols;
    const int c = (int)(e - r2 * (unsigned)J.cols);
    const int tap = (int)(r2 / (unsigned)J.rows);
    const int r = (int)(r2 - (unsigned)tap * (unsigned)J.rows);
    int packed_ci = J.transpose ? r : c;
    int co = J.transpose ? c : r;
    int ci = -1;
#pragma unroll
    for (int s = 0; s < 3; ++s)
      if (s < J.nseg && packed_ci >= J.seg_base[s] && packed_ci < J.seg_base[s] + J.seg_cnt[s])
        ci = J.seg_off[s] + packed_ci - J.seg_base[s];
    float v = 0.f;
    if (ci >= 0 && co < J.Cout) v = J.w[((size_t)tap * J.Cin + ci) * J.Cout + co];
    out[e] = __float2bfloat16_rn(v);
  }
}

// ---- host side ----------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled g_encode = nullptr;

int of_tc_get_encode(void** fn) {
  if (!g_encode) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) {
      of_set_error("cuTensorMapEncodeTiled unavailable (%s)", cudaGetErrorString(e));
      return OF_ERR_STATE;
    }
    g_encode = (PFN_encodeTiled)p;
  }
  *fn = (void*)g_encode;
  return 0;
}

// bf16 activation tensor as a 4-D map {C, W, H, N} with explicit element strides between consecutive
// W / H / N coordinates (dense NHWC: stride_elems, W*stride_elems, H*W*stride_elems; an overlapping
// window view is legal: strides only have to be multiples of 16 bytes), box {64, box_w*sw, box_h*sh, 1},
// elementStrides {1, sw, sh, 1}, SWIZZLE_128B.
int of_tc_make_view_map(CUtensorMap* m, const void* base, int c_used, long long w_stride, long long h_stride,
                        long long n_stride, int N, int H, int W, int box_w, int box_h, int sw, int sh) {
  void* fn;
  if (int r = of_tc_get_encode(&fn)) return r;
  OF_REQUIRE(((uintptr_t)base & 15) == 0 && (w_stride * 2) % 16 == 0 && (h_stride * 2) % 16 == 0 &&
                 (n_stride * 2) % 16 == 0,
             OF_ERR_ARG, "TMA needs 16-byte aligned base/strides (channels %% 8 == 0)");
  cuuint64_t dims[4] = {(cuuint64_t)c_used, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)w_stride * 2, (cuuint64_t)h_stride * 2, (cuuint64_t)n_stride * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)(box_w * sw), (cuuint32_t)(box_h * sh), 1};
  cuuint32_t es[4] = {1, (cuuint32_t)sw, (cuuint32_t)sh, 1};
  CUresult r = ((PFN_encodeTiled)fn)(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides,
                                     box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  OF_REQUIRE(r == CUDA_SUCCESS, OF_ERR_STATE, "cuTensorMapEncodeTiled(activation) failed: %d", (int)r);
  return 0;
}

int of_tc_make_act_map(CUtensorMap* m, const void* base, int c_used, long long stride_elems, int N, int H, int W,
                       int box_w, int box_h, int estride) {
  return of_tc_make_view_map(m, base, c_used, stride_elems, (long long)W * stride_elems,
                             (long long)H * W * stride_elems, N, H, W, box_w, box_h, estride, estride);
}

// packed weights bf16 [taps][rows][cols] as a 3-D map {cols, rows, taps}, box {64, box_rows, 1}
int of_tc_make_weight_map(CUtensorMap* m, const void* base, int taps, int rows, int cols, int box_rows) {
  void* fn;
  if (int r = of_tc_get_encode(&fn)) return r;
  OF_REQUIRE(((uintptr_t)base & 15) == 0 && cols % 8 == 0, OF_ERR_ARG, "packed weights must be 16-byte aligned");
  cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)taps};
  cuuint64_t strides[2] = {(cuuint64_t)cols * 2, (cuuint64_t)rows * cols * 2};
  cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  cuuint32_t es[3] = {1, 1, 1};
  CUresult r = ((PFN_encodeTiled)fn)(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides,
                                     box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  OF_REQUIRE(r == CUDA_SUCCESS, OF_ERR_STATE, "cuTensorMapEncodeTiled(weights) failed: %d", (int)r);
  return 0;
}

static void pick_tile(int Wo, int* TW, int* TH) {
  // 128 output pixels per tile; prefer the widest TW that divides Wo, else 16x8
  const int cands[5] = {16, 32, 8, 64, 128};
  for (int k = 0; k < 5; ++k)
    if (Wo % cands[k] == 0) { *TW = cands[k]; *TH = 128 / cands[k]; return; }
  *TW = 16; *TH = 8;
}

int of_conv_tc3_try(const of_tc_conv* c, void* stream);   // conv_tc3.cu

extern "C" {

int of_conv_tc_pack_weights(const float* w_hwio, void* packed, int taps, int Cin, int Cout, int rows, int cols,
                            int transpose, int nseg, const int* seg_off, const int* seg_cnt, const int* seg_base,
                            void* stream) {
  OF_REQUIRE(w_hwio && packed && seg_off && seg_cnt && seg_base, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(nseg >= 1 && nseg <= 3 && taps > 0 && rows > 0 && cols > 0 && cols % 64 == 0, OF_ERR_ARG,
             "bad packing geometry");
  PackSeg seg;
  seg.nseg = nseg;
  for (int s = 0; s < 3; ++s) {
    seg.off[s] = s < nseg ? seg_off[s] : 0;
    seg.cnt[s] = s < nseg ? seg_cnt[s] : 0;
    seg.base[s] = s < nseg ? seg_base[s] : 0;
  }
  size_t total = (size_t)taps * rows * cols;
  size_t blocks = (total + 255) / 256;
  if (blocks > (size_t)OF_NUM_SMS * 8) blocks = (size_t)OF_NUM_SMS * 8;
  pack_weights_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(w_hwio, (__nv_bfloat16*)packed, taps, Cin, Cout,
                                                                     rows, cols, transpose, seg);
  return of_launch_check("pack_weights_kernel");
}

int of_conv_tc_pack_weights_batch(const of_pack_job* jobs_device, int njobs, int total_blocks, void* stream) {
  OF_REQUIRE(jobs_device && njobs > 0 && total_blocks > 0, OF_ERR_ARG, "bad job table");
  pack_weights_batch_kernel<<<total_blocks, 256, 0, (cudaStream_t)stream>>>(jobs_device, njobs);
  return of_launch_check("pack_weights_batch_kernel");
}

int of_conv_tc_run(const of_tc_conv* c, void* stream) {
  OF_REQUIRE(c && c->src[0] && c->wpacked && c->out, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(c->num_src == 1 || c->num_src == 2, OF_ERR_ARG, "num_src must be 1 or 2");
  OF_REQUIRE(c->n_mma >= 16 && c->n_mma <= 256 && c->n_mma % 16 == 0 && c->n_valid <= c->n_mma && c->n_valid > 0,
             OF_ERR_SHAPE, "n_mma %d / n_valid %d unsupported", c->n_mma, c->n_valid);
  OF_REQUIRE(c->n_rows >= c->n_valid && c->k_pad % 64 == 0 && of_valid_dtype(c->out_dtype), OF_ERR_SHAPE,
             "bad packed weight dims / out dtype");
  for (int s = 0; s < c->num_src; ++s) OF_REQUIRE(c->src[s] && c->src_c[s] > 0, OF_ERR_ARG, "bad source %d", s);
  of_prof_note(2.0 * c->N * c->Ho * c->Wo * (double)(c->src_c[0] + (c->num_src == 2 ? c->src_c[1] : 0)) * c->n_valid *
               c->KH * c->KW);
  if (int r = of_conv_tc3_try(c, stream)) return r < 0 || r > 1 ? r : 0;    // 3x3 / stride-1 fast path
  OF_REQUIRE(c->out2 == nullptr, OF_ERR_UNSUPPORTED, "split outputs are a feature of the 3x3 / stride-1 kernel (Cout <= 128)");
  OF_REQUIRE(c->add_mode == 0, OF_ERR_UNSUPPORTED, "add_mode 1 is a feature of the 3x3 / stride-1 kernel (Cout <= 128)");
  OF_REQUIRE(c->sign_out == nullptr && !(c->mask != nullptr && c->mask_dtype == 2), OF_ERR_UNSUPPORTED,
             "sign-bit activation masks are a feature of the 3x3 / stride-1 kernel (Cout <= 128)");
  OF_REQUIRE(c->num_src == 1 || c->num_src == 2, OF_ERR_ARG, "num_src must be 1 or 2");
  OF_REQUIRE(c->stride == 1 || c->stride == 2, OF_ERR_UNSUPPORTED, "stride %d unsupported", c->stride);
  OF_REQUIRE(c->n_mma >= 16 && c->n_mma <= 256 && c->n_mma % 16 == 0 && c->n_valid <= c->n_mma && c->n_valid > 0,
             OF_ERR_SHAPE, "n_mma %d / n_valid %d unsupported", c->n_mma, c->n_valid);
  OF_REQUIRE(c->n_rows >= c->n_valid && c->k_pad % 64 == 0, OF_ERR_SHAPE, "bad packed weight dims");
  OF_REQUIRE(of_valid_dtype(c->out_dtype), OF_ERR_ARG, "bad out dtype");
  ConvTcParams P;
  memset(&P, 0, sizeof(P));
  P.num_src = c->num_src;
  int kb = 0;
  for (int s = 0; s < c->num_src; ++s) {
    OF_REQUIRE(c->src[s] && c->src_c[s] > 0, OF_ERR_ARG, "bad source %d", s);
    P.kchunks[s] = (c->src_c[s] + 63) / 64;
    P.kbase[s] = kb;
    kb += P.kchunks[s] * 64;
  }
  OF_REQUIRE(kb <= c->k_pad, OF_ERR_SHAPE, "packed K %d smaller than the sources' %d", c->k_pad, kb);
  P.taps = c->KH * c->KW; P.KW = c->KW; P.pad_t = c->pad_t; P.pad_l = c->pad_l;
  P.stride_w = c->stride_w > 0 ? c->stride_w : c->stride;
  P.stride_h = c->stride_h > 0 ? c->stride_h : c->stride;
  OF_REQUIRE(P.stride_w >= 1 && P.stride_w <= 2 && P.stride_h >= 1 && P.stride_h <= 2, OF_ERR_UNSUPPORTED,
             "conv strides %d x %d unsupported", P.stride_h, P.stride_w);
  const bool view = c->view_w_stride > 0;
  OF_REQUIRE(!view || c->num_src == 1, OF_ERR_ARG, "a strided view takes one source");
  P.sign = c->sign >= 0 ? 1 : -1;
  P.NB = c->N; P.Ho = c->Ho; P.Wo = c->Wo;
  pick_tile(c->Wo, &P.TW, &P.TH);
  P.tiles_w = (c->Wo + P.TW - 1) / P.TW;
  P.tiles_h = (c->Ho + P.TH - 1) / P.TH;
  long long mt = (long long)c->N * P.tiles_w * P.tiles_h;
  OF_REQUIRE(mt > 0 && mt < (1ll << 30), OF_ERR_SHAPE, "bad output tiling");
  P.m_tiles = (int)mt;
  P.n_mma = c->n_mma; P.n_valid = c->n_valid;
  P.b_tile_bytes = ((c->n_mma * 128) + 1023) / 1024 * 1024;
  int stage_bytes = A_TILE_BYTES + P.b_tile_bytes;
  int cols = 2 * c->n_mma;
  P.tmem_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : cols <= 256 ? 256 : 512;
  // several CTAs per SM where TMEM allows it (the kernel needs 58 registers): a second CTA's TMA stream and MMAs
  // fill the bubbles of the first; the stage ring is split between them
  int ctas = 512 / P.tmem_cols;
  {
    const char* e = of_tune_env("OF_TC_CTAS");
    const int cap = e ? atoi(e) : 2;
    if (ctas > cap) ctas = cap;
    if (ctas < 1) ctas = 1;
    while (ctas > 1 && ((200 * 1024) / ctas) / stage_bytes < 3) --ctas;
    if (P.m_tiles <= OF_NUM_SMS) ctas = 1;
  }
  P.stages = ((200 * 1024) / ctas) / stage_bytes;
  if (P.stages > MAX_STAGES) P.stages = MAX_STAGES;
  P.out = c->out; P.out_dtype = c->out_dtype; P.out_stride = c->out_stride;
  P.bias = c->bias; P.act = c->act;
  P.add = c->add; P.add_dtype = c->add_dtype; P.add_stride = c->add_stride;
  P.mask = c->mask; P.mask_dtype = c->mask_dtype; P.mask_stride = c->mask_stride; P.mask_act = c->mask_act;

  CUtensorMap tmA[2], tmB;
  memset(tmA, 0, sizeof(tmA));
  for (int s = 0; s < c->num_src; ++s) {
    long long ws_ = view ? c->view_w_stride : c->src_stride[s];
    long long hs_ = view ? c->view_h_stride : (long long)c->W * c->src_stride[s];
    long long ns_ = view ? c->view_n_stride : (long long)c->H * c->W * c->src_stride[s];
    if (int r = of_tc_make_view_map(&tmA[s], c->src[s], c->src_c[s], ws_, hs_, ns_, c->N, c->H, c->W, P.TW, P.TH,
                                    P.stride_w, P.stride_h))
      return r;
  }
  if (c->num_src == 1) tmA[1] = tmA[0];
  if (int r = of_tc_make_weight_map(&tmB, c->wpacked, P.taps, c->n_rows, c->k_pad, c->n_mma)) return r;

  size_t smem = 1024 + (size_t)P.stages * stage_bytes + (2 * MAX_STAGES + 4) * sizeof(uint64_t) + 16;
  constexpr int G_R = TC_SPEC(OF_ACT_RELU, 0, 0, 0), G_0 = TC_SPEC(0, 0, 0, 0), G_A = TC_SPEC(0, 1, 0, 0),
                G_MR = TC_SPEC(0, 0, 1, OF_ACT_RELU), G_AMR = TC_SPEC(0, 1, 1, OF_ACT_RELU);
  int spec = -1;
  if (c->out_dtype == OF_BF16 && c->n_valid == c->n_mma && (c->out_stride % 16) == 0 && ((uintptr_t)c->out & 31) == 0 &&
      (c->add == nullptr || (c->add_dtype == OF_BF16 && (c->add_stride % 16) == 0 && ((uintptr_t)c->add & 31) == 0)) &&
      (c->mask == nullptr || (c->mask_dtype == OF_BF16 && (c->mask_stride % 16) == 0 && ((uintptr_t)c->mask & 31) == 0)) &&
      !of_tune_env("OF_TC_NO_SPEC"))
    spec = TC_SPEC(c->act, c->add != nullptr ? 1 : 0, c->mask != nullptr ? 1 : 0, c->mask != nullptr ? c->mask_act : 0);
  static bool attr_set = false;
  if (!attr_set) {
    OF_CUDA(cudaFuncSetAttribute(conv_tc_kernel<-1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    OF_CUDA(cudaFuncSetAttribute(conv_tc_kernel<G_R>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));   // + 1 KB static (bias)
    OF_CUDA(cudaFuncSetAttribute(conv_tc_kernel<G_0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));   // + 1 KB static (bias)
    OF_CUDA(cudaFuncSetAttribute(conv_tc_kernel<G_A>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));   // + 1 KB static (bias)
    OF_CUDA(cudaFuncSetAttribute(conv_tc_kernel<G_MR>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));   // + 1 KB static (bias)
    OF_CUDA(cudaFuncSetAttribute(conv_tc_kernel<G_AMR>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    attr_set = true;
  }
  int grid = P.m_tiles < ctas * OF_NUM_SMS ? P.m_tiles : ctas * OF_NUM_SMS;
#define TC_GO(S) OF_CUDA(of_launch_pdl(conv_tc_kernel<S>, dim3(grid), dim3(TC_THREADS), smem, (cudaStream_t)stream, tmA[0], tmA[1], tmB, P))
  if (spec == G_R) TC_GO(G_R);
  else if (spec == G_0) TC_GO(G_0);
  else if (spec == G_A) TC_GO(G_A);
  else if (spec == G_MR) TC_GO(G_MR);
  else if (spec == G_AMR) TC_GO(G_AMR);
  else TC_GO(-1);
#undef TC_GO
  // timeline names: the stem (strided view of the packed image, Cin = 3) is reported apart from the Cin >= 64 layers
  return of_launch_check(view ? "conv_tc_kernel<stem>" : "conv_tc_kernel");
}

}  // extern "C"
