// K8 on the tensor cores: cost volume (model.py:29-42, max_disp = 3) forward and backward as BANDED
// GEMMs on tcgen05 / TMEM, bf16 operands, fp32 accumulation.
//
// A tile is 128 consecutive pixels of one image row (b, y, x0 .. x0+127); its "neighbour window" is the
// 144 pixels x0-3 .. x0+140 of another row (134 used; TMA zero fill outside the image == tf.pad).
//
//   forward   for each row displacement i (7 steps):   D[m][n] = sum_c f1[y, x0+m, c] * f2[y+i-3, x0-3+n, c]
//             cv[y, x0+m, i*7 + j] = D[m][m + j]       (the band of a 128 x 144 product, K = C)
//   backward  d_f1[y, x0+m, :] = sum_i  G_i[m][:] . f2[y+i-3, window, :]   with  G_i[m][m+j] = g[y, x0+m, i*7+j]
//             d_f2[y, x0+m, :] = sum_i  G'_i[m][:] . f1[y-i+3, window, :]  with  G'_i[m][n] = g[y-i+3, x0-3+n, i*7 + m-n+6]
//             (a 128 x 144 band matrix times a 144 x C feature window, accumulated over the 7 steps in TMEM)
//
// The dense products do 144/7 = 20x the useful multiply-adds, but the tensor pipe is > 60x faster than the
// FMA path the SIMT kernels (costvol7.cu) use, and every feature value is read from L2 once per step
// instead of being staged through shared memory by LSU instructions.
//
// Warp roles (320 threads): warp 0 TMA producer, warp 1 MMA issuer (+ TMEM owner), warps 2-5 band-matrix
// builders (backward) / idle (forward), warps 6-9 epilogue.  mbarrier rings between them; two TMEM
// accumulators so the epilogue of one tile / step overlaps the MMAs of the next.
#include <stdlib.h>
#include "common.cuh"
#include "tc_common.cuh"

// 32 bytes in one instruction = one full L2 sector (p 32-byte aligned)
__device__ __forceinline__ void st256(void* p, const uint32_t (&pk)[8]) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]),
               "r"(pk[3]), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7])
               : "memory");
}

#define CT_THREADS 320
#define CT_M 128
#define CT_NB 144
#define CT_G_BYTES 49152      // band matrix: 3 K-chunks x 128 rows x 128 B (K-major, 128-byte swizzle)
#define CT_ROW_BYTES 18432    // one 64-channel chunk of a 144-pixel feature window
#define CT_A_BYTES 16384      // one 64-channel chunk of the 128-pixel f1 tile (forward)
#define CT_MAX_STAGES 4

int of_tc_make_view_map(CUtensorMap* m, const void* base, int c_used, long long w_stride, long long h_stride,
                        long long n_stride, int N, int H, int W, int box_w, int box_h, int sw, int sh);

struct CvTcParams {
  int B, H, W, C, nchunks, tiles_x;
  int ntiles;
  int stages, stage_bytes, a_bufs;
  int tmem_cols;
  int from_f2;                          // backward: 1 = d_f1 (features = f2), 0 = d_f2 (features = f1)
  int cv_stride;                        // pixel stride (elements) of cv / d_cv
  const __nv_bfloat16* d_cv;            // backward
  __nv_bfloat16* out;                   // backward: (B,H,W,C); forward: cv (B,H,W,cv_stride)
  const __nv_bfloat16* add;             // backward, nullable
};

__device__ __forceinline__ void ct_decode(const CvTcParams& P, int tile, int& b, int& y, int& x0) {
  int tx = tile % P.tiles_x; int r = tile / P.tiles_x;
  y = r % P.H; b = r / P.H; x0 = tx * CT_M;
}

// band-matrix element (row m, K index k) of a K-major SWIZZLE_128B tile set: chunk = k / 64
__device__ __forceinline__ void g_store(uint8_t* G, int m, int k, uint16_t v) {
  uint8_t* p = G + (k >> 6) * 16384 + m * 128 + ((((k & 63) >> 3) ^ (m & 7)) << 4) + (k & 7) * 2;
  *reinterpret_cast<uint16_t*>(p) = v;
}

// ------------------------------------------------------------------------------ backward --
__global__ void __launch_bounds__(CT_THREADS, 1)
cvtc_bwd_kernel(const __grid_constant__ CUtensorMap tmF, const __grid_constant__ CvTcParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* b_full = reinterpret_cast<uint64_t*>(base + (size_t)P.stages * P.stage_bytes);
  uint64_t* g_full = b_full + CT_MAX_STAGES;
  uint64_t* empty = g_full + CT_MAX_STAGES;
  uint64_t* tmem_full = empty + CT_MAX_STAGES;     // [2]
  uint64_t* tmem_empty = tmem_full + 2;            // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P.stages; ++s) { tc::mbar_init(&b_full[s], 1); tc::mbar_init(&g_full[s], 128); tc::mbar_init(&empty[s], 1); }
    for (int k = 0; k < 2; ++k) { tc::mbar_init(&tmem_full[k], 1); tc::mbar_init(&tmem_empty[k], 4); }
    tc::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) tc::prefetch_tmap(&tmF);
  if (warp == 1) { tc::tmem_alloc(tmem_slot, (uint32_t)P.tmem_cols); tc::tmem_relinquish(); }
  // the band matrices are zero outside the band for the whole kernel: clear them once
  for (int s = 0; s < P.stages; ++s)
    for (int i = threadIdx.x; i < CT_G_BYTES / 16; i += CT_THREADS)
      reinterpret_cast<uint4*>(base + (size_t)s * P.stage_bytes)[i] = make_uint4(0, 0, 0, 0);
  tc::fence_proxy_async();
  if (threadIdx.x == 0) OF_PDL_TRIGGER();
  OF_PDL_WAIT();                                     // nothing above touches global memory
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer: one feature-row window per step =====================
    if (OF_ONE_LANE()) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x) {
        int b, y, x0; ct_decode(P, tile, b, y, x0);
        for (int i = 0; i < 7; ++i) {
          const int ys = P.from_f2 ? y + i - 3 : y - i + 3;
          tc::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* dst = base + (size_t)stage * P.stage_bytes + CT_G_BYTES;
          tc::mbar_expect_tx(&b_full[stage], (uint32_t)(P.nchunks * CT_ROW_BYTES));
          for (int c = 0; c < P.nchunks; ++c)
            tc::tma_load_4d(dst + (size_t)c * CT_ROW_BYTES, &tmF, &b_full[stage], c * 64, x0 - 3, ys, b);
          if (++stage == P.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (OF_ONE_LANE()) {
      const uint32_t idesc = tc::idesc_bf16_f32(128, P.C, 0, 1);
      int stage = 0; uint32_t phase = 0; int it = 0;
      for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x, ++it) {
        const int buf = it & 1; const uint32_t use = (uint32_t)(it >> 1);
        tc::mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);
        tc::fence_after_sync();
        const uint32_t d = tmem_base + (uint32_t)(buf * P.C);
        for (int i = 0; i < 7; ++i) {
          tc::mbar_wait(&b_full[stage], phase);
          tc::mbar_wait(&g_full[stage], phase);
          tc::fence_after_sync();
          const uint32_t gaddr = tc::smem_u32(base + (size_t)stage * P.stage_bytes);
          const uint32_t baddr = gaddr + CT_G_BYTES;
#pragma unroll
          for (int s = 0; s < CT_NB / 16; ++s) {
            const uint64_t adesc = tc::desc_sw128_kmajor(gaddr + (uint32_t)((s >> 2) * 16384 + (s & 3) * 32));
            const uint64_t bdesc = tc::desc_sw128_mnmajor(baddr + (uint32_t)(s * 2048), CT_ROW_BYTES);
            tc::mma_bf16(d, adesc, bdesc, idesc, (i | s) ? 1u : 0u);
          }
          tc::mma_commit(&empty[stage]);
          if (++stage == P.stages) { stage = 0; phase ^= 1; }
        }
        tc::mma_commit(&tmem_full[buf]);
      }
    }
  } else if (warp < 6) {
    // ===================== band-matrix builders (128 threads) =====================
    const int tb = threadIdx.x - 64;
    int stage = 0; uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x) {
      int b, y, x0; ct_decode(P, tile, b, y, x0);
      // All gradient values this thread needs for the tile's 7 steps are fetched up front (the loads are
      // independent, their latency overlaps the pipeline); per step only shared-memory stores remain.
      //   from_f2 : own pixel (y, x0 + tb), all 49 values (7 x 16 B)
      //   else    : window column tb (and tb + 128 for tb < 6) of row y - i + 3, the 7 values of step i
      //             = 14 B at byte offset 14 i -> the four 32-bit words from word (14 i - 2 (i & 1)) / 4
      uint32_t w[28], wa[7][4], wb[7][4];
      if (P.from_f2) {
        const int x = x0 + tb;
        if (x < P.W) {
          const uint4* src = reinterpret_cast<const uint4*>(P.d_cv + (((size_t)b * P.H + y) * P.W + x) * P.cv_stride);
#pragma unroll
          for (int q = 0; q < 7; ++q) {                 // 49 bf16 = 98 B: the first 7 x 16 B cover them
            uint4 v = __ldg(src + q);
            w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
          }
        } else {
#pragma unroll
          for (int q = 0; q < 28; ++q) w[q] = 0;
        }
      } else {
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          const int ys = y - i + 3;
          const bool row_ok = ys >= 0 && ys < P.H;
          const int wofs = (14 * i - 2 * (i & 1)) / 4;
          const int xa = x0 - 3 + tb, xb = xa + 128;
          const bool oka = row_ok && xa >= 0 && xa < P.W;
          const bool okb = row_ok && tb < 6 && xb < P.W;
          // the 4-word window starts at word wofs of the pixel's 128-byte row: one or two 16-byte chunks instead of
          // four 4-byte loads (every request is a sector operation in L2, and this kernel is L2-bound)
          const int q0 = wofs >> 2, sh = wofs & 3;
          const uint4* pa = reinterpret_cast<const uint4*>(P.d_cv + (((size_t)b * P.H + (row_ok ? ys : 0)) * P.W + (oka ? xa : 0)) * P.cv_stride) + q0;
          const uint4* pb = reinterpret_cast<const uint4*>(P.d_cv + (((size_t)b * P.H + (row_ok ? ys : 0)) * P.W + (okb ? xb : 0)) * P.cv_stride) + q0;
          const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
          const uint4 alo = oka ? __ldg(pa) : z4, ahi = (oka && sh) ? __ldg(pa + 1) : z4;
          const uint4 blo = okb ? __ldg(pb) : z4, bhi = (okb && sh) ? __ldg(pb + 1) : z4;
          const uint32_t ta[8] = {alo.x, alo.y, alo.z, alo.w, ahi.x, ahi.y, ahi.z, ahi.w};
          const uint32_t tb8[8] = {blo.x, blo.y, blo.z, blo.w, bhi.x, bhi.y, bhi.z, bhi.w};
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            wa[i][q] = ta[sh + q];
            wb[i][q] = tb8[sh + q];
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 7; ++i) {
        // (one waiter + one arrival per builder warp instead of 128 was tried in round 2: d_f1 355 -> 451 us, d_f2 481 ->
        //  578 us - the extra __syncwarp pair per step costs more than the spinning it saves)
        tc::mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* G = base + (size_t)stage * P.stage_bytes;
        if (P.from_f2) {
#pragma unroll
          for (int j = 0; j < 7; ++j) {
            const int e = i * 7 + j;
            const uint16_t v = (uint16_t)((e & 1) ? (w[e >> 1] >> 16) : (w[e >> 1] & 0xffffu));
            g_store(G, tb, tb + j, v);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 7; ++j) {
            const int h = (i & 1) + j;                               // half-word inside the 4-word window
            const uint16_t va = (uint16_t)((h & 1) ? (wa[i][h >> 1] >> 16) : (wa[i][h >> 1] & 0xffffu));
            const int ma = tb - 6 + j;
            if (ma >= 0) g_store(G, ma, tb, va);                     // ma <= tb <= 127
            if (tb < 6) {
              const uint16_t vb = (uint16_t)((h & 1) ? (wb[i][h >> 1] >> 16) : (wb[i][h >> 1] & 0xffffu));
              const int mb = tb + 128 - 6 + j;
              if (mb < CT_M) g_store(G, mb, tb + 128, vb);
            }
          }
        }
        tc::fence_proxy_async();                                    // generic-proxy stores -> visible to the MMA
        tc::mbar_arrive(&g_full[stage]);
        if (++stage == P.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue warps 6..9 =====================
    const int q = warp & 3;
    const int m = q * 32 + lane;
    int it = 0;
    for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x, ++it) {
      const int buf = it & 1; const uint32_t use = (uint32_t)(it >> 1);
      int b, y, x0; ct_decode(P, tile, b, y, x0);
      const int x = x0 + m;
      const bool valid = x < P.W;
      const size_t off = (((size_t)b * P.H + y) * P.W + (valid ? x : 0)) * P.C;
      tc::mbar_wait(&tmem_full[buf], use & 1);
      tc::fence_after_sync();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * P.C);
      for (int c0 = 0; c0 < P.C; c0 += 16) {
        __syncwarp();
        uint32_t rr[16];
        tc::tmem_ld16(taddr + (uint32_t)c0, rr);
        tc::tmem_ld_wait();
        if (valid) {
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rr[j]);
          if (P.add != nullptr) {
            const uint4* ap = reinterpret_cast<const uint4*>(P.add + off + c0);
            uint4 a0 = __ldg(ap), a1 = __ldg(ap + 1);
            const uint32_t aw[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[2 * j] += __uint_as_float(aw[j] << 16);
              v[2 * j + 1] += __uint_as_float(aw[j] & 0xffff0000u);
            }
          }
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            __nv_bfloat162 t2 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
            pk[j] = *reinterpret_cast<uint32_t*>(&t2);
          }
          uint4* op = reinterpret_cast<uint4*>(P.out + off + c0);
          if ((reinterpret_cast<uintptr_t>(op) & 31) == 0) {
            st256(op, pk);                               // one full 32-byte L2 sector per instruction
          } else {
            op[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            op[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
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
  if (warp == 1) tc::tmem_dealloc(tmem_base, (uint32_t)P.tmem_cols);
}

// ------------------------------------------------------------------------------- forward --
// shared memory: [a_bufs][nchunks][16 KB] f1 tiles, then [stages][nchunks][18 KB] f2 row windows
__global__ void __launch_bounds__(CT_THREADS, 1)
cvtc_fwd_kernel(const __grid_constant__ CUtensorMap tmF1, const __grid_constant__ CUtensorMap tmF2,
                const __grid_constant__ CvTcParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int a_bytes = P.nchunks * CT_A_BYTES;
  uint8_t* a_tiles = base;
  uint8_t* b_ring = base + (size_t)P.a_bufs * a_bytes;
  uint64_t* b_full = reinterpret_cast<uint64_t*>(b_ring + (size_t)P.stages * P.stage_bytes);
  uint64_t* b_empty = b_full + CT_MAX_STAGES;
  uint64_t* a_full = b_empty + CT_MAX_STAGES;      // [2]
  uint64_t* a_empty = a_full + 2;                  // [2]
  uint64_t* tmem_full = a_empty + 2;               // [2]
  uint64_t* tmem_empty = tmem_full + 2;            // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P.stages; ++s) { tc::mbar_init(&b_full[s], 1); tc::mbar_init(&b_empty[s], 1); }
    for (int k = 0; k < 2; ++k) {
      tc::mbar_init(&a_full[k], 1); tc::mbar_init(&a_empty[k], 1);
      tc::mbar_init(&tmem_full[k], 1); tc::mbar_init(&tmem_empty[k], 4);
    }
    tc::fence_barrier_init();
  }
  if (warp == 0 && lane == 0) { tc::prefetch_tmap(&tmF1); tc::prefetch_tmap(&tmF2); }
  if (warp == 1) { tc::tmem_alloc(tmem_slot, 512); tc::tmem_relinquish(); }
  if (threadIdx.x == 0) OF_PDL_TRIGGER();
  OF_PDL_WAIT();                                     // nothing above touches global memory
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (OF_ONE_LANE()) {
      int stage = 0; uint32_t phase = 0; int it = 0;
      for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x, ++it) {
        int b, y, x0; ct_decode(P, tile, b, y, x0);
        const int ab = it % P.a_bufs; const uint32_t ause = (uint32_t)(it / P.a_bufs);
        tc::mbar_wait(&a_empty[ab], (ause & 1) ^ 1);
        tc::mbar_expect_tx(&a_full[ab], (uint32_t)a_bytes);
        for (int c = 0; c < P.nchunks; ++c)
          tc::tma_load_4d(a_tiles + (size_t)ab * a_bytes + (size_t)c * CT_A_BYTES, &tmF1, &a_full[ab], c * 64, x0, y, b);
        for (int i = 0; i < 7; ++i) {
          tc::mbar_wait(&b_empty[stage], phase ^ 1);
          uint8_t* dst = b_ring + (size_t)stage * P.stage_bytes;
          tc::mbar_expect_tx(&b_full[stage], (uint32_t)(P.nchunks * CT_ROW_BYTES));
          for (int c = 0; c < P.nchunks; ++c)
            tc::tma_load_4d(dst + (size_t)c * CT_ROW_BYTES, &tmF2, &b_full[stage], c * 64, x0 - 3, y + i - 3, b);
          if (++stage == P.stages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (OF_ONE_LANE()) {
      const uint32_t idesc = tc::idesc_bf16_f32(128, CT_NB, 0, 0);
      int stage = 0; uint32_t phase = 0; int it = 0; uint32_t step = 0;
      for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x, ++it) {
        const int ab = it % P.a_bufs; const uint32_t ause = (uint32_t)(it / P.a_bufs);
        tc::mbar_wait(&a_full[ab], ause & 1);
        tc::fence_after_sync();
        const uint32_t aaddr = tc::smem_u32(a_tiles + (size_t)ab * a_bytes);
        for (int i = 0; i < 7; ++i, ++step) {
          const int buf = (int)(step & 1); const uint32_t use = step >> 1;
          tc::mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);
          tc::mbar_wait(&b_full[stage], phase);
          tc::fence_after_sync();
          const uint32_t baddr = tc::smem_u32(b_ring + (size_t)stage * P.stage_bytes);
          const uint32_t d = tmem_base + (uint32_t)(buf * CT_NB);
          for (int c = 0; c < P.nchunks; ++c) {
            const uint64_t adesc = tc::desc_sw128_kmajor(aaddr + (uint32_t)(c * CT_A_BYTES));
            const uint64_t bdesc = tc::desc_sw128_kmajor(baddr + (uint32_t)(c * CT_ROW_BYTES));
#pragma unroll
            for (int k = 0; k < 4; ++k)
              tc::mma_bf16(d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (c | k) ? 1u : 0u);
          }
          tc::mma_commit(&b_empty[stage]);
          tc::mma_commit(&tmem_full[buf]);
          if (++stage == P.stages) { stage = 0; phase ^= 1; }
        }
        tc::mma_commit(&a_empty[ab]);
      }
    }
  } else if (warp >= 6) {
    // (a third TMEM accumulator in flight - 3 x 144 columns - was tried in round 2, call 60: 268.7 vs 269.0 us per step,
    // no change; the kernel is bound by the MMAs of the banded formulation, not by the accumulator hand-over)
    // epilogue: row m of step i needs D[m][m + j], j = 0..6 - columns q*32 + lane + j of this warp's
    // 48-column slice; a 5-stage predicated shift network brings them to registers 0..6
    const int q = warp & 3;
    const int m = q * 32 + lane;
    uint32_t step = 0;
    for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x) {
      int b, y, x0; ct_decode(P, tile, b, y, x0);
      const int x = x0 + m;
      const bool valid = x < P.W;
      uint32_t ow[25];                                       // 49 bf16 results, packed in pairs
#pragma unroll
      for (int k = 0; k < 25; ++k) ow[k] = 0;
#pragma unroll
      for (int i = 0; i < 7; ++i, ++step) {
        const int buf = (int)(step & 1); const uint32_t use = step >> 1;
        tc::mbar_wait(&tmem_full[buf], use & 1);
        tc::fence_after_sync();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * CT_NB + q * 32);
        uint32_t r0[16], r1[16], r2[16];
        __syncwarp();
        tc::tmem_ld16(taddr, r0);
        tc::tmem_ld16(taddr + 16, r1);
        tc::tmem_ld16(taddr + 32, r2);
        tc::tmem_ld_wait();
        tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&tmem_empty[buf]);    // values are in registers: release the accumulator
        uint32_t r[48];
#pragma unroll
        for (int k = 0; k < 16; ++k) { r[k] = r0[k]; r[16 + k] = r1[k]; r[32 + k] = r2[k]; }
        if (lane & 16) {
#pragma unroll
          for (int k = 0; k < 22; ++k) r[k] = r[k + 16];
        }
        if (lane & 8) {
#pragma unroll
          for (int k = 0; k < 14; ++k) r[k] = r[k + 8];
        }
        if (lane & 4) {
#pragma unroll
          for (int k = 0; k < 10; ++k) r[k] = r[k + 4];
        }
        if (lane & 2) {
#pragma unroll
          for (int k = 0; k < 8; ++k) r[k] = r[k + 2];
        }
        if (lane & 1) {
#pragma unroll
          for (int k = 0; k < 7; ++k) r[k] = r[k + 1];
        }
#pragma unroll
        for (int j = 0; j < 7; ++j) {
          const int e = i * 7 + j;
          const uint32_t h = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(__uint_as_float(r[j])));
          ow[e >> 1] |= (e & 1) ? (h << 16) : h;
        }
      }
      if (valid) {
        __nv_bfloat16* o = P.out + (((size_t)b * P.H + y) * P.W + x) * P.cv_stride;
        uint4* o4 = reinterpret_cast<uint4*>(o);
#pragma unroll
        for (int k = 0; k < 6; k += 2) {
          if ((reinterpret_cast<uintptr_t>(o4) & 31) == 0) {
            const uint32_t w8[8] = {ow[4 * k], ow[4 * k + 1], ow[4 * k + 2], ow[4 * k + 3],
                                    ow[4 * k + 4], ow[4 * k + 5], ow[4 * k + 6], ow[4 * k + 7]};
            st256(o4 + k, w8);
          } else {
            o4[k] = make_uint4(ow[4 * k], ow[4 * k + 1], ow[4 * k + 2], ow[4 * k + 3]);
            o4[k + 1] = make_uint4(ow[4 * k + 4], ow[4 * k + 5], ow[4 * k + 6], ow[4 * k + 7]);
          }
        }
        *reinterpret_cast<uint16_t*>(o + 48) = (uint16_t)(ow[24] & 0xffffu);
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------------------------- host --
#define CT_SMEM_BUDGET (220 * 1024)

static bool cvtc_applicable(const void* a, const void* b, const void* c, int C, int cv_stride) {
  if (of_tune_env("OF_CV_NO_TC")) return false;
  if (C % 64 != 0 || C > 256 || cv_stride != 64) return false;
  return (((uintptr_t)a | (uintptr_t)b | (uintptr_t)c) & 15) == 0;
}

static int cvtc_set_attr() {
  static bool set = false;
  if (!set) {
    OF_CUDA(cudaFuncSetAttribute(cvtc_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    OF_CUDA(cudaFuncSetAttribute(cvtc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    set = true;
  }
  return 0;
}

static void cvtc_common(CvTcParams& P, int B, int H, int W, int C, int cv_stride) {
  memset(&P, 0, sizeof(P));
  P.B = B; P.H = H; P.W = W; P.C = C; P.nchunks = C / 64;
  P.tiles_x = (W + CT_M - 1) / CT_M;
  P.ntiles = B * H * P.tiles_x;
  P.cv_stride = cv_stride;
}

// Returns 1 if handled, 0 if not applicable (caller falls back to costvol7.cu), other values = error.
int cvtc_fwd_try(const void* f1, const void* f2, void* cv, int B, int H, int W, int C, int cv_stride, cudaStream_t st) {
  if (!cvtc_applicable(f1, f2, cv, C, cv_stride)) return 0;
  if ((long long)B * H * ((W + CT_M - 1) / CT_M) >= (1ll << 30)) return 0;
  if (int r = cvtc_set_attr()) return r;
  CvTcParams P;
  cvtc_common(P, B, H, W, C, cv_stride);
  P.out = (__nv_bfloat16*)cv;
  P.a_bufs = C <= 128 ? 2 : 1;
  P.stage_bytes = P.nchunks * CT_ROW_BYTES;
  P.stages = (CT_SMEM_BUDGET - P.a_bufs * P.nchunks * CT_A_BYTES) / P.stage_bytes;
  if (P.stages > CT_MAX_STAGES) P.stages = CT_MAX_STAGES;
  if (P.stages < 1) return 0;
  CUtensorMap tm1, tm2;
  if (int r = of_tc_make_view_map(&tm1, f1, C, C, (long long)W * C, (long long)H * W * C, B, H, W, CT_M, 1, 1, 1)) return r;
  if (int r = of_tc_make_view_map(&tm2, f2, C, C, (long long)W * C, (long long)H * W * C, B, H, W, CT_NB, 1, 1, 1)) return r;
  size_t smem = 1024 + (size_t)P.a_bufs * P.nchunks * CT_A_BYTES + (size_t)P.stages * P.stage_bytes + 256;
  int grid = P.ntiles < OF_NUM_SMS ? P.ntiles : OF_NUM_SMS;
  OF_CUDA(of_launch_pdl(cvtc_fwd_kernel, dim3(grid), dim3(CT_THREADS), smem, st, tm1, tm2, P));
  if (int r = of_launch_check("cvtc_fwd_kernel")) return r;
  return 1;
}

int cvtc_bwd_try(const void* f1, const void* f2, const void* d_cv, void* d_f1, void* d_f2, const void* d_f1_add, int B,
                 int H, int W, int C, int cv_stride, cudaStream_t st) {
  if (!cvtc_applicable(f1, f2, d_cv, C, cv_stride)) return 0;
  if ((((uintptr_t)d_f1 | (uintptr_t)d_f2 | (uintptr_t)d_f1_add) & 15) != 0) return 0;
  if ((long long)B * H * ((W + CT_M - 1) / CT_M) >= (1ll << 30)) return 0;
  if (int r = cvtc_set_attr()) return r;
  CvTcParams P;
  cvtc_common(P, B, H, W, C, cv_stride);
  P.d_cv = (const __nv_bfloat16*)d_cv;
  P.stage_bytes = CT_G_BYTES + P.nchunks * CT_ROW_BYTES;
  P.stages = CT_SMEM_BUDGET / P.stage_bytes;
  if (P.stages > CT_MAX_STAGES) P.stages = CT_MAX_STAGES;
  if (P.stages < 1) return 0;
  P.tmem_cols = 2 * C < 32 ? 32 : 2 * C;                   // 128 / 256 / 512
  size_t smem = 1024 + (size_t)P.stages * P.stage_bytes + 256;
  int grid = P.ntiles < OF_NUM_SMS ? P.ntiles : OF_NUM_SMS;
  CUtensorMap tm;
  // d_f1: features = f2
  if (int r = of_tc_make_view_map(&tm, f2, C, C, (long long)W * C, (long long)H * W * C, B, H, W, CT_NB, 1, 1, 1)) return r;
  P.from_f2 = 1; P.out = (__nv_bfloat16*)d_f1; P.add = (const __nv_bfloat16*)d_f1_add;
  OF_CUDA(of_launch_pdl(cvtc_bwd_kernel, dim3(grid), dim3(CT_THREADS), smem, st, tm, P));
  if (int r = of_launch_check("cvtc_bwd_kernel<d_f1>")) return r;
  // d_f2: features = f1
  if (int r = of_tc_make_view_map(&tm, f1, C, C, (long long)W * C, (long long)H * W * C, B, H, W, CT_NB, 1, 1, 1)) return r;
  P.from_f2 = 0; P.out = (__nv_bfloat16*)d_f2; P.add = nullptr;
  OF_CUDA(of_launch_pdl(cvtc_bwd_kernel, dim3(grid), dim3(CT_THREADS), smem, st, tm, P));
  if (int r = of_launch_check("cvtc_bwd_kernel<d_f2>")) return r;
  return 1;
}
