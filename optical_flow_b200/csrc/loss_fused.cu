// K6 fast path: fused 4-scale resize + warp + L1 photometric loss, forward and d_flow, ONE launch
// (loss.py:5-32 with model.py:55-73 warp semantics), for H, W multiples of 16.
//
// Dataflow through L2 inside one persistent (co-resident) grid:
//   BUILD units  stream the full-resolution 6-channel batch exactly once: a TMA bulk copy
//                (cp.async.bulk + mbarrier, 3-stage ring, one producer warp) lands 16 rows x 64 pixels in shared memory;
//                every resized pixel of all four scales that this 16x16-aligned patch owns (the
//                central 2x2 mean tf.image.resize computes at scale 2^k) is formed from shared
//                memory and written to a float4 pyramid that is kept in L2 (evict_last), the
//                image lines are streamed (evict_first).
//   GATHER units (32x32 output pixels of one scale; the flow tile arrives through the same TMA
//                ring, so its HBM latency is hidden) read flow, the frame-1 pyramid pixel and
//                the 4 bilinear taps of the frame-2 pyramid (one LDG.128 per tap, lanes mapped
//                so a warp's taps fall in few cache lines even for the reference's transposed
//                gather), and write d_flow and the loss partial.
//   A per-image counter (release/acquire) orders gather(b) after build(b); unit order is
//   build(0), build(1), gather(0), build(2), gather(1), ... so the HBM stream of image b+1
//   overlaps the L2-bound gather of image b and nobody spins in steady state.
// HBM traffic = image pair once + flows + d_flows (the algorithmic bytes of SURVEY 8d).
#include <stdlib.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "loss_fused_common.cuh"

#define LF_MAX_SCALES 4
#define LF_ROWS 32           // full-resolution rows per build unit
#define LF_COLS 32           // full-resolution columns per build unit
#define LF_ROW_BYTES (LF_COLS * 24)
#define LF_STAGE_BYTES (LF_ROWS * LF_ROW_BYTES)   // 24576: one TMA box {192 floats, 32 rows}
#define LF_STAGES 4
#define LF_CONSUMERS 256
#define LF_THREADS (LF_CONSUMERS + 64)   // 8 consumer warps + TMA producer warp + publisher warp
#define LF_GT_I 32           // gather tile: 32 x 32 output pixels, 4 per thread
#define LF_GT_J 32
#define LF_FLOW_ROW_BYTES 272   // 32 float2 + 16 B pad (bank spread; keeps 16-byte TMA alignment)
#define LF_CTAS_PER_SM 2
#define LF_MAX_GRID (OF_NUM_SMS * 4)
#define LF_HDR_BYTES 65536   // [0,16K): ticket, per-image counters, per-CTA partials; [16K,32K): generic kernel's header;
                             // [32K,64K): loss_fused2.cu's per-CTA relay flags (one 128-byte line per CTA)

struct LfScale {
  const float* flow;
  float* d_flow;
  float4* p1;      // resized frame 1 (B,h,w) as (r,g,b,0)
  float4* p2;      // resized frame 2
  int h, w;
  int gt_i, gt_j;  // gather tiles along i (rows) and j (columns)
  int g_begin;     // first gather unit of this scale within an image
  float inv_n;     // 1 / (B*h*w*3*num_scales)
};

struct LfMaps {
  CUtensorMap img;                    // (B, H, W/16, 96) f32 (16-pixel segments), box {96, 2, 32, 1}: a 32 x 32 pixel patch
  CUtensorMap flow[LF_MAX_SCALES];    // (B, h, w*2) f32, box {68, 32, 1}: 32 x 32 flow vectors (+ pad)
};

struct LfParams {
  LfScale sc[LF_MAX_SCALES];
  int ns;
  const float* imgs;
  float* loss;
  unsigned* ticket;      // [0]
  unsigned* done;        // [B] build units finished per image
  double* partials;      // [grid]
  int B, H, W, mode;
  int bands, chunks;     // H/16, ceil(W/64)
  int nb, ng;            // build / gather units per image
  int lead;              // images built ahead of the first gather
  int prows, pcols;      // patch shape (32 x 32; other shapes only for the dbg streaming experiment)
  int dbg;               // profiling only: 1 skip build math, 2 skip gather math, 4 skip flag polling
  int total_units;
};

__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(LF_CONSUMERS) : "memory"); }

// unit index -> (is_build, image, local index).  Order: build(0..L-1), then pairs
// [build(L+k), gather(k)] for k < B-L, then gather(B-L..B-1): the pyramid of an image is complete
// L = lead images' worth of work before anybody gathers from it.
__device__ __forceinline__ void decode_unit(const LfParams& P, int u, bool* build, int* b, int* local) {
  const int head = P.lead * P.nb;
  if (u < head) { *build = true; *b = u / P.nb; *local = u - *b * P.nb; return; }
  int v = u - head;
  const int per = P.nb + P.ng, pairs = P.B - P.lead;
  if (v < pairs * per) {
    int k = v / per, r = v - k * per;
    if (r < P.nb) { *build = true; *b = P.lead + k; *local = r; }
    else { *build = false; *b = k; *local = r - P.nb; }
    return;
  }
  v -= pairs * per;
  int k = v / P.ng;
  *build = false; *b = pairs + k; *local = v - k * P.ng;
}

// one resized pixel of BOTH frames from the staged patch: the central 2x2 block starting at
// (r0, c0).  VEC: c0 is even (scale 0) so the 2 pixels x 6 floats are 16-byte aligned.
__device__ __forceinline__ void lf_build_pixel(const bool VEC, const float* T, int r0, int c0, float4* o1,
                                               float4* o2) {
  const float* a = T + r0 * (LF_ROW_BYTES / 4) + c0 * 6;
  const float* c = a + (LF_ROW_BYTES / 4);
  float ra[12], rc[12];
  if (VEC) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float4 u = reinterpret_cast<const float4*>(a)[k], v = reinterpret_cast<const float4*>(c)[k];
      ra[4 * k] = u.x; ra[4 * k + 1] = u.y; ra[4 * k + 2] = u.z; ra[4 * k + 3] = u.w;
      rc[4 * k] = v.x; rc[4 * k + 1] = v.y; rc[4 * k + 2] = v.z; rc[4 * k + 3] = v.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      float2 u = reinterpret_cast<const float2*>(a)[k], v = reinterpret_cast<const float2*>(c)[k];
      ra[2 * k] = u.x; ra[2 * k + 1] = u.y;
      rc[2 * k] = v.x; rc[2 * k + 1] = v.y;
    }
  }
  *o1 = make_float4(lf_mean4(ra[0], ra[6], rc[0], rc[6]), lf_mean4(ra[1], ra[7], rc[1], rc[7]),
                    lf_mean4(ra[2], ra[8], rc[2], rc[8]), 0.f);
  *o2 = make_float4(lf_mean4(ra[3], ra[9], rc[3], rc[9]), lf_mean4(ra[4], ra[10], rc[4], rc[10]),
                    lf_mean4(ra[5], ra[11], rc[5], rc[11]), 0.f);
}

__global__ void __launch_bounds__(LF_THREADS, LF_CTAS_PER_SM) loss_fused_kernel(const __grid_constant__ LfParams P, const __grid_constant__ LfMaps M) {
  extern __shared__ __align__(128) uint8_t lf_smem[];
  uint8_t* tiles = lf_smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(lf_smem + LF_STAGES * LF_STAGE_BYTES);
  uint64_t* empty_bar = full_bar + LF_STAGES;
  uint64_t* pub_bar = empty_bar + LF_STAGES;     // consumers -> publisher warp
  __shared__ double s_part[LF_CONSUMERS / 32];
  __shared__ bool s_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < LF_STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], LF_CONSUMERS / 32 + 1);   // 8 consumer warps + the publisher
      tc::mbar_init(&pub_bar[s], LF_CONSUMERS / 32);
    }
    tc::fence_barrier_init();
  }
  __syncthreads();

  float acc = 0.f;
  if (warp == LF_CONSUMERS / 32 && lane == 0) {
    tc::prefetch_tmap(&M.img);
    for (int k = 0; k < P.ns; ++k) tc::prefetch_tmap(&M.flow[k]);
  }
  if (warp == LF_CONSUMERS / 32) {
    // ============ producer warp: TMA bulk loads (image patches for build units, flow tiles for
    // gather units) into the shared-memory ring; one stage per unit ============
    const uint64_t pol = policy_evict_first();
    int stage = 0; uint32_t phase = 0;
    for (int u = blockIdx.x; u < P.total_units; u += gridDim.x) {
      bool build; int b, local;
      decode_unit(P, u, &build, &b, &local);
      uint8_t* dst = tiles + (size_t)stage * LF_STAGE_BYTES;
      if (lane == 0) tc::mbar_wait(&empty_bar[stage], phase ^ 1);
      __syncwarp();
      if (lane == 0) {
        if (build) {
          int band = local / P.chunks, cx = local - band * P.chunks;
          tc::mbar_expect_tx(&full_bar[stage], (uint32_t)(P.prows * P.pcols * 24));   // OOB is zero-filled
          tma_load_4d_hint(dst, &M.img, &full_bar[stage], 0, cx * (P.pcols / 16), band * P.prows, b, pol);
        } else {
          int s = 0;
#pragma unroll
          for (int k = 1; k < LF_MAX_SCALES; ++k)
            if (k < P.ns && local >= P.sc[k].g_begin) s = k;
          const LfScale& S = P.sc[s];
          int lt = local - S.g_begin;
          int ti = lt / S.gt_j, tj = lt - ti * S.gt_j;
          bool staged = (S.w & 1) == 0;
          tc::mbar_expect_tx(&full_bar[stage], staged ? (uint32_t)(LF_GT_I * LF_FLOW_ROW_BYTES) : 0u);
          if (staged) tma_load_3d_hint(dst, &M.flow[s], &full_bar[stage], tj * LF_GT_J * 2, ti * LF_GT_I, b, pol);
        }
      }
      if (++stage == LF_STAGES) { stage = 0; phase ^= 1; }
    }
  } else if (warp == LF_CONSUMERS / 32 + 1) {
    // ============ publisher warp: one thread turns "all 8 consumer warps finished build unit n"
    // (mbarrier, CTA-scope release/acquire) into a gpu-scope release of done[image]: the ~1 us
    // membar stalls nobody who computes.  Units that completed meanwhile share one fence. ============
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      int u = blockIdx.x;
      while (u < P.total_units) {
        tc::mbar_wait(&pub_bar[stage], phase);
        int imgs_done[LF_STAGES]; int k = 0;
        do {
          bool build; int b, local;
          decode_unit(P, u, &build, &b, &local);
          if (build) imgs_done[k++] = b;
          tc::mbar_arrive(&empty_bar[stage]);
          if (++stage == LF_STAGES) { stage = 0; phase ^= 1; }
          u += gridDim.x;
        } while (u < P.total_units && k < LF_STAGES && mbar_test(&pub_bar[stage], phase));
        if (k) {
          asm volatile("fence.acq_rel.gpu;" ::: "memory");
          for (int i = 0; i < k; ++i)
            asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(P.done + imgs_done[i]) : "memory");
        }
      }
    }
  } else {
    // ===================== consumers =====================
    const uint64_t keep = policy_evict_last();
    int stage = 0; uint32_t phase = 0;
    int confirmed = -1;                 // images whose pyramid this CTA has already seen complete
    for (int u = blockIdx.x; u < P.total_units; u += gridDim.x) {
      bool build; int b, local;
      decode_unit(P, u, &build, &b, &local);
      const uint8_t* Tb = tiles + (size_t)stage * LF_STAGE_BYTES;
      if (build) {
        int band = local / P.chunks, cx = local - band * P.chunks;
        const int wc = min(LF_COLS, P.W - cx * LF_COLS), hr = min(LF_ROWS, P.H - band * LF_ROWS);
        if (lane == 0) tc::mbar_wait(&full_bar[stage], phase);   // one poller per warp
        __syncwarp();
        const float* T = reinterpret_cast<const float*>(Tb);
        // items enumerated on the full 32 x 32 patch grid (shifts, no divisions); rows / columns beyond
        // the image (last band / chunk) are skipped.  scale s: (32 >> (s+1))^2 resized pixels.
        int t = (OF_DBG(P) & 1) ? 1 << 20 : tid;
#pragma unroll
        for (int s = 0; s < LF_MAX_SCALES; ++s) {
          const int lc = 4 - s;                         // log2(columns at scale s)
          const int n_s = (LF_ROWS >> (s + 1)) << lc;
          if (s < P.ns) {
            const LfScale& S = P.sc[s];
            const int bs = 2 << s;
            for (; t < n_s; t += LF_CONSUMERS) {
              int orow = t >> lc, ocol = t & ((1 << lc) - 1);
              if (ocol * bs < wc && orow * bs < hr) {
                float4 o1, o2;
                lf_build_pixel(s == 0, T, bs * orow + (bs >> 1) - 1, bs * ocol + (bs >> 1) - 1, &o1, &o2);
                unsigned pix = ((unsigned)b * S.h + band * (LF_ROWS / bs) + orow) * S.w + cx * (LF_COLS / bs) + ocol;
                st_f4_policy(S.p1 + pix, o1, keep);
                st_f4_policy(S.p2 + pix, o2, keep);
              }
            }
          }
          t -= n_s;
        }
        // warps are decoupled (no CTA barrier): each warp hands the unit to the publisher warp (which
        // makes the pyramid stores visible gpu-wide off the critical path) and frees the ring slot
        __syncwarp();
        if (lane == 0) { tc::mbar_arrive(&pub_bar[stage]); tc::mbar_arrive(&empty_bar[stage]); }
      } else {
        // ---- gather unit: 32 x 32 output pixels of one scale, 4 per thread (2 in flight) ----
        if (b > confirmed && !(OF_DBG(P) & 4)) {           // first gather unit of image b in this CTA: wait for its pyramid
          if (lane == 0)
          {
            while (ld_relaxed(&P.done[b]) < (unsigned)P.nb) __nanosleep(200);
            asm volatile("fence.acq_rel.gpu;" ::: "memory");
          }
          __syncwarp();
          confirmed = b;
        }
        int s = 0;
#pragma unroll
        for (int k = 1; k < LF_MAX_SCALES; ++k)
          if (k < P.ns && local >= P.sc[k].g_begin) s = k;
        const LfScale& S = P.sc[s];
        int lt = local - S.g_begin;
        int ti = lt / S.gt_j, tj = lt - ti * S.gt_j;
        int di, dj;
        if (P.mode == OF_WARP_REF_IJ) { di = (warp & 1) * 8 + (lane & 7); dj = (warp >> 1) * 4 + (lane >> 3); }
        else { di = warp * 2 + (lane >> 4); dj = lane & 15; }
        const int h = S.h, w = S.w;
        const bool staged = (w & 1) == 0;
        const unsigned img_px = (unsigned)b * h * w;
        const float4* q2 = S.p2 + img_px;
        const float4* q1 = S.p1 + img_px;
        const float2* gflow = reinterpret_cast<const float2*>(S.flow) + img_px;
        float2* gdflow = S.d_flow ? reinterpret_cast<float2*>(S.d_flow) + img_px : nullptr;
        if (lane == 0) tc::mbar_wait(&full_bar[stage], phase);   // one poller per warp
        __syncwarp();
#pragma unroll 1
        for (int q = (OF_DBG(P) & 2) ? 2 : 0; q < 2; ++q) {
          const int lj = dj + 16 * q, j = tj * LF_GT_J + lj;
          int ii[2]; bool ok[2]; float2 fl[2];
#pragma unroll
          for (int p = 0; p < 2; ++p) {
            const int li = di + 16 * p;
            ii[p] = ti * LF_GT_I + li;
            ok[p] = ii[p] < h && j < w;
            fl[p] = make_float2(0.f, 0.f);
            if (ok[p])
              fl[p] = staged ? *reinterpret_cast<const float2*>(Tb + li * LF_FLOW_ROW_BYTES + lj * 8)
                             : __ldg(gflow + ii[p] * w + j);
          }
          float ax[2], ay[2], bx[2], by[2];
          float4 v00[2], v01[2], v10[2], v11[2], ref[2];
#pragma unroll
          for (int p = 0; p < 2; ++p) {
            if (!ok[p]) continue;
            // model.py:65-71 + transformations.py:93-125 (same op order as loss.cu / warp.cu)
            float x, y;
            if (P.mode == OF_WARP_REF_IJ) { x = __fadd_rn((float)ii[p], fl[p].x); y = __fadd_rn((float)j, fl[p].y); }
            else { x = __fadd_rn((float)j, fl[p].x); y = __fadd_rn((float)ii[p], fl[p].y); }
            int xf = __float2int_rd(x), yf = __float2int_rd(y);
            int x0 = lf_clamp(xf, 0, w - 1), x1 = lf_clamp(xf == 0x7fffffff ? xf : xf + 1, 0, w - 1);
            int y0 = lf_clamp(yf, 0, h - 1), y1 = lf_clamp(yf == 0x7fffffff ? yf : yf + 1, 0, h - 1);
            ax[p] = __fsub_rn((float)x1, x);
            ay[p] = __fsub_rn((float)y1, y);
            bx[p] = __fsub_rn(1.f, ax[p]);
            by[p] = __fsub_rn(1.f, ay[p]);
            v00[p] = ld_f4_policy(q2 + y0 * w + x0, keep);
            v01[p] = ld_f4_policy(q2 + y1 * w + x0, keep);
            v10[p] = ld_f4_policy(q2 + y0 * w + x1, keep);
            v11[p] = ld_f4_policy(q2 + y1 * w + x1, keep);
            ref[p] = ld_f4_policy(q1 + ii[p] * w + j, keep);
          }
#pragma unroll
          for (int p = 0; p < 2; ++p) {
            if (!ok[p]) continue;
            float w00 = ax[p] * ay[p], w01 = ax[p] * by[p], w10 = bx[p] * ay[p], w11 = bx[p] * by[p];
            float wx = w00 * v00[p].x + w01 * v01[p].x + w10 * v10[p].x + w11 * v11[p].x;
            float wy = w00 * v00[p].y + w01 * v01[p].y + w10 * v10[p].y + w11 * v11[p].y;
            float wz = w00 * v00[p].z + w01 * v01[p].z + w10 * v10[p].z + w11 * v11[p].z;
            float ex = ref[p].x - wx, ey = ref[p].y - wy, ez = ref[p].z - wz;                 // loss.py:28
            acc += (fabsf(ex) + fabsf(ey) + fabsf(ez)) * S.inv_n;
            if (gdflow != nullptr) {
              float gxx = ay[p] * (v10[p].x - v00[p].x) + by[p] * (v11[p].x - v01[p].x);
              float gxy = ay[p] * (v10[p].y - v00[p].y) + by[p] * (v11[p].y - v01[p].y);
              float gxz = ay[p] * (v10[p].z - v00[p].z) + by[p] * (v11[p].z - v01[p].z);
              float gyx = ax[p] * (v01[p].x - v00[p].x) + bx[p] * (v11[p].x - v10[p].x);
              float gyy = ax[p] * (v01[p].y - v00[p].y) + bx[p] * (v11[p].y - v10[p].y);
              float gyz = ax[p] * (v01[p].z - v00[p].z) + bx[p] * (v11[p].z - v10[p].z);
              float sx_ = -lf_sgn(ex) * S.inv_n, sy_ = -lf_sgn(ey) * S.inv_n, sz_ = -lf_sgn(ez) * S.inv_n;
              float d0 = sx_ * gxx + sy_ * gxy + sz_ * gxz;
              float d1 = sx_ * gyx + sy_ * gyy + sz_ * gyz;
              gdflow[ii[p] * w + j] = make_float2(d0, d1);
            }
          }
        }
        __syncwarp();                  // flow tile no longer read by this warp
        if (lane == 0) { tc::mbar_arrive(&pub_bar[stage]); tc::mbar_arrive(&empty_bar[stage]); }
      }
      if (++stage == LF_STAGES) { stage = 0; phase ^= 1; }
    }
  }

  // deterministic reduction: per-CTA double partial, last CTA adds them in index order
  double d = (double)acc;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
  if (lane == 0 && warp < LF_CONSUMERS / 32) s_part[warp] = d;
  __syncthreads();
  if (tid == 0) {
    double cta = 0.0;
#pragma unroll
    for (int k = 0; k < LF_CONSUMERS / 32; ++k) cta += s_part[k];
    P.partials[blockIdx.x] = cta;
    __threadfence();
    unsigned t = atomicAdd(P.ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    // every CTA has finished all of its units: re-arm the per-image counters and the ticket, and add
    // the per-CTA partials in a fixed (deterministic) order - in parallel: a one-thread loop over
    // ~300 L2 round trips costs more than the rest of the kernel
    for (int b = tid; b < P.B; b += LF_THREADS) P.done[b] = 0u;
    __shared__ double s_red[LF_CONSUMERS];
    if (tid < LF_CONSUMERS) {
      __threadfence();
      double v = 0.0;
      for (unsigned i = tid; i < gridDim.x; i += LF_CONSUMERS) v += reinterpret_cast<volatile double*>(P.partials)[i];
      s_red[tid] = v;
    }
    __syncthreads();
    for (int o = LF_CONSUMERS / 2; o > 0; o >>= 1) {
      if (tid < o) s_red[tid] += s_red[tid + o];
      __syncthreads();
    }
    if (tid == 0) {
      *P.loss = (float)s_red[0];
      *P.ticket = 0u;
    }
  }
}

// ---- host side ------------------------------------------------------------------------------
int of_tc_get_encode(void** fn);   // conv_tc.cu: cuTensorMapEncodeTiled through the runtime's driver entry point

typedef CUresult (*LfEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// (B, rows, row_floats) f32 tensor, box {box_floats, box_rows, 1}, no swizzle, zero fill outside
static int lf_make_map(CUtensorMap* m, const void* base, int B, int rows, long long row_floats, int box_floats,
                       int box_rows) {
  void* fn;
  if (int r = of_tc_get_encode(&fn)) return r;
  OF_REQUIRE(((uintptr_t)base & 15) == 0 && (row_floats * 4) % 16 == 0, OF_ERR_ARG,
             "TMA needs 16-byte aligned base and row pitch");
  cuuint64_t dims[3] = {(cuuint64_t)row_floats, (cuuint64_t)rows, (cuuint64_t)B};
  cuuint64_t strides[2] = {(cuuint64_t)row_floats * 4, (cuuint64_t)rows * row_floats * 4};
  cuuint32_t box[3] = {(cuuint32_t)box_floats, (cuuint32_t)box_rows, 1};
  cuuint32_t es[3] = {1, 1, 1};
  CUresult r = ((LfEncodeTiled)fn)(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(base), dims, strides, box,
                                   es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  OF_REQUIRE(r == CUDA_SUCCESS, OF_ERR_STATE, "cuTensorMapEncodeTiled(loss) failed: %d", (int)r);
  return 0;
}

size_t of_lf_workspace_bytes(int B, int H, int W, int num_scales) {
  size_t px = 0;
  for (int s = 0; s < num_scales && s < LF_MAX_SCALES; ++s) px += (size_t)B * (H >> (s + 1)) * (W >> (s + 1));
  return LF_HDR_BYTES + px * 2 * sizeof(float4);
}

bool of_lf_supported(int B, int H, int W, int num_scales, const int* hs, const int* ws) {
  if (num_scales < 1 || num_scales > LF_MAX_SCALES || H % 16 || W % 16 || B < 1 || B > 2048) return false;
  for (int s = 0; s < num_scales; ++s)
    if (hs[s] != (H >> (s + 1)) || ws[s] != (W >> (s + 1))) return false;
  return true;
}

int of_lf_run(const float* imgs, const float* const* flows, float* const* d_flows, const int* hs, const int* ws,
              int num_scales, float* loss, int B, int H, int W, int mode, void* workspace, void* stream) {
  LfParams P;
  memset(&P, 0, sizeof(P));
  uint8_t* wsb = reinterpret_cast<uint8_t*>(workspace);
  OF_REQUIRE(((uintptr_t)wsb & 15) == 0 && ((uintptr_t)imgs & 15) == 0, OF_ERR_ARG,
             "workspace and images must be 16-byte aligned");
  P.ticket = reinterpret_cast<unsigned*>(wsb);
  P.done = reinterpret_cast<unsigned*>(wsb + 64);                         // B <= 2048 -> 8 KB
  P.partials = reinterpret_cast<double*>(wsb + 64 + 2048 * 4);            // LF_MAX_GRID doubles
  float4* pyr = reinterpret_cast<float4*>(wsb + LF_HDR_BYTES);
  int ng = 0;
  for (int s = 0; s < num_scales; ++s) {
    LfScale& S = P.sc[s];
    S.flow = flows[s];
    S.d_flow = d_flows ? d_flows[s] : nullptr;
    S.h = hs[s]; S.w = ws[s];
    size_t px = (size_t)B * S.h * S.w;
    S.p1 = pyr; S.p2 = pyr + px; pyr += 2 * px;
    S.gt_i = (S.h + LF_GT_I - 1) / LF_GT_I;
    S.gt_j = (S.w + LF_GT_J - 1) / LF_GT_J;
    S.g_begin = ng;
    ng += S.gt_i * S.gt_j;
    S.inv_n = (float)(1.0 / ((double)B * S.h * S.w * 3.0 * num_scales));
  }
  P.ns = num_scales;
  P.imgs = imgs; P.loss = loss;
  P.B = B; P.H = H; P.W = W; P.mode = mode;
  P.prows = LF_ROWS; P.pcols = LF_COLS;
  if (const char* e = of_tune_env("OF_LF_SHAPE")) {            // streaming experiment (use with OF_LF_DBG=7)
    int r_ = 0, c_ = 0;
    if (sscanf(e, "%d,%d", &r_, &c_) == 2 && r_ > 0 && c_ > 0 && c_ % 16 == 0 && r_ * c_ * 24 <= LF_STAGE_BYTES) {
      P.prows = r_; P.pcols = c_;
    }
  }
  P.bands = (H + P.prows - 1) / P.prows;
  P.chunks = (W + P.pcols - 1) / P.pcols;
  P.nb = P.bands * P.chunks;
  P.ng = ng;
  P.lead = B < 4 ? B : 4;   // measured on B200 at C2: lead 1 / 2 / 3 / 4 -> 84 / 64 / 55 / 53 us
  if (const char* e = of_tune_env("OF_LF_DBG")) P.dbg = atoi(e);
  if (const char* e = of_tune_env("OF_LF_LEAD")) { int v = atoi(e); if (v >= 1 && v <= B) P.lead = v; }
  long long total = (long long)B * (P.nb + P.ng);
  OF_REQUIRE(total < (1ll << 30), OF_ERR_SHAPE, "too many loss work units");
  P.total_units = (int)total;

  LfMaps M;
  memset(&M, 0, sizeof(M));
  {
    void* fn;
    if (int r = of_tc_get_encode(&fn)) return r;
    cuuint64_t dims[4] = {96, (cuuint64_t)(W / 16), (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t strides[3] = {384, (cuuint64_t)W * 24, (cuuint64_t)H * W * 24};
    cuuint32_t box[4] = {96, (cuuint32_t)(P.pcols / 16), (cuuint32_t)P.prows, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = ((LfEncodeTiled)fn)(&M.img, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(imgs), dims, strides,
                                     box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    OF_REQUIRE(r == CUDA_SUCCESS, OF_ERR_STATE, "cuTensorMapEncodeTiled(images) failed: %d", (int)r);
  }
  for (int s = 0; s < num_scales; ++s) {
    M.flow[s] = M.img;                                    // unstaged scales never touch their map
    if ((ws[s] & 1) == 0)
      if (int r = lf_make_map(&M.flow[s], flows[s], B, hs[s], (long long)ws[s] * 2, LF_FLOW_ROW_BYTES / 4, LF_GT_I))
        return r;
  }
  for (int s = num_scales; s < LF_MAX_SCALES; ++s) M.flow[s] = M.img;

  const size_t smem = (size_t)LF_STAGES * LF_STAGE_BYTES + 3 * LF_STAGES * sizeof(uint64_t);
  static int max_ctas = 0;
  if (max_ctas == 0) {
    OF_CUDA(cudaFuncSetAttribute(loss_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0, dev = 0, sms = 0;
    OF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, loss_fused_kernel, LF_THREADS, smem));
    OF_CUDA(cudaGetDevice(&dev));
    OF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    OF_REQUIRE(per_sm >= 1, OF_ERR_STATE, "fused loss kernel does not fit on an SM");
    max_ctas = per_sm * sms;
    if (max_ctas > LF_MAX_GRID) max_ctas = LF_MAX_GRID;
  }
  int grid = P.total_units < max_ctas ? P.total_units : max_ctas;
  if (const char* e = of_tune_env("OF_LF_GRID")) { int v = atoi(e); if (v >= 1 && v < grid) grid = v; }
  // co-residency is required (gather units wait on build units of other CTAs): cooperative launch
  // makes the runtime enforce it
  void* args[2] = {(void*)&P, (void*)&M};
  OF_CUDA(cudaLaunchCooperativeKernel((const void*)loss_fused_kernel, dim3(grid), dim3(LF_THREADS), args, smem,
                                      (cudaStream_t)stream));
  return of_launch_check("loss_fused_kernel");
}
