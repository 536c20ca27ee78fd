// K6 fast path, warp-specialised generation: fused 4-scale resize + warp + L1 photometric loss, forward and d_flow,
// ONE launch (loss.py:5-32 with model.py:55-73 warp semantics), for H, W multiples of 16.
//
// Same dataflow through L2 as loss_fused.cu (the image pair is streamed from HBM exactly once, the float4 pyramid of
// both frames lives in the 126 MB L2 between its producer and its consumers), but the two halves of the work no
// longer take turns inside a CTA.  One persistent CTA per SM runs, side by side,
//   warp 0          TMA producer: 32 x 32 pixel patches of the 6-channel pair into an 8-stage shared-memory ring
//                   (24 KB per stage, mbarrier full / empty per stage, evict_first);
//   warp 1          watcher: the only poller of the global per-image "pyramid complete" counters; it republishes
//                   them to a shared-memory word the gather warps spin on (no L2 hot spot);
//   warps 2..5      BUILD: each warp owns every fourth patch of its CTA: all four pyramid levels of both frames
//                   (tf.image.resize at 2^k = the central 2x2 mean) from shared memory -> float4 texels, evict_last;
//                   then ONE gpu-scope release + counter increment per patch;
//   warps 6..       GATHER: each warp owns its own 128-pixel output tiles (static round-robin over the whole grid, so
//                   the loss sum stays deterministic): flow prefetched one tile ahead, 4 bilinear taps of frame 2
//                   and the frame-1 texel as LDG.128 from the L2-resident pyramid, |diff| -> loss partial, sign ->
//                   d_flow.  A gather tile of image b only waits until all patches of image b are built.
// So the HBM stream (BUILD) and the L2-latency-bound taps (GATHER) overlap on every SM for the whole kernel; only the
// first image's build and the last image's gather are exposed.  No CTA-wide barrier after start-up.
// HBM traffic = image pair once + flows + d_flows (the algorithmic bytes of SURVEY 8d).
#include "common.cuh"
#include "tc_common.cuh"
#include "loss_fused_common.cuh"

#define L2F_MAX_SCALES 4
#define L2F_ROWS 32
#define L2F_COLS 32
#define L2F_ROW_FLOATS (L2F_COLS * 6)
#define L2F_STAGE_BYTES (L2F_ROWS * L2F_COLS * 24)      // 24576: one TMA box {96 floats, 2 segments, 32 rows}
#define L2F_STAGES 8
#define L2F_BUILD_WARPS 4
#ifndef L2F_GATHER_WARPS
#define L2F_GATHER_WARPS 18
#endif
#define L2F_FIRST_BUILD 2
#define L2F_FIRST_GATHER (L2F_FIRST_BUILD + L2F_BUILD_WARPS)
#define L2F_WARPS (L2F_FIRST_GATHER + L2F_GATHER_WARPS)
#define L2F_THREADS (L2F_WARPS * 32)
#define L2F_HDR_BYTES 32768      // same header layout as loss_fused.cu: ticket, per-image counters, per-CTA partials

struct L2fScale {
  const float* flow;
  float* d_flow;
  float4* p1;      // resized frame 1 (B,h,w) as (r,g,b,0)
  float4* p2;      // resized frame 2
  int h, w;
  int t_i, t_j;    // gather tiles along i (rows) and j (columns)
  int u_begin;     // first gather tile of this scale within an image
  float inv_n;     // 1 / (B*h*w*3*num_scales)
};

struct L2fParams {
  L2fScale sc[L2F_MAX_SCALES];
  int ns;
  float* loss;
  unsigned* ticket;      // [0]
  unsigned* done;        // [B] patches built per image
  double* partials;      // [grid]
  int B, H, W, mode;
  int chunks;            // patches per row of patches
  int nb;                // patches per image
  int ngu;               // gather tiles per image
  int tile_i, tile_j;    // gather tile shape (REF_IJ 8 x 16, XY 4 x 32)
  int total_build, total_gather;
  int dbg;               // -DOF_PROFILE_HOOKS only: 1 no build math / stores, 2 no gather, 4 gather does not wait for
                         // the builds, 8 no release fence + counter, 16 no image stream at all (with 4: gather only)
};

__device__ __forceinline__ void st_release_cta_shared(unsigned* p, unsigned v) {
  asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" ::"r"(tc::smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_cta_shared(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(tc::smem_u32(p)) : "memory");
  return v;
}

// one resized pixel of BOTH frames from the staged patch: the central 2x2 block whose top-left is (r0, c0).
// VEC: c0 is even (scale 0), so the 2 pixels x 6 floats are 16-byte aligned.  Same op order as loss_fused.cu.
__device__ __forceinline__ void l2f_build_pixel(const bool VEC, const float* T, int r0, int c0, float4* o1, float4* o2) {
  const float* a = T + r0 * L2F_ROW_FLOATS + c0 * 6;
  const float* c = a + L2F_ROW_FLOATS;
  float ra[12], rc[12];
  if (VEC) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float4 u = reinterpret_cast<const float4*>(a)[k], v = reinterpret_cast<const float4*>(c)[k];
      ra[4 * k] = u.x; ra[4 * k + 1] = u.y; ra[4 * k + 2] = u.z; ra[4 * k + 3] = u.w;
      rc[4 * k] = v.x; rc[4 * k + 1] = v.y; rc[4 * k + 2] = v.z; rc[4 * k + 3] = v.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const float2 u = reinterpret_cast<const float2*>(a)[k], v = reinterpret_cast<const float2*>(c)[k];
      ra[2 * k] = u.x; ra[2 * k + 1] = u.y;
      rc[2 * k] = v.x; rc[2 * k + 1] = v.y;
    }
  }
  *o1 = make_float4(lf_mean4(ra[0], ra[6], rc[0], rc[6]), lf_mean4(ra[1], ra[7], rc[1], rc[7]),
                    lf_mean4(ra[2], ra[8], rc[2], rc[8]), 0.f);
  *o2 = make_float4(lf_mean4(ra[3], ra[9], rc[3], rc[9]), lf_mean4(ra[4], ra[10], rc[4], rc[10]),
                    lf_mean4(ra[5], ra[11], rc[5], rc[11]), 0.f);
}

// gather tile index -> (image, scale, tile row, tile column)
struct L2fTile { int b, s, ti, tj; };
__device__ __forceinline__ L2fTile l2f_decode(const L2fParams& P, int u) {
  L2fTile t;
  t.b = u / P.ngu;
  const int local = u - t.b * P.ngu;
  t.s = 0;
#pragma unroll
  for (int k = 1; k < L2F_MAX_SCALES; ++k)
    if (k < P.ns && local >= P.sc[k].u_begin) t.s = k;
  const int lt = local - P.sc[t.s].u_begin, tj_n = P.sc[t.s].t_j;
  t.ti = lt / tj_n;
  t.tj = lt - t.ti * tj_n;
  return t;
}

// pixel q (0..3) of this lane within a tile
__device__ __forceinline__ void l2f_pixel(const L2fParams& P, const L2fTile& t, int lane, int q, int* i, int* j) {
  if (P.mode == OF_WARP_REF_IJ) {       // 8 x 16 tile; an instruction covers 8 rows x 4 columns: the transposed gather
    *i = t.ti * 8 + (lane & 7);         // (source column = i) then reads 8 consecutive texels of 4 source rows
    *j = t.tj * 16 + (lane >> 3) + 4 * q;
  } else {                              // 4 x 32 tile, one output row per instruction
    *i = t.ti * 4 + q;
    *j = t.tj * 32 + lane;
  }
}

__device__ __forceinline__ void l2f_load_flows(const L2fParams& P, const L2fTile& t, int lane, float2 (&fl)[4]) {
  const L2fScale& S = P.sc[t.s];
  const float2* gflow = reinterpret_cast<const float2*>(S.flow) + (size_t)t.b * S.h * S.w;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    int i, j;
    l2f_pixel(P, t, lane, q, &i, &j);
    fl[q] = (i < S.h && j < S.w) ? __ldg(gflow + i * S.w + j) : make_float2(0.f, 0.f);
  }
}

__global__ void __launch_bounds__(L2F_THREADS, 1)
loss_ws_kernel(const __grid_constant__ L2fParams P, const __grid_constant__ CUtensorMap img_map) {
  extern __shared__ __align__(128) uint8_t l2f_smem[];
  uint8_t* tiles = l2f_smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(l2f_smem + L2F_STAGES * L2F_STAGE_BYTES);
  uint64_t* empty_bar = full_bar + L2F_STAGES;
  __shared__ double s_part[L2F_GATHER_WARPS];
  __shared__ unsigned s_built;          // images whose pyramid is complete and visible (watcher -> gather warps)
  __shared__ bool s_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < L2F_STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    s_built = 0u;
    tc::fence_barrier_init();
  }
  __syncthreads();

  float acc = 0.f;
  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0 && !(OF_DBG(P) & 16)) {
      tc::prefetch_tmap(&img_map);
      const uint64_t pol = policy_evict_first();
      int k = 0;
      for (int p = blockIdx.x; p < P.total_build; p += gridDim.x, ++k) {
        const int stage = k % L2F_STAGES;
        const uint32_t phase = (uint32_t)(k / L2F_STAGES) & 1u;
        tc::mbar_wait(&empty_bar[stage], phase ^ 1u);
        const int b = p / P.nb, local = p - b * P.nb;
        const int band = local / P.chunks, cx = local - band * P.chunks;
        tc::mbar_expect_tx(&full_bar[stage], (uint32_t)L2F_STAGE_BYTES);          // OOB rows / columns are zero-filled
        tma_load_4d_hint(tiles + (size_t)stage * L2F_STAGE_BYTES, &img_map, &full_bar[stage], 0, cx * (L2F_COLS / 16),
                         band * L2F_ROWS, b, pol);
      }
    }
  } else if (warp == 1) {
    // ===================== watcher =====================
    if (lane == 0 && !(OF_DBG(P) & (4 | 8 | 16))) {
      for (int b = 0; b < P.B; ++b) {
        while (ld_relaxed(&P.done[b]) < (unsigned)P.nb) __nanosleep(128);
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
        st_release_cta_shared(&s_built, (unsigned)(b + 1));
      }
    }
  } else if (warp < L2F_FIRST_GATHER) {
    // ===================== BUILD =====================
    const int bw = warp - L2F_FIRST_BUILD;
    const uint64_t keep = policy_evict_last();
    int k = 0;
    for (int p = blockIdx.x; p < ((OF_DBG(P) & 16) ? 0 : P.total_build); p += gridDim.x, ++k) {
      if (k % L2F_BUILD_WARPS != bw) continue;
      const int stage = k % L2F_STAGES;
      const uint32_t phase = (uint32_t)(k / L2F_STAGES) & 1u;
      const int b = p / P.nb, local = p - b * P.nb;
      const int band = local / P.chunks, cx = local - band * P.chunks;
      const int wc = min(L2F_COLS, P.W - cx * L2F_COLS), hr = min(L2F_ROWS, P.H - band * L2F_ROWS);
      if (lane == 0) tc::mbar_wait(&full_bar[stage], phase);
      __syncwarp();
      const float* T = reinterpret_cast<const float*>(tiles + (size_t)stage * L2F_STAGE_BYTES);
#pragma unroll
      for (int s = 0; s < L2F_MAX_SCALES; ++s) {
        if (s < P.ns && !(OF_DBG(P) & 1)) {
          const L2fScale& S = P.sc[s];
          const int lc = 4 - s, bs = 2 << s;                 // log2(columns at scale s), block size
          const int n_s = (L2F_ROWS >> (s + 1)) << lc;
          const unsigned base = ((unsigned)b * S.h + band * (L2F_ROWS / bs)) * S.w + cx * (L2F_COLS / bs);
          for (int t = lane; t < n_s; t += 32) {
            const int orow = t >> lc, ocol = t & ((1 << lc) - 1);
            if (ocol * bs < wc && orow * bs < hr) {
              float4 o1, o2;
              l2f_build_pixel(s == 0, T, bs * orow + (bs >> 1) - 1, bs * ocol + (bs >> 1) - 1, &o1, &o2);
              const unsigned pix = base + orow * S.w + ocol;
              st_f4_policy(S.p1 + pix, o1, keep);
              st_f4_policy(S.p2 + pix, o2, keep);
            }
          }
        }
      }
      __syncwarp();                       // every lane's shared-memory reads and pyramid stores are issued
      if (lane == 0) {
        tc::mbar_arrive(&empty_bar[stage]);                       // the ring slot refills while the release drains
        if (!(OF_DBG(P) & 8)) {
          asm volatile("fence.acq_rel.gpu;" ::: "memory");        // the warp's stores, cumulatively, before the count
          asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(P.done + b) : "memory");
        }
      }
    }
  } else {
    // ===================== GATHER =====================
    const uint64_t keep = policy_evict_last();
    const int gw = (warp - L2F_FIRST_GATHER) * gridDim.x + blockIdx.x;
    const int tgw = L2F_GATHER_WARPS * gridDim.x;
    int confirmed = 0;                  // images known complete
    float2 nfl[4];
    L2fTile nt;
    const int total_gather = (OF_DBG(P) & 2) ? 0 : P.total_gather;
    if (gw < total_gather) { nt = l2f_decode(P, gw); l2f_load_flows(P, nt, lane, nfl); }
    for (int u = gw; u < total_gather; u += tgw) {
      const L2fTile t = nt;
      float2 fl[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) fl[q] = nfl[q];
      if (u + tgw < total_gather) { nt = l2f_decode(P, u + tgw); l2f_load_flows(P, nt, lane, nfl); }   // in flight
      if (t.b >= confirmed && !(OF_DBG(P) & 4)) {
        unsigned v;
        while ((v = ld_acquire_cta_shared(&s_built)) <= (unsigned)t.b) __nanosleep(64);
        confirmed = (int)v;
      }
      const L2fScale& S = P.sc[t.s];
      const int h = S.h, w = S.w;
      const unsigned img_px = (unsigned)t.b * h * w;
      const float4* q2 = S.p2 + img_px;
      const float4* q1 = S.p1 + img_px;
      float2* gdflow = S.d_flow ? reinterpret_cast<float2*>(S.d_flow) + img_px : nullptr;
      const float inv_n = S.inv_n;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        int ii[2], jj[2]; bool ok[2];
        float ax[2], ay[2], bx[2], by[2];
        float4 v00[2], v01[2], v10[2], v11[2], ref[2];
#pragma unroll
        for (int p = 0; p < 2; ++p) {
          l2f_pixel(P, t, lane, 2 * half + p, &ii[p], &jj[p]);
          ok[p] = ii[p] < h && jj[p] < w;
          if (!ok[p]) continue;
          // model.py:65-71 + transformations.py:93-125 (same op order as loss.cu / warp.cu)
          const float2 f = fl[2 * half + p];
          float x, y;
          if (P.mode == OF_WARP_REF_IJ) { x = __fadd_rn((float)ii[p], f.x); y = __fadd_rn((float)jj[p], f.y); }
          else { x = __fadd_rn((float)jj[p], f.x); y = __fadd_rn((float)ii[p], f.y); }
          const int xf = __float2int_rd(x), yf = __float2int_rd(y);
          const int x0 = lf_clamp(xf, 0, w - 1), x1 = lf_clamp(xf == 0x7fffffff ? xf : xf + 1, 0, w - 1);
          const int y0 = lf_clamp(yf, 0, h - 1), y1 = lf_clamp(yf == 0x7fffffff ? yf : yf + 1, 0, h - 1);
          ax[p] = __fsub_rn((float)x1, x);
          ay[p] = __fsub_rn((float)y1, y);
          bx[p] = __fsub_rn(1.f, ax[p]);
          by[p] = __fsub_rn(1.f, ay[p]);
          v00[p] = ld_f4_policy(q2 + y0 * w + x0, keep);
          v01[p] = ld_f4_policy(q2 + y1 * w + x0, keep);
          v10[p] = ld_f4_policy(q2 + y0 * w + x1, keep);
          v11[p] = ld_f4_policy(q2 + y1 * w + x1, keep);
          ref[p] = ld_f4_policy(q1 + ii[p] * w + jj[p], keep);
        }
#pragma unroll
        for (int p = 0; p < 2; ++p) {
          if (!ok[p]) continue;
          const float w00 = ax[p] * ay[p], w01 = ax[p] * by[p], w10 = bx[p] * ay[p], w11 = bx[p] * by[p];
          const float wx = w00 * v00[p].x + w01 * v01[p].x + w10 * v10[p].x + w11 * v11[p].x;
          const float wy = w00 * v00[p].y + w01 * v01[p].y + w10 * v10[p].y + w11 * v11[p].y;
          const float wz = w00 * v00[p].z + w01 * v01[p].z + w10 * v10[p].z + w11 * v11[p].z;
          const float ex = ref[p].x - wx, ey = ref[p].y - wy, ez = ref[p].z - wz;               // loss.py:28
          acc += (fabsf(ex) + fabsf(ey) + fabsf(ez)) * inv_n;
          if (gdflow != nullptr) {
            const float gxx = ay[p] * (v10[p].x - v00[p].x) + by[p] * (v11[p].x - v01[p].x);
            const float gxy = ay[p] * (v10[p].y - v00[p].y) + by[p] * (v11[p].y - v01[p].y);
            const float gxz = ay[p] * (v10[p].z - v00[p].z) + by[p] * (v11[p].z - v01[p].z);
            const float gyx = ax[p] * (v01[p].x - v00[p].x) + bx[p] * (v11[p].x - v10[p].x);
            const float gyy = ax[p] * (v01[p].y - v00[p].y) + bx[p] * (v11[p].y - v10[p].y);
            const float gyz = ax[p] * (v01[p].z - v00[p].z) + bx[p] * (v11[p].z - v10[p].z);
            const float sx_ = -lf_sgn(ex) * inv_n, sy_ = -lf_sgn(ey) * inv_n, sz_ = -lf_sgn(ez) * inv_n;
            const float d0 = sx_ * gxx + sy_ * gxy + sz_ * gxz;
            const float d1 = sx_ * gyx + sy_ * gyy + sz_ * gyz;
            gdflow[ii[p] * w + jj[p]] = make_float2(d0, d1);
          }
        }
      }
    }
  }

  // deterministic reduction: per-warp double (static tile assignment), per-CTA sum in warp order, the last CTA adds
  // the per-CTA partials in index order
  double d = (double)acc;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
  if (lane == 0 && warp >= L2F_FIRST_GATHER) s_part[warp - L2F_FIRST_GATHER] = d;
  __syncthreads();
  if (tid == 0) {
    double cta = 0.0;
#pragma unroll
    for (int k = 0; k < L2F_GATHER_WARPS; ++k) cta += s_part[k];
    P.partials[blockIdx.x] = cta;
    __threadfence();
    const unsigned t = atomicAdd(P.ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    // every CTA has finished: re-arm the per-image counters and the ticket, add the partials in a fixed order
    for (int b = tid; b < P.B; b += L2F_THREADS) P.done[b] = 0u;
    if (warp == 0) {
      __threadfence();
      double v = 0.0;
      for (unsigned i = lane; i < gridDim.x; i += 32) v += reinterpret_cast<volatile double*>(P.partials)[i];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) {
        *P.loss = (float)v;
        *P.ticket = 0u;
      }
    }
  }
}

// ---- host side ------------------------------------------------------------------------------
int of_tc_get_encode(void** fn);   // conv_tc.cu: cuTensorMapEncodeTiled through the runtime's driver entry point

typedef CUresult (*L2fEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int of_lf2_run(const float* imgs, const float* const* flows, float* const* d_flows, const int* hs, const int* ws,
               int num_scales, float* loss, int B, int H, int W, int mode, void* workspace, void* stream) {
  L2fParams P;
  memset(&P, 0, sizeof(P));
  uint8_t* wsb = reinterpret_cast<uint8_t*>(workspace);
  OF_REQUIRE(((uintptr_t)wsb & 15) == 0 && ((uintptr_t)imgs & 15) == 0, OF_ERR_ARG,
             "workspace and images must be 16-byte aligned");
  P.ticket = reinterpret_cast<unsigned*>(wsb);
  P.done = reinterpret_cast<unsigned*>(wsb + 64);                         // B <= 2048 -> 8 KB
  P.partials = reinterpret_cast<double*>(wsb + 64 + 2048 * 4);
  float4* pyr = reinterpret_cast<float4*>(wsb + L2F_HDR_BYTES);
  P.tile_i = mode == OF_WARP_REF_IJ ? 8 : 4;
  P.tile_j = mode == OF_WARP_REF_IJ ? 16 : 32;
  int ngu = 0;
  for (int s = 0; s < num_scales; ++s) {
    L2fScale& S = P.sc[s];
    S.flow = flows[s];
    S.d_flow = d_flows ? d_flows[s] : nullptr;
    S.h = hs[s]; S.w = ws[s];
    const size_t px = (size_t)B * S.h * S.w;
    S.p1 = pyr; S.p2 = pyr + px; pyr += 2 * px;
    S.t_i = (S.h + P.tile_i - 1) / P.tile_i;
    S.t_j = (S.w + P.tile_j - 1) / P.tile_j;
    S.u_begin = ngu;
    ngu += S.t_i * S.t_j;
    S.inv_n = (float)(1.0 / ((double)B * S.h * S.w * 3.0 * num_scales));
  }
  P.ns = num_scales;
  P.loss = loss;
  P.B = B; P.H = H; P.W = W; P.mode = mode;
  const int bands = (H + L2F_ROWS - 1) / L2F_ROWS;
  P.chunks = (W + L2F_COLS - 1) / L2F_COLS;
  P.nb = bands * P.chunks;
  P.ngu = ngu;
  const long long tb = (long long)B * P.nb, tg = (long long)B * ngu;
  OF_REQUIRE(tb < (1ll << 30) && tg < (1ll << 30), OF_ERR_SHAPE, "too many loss work units");
  P.total_build = (int)tb;
  P.total_gather = (int)tg;
  if (const char* e = of_tune_env("OF_LF2_DBG")) P.dbg = atoi(e);

  CUtensorMap img_map;
  memset(&img_map, 0, sizeof(img_map));
  {
    void* fn;
    if (int r = of_tc_get_encode(&fn)) return r;
    // (B, H, W/16, 96 floats): 16-pixel segments of a row; box {96, 2, 32, 1} = a 32 x 32 pixel patch
    cuuint64_t dims[4] = {96, (cuuint64_t)(W / 16), (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t strides[3] = {384, (cuuint64_t)W * 24, (cuuint64_t)H * W * 24};
    cuuint32_t box[4] = {96, (cuuint32_t)(L2F_COLS / 16), (cuuint32_t)L2F_ROWS, 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = ((L2fEncodeTiled)fn)(&img_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(imgs), dims, strides,
                                      box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    OF_REQUIRE(r == CUDA_SUCCESS, OF_ERR_STATE, "cuTensorMapEncodeTiled(images) failed: %d", (int)r);
  }

  const size_t smem = (size_t)L2F_STAGES * L2F_STAGE_BYTES + 2 * L2F_STAGES * sizeof(uint64_t);
  static int max_ctas = 0;
  if (max_ctas == 0) {
    OF_CUDA(cudaFuncSetAttribute(loss_ws_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0, dev = 0, sms = 0;
    OF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, loss_ws_kernel, L2F_THREADS, smem));
    OF_CUDA(cudaGetDevice(&dev));
    OF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    OF_REQUIRE(per_sm >= 1, OF_ERR_STATE, "fused loss kernel does not fit on an SM");
    max_ctas = sms;                                       // one persistent CTA per SM
  }
  int grid = P.total_build < max_ctas ? P.total_build : max_ctas;
  if (grid > (L2F_HDR_BYTES - 64 - 2048 * 4) / 8) grid = (L2F_HDR_BYTES - 64 - 2048 * 4) / 8;
  // co-residency is required (gather warps wait on patches built by other CTAs): cooperative launch enforces it
  void* args[2] = {(void*)&P, (void*)&img_map};
  OF_CUDA(cudaLaunchCooperativeKernel((const void*)loss_ws_kernel, dim3(grid), dim3(L2F_THREADS), args, smem,
                                      (cudaStream_t)stream));
  return of_launch_check("loss_ws_kernel");
}
