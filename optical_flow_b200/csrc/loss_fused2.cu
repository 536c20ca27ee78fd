// K6 fast path, warp-specialised generation: fused 4-scale resize + warp + L1 photometric loss, forward and d_flow,
// ONE launch (loss.py:5-32 with model.py:55-73 warp semantics), for H, W multiples of 16.
//
// Same dataflow through L2 as loss_fused.cu (the image pair is streamed from HBM exactly once, the float4 pyramid of
// both frames lives in the 126 MB L2 between its producer and its consumers), but the two halves of the work no
// longer take turns inside a CTA.  One persistent CTA per SM runs, side by side,
//   warp 0          TMA producer: 16 x 32 pixel patches of the 6-channel pair into a 16-stage shared-memory ring
//                   (12 KB per stage, mbarrier full / built / empty per stage, evict_first);
//   warp 1          watcher: the only poller of the global per-image "pyramid complete" counters; it republishes
//                   them to a shared-memory word the gather warps spin on (no L2 hot spot);
//   warp 2          publisher: follows the patches of its CTA in order (per-builder-warp progress words) and at every
//                   image boundary issues ONE gpu-scope release fence + counter add for all the patches of that image
//                   this CTA built - the ~1-2 us fence stalls nobody who computes;
//   warps 3..10     BUILD: each warp owns every eighth patch of its CTA: all four pyramid levels of both frames
//                   (tf.image.resize at 2^k = the central 2x2 mean) from shared memory -> float4 texels, evict_last;
//   warps 11..      GATHER: each warp owns its own 128-pixel output tiles (static round-robin over the whole grid, so
//                   the loss sum stays deterministic): flow prefetched one tile ahead, 4 bilinear taps of frame 2
//                   and the frame-1 texel as LDG.128 from the L2-resident pyramid, |diff| -> loss partial, sign ->
//                   d_flow.  A gather tile of image b only waits until all patches of image b are built.
// So the HBM stream (BUILD) and the L2-latency-bound taps (GATHER) overlap on every SM for the whole kernel; only the
// first image's build and the last image's gather are exposed.  No CTA-wide barrier after start-up.
// Measured dissection on B200 at C2 (profile-hooks build, OF_LF2_DBG): launch + final reduction 4.1 us, the TMA stream
// alone 14.6 us more (6.5 TB/s); everything else has to hide under that stream.
// HBM traffic = image pair once + flows + d_flows (the algorithmic bytes of SURVEY 8d).
#include "common.cuh"
#include "tc_common.cuh"
#include "loss_fused_common.cuh"

#define L2F_MAX_SCALES 4
#define L2F_ROWS 16
#define L2F_COLS 32
#define L2F_ROW_FLOATS (L2F_COLS * 6)
#define L2F_STAGE_BYTES (L2F_ROWS * L2F_COLS * 24)      // 12288: one TMA box {96 floats, 2 segments, 16 rows}
#define L2F_STAGES 16
#ifndef L2F_BUILD_WARPS
#define L2F_BUILD_WARPS 8
#endif
#ifndef L2F_GATHER_WARPS
#define L2F_GATHER_WARPS 17
#endif
#ifndef L2F_GATHER_WAIT
#define L2F_GATHER_WAIT 0        // how gather warps wait for an image: 0 poll s_built (nanosleep); 1 park on the image's
#endif                           // mbarrier with a time hint (measured 6 us SLOWER at C2: 56.5 vs 50.1 us - late wake-ups)
#define L2F_IMG_BARS 32          // per-image "pyramid complete" mbarriers (image b -> slot b % 32, phase b / 32)
#define L2F_FIRST_BUILD 3
#define L2F_FIRST_GATHER (L2F_FIRST_BUILD + L2F_BUILD_WARPS)
#define L2F_WARPS (L2F_FIRST_GATHER + L2F_GATHER_WARPS)
#define L2F_THREADS (L2F_WARPS * 32)
#define L2F_FLAGS_OFF 32768      // relay flags: one 128-byte line per CTA in [32 KB, 64 KB) of the header ([16 KB, 32 KB) belongs
#define L2F_FLAGS_BYTES 32768    // to the generic kernel and loss_fused.cu's partials reach past 12 KB)
#define L2F_HDR_BYTES 65536      // same header layout as loss_fused.cu: ticket, per-image counters, per-CTA partials
#define L2F_PUB_BATCH 8          // images one publisher fence can cover

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
  unsigned* flags;       // [grid][32 words = one 128-byte line]: images complete, relayed to CTA c by CTA 0's watcher
  int B, H, W;
  int chunks;            // patches per row of patches
  int nb;                // patches per image
  int ngu;               // gather tiles per image
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

// bounded mbarrier wait: suspends the thread in hardware until the phase completes or ~`ns` have passed
__device__ __forceinline__ bool mbar_try_wait_ns(uint64_t* bar, uint32_t parity, uint32_t ns) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(tc::smem_u32(bar)), "r"(parity), "r"(ns)
      : "memory");
  return ok != 0;
}

// lf_mean4 on two channels at once (sm_100 packed fp32x2: FADD2 / FFMA2, each lane rounds exactly like the scalar
// FADD / FFMA the compiler emits for lf_mean4, so the texels are bit-identical to loss_fused.cu / loss.cu)
__device__ __forceinline__ float2 l2f_mean4_2(float2 a, float2 b, float2 c, float2 d) {
  const float2 h = make_float2(0.5f, 0.5f);
  const float2 top = __ffma2_rn(__fadd2_rn(b, make_float2(-a.x, -a.y)), h, a);
  const float2 bot = __ffma2_rn(__fadd2_rn(d, make_float2(-c.x, -c.y)), h, c);
  return __ffma2_rn(__fadd2_rn(bot, make_float2(-top.x, -top.y)), h, top);
}

// one resized pixel of BOTH frames from the staged patch: the central 2x2 block whose top-left is (r0, c0).
// VEC: c0 is even (scale 0), so the 2 pixels x 6 floats are 16-byte aligned.  Channel pairs (0,1) (2,3) (4,5) of the
// left / right pixel sit in aligned register pairs of the loaded vectors.
__device__ __forceinline__ void l2f_build_pixel(const bool VEC, const float* T, int r0, int c0, float4* o1, float4* o2) {
  const float* a = T + r0 * L2F_ROW_FLOATS + c0 * 6;
  const float* c = a + L2F_ROW_FLOATS;
  float2 al[3], ar[3], cl[3], cr[3];                 // [channel pair] of the left / right pixel, rows a / c
  if (VEC) {
    const float4 u0 = reinterpret_cast<const float4*>(a)[0], u1 = reinterpret_cast<const float4*>(a)[1],
                 u2 = reinterpret_cast<const float4*>(a)[2];
    const float4 v0 = reinterpret_cast<const float4*>(c)[0], v1 = reinterpret_cast<const float4*>(c)[1],
                 v2 = reinterpret_cast<const float4*>(c)[2];
    al[0] = make_float2(u0.x, u0.y); al[1] = make_float2(u0.z, u0.w); al[2] = make_float2(u1.x, u1.y);
    ar[0] = make_float2(u1.z, u1.w); ar[1] = make_float2(u2.x, u2.y); ar[2] = make_float2(u2.z, u2.w);
    cl[0] = make_float2(v0.x, v0.y); cl[1] = make_float2(v0.z, v0.w); cl[2] = make_float2(v1.x, v1.y);
    cr[0] = make_float2(v1.z, v1.w); cr[1] = make_float2(v2.x, v2.y); cr[2] = make_float2(v2.z, v2.w);
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      al[k] = reinterpret_cast<const float2*>(a)[k]; ar[k] = reinterpret_cast<const float2*>(a)[3 + k];
      cl[k] = reinterpret_cast<const float2*>(c)[k]; cr[k] = reinterpret_cast<const float2*>(c)[3 + k];
    }
  }
#ifdef L2F_SCALAR_BUILD
  const float2 m01 = make_float2(lf_mean4(al[0].x, ar[0].x, cl[0].x, cr[0].x), lf_mean4(al[0].y, ar[0].y, cl[0].y, cr[0].y));
  const float2 m23 = make_float2(lf_mean4(al[1].x, ar[1].x, cl[1].x, cr[1].x), lf_mean4(al[1].y, ar[1].y, cl[1].y, cr[1].y));
  const float2 m45 = make_float2(lf_mean4(al[2].x, ar[2].x, cl[2].x, cr[2].x), lf_mean4(al[2].y, ar[2].y, cl[2].y, cr[2].y));
#else
  const float2 m01 = l2f_mean4_2(al[0], ar[0], cl[0], cr[0]);
  const float2 m23 = l2f_mean4_2(al[1], ar[1], cl[1], cr[1]);
  const float2 m45 = l2f_mean4_2(al[2], ar[2], cl[2], cr[2]);
#endif
  *o1 = make_float4(m01.x, m01.y, m23.x, 0.f);
  *o2 = make_float4(m23.y, m45.x, m45.y, 0.f);
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

// Gather tile geometry: 64 pixels per warp tile, two per lane (both in flight together).  REF_IJ: 8 x 8 pixels, an
// instruction covers 8 rows x 4 columns - the transposed gather (source column = i) then reads 8 consecutive texels
// of 4 source rows.  XY: 2 x 32, one output row per instruction.  Small tiles on purpose: a gather tile is one L2
// round trip (half of it to the far die's L2 partition), and an image's tiles can only start once the whole image is
// built, so the tiles of one image should cover ALL gather warps of the grid (C2: 2 550 tiles vs 2 516 warps).
// Pixel q of a lane sits at a constant offset (di(q), dj(q)) from the lane's first pixel.
template <int MODE> struct L2fGeom;
template <> struct L2fGeom<OF_WARP_REF_IJ> {
  static constexpr int TI = 8, TJ = 8;
  __device__ static __forceinline__ void first(const L2fTile& t, int lane, int* i, int* j) {
    *i = t.ti * 8 + (lane & 7);
    *j = t.tj * 8 + (lane >> 3);
  }
  __device__ static __forceinline__ int di(int) { return 0; }
  __device__ static __forceinline__ int dj(int q) { return 4 * q; }
};
template <> struct L2fGeom<OF_WARP_XY> {
  static constexpr int TI = 2, TJ = 32;
  __device__ static __forceinline__ void first(const L2fTile& t, int lane, int* i, int* j) {
    *i = t.ti * 2;
    *j = t.tj * 32 + lane;
  }
  __device__ static __forceinline__ int di(int q) { return q; }
  __device__ static __forceinline__ int dj(int) { return 0; }
};

// Ragged tiles (h or w not a multiple of the tile) are handled without a second code path: a pixel outside the image
// is CLAMPED onto the last row / column (all of its loads are then valid) and its contribution is masked at the end.
template <int MODE>
__device__ __forceinline__ void l2f_load_flows(const L2fParams& P, const L2fTile& t, int lane, float2& fl0, float2& fl1) {
  using G = L2fGeom<MODE>;
  const L2fScale& S = P.sc[t.s];
  int i0, j0;
  G::first(t, lane, &i0, &j0);
  const float2* g = reinterpret_cast<const float2*>(S.flow) + (size_t)t.b * S.h * S.w;
  const int hm = S.h - 1, wm = S.w - 1;
  fl0 = __ldg(g + min(i0 + G::di(0), hm) * S.w + min(j0 + G::dj(0), wm));
  fl1 = __ldg(g + min(i0 + G::di(1), hm) * S.w + min(j0 + G::dj(1), wm));
}

// one pixel of a gather tile in two steps, so a thread keeps the taps of two pixels in flight: l2f_issue computes the
// sampling position and sends out the 5 loads, l2f_finish does the arithmetic.
struct L2fPix {
  float ax, ay, bx, by;
  float4 v00, v01, v10, v11, ref;
  int off;            // pixel offset within the image; bit 31 set = outside the image (masked)
};

template <int MODE>
__device__ __forceinline__ void l2f_issue(L2fPix& r, const float4* __restrict__ q1, const float4* __restrict__ q2,
                                          int h, int w, uint64_t keep, int i_raw, int j_raw, float2 f) {
  const int i = min(i_raw, h - 1), j = min(j_raw, w - 1);
  r.off = (i * w + j) | ((i_raw < h && j_raw < w) ? 0 : (int)0x80000000);
  // model.py:65-71 + transformations.py:93-125 (same op order as loss.cu / warp.cu)
  float x, y;
  if (MODE == OF_WARP_REF_IJ) { x = __fadd_rn((float)i, f.x); y = __fadd_rn((float)j, f.y); }
  else { x = __fadd_rn((float)j, f.x); y = __fadd_rn((float)i, f.y); }
  const int xf = __float2int_rd(x), yf = __float2int_rd(y);
  // clip(floor + 1) with the +1 saturating (floor may be INT_MAX for a saturated cast)
  const int x0 = lf_clamp(xf, 0, w - 1), x1 = lf_clamp(min(xf, 0x7ffffffe) + 1, 0, w - 1);
  const int y0 = lf_clamp(yf, 0, h - 1), y1 = lf_clamp(min(yf, 0x7ffffffe) + 1, 0, h - 1);
  r.ax = __fsub_rn((float)x1, x);
  r.ay = __fsub_rn((float)y1, y);
  r.bx = __fsub_rn(1.f, r.ax);
  r.by = __fsub_rn(1.f, r.ay);
  const float4* r0 = q2 + y0 * w;
  const float4* r1 = q2 + y1 * w;
  r.v00 = ld_f4_policy(r0 + x0, keep);
  r.v01 = ld_f4_policy(r1 + x0, keep);
  r.v10 = ld_f4_policy(r0 + x1, keep);
  r.v11 = ld_f4_policy(r1 + x1, keep);
  r.ref = ld_f4_policy(q1 + (r.off & 0x7fffffff), keep);
}

__device__ __forceinline__ void l2f_finish(const L2fPix& r, float2* __restrict__ gdflow, float inv_n, float& acc) {
  const bool inside = r.off >= 0;
  const float w00 = r.ax * r.ay, w01 = r.ax * r.by, w10 = r.bx * r.ay, w11 = r.bx * r.by;
  const float wx = w00 * r.v00.x + w01 * r.v01.x + w10 * r.v10.x + w11 * r.v11.x;
  const float wy = w00 * r.v00.y + w01 * r.v01.y + w10 * r.v10.y + w11 * r.v11.y;
  const float wz = w00 * r.v00.z + w01 * r.v01.z + w10 * r.v10.z + w11 * r.v11.z;
  const float ex = r.ref.x - wx, ey = r.ref.y - wy, ez = r.ref.z - wz;               // loss.py:28
  const float term = (fabsf(ex) + fabsf(ey) + fabsf(ez)) * inv_n;
  acc += inside ? term : 0.f;
  if (gdflow != nullptr && inside) {
    const float gxx = r.ay * (r.v10.x - r.v00.x) + r.by * (r.v11.x - r.v01.x);
    const float gxy = r.ay * (r.v10.y - r.v00.y) + r.by * (r.v11.y - r.v01.y);
    const float gxz = r.ay * (r.v10.z - r.v00.z) + r.by * (r.v11.z - r.v01.z);
    const float gyx = r.ax * (r.v01.x - r.v00.x) + r.bx * (r.v11.x - r.v10.x);
    const float gyy = r.ax * (r.v01.y - r.v00.y) + r.bx * (r.v11.y - r.v10.y);
    const float gyz = r.ax * (r.v01.z - r.v00.z) + r.bx * (r.v11.z - r.v10.z);
    const float sx_ = -lf_sgn(ex) * inv_n, sy_ = -lf_sgn(ey) * inv_n, sz_ = -lf_sgn(ez) * inv_n;
    const float d0 = sx_ * gxx + sy_ * gxy + sz_ * gxz;
    const float d1 = sx_ * gyx + sy_ * gyy + sz_ * gyz;
    gdflow[r.off] = make_float2(d0, d1);
  }
}

template <int MODE>
__device__ __forceinline__ void l2f_gather_tile(const float4* __restrict__ q1, const float4* __restrict__ q2,
                                                float2* __restrict__ gdflow, int h, int w, float inv_n, uint64_t keep,
                                                int i0, int j0, float2 f0, float2 f1, float& acc) {
  using G = L2fGeom<MODE>;
  L2fPix a, b;
  l2f_issue<MODE>(a, q1, q2, h, w, keep, i0 + G::di(0), j0 + G::dj(0), f0);
  l2f_issue<MODE>(b, q1, q2, h, w, keep, i0 + G::di(1), j0 + G::dj(1), f1);
  l2f_finish(a, gdflow, inv_n, acc);
  l2f_finish(b, gdflow, inv_n, acc);
}

template <int MODE>
__global__ void __launch_bounds__(L2F_THREADS, 1)
loss_ws_kernel(const __grid_constant__ L2fParams P, const __grid_constant__ CUtensorMap img_map) {
  extern __shared__ __align__(128) uint8_t l2f_smem[];
  uint8_t* tiles = l2f_smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(l2f_smem + L2F_STAGES * L2F_STAGE_BYTES);
  uint64_t* empty_bar = full_bar + L2F_STAGES;      // builder warp -> producer
  __shared__ double s_part[L2F_GATHER_WARPS];
  __shared__ unsigned s_progress[L2F_BUILD_WARPS];   // builder warp w: 1 + index of its latest finished local patch
  __shared__ unsigned s_built;          // images whose pyramid is complete and visible (watcher -> gather warps)
  __shared__ __align__(8) uint64_t img_bar[L2F_IMG_BARS];   // image b complete: phase b / 32 of slot b % 32 (gather warps
                                                            // sleep on it in hardware instead of polling s_built)
  __shared__ bool s_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < L2F_STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    for (int w = 0; w < L2F_BUILD_WARPS; ++w) s_progress[w] = 0u;
    for (int i = 0; i < L2F_IMG_BARS; ++i) tc::mbar_init(&img_bar[i], 1);
    s_built = 0u;
    tc::fence_barrier_init();
  }
  __syncthreads();
  const bool stream_on = !(OF_DBG(P) & 16);
  const int n_build = stream_on ? P.total_build : 0;

  float acc = 0.f;
  if (warp == 0) {
    // ===================== TMA producer =====================
    if (OF_ONE_LANE()) {
      tc::prefetch_tmap(&img_map);
      const uint64_t pol = policy_evict_first();
      int k = 0;
      for (int p = blockIdx.x; p < n_build; p += gridDim.x, ++k) {
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
    // Every watcher polls the per-image counters itself (default).  -DL2F_RELAY: only CTA 0's watcher polls them and
    // relays "image b complete" to one 128-byte flag line per CTA, which that CTA's watcher polls (one poller per line).
    // Measured at C2: the relay is SLOWER (52.2 vs 50.4 us; skeleton without gather 35.2 vs 29.0 us) - the 5 us the
    // pollers add over a run without them (24.1 us) is the completion chain's latency at the tail (fence -> counter ->
    // poll -> fence), not an L2 hot spot, and the relay adds one more hop to that chain.
    if (!(OF_DBG(P) & (4 | 8 | 16 | 64))) {                     // 64 (with 2): fences and counters on, nobody polls them
#ifndef L2F_RELAY
      if (lane == 0)
        for (int b = 0; b < P.B; ++b) {
          while (ld_relaxed(&P.done[b]) < (unsigned)P.nb) __nanosleep(400);
          asm volatile("fence.acq_rel.gpu;" ::: "memory");
          st_release_cta_shared(&s_built, (unsigned)(b + 1));
          tc::mbar_arrive(&img_bar[b % L2F_IMG_BARS]);
        }
#else
      if (blockIdx.x == 0) {
        for (int b = 0; b < P.B; ++b) {
          if (lane == 0) {
            while (ld_relaxed(&P.done[b]) < (unsigned)P.nb) __nanosleep(100);
            asm volatile("fence.acq_rel.gpu;" ::: "memory");     // acquire the publishers' releases, release the relay
          }
          __syncwarp();
          for (unsigned c = 1 + lane; c < gridDim.x; c += 32)
            asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(P.flags + (size_t)c * 32), "r"((unsigned)(b + 1)) : "memory");
          if (lane == 0) {
            st_release_cta_shared(&s_built, (unsigned)(b + 1));
            tc::mbar_arrive(&img_bar[b % L2F_IMG_BARS]);
          }
          __syncwarp();
        }
      } else if (lane == 0) {
        const unsigned* flag = P.flags + (size_t)blockIdx.x * 32;
        for (int b = 0; b < P.B;) {
          unsigned v;
          while ((v = ld_relaxed(flag)) <= (unsigned)b) __nanosleep(200);
          asm volatile("fence.acq_rel.gpu;" ::: "memory");
          for (; b < (int)v; ++b) {                              // several images may have completed since the last look
            st_release_cta_shared(&s_built, (unsigned)(b + 1));
            tc::mbar_arrive(&img_bar[b % L2F_IMG_BARS]);
          }
        }
      }
#endif
    }
  } else if (warp == 2) {
    // ===================== publisher =====================
    if (lane == 0) {
      int k = 0;
      int p = blockIdx.x;
      int cur_cnt = 0;                             // patches of the current image counted so far
      int pub_img[L2F_PUB_BATCH], pub_cnt[L2F_PUB_BATCH];
      while (p < n_build) {
        int npub = 0;
        bool first = true;
        // take every patch that is already built (blocking only for the first), in order: local patch k belongs to
        // builder warp k % 8, whose progress word passes k when its stores are issued
        while (p < n_build && npub < L2F_PUB_BATCH) {
          const unsigned* word = &s_progress[k % L2F_BUILD_WARPS];
          if (first) { while (ld_acquire_cta_shared(word) <= (unsigned)k) __nanosleep(40); }
          else if (ld_acquire_cta_shared(word) <= (unsigned)k) break;
          first = false;
          const int b = p / P.nb;
          ++cur_cnt;
          p += gridDim.x; ++k;
          if (p >= n_build || p / P.nb != b) {                     // that was this CTA's last patch of image b
            pub_img[npub] = b; pub_cnt[npub] = cur_cnt; ++npub; cur_cnt = 0;
          }
        }
        if (npub > 0 && !(OF_DBG(P) & 8)) {
          asm volatile("fence.acq_rel.gpu;" ::: "memory");         // the builders' stores (observed through the built
          for (int i = 0; i < npub; ++i)                           // barriers), cumulatively, before the counts
            asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(P.done + pub_img[i]), "r"((unsigned)pub_cnt[i])
                         : "memory");
        }
      }
    }
  } else if (warp < L2F_FIRST_GATHER) {
    // ===================== BUILD =====================
    const int bw = warp - L2F_FIRST_BUILD;
    const uint64_t keep = policy_evict_last();
    // patch index -> (image b, band, chunk cx) incrementally: the warp's patches are `step` apart
    const int step = L2F_BUILD_WARPS * (int)gridDim.x;
    const int bands = P.nb / P.chunks;
    const int step_b = step / P.nb, step_l = step - step_b * P.nb;
    const int step_band = step_l / P.chunks, step_cx = step_l - step_band * P.chunks;
    int p = blockIdx.x + bw * (int)gridDim.x;
    int b = p / P.nb, band, cx;
    { const int local = p - b * P.nb; band = local / P.chunks; cx = local - band * P.chunks; }
    const int ocol0 = lane & 15, or0 = lane >> 4;          // scale 0: 8 x 16 resized pixels, 4 rows per lane
    const int orow1 = lane >> 3, ocol1 = lane & 7;         // scale 1: 4 x 8, one per lane
    for (int k = bw; p < n_build; p += step, k += L2F_BUILD_WARPS) {
      const int stage = k % L2F_STAGES;
      const uint32_t phase = (uint32_t)(k / L2F_STAGES) & 1u;
      const int wc = min(L2F_COLS, P.W - cx * L2F_COLS), hr = min(L2F_ROWS, P.H - band * L2F_ROWS);
#ifdef L2F_NO_FULL_PATH
      const bool full = false;
#else
      const bool full = wc == L2F_COLS && hr == L2F_ROWS;  // interior patch: no ragged-edge predicates (warp-uniform)
#endif
      if (lane == 0) tc::mbar_wait(&full_bar[stage], phase);
      __syncwarp();
      const float* T = reinterpret_cast<const float*>(tiles + (size_t)stage * L2F_STAGE_BYTES);
      if (!(OF_DBG(P) & 1)) {
        // scale 0: two pixels at a time (independent chains)
        if (P.ns > 0) {
          const L2fScale& S = P.sc[0];
          const unsigned pix0 = ((unsigned)b * S.h + band * (L2F_ROWS / 2) + or0) * S.w + cx * (L2F_COLS / 2) + ocol0;
          float4* q1 = S.p1 + pix0;
          float4* q2 = S.p2 + pix0;
          const unsigned rstep = 2u * (unsigned)S.w;
          const float* Tl = T + (2 * or0) * L2F_ROW_FLOATS + (2 * ocol0) * 6;
          if (full) {
#pragma unroll 2
            for (int it = 0; it < 4; ++it) {
              float4 o1, o2;
              l2f_build_pixel(true, Tl + it * (4 * L2F_ROW_FLOATS), 0, 0, &o1, &o2);
              st_f4_policy(q1 + it * rstep, o1, keep);
              st_f4_policy(q2 + it * rstep, o2, keep);
            }
          } else {
#pragma unroll 2
            for (int it = 0; it < 4; ++it) {
              if (ocol0 * 2 < wc && (or0 + 2 * it) * 2 < hr) {
                float4 o1, o2;
                l2f_build_pixel(true, Tl + it * (4 * L2F_ROW_FLOATS), 0, 0, &o1, &o2);
                st_f4_policy(q1 + it * rstep, o1, keep);
                st_f4_policy(q2 + it * rstep, o2, keep);
              }
            }
          }
        }
        // scale 1: 4 x 8 pixels, one per lane
        if (P.ns > 1) {
          const L2fScale& S = P.sc[1];
          if (full || (ocol1 * 4 < wc && orow1 * 4 < hr)) {
            float4 o1, o2;
            l2f_build_pixel(false, T, 4 * orow1 + 1, 4 * ocol1 + 1, &o1, &o2);
            const unsigned pix = ((unsigned)b * S.h + band * (L2F_ROWS / 4) + orow1) * S.w + cx * (L2F_COLS / 4) + ocol1;
            st_f4_policy(S.p1 + pix, o1, keep);
            st_f4_policy(S.p2 + pix, o2, keep);
          }
        }
        // scales 2 and 3 in one pass: lanes 0-7 -> the 2 x 4 pixels of scale 2, lanes 8-9 -> the 1 x 2 of scale 3
        if (P.ns > 2 && lane < 10) {
          const int sidx = lane < 8 ? 2 : 3;
          if (sidx < P.ns) {
            const L2fScale& S = P.sc[sidx];
            const int bs = 2 << sidx;
            const int t = lane < 8 ? lane : lane - 8;
            const int lc = 4 - sidx;                                 // log2(columns of this scale per patch)
            const int orow = t >> lc, ocol = t & ((1 << lc) - 1);
            if (full || (ocol * bs < wc && orow * bs < hr)) {
              float4 o1, o2;
              l2f_build_pixel(false, T, bs * orow + (bs >> 1) - 1, bs * ocol + (bs >> 1) - 1, &o1, &o2);
              const unsigned pix = ((unsigned)b * S.h + band * (L2F_ROWS / bs) + orow) * S.w + cx * (L2F_COLS / bs) + ocol;
              st_f4_policy(S.p1 + pix, o1, keep);
              st_f4_policy(S.p2 + pix, o2, keep);
            }
          }
        }
      }
      __syncwarp();                       // every lane's shared-memory reads and pyramid stores are issued
      if (lane == 0) {
        tc::mbar_arrive(&empty_bar[stage]);                        // the ring slot refills
        st_release_cta_shared(&s_progress[bw], (unsigned)(k + 1)); // the publisher follows
      }
      // next patch of this warp
      b += step_b; band += step_band; cx += step_cx;
      if (cx >= P.chunks) { cx -= P.chunks; ++band; }
      if (band >= bands) { band -= bands; ++b; }
    }
  } else {
    // ===================== GATHER =====================
    using G = L2fGeom<MODE>;
    const uint64_t keep = policy_evict_last();
    const int gw = (warp - L2F_FIRST_GATHER) * gridDim.x + blockIdx.x;
    const int tgw = L2F_GATHER_WARPS * gridDim.x;
    const int total_gather = (OF_DBG(P) & 2) ? 0 : P.total_gather;
    int confirmed = 0;                  // images known complete
    float2 n0 = make_float2(0.f, 0.f), n1 = n0;
    L2fTile nt;
    if (gw < total_gather) { nt = l2f_decode(P, gw); l2f_load_flows<MODE>(P, nt, lane, n0, n1); }
    for (int u = gw; u < total_gather; u += tgw) {
      const L2fTile t = nt;
      const float2 f0 = n0, f1 = n1;
      if (u + tgw < total_gather) { nt = l2f_decode(P, u + tgw); l2f_load_flows<MODE>(P, nt, lane, n0, n1); }   // in flight
      if (t.b >= confirmed && !(OF_DBG(P) & 4)) {
        // s_built is the truth; the mbarrier only parks the warp (images complete in order, so slot b % 32 is in
        // phase b / 32 or already past it; an early return just re-reads s_built)
        unsigned v;
#if L2F_GATHER_WAIT == 0
        while ((v = ld_acquire_cta_shared(&s_built)) <= (unsigned)t.b) __nanosleep(64);
#else
        while ((v = ld_acquire_cta_shared(&s_built)) <= (unsigned)t.b)
          mbar_try_wait_ns(&img_bar[t.b % L2F_IMG_BARS], (uint32_t)(t.b / L2F_IMG_BARS) & 1u, 4000u);
#endif
        confirmed = (int)v;
      }
      const L2fScale& S = P.sc[t.s];
      const int h = S.h, w = S.w;
      const unsigned img_px = (unsigned)t.b * h * w;
      const float4* q2 = S.p2 + img_px;
      const float4* q1 = S.p1 + img_px;
      float2* gdflow = S.d_flow ? reinterpret_cast<float2*>(S.d_flow) + img_px : nullptr;
      int i0, j0;
      G::first(t, lane, &i0, &j0);
      l2f_gather_tile<MODE>(q1, q2, gdflow, h, w, S.inv_n, keep, i0, j0, f0, f1, acc);
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
    for (unsigned c = tid; c < gridDim.x; c += L2F_THREADS) P.flags[(size_t)c * 32] = 0u;
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
  P.partials = reinterpret_cast<double*>(wsb + 64 + 2048 * 4);            // [8256, 12288): up to 504 CTAs
  float4* pyr = reinterpret_cast<float4*>(wsb + L2F_HDR_BYTES);
  const int tile_i = mode == OF_WARP_REF_IJ ? L2fGeom<OF_WARP_REF_IJ>::TI : L2fGeom<OF_WARP_XY>::TI;
  const int tile_j = mode == OF_WARP_REF_IJ ? L2fGeom<OF_WARP_REF_IJ>::TJ : L2fGeom<OF_WARP_XY>::TJ;
  int ngu = 0;
  for (int s = 0; s < num_scales; ++s) {
    L2fScale& S = P.sc[s];
    S.flow = flows[s];
    S.d_flow = d_flows ? d_flows[s] : nullptr;
    S.h = hs[s]; S.w = ws[s];
    const size_t px = (size_t)B * S.h * S.w;
    S.p1 = pyr; S.p2 = pyr + px; pyr += 2 * px;
    S.t_i = (S.h + tile_i - 1) / tile_i;
    S.t_j = (S.w + tile_j - 1) / tile_j;
    S.u_begin = ngu;
    ngu += S.t_i * S.t_j;
    S.inv_n = (float)(1.0 / ((double)B * S.h * S.w * 3.0 * num_scales));
  }
  P.ns = num_scales;
  P.loss = loss;
  P.flags = reinterpret_cast<unsigned*>(wsb + L2F_FLAGS_OFF);
  P.B = B; P.H = H; P.W = W;
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
    // (B, H, W/16, 96 floats): 16-pixel segments of a row; box {96, 2, 16, 1} = a 16 x 32 pixel patch
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
    OF_CUDA(cudaFuncSetAttribute(loss_ws_kernel<OF_WARP_REF_IJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    OF_CUDA(cudaFuncSetAttribute(loss_ws_kernel<OF_WARP_XY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0, dev = 0, sms = 0;
    OF_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, loss_ws_kernel<OF_WARP_REF_IJ>, L2F_THREADS, smem));
    OF_CUDA(cudaGetDevice(&dev));
    OF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    OF_REQUIRE(per_sm >= 1, OF_ERR_STATE, "fused loss kernel does not fit on an SM");
    max_ctas = sms;                                       // one persistent CTA per SM
  }
  int grid = P.total_build < max_ctas ? P.total_build : max_ctas;
  if (grid > L2F_FLAGS_BYTES / 128) grid = L2F_FLAGS_BYTES / 128;
  if (grid > (16384 - 64 - 2048 * 4) / 8) grid = (16384 - 64 - 2048 * 4) / 8;          // partials stay inside the first 16 KB
  // co-residency is required (gather warps wait on patches built by other CTAs): cooperative launch enforces it
  void* args[2] = {(void*)&P, (void*)&img_map};
  const void* fn = mode == OF_WARP_REF_IJ ? (const void*)loss_ws_kernel<OF_WARP_REF_IJ> : (const void*)loss_ws_kernel<OF_WARP_XY>;
  OF_CUDA(cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(L2F_THREADS), args, smem, (cudaStream_t)stream));
  return of_launch_check("loss_ws_kernel");
}
