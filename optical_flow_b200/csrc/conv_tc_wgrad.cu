// K10 (wgrad): weight gradient of Conv2D as a tcgen05 / TMEM GEMM with the PIXELS as the
// reduction dimension:  dW[tap][ci][co] = sum_pix x[pix (+) tap][ci] * dy[pix][co].
//
// Both operands are NHWC activations, i.e. "MN-major" for this GEMM (the contiguous run is
// channels = M for x / N for dy; K = pixels runs over 128-byte rows).  A TMA box
// {64 channels, TWk, THk, 1} (TWk*THk = 64 pixels) lands in shared memory as 64 K-rows x 128 B
// with SWIZZLE_128B, exactly the canonical MN-major UMMA operand; the x box is shifted by the
// filter tap (zero fill outside the image == SAME padding, elementStrides for stride 2).
// One CTA owns 128 input channels (two 64-channel chunks, possibly from two tensors - the
// decoder concat), all Cout columns, a group of taps (one TMEM accumulator per tap: the dy
// tile is loaded once per K step and reused by every tap) and a slice of the pixel range
// (split-K).  Partial results are reduced into dW with red.global.add.f32.
// Bias gradient (db[co] = sum_pix dy[pix][co]) rides along: one job per K tile (rotating) issues one more MMA group
// against a constant all-ones tile into a spare TMEM accumulator - the dy tile is already in shared memory, so the
// separate column-sum pass over dy (1.25 ms / step) disappears.  Default form (round 2): TRANSPOSED - the dy tile is the
// A operand (M = its channels), ones the B operand, N = 16; see of_conv_tc_wgrad for the measurements.
#include "common.cuh"
#include "tc_common.cuh"

#include <atomic>

#define WG_THREADS 192
#define WG_CHUNK_BYTES 8192   // 64 pixels x 64 bf16
#define WG_MAX_TAPS 9
#define WG_ONES_BYTES 4096   // all-ones A operand: 2 chunks x 16 K-rows x 128 B (reused for every K slice)

int of_tc_make_act_map(CUtensorMap* m, const void* base, int c_used, long long stride_elems, int N, int H, int W,
                       int box_w, int box_h, int estride);
int of_tc_make_view_map(CUtensorMap* m, const void* base, int c_used, long long w_stride, long long h_stride,
                        long long n_stride, int N, int H, int W, int box_w, int box_h, int sw, int sh);

#define WG_MAX_CHUNKS 16
#define WG_MAX_JOBS 48
struct WgradParams {
  // 64-channel chunks of the (concatenated) input; M tile m uses chunks 2m and 2m+1
  int chunk_map[WG_MAX_CHUNKS];   // which x tensor map (0/1); -1 = none (zero)
  int chunk_c0[WG_MAX_CHUNKS];    // channel coordinate inside that tensor
  int chunk_ci[WG_MAX_CHUNKS];    // original ci of the chunk's first channel
  int chunk_cnt[WG_MAX_CHUNKS];   // valid channels in the chunk
  // job (blockIdx.x) = (M tile, tap group)
  int job_mt[WG_MAX_JOBS], job_tap0[WG_MAX_JOBS], job_ntaps[WG_MAX_JOBS];
  int KW, pad_t, pad_l, stride_w, stride_h;
  int NB, Ho, Wo, TW, TH, tiles_w, tiles_h;
  int ktiles, ktiles_per_split, kt_step;
  // dynamic K-tile scheduling: the CTAs of a job claim `claim` consecutive K tiles at a time from the job's counter in
  // g_wg_ctr[ctr_set] (0 = static ranges as described by ktiles_per_split / kt_step)
  int dyn, claim, ctr_set;
  int dbg;                        // -DOF_PROFILE_HOOKS only (OF_WG_DBG): 1 = the epilogue reads TMEM but issues no atomics
  int n_mma, n_valid, n_chunks;   // dy: MMA N, real Cout, 64-channel chunks
  int Cin, Cout;
  int stages, stage_bytes, tmem_cols;
  float* dw;
  float* db;                      // nullable: fused bias gradient (one job per pixel range)
  int k_outer;                    // N-stacked main loop: K slice outermost (option OF_WG_KOUTER)
  int bias_t;                     // 1: transposed bias MMAs - A = the dy tile (M = its channels), B = ones, N = 16
  int bias_m;                     // M of the ones-MMAs of the bias gradient (64 or 128; only row 0 is read)
  int bias_range;                 // 1: ONE job per split-K range computes the bias gradient of all its tiles (static ranges only)
  // Cin <= 64 (one chunk): the two 64-row halves of an accumulator hold two different TAPS of that chunk
  // instead of two channel chunks - no zero half, half the MMAs.  job_tap0 / job_ntaps then count accumulators.
  int pack2, taps;
  // 3x3 / stride 1 with >= 128 input channels per M tile: a job is one kw column; ONE TMA box
  // {64, TW, TH + 2} per chunk (the "slab") serves its three kh taps as descriptors at row offsets
  // kh * TW * 128 B (multiples of the 1024-byte swizzle atom) - half the L2 -> SM bytes of tap-by-tap loads.
  int slab, slab_bytes;
  // 3x3 / stride 1 with Cin <= 64: pack2 fed from THREE slabs (one per kw, all in one CTA) instead of ten
  // tap-by-tap boxes - half the L2 -> SM bytes per K step.  Accumulators: t = 0..2 pair (kh 0, kh 1) of kw = t
  // (halves TW rows apart inside a slab), t = 3 pairs kh 2 of kw 0 / kw 1 (halves one slab apart), t = 4 holds
  // kh 2 of kw 2 in its lower half (the upper half multiplies whatever follows in the stage and is discarded).
  int slab2;
  // 3x3 / stride 1 with Cout <= 128 ("N-stacked"): shift dy instead of x along kh.  dW[kh][kw] = sum_q x[q + (0, kw-1)]
  // * dy[q - (kh-1, 0)] (x is zero outside the image, so q over the image rows is complete): the B operand is ONE dy
  // slab {64, TW, TH + 2} whose three kh-shifted views are the three 64-column chunks of an N = 192 MMA (chunk j at
  // +j rows = kh 2 - j), the A operand x boxes without halo.  A third of the MMA instructions, each 3x wider - these
  // narrow layers are bound by the per-instruction cost, not by the math.  Cin <= 64: one job, accumulators
  // (kw0 | kw1) and (kw2 | junk); otherwise a job is (M tile, kw) with one accumulator per 64-column dy chunk
  // (Cout 96 / 128: two dy slabs, two N = 192 MMAs per K slice instead of three N = Cout ones: -12..15 %).
  int nstack, nstack_one;
  // nstack_one split over two jobs ((kw0 | kw1) and (kw2 | junk)): 256 TMEM columns each, so TWO CTAs fit an SM
  int one_split;
};

// Dynamic K-tile scheduling (option OF_WG_DYNAMIC; OFF by default - measured slower).  With static, equal pixel ranges the
// equal-work CTAs of a launch finish up to 35 % apart (ncu, 128->128: SM-active cycles avg 342k, max 461k).  Hypothesis:
// the kernel ends with its slowest CTA and the jobs of a range drift so far apart that the later ones miss L2.  Test: hand
// the K tiles out at run time - one counter per job, a CTA claims `claim` consecutive tiles per atomic (the next claim is in
// flight while the current tiles load), every CTA works at the same frontier of the tensor and all finish together.
// Result (round 2, call 44, same box): wgrad per step 2 409 us static -> 2 510 us dynamic (claim 1 / 2 / 4 alike, 8 worse);
// 128->128 254 -> 261 us, 16 x 96 x 320 64->64 66 -> 75 us.  So the spread is not lost time: a CTA that finishes early
// leaves its share of the L2 / shared-memory / tensor throughput to the others, the launch is bound by the aggregate.
// The sums are added with red.global.add in either mode, so which CTA takes which tile does not change the result beyond
// the summation order.  Counter sets: one per launch, round robin (concurrent launches on different streams never share
// one); the last CTA to finish zeroes the set for its next use.
#define WG_CTR_SETS 256
#define WG_CTR_STRIDE (WG_MAX_JOBS + 1)          // [0, WG_MAX_JOBS): next K tile of the job; [WG_MAX_JOBS]: finished CTAs
static __device__ unsigned g_wg_ctr[WG_CTR_SETS * WG_CTR_STRIDE];

__global__ void __launch_bounds__(WG_THREADS, 1)
conv_tc_wgrad_kernel(const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1,
                     const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ WgradParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* ones = tiles + (size_t)P.stages * P.stage_bytes;            // 1024-aligned (stage_bytes is a multiple of 8 KB)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ones + WG_ONES_BYTES);
  uint64_t* empty_bar = full_bar + 8;
  uint64_t* acc_full = empty_bar + 8;
  uint64_t* meta_bar = acc_full + 1;                       // producer -> epilogue: s_meta is written
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(meta_bar + 1);
  int* s_kt = reinterpret_cast<int*>(tmem_base_slot + 2);   // K tile held by each stage (-1: no more tiles)
  int* s_meta = s_kt + 8;                                   // [0] tiles this CTA processed, [1] any of them with the bias MMAs
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P.stages; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
    tc::mbar_init(acc_full, 1);
    tc::mbar_init(meta_bar, 1);
    tc::fence_barrier_init();
  }
  if (warp == 4 && lane == 0) { tc::prefetch_tmap(&tmX0); tc::prefetch_tmap(&tmX1); tc::prefetch_tmap(&tmDY); }
  // the bias gradient of a K tile is computed by ONE of the jobs that walk it, rotating PER K TILE (tile kt: job
  // kt % jobs): every CTA carries 1 / jobs of the extra MMAs.  (Rotating per pixel range gave a third of the CTAs all of them.)
  const bool do_bias = P.db != nullptr;
#define WG_BIAS_TILE(kt_) (P.bias_range ? (blockIdx.x == blockIdx.y % gridDim.x) : (((kt_) % (int)gridDim.x) == (int)blockIdx.x))
  if (do_bias) {
    // bf16 1.0 = 0x3f80; all-ones is invariant under the 128-byte swizzle
    for (int i = threadIdx.x; i < WG_ONES_BYTES / 16; i += WG_THREADS)
      reinterpret_cast<uint4*>(ones)[i] = make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);
    tc::fence_proxy_async();                               // generic-proxy writes -> visible to the MMA's reads
  }
  if (warp == 5) { tc::tmem_alloc(tmem_base_slot, (uint32_t)P.tmem_cols); tc::tmem_relinquish(); }
  if (threadIdx.x == 0) OF_PDL_TRIGGER();
  OF_PDL_WAIT();                                     // nothing above touches global memory
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_base_slot;

  // job is the FAST grid index: the CTAs that walk the same pixel range (one per tap group / M tile) are
  // scheduled together, so x and dy come from HBM once and from L2 for the other jobs
  const int job = blockIdx.x;
  const int mt = P.job_mt[job], tap0 = P.job_tap0[job], ntaps = P.job_ntaps[job];
  // split-K ranges: contiguous by default; P.kt_step > 1 = INTERLEAVED (split s walks K tiles s, s + splits, ...), so
  // every CTA's pixels are spread over the whole tensor (both dies' memory) instead of one contiguous 1/splits of it -
  // ncu showed equal-work CTAs finishing 35 % apart with contiguous ranges (SM active min / avg / max 0 / 342k / 461k cycles)
  const int kt_step = P.kt_step;
  const int k_begin = kt_step > 1 ? (int)blockIdx.y : blockIdx.y * P.ktiles_per_split;
  const int k_end = kt_step > 1 ? P.ktiles : min(P.ktiles, k_begin + P.ktiles_per_split);
  const int a_bytes = P.nstack ? (P.nstack_one ? (P.one_split ? (tap0 == 0 ? 2 : 1) : 3) : 2) * WG_CHUNK_BYTES
                      : P.slab2 ? 3 * P.slab_bytes : P.slab ? 2 * P.slab_bytes : ntaps * 2 * WG_CHUNK_BYTES;
  const int tx_bytes = a_bytes + (P.nstack ? P.n_chunks * P.slab_bytes : P.n_chunks * WG_CHUNK_BYTES);

  if (warp == 4) {
    if (OF_ONE_LANE()) {
      int stage = 0; uint32_t phase = 0;
      unsigned* ctr = g_wg_ctr + (size_t)P.ctr_set * WG_CTR_STRIDE;
      int cur = k_begin, cur_end = k_end;                  // static: [k_begin, k_end) in steps of kt_step
      unsigned nxt = 0;
      if (P.dyn) { cur = cur_end = 0; nxt = atomicAdd(&ctr[job], (unsigned)P.claim); }
      int n_tiles = 0, any_bias = 0;
      for (;;) {
        int kt;
        if (P.dyn) {
          if (cur == cur_end) {                            // take the claim that was in flight, put the next one in flight
            cur = (int)min(nxt, (unsigned)P.ktiles);
            cur_end = min(cur + P.claim, P.ktiles);
            if (cur < P.ktiles) nxt = atomicAdd(&ctr[job], (unsigned)P.claim);
          }
          kt = cur < cur_end ? cur++ : -1;
        } else {
          kt = cur < cur_end ? cur : -1;
          cur += kt_step;
        }
        tc::mbar_wait(&empty_bar[stage], phase ^ 1);
        s_kt[stage] = kt;                                  // read by the issuer after its full_bar wait (arrive = release)
        if (kt < 0) { tc::mbar_arrive(&full_bar[stage]); break; }
        ++n_tiles; any_bias |= WG_BIAS_TILE(kt) ? 1 : 0;
        int tw_i = kt % P.tiles_w; int r = kt / P.tiles_w;
        int th_i = r % P.tiles_h; int n = r / P.tiles_h;
        int w0 = tw_i * P.TW, h0 = th_i * P.TH;
        uint8_t* base = tiles + (size_t)stage * P.stage_bytes;
        if ((OF_DBG(P) & 2) && n_tiles > P.stages) {       // profile hook: MMAs only (stages keep their first tiles)
          tc::mbar_arrive(&full_bar[stage]);
          if (++stage == P.stages) { stage = 0; phase ^= 1; }
          continue;
        }
        tc::mbar_expect_tx(&full_bar[stage], (uint32_t)tx_bytes);
        if (P.nstack) {
          const int nbox = P.nstack_one ? (P.one_split ? (tap0 == 0 ? 2 : 1) : 3) : 2;
          for (int c = 0; c < nbox; ++c) {
            const int chunk = P.nstack_one ? 0 : mt * 2 + c;
            const int kw = P.nstack_one ? (P.one_split ? tap0 + c : c) : tap0;
            const int cm = P.chunk_map[chunk];
            const CUtensorMap* tm = (cm == 1) ? &tmX1 : &tmX0;
            int cc = cm < 0 ? (1 << 20) : P.chunk_c0[chunk];
            tc::tma_load_4d(base + (size_t)c * WG_CHUNK_BYTES, tm, &full_bar[stage], cc, w0 + kw - P.pad_l, h0, n);
          }
          for (int c = 0; c < P.n_chunks; ++c)
            tc::tma_load_4d(base + a_bytes + (size_t)c * P.slab_bytes, &tmDY, &full_bar[stage], c * 64, w0, h0 - 1, n);
          if (++stage == P.stages) { stage = 0; phase ^= 1; }
          continue;
        }
        if (P.slab2) {
          for (int kw = 0; kw < 3; ++kw)
            tc::tma_load_4d(base + (size_t)kw * P.slab_bytes, &tmX0, &full_bar[stage], P.chunk_c0[0],
                            w0 + kw - P.pad_l, h0 - P.pad_t, n);
        } else if (P.slab) {
          for (int c = 0; c < 2; ++c) {
            const int cm = P.chunk_map[mt * 2 + c];
            const CUtensorMap* tm = (cm == 1) ? &tmX1 : &tmX0;
            int cc = cm < 0 ? (1 << 20) : P.chunk_c0[mt * 2 + c];
            tc::tma_load_4d(base + (size_t)c * P.slab_bytes, tm, &full_bar[stage], cc, w0 + tap0 - P.pad_l, h0 - P.pad_t, n);
          }
        } else
        for (int t = 0; t < ntaps; ++t) {
          for (int c = 0; c < 2; ++c) {
            const int tap = P.pack2 ? 2 * (tap0 + t) + c : tap0 + t;
            const int chunk = P.pack2 ? 0 : mt * 2 + c;
            int kh = tap / P.KW, kw = tap - kh * P.KW;
            int cw = w0 * P.stride_w + kw - P.pad_l, ch = h0 * P.stride_h + kh - P.pad_t;
            // a half with no source reads channel coordinate far outside the tensor -> TMA zero fill
            const int cm = tap < P.taps ? P.chunk_map[chunk] : -1;
            const CUtensorMap* tm = (cm == 1) ? &tmX1 : &tmX0;
            int cc = cm < 0 ? (1 << 20) : P.chunk_c0[chunk];
            tc::tma_load_4d(base + (size_t)(t * 2 + c) * WG_CHUNK_BYTES, tm, &full_bar[stage], cc, cw, ch, n);
          }
        }
        for (int c = 0; c < P.n_chunks; ++c)
          tc::tma_load_4d(base + a_bytes + (size_t)c * WG_CHUNK_BYTES, &tmDY, &full_bar[stage], c * 64, w0, h0, n);
        if (++stage == P.stages) { stage = 0; phase ^= 1; }
      }
      s_meta[0] = n_tiles; s_meta[1] = any_bias;
      tc::mbar_arrive(meta_bar);
      if (P.dyn) {
        // every claim of this CTA has returned (its value decided the loop's end); the last CTA to get here resets the set
        __threadfence();
        if (atomicAdd(&ctr[WG_MAX_JOBS], 1u) == gridDim.x * gridDim.y - 1) {
          for (int j = 0; j < (int)gridDim.x; ++j) ctr[j] = 0;
          ctr[WG_MAX_JOBS] = 0;
        }
      }
    }
  } else if (warp == 5) {
    if (OF_ONE_LANE()) {
      const uint32_t idesc = tc::idesc_bf16_f32(128, P.n_mma, 1, 1);
      int stage = 0; uint32_t phase = 0;
      bool bias_acc = false, first = true;
      for (;; first = false) {
        tc::mbar_wait(&full_bar[stage], phase);
        const int kt = s_kt[stage];
        if (kt < 0) break;
        const bool bias_now = do_bias && WG_BIAS_TILE(kt);
        tc::fence_after_sync();
        uint32_t base = tc::smem_u32(tiles + (size_t)stage * P.stage_bytes);
        uint32_t b_addr = base + (uint32_t)a_bytes;
        if (P.nstack) {
          const uint32_t idesc_n = tc::idesc_bf16_f32(128, 192, 1, 1);
          const uint32_t row = (uint32_t)(P.TW * 128);
          if (P.k_outer && !((OF_DBG(P) & 4) && !first)) {
            // K slice outermost: consecutive MMAs go to DIFFERENT accumulators (A/B of the accumulate dependency)
            for (int a = 0; a < ntaps; ++a) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                uint64_t adesc = tc::desc_sw128_mnmajor(base + (uint32_t)(a * 2 * WG_CHUNK_BYTES) + k * 2048, WG_CHUNK_BYTES);
                for (int c = 0; c < P.n_chunks; ++c) {
                  uint64_t bdesc = tc::desc_sw128_mnmajor(b_addr + (uint32_t)(c * P.slab_bytes) + k * 2048, row);
                  tc::mma_bf16(tmem_base + (uint32_t)((a * P.n_chunks + c) * 192), adesc, bdesc, idesc_n,
                               (!first || k > 0) ? 1u : 0u);
                }
              }
            }
          } else
          for (int a = 0; a < ((OF_DBG(P) & 4) && !first ? 0 : ntaps); ++a)      // profile hook 4: loads only
            for (int c = 0; c < P.n_chunks; ++c) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                uint64_t adesc = tc::desc_sw128_mnmajor(base + (uint32_t)(a * 2 * WG_CHUNK_BYTES) + k * 2048, WG_CHUNK_BYTES);
                uint64_t bdesc = tc::desc_sw128_mnmajor(b_addr + (uint32_t)(c * P.slab_bytes) + k * 2048, row);
                tc::mma_bf16(tmem_base + (uint32_t)((a * P.n_chunks + c) * 192), adesc, bdesc, idesc_n,
                             (!first || k > 0) ? 1u : 0u);
              }
            }
          if (bias_now && P.bias_t) {
            // transposed: D[channel][0..15] += dy tile^T (A, MN-major: M = 2 x 64 channels, one slab apart) x ones (B, N = 16);
            // an odd last chunk is paired with the ones tile (rows 64..127 are not read back)
            const uint32_t idesc_t = tc::idesc_bf16_f32(128, 16, 1, 1);
            const uint32_t ones_a = tc::smem_u32(ones);
            const uint64_t odesc = tc::desc_sw128_mnmajor(ones_a, 2048);
            for (int p = 0; 2 * p < P.n_chunks; ++p) {
              const uint32_t a0 = b_addr + (uint32_t)(2 * p * P.slab_bytes) + row;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint32_t lbo = (2 * p + 1 < P.n_chunks) ? (uint32_t)P.slab_bytes : ones_a - (a0 + k * 2048);
                tc::mma_bf16(tmem_base + (uint32_t)(ntaps * P.n_chunks * 192 + 16 * p),
                             tc::desc_sw128_mnmajor(a0 + k * 2048, lbo), odesc, idesc_t, (bias_acc || k > 0) ? 1u : 0u);
              }
            }
            bias_acc = true;
          } else if (bias_now) {
            // kh = 1 (one row into the slab) is dy unshifted; the chunks are one slab apart
            const uint32_t idesc_b = tc::idesc_bf16_f32(P.bias_m, P.n_chunks * 64, 1, 1);
            const uint64_t odesc = tc::desc_sw128_mnmajor(tc::smem_u32(ones), 2048);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              uint64_t bdesc = tc::desc_sw128_mnmajor(b_addr + row + k * 2048, (uint32_t)P.slab_bytes);
              tc::mma_bf16(tmem_base + (uint32_t)(ntaps * P.n_chunks * 192), odesc, bdesc, idesc_b,
                           (bias_acc || k > 0) ? 1u : 0u);
            }
            bias_acc = true;
          }
          tc::mma_commit(&empty_bar[stage]);
          if (++stage == P.stages) { stage = 0; phase ^= 1; }
          continue;
        }
        // tap outermost: the K slices of one accumulator back to back (interleaving the accumulators was 14 % slower)
        for (int t = 0; t < ntaps; ++t) {
          uint32_t a_addr = P.slab ? base + (uint32_t)(t * P.TW * 128) : base + (uint32_t)(t * 2 * WG_CHUNK_BYTES);
          uint32_t a_lbo = P.slab ? (uint32_t)P.slab_bytes : (uint32_t)WG_CHUNK_BYTES;
          if (P.slab2) {
            const uint32_t row = (uint32_t)(P.TW * 128);
            if (t < 3) { a_addr = base + (uint32_t)(t * P.slab_bytes); a_lbo = row; }
            else if (t == 3) { a_addr = base + 2 * row; a_lbo = (uint32_t)P.slab_bytes; }
            else { a_addr = base + (uint32_t)(2 * P.slab_bytes) + 2 * row; a_lbo = row; }
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            // 16 pixels (K) per MMA = 16 rows of 128 B = 2048 B
            uint64_t adesc = tc::desc_sw128_mnmajor(a_addr + k * 2048, a_lbo);
            uint64_t bdesc = tc::desc_sw128_mnmajor(b_addr + k * 2048, WG_CHUNK_BYTES);
            tc::mma_bf16(tmem_base + (uint32_t)(t * P.n_mma), adesc, bdesc, idesc, (!first || k > 0) ? 1u : 0u);
          }
        }
        if (bias_now && P.bias_t) {
          const uint32_t idesc_t = tc::idesc_bf16_f32(128, 16, 1, 1);
          const uint32_t ones_a = tc::smem_u32(ones);
          const uint64_t odesc = tc::desc_sw128_mnmajor(ones_a, 2048);
          for (int p = 0; 2 * p < P.n_chunks; ++p) {
            const uint32_t a0 = b_addr + (uint32_t)(2 * p * WG_CHUNK_BYTES);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint32_t lbo = (2 * p + 1 < P.n_chunks) ? (uint32_t)WG_CHUNK_BYTES : ones_a - (a0 + k * 2048);
              tc::mma_bf16(tmem_base + (uint32_t)(ntaps * P.n_mma + 16 * p), tc::desc_sw128_mnmajor(a0 + k * 2048, lbo), odesc,
                           idesc_t, (bias_acc || k > 0) ? 1u : 0u);
            }
          }
          bias_acc = true;
        } else if (bias_now) {
          const uint64_t odesc = tc::desc_sw128_mnmajor(tc::smem_u32(ones), 2048);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            uint64_t bdesc = tc::desc_sw128_mnmajor(b_addr + k * 2048, WG_CHUNK_BYTES);
            tc::mma_bf16(tmem_base + (uint32_t)(ntaps * P.n_mma), odesc, bdesc, tc::idesc_bf16_f32(P.bias_m, P.n_mma, 1, 1),
                         (bias_acc || k > 0) ? 1u : 0u);
          }
          bias_acc = true;
        }
        tc::mma_commit(&empty_bar[stage]);
        if (++stage == P.stages) { stage = 0; phase ^= 1; }
      }
      tc::mma_commit(acc_full);
    }
  } else {
    // epilogue: row = input channel (M), columns = output channels
    tc::mbar_wait(meta_bar, 0);
    const bool has_work = s_meta[0] > 0;
    if (has_work) {
      tc::mbar_wait(acc_full, 0);
      tc::fence_after_sync();
      int row = warp * 32 + lane;
      if (P.nstack) {
        const int chunk = P.nstack_one ? 0 : mt * 2 + (row >> 6), rcn = row & 63;
        const bool rok = P.chunk_map[chunk] >= 0 && rcn < P.chunk_cnt[chunk];
        const int cin = P.chunk_ci[chunk] + rcn;
        for (int a = 0; a < ntaps; ++a) {
          const int kw = P.nstack_one ? (P.one_split ? tap0 : 2 * a) + (row >> 6) : tap0;
          const bool valid = rok && kw < 3;
          const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(a * P.n_chunks * 192);
          for (int c0 = 0; c0 < P.n_chunks * 192; c0 += 16) {
            const int kh = 2 - ((c0 % 192) >> 6), co0 = (c0 / 192) * 64 + (c0 & 63);
            if (co0 >= P.n_valid) continue;                       // warp-uniform
            __syncwarp();
            uint32_t rr[16];
            tc::tmem_ld16(taddr + (uint32_t)c0, rr);
            tc::tmem_ld_wait();
            if (!valid || (OF_DBG(P) & 1)) continue;
            float* dst = P.dw + ((size_t)(kh * 3 + kw) * P.Cin + cin) * P.Cout + co0;
            const int nv = min(16, P.n_valid - co0);
            if ((P.Cout % 4) == 0 && nv == 16) {
#pragma unroll
              for (int j = 0; j < 4; ++j)
                red_add_v4(dst + 4 * j, make_float4(__uint_as_float(rr[4 * j]), __uint_as_float(rr[4 * j + 1]),
                                                    __uint_as_float(rr[4 * j + 2]), __uint_as_float(rr[4 * j + 3])));
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < nv) red_add(dst + j, __uint_as_float(rr[j]));
            }
          }
        }
      }
      const int c = P.pack2 ? 0 : mt * 2 + (row >> 6), rc = row & 63;
      const bool row_ok = P.chunk_map[c] >= 0 && rc < P.chunk_cnt[c];
      int ci = P.chunk_ci[c] + rc;
      const bool vec4 = (P.Cout % 4) == 0;
      for (int t = 0; t < (P.nstack ? 0 : ntaps); ++t) {
        int tap = P.slab ? t * P.KW + tap0 : P.pack2 ? 2 * (tap0 + t) + (row >> 6) : tap0 + t;
        if (P.slab2) tap = t < 3 ? (row >> 6) * 3 + t : t == 3 ? 6 + (row >> 6) : 8 + (row >> 6);
        const bool valid = row_ok && tap < P.taps;
        float* dst_row = P.dw + ((size_t)(valid ? tap : 0) * P.Cin + (valid ? ci : 0)) * P.Cout;
        uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(t * P.n_mma);
        for (int c0 = 0; c0 < P.n_mma; c0 += 16) {
          __syncwarp();
          uint32_t rr[16];
          tc::tmem_ld16(taddr + (uint32_t)c0, rr);
          tc::tmem_ld_wait();
          if (valid && c0 < P.n_valid && !(OF_DBG(P) & 1)) {
            int nv = min(16, P.n_valid - c0);
            if (vec4 && nv == 16) {
#pragma unroll
              for (int j = 0; j < 4; ++j)
                red_add_v4(dst_row + c0 + 4 * j,
                           make_float4(__uint_as_float(rr[4 * j]), __uint_as_float(rr[4 * j + 1]),
                                       __uint_as_float(rr[4 * j + 2]), __uint_as_float(rr[4 * j + 3])));
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (j < nv) red_add(dst_row + c0 + j, __uint_as_float(rr[j]));
            }
          }
        }
      }
      const bool had_bias = do_bias && s_meta[1] != 0;     // did this CTA run any bias tile?
      if (had_bias && P.bias_t) {
        // transposed accumulators: row = channel (chunk pair p: 2 x 64 rows), every column holds the sum
        const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) +
                               (uint32_t)(ntaps * (P.nstack ? P.n_chunks * 192 : P.n_mma));
        for (int p = 0; 2 * p < P.n_chunks; ++p) {
          __syncwarp();
          uint32_t rr[16];
          tc::tmem_ld16(taddr + (uint32_t)(16 * p), rr);
          tc::tmem_ld_wait();
          const int ch = 2 * p + (row >> 6), co = ch * 64 + (row & 63);
          if (ch < P.n_chunks && co < P.n_valid) red_add(P.db + co, __uint_as_float(rr[0]));
        }
      } else if (had_bias && warp == 0) {
        // every row of the ones-accumulator holds the column sums; row 0 (lane 0) reduces them into db
        uint32_t taddr = tmem_base + (uint32_t)(ntaps * (P.nstack ? P.n_chunks * 192 : P.n_mma));
        for (int c0 = 0; c0 < P.n_mma; c0 += 16) {
          __syncwarp();
          uint32_t rr[16];
          tc::tmem_ld16(taddr + (uint32_t)c0, rr);
          tc::tmem_ld_wait();
          if (lane == 0) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c0 + j < P.n_valid) red_add(P.db + c0 + j, __uint_as_float(rr[j]));
          }
        }
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 5) tc::tmem_dealloc(tmem_base, (uint32_t)P.tmem_cols);
}

// ---- CTA-pair variant (cta_group::2) of the N-stacked mode with two dy chunks (Cout 65..128, Cin > 64) ----------------
// Two jobs (M tile, kw) of one pixel range run as a CTA pair on the two SMs of a TPC: each CTA loads its own x tile
// (A, 128 rows = 128 input channels, 16 KB per K tile) and ONE of the two dy slabs (CTA r: output channels 64 r ..); an
// MMA of M = 256 takes B's columns half from each CTA.  Per K slice of 16 pixels: N = 256 (per CTA the kh = 2 and kh = 1
// views of its slab) and N = 128 (kh = 0).  Accumulator columns of BOTH CTAs (rows = the CTA's own input channels):
//   [  0, 256): [kh2 | kh1] of chunk 0, [kh2 | kh1] of chunk 1  (64 columns each)     [256, 384): kh0 of chunk 0 | chunk 1
// Per SM and K tile 28 KB of loads / 14 KB of MMA operand reads per K slice instead of 40 KB / 20 KB.  No bias gradient
// here: the launch that handles the unpaired job(s) of the layer computes it.
__global__ void __launch_bounds__(WG_THREADS, 1)
conv_tc_wgrad_pair_kernel(const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1,
                          const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ WgradParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(tiles + (size_t)P.stages * P.stage_bytes);
  uint64_t* empty_bar = full_bar + 8;
  uint64_t* acc_full = empty_bar + 8;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(acc_full + 1);
  int* s_kt = reinterpret_cast<int*>(tmem_base_slot + 2);
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = tc::cluster_ctarank();               // cluster (2, 1, 1): rank = blockIdx.x & 1

  if (threadIdx.x == 0) {
    for (int s = 0; s < P.stages; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
    tc::mbar_init(acc_full, 1);
    tc::fence_barrier_init();
  }
  if (warp == 4 && tc::elect_one()) { tc::prefetch_tmap(&tmX0); tc::prefetch_tmap(&tmX1); tc::prefetch_tmap(&tmDY); }
  if (warp == 5) { tc::tmem_alloc_2cta(tmem_base_slot, 512u); tc::tmem_relinquish_2cta(); }
  if (threadIdx.x == 0) OF_PDL_TRIGGER();
  OF_PDL_WAIT();
  tc::fence_before_sync();
  tc::cluster_sync();                                        // both CTAs' barriers are initialised, TMEM is allocated
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_base_slot;

  const int job = (int)(blockIdx.x & ~1u) + (int)rank;
  const int mt = P.job_mt[job], kw = P.job_tap0[job];
  const int kt_step = P.kt_step;
  const int k_begin = kt_step > 1 ? (int)blockIdx.y : blockIdx.y * P.ktiles_per_split;
  const int k_end = kt_step > 1 ? P.ktiles : min(P.ktiles, k_begin + P.ktiles_per_split);
  const int a_bytes = 2 * WG_CHUNK_BYTES;
  const uint32_t tx_pair = 2u * (uint32_t)(a_bytes + P.slab_bytes);

  if (warp == 4) {
    if (tc::elect_one()) {
      int stage = 0; uint32_t phase = 0;
      for (int kt = k_begin;; kt += kt_step) {
        tc::mbar_wait(&empty_bar[stage], phase ^ 1);
        if (kt >= k_end) {
          if (rank == 0) { s_kt[stage] = -1; tc::mbar_arrive(&full_bar[stage]); }
          break;
        }
        if (rank == 0) { s_kt[stage] = kt; tc::mbar_expect_tx(&full_bar[stage], tx_pair); }
        int tw_i = kt % P.tiles_w; int r = kt / P.tiles_w;
        int th_i = r % P.tiles_h; int n = r / P.tiles_h;
        int w0 = tw_i * P.TW, h0 = th_i * P.TH;
        uint8_t* base = tiles + (size_t)stage * P.stage_bytes;
        for (int c = 0; c < 2; ++c) {
          const int chunk = mt * 2 + c;
          const int cm = P.chunk_map[chunk];
          const CUtensorMap* tm = (cm == 1) ? &tmX1 : &tmX0;
          int cc = cm < 0 ? (1 << 20) : P.chunk_c0[chunk];
          tc::tma_load_4d_2cta(base + (size_t)c * WG_CHUNK_BYTES, tm, &full_bar[stage], cc, w0 + kw - P.pad_l, h0, n);
        }
        tc::tma_load_4d_2cta(base + a_bytes, &tmDY, &full_bar[stage], (int)rank * 64, w0, h0 - 1, n);
        if (++stage == P.stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 5) {
    if (rank == 0 && tc::elect_one()) {
      const uint32_t idesc1 = tc::idesc_bf16_f32(256, 256, 1, 1), idesc2 = tc::idesc_bf16_f32(256, 128, 1, 1);
      const uint32_t row = (uint32_t)(P.TW * 128);
      int stage = 0; uint32_t phase = 0;
      for (bool first = true;; first = false) {
        tc::mbar_wait(&full_bar[stage], phase);
        if (s_kt[stage] < 0) break;
        tc::fence_after_sync();
        const uint32_t base = tc::smem_u32(tiles + (size_t)stage * P.stage_bytes);
        const uint32_t b_addr = base + (uint32_t)a_bytes;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          tc::mma_bf16_2cta(tmem_base, tc::desc_sw128_mnmajor(base + k * 2048, WG_CHUNK_BYTES),
                            tc::desc_sw128_mnmajor(b_addr + k * 2048, row), idesc1, (!first || k > 0) ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          tc::mma_bf16_2cta(tmem_base + 256u, tc::desc_sw128_mnmajor(base + k * 2048, WG_CHUNK_BYTES),
                            tc::desc_sw128_mnmajor(b_addr + 2 * row + k * 2048, row), idesc2, (!first || k > 0) ? 1u : 0u);
        tc::mma_commit_2cta(&empty_bar[stage]);
        if (++stage == P.stages) { stage = 0; phase ^= 1; }
      }
      tc::mma_commit_2cta(acc_full);
    }
  } else {
    tc::mbar_wait(acc_full, 0);
    tc::fence_after_sync();
    const int lane = threadIdx.x & 31;
    const int row = warp * 32 + lane;
    const int chunk = mt * 2 + (row >> 6), rcn = row & 63;
    const bool valid = P.chunk_map[chunk] >= 0 && rcn < P.chunk_cnt[chunk];
    const int cin = P.chunk_ci[chunk] + rcn;
    const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16);
    for (int c0 = 0; c0 < 384; c0 += 16) {
      int kh, co0;
      if (c0 < 256) { kh = 2 - ((c0 & 127) >> 6); co0 = (c0 >> 7) * 64 + (c0 & 63); }
      else          { kh = 0; co0 = ((c0 - 256) >> 6) * 64 + (c0 & 63); }
      if (co0 >= P.n_valid) continue;                        // warp-uniform
      __syncwarp();
      uint32_t rr[16];
      tc::tmem_ld16(taddr + (uint32_t)c0, rr);
      tc::tmem_ld_wait();
      if (!valid) continue;
      float* dst = P.dw + ((size_t)(kh * 3 + kw) * P.Cin + cin) * P.Cout + co0;
      const int nv = min(16, P.n_valid - co0);
      if ((P.Cout % 4) == 0 && nv == 16) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          red_add_v4(dst + 4 * j, make_float4(__uint_as_float(rr[4 * j]), __uint_as_float(rr[4 * j + 1]),
                                              __uint_as_float(rr[4 * j + 2]), __uint_as_float(rr[4 * j + 3])));
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          if (j < nv) red_add(dst + j, __uint_as_float(rr[j]));
      }
    }
  }
  tc::fence_before_sync();
  tc::cluster_sync();                                        // both epilogues have read their TMEM, no arrive is in flight
  if (warp == 5) tc::tmem_dealloc_2cta(tmem_base, 512u);
}

extern "C" {

int of_conv_tc_wgrad(const of_tc_wgrad* c, void* stream) {
  OF_REQUIRE(c && c->x[0] && c->dy && c->dw, OF_ERR_ARG, "null pointer argument");
  OF_REQUIRE(c->num_src == 1 || c->num_src == 2, OF_ERR_ARG, "num_src must be 1 or 2");
  OF_REQUIRE(c->stride == 1 || c->stride == 2, OF_ERR_UNSUPPORTED, "stride %d unsupported", c->stride);
  OF_REQUIRE(c->Cout >= 1 && c->Cout <= 256 && c->dy_c >= c->Cout, OF_ERR_SHAPE, "Cout %d unsupported", c->Cout);
  of_prof_note(2.0 * c->N * c->Ho * c->Wo * (double)c->Cin * c->Cout * c->KH * c->KW);
  cudaStream_t st = (cudaStream_t)stream;
  const int taps = c->KH * c->KW;
  OF_CUDA(cudaMemsetAsync(c->dw, 0, (size_t)taps * c->Cin * c->Cout * sizeof(float), st));
  if (c->db) OF_CUDA(cudaMemsetAsync(c->db, 0, (size_t)c->Cout * sizeof(float), st));

  WgradParams P;
  memset(&P, 0, sizeof(P));
  P.KW = c->KW; P.pad_t = c->pad_t; P.pad_l = c->pad_l;
  P.stride_w = c->stride_w > 0 ? c->stride_w : c->stride;
  P.stride_h = c->stride_h > 0 ? c->stride_h : c->stride;
  const bool view = c->view_w_stride > 0;
  OF_REQUIRE(!view || c->num_src == 1, OF_ERR_ARG, "a strided view takes one source");
  P.NB = c->N; P.Ho = c->Ho; P.Wo = c->Wo;
  const int cands[4] = {16, 32, 8, 64};
  P.TW = 16; P.TH = 4;
  for (int k = 0; k < 4; ++k)
    if (c->Wo % cands[k] == 0) { P.TW = cands[k]; P.TH = 64 / cands[k]; break; }
  // A/B of the slab halo (same MMAs, different L2 -> SM bytes per K tile: 8x8 36 KB, 16x4 40 KB, 32x2 48 KB at 128->128).
  // Round 2, call 45: 258.7 / 254.0 / 268.1 us - +20 % bytes cost 5 %, -10 % bytes nothing: not the L2 -> SM path.
  if (const char* e = of_tune_env("OF_WG_TW")) {
    int v = atoi(e);
    if ((v == 8 || v == 16 || v == 32) && c->Wo % v == 0 && c->Ho % (64 / v) == 0) { P.TW = v; P.TH = 64 / v; }
  }
  P.tiles_w = (c->Wo + P.TW - 1) / P.TW;
  P.tiles_h = (c->Ho + P.TH - 1) / P.TH;
  long long kt = (long long)c->N * P.tiles_w * P.tiles_h;
  OF_REQUIRE(kt > 0 && kt < (1ll << 30), OF_ERR_SHAPE, "bad pixel tiling");
  P.ktiles = (int)kt;
  P.n_valid = c->Cout;
  P.n_mma = (c->Cout + 15) / 16 * 16;
  P.n_chunks = (P.n_mma + 63) / 64;
  P.Cin = c->Cin; P.Cout = c->Cout;
  P.dw = c->dw;
  P.db = c->db;
  // Bias gradient: MMAs against an all-ones tile into a spare accumulator.  Default (round 2, call 53): TRANSPOSED - A = the
  // unshifted dy tile (MN-major, M = its channels, two 64-channel chunks per MMA), B = ones, N = 16: 1 / 8 of the tensor
  // time of the first form (A = ones with 128 identical rows, B = dy, N = Cout; option OF_WG_BIAS_ONES_A).  Measured:
  // 16 x 96 x 320 64->64 69.0 -> 62.4 us (no bias gradient at all: 61.0), 96->64 140.6 -> 131.7 (126.4), 128->128
  // 261 -> 262 (236: with one CTA per SM the cost is the switch of accumulator, not the MMA's size); step: wgrad
  // 2 514 -> 2 452 us.  The bias MMAs still cost 3-10 % of the kernel.  Three SIMT replacements by the four epilogue warps,
  // which idle during the main loop, were built and measured (calls 48-50) - all slower, none kept: column sums from the
  // shared-memory ring (stage released by MMA commit + 4 warp arrives): 128->128 261 -> 247 us but 96->64 136 -> 166 us,
  // 64->64 66 -> 79 us, wgrad per step 2 433 -> 2 541 us; the same with one polling lane per warp or one bias job per
  // range: no better; column sums read once more through L2 with __ldg: 2-3x slower (one tile's loads in flight at a time).
  // M = 64 for the ones-A form (OF_WG_BIAS_M64): no change.
  const bool mma_bias = c->db != nullptr;
  P.bias_m = of_tune_env("OF_WG_BIAS_M64") ? 64 : 128;
  P.bias_t = of_tune_env("OF_WG_BIAS_ONES_A") ? 0 : 1;
  P.k_outer = of_tune_env("OF_WG_KOUTER") ? 1 : 0;
  const int bias_cols_t = 32;                                   // 16 columns per pair of dy chunks, n_chunks <= 4
  OF_REQUIRE(!mma_bias || P.n_mma + (P.bias_t ? bias_cols_t : P.n_mma) <= 512, OF_ERR_UNSUPPORTED,
             "fused bias gradient needs a spare TMEM accumulator (Cout <= 256 -> n_mma <= 256)");
  // taps per pass: one TMEM accumulator (n_mma columns) per tap, one more for the fused bias gradient
  int tpp = (512 - (mma_bias ? (P.bias_t ? bias_cols_t : P.n_mma) : 0)) / P.n_mma;
  if (tpp > taps) tpp = taps;
  // keep >= 2 pipeline stages inside 200 KB of shared memory
  while (tpp > 1 && 2 * (tpp * 2 * WG_CHUNK_BYTES + P.n_chunks * WG_CHUNK_BYTES) > 200 * 1024) --tpp;
  int passes = (taps + tpp - 1) / tpp;
  tpp = (taps + passes - 1) / passes;   // balance the groups

  // (slab mode is decided before the tensor maps: its x boxes are TH + 2 rows tall)
  int x_chunks = 0;
  for (int s_ = 0; s_ < c->num_src; ++s_) x_chunks += (c->x_valid[s_] + 63) / 64;
  P.slab = (c->KH == 3 && c->KW == 3 && P.stride_w == 1 && P.stride_h == 1 && !view && x_chunks >= 2 &&
            (512 - (mma_bias ? (P.bias_t ? bias_cols_t : P.n_mma) : 0)) / P.n_mma >= 3 && P.TW >= 8 && P.TW <= 32 && !of_tune_env("OF_WG_NO_SLAB"))
               ? 1 : 0;
  P.slab_bytes = (P.TH + 2) * P.TW * 128;
  if (P.slab) { tpp = 3; passes = 3; }
  P.slab2 = (x_chunks == 1 && !of_tune_env("OF_WG_NO_PACK2") && c->KH == 3 && c->KW == 3 && P.stride_w == 1 &&
             P.stride_h == 1 && !view && P.TW >= 8 && P.TW <= 32 && 5 * P.n_mma + (mma_bias ? (P.bias_t ? bias_cols_t : P.n_mma) : 0) <= 512 &&
             !of_tune_env("OF_WG_NO_SLAB2"))
                ? 1 : 0;
  P.nstack = (c->KH == 3 && c->KW == 3 && P.stride_w == 1 && P.stride_h == 1 && !view &&
              (P.n_chunks == 1 || (P.n_chunks == 2 && x_chunks >= 2 && !of_tune_env("OF_WG_NO_NSTACK2"))) &&
              P.TW >= 8 && P.TW <= 32 && !of_tune_env("OF_WG_NO_NSTACK"))
                 ? 1 : 0;
  P.nstack_one = (P.nstack && x_chunks == 1) ? 1 : 0;
  P.one_split = (P.nstack_one && P.n_chunks == 1 && !of_tune_env("OF_WG_NO_ONE_SPLIT")) ? 1 : 0;
  if (P.nstack) { P.slab = 0; P.slab2 = 0; }
  CUtensorMap tmX[2], tmDY;
  memset(tmX, 0, sizeof(tmX));
  for (int s = 0; s < c->num_src; ++s) {
    long long ws_ = view ? c->view_w_stride : c->x_stride[s];
    long long hs_ = view ? c->view_h_stride : (long long)c->W * c->x_stride[s];
    long long ns_ = view ? c->view_n_stride : (long long)c->H * c->W * c->x_stride[s];
    if (int r = of_tc_make_view_map(&tmX[s], c->x[s], c->x_c[s], ws_, hs_, ns_, c->N, c->H, c->W, P.TW,
                                    (P.slab || P.slab2) ? P.TH + 2 : P.TH, P.stride_w, P.stride_h))
      return r;
  }
  if (c->num_src == 1) tmX[1] = tmX[0];
  if (int r = of_tc_make_act_map(&tmDY, c->dy, c->dy_c, c->dy_stride, c->N, c->Ho, c->Wo, P.TW,
                                 P.nstack ? P.TH + 2 : P.TH, 1))
    return r;

  // list of 64-channel chunks over the sources (original ci offsets follow the concat order)
  int nch = 0;
  for (int k = 0; k < WG_MAX_CHUNKS; ++k) P.chunk_map[k] = -1;
  for (int s = 0; s < c->num_src; ++s)
    for (int c0 = 0; c0 < c->x_c[s]; c0 += 64) {
      int valid_c = c->x_valid[s] - c0;
      if (valid_c <= 0) break;
      OF_REQUIRE(nch < WG_MAX_CHUNKS, OF_ERR_SHAPE, "too many input channels");
      P.chunk_map[nch] = s; P.chunk_c0[nch] = c0; P.chunk_ci[nch] = c->x_ci0[s] + c0;
      P.chunk_cnt[nch] = valid_c < 64 ? valid_c : 64;
      ++nch;
    }
  static bool attr_set = false;
  if (!attr_set) {
    OF_CUDA(cudaFuncSetAttribute(conv_tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_set = true;
  }
  P.taps = taps;
  P.pack2 = (nch == 1 && taps > 1 && !P.nstack && !of_tune_env("OF_WG_NO_PACK2")) ? 1 : 0;
  const int units = P.pack2 ? (taps + 1) / 2 : taps;       // accumulators needed per M tile
  if (P.pack2) {
    tpp = (512 - (mma_bias ? (P.bias_t ? bias_cols_t : P.n_mma) : 0)) / P.n_mma;
    if (tpp > units) tpp = units;
    while (tpp > 1 && 2 * (tpp * 2 * WG_CHUNK_BYTES + P.n_chunks * WG_CHUNK_BYTES) > 200 * 1024) --tpp;
    passes = (units + tpp - 1) / tpp;
    tpp = (units + passes - 1) / passes;
  }
  if (P.slab2) { tpp = 5; passes = 1; }
  int m_tiles = (nch + 1) / 2;
  int jobs = 0;
  if (P.nstack) {
    tpp = (P.nstack_one && !P.one_split) ? 2 : 1;
    for (int mt = 0; mt < m_tiles; ++mt)
      for (int kw = 0; kw < (P.nstack_one ? (P.one_split ? 4 : 1) : 3); kw += (P.one_split ? 2 : 1)) {
        OF_REQUIRE(jobs < WG_MAX_JOBS, OF_ERR_SHAPE, "too many wgrad jobs");
        P.job_mt[jobs] = mt; P.job_tap0[jobs] = kw; P.job_ntaps[jobs] = tpp;
        ++jobs;
      }
  } else
  for (int mt = 0; mt < m_tiles; ++mt)
    for (int p = 0; p < passes; ++p) {
      int t0 = P.slab ? p : p * tpp;                  // slab: the job's kw; its three taps are kh = 0..2
      int nt = P.slab ? 3 : (units - t0 < tpp ? units - t0 : tpp);
      if (nt <= 0) break;
      OF_REQUIRE(jobs < WG_MAX_JOBS, OF_ERR_SHAPE, "too many wgrad jobs");
      P.job_mt[jobs] = mt; P.job_tap0[jobs] = t0; P.job_ntaps[jobs] = nt;
      ++jobs;
    }
  // CTA pairs (conv_tc_wgrad_pair_kernel) for the wide N-stacked layers at the large levels: the jobs are paired two by
  // two in one launch; the unpaired job (and, when the bias gradient is wanted, at least one job) goes through the
  // one-CTA kernel below, which also computes the bias gradient.
  const int pair_min = of_tune_env("OF_WG_PAIR_MIN") ? atoi(of_tune_env("OF_WG_PAIR_MIN")) : 4096;
  if (P.nstack && !P.nstack_one && P.n_chunks == 2 && P.ktiles >= pair_min && of_tune_env("OF_WG_PAIR")) {
    const int n_single = (jobs & 1) ? 1 : (c->db ? 2 : 0);
    const int n_pairjobs = (jobs - n_single) / 2;
    if (n_pairjobs >= 1) {
      static bool pair_attr_set = false;
      if (!pair_attr_set) {
        OF_CUDA(cudaFuncSetAttribute(conv_tc_wgrad_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        pair_attr_set = true;
      }
      WgradParams PP = P;
      PP.db = nullptr;
      PP.stage_bytes = 2 * WG_CHUNK_BYTES + P.slab_bytes;
      PP.stages = (200 * 1024) / PP.stage_bytes;
      if (PP.stages > 8) PP.stages = 8;
      PP.tmem_cols = 512;
      cudaLaunchAttribute pat[2];
      cudaLaunchConfig_t pcfg = {};
      pcfg.blockDim = dim3(WG_THREADS);
      pcfg.dynamicSmemBytes = 1024 + (size_t)PP.stages * PP.stage_bytes + 256;
      pcfg.stream = st;
      pat[0].id = cudaLaunchAttributeClusterDimension;
      pat[0].val.clusterDim.x = 2; pat[0].val.clusterDim.y = 1; pat[0].val.clusterDim.z = 1;
      pat[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      pat[1].val.programmaticStreamSerializationAllowed = of_tune_env("OF_NO_PDL") ? 0 : 1;
      pcfg.attrs = pat; pcfg.numAttrs = 2;
      int psplits = (OF_NUM_SMS / 2) / n_pairjobs;               // one wave of CTA pairs
      pcfg.gridDim = dim3(2 * n_pairjobs, psplits);
      int max_clusters = 0;
      if (cudaOccupancyMaxActiveClusters(&max_clusters, conv_tc_wgrad_pair_kernel, &pcfg) == cudaSuccess && max_clusters >= 1) {
        if (psplits * n_pairjobs > max_clusters) psplits = max_clusters / n_pairjobs;
      } else {
        cudaGetLastError();
      }
      if (psplits > PP.ktiles) psplits = PP.ktiles;
      if (psplits < 1) psplits = 1;
      PP.ktiles_per_split = (PP.ktiles + psplits - 1) / psplits;
      psplits = (PP.ktiles + PP.ktiles_per_split - 1) / PP.ktiles_per_split;
      PP.kt_step = of_tune_env("OF_WG_CONTIGUOUS") ? 1 : psplits;
      pcfg.gridDim = dim3(2 * n_pairjobs, psplits);
      OF_CUDA(cudaLaunchKernelEx(&pcfg, conv_tc_wgrad_pair_kernel, tmX[0], tmX[1], tmDY, PP));
      if (int r = of_launch_check("conv_tc_wgrad_pair_kernel")) return r;
      if (n_single == 0) return 0;
      for (int j = 0; j < n_single; ++j) {
        P.job_mt[j] = P.job_mt[jobs - n_single + j]; P.job_tap0[j] = P.job_tap0[jobs - n_single + j];
        P.job_ntaps[j] = P.job_ntaps[jobs - n_single + j];
      }
      jobs = n_single;
    }
  }
  // ONE wave of one CTA per SM (never round up: 99 x 3 = 297 CTAs leaves one CTA for a wave of its own); a second
  // wave only repeats the prologue and doubles the split-K atomics (measured 3-15 % slower)
  int splits = OF_NUM_SMS / jobs;
  // N-stacked, several input chunks, Cout <= 64: 256 TMEM columns and a 28 KB stage - TWO CTAs fit an SM and fill
  // each other's bubbles (96->64 at N8 192x640: 198 -> 150 us)
  // (also the stem's strided-view wgrad: its overlapping-window boxes fill shared memory at ~20 B/clk per CTA)
  const int acc_cols = P.nstack ? tpp * P.n_chunks * 192 + (mma_bias ? (P.bias_t ? bias_cols_t : P.n_chunks * 64) : 0) : (tpp * P.n_mma + (mma_bias ? (P.bias_t ? bias_cols_t : P.n_mma) : 0));
  const bool two_per_sm = ((P.nstack && (!P.nstack_one || P.one_split) && P.n_chunks == 1) || (view && acc_cols <= 256)) &&
                          !of_tune_env("OF_WG_NO_2CTA");
  if (two_per_sm) splits = (2 * OF_NUM_SMS) / jobs;
  if (const char* e = of_tune_env("OF_WG_WAVES")) splits = (atoi(e) * OF_NUM_SMS) / jobs;
  P.stage_bytes = (P.slab2 ? 3 * P.slab_bytes : P.slab ? 2 * P.slab_bytes : tpp * 2 * WG_CHUNK_BYTES) +
                  P.n_chunks * WG_CHUNK_BYTES;
  if (P.nstack) P.stage_bytes = ((P.nstack_one && !P.one_split) ? 3 : 2) * WG_CHUNK_BYTES + P.n_chunks * P.slab_bytes;
  P.stages = (200 * 1024) / P.stage_bytes;
  if (P.stages > 8) P.stages = 8;
  if (two_per_sm) { int cap = (100 * 1024) / P.stage_bytes; if (cap < 2) cap = 2; if (P.stages > cap) P.stages = cap; }
  if (const char* e = of_tune_env("OF_WG_STAGES")) { int v = atoi(e); if (v >= 1 && v < P.stages) P.stages = v; }
  OF_REQUIRE(P.stages >= 1, OF_ERR_SHAPE, "wgrad stage does not fit shared memory");
  int cols = (tpp * P.n_mma + (mma_bias ? (P.bias_t ? bias_cols_t : P.n_mma) : 0));
  if (P.nstack) cols = tpp * P.n_chunks * 192 + (mma_bias ? (P.bias_t ? bias_cols_t : P.n_chunks * 64) : 0);
  P.tmem_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : cols <= 256 ? 256 : 512;
  size_t smem = 1024 + (size_t)P.stages * P.stage_bytes + WG_ONES_BYTES + 256;   // barriers, TMEM slot, s_kt, s_meta

  // The jobs of one pixel range (kw columns / M tiles) read the same dy slabs and nearly the same x tiles; ncu shows 1 018 MB
  // of DRAM reads per 128->128 launch against 504 MB of operands and an L2 read hit rate of 40 %.  Option OF_WG_CLUSTER
  // launches them as a thread-block CLUSTER (jobs, 1, 1) - co-scheduled on one GPC, same L2 partition, same start time.
  // Measured (round 2, call 41): SLOWER - wgrad 2.39 -> 2.55 ms per step, 128->128 255 -> 262 us, 96->64 131 -> 187 us -
  // so it is off by default.
  cudaLaunchAttribute at[2];
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(WG_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = of_tune_env("OF_NO_PDL") ? 0 : 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  int cluster = (jobs >= 2 && jobs <= 8 && of_tune_env("OF_WG_CLUSTER")) ? jobs : 1;
  if (cluster > 1) {
    at[1].id = cudaLaunchAttributeClusterDimension;
    at[1].val.clusterDim.x = (unsigned)cluster; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
    cfg.numAttrs = 2;
    cfg.gridDim = dim3(jobs, splits);
    int max_clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&max_clusters, conv_tc_wgrad_kernel, &cfg) == cudaSuccess && max_clusters >= 1) {
      if (splits > max_clusters) splits = max_clusters;            // still ONE wave
    } else {
      cudaGetLastError();
      cluster = 1; cfg.numAttrs = 1;
    }
  }
  if (splits > P.ktiles) splits = P.ktiles;
  if (splits < 1) splits = 1;
  P.ktiles_per_split = (P.ktiles + splits - 1) / splits;
  splits = (P.ktiles + P.ktiles_per_split - 1) / P.ktiles_per_split;
  if (const char* e = of_tune_env("OF_WG_DBG")) P.dbg = atoi(e);
  P.kt_step = of_tune_env("OF_WG_CONTIGUOUS") ? 1 : splits;     // interleaved K tiles (every split has >= 1 tile: splits <= ktiles)
  // dynamic K-tile scheduling (see g_wg_ctr; opt-in): ~16 claims per CTA, at most 4 tiles each
  static std::atomic<unsigned> launch_seq{0};
  P.dyn = of_tune_env("OF_WG_DYNAMIC") ? 1 : 0;
  P.bias_range = (!P.dyn && of_tune_env("OF_WG_BIAS_RANGE")) ? 1 : 0;
  P.ctr_set = (int)(launch_seq.fetch_add(1) % WG_CTR_SETS);
  P.claim = P.ktiles / (splits * 16);
  if (P.claim < 1) P.claim = 1;
  if (P.claim > 4) P.claim = 4;
  if (const char* e = of_tune_env("OF_WG_CLAIM")) { int v = atoi(e); if (v >= 1 && v <= 64) P.claim = v; }
  cfg.gridDim = dim3(jobs, splits);
  OF_CUDA(cudaLaunchKernelEx(&cfg, conv_tc_wgrad_kernel, tmX[0], tmX[1], tmDY, P));
  if (int r = of_launch_check(view ? "conv_tc_wgrad_kernel<stem>" : "conv_tc_wgrad_kernel")) return r;
  return 0;
}

}  // extern "C"
