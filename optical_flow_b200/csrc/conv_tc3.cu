// K10 fast path: 3x3 / stride-1 convolution (fprop and dgrad) as a tcgen05 / TMEM implicit GEMM with
// operand REUSE in shared memory - the layers that hold > 90 % of the conv FLOPs of the step
// (model.py:104-114 decoder chains, resnet/models.py:17-24 residual blocks).
//
// conv_tc.cu reloads a 128-pixel x 64-channel A tile and a weight tile for each of the 9 taps:
// 288 KB of L2 -> SM traffic per 128 pixels per 64 channels at Cout = 128, which bounds the kernel by
// L2 bandwidth (~650 TFLOP/s measured).  Here
//   * one CTA step covers a SUPER-TILE of 256 output pixels (TW x 2*TH, two 128-row accumulators),
//   * per (source, 64-channel chunk, kw) ONE TMA box {64, TW, 2*TH + 2} - the "slab" - lands in
//     shared memory; the three kh taps of both halves are UMMA descriptors into the same slab at row
//     offsets that are multiples of 8 rows (1024 B), so the 128-byte swizzle phase is preserved,
//   * every weight tile (tap, chunk) is loaded once per super-tile and used by both halves.
// L2 -> SM traffic per 128 pixels per chunk: (3 * 36 KB + 9 * 16 KB) / 2 = 126 KB (2.3x less).
// Two rings (A slabs, B weight tiles) with mbarrier full/empty pairs; D = 2 halves x 2 buffers x
// n_mma TMEM columns (<= 512), so the epilogue of super-tile i overlaps the MMAs of i+1.
// Warp roles: warps 0-7 epilogue (warp & 3 = TMEM lane quarter, warp >> 2 = half of the super-tile),
// warp 8 TMA producer, warp 9 MMA issuer.
// Epilogue: with 4 epilogue warps the kernel was epilogue-bound (mainloop alone 1350 TFLOP/s at
// Cout = 128, 880 with the epilogue); 8 warps, one half each, bring it to ~1200.  An optional variant
// (COAL) sends bf16 outputs through a per-warp swizzled staging tile (32 pixels x 64 channels, 4 KB) so
// 8 lanes cover one 128-byte pixel row per st.global.v4, and brings the dgrad activation mask in the
// same way; it measured slightly slower than the direct row-per-thread stores and is off by default.
#include <stdlib.h>
#include "common.cuh"
#include "tc_common.cuh"
#include "conv_epi.cuh"

#define T3_THREADS 320
#define T3_EPI_WARPS 8
#define T3_WSTG_BYTES 4096
#define T3_MAX_A 5
#define T3_MAX_B 12

int of_tc_make_view_map(CUtensorMap* m, const void* base, int c_used, long long w_stride, long long h_stride,
                        long long n_stride, int N, int H, int W, int box_w, int box_h, int sw, int sh);
int of_tc_make_weight_map(CUtensorMap* m, const void* base, int taps, int rows, int cols, int box_rows);

struct Conv3Params {
  int num_src;
  int kchunks[2], kbase[2];
  int klast_slab, klast_n;                 // -DT3_KSKIP: slabs sl with sl % slabs_per_tile >= klast_slab issue klast_n K steps
  int pad_t, pad_l, sign;
  int NB, Ho, Wo;
  int TW, TH, tiles_w, tiles_h, m_tiles;   // super-tile = TW x 2*TH output pixels
  int n_mma;
  int na, nb;                              // ring depths
  int slab_bytes, btile_bytes;
  int tmem_cols;
  int tw_shift;                            // log2(TW)
  int coal;                                // bf16 output through the per-warp staging tile (coalesced stores)
  int dbg;                                 // profiling only: 1 = no epilogue memory traffic, 2 = no epilogue at all
  ConvEpi E;
};

// MINB = 2: two CTAs per SM (narrow layers, <= 256 TMEM columns and half the rings each): the second CTA's TMA
// stream and MMAs fill the first one's bubbles
// SPEC >= 0: the epilogue's options are COMPILE-TIME constants (bit field below) - the run-time-flagged epilogue spent
// 37 uniform branches per 16-column chunk and the eight epilogue warps were busy for the whole kernel (ncu, round 2:
// 83 % of the warp samples in epilogue code, tensor pipe 61 % active at 128->128); the host picks a specialisation when
// the launch matches one and falls back to SPEC = -1 otherwise.
//   bits 0-1 activation, bit 2 add (bf16), bit 3 sign_out, bits 4-5 mask kind (1 = sign bits, 2 = bf16 tensor),
//   bits 6-7 mask activation, bit 8 split output, bit 9 the add comes AFTER the activation and is followed by ReLU
//   (add_mode 1);  always: bf16 output, full 16-channel chunks, 32-byte aligned rows
#define T3_SPEC(act, add, sign, mkind, mact, split) ((act) | ((add) << 2) | ((sign) << 3) | ((mkind) << 4) | ((mact) << 6) | ((split) << 8))
#define T3_SPEC_POST (1 << 9)
// KS: K steps (16 channels each) issued per 64-channel chunk.  4 by default; 2 for a single source of <= 32 channels (the
// Cout = 2 flow heads' fprop and the dgrad of the 64->32 layers): the upper half of the chunk is TMA zero fill, and a
// compile-time count costs the issuing thread nothing (the run-time skip of round 2 made every layer 6-15 % slower).
// KS = 42: 4 per chunk and 2 on the last chunk of a single source with 96 (160, ...) channels - compiled only into the
// two specialisations those layers use (fprop 96->64, dgrad of 128->96).
template <bool COAL, int MINB = 1, int SPEC = -1, int KS = 4>
__global__ void __launch_bounds__(T3_THREADS, MINB)
conv_tc3_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                const __grid_constant__ CUtensorMap tmB, const __grid_constant__ Conv3Params P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* a_ring = base;
  uint8_t* b_ring = base + (size_t)P.na * P.slab_bytes;
  uint8_t* staging = b_ring + (size_t)P.nb * P.btile_bytes;           // one 4 KB tile per epilogue warp
  uint64_t* a_full = reinterpret_cast<uint64_t*>(staging + (COAL ? T3_EPI_WARPS * T3_WSTG_BYTES : 0));
  uint64_t* a_empty = a_full + T3_MAX_A;
  uint64_t* b_full = a_empty + T3_MAX_A;
  uint64_t* b_empty = b_full + T3_MAX_B;
  uint64_t* tmem_full = b_empty + T3_MAX_B;    // [2]
  uint64_t* tmem_empty = tmem_full + 2;        // [2]
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  __shared__ float s_bias[128];                     // n_mma <= 128: bias read once, not per chunk per pixel

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) OF_PDL_TRIGGER();
  if (threadIdx.x == 0) {
    for (int s = 0; s < P.na; ++s) { tc::mbar_init(&a_full[s], 1); tc::mbar_init(&a_empty[s], 1); }
    for (int s = 0; s < P.nb; ++s) { tc::mbar_init(&b_full[s], 1); tc::mbar_init(&b_empty[s], 1); }
    for (int b = 0; b < 2; ++b) { tc::mbar_init(&tmem_full[b], 1); tc::mbar_init(&tmem_empty[b], T3_EPI_WARPS); }
    tc::fence_barrier_init();
  }
  if (warp == 8 && lane == 0) {
    tc::prefetch_tmap(&tmA0);
    if (P.num_src > 1) tc::prefetch_tmap(&tmA1);
    tc::prefetch_tmap(&tmB);
  }
  if (warp == 9) {
    tc::tmem_alloc(tmem_base_slot, (uint32_t)P.tmem_cols);
    tc::tmem_relinquish();
  }
  OF_PDL_WAIT();                                     // first global access below: the bias
  if (threadIdx.x < 128) s_bias[threadIdx.x] = (P.E.bias && (int)threadIdx.x < P.E.n_valid) ? P.E.bias[threadIdx.x] : 0.f;
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *tmem_base_slot;
  const int SH = 2 * P.TH;                                        // super-tile height
  // slab start row relative to the super-tile's first output row, and the slab row of tap kh
  const int slab_dh = P.sign > 0 ? -P.pad_t : -(2 - P.pad_t);

  if (warp == 8) {
    // ===================== TMA producer =====================
    if (OF_ONE_LANE()) {
      int as = 0, bs = 0; uint32_t aph = 0, bph = 0;
      for (int tile = blockIdx.x; tile < P.m_tiles; tile += gridDim.x) {
        int tw_i = tile % P.tiles_w; int r = tile / P.tiles_w;
        int th_i = r % P.tiles_h; int n = r / P.tiles_h;
        int w0 = tw_i * P.TW, h0 = th_i * SH;
        for (int s = 0; s < P.num_src; ++s) {
          const CUtensorMap* tmA = s == 0 ? &tmA0 : &tmA1;
          for (int kc = 0; kc < P.kchunks[s]; ++kc) {
            for (int kw = 0; kw < 3; ++kw) {
              tc::mbar_wait(&a_empty[as], aph ^ 1);
              if ((OF_DBG(P) & 32) && tile != (int)blockIdx.x) {
                tc::mbar_arrive(&a_full[as]);         // profiling only: activation slabs "free"
              } else {
              tc::mbar_expect_tx(&a_full[as], (uint32_t)P.slab_bytes);
              tc::tma_load_4d(a_ring + (size_t)as * P.slab_bytes, tmA, &a_full[as], kc * 64,
                              w0 + P.sign * (kw - P.pad_l), h0 + slab_dh, n);
              }
              if (++as == P.na) { as = 0; aph ^= 1; }
              for (int kh = 0; kh < 3; ++kh) {
                tc::mbar_wait(&b_empty[bs], bph ^ 1);
                if ((OF_DBG(P) & 8) && tile != (int)blockIdx.x) {
                  tc::mbar_arrive(&b_full[bs]);       // profiling only: weight tiles "free" after the first output tile
                } else if ((OF_DBG(P) & 16) && tile != (int)blockIdx.x && kh > 0) {
                  tc::mbar_arrive(&b_full[bs]);       // profiling only: a third of the weight traffic
                } else {
                tc::mbar_expect_tx(&b_full[bs], (uint32_t)(P.n_mma * 128));
                tc::tma_load_3d(b_ring + (size_t)bs * P.btile_bytes, &tmB, &b_full[bs], P.kbase[s] + kc * 64, 0,
                                kh * 3 + kw);
                }
                if (++bs == P.nb) { bs = 0; bph ^= 1; }
              }
            }
          }
        }
      }
    }
  } else if (warp == 9) {
    // ===================== MMA issuer =====================
    // (skipping the all-zero K steps of a 32- / 96-channel source was tried: the extra predicate per MMA in THIS thread's
    //  instruction stream made every layer 6-15 % slower - the issuing thread is on the critical path)
    if (OF_ONE_LANE()) {
      const uint32_t idesc = tc::idesc_bf16_f32(128, P.n_mma, 0, 0);
      constexpr bool LH = (KS == 42);                    // 4 K steps per chunk, 2 on the slabs of the last chunk
      constexpr int KSN = LH ? 4 : KS;
      const int slabs_per_tile = 3 * (P.kchunks[0] + (P.num_src > 1 ? P.kchunks[1] : 0));
      int as = 0, bs = 0; uint32_t aph = 0, bph = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < P.m_tiles; tile += gridDim.x, ++it) {
        int buf = it & 1;
        uint32_t use = (uint32_t)(it >> 1);
        tc::mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);
        tc::fence_after_sync();
        const uint32_t d0 = tmem_base + (uint32_t)(buf * 2 * P.n_mma);
        for (int sl = 0; sl < slabs_per_tile; ++sl) {
          tc::mbar_wait(&a_full[as], aph);
          tc::fence_after_sync();
          const uint32_t slab = tc::smem_u32(a_ring + (size_t)as * P.slab_bytes);
          for (int kh = 0; kh < 3; ++kh) {
            tc::mbar_wait(&b_full[bs], bph);
            tc::fence_after_sync();
            const uint64_t bdesc = tc::desc_sw128_kmajor(tc::smem_u32(b_ring + (size_t)bs * P.btile_bytes));
            const int tap_row = P.sign > 0 ? kh : 2 - kh;
            if (OF_DBG(P) & 128) {
              // timing experiment (profile-hooks build, results are garbage): operands SWAPPED - the weight tile is the
              // A operand (M = Cout rows), the 256 pixels of the super-tile are ONE N = 256 B operand (the two halves are
              // contiguous in the slab): half the MMA instructions, each twice as wide
              const uint32_t idesc_sw = tc::idesc_bf16_f32(P.n_mma <= 64 ? 64 : 128, 256, 0, 0);
              const uint64_t pdesc = tc::desc_sw128_kmajor(slab + (uint32_t)(tap_row * P.TW * 128));
#pragma unroll
              for (int k = 0; k < 4; ++k)
                tc::mma_bf16(tmem_base + (uint32_t)(buf * 256), bdesc + (uint64_t)(2 * k), pdesc + (uint64_t)(2 * k), idesc_sw,
                             (sl | kh | k) ? 1u : 0u);
            } else if (LH && sl >= P.klast_slab) {
              // KS == 42: the LAST 64-channel chunk of a single source with 32 channels (mod 64) holds zeros in its upper
              // K steps (TMA zero fill): its slabs issue only the two real ones
#pragma unroll
              for (int hh = 0; hh < 2; ++hh) {
                const uint64_t adesc = tc::desc_sw128_kmajor(slab + (uint32_t)((tap_row + hh * P.TH) * P.TW * 128));
#pragma unroll
                for (int k = 0; k < 2; ++k)
                  tc::mma_bf16(d0 + (uint32_t)(hh * P.n_mma), adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                               (sl | kh | k) ? 1u : 0u);
              }
            } else {
#pragma unroll
              for (int hh = 0; hh < 2; ++hh) {
                const uint64_t adesc = tc::desc_sw128_kmajor(slab + (uint32_t)((tap_row + hh * P.TH) * P.TW * 128));
#pragma unroll
                for (int k = 0; k < KSN; ++k)
                  tc::mma_bf16(d0 + (uint32_t)(hh * P.n_mma), adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc,
                               (sl | kh | k) ? 1u : 0u);
              }
            }
            tc::mma_commit(&b_empty[bs]);
            if (++bs == P.nb) { bs = 0; bph ^= 1; }
          }
          tc::mma_commit(&a_empty[as]);
          if (sl == slabs_per_tile - 1) tc::mma_commit(&tmem_full[buf]);
          if (++as == P.na) { as = 0; aph ^= 1; }
        }
      }
    }
  } else {
    // ===================== epilogue warps 0..7 =====================
    const int q = warp & 3, hh = warp >> 2;             // TMEM lane quarter, half of the super-tile
    uint8_t* wstg = staging + (size_t)warp * T3_WSTG_BYTES;
    const int r_sub = lane >> 3, u = lane & 7;          // copy phase: 8 lanes x 16 B cover one 128-byte pixel row
    const int row = q * 32 + lane;                      // accumulator row == TMEM lane == pixel in the half
    const int nch = (P.E.n_valid + 15) >> 4;            // 16-column chunks that hold real channels
    const bool mask_coal = COAL && P.E.mask != nullptr;   // host: COAL implies a bf16, 16-byte-aligned mask
    int it = 0;
    for (int tile = blockIdx.x; tile < P.m_tiles; tile += gridDim.x, ++it) {
      int buf = it & 1;
      uint32_t use = (uint32_t)(it >> 1);
      int tw_i = tile % P.tiles_w; int r = tile / P.tiles_w;
      int th_i = r % P.tiles_h; int n = r / P.tiles_h;
      const int h_base = th_i * SH + hh * P.TH, w_base = tw_i * P.TW;
      const int ph = h_base + (row >> P.tw_shift), pw = w_base + (row & (P.TW - 1));
      const bool valid = ph < P.Ho && pw < P.Wo && !(OF_DBG(P) & 1);
      const size_t pix = ((size_t)n * P.Ho + ph) * P.Wo + pw;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((buf * 2 + hh) * P.n_mma);

      // copy-phase pixel of (i, lane): staging row i*4 + r_sub
      auto copy_pix = [&](int i, size_t& cpix) -> bool {
        int rr_ = q * 32 + i * 4 + r_sub;
        int ch_ = h_base + (rr_ >> P.tw_shift), cw_ = w_base + (rr_ & (P.TW - 1));
        cpix = ((size_t)n * P.Ho + ch_) * P.Wo + cw_;
        return ch_ < P.Ho && cw_ < P.Wo && !(OF_DBG(P) & 1);
      };
      uint4 mreg[8];
      auto load_mask = [&](int g) {                     // coalesced: 4 pixel rows x 128 B per instruction
        const int ch = g * 64 + u * 8;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          size_t cpix;
          bool ok = copy_pix(i, cpix) && ch + 8 <= P.E.n_valid;
          mreg[i] = ok ? __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const __nv_bfloat16*>(P.E.mask) +
                                                              cpix * P.E.mask_stride + ch))
                       : make_uint4(0, 0, 0, 0);
        }
      };
      if (mask_coal && !(OF_DBG(P) & 2)) load_mask(0);       // in flight while the accumulator is still being produced
      // sign-bit activation mask (16 B / pixel, written by the producing layer's fprop): one load per row
      const bool mask_bits = !COAL && P.E.mask != nullptr && P.E.mask_dtype == OF_MASK_BITS;
      uint64_t mb_lo = 0, mb_hi = 0, sb_lo = 0, sb_hi = 0;
      if (mask_bits && valid) {
        const uint4 mb = __ldg(reinterpret_cast<const uint4*>(P.E.mask) + pix);
        mb_lo = (uint64_t)mb.x | ((uint64_t)mb.y << 32);
        mb_hi = (uint64_t)mb.z | ((uint64_t)mb.w << 32);
      }

      tc::mbar_wait(&tmem_full[buf], use & 1);
      tc::fence_after_sync();
      if constexpr (SPEC >= 0) {
        constexpr int ACT = SPEC & 3, MACT = (SPEC >> 6) & 3;
        constexpr bool ADD = (SPEC >> 2) & 1, SIGN = (SPEC >> 3) & 1, MBITS = ((SPEC >> 4) & 3) == 1, SPLIT = (SPEC >> 8) & 1;
        constexpr bool MTEN = ((SPEC >> 4) & 3) == 2, POST = (SPEC >> 9) & 1;
        uint32_t rr[2][16];
        uint4 araw[2][2], mraw[2][2];
        const int nchunks = P.n_mma >> 4;
        const __nv_bfloat16* addp = reinterpret_cast<const __nv_bfloat16*>(P.E.add) + pix * P.E.add_stride;
        const __nv_bfloat16* maskp = reinterpret_cast<const __nv_bfloat16*>(P.E.mask) + pix * P.E.mask_stride;
        auto issue = [&](int c, int slot) {
          __syncwarp();                                  // tcgen05.ld is .sync.aligned: reconverge first
          tc::tmem_ld16(taddr + (uint32_t)(c * 16), rr[slot]);
          if (ADD && valid) ld_global_256(addp + c * 16, araw[slot][0], araw[slot][1]);
          if (MTEN && valid) ld_global_256(maskp + c * 16, mraw[slot][0], mraw[slot][1]);
        };
        auto finish = [&](int c, int slot) {
          __syncwarp();
          tc::tmem_ld_wait();
          if (c + 1 < nchunks) issue(c + 1, slot ^ 1);
          if (!valid) return;
          const int c0 = c * 16;
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rr[slot][j]) + s_bias[c0 + j];
          float ta[16];
          if (ADD) {
            const uint4 a4[4] = {araw[slot][0], araw[slot][1], make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
            unpack16(a4, OF_BF16, ta);
          }
          if (ADD && !POST) {
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] += ta[j];
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = apply_act(v[j], ACT);
          if (SIGN) {
            uint32_t m16 = 0;
#pragma unroll
            for (int j = 0; j < 16; ++j) m16 |= (v[j] > 0.f ? 1u : 0u) << j;
            if (c < 4) sb_lo |= (uint64_t)m16 << (16 * c); else sb_hi |= (uint64_t)m16 << (16 * (c - 4));
          }
          if (ADD && POST) {                             // resnet/models.py:30-35: relu(relu(bn2(conv2)) + shortcut)
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j] + ta[j], 0.f);
          }
          if (MTEN) {
            float t[16];
            const uint4 m4[4] = {mraw[slot][0], mraw[slot][1], make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
            unpack16(m4, OF_BF16, t);
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] *= act_slope(t[j], MACT);
          }
          if (MBITS) {
            const uint32_t m16 = (uint32_t)((c < 4 ? mb_lo >> (16 * c) : mb_hi >> (16 * (c - 4))) & 0xffffu);
            constexpr float s0 = MACT == OF_ACT_RELU ? 0.f : (MACT == OF_ACT_LEAKY ? 0.3f : 1.f);
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] *= ((m16 >> j) & 1u) ? 1.f : s0;
          }
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            __nv_bfloat162 t2 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
            pk[j] = *reinterpret_cast<uint32_t*>(&t2);
          }
          __nv_bfloat16* o = (SPLIT && c0 >= P.E.split)
                                 ? reinterpret_cast<__nv_bfloat16*>(P.E.out2) + pix * P.E.out2_stride + (c0 - P.E.split)
                                 : reinterpret_cast<__nv_bfloat16*>(P.E.out) + pix * P.E.out_stride + c0;
          st_global_256(o, pk);
        };
        issue(0, 0);
#pragma unroll 1
        for (int c = 0; c < nchunks; c += 2) {
          finish(c, 0);
          if (c + 1 < nchunks) finish(c + 1, 1);
        }
        if (SIGN && valid)
          reinterpret_cast<uint4*>(P.E.sign_out)[pix] = make_uint4((uint32_t)sb_lo, (uint32_t)(sb_lo >> 32), (uint32_t)sb_hi,
                                                                    (uint32_t)(sb_hi >> 32));
      } else
      if (!(OF_DBG(P) & 2)) {
        // software pipeline over 16-column chunks: the TMEM load of chunk c+1 and the global loads of
        // its add (and row-wise mask) operands are in flight while chunk c is finished
        uint32_t rr[2][16];
        // raw add / mask operands of one chunk (bf16: 2 x uint4, f32: 4); COAL: add is bf16, mask comes via staging
        uint4 araw[2][COAL ? 2 : 4], mraw[2][COAL ? 1 : 4];
        const int nchunks = COAL ? nch : (P.n_mma >> 4);
        auto issue = [&](int c, int slot) {
          __syncwarp();                                  // tcgen05.ld is .sync.aligned: reconverge first
          tc::tmem_ld16(taddr + (uint32_t)(c * 16), rr[slot]);
          if (valid && c * 16 + 16 <= P.E.n_valid) {
            if (P.E.add) {
              const int nq = (COAL || P.E.add_dtype == OF_BF16) ? 2 : 4;
              const uint4* qd = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(P.E.add) +
                                                               (pix * P.E.add_stride + c * 16) * (nq == 2 ? 2 : 4));
              if (nq == 2 && (reinterpret_cast<uintptr_t>(qd) & 31) == 0) {
                ld_global_256(qd, araw[slot][0], araw[slot][1]);      // one full sector per request
              } else {
#pragma unroll
                for (int k = 0; k < (COAL ? 2 : 4); ++k) if (k < nq) araw[slot][k] = __ldg(qd + k);
              }
            }
            if (!COAL && P.E.mask && !mask_bits) {
              const int nq = P.E.mask_dtype == OF_BF16 ? 2 : 4;
              const uint4* qd = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(P.E.mask) +
                                                               (pix * P.E.mask_stride + c * 16) * (P.E.mask_dtype == OF_BF16 ? 2 : 4));
              if (nq == 2 && (reinterpret_cast<uintptr_t>(qd) & 31) == 0) {
                ld_global_256(qd, mraw[slot][0], mraw[slot][1]);
              } else {
#pragma unroll
                for (int k = 0; k < 4; ++k) if (k < nq) mraw[slot][k] = __ldg(qd + k);
              }
            }
          }
        };
        auto finish = [&](int c, int slot) {             // slot is a compile-time constant at both call sites
          __syncwarp();
          tc::tmem_ld_wait();                            // chunk c has landed (only one tcgen05.ld is ever pending here)
          if (c + 1 < nchunks) issue(c + 1, slot ^ 1);
          const int c0 = c * 16;
          if constexpr (COAL) {
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rr[slot][j]) + s_bias[c0 + j];
            if (P.E.add && valid) {
              float t[16];
              uint4 a4[4] = {araw[slot][0], araw[slot][1], make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
              unpack16(a4, OF_BF16, t);
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] += t[j];
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = apply_act(v[j], P.E.act);
            // own row of the swizzled staging tile: 16-byte units u0, u0 + 1 of the 128-byte row
            uint8_t* rowp = wstg + lane * 128;
            const int u0 = (c0 & 63) >> 3;
            uint4* p0 = reinterpret_cast<uint4*>(rowp + ((u0 ^ (lane & 7)) << 4));
            uint4* p1 = reinterpret_cast<uint4*>(rowp + (((u0 + 1) ^ (lane & 7)) << 4));
            if (P.E.mask) {
              float t[16];
              uint4 m[4] = {*p0, *p1, make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
              unpack16(m, OF_BF16, t);
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] *= act_slope(t[j], P.E.mask_act);
            }
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              __nv_bfloat162 t2 = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
              pk[j] = *reinterpret_cast<uint32_t*>(&t2);
            }
            *p0 = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            *p1 = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            return;
          } else {
          if (!(valid && c0 < P.E.n_valid)) return;
          if (c0 + 16 <= P.E.n_valid && (P.E.out_stride & 7) == 0 && (P.E.add == nullptr || (P.E.add_stride & 7) == 0) &&
              (P.E.mask == nullptr || (P.E.mask_stride & 7) == 0)) {
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(rr[slot][j]) + s_bias[c0 + j];
            if (P.E.add && !P.E.add_post) {
              float t[16];
              unpack16(araw[slot], P.E.add_dtype, t);
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] += t[j];
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = apply_act(v[j], P.E.act);
            if (P.E.sign_out) {
              uint32_t m16 = 0;
#pragma unroll
              for (int j = 0; j < 16; ++j) m16 |= (v[j] > 0.f ? 1u : 0u) << j;
              if (c < 4) sb_lo |= (uint64_t)m16 << (16 * c); else sb_hi |= (uint64_t)m16 << (16 * (c - 4));
            }
            if (P.E.add && P.E.add_post) {               // resnet/models.py:30-35: relu(relu(bn2(conv2)) + shortcut)
              float t[16];
              unpack16(araw[slot], P.E.add_dtype, t);
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j] + t[j], 0.f);
            }
            if (mask_bits) {
              const uint32_t m16 = (uint32_t)((c < 4 ? mb_lo >> (16 * c) : mb_hi >> (16 * (c - 4))) & 0xffffu);
              const float s0 = P.E.mask_act == OF_ACT_RELU ? 0.f : (P.E.mask_act == OF_ACT_LEAKY ? 0.3f : 1.f);
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] *= ((m16 >> j) & 1u) ? 1.f : s0;
            } else if (P.E.mask) {
              float t[16];
              unpack16(mraw[slot], P.E.mask_dtype, t);
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] *= act_slope(t[j], P.E.mask_act);
            }
            // profiling only (dbg 64): all the epilogue math, no global store
            if (!(OF_DBG(P) & 64) || v[0] + v[7] + v[15] == 1.2345e-30f) store16(P.E, pix, c0, v);
          } else {
            epilogue_store16(P.E, pix, c0, rr[slot]);    // ragged channel tail / unaligned strides: generic path
          }
          }
        };
        auto begin_group = [&](int g) {                  // mask of group g: registers -> staging, prefetch group g+1
          if (!mask_coal) return;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            int rs = i * 4 + r_sub;
            *reinterpret_cast<uint4*>(wstg + rs * 128 + ((u ^ (rs & 7)) << 4)) = mreg[i];
          }
          __syncwarp();
          if ((g + 1) * 64 < P.E.n_valid) load_mask(g + 1);
        };
        auto flush_group = [&](int g) {                  // staging -> global, 4 full pixel rows per instruction
          __syncwarp();
          const int ch = g * 64 + u * 8;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            int rs = i * 4 + r_sub;
            size_t cpix;
            bool ok = copy_pix(i, cpix) && ch + 8 <= P.E.n_valid;
            uint4 d = *reinterpret_cast<const uint4*>(wstg + rs * 128 + ((u ^ (rs & 7)) << 4));
            if (ok) *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(P.E.out) + cpix * P.E.out_stride + ch) = d;
          }
          __syncwarp();
        };
        issue(0, 0);
#pragma unroll 1
        for (int c = 0; c < nchunks; c += 2) {
          if (COAL && (c & 3) == 0) begin_group(c >> 2);
          finish(c, 0);
          if (c + 1 < nchunks) finish(c + 1, 1);
          const int last = c + 1 < nchunks ? c + 1 : c;
          if (COAL && ((last & 3) == 3 || last == nchunks - 1)) flush_group(last >> 2);
        }
        if (!COAL && P.E.sign_out && valid)
          reinterpret_cast<uint4*>(P.E.sign_out)[pix] = make_uint4((uint32_t)sb_lo, (uint32_t)(sb_lo >> 32), (uint32_t)sb_hi,
                                                                    (uint32_t)(sb_hi >> 32));
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&tmem_empty[buf]);
    }
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 9) tc::tmem_dealloc(tmem_base, (uint32_t)P.tmem_cols);
}

// Returns 1 if handled, 0 if the shape is not this kernel's (caller falls back to conv_tc_kernel), < 0 / > 0 on error.
int of_conv_tc3_try(const of_tc_conv* c, void* stream) {
  const bool view = c->view_w_stride > 0;
  if (view || c->KH != 3 || c->KW != 3 || c->stride != 1 || (c->stride_w > 1) || (c->stride_h > 1)) return 0;
  if (c->n_mma > 128) return 0;                         // 2 halves x 2 buffers x n_mma TMEM columns <= 512
  if (of_tune_env("OF_CONV_NO_SLAB")) return 0;
  Conv3Params P;
  memset(&P, 0, sizeof(P));
  P.num_src = c->num_src;
  int kb = 0;
  for (int s = 0; s < c->num_src; ++s) {
    P.kchunks[s] = (c->src_c[s] + 63) / 64;
    P.kbase[s] = kb;
    kb += P.kchunks[s] * 64;
  }
  OF_REQUIRE(kb <= c->k_pad, OF_ERR_SHAPE, "packed K %d smaller than the sources' %d", c->k_pad, kb);
  // (T3_KSKIP) single source whose last chunk holds exactly 32 real channels: its 3 slabs (kw) issue 2 K steps
  P.klast_slab = 1 << 30; P.klast_n = 4;
  if (c->num_src == 1 && c->src_c[0] - (P.kchunks[0] - 1) * 64 <= 32) { P.klast_slab = 3 * (P.kchunks[0] - 1); P.klast_n = 2; }
  P.pad_t = c->pad_t; P.pad_l = c->pad_l; P.sign = c->sign >= 0 ? 1 : -1;
  P.NB = c->N; P.Ho = c->Ho; P.Wo = c->Wo;
  // narrow tiles keep the halo overhead (2*TH + 2) / (2*TH) small; TW must be a multiple of 8 rows
  P.TW = (c->Wo % 16 == 0 || c->Wo % 8 != 0) ? 16 : 8;
  P.TH = 128 / P.TW;
  P.tiles_w = (c->Wo + P.TW - 1) / P.TW;
  P.tiles_h = (c->Ho + 2 * P.TH - 1) / (2 * P.TH);
  long long mt = (long long)c->N * P.tiles_w * P.tiles_h;
  OF_REQUIRE(mt > 0 && mt < (1ll << 30), OF_ERR_SHAPE, "bad output tiling");
  P.m_tiles = (int)mt;
  P.n_mma = c->n_mma;
  P.slab_bytes = (2 * P.TH + 2) * P.TW * 128;
  P.btile_bytes = ((c->n_mma * 128) + 1023) / 1024 * 1024;
  P.tw_shift = P.TW == 16 ? 4 : 3;
  // coalesced bf16 output through the per-warp staging tile: opt-in (OF_T3_COAL=1).  Measured on B200
  // at N8 192x640: 128->128 242 us staged vs 239 us direct, 64->32 111 vs 97 - with 8 epilogue warps the
  // row-per-thread stores are no longer the limit and the staging round trip costs more than it saves.
  const bool mask_coal_ok = c->mask == nullptr || (c->mask_dtype == OF_BF16 && (c->mask_stride % 8) == 0 &&
                                                   ((uintptr_t)c->mask & 15) == 0);
  const bool add_coal_ok = c->add == nullptr || (c->add_dtype == OF_BF16 && (c->add_stride % 8) == 0 &&
                                                 ((uintptr_t)c->add & 15) == 0 && (c->n_valid % 16) == 0);
  P.coal = c->out_dtype == OF_BF16 && (c->out_stride % 8) == 0 && ((uintptr_t)c->out & 15) == 0 && c->n_valid >= 8 &&
           (c->n_valid % 8) == 0 && mask_coal_ok && add_coal_ok;
  {
    const char* e = of_tune_env("OF_T3_COAL");                 // 1 = always, 2 = only with a mask operand (dgrad)
    const int mode = e ? atoi(e) : 0;
    P.coal = P.coal && (mode == 1 || (mode == 2 && c->mask != nullptr));
  }
  if (const char* e = of_tune_env("OF_T3_DBG")) P.dbg = atoi(e);
  const bool bits = c->sign_out != nullptr || (c->mask != nullptr && c->mask_dtype == OF_MASK_BITS);
  if (bits) {
    OF_REQUIRE((c->n_valid % 16) == 0 && c->n_valid <= 128 && (c->out_stride % 8) == 0 &&
                   (((uintptr_t)c->sign_out | (uintptr_t)(c->mask_dtype == OF_MASK_BITS ? c->mask : nullptr)) & 15) == 0 &&
                   (c->add == nullptr || (c->add_stride % 8) == 0),
               OF_ERR_UNSUPPORTED, "sign-bit masks need Cout %% 16 == 0, Cout <= 128 and 16-byte aligned rows");
    P.coal = 0;
  }
  // ring depths: TMA latency (~2k cycles loaded) must be covered by the MMAs of the slots in flight.
  // One slab feeds 24 MMAs (~1.6k cycles at Cout = 128), one weight tile 8 (~0.5k): 3 slabs + >= 4 tiles.
  const int stg_bytes = P.coal ? T3_EPI_WARPS * T3_WSTG_BYTES : 0;
  const int budget = 206 * 1024 - stg_bytes;
  P.na = 3;
  P.nb = (budget - P.na * P.slab_bytes) / P.btile_bytes;
  if (P.nb < 4) { P.na = 2; P.nb = (budget - P.na * P.slab_bytes) / P.btile_bytes; }
  if (P.nb > T3_MAX_B) P.nb = T3_MAX_B;
  if (const char* e = of_tune_env("OF_T3_RINGS")) {            // tuning hook: "na,nb"
    int a_ = 0, b_ = 0;
    if (sscanf(e, "%d,%d", &a_, &b_) == 2 && a_ >= 1 && a_ <= T3_MAX_A && b_ >= 3 && b_ <= T3_MAX_B &&
        a_ * P.slab_bytes + b_ * P.btile_bytes <= budget) { P.na = a_; P.nb = b_; }
  }
  // Two CTAs per SM for the narrow layers (measured at N8 192x640: 64->32 96 -> 64 us, 96->64 183 -> 153 us; 64->64
  // and the Cout >= 96 layers do not gain): 4 * n_mma <= 256 TMEM columns each, rings cut to ~100 KB, 96 registers.
  int two = 0;
  {
    const int kch = kb / 64;
    int b_ = (100 * 1024 - stg_bytes - 2 * P.slab_bytes) / P.btile_bytes;
    if (b_ > T3_MAX_B) b_ = T3_MAX_B;
    const char* e = of_tune_env("OF_T3_2CTA");                  // 0 = never, 1 = every layer with n_mma <= 64
    // (round 2, after the elect.sync issuer: the multi-chunk n_mma <= 64 layers are faster with ONE CTA - 96->64 at N8
    //  192x640 139 vs 145 us, at 96x320 40.4 vs 45.8 us - so only n_mma <= 32 takes two; ring depths 2..5 x 3..12 move
    //  no layer by more than 3 %, scripts/gpu_r2_run19.sh)
    const bool want = e ? (atoi(e) > 0 && c->n_mma <= 64) : (c->n_mma <= 32);
    (void)kch;
    if (want && !P.coal && b_ >= 3 && P.m_tiles > OF_NUM_SMS && !(OF_DBG(P) & 128)) { two = 1; P.na = 2; P.nb = b_; }
  }
  if (P.nb < 3) return 0;
  int cols = 4 * c->n_mma;
  P.tmem_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : cols <= 256 ? 256 : 512;
  if (OF_DBG(P) & 128) P.tmem_cols = 512;              // swapped-operand timing hook: two 256-column accumulators
  P.E.out = c->out; P.E.out_dtype = c->out_dtype; P.E.out_stride = c->out_stride;
  P.E.bias = c->bias; P.E.act = c->act;
  P.E.add = c->add; P.E.add_dtype = c->add_dtype; P.E.add_stride = c->add_stride;
  P.E.mask = c->mask; P.E.mask_dtype = c->mask_dtype; P.E.mask_stride = c->mask_stride; P.E.mask_act = c->mask_act;
  P.E.n_valid = c->n_valid;
  P.E.sign_out = c->sign_out;
  P.E.add_post = c->add_mode == 1 ? 1 : 0;
  if (P.E.add_post) {
    OF_REQUIRE(c->add != nullptr && c->out_dtype == OF_BF16 && (c->n_valid % 16) == 0 && (c->out_stride % 8) == 0 &&
                   (c->add_stride % 8) == 0 && c->mask == nullptr && c->out2 == nullptr,
               OF_ERR_UNSUPPORTED, "add_mode 1 needs bf16 output, Cout %% 16 == 0, aligned rows, no mask / split output");
    P.coal = 0;
  }
  if (c->out2 != nullptr) {
    OF_REQUIRE(c->out_split > 0 && (c->out_split % 16) == 0 && c->out_split < c->n_valid && (c->n_valid % 16) == 0 &&
                   c->out_dtype == OF_BF16 && (c->out_stride % 8) == 0 && (c->out2_stride % 8) == 0 &&
                   (((uintptr_t)c->out | (uintptr_t)c->out2) & 15) == 0 && c->add == nullptr && c->mask == nullptr,
               OF_ERR_UNSUPPORTED, "split output needs bf16, 16-aligned column split and no add / mask operand");
    P.E.out2 = c->out2; P.E.out2_stride = c->out2_stride; P.E.split = c->out_split;
    P.coal = 0;
  }

  CUtensorMap tmA[2], tmB;
  memset(tmA, 0, sizeof(tmA));
  for (int s = 0; s < c->num_src; ++s)
    if (int r = of_tc_make_view_map(&tmA[s], c->src[s], c->src_c[s], c->src_stride[s], (long long)c->W * c->src_stride[s],
                                    (long long)c->H * c->W * c->src_stride[s], c->N, c->H, c->W, P.TW, 2 * P.TH + 2, 1, 1))
      return r;
  if (c->num_src == 1) tmA[1] = tmA[0];
  if (int r = of_tc_make_weight_map(&tmB, c->wpacked, 9, c->n_rows, c->k_pad, c->n_mma)) return r;

  size_t smem = 1024 + (size_t)P.na * P.slab_bytes + (size_t)P.nb * P.btile_bytes + stg_bytes +
                (2 * T3_MAX_A + 2 * T3_MAX_B + 4) * sizeof(uint64_t) + 16;
  // compile-time epilogue specialisations the train step uses (anything else runs the run-time-flagged epilogue)
  constexpr int S_FL = T3_SPEC(OF_ACT_LEAKY, 0, 1, 0, 0, 0), S_FR = T3_SPEC(OF_ACT_RELU, 0, 1, 0, 0, 0);   // fprop + sign bits
  constexpr int S_DML = T3_SPEC(0, 0, 0, 1, OF_ACT_LEAKY, 0), S_DMR = T3_SPEC(0, 0, 0, 1, OF_ACT_RELU, 0);  // dgrad * act'(bits)
  constexpr int S_DA = T3_SPEC(0, 1, 0, 0, 0, 0), S_D0 = T3_SPEC(0, 0, 0, 0, 0, 0), S_DS = T3_SPEC(0, 0, 0, 0, 0, 1);
  constexpr int S_FL0 = T3_SPEC(OF_ACT_LEAKY, 0, 0, 0, 0, 0), S_FR0 = T3_SPEC(OF_ACT_RELU, 0, 0, 0, 0, 0);   // fprop, no sign bits
  constexpr int S_FRP = T3_SPEC(OF_ACT_RELU, 1, 1, 0, 0, 0) | T3_SPEC_POST;      // conv2 + shortcut + ReLU (fuse_residual_epilogues)
  constexpr int S_DAMR = T3_SPEC(0, 1, 0, 2, OF_ACT_RELU, 0);                   // dgrad + skip, * relu'(previous block's output)
  int spec = -1;
  if (!P.coal && P.dbg == 0 && c->out_dtype == OF_BF16 && c->n_valid == c->n_mma && (c->out_stride % 16) == 0 &&
      ((uintptr_t)c->out & 31) == 0 &&
      (c->add == nullptr || (c->add_dtype == OF_BF16 && (c->add_stride % 16) == 0 && ((uintptr_t)c->add & 31) == 0)) &&
      (c->mask == nullptr || c->mask_dtype == OF_MASK_BITS ||
       (c->mask_dtype == OF_BF16 && (c->mask_stride % 16) == 0 && ((uintptr_t)c->mask & 31) == 0)) &&
      (c->out2 == nullptr || ((c->out2_stride % 16) == 0 && ((uintptr_t)c->out2 & 31) == 0 && (c->out_split % 16) == 0)) &&
      !of_tune_env("OF_T3_NO_SPEC"))
    spec = T3_SPEC(c->act, c->add != nullptr ? 1 : 0, c->sign_out != nullptr ? 1 : 0,
                   c->mask == nullptr ? 0 : (c->mask_dtype == OF_MASK_BITS ? 1 : 2), c->mask != nullptr ? c->mask_act : 0,
                   c->out2 != nullptr ? 1 : 0) | (P.E.add_post ? T3_SPEC_POST : 0);
  static bool attr_set = false;
  if (!attr_set) {
    OF_CUDA(cudaFuncSetAttribute(conv_tc3_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    OF_CUDA(cudaFuncSetAttribute(conv_tc3_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    OF_CUDA(cudaFuncSetAttribute(conv_tc3_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    OF_CUDA(cudaFuncSetAttribute(conv_tc3_kernel<false, 2, S_FL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    OF_CUDA(cudaFuncSetAttribute(conv_tc3_kernel<false, 2, -1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
    OF_CUDA(cudaFuncSetAttribute(conv_tc3_kernel<false, 1, -1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    OF_CUDA(cudaFuncSetAttribute(conv_tc3_kernel<false, 1, S_DML, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    OF_CUDA(cudaFuncSetAttribute(conv_tc3_kernel<false, 1, S_FL, 42>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    OF_CUDA(cudaFuncSetAttribute(conv_tc3_kernel<false, 1, S_DML, 42>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
#define T3_ATTR(S) OF_CUDA(cudaFuncSetAttribute(conv_tc3_kernel<false, 1, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024))
    T3_ATTR(S_FL); T3_ATTR(S_FR); T3_ATTR(S_DML); T3_ATTR(S_DMR); T3_ATTR(S_DA); T3_ATTR(S_D0); T3_ATTR(S_DS);
    T3_ATTR(S_FRP); T3_ATTR(S_DAMR); T3_ATTR(S_FL0); T3_ATTR(S_FR0);
#undef T3_ATTR
    attr_set = true;
  }
  int grid = P.m_tiles < OF_NUM_SMS ? P.m_tiles : OF_NUM_SMS;
#define T3_GO(MINB_, S) OF_CUDA(of_launch_pdl(conv_tc3_kernel<false, MINB_, S>, dim3(grid), dim3(T3_THREADS), smem, (cudaStream_t)stream, tmA[0], tmA[1], tmB, P))
  if (two) grid = P.m_tiles < 2 * OF_NUM_SMS ? P.m_tiles : 2 * OF_NUM_SMS;
  // half-K variants (KS = 2): one source of <= 32 channels
  const bool half_k = c->num_src == 1 && c->src_c[0] <= 32 && !P.coal && !of_tune_env("OF_T3_NO_HALFK");
  const bool last_half = P.klast_n == 2 && P.kchunks[0] >= 2 && !P.coal && !of_tune_env("OF_T3_NO_HALFK");
#define T3_GOK(MINB_, S) OF_CUDA(of_launch_pdl(conv_tc3_kernel<false, MINB_, S, 2>, dim3(grid), dim3(T3_THREADS), smem, (cudaStream_t)stream, tmA[0], tmA[1], tmB, P))
  if (half_k && two && spec != S_FL) T3_GOK(2, -1);
  else if (half_k && !two && spec == S_DML) T3_GOK(1, S_DML);
  else if (half_k && !two && spec == -1) T3_GOK(1, -1);
  else if (last_half && !two && spec == S_FL) OF_CUDA(of_launch_pdl(conv_tc3_kernel<false, 1, S_FL, 42>, dim3(grid), dim3(T3_THREADS), smem, (cudaStream_t)stream, tmA[0], tmA[1], tmB, P));
  else if (last_half && !two && spec == S_DML) OF_CUDA(of_launch_pdl(conv_tc3_kernel<false, 1, S_DML, 42>, dim3(grid), dim3(T3_THREADS), smem, (cudaStream_t)stream, tmA[0], tmA[1], tmB, P));
  else
  if (two) {
    if (spec == S_FL) T3_GO(2, S_FL);
    else OF_CUDA(of_launch_pdl(conv_tc3_kernel<false, 2>, dim3(grid), dim3(T3_THREADS), smem, (cudaStream_t)stream, tmA[0], tmA[1], tmB, P));
  } else
  if (P.coal) OF_CUDA(of_launch_pdl(conv_tc3_kernel<true>, dim3(grid), dim3(T3_THREADS), smem, (cudaStream_t)stream, tmA[0], tmA[1], tmB, P));
  else if (spec == S_FL) T3_GO(1, S_FL);
  else if (spec == S_FR) T3_GO(1, S_FR);
  else if (spec == S_DML) T3_GO(1, S_DML);
  else if (spec == S_DMR) T3_GO(1, S_DMR);
  else if (spec == S_DA) T3_GO(1, S_DA);
  else if (spec == S_D0) T3_GO(1, S_D0);
  else if (spec == S_DS) T3_GO(1, S_DS);
  else if (spec == S_FL0) T3_GO(1, S_FL0);
  else if (spec == S_FR0) T3_GO(1, S_FR0);
  else if (spec == S_FRP) T3_GO(1, S_FRP);
  else if (spec == S_DAMR) T3_GO(1, S_DAMR);
  else OF_CUDA(of_launch_pdl(conv_tc3_kernel<false>, dim3(grid), dim3(T3_THREADS), smem, (cudaStream_t)stream, tmA[0], tmA[1], tmB, P));
#undef T3_GO
#undef T3_GOK
  if (int r = of_launch_check(c->n_valid < 16 ? "conv_tc3_kernel<head>" : "conv_tc3_kernel")) return r;   // Cout = 2 flow heads apart
  return 1;
}
