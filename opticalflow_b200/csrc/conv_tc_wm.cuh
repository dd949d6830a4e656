// Third-generation tcgen05 3x3 convolution: halo tile as in conv_tc_halo.cuh, but with the operand roles swapped --
// the WEIGHTS are the M operand (M = 128 output channels, zero/garbage rows above CoutP are never read back) and the
// PIXELS are the N operand (N = 8*rows <= 256 pixels of a rows x 8 tile).  Measured on B200: one tcgen05.mma costs
// >= ~100-130 cycles to issue/execute whatever its shape, so the tensor pipe only fills when every instruction is as
// large as the ISA allows (128 x 256 x 16); with pixels on M a Cout = 128 layer tops out near 40% of peak.
//   D^T[cout, pixel] += W[tap][chunk][cout, 0:64] * X[pixel shifted by tap, 0:64]^T
//   * X: one TMA box per 64-channel chunk = the (rows+2) x 10 pixel halo (dilation d: sub-lattice via elementStrides);
//     tap (ky,kx) = N-side descriptor start shifted by (ky*10 + kx) 128-byte rows, SBO = 10 rows.
//   * W: one TMA box [CoutP x 64] per (tap, chunk) through a ring (or resident when the whole layer fits).
//   * accumulator: TMEM lanes = output channels, columns = pixels; double buffered (2 x N columns).
//   * epilogue: thread = one output channel; tcgen05.ld brings 16 pixels of it; a warp stores 32 consecutive channels
//     of one pixel per instruction (64 contiguous bytes in the NHWC concat slice).
#pragma once
#include "conv_tc_halo.cuh"

namespace pwc {

struct WmParams {
  int N, H, W, dil;
  int xsplit;                                     // dilation > 8 (TMA element strides stop at 8): W viewed as (W/dil, dil), 5-D maps
  int dy, kyr;                                    // y step of the pixel lattice and tile rows per ky tap: (dil, 1) = d x d
                                                  // sub-lattices; (1, dil) = x phases only, dilation in y by in-tile row shift
  int cout, coutp, kc, nchunk, rows, npx;       // npx = 8*rows = MMA N
  int ksplit, cps;                                // split-K: ksplit CTAs share one pixel tile, cps chunks each; partial sums
  float* scratch; int* counters;                  // meet in an fp32 scratch (red.add), the last CTA to arrive finishes the tile
  int ks2pass;                                    // deterministic form: every CTA stores its partial tile to its own scratch
                                                  // slab [ks][item][pixel][128 ch]; wm_splitk_reduce_kernel sums the slabs in
                                                  // order and does bias / LeakyReLU / 16-bit (hi | lo) stores
  int x_stages;                                   // halo-tile pipeline depth (2 for big tiles, up to 4 for small ones)
  int cin_span, tps, rep, cg;                   // tps: taps per weight stage (1|3|9); rep: weight-tile replicas on M
                                                // (lane groups rep*cg <= 4 hold the same channels -> epilogue splits pixels)
  int tiles_x, tiles_y, tiles_per_img, num_items;
  uint32_t x_bytes_tx, x_bytes, w_bytes_tx, w_bytes, w_stages, w_resident, w_pad, o_bytes, tmem_cols;
  int tma_store;                                  // 16-bit output through a shared-memory staging tile + TMA store
  int o_nbuf; uint32_t o_half;                    // staging buffers (1 | 2) of o_half bytes each
  int wstore, o_rs;                               // per-warp stores: every epilogue warp owns 2 buffers of o_rs rows x 8 px x 32 ch
  int fp16, dbg;
  // split precision (conv_mode 'mixed'): an fp32-grade tensor is stored as two 16-bit planes of the same pixel row, hi = RN16(v)
  // at channel c and lo = RN16(v - hi) at channel c + lo_off.  split != 0: nchunk counts VIRTUAL chunks, three per logical
  // 64-channel chunk -- (x_hi, W_hi), (x_hi, W_lo), (x_lo, W_hi) -- whose products meet in the same fp32 accumulator (the
  // dropped x_lo * W_lo term is 2^-22 relative).  split == 2: two virtual chunks, (x_hi, W_hi), (x_lo, W_hi): split activations,
  // single-precision weights.  out_lo_off != 0: the epilogue stores both planes of its output.
  int split, lo_off, out_lo_off;
  int up_lo_off, up_dup_off;                      // fused upfeat outputs (out_f32 == 2): lo plane / second hi copy, 0 = none
  // tap-lane stacking (kMode 4; Cout <= 64): the 128 MMA rows hold tl LANES of output channels, lane l = the weights of kernel
  // row ky = l (tl = 3, Cout <= 32) or of ky = 0 | 1 (tl = 2, Cout <= 64, ky = 2 follows as single-lane slots); one MMA per
  // kx covers all lanes at once on N = 8 * (rows + 2) halo pixels (3 / 6 MMA groups per chunk instead of 9); lane l's
  // accumulator row r + l holds its part of output row r, and the epilogue warps of a channel group add the lanes up through
  // the fp32 tile S (each warp takes another pixel block per phase, the last phase finishes the block).
  int tl, tl_bc;                                  // tl_bc: channels per output store box (16 for Cout <= 16, else 32)
  // head rows (tl == 3 only): the fourth lane quarter of the 128 MMA rows, idle otherwise, holds nhead <= 10 extra output channels
  // of ANOTHER layer that reads the same input -- row 96 + ky * nhead + h = that layer's weights of kernel row ky -- and the fourth
  // epilogue warp adds the three ky rows up (row offset undone as for the lanes) and stores them RAW (no bias, no activation) as
  // fp32 [pixel][nhead] to head_out: predict_flow / upfeat over the 533 concat channels conv{2,3}_4 reads anyway (their own pass
  // over the concat buffer is bound by its halo-tile traffic: 71 us for 2 output channels).  The consumer (a tap-stacked head over
  // the remaining 32 channels) adds `partial` before its activation.
  int nhead; float* head_out;
  const float* partial; int partial_pitch;
  // stride 2 (s2): H, W are the OUTPUT dims; the input is read through a pair view (C, col parity, W_in/2, H_in, N): one stage
  // = two dense sub-tiles of (2*rows + 1) input rows x 9 column pairs, x_sub bytes apart -- [0] the odd columns, [1] the even
  // ones -- so tap (ky, kx) is again a shifted start address: row ky (SBO = 2 tile rows = 18 pixel rows), pair (kx ? 1 : 0), sub-tile
  // (kx == 1).  Every input pixel enters shared memory once, as whole 2*kc-byte rows (the per-tap kernel gathers it 2.25 times
  // through element strides, at ~12 B/cycle/SM).
  int s2; uint32_t x_sub;
  // xdup: a virtual chunk that multiplies the same x tile as the one before it (split == 3: (x_hi, W_lo) after (x_hi, W_hi);
  // fused: [W_lo | 0] after [W_hi | W_hi] on the [hi | lo] tile) reuses the tile in shared memory instead of loading it again
  int xdup, fused;
  unsigned long long skip[9];                    // bit ch of skip[tap]: weight block (tap, chunk ch) is all zero -> no load, no MMA
  int has_kk, kk_mask[9], kk_first_tap, kk_first_kk;   // single-chunk layers: bit kk of kk_mask[tap] set <=> the K=16 step kk of
                                                  // tap's weight tile is all zero (space-to-depth block structure) -> no MMA
  int has_skip, out_groups;                       // out_groups g > 1: CoutP = g groups written group_pitch apart (5-D TMA store)
  int ts, ntaps, s_stride;                        // tap-stacked mode (Cout*9 <= 128): M rows = (tap, cout), ONE MMA group per
  uint32_t s_bytes;                               // chunk on the unshifted halo tile, taps summed in the epilogue via smem S
  const float* bias; float slope;
  void* out; long out_pitch; int out_coff;
  void* out2; long out2_pitch; int out2_coff; int n_split;
  int out_f32;
  const float* residual; long res_pitch;
};

// debug timeline of CTA 0 (PWC_TC_HALO_BO bit 5): [role 0 producer | 1 mma | 2 epilogue][item index][4 marks]
__device__ unsigned long long g_wm_trace[3 * 64 * 4];
__device__ __forceinline__ void wm_mark(bool on, int role, int idx, int k) {
  if (on && idx < 64) g_wm_trace[(role * 64 + idx) * 4 + k] = clock64();
}

__device__ __forceinline__ void named_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// Epilogue staging of 8 accumulator columns (8 pixels of one tile row) of one thread (= one output channel): bias, LeakyReLU,
// 16-bit rounding, one staging row per pixel (`stride` 16-bit elements apart); LO also stores the rounding residual (the lo
// plane of a split output).  The 16-bit format, the lo plane and the activation form are compile-time here: with run-time flags
// and a per-value bounds test every value was two branches plus a serial FADD-FMUL-FMNMX-F2FP-STS chain, ~45 cycles per value
// with one epilogue warp per scheduler (ncu source page of dc_conv2: 13 k cycles of epilogue per 256-pixel tile against 9 k of
// MMAs); branch-free, the eight chains interleave.
template <bool F16, bool LO, bool MAXFORM>
__device__ __forceinline__ void wm_stage8(const uint32_t* v, float bias, float slope, uint16_t* st, uint16_t* st_lo, int stride) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float a = __uint_as_float(v[i]) + bias;
    const float val = MAXFORM ? fmaxf(a, a * slope) : lrelu(a, slope);      // MAXFORM: 0 <= slope <= 1
    if (F16) {
      const __half hv = __float2half_rn(val);
      st[i * stride] = __half_as_ushort(hv);
      if (LO) st_lo[i * stride] = __half_as_ushort(__float2half_rn(val - __half2float(hv)));
    } else {
      const __nv_bfloat16 hv = __float2bfloat16_rn(val);
      st[i * stride] = __bfloat16_as_ushort(hv);
      if (LO) st_lo[i * stride] = __bfloat16_as_ushort(__float2bfloat16_rn(val - __bfloat162float(hv)));
    }
  }
}

// ncol (a multiple of 8, <= 8 * NB) columns held in v[8 * NB]: one warp-uniform test per 8 columns, one format dispatch per call.
// STRIDE > 0: compile-time staging stride, else `stride`.
template <int NB, int STRIDE, bool MAXFORM>
__device__ __forceinline__ void wm_stage(const uint32_t* v, int ncol, int fp16, bool lo, float bias, float slope,
                                         uint16_t* st, uint16_t* st_lo, int stride) {
  const int s = STRIDE > 0 ? STRIDE : stride;
  if (fp16) {
    if (lo) {
#pragma unroll
      for (int b = 0; b < NB; ++b) if (b * 8 < ncol) wm_stage8<true, true, MAXFORM>(v + b * 8, bias, slope, st + b * 8 * s, st_lo + b * 8 * s, s);
    } else {
#pragma unroll
      for (int b = 0; b < NB; ++b) if (b * 8 < ncol) wm_stage8<true, false, MAXFORM>(v + b * 8, bias, slope, st + b * 8 * s, st_lo, s);
    }
  } else {
    if (lo) {
#pragma unroll
      for (int b = 0; b < NB; ++b) if (b * 8 < ncol) wm_stage8<false, true, MAXFORM>(v + b * 8, bias, slope, st + b * 8 * s, st_lo + b * 8 * s, s);
    } else {
#pragma unroll
      for (int b = 0; b < NB; ++b) if (b * 8 < ncol) wm_stage8<false, false, MAXFORM>(v + b * 8, bias, slope, st + b * 8 * s, st_lo, s);
    }
  }
}

// kMode = 0: the plain 9-tap path with nothing extra on the MMA issue path (the big layers run at > 90% of the tensor
// peak only while the issuing thread does no more than descriptor arithmetic); the tap-stacked mode and the zero-block /
// K-step skip lists are separate instances.
// kMode: 0 plain | 1 tap-stacked | 2 zero-block skip list | 3 K-step skip masks | 4 tap lanes | 5 stride 2 (one instance each: see above)
template <int kMode>
__global__ void __launch_bounds__(kHaloThreads, 1)
conv_tc_wm_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                  const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmO2, const WmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_addr(smem_raw) + 1023u) & ~1023u;
  const uint32_t x_base = base;                                   // x_stages (2..4) halo buffers
  const uint32_t w_base = base + (uint32_t)p.x_stages * p.x_bytes;                  // weight ring (or the resident weight set)
  const uint32_t nw = p.w_resident ? (uint32_t)(9 * p.nchunk) : p.w_stages * p.tps;
  const uint32_t o_base = w_base + nw * p.w_bytes + p.w_pad;      // w_pad: the M=128 MMA reads 128 rows from a CoutP-row tile
  const uint32_t s_base = o_base + p.o_bytes;                     // S: per-tap partial sums [9*cout][s_stride] fp32 (TS mode)
  const uint32_t bar_base = s_base + p.s_bytes;                   // o: output staging tile (half an item)
  auto x_full = [&](int i) { return bar_base + 8u * i; };
  auto x_empty = [&](int i) { return bar_base + 8u * (4 + i); };
  auto t_full = [&](int i) { return bar_base + 8u * (8 + i); };
  auto t_empty = [&](int i) { return bar_base + 8u * (10 + i); };
  const uint32_t w_all = bar_base + 8u * 12;
  auto w_full = [&](int s) { return bar_base + 8u * (13 + s); };
  auto w_empty = [&](int s) { return bar_base + 8u * (13 + p.w_stages + s); };
  const uint32_t tmem_slot = bar_base + 8u * (13 + 2 * p.w_stages);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - ptx::smem_addr(smem_raw)));
  volatile int* ks_flag = reinterpret_cast<volatile int*>(smem_raw + (tmem_slot + 8 - ptx::smem_addr(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int d = p.dil;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmX); ptx::prefetch_tmap(&tmW);
    for (int i = 0; i < 4; ++i) { ptx::mbar_init(x_full(i), 1); ptx::mbar_init(x_empty(i), 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(t_full(i), 1); ptx::mbar_init(t_empty(i), 4); }
    ptx::mbar_init(w_all, 1);
    for (uint32_t s = 0; s < p.w_stages; ++s) { ptx::mbar_init(w_full(s), 1); ptx::mbar_init(w_empty(s), 1); }
    ptx::fence_barrier_init();
    ptx::fence_proxy_async();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, p.tmem_cols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  auto decode = [&](int item, int& n, int& ry, int& rx, int& ty0, int& tx0) {
    n = item / p.tiles_per_img;
    int r = item - n * p.tiles_per_img;
    const int per_phase = p.tiles_x * p.tiles_y;
    const int ph = r / per_phase;
    r -= ph * per_phase;
    if (p.dy == 1) { ry = 0; rx = ph; } else { ry = ph / d; rx = ph - ry * d; }
    ty0 = (r / p.tiles_x) * p.rows; tx0 = (r % p.tiles_x) * 8;
  };

  // chunk ch multiplies the x tile chunk ch - 1 loaded (WmParams::xdup); a CTA's first chunk always loads
  auto reuse_x = [&](int ch, int ch_lo) { return (kMode == 5) && p.xdup && ch != ch_lo && (p.fused ? (ch & 1) != 0 : (ch % 3) == 1); };

  if (warp == 0) {
    // ===================== TMA producer (whole warp converged; one elected lane issues) =====================
    if (p.w_resident && ptx::elect_one()) {
      ptx::mbar_expect_tx(w_all, 9u * p.nchunk * p.w_bytes_tx * p.rep);
      for (int j = 0; j < 9 * p.nchunk; ++j)
        for (int r = 0; r < p.rep; ++r)
          ptx::tma_load_2d(w_base + j * p.w_bytes + r * p.cg * 32 * p.kc * 2, &tmW, w_all, 0, j * p.coutp);
    }
    __syncwarp();
    pdl_wait();                                          // weights above do not depend on the previous kernel; activations do
    pdl_launch_dependents();
    int xb = 0; uint32_t xph = 0; uint32_t s = 0, wph = 0;
    for (int vitem = blockIdx.x; vitem < p.num_items * p.ksplit; vitem += gridDim.x) {
      const int item = vitem / p.ksplit, ks = vitem - item * p.ksplit;
      const int ch_lo = ks * p.cps, ch_hi = min(p.nchunk, ch_lo + p.cps);
      int n, ry, rx, ty0, tx0;
      decode(item, n, ry, rx, ty0, tx0);
      const int ya = ry + p.dy * (ty0 - p.kyr), xa = rx + d * (tx0 - 1);
      const bool trp = (p.dbg & 32) && blockIdx.x == 0 && lane == 0;
      const int it_idx = (vitem - (int)blockIdx.x) / (int)gridDim.x;
      for (int ch = ch_lo; ch < ch_hi; ++ch) {
        if (ch == ch_lo) wm_mark(trp, 0, it_idx, 0);
        if (!reuse_x(ch, ch_lo)) {
          ptx::mbar_wait(x_empty(xb), xph ^ 1);
          if (ch == ch_lo) wm_mark(trp, 0, it_idx, 1);
          if (ptx::elect_one()) {
            ptx::mbar_expect_tx(x_full(xb), p.x_bytes_tx);
            const int xc = ((kMode == 5) && p.fused) ? 0 : p.split ? (ch / p.split) * p.kc + ((ch % p.split) == p.split - 1 ? p.lo_off : 0) : ch * p.kc;
            if ((kMode == 5)) {
              ptx::tma_load_5d(x_base + xb * p.x_bytes, &tmX, x_full(xb), xc, 1, tx0 - 1, 2 * ty0 - 1, n);
              ptx::tma_load_5d(x_base + xb * p.x_bytes + p.x_sub, &tmX, x_full(xb), xc, 0, tx0 - 1, 2 * ty0 - 1, n);
            } else if (p.xsplit) ptx::tma_load_5d(x_base + xb * p.x_bytes, &tmX, x_full(xb), xc, rx, tx0 - 1, ya, n);
            else ptx::tma_load_4d(x_base + xb * p.x_bytes, &tmX, x_full(xb), xc, xa, ya, n);
          }
          __syncwarp();
          if (++xb == p.x_stages) { xb = 0; xph ^= 1; }
        }
        if (!p.w_resident) {
          for (int tap = 0; tap < (kMode == 1 ? 1 : kMode == 4 ? p.ntaps : 9); tap += p.tps) {
            if ((kMode == 2) && ((p.skip[tap] >> ch) & 1ull)) continue;
            ptx::mbar_wait(w_empty(s), wph ^ 1);
            if (ptx::elect_one()) {
              ptx::mbar_expect_tx(w_full(s), p.w_bytes_tx * p.tps);
              for (int q = 0; q < p.tps; ++q)
                ptx::tma_load_2d(w_base + (s * p.tps + q) * p.w_bytes, &tmW, w_full(s), 0, ((tap + q) * p.nchunk + ch) * p.coutp);
            }
            __syncwarp();
            if (++s == p.w_stages) { s = 0; wph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: whole warp converged, one elected lane issues =====================
    uint32_t idesc = ptx::make_idesc_bf16(128, p.npx);
    if (p.fp16) idesc &= ~((7u << 7) | (7u << 10));          // A/B format 0 = F16
    const uint32_t row_bytes = p.kc * 2;
    const uint32_t ltype = (p.kc == 64) ? 2u : (p.kc == 32) ? 4u : 6u;
    const uint32_t sbo_x = ((kMode == 1) ? 8u : (kMode == 5) ? 18u : 10u) * row_bytes;   // TS: all halo pixels in order; else 8 of every 10
                                                                                  // (stride 2: of every second 9-pair row)
    int xb = 0; uint32_t xph = 0; uint32_t s = 0, wph = 0;
    int acc = 0; uint32_t tph = 0;
    if (p.w_resident) ptx::mbar_wait(w_all, 0);
    for (int vitem = blockIdx.x; vitem < p.num_items * p.ksplit; vitem += gridDim.x) {
      const int item = vitem / p.ksplit, ks = vitem - item * p.ksplit;
      const int ch_lo = ks * p.cps, ch_hi = min(p.nchunk, ch_lo + p.cps);
      const bool trm = (p.dbg & 32) && blockIdx.x == 0 && lane == 0;
      const int it_idx = (vitem - (int)blockIdx.x) / (int)gridDim.x;
      wm_mark(trm, 1, it_idx, 0);
      ptx::mbar_wait(t_empty(acc), tph ^ 1);
      ptx::tc_fence_after();
      wm_mark(trm, 1, it_idx, 1);
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * ((kMode == 1) ? 256 : p.npx));
      bool first = true;                                   // the first MMA of an item overwrites the accumulator
      for (int ch = ch_lo; ch < ch_hi; ++ch) {
        const bool keep_x = (ch + 1 < ch_hi) && reuse_x(ch + 1, ch_lo);   // the next chunk multiplies this tile again
        if (!reuse_x(ch, ch_lo)) {
          ptx::mbar_wait(x_full(xb), xph);
          ptx::tc_fence_after();
        }
        if (ch == ch_lo) wm_mark(trm, 1, it_idx, 2);
        const uint32_t sx = x_base + xb * p.x_bytes;
        const int nkk = ((kMode == 5) && p.fused) ? ((ch & 1) ? p.kc >> 5 : p.kc >> 4)       // [W_lo | 0]: the hi half of the K range only
                                : min(p.kc, p.cin_span - (p.split ? ch / p.split : ch) * p.kc + 15) >> 4;   // K=16 steps holding real channels
        const uint64_t dx0 = ptx::make_kmajor_desc(sx, sbo_x, ltype);
        int ftap = 0;                                        // first tap actually issued for this chunk
        if ((kMode == 2)) while (ftap < 8 && ((p.skip[ftap] >> ch) & 1ull)) ++ftap;
        auto issue_tap = [&](uint32_t sw, int tap) {
          const int ky = (kMode == 4) ? (tap < 3 ? 0 : 2) : tap / 3, kx = tap - 3 * (tap / 3);   // kMode 4: slots 3..5 = ky 2, one lane
          const uint64_t dw = ptx::make_kmajor_desc(sw, 8 * row_bytes, ltype);
          const uint32_t xoff = (kMode == 5) ? (uint32_t)(ky * 9 + (kx ? 1 : 0)) * row_bytes + (kx == 1 ? p.x_sub : 0u)
                                     : (uint32_t)(ky * p.kyr * 10 + kx) * row_bytes;
          const uint64_t dx = dx0 + (uint64_t)(xoff >> 4);
          if ((kMode == 3)) {
            const int km = p.kk_mask[tap];
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              if (kk < nkk && !((km >> kk) & 1))
                ptx::umma_bf16(d_tmem, dw + 2 * kk, dx + 2 * kk, idesc, (tap == p.kk_first_tap && kk == p.kk_first_kk) ? 0u : 1u);
            return;
          }
          ptx::umma_bf16(d_tmem, dw, dx, idesc, (first && tap == ftap) ? 0u : 1u);
          if (nkk > 1) ptx::umma_bf16(d_tmem, dw + 2, dx + 2, idesc, 1u);
          if (nkk > 2) ptx::umma_bf16(d_tmem, dw + 4, dx + 4, idesc, 1u);
          if (nkk > 3) ptx::umma_bf16(d_tmem, dw + 6, dx + 6, idesc, 1u);
        };
        if (p.w_resident) {
          if (ptx::elect_one()) {
#pragma unroll 1
            for (int tap = 0; tap < 9; ++tap) issue_tap(w_base + (tap * p.nchunk + ch) * p.w_bytes, tap);
            if (!keep_x) ptx::umma_commit(x_empty(xb));
            if (ch == ch_hi - 1) ptx::umma_commit(t_full(acc));
          }
          __syncwarp();
        } else {
          for (int tap0 = 0; tap0 < (kMode == 1 ? 1 : kMode == 4 ? p.ntaps : 9); tap0 += p.tps) {
            if ((kMode == 2) && ((p.skip[tap0] >> ch) & 1ull)) continue;
            ptx::mbar_wait(w_full(s), wph);
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
              const uint32_t sw0 = w_base + s * p.tps * p.w_bytes;
              issue_tap(sw0, tap0);
              if (p.tps == 3) { issue_tap(sw0 + p.w_bytes, tap0 + 1); issue_tap(sw0 + 2 * p.w_bytes, tap0 + 2); }
              ptx::umma_commit(w_empty(s));
            }
            __syncwarp();
            if (++s == p.w_stages) { s = 0; wph ^= 1; }
          }
          if (ptx::elect_one()) {
            if (!keep_x) ptx::umma_commit(x_empty(xb));
            if (ch == ch_hi - 1) ptx::umma_commit(t_full(acc));
          }
          __syncwarp();
        }
        if (ch == ch_hi - 1) wm_mark(trm, 1, it_idx, 3);
        if (!(kMode == 2) || !((p.skip[ftap] >> ch) & 1ull)) first = false;   // some tap of this chunk was issued
        if (!keep_x && ++xb == p.x_stages) { xb = 0; xph ^= 1; }
      }
      if (++acc == 2) { acc = 0; tph ^= 1; }
    }
  } else {
    // ===================== epilogue warps 2..5: thread = output channel =====================
    pdl_wait();
    const int lg = warp & 3;
    const int part = lg / p.cg;                          // which pixel part this warp stores (weight replicas on M)
    const int co = (lg - part * p.cg) * 32 + lane;
    const bool co_live = p.out_f32 ? (co < p.cout) : (co < p.coutp);
    const bool warp_live = (part < p.rep) && (lg - part * p.cg) * 32 < (p.out_f32 ? p.cout : p.coutp);
    const float bias = (co < p.cout) ? __ldg(p.bias + co) : 0.f;
    uint8_t* o_gen = smem_raw + (o_base - ptx::smem_addr(smem_raw));
    const bool issuer = (threadIdx.x == 64);
    int obuf = 0, wbuf = 0;
    int acc = 0; uint32_t tph = 0;
    for (int vitem = blockIdx.x; vitem < p.num_items * p.ksplit; vitem += gridDim.x) {
      const int item = vitem / p.ksplit, ks = vitem - item * p.ksplit;
      const int ch_lo = ks * p.cps, ch_hi = min(p.nchunk, ch_lo + p.cps);
      int n, ry, rx, ty0, tx0;
      decode(item, n, ry, rx, ty0, tx0);
      const bool second = (n >= p.n_split) && (p.out2 != nullptr) && (p.out_f32 != 2);
      const int nn = second ? n - p.n_split : n;
      const bool tre = (p.dbg & 32) && blockIdx.x == 0 && threadIdx.x == 128;     // warp 4 = lane group 0 (always live)
      const int it_idx = (vitem - (int)blockIdx.x) / (int)gridDim.x;
      wm_mark(tre, 2, it_idx, 0);
      ptx::mbar_wait(t_full(acc), tph);
      ptx::tc_fence_after();
      wm_mark(tre, 2, it_idx, 1);
      const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(acc * ((kMode == 1) ? 256 : p.npx));
      if ((kMode == 1)) {
        // ---- tap-stacked: TMEM lane m = tap*cout + co holds that tap's partial sums over the whole halo tile ----
        float* S = reinterpret_cast<float*>(smem_raw + (s_base - ptx::smem_addr(smem_raw)));
        const int m = lg * 32 + lane, mrows = 9 * p.cout;
        if (lg * 32 < mrows) {
          float4* srow = reinterpret_cast<float4*>(S + (size_t)m * p.s_stride);
          for (int c0 = 0; c0 < p.npx; c0 += 16) {
            uint32_t v[16];
            ptx::tmem_ld16(taddr + c0, v);
            ptx::tmem_ld_wait();
            if (m < mrows) {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                srow[(c0 >> 2) + i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                                  __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
            }
          }
        }
        ptx::tc_fence_before();
        named_bar(2, 128);
        if (lane == 0) ptx::mbar_arrive(t_empty(acc));       // accumulator is in S now: the next item may start
        // thread = output pixel, all Cout channels of it in turn (the (channel, row, pixel) enumeration this replaces paid two
        // integer divisions and a 64-bit address per VALUE; neighbouring lanes still read neighbouring words of S)
        const int et = threadIdx.x - 64;
        const int npix = p.rows * 8;
        for (int pxi = et; pxi < npix; pxi += 128) {
          const int px = pxi & 7, py = pxi >> 3;
          const int oy = ty0 + py, ox = tx0 + px;
          if (oy >= p.H || ox >= p.W) continue;
          const long pix = ((long)n * p.H + oy) * p.W + ox;
          const float* sp0 = S + py * 10 + px;
          const float* part = p.partial ? p.partial + pix * p.partial_pitch : nullptr;
          float* o32 = reinterpret_cast<float*>(p.out) + pix * p.out_pitch + p.out_coff;
          const float* res = p.residual ? p.residual + pix * p.res_pitch : nullptr;
          const long up0 = (((long)n * 2 * p.H + 2 * oy) * (2 * p.W) + 2 * ox) * p.out2_pitch + p.out2_coff;
          for (int co2 = 0; co2 < p.cout; ++co2) {
            const float* sp = sp0 + (size_t)co2 * p.s_stride;
            float a = __ldg(p.bias + co2);
#pragma unroll
            for (int tap = 0; tap < 9; ++tap)
              a += sp[(size_t)tap * p.cout * p.s_stride + (tap / 3) * 10 + (tap % 3)];
            if (part) a += __ldg(part + co2);              // the other input channels' share (WmParams::nhead)
            float val = lrelu(a, p.slope);
            if ((p.out_f32 & 3) == 2 && co2 >= 2) {
              const int k = co2 - 2, j = k & 1, b = (k >> 1) & 1, aa = k >> 2;
              const long up = up0 + ((long)aa * (2 * p.W) + b) * p.out2_pitch + j;
              const uint16_t hb = pack1_h16(val, p.fp16);
              reinterpret_cast<uint16_t*>(p.out2)[up] = hb;
              if (p.up_lo_off) reinterpret_cast<uint16_t*>(p.out2)[up + p.up_lo_off] = pack1_h16(val - unpack1_h16(hb, p.fp16), p.fp16);
              if (p.up_dup_off) reinterpret_cast<uint16_t*>(p.out2)[up + p.up_dup_off] = hb;
            } else {
              if (res) val += res[co2];
              o32[co2] = val;
            }
          }
        }
        named_bar(2, 128);                                   // S may be overwritten by the next item
        if (++acc == 2) { acc = 0; tph ^= 1; }
        continue;
      }
      if ((kMode == 4)) {
        // ---- tap lanes: warp = (lane l of the MMA rows, 32-channel group cgi); phase ph handles pixel block (l + ph) % G ----
        const int G = p.tl, l = lg / p.cg, cgi = lg - l * p.cg;
        const int bh = p.rows / G;                           // tile rows per pixel block
        if (l < G) {
          float* S = reinterpret_cast<float*>(smem_raw + (s_base - ptx::smem_addr(smem_raw))) + (size_t)cgi * p.rows * 8 * 32 + lane;
          const uint32_t my_o = o_base + (uint32_t)lg * 2u * p.o_half;
          for (int ph = 0; ph < G; ++ph) {
            const int b = (l + ph) % G;
            for (int r0 = 0; r0 < bh; r0 += p.o_rs) {
              const int row = b * bh + r0, ncol = p.o_rs * 8;
              uint32_t v[4][16];
#pragma unroll
              for (int q = 0; q < 4; ++q) {                   // lane l: output row r sits at r + l; never read past the
                if (q * 16 + 16 <= ncol) ptx::tmem_ld16(taddr + (row + l) * 8 + q * 16, v[q]);   // accumulator's last column
                else if (q * 16 < ncol) ptx::tmem_ld8(taddr + (row + l) * 8 + q * 16, *reinterpret_cast<uint32_t(*)[8]>(&v[q][0]));
              }
              ptx::tmem_ld_wait();
              float* sp = S + (size_t)row * 8 * 32;
              uint32_t* vf = &v[0][0];
              if (ph < G - 1) {
#pragma unroll
                for (int b8 = 0; b8 < 8; ++b8)
                  if (b8 * 8 < ncol) {                         // warp-uniform, once per 8 columns
                    if (ph == 0) {
#pragma unroll
                      for (int i = 0; i < 8; ++i) sp[(b8 * 8 + i) * 32] = __uint_as_float(vf[b8 * 8 + i]);
                    } else {
#pragma unroll
                      for (int i = 0; i < 8; ++i) sp[(b8 * 8 + i) * 32] += __uint_as_float(vf[b8 * 8 + i]);
                    }
                  }
              } else {
                const uint32_t so = my_o + (uint32_t)wbuf * p.o_half;
                const uint32_t so_lo = my_o + (uint32_t)(wbuf ^ 1) * p.o_half;     // split output: the other buffer takes the lo plane
                if (lane == 0) { if (p.out_lo_off) ptx::bulk_wait_read0(); else ptx::bulk_wait_read1(); }
                __syncwarp();
                const int bc = p.tl_bc;                      // staged pixel = bc channels (the store box)
                uint16_t* st = reinterpret_cast<uint16_t*>(smem_raw + (so - ptx::smem_addr(smem_raw))) + lane;
                uint16_t* st_lo = reinterpret_cast<uint16_t*>(smem_raw + (so_lo - ptx::smem_addr(smem_raw))) + lane;
                const float bia = (cgi * 32 + lane < p.cout) ? __ldg(p.bias + cgi * 32 + lane) : 0.f;
                if (lane < bc) {
#pragma unroll
                  for (int b8 = 0; b8 < 8; ++b8)
                    if (b8 * 8 < ncol) {
#pragma unroll
                      for (int i = 0; i < 8; ++i)
                        vf[b8 * 8 + i] = __float_as_uint(__uint_as_float(vf[b8 * 8 + i]) + sp[(b8 * 8 + i) * 32]);
                    }
                  if (bc == 16) wm_stage<8, 16, true>(vf, ncol, p.fp16, p.out_lo_off != 0, bia, p.slope, st, st_lo, 16);
                  else wm_stage<8, 32, true>(vf, ncol, p.fp16, p.out_lo_off != 0, bia, p.slope, st, st_lo, 32);
                }
                ptx::fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                  ptx::tma_store_4d(second ? &tmO2 : &tmO, so, cgi * 32, rx + d * tx0, ry + p.dy * (ty0 + row), nn);
                  if (p.out_lo_off)
                    ptx::tma_store_4d(second ? &tmO2 : &tmO, so_lo, cgi * 32 + p.out_lo_off, rx + d * tx0, ry + p.dy * (ty0 + row), nn);
                  ptx::bulk_commit();
                }
                if (!p.out_lo_off) wbuf ^= 1;
              }
            }
            if (ph == G - 1) {                               // every TMEM read of this item is done
              ptx::tc_fence_before();
              __syncwarp();
              if (lane == 0) ptx::mbar_arrive(t_empty(acc));
            }
            named_bar(3 + cgi, 32 * G);                      // the block's partial sums are visible to the next phase's warp
          }
        } else {
          if (p.nhead > 0 && l == G) {
            // ---- head rows: TMEM lane 96 + ky * nhead + h; output row r of kernel row ky sits at accumulator row r + ky.
            //      Phase 1: every lane drops its row's value of the step's 48 pixels into hs[pixel][lane]; phase 2: the 48 * nhead
            //      (pixel, channel) sums of three kernel rows are dealt over ALL lanes (two shuffles per value on nhead live lanes
            //      made this warp the slowest of the tile: conv2_4 91 -> 145 us) and stored as contiguous fp32 runs ----
            const int NH = p.nhead, kyt = lane / NH;
            float* hs = reinterpret_cast<float*>(smem_raw + (((tmem_slot + 16u + 15u) & ~15u) - ptx::smem_addr(smem_raw)));
            for (int r0 = 0; r0 < p.rows; r0 += 6) {         // six tile rows per step (rows % 6 == 0): ONE wait for the 64 columns
              uint32_t v[4][16];                             // = accumulator rows r0 .. r0 + 7 that the three kernel rows need
#pragma unroll
              for (int q = 0; q < 4; ++q) ptx::tmem_ld16(taddr + r0 * 8 + q * 16, v[q]);
              ptx::tmem_ld_wait();
              const uint32_t* vf = &v[0][0];
              if (kyt == 0) {
#pragma unroll
                for (int i = 0; i < 48; ++i) hs[i * 32 + lane] = __uint_as_float(vf[i]);
              } else if (kyt == 1) {
#pragma unroll
                for (int i = 0; i < 48; ++i) hs[i * 32 + lane] = __uint_as_float(vf[i + 8]);
              } else if (kyt == 2) {
#pragma unroll
                for (int i = 0; i < 48; ++i) hs[i * 32 + lane] = __uint_as_float(vf[i + 16]);
              }
              __syncwarp();
              for (int idx = lane; idx < 48 * NH; idx += 32) {
                const int i = idx / NH, hh = idx - i * NH;
                const float tot = hs[i * 32 + hh] + hs[i * 32 + NH + hh] + hs[i * 32 + 2 * NH + hh];
                const int oy = ty0 + r0 + (i >> 3), ox = tx0 + (i & 7);
                if (oy < p.H && ox < p.W) p.head_out[(((long)n * p.H + oy) * p.W + ox) * NH + hh] = tot;
              }
              __syncwarp();
            }
          }
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(t_empty(acc));
        }
        if (++acc == 2) { acc = 0; tph ^= 1; }
        continue;
      }
      if (p.wstore) {
        // ---- per-warp stores: warp (channel group gc, pixel part) stages [o_rs rows x 8 px][32 ch] and issues its own
        //      TMA store (box 32 ch x 8 px x o_rs rows); no CTA-wide barrier, two staging buffers per warp ----
        const int gc = lg - part * p.cg;                     // 32-channel group of this warp
        const int R = p.rows / p.rep;                        // tile rows of this warp's part
        // partial sums of this CTA -> scratch [item][pixel][128 ch] (split-K only); lanes = consecutive channels
        float* scr = p.scratch + ((size_t)item * p.npx + (size_t)part * R * 8) * 128 + gc * 32 + lane;
        bool finish = true;
        if (p.ksplit > 1) {
          if (warp_live) {
            for (int c0 = 0; c0 < R * 8; c0 += 64) {
              uint32_t v[4][16];
#pragma unroll
              for (int q = 0; q < 4; ++q)
                if (c0 + q * 16 < R * 8) ptx::tmem_ld16(taddr + part * R * 8 + c0 + q * 16, v[q]);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int q = 0; q < 4; ++q)
                if (c0 + q * 16 < R * 8) {
#pragma unroll
                  for (int i = 0; i < 16; ++i)
                    if (c0 + q * 16 + i < R * 8) {
                      if (p.ks2pass) __stcg(scr + (size_t)ks * p.num_items * p.npx * 128 + (size_t)(c0 + q * 16 + i) * 128, __uint_as_float(v[q][i]));
                      else atomicAdd(scr + (size_t)(c0 + q * 16 + i) * 128, __uint_as_float(v[q][i]));
                    }
                }
            }
          }
          // accumulator drained: hand it back before the cross-CTA rendezvous
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(t_empty(acc));
          if (p.ks2pass) {                                   // the reduce kernel finishes the tile
            if (++acc == 2) { acc = 0; tph ^= 1; }
            continue;
          }
          __threadfence();
          named_bar(2, 128);
          if (threadIdx.x == 64) {
            const int old = atomicAdd(p.counters + item, 1);
            *ks_flag = (old == p.ksplit - 1) ? 1 : 0;
            if (old == p.ksplit - 1) p.counters[item] = 0;   // ready for the next forward
          }
          named_bar(2, 128);
          finish = (*ks_flag != 0);
          if (finish) __threadfence();
          named_bar(2, 128);                                 // flag may be rewritten for the next item
        }
        if (warp_live && finish) {
          const uint32_t my_o = o_base + (uint32_t)lg * 2u * p.o_half;
          int c0g = gc * 32, grp = 0;
          if (p.out_groups > 1) { const int cgp = p.coutp / p.out_groups; grp = c0g / cgp; c0g -= grp * cgp; }
          for (int r0 = 0; r0 < R; r0 += p.o_rs) {
            const uint32_t so = my_o + (uint32_t)wbuf * p.o_half;
            const uint32_t so_lo = my_o + (uint32_t)(wbuf ^ 1) * p.o_half;   // split output: the other buffer takes the lo plane
            // (four buffers per warp, so that a split output's stores need not drain before the next step, were measured: no gain
            // on the pyramid layers and the smaller weight ring cost conv4_1 24.8 -> 29.1 us)
            if (lane == 0) { if (p.out_lo_off) ptx::bulk_wait_read0(); else ptx::bulk_wait_read1(); }   // buffer(s) released
            __syncwarp();
            uint16_t* st = reinterpret_cast<uint16_t*>(smem_raw + (so - ptx::smem_addr(smem_raw))) + lane;
            uint16_t* st_lo = reinterpret_cast<uint16_t*>(smem_raw + (so_lo - ptx::smem_addr(smem_raw))) + lane;
            const int col0 = (part * R + r0) * 8, ncol = p.o_rs * 8;
            // all TMEM loads of the step (<= 64 columns) are issued before the single wait: their latency (several
            // hundred cycles while the next item's MMAs are writing TMEM) is paid once per step, not once per 16 columns
            uint32_t v[4][16];
            if (p.ksplit > 1) {                              // last CTA of the tile: the sums are complete in the scratch
#pragma unroll
              for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int i = 0; i < 16; ++i)
                  if (q * 16 + i < ncol) {
                    float* sp = scr + (size_t)(r0 * 8 + q * 16 + i) * 128;
                    v[q][i] = __float_as_uint(__ldcg(sp));
                    __stcg(sp, 0.f);                         // leave the scratch zeroed for the next forward
                  }
            } else {
#pragma unroll
              for (int q = 0; q < 4; ++q)
                if (q * 16 < ncol) ptx::tmem_ld16(taddr + col0 + q * 16, v[q]);
              ptx::tmem_ld_wait();
            }
            wm_stage<8, 32, true>(&v[0][0], ncol, p.fp16, p.out_lo_off != 0, bias, p.slope, st, st_lo, 32);   // 0 <= slope <= 1
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              const int oy = ry + p.dy * (ty0 + part * R + r0), ox = rx + d * tx0;
              if (p.out_groups > 1) ptx::tma_store_5d(&tmO, so, c0g, grp, ox, oy, nn);
              else if (p.xsplit) ptx::tma_store_5d(&tmO, so, c0g, rx, tx0, oy, nn);
              else ptx::tma_store_4d(second ? &tmO2 : &tmO, so, c0g, ox, oy, nn);
              if (p.out_lo_off) ptx::tma_store_4d(second ? &tmO2 : &tmO, so_lo, c0g + p.out_lo_off, ox, oy, nn);
              ptx::bulk_commit();
            }
            if (!p.out_lo_off) wbuf ^= 1;
          }
        }
        if (p.ksplit > 1) {                                  // t_empty was signalled above
          wm_mark(tre, 2, it_idx, 3);
          if (++acc == 2) { acc = 0; tph ^= 1; }
          continue;
        }
        wm_mark(tre, 2, it_idx, 3);
      } else if (p.tma_store) {
        const int hpx = p.npx >> 1;                        // pixels per half item (rows/2 tile rows)
        for (int half = 0; half < 2; ++half) {
          // two staging buffers: the TMA store of the previous half keeps reading its buffer while this one is filled
          const uint32_t ooff = (p.o_nbuf == 2) ? (uint32_t)obuf * p.o_half : 0u;
          if (p.o_nbuf == 2) {
            if (issuer) ptx::bulk_wait_read1();            // the store issued two halves ago has released this buffer
            named_bar(2, 128);
          }
          if (warp_live) {
            uint16_t* st = reinterpret_cast<uint16_t*>(o_gen + ooff) + co;
            const int ppx = ((hpx / p.rep) + 7) & ~7;         // pixels per part
            const int cend = min(hpx, (part + 1) * ppx);
            for (int c0 = part * ppx; c0 < cend; c0 += 16) {
              uint32_t v[16];
              ptx::tmem_ld16(taddr + half * hpx + c0, v);
              ptx::tmem_ld_wait();
              if (co_live)                                   // c0, cend: multiples of 8
                wm_stage<2, 0, false>(v, cend - c0, p.fp16, false, bias, p.slope, st + (size_t)c0 * p.coutp, st, p.coutp);
            }
          }
          ptx::fence_proxy_async();                        // generic-proxy smem writes -> visible to the TMA store
          named_bar(2, 128);
          if (issuer) {
            if (p.out_groups > 1)
              ptx::tma_store_5d(&tmO, o_base + ooff, 0, 0, rx + d * tx0, ry + p.dy * (ty0 + half * (p.rows >> 1)), nn);
            else
              ptx::tma_store_4d(second ? &tmO2 : &tmO, o_base + ooff, 0, rx + d * tx0, ry + p.dy * (ty0 + half * (p.rows >> 1)), nn);
            ptx::bulk_commit();
            if (p.o_nbuf != 2) ptx::bulk_wait_read0();     // single staging tile: wait until it may be overwritten
          }
          if (p.o_nbuf != 2) named_bar(2, 128);
          obuf ^= 1;
          wm_mark(tre, 2, it_idx, 2 + half);
        }
      } else if (warp_live) {
        for (int c0 = 0; c0 < p.npx; c0 += 16) {           // 16 pixels = 2 tile rows of 8
          uint32_t v[16];
          ptx::tmem_ld16(taddr + c0, v);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int py = (c0 + i) >> 3, px = (c0 + i) & 7;
            const int oy = ry + p.dy * (ty0 + py), ox = rx + d * (tx0 + px);
            if (co_live && oy < p.H && ox < p.W) {
              const long pix = ((long)nn * p.H + oy) * p.W + ox;
              float val = lrelu(__uint_as_float(v[i]) + bias, p.slope);
              if (p.out_f32 == 2 && co >= 2) {
                // fused upfeat (ConvTranspose 4x4 s2 as 4 phase 3x3 convs): channel 2 + 2*(2a+b) + j -> pixel (2y+a, 2x+b)
                const int k = co - 2, j = k & 1, b = (k >> 1) & 1, a = k >> 2;
                const long up = (((long)n * 2 * p.H + 2 * oy + a) * (2 * p.W) + 2 * ox + b) * p.out2_pitch + p.out2_coff + j;
                reinterpret_cast<uint16_t*>(p.out2)[up] = pack1_h16(val, p.fp16);
              } else if (p.out_f32) {
                if (p.residual) val += p.residual[pix * p.res_pitch + co];
                reinterpret_cast<float*>(p.out)[pix * p.out_pitch + p.out_coff + co] = val;
              } else {
                uint16_t* o = second ? reinterpret_cast<uint16_t*>(p.out2) + pix * p.out2_pitch + p.out2_coff
                                     : reinterpret_cast<uint16_t*>(p.out) + pix * p.out_pitch + p.out_coff;
                uint16_t bits;
                if (p.fp16) { __half hv = __float2half_rn(val); bits = *reinterpret_cast<uint16_t*>(&hv); }
                else { __nv_bfloat16 hv = __float2bfloat16_rn(val); bits = *reinterpret_cast<uint16_t*>(&hv); }
                o[co] = bits;
              }
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(t_empty(acc));
      if (++acc == 2) { acc = 0; tph ^= 1; }
    }
    if (issuer || (p.wstore && lane == 0)) ptx::bulk_wait0();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, p.tmem_cols);
}

// Second pass of the deterministic split-K (WmParams::ks2pass): out[pixel][c] = act(bias[c] + sum_ks partial[ks][item][pixel][c]),
// slabs summed in ks order; thread = (pixel of an item, 4 output channels).  dil == 1 layers only.
__global__ void __launch_bounds__(256) wm_splitk_reduce_kernel(const WmParams p) {
  pdl_wait();
  pdl_launch_dependents();
  const int cq = p.coutp >> 2;
  const long total = (long)p.num_items * p.npx * cq;
  const size_t slab = (size_t)p.num_items * p.npx * 128;
  for (long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
    const int c4 = (int)(e % cq); const long r = e / cq;
    const int pix = (int)(r % p.npx), item = (int)(r / p.npx);
    const int n = item / p.tiles_per_img, t = item - n * p.tiles_per_img;
    const int oy = (t / p.tiles_x) * p.rows + (pix >> 3), ox = (t % p.tiles_x) * 8 + (pix & 7);
    if (oy >= p.H || ox >= p.W) continue;
    const float* src = p.scratch + ((size_t)item * p.npx + pix) * 128 + 4 * c4;
    float4 a = __ldcg(reinterpret_cast<const float4*>(src));
    for (int ks = 1; ks < p.ksplit; ++ks) {
      const float4 b = __ldcg(reinterpret_cast<const float4*>(src + ks * slab));
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    const int c = 4 * c4;
    float v[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[i] += (c + i < p.cout) ? __ldg(p.bias + c + i) : 0.f;
      v[i] = fmaxf(v[i], v[i] * p.slope);
    }
    const bool second = (n >= p.n_split) && (p.out2 != nullptr);
    const int nn = second ? n - p.n_split : n;
    uint16_t* o = second ? reinterpret_cast<uint16_t*>(p.out2) + (((long)nn * p.H + oy) * p.W + ox) * p.out2_pitch + p.out2_coff
                         : reinterpret_cast<uint16_t*>(p.out) + (((long)nn * p.H + oy) * p.W + ox) * p.out_pitch + p.out_coff;
    const uint32_t h0 = pack2_h16(v[0], v[1], p.fp16), h1 = pack2_h16(v[2], v[3], p.fp16);
    *reinterpret_cast<uint2*>(o + c) = make_uint2(h0, h1);
    if (p.out_lo_off) {
      const float2 f0 = unpack2_h16(h0, p.fp16), f1 = unpack2_h16(h1, p.fp16);
      *reinterpret_cast<uint2*>(o + c + p.out_lo_off) =
          make_uint2(pack2_h16(v[0] - f0.x, v[1] - f0.y, p.fp16), pack2_h16(v[2] - f1.x, v[3] - f1.y, p.fp16));
    }
  }
}

}  // namespace pwc
