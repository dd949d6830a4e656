// 16-bit implicit-GEMM 3x3 convolution (stride 1, dilation d <= 8) on tcgen05 with a HALO tile of the input kept in
// shared memory and reused by all 9 filter taps -- the second-generation kernel behind pwc_tc_conv_* for the decoder
// (/root/reference/models/PWCNet.py:202-264), the context network (PWCNet.py:126-132,267-268), predict_flow
// (PWCNet.py:32-33) and the stride-1 pyramid layers (PWCNet.py:52-69).
//
// Why: with one TMA box per (tap, channel chunk) the A operand is fetched 9 times from L2 and a 128x128 tile needs
// ~128 B/clk/SM of operand traffic -- more than the L2 can deliver, so the tensor pipe starves (measured: N=64/32
// layers at 120-240 TFLOP/s).  Here, per 64-channel chunk, ONE TMA box brings the (16+2)x(8h+2)-pixel halo of the
// output tile; tap (ky,kx) is the same shared-memory tile read through a UMMA descriptor whose start address is shifted
// by (ky*pitch + kx) pixel rows (128-byte rows, SWIZZLE_128B pattern anchored at the 1024-byte aligned tile base).
//   * output tile: 16 rows x 8 columns per "half" (M = 128: 8-pixel row segments are the 8-row core-matrix groups, the
//     group stride SBO is the halo row pitch), 1 or 2 halves per CTA sharing every weight tile (B traffic / 2).
//   * dilation d: the tile is a tile of the d x d sub-lattice (TMA elementStrides = d), so a dilated convolution is
//     d*d ordinary ones and keeps the halo reuse.
//   * weights [tap][chunk][CoutP][64]: streamed through a ring of stages, or kept resident in shared memory for the
//     whole kernel when they fit (the thin pyramid layers).
//   * optional split of the output channels over CTAs (small pyramid levels: more CTAs than tiles).
//   * warp roles as in conv_tc.cu: warp 0 TMA producer, warp 1 MMA issuer + TMEM owner, warps 2-5 epilogue; TMEM
//     accumulators double buffered.
#pragma once
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "tmap.cuh"
#include <cuda_fp16.h>

namespace pwc {

constexpr int kHaloTH = 16;          // output rows per tile
constexpr int kHaloThreads = 192;

struct HaloParams {
  int N, H, W, dil;
  int cout, coutp, kc, nchunk, halves, nsplit, nsl;
  int tiles_x, tiles_y, tiles_per_img, num_items;
  int pitch_px;                      // halo row pitch in pixels = 8*halves + 2
  uint32_t a_bytes_tx, a_bytes, b_bytes_tx, b_bytes, b_stages, b_resident, tmem_cols;
  int fp16, bo_mode;
  const float* bias; float slope;
  void* out; long out_pitch; int out_coff;
  void* out2; long out2_pitch; int out2_coff; int n_split;
  int out_f32;
  const float* residual; long res_pitch;
};

// Debug timeline (PWC_TC_HALO_BO bit 16): CTA 0's producer and MMA threads log (tag << 56 | clock64) events.
__device__ unsigned long long g_halo_trace[8192];
__device__ __forceinline__ void trace_ev(bool on, int& idx, int base, int tag) {
  if (on && idx < 4096) { g_halo_trace[base + idx] = ((unsigned long long)tag << 56) | (clock64() & 0x00ffffffffffffffULL); ++idx; }
}

__device__ __forceinline__ uint64_t halo_desc(uint32_t saddr, uint32_t sbo_bytes, uint32_t layout_type, int bo_mode) {
  uint64_t d = ptx::make_kmajor_desc(saddr, sbo_bytes, layout_type);
  // Measured on B200: the swizzle XOR is taken from the ABSOLUTE shared-memory address bits, so a start address that is
  // shifted by whole pixel rows inside a 1024-byte anchored pattern needs base_offset = 0 (bo_mode 1 = experiment only).
  if (bo_mode & 1) d |= (uint64_t)((saddr >> 7) & 7u) << 49;
  return d;
}

__global__ void __launch_bounds__(kHaloThreads, 1)
conv_tc_halo_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const HaloParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_addr(smem_raw) + 1023u) & ~1023u;
  const uint32_t a_base = base;                                   // 2 halo buffers
  const uint32_t b_base = base + 2u * p.a_bytes;                  // weight stages (or the resident weight set)
  const uint32_t nb = p.b_resident ? (uint32_t)(9 * p.nchunk) : p.b_stages;
  const uint32_t bar_base = b_base + nb * p.b_bytes;
  auto a_full = [&](int i) { return bar_base + 8u * i; };
  auto a_empty = [&](int i) { return bar_base + 8u * (2 + i); };
  auto t_full = [&](int i) { return bar_base + 8u * (4 + i); };
  auto t_empty = [&](int i) { return bar_base + 8u * (6 + i); };
  const uint32_t b_all = bar_base + 8u * 8;
  auto b_full = [&](int s) { return bar_base + 8u * (9 + s); };
  auto b_empty = [&](int s) { return bar_base + 8u * (9 + p.b_stages + s); };
  const uint32_t tmem_slot = bar_base + 8u * (9 + 2 * p.b_stages);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - ptx::smem_addr(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int d = p.dil;
  const int twp = 8 * p.halves;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA); ptx::prefetch_tmap(&tmB);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(a_full(i), 1); ptx::mbar_init(a_empty(i), 1);
      ptx::mbar_init(t_full(i), 1); ptx::mbar_init(t_empty(i), 4);
    }
    ptx::mbar_init(b_all, 1);
    for (uint32_t s = 0; s < p.b_stages; ++s) { ptx::mbar_init(b_full(s), 1); ptx::mbar_init(b_empty(s), 1); }
    ptx::fence_barrier_init();
    ptx::fence_proxy_async();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, p.tmem_cols);
  ptx::tc_fence_before();
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  // item -> (image n, lattice phase (ry,rx), tile origin in lattice coordinates, output-channel slice)
  auto decode = [&](int item, int& n, int& ry, int& rx, int& ty0, int& tx0, int& sl) {
    const int tile = item / p.nsplit;
    sl = item - tile * p.nsplit;
    n = tile / p.tiles_per_img;
    int r = tile - n * p.tiles_per_img;
    const int per_phase = p.tiles_x * p.tiles_y;
    const int ph = r / per_phase;
    r -= ph * per_phase;
    ry = ph / d; rx = ph - ry * d;
    ty0 = (r / p.tiles_x) * kHaloTH; tx0 = (r % p.tiles_x) * twp;
  };

  if (warp == 0) {
    // ===================== TMA producer (whole warp converged; one elected lane issues) =====================
    {
      if (p.b_resident && ptx::elect_one()) {
        ptx::mbar_expect_tx(b_all, 9u * p.nchunk * p.b_bytes_tx);
        for (int j = 0; j < 9 * p.nchunk; ++j) ptx::tma_load_2d(b_base + j * p.b_bytes, &tmB, b_all, 0, j * p.coutp);
      }
      __syncwarp();
      int ab = 0; uint32_t aph = 0; uint32_t s = 0, bph = 0;
      const bool tr = (p.bo_mode & 16) && blockIdx.x == 0; int ti = 0;
      for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
        int n, ry, rx, ty0, tx0, sl;
        decode(item, n, ry, rx, ty0, tx0, sl);
        const int ya = ry + d * (ty0 - 1), xa = rx + d * (tx0 - 1);
        for (int ch = 0; ch < p.nchunk; ++ch) {
          ptx::mbar_wait(a_empty(ab), aph ^ 1);
          if (ptx::elect_one()) {
            trace_ev(tr, ti, 4096, 10);
            if (p.bo_mode & 4) {                                   // timing experiment: no A traffic
              ptx::mbar_arrive(a_full(ab));
            } else {
              ptx::mbar_expect_tx(a_full(ab), p.a_bytes_tx);
              ptx::tma_load_4d(a_base + ab * p.a_bytes, &tmA, a_full(ab), ch * p.kc, xa, ya, n);
            }
          }
          __syncwarp();
          if (++ab == 2) { ab = 0; aph ^= 1; }
          if (!p.b_resident) {
            for (int tap = 0; tap < 9; ++tap) {
              ptx::mbar_wait(b_empty(s), bph ^ 1);
              if (ptx::elect_one()) {
                trace_ev(tr, ti, 4096, 12);
                ptx::mbar_expect_tx(b_full(s), p.b_bytes_tx);
                ptx::tma_load_2d(b_base + s * p.b_bytes, &tmB, b_full(s), 0, (tap * p.nchunk + ch) * p.coutp + sl * p.nsl);
              }
              __syncwarp();
              if (++s == p.b_stages) { s = 0; bph ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: the whole warp runs the loop converged, ONE elected lane issues =====
    // (issuing from a divergent `if (lane == 0)` makes ptxas wrap every UTCHMMA in a per-thread serialisation loop:
    //  measured ~110 cycles per instruction instead of the 16-128 the tensor pipe needs)
    {
      uint32_t idesc = ptx::make_idesc_bf16((p.bo_mode & 8) ? 64 : 128, p.nsl);   // &8: timing experiment (M=64)
      if (p.fp16) idesc &= ~((7u << 7) | (7u << 10));          // A/B format 0 = F16
      const uint32_t row_bytes = p.kc * 2;
      const uint32_t ltype = (p.kc == 64) ? 2u : (p.kc == 32) ? 4u : 6u;
      const uint32_t sbo_a = p.pitch_px * row_bytes;
      int ab = 0; uint32_t aph = 0; uint32_t s = 0, bph = 0;
      int acc = 0; uint32_t tph = 0;
      if (p.b_resident) ptx::mbar_wait(b_all, 0);
      const bool tr = (p.bo_mode & 16) && blockIdx.x == 0; int ti = 0;
      for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
        ptx::mbar_wait(t_empty(acc), tph ^ 1);
        ptx::tc_fence_after();
        if (lane == 0) trace_ev(tr, ti, 0, 1);
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.halves * p.nsl);
        for (int ch = 0; ch < p.nchunk; ++ch) {
          ptx::mbar_wait(a_full(ab), aph);
          ptx::tc_fence_after();
          if (lane == 0) trace_ev(tr, ti, 0, 2);
          const uint32_t sa = a_base + ab * p.a_bytes;
          for (int tap = 0; tap < 9; ++tap) {
            const int ky = tap / 3, kx = tap - 3 * ky;
            uint32_t sb;
            if (p.b_resident) {
              sb = b_base + (tap * p.nchunk + ch) * p.b_bytes;
            } else {
              ptx::mbar_wait(b_full(s), bph);
              ptx::tc_fence_after();
              if (lane == 0) trace_ev(tr, ti, 0, 3);
              sb = b_base + s * p.b_bytes;
            }
            const uint64_t db = ptx::make_kmajor_desc(sb, 8 * row_bytes, ltype);
            if (ptx::elect_one()) {
              for (int h = 0; h < p.halves; ++h) {
                const uint32_t sah = sa + ((p.bo_mode & 2) ? 0u : (uint32_t)(ky * p.pitch_px + kx + 8 * h) * row_bytes);
                const uint64_t da = halo_desc(sah, sbo_a, ltype, p.bo_mode);
                for (int kk = 0; kk < p.kc / 16; ++kk)
                  ptx::umma_bf16(d_tmem + h * p.nsl, da + 2 * kk, db + 2 * kk, idesc, (ch | tap | kk) ? 1u : 0u);
              }
              if (!p.b_resident) ptx::umma_commit(b_empty(s));
              if (tap == 8) ptx::umma_commit(a_empty(ab));
              if (tap == 8 && ch == p.nchunk - 1) ptx::umma_commit(t_full(acc));
              trace_ev(tr, ti, 0, 4);
            }
            __syncwarp();
            if (!p.b_resident) { if (++s == p.b_stages) { s = 0; bph ^= 1; } }
          }
          if (++ab == 2) { ab = 0; aph ^= 1; }
        }
        if (++acc == 2) { acc = 0; tph ^= 1; }
      }
    }
  } else {
    // ===================== epilogue warps 2..5 =====================
    const int lg = warp & 3;
    const int m = lg * 32 + lane;
    const int py = m >> 3, pxl = m & 7;
    int acc = 0; uint32_t tph = 0;
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
      int n, ry, rx, ty0, tx0, sl;
      decode(item, n, ry, rx, ty0, tx0, sl);
      const bool second = (n >= p.n_split) && (p.out2 != nullptr);
      const int nn = second ? n - p.n_split : n;
      const int oy = ry + d * (ty0 + py);
      ptx::mbar_wait(t_full(acc), tph);
      ptx::tc_fence_after();
      for (int h = 0; h < p.halves; ++h) {
        const int ox = rx + d * (tx0 + pxl + 8 * h);
        const bool live = (oy < p.H) && (ox < p.W);
        const long pix = ((long)nn * p.H + oy) * p.W + ox;
        const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)((acc * p.halves + h) * p.nsl);
        for (int c0 = 0; c0 < p.nsl; c0 += 16) {
          uint32_t v[16];
          ptx::tmem_ld16(taddr + c0, v);
          ptx::tmem_ld_wait();
          if (live) {
            const int cb = sl * p.nsl + c0;
            float f[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int co = cb + i;
              const float b = (co < p.cout) ? __ldg(p.bias + co) : 0.f;
              f[i] = lrelu(__uint_as_float(v[i]) + b, p.slope);
            }
            if (p.out_f32) {
              float* o = reinterpret_cast<float*>(p.out) + pix * p.out_pitch + p.out_coff;
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const int co = cb + i;
                if (co < p.cout) {
                  float val = f[i];
                  if (p.residual) val += p.residual[pix * p.res_pitch + co];
                  o[co] = val;
                }
              }
            } else {
              uint16_t* o = second ? reinterpret_cast<uint16_t*>(p.out2) + pix * p.out2_pitch + p.out2_coff
                                   : reinterpret_cast<uint16_t*>(p.out) + pix * p.out_pitch + p.out_coff;
              uint32_t w[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                if (p.fp16) {
                  __half2 hh = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
                  w[i] = *reinterpret_cast<uint32_t*>(&hh);
                } else {
                  __nv_bfloat162 hh = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
                  w[i] = *reinterpret_cast<uint32_t*>(&hh);
                }
              }
              uint4* o4 = reinterpret_cast<uint4*>(o + cb);
              o4[0] = make_uint4(w[0], w[1], w[2], w[3]);
              o4[1] = make_uint4(w[4], w[5], w[6], w[7]);
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(t_empty(acc));
      if (++acc == 2) { acc = 0; tph ^= 1; }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, p.tmem_cols);
}

}  // namespace pwc
