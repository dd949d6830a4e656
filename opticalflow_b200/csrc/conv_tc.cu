// bf16 implicit-GEMM 3x3 convolution on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM, operands staged
// by TMA), fused bias + LeakyReLU epilogue that writes straight into a channel slice of the concat buffer.
// Replaces nn.Conv2d + nn.LeakyReLU of conv() (/root/reference/models/PWCNet.py:26-30) for the decoder
// (PWCNet.py:202-264), the context network (PWCNet.py:126-132,267-268) and the stride-1 pyramid layers.
//
// GEMM view: D[128 pixels, Cout] += A[128 pixels, KC] * W[Cout, KC]^T for each (tap, KC-channel chunk).
//   * M tile = 8x16 output pixels of one image.  For tap (ky,kx) the A tile is the SAME box of the NHWC tensor shifted
//     by ((ky-1)*dil, (kx-1)*dil): one 4-D TMA box (KC, 16, 8, 1); out-of-image coordinates are zero-filled by TMA,
//     which implements the convolution padding.  The box lands in smem as 128 rows of KC bf16 = the K-major
//     SWIZZLE_{128,64,32}B layout tcgen05 consumes.
//   * weights are pre-packed [tap][chunk][CoutP][KC] bf16: one 2-D TMA box (KC, CoutP) per (tap, chunk).
//   * warp roles (192 threads): warp0 = TMA producer, warp1 = TMEM allocator + single-thread MMA issuer,
//     warps2-5 = epilogue (tcgen05.ld -> bias/LeakyReLU -> bf16/fp32 global stores).  Persistent over tiles, TMEM
//     accumulator double-buffered so the epilogue of tile i overlaps the MMAs of tile i+1.
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "tmap.cuh"
#include "conv_tc_halo.cuh"
#include "conv_tc_wm.cuh"
#include "../../include/pwc_b200.h"

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <new>

namespace pwc {

constexpr int kTcTH = 8, kTcTW = 16, kTcM = 128, kTcThreads = 192;

struct TcConvParams {
  int N, Hin, Win, Hout, Wout, stride, dil;
  int cout, coutp, kc, nchunk, n_split;
  int tiles_x, tiles_y, num_tiles, stages;
  uint32_t a_bytes, stage_bytes, tmem_cols;
  const float* bias; float slope;
  void* out; long out_pitch; int out_coff;
  void* out2; long out2_pitch; int out2_coff;
  int out_f32;
  const float* residual; long res_pitch;
  int fp16;                                       // 16-bit operands / outputs are IEEE half instead of bf16
  int split, lo_off, out_lo_off;                  // split precision: see WmParams::split (conv_tc_wm.cuh)
  uint32_t w_resident, w_tile;                    // the whole packed weight set (9 * nchunk tiles of w_tile bytes) stays in shared memory
                                                  // behind the A ring: a stage is the A box alone (streamed, the weight tile is half of
                                                  // every stage's bytes and the L2 -> SM ingest, ~30 B/cycle, is what bounds these layers)
  int xdup;                                       // split == 3: the (x_hi, W_lo) chunk reuses the A box of the (x_hi, W_hi) chunk before it
  int pairs;                                      // stride 2 without TMA element strides (which fetch the skipped pixels too: half of
                                                  // the ingest): tmA0 / tmA1 = the even / odd input rows as (C, column parity, W/2,
                                                  // rows, N); tap (ky, kx) = a dense box of map (ky != 1), parity (kx != 1), shifted by
                                                  // -1 pair / row for kx == 0 / ky == 0
  int fused;                                      // split input whose planes are adjacent and fit one swizzle row ([hi | lo] <= 64
                                                  // channels): kc counts BOTH planes, one stage = the A box + two weight tiles
                                                  // [W_hi | W_hi] and [W_lo | 0]; the second runs on the hi half only
};

// 16 output channels of one pixel: 16-bit rounding and 2 x 16-byte stores; LO: the rounding residuals go to the lo plane
template <bool F16>
__device__ __forceinline__ void tc_store16(const float* f, uint16_t* o, int lo_off) {
  uint32_t w[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) w[i] = pack2_h16(f[2 * i], f[2 * i + 1], F16 ? 1 : 0);
  uint4* o4 = reinterpret_cast<uint4*>(o);
  o4[0] = make_uint4(w[0], w[1], w[2], w[3]);
  o4[1] = make_uint4(w[4], w[5], w[6], w[7]);
  if (lo_off) {                                   // lo plane: the exact remainders of the hi roundings
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float2 hv = unpack2_h16(w[i], F16 ? 1 : 0);
      w[i] = pack2_h16(f[2 * i] - hv.x, f[2 * i + 1] - hv.y, F16 ? 1 : 0);
    }
    uint4* l4 = reinterpret_cast<uint4*>(o + lo_off);
    l4[0] = make_uint4(w[0], w[1], w[2], w[3]);
    l4[1] = make_uint4(w[4], w[5], w[6], w[7]);
  }
}

__global__ void __launch_bounds__(kTcThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmB, const TcConvParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-B aligned base (SWIZZLE_128B atoms)
  const uint32_t base = (ptx::smem_addr(smem_raw) + 1023u) & ~1023u;
  const uint32_t w_base = base + p.stages * p.stage_bytes;           // resident weight set (w_resident)
  const uint32_t bar_base = w_base + (p.w_resident ? 9u * p.nchunk * p.w_tile : 0u);
  const uint32_t w_all = bar_base + 8u * (2 * p.stages + 6);
  // bias, zero padded to coutp, in shared memory: the epilogue thread (= one pixel) needs every channel's bias, and a global
  // load per value put an L2 round trip on the one epilogue warp of each scheduler (ncu: the top stall of the stride-2 layers)
  float* sbias = reinterpret_cast<float*>(smem_raw + (bar_base + 8u * (2 * p.stages + 8) - ptx::smem_addr(smem_raw)));
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (p.stages + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * p.stages + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * p.stages + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * p.stages + 4);
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - ptx::smem_addr(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ksteps = p.fused ? 9 : 9 * p.nchunk;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA0); ptx::prefetch_tmap(&tmA1); ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < p.stages; ++s) { ptx::mbar_init(full_bar(s), 1); ptx::mbar_init(empty_bar(s), 1); }
    for (int a = 0; a < 2; ++a) { ptx::mbar_init(tfull_bar(a), 1); ptx::mbar_init(tempty_bar(a), 4); }
    ptx::mbar_init(w_all, 1);
    ptx::fence_barrier_init();
    ptx::fence_proxy_async();
    if (p.w_resident) {                                // weights do not depend on the previous kernel: before the PDL wait
      ptx::mbar_expect_tx(w_all, 9u * p.nchunk * (uint32_t)p.coutp * p.kc * 2u);
      for (int j = 0; j < 9 * p.nchunk; ++j) ptx::tma_load_2d(w_base + j * p.w_tile, &tmB, w_all, 0, j * p.coutp);
    }
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, p.tmem_cols);
  ptx::tc_fence_before();
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ===================== TMA producer (whole warp converged; one elected lane issues) =====================
    {
      int s = 0; uint32_t ph = 0;
      for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x) {
        const int n = t / (p.tiles_x * p.tiles_y);
        const int r = t - n * (p.tiles_x * p.tiles_y);
        const int ty0 = (r / p.tiles_x) * kTcTH, tx0 = (r % p.tiles_x) * kTcTW;
        const CUtensorMap* tmA = (n >= p.n_split) ? &tmA1 : &tmA0;
        const int nn = (n >= p.n_split) ? n - p.n_split : n;
        for (int tap = 0; tap < 9; ++tap) {
          const int ky = tap / 3, kx = tap - 3 * ky;
          const int y = ty0 * p.stride + (ky - 1) * p.dil, x = tx0 * p.stride + (kx - 1) * p.dil;
          if (p.fused) {
            ptx::mbar_wait(empty_bar(s), ph ^ 1);
            if (ptx::elect_one()) {
              const uint32_t sa = base + s * p.stage_bytes, bb = (uint32_t)p.coutp * p.kc * 2;
              ptx::mbar_expect_tx(full_bar(s), p.a_bytes + (p.w_resident ? 0u : 2 * bb));
              if (p.pairs) ptx::tma_load_5d(sa, ky == 1 ? &tmA0 : &tmA1, full_bar(s), 0, kx != 1, tx0 - (kx == 0), ty0 - (ky == 0), nn);
              else ptx::tma_load_4d(sa, tmA, full_bar(s), 0, x, y, nn);
              if (!p.w_resident) {
                ptx::tma_load_2d(sa + p.a_bytes, &tmB, full_bar(s), 0, (tap * 2) * p.coutp);
                ptx::tma_load_2d(sa + p.a_bytes + bb, &tmB, full_bar(s), 0, (tap * 2 + 1) * p.coutp);
              }
            }
            __syncwarp();
            if (++s == p.stages) { s = 0; ph ^= 1; }
            continue;
          }
          for (int ch = 0; ch < p.nchunk; ++ch) {
            if (p.xdup && ch % 3 == 1) continue;       // resident weights: the MMA warp reuses the stage of chunk ch - 1
            ptx::mbar_wait(empty_bar(s), ph ^ 1);
            if (ptx::elect_one()) {
              const uint32_t sa = base + s * p.stage_bytes;
              ptx::mbar_expect_tx(full_bar(s), p.a_bytes + (p.w_resident ? 0u : (uint32_t)p.coutp * p.kc * 2));
              const int xc = p.split ? (ch / p.split) * p.kc + ((ch % p.split) == p.split - 1 ? p.lo_off : 0) : ch * p.kc;
              if (p.pairs) ptx::tma_load_5d(sa, ky == 1 ? &tmA0 : &tmA1, full_bar(s), xc, kx != 1, tx0 - (kx == 0), ty0 - (ky == 0), nn);
              else ptx::tma_load_4d(sa, tmA, full_bar(s), xc, x, y, nn);
              if (!p.w_resident) ptx::tma_load_2d(sa + p.a_bytes, &tmB, full_bar(s), 0, (tap * p.nchunk + ch) * p.coutp);
            }
            __syncwarp();
            if (++s == p.stages) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: whole warp converged, ONE elected lane issues (see conv_tc_halo.cuh) =====
    {
      uint32_t idesc = ptx::make_idesc_bf16(kTcM, p.coutp);
      if (p.fp16) idesc &= ~((7u << 7) | (7u << 10));          // A/B format 0 = F16
      const uint32_t row_bytes = p.kc * 2;
      const uint32_t ltype = (p.kc == 64) ? 2u : (p.kc == 32) ? 4u : 6u;
      int s = 0; uint32_t ph = 0;
      int acc = 0; uint32_t aph = 0;
      if (p.w_resident) { ptx::mbar_wait(w_all, 0); ptx::tc_fence_after(); }
      for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x) {
        ptx::mbar_wait(tempty_bar(acc), aph ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.coutp);
        for (int k = 0; k < ksteps; ++k) {
          // k = tap * nchunk + ch (fused: k = tap).  xdup: chunk ch % 3 == 1 multiplies the stage chunk ch - 1 left in place
          const int ch = p.fused ? 0 : k % p.nchunk;
          const bool reuse = p.xdup && ch % 3 == 1, keep = p.xdup && ch % 3 == 0;
          if (!reuse) {
            ptx::mbar_wait(full_bar(s), ph);
            ptx::tc_fence_after();
          }
          if (ptx::elect_one()) {
            const uint32_t sa = base + s * p.stage_bytes;
            const uint32_t sb = p.w_resident ? w_base + (uint32_t)(p.fused ? 2 * k : k) * p.w_tile : sa + p.a_bytes;
            const uint64_t da = ptx::make_kmajor_desc(sa, 8 * row_bytes, ltype);
            const uint64_t db = ptx::make_kmajor_desc(sb, 8 * row_bytes, ltype);
            for (int kk = 0; kk < p.kc / 16; ++kk)
              ptx::umma_bf16(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (k | kk) ? 1u : 0u);
            if (p.fused) {                            // x_hi * W_lo: the hi half of the box against the second weight tile
              const uint64_t db2 = ptx::make_kmajor_desc(sb + (p.w_resident ? p.w_tile : (uint32_t)p.coutp * row_bytes), 8 * row_bytes, ltype);
              for (int kk = 0; kk < p.kc / 32; ++kk) ptx::umma_bf16(d_tmem, da + 2 * kk, db2 + 2 * kk, idesc, 1u);
            }
            if (!keep) ptx::umma_commit(empty_bar(s));          // frees the smem stage when these MMAs retire
            if (k == ksteps - 1) ptx::umma_commit(tfull_bar(acc));   // accumulator complete -> epilogue
          }
          __syncwarp();
          if (!keep && ++s == p.stages) { s = 0; ph ^= 1; }
        }
        if (++acc == 2) { acc = 0; aph ^= 1; }
      }
    }
  } else {
    // ===================== epilogue warps 2..5 =====================
    const int lg = warp & 3;                       // TMEM lane group this warp may access
    const int m = lg * 32 + lane;                  // pixel within the tile
    const int py = m >> 4, px = m & 15;
    int acc = 0; uint32_t aph = 0;
    for (int i = threadIdx.x - 64; i < p.coutp; i += 128) sbias[i] = (i < p.cout) ? __ldg(p.bias + i) : 0.f;
    asm volatile("bar.sync 1, 128;" ::: "memory");
    for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x) {
      const int n = t / (p.tiles_x * p.tiles_y);
      const int r = t - n * (p.tiles_x * p.tiles_y);
      const int oy = (r / p.tiles_x) * kTcTH + py, ox = (r % p.tiles_x) * kTcTW + px;
      const bool live = (oy < p.Hout) && (ox < p.Wout);
      const bool second = (n >= p.n_split) && (p.out2 != nullptr);
      const int nn = second ? n - p.n_split : n;
      const long pix = ((long)nn * p.Hout + oy) * p.Wout + ox;
      ptx::mbar_wait(tfull_bar(acc), aph);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(acc * p.coutp);
      for (int c0 = 0; c0 < p.coutp; c0 += 16) {
        uint32_t v[16];
        ptx::tmem_ld16(taddr + c0, v);
        ptx::tmem_ld_wait();
        if (live) {
          float f[16];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 b4 = *reinterpret_cast<const float4*>(sbias + c0 + 4 * q);
            f[4 * q + 0] = lrelu(__uint_as_float(v[4 * q + 0]) + b4.x, p.slope);
            f[4 * q + 1] = lrelu(__uint_as_float(v[4 * q + 1]) + b4.y, p.slope);
            f[4 * q + 2] = lrelu(__uint_as_float(v[4 * q + 2]) + b4.z, p.slope);
            f[4 * q + 3] = lrelu(__uint_as_float(v[4 * q + 3]) + b4.w, p.slope);
          }
          if (p.out_f32) {
            float* o = reinterpret_cast<float*>(p.out) + pix * p.out_pitch + p.out_coff;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int co = c0 + i;
              if (co < p.cout) {
                float val = f[i];
                if (p.residual) val += p.residual[pix * p.res_pitch + co];
                o[co] = val;
              }
            }
          } else {
            uint16_t* o = second ? reinterpret_cast<uint16_t*>(p.out2) + pix * p.out2_pitch + p.out2_coff
                                 : reinterpret_cast<uint16_t*>(p.out) + pix * p.out_pitch + p.out_coff;
            if (p.fp16) tc_store16<true>(f, o + c0, p.out_lo_off);
            else tc_store16<false>(f, o + c0, p.out_lo_off);
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(tempty_bar(acc));
      if (++acc == 2) { acc = 0; aph ^= 1; }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, p.tmem_cols);
}

}  // namespace pwc

// ------------------------------------------------------------------------------------------------------------------
// host side: descriptor object (tensor maps + parameters), launch, weight packing
// ------------------------------------------------------------------------------------------------------------------
struct pwc_tc_conv {
  CUtensorMap tmA0, tmA1, tmB, tmO, tmO2;
  pwc::TcConvParams p;
  pwc::HaloParams hp;
  pwc::WmParams wp;
  int halo;                 // 2: conv_tc_wm_kernel, 1: conv_tc_halo_kernel, 0: conv_tc_kernel (per-tap boxes)
  size_t smem_bytes;
  int grid;
  void* scratch = nullptr;  // split-K partial sums + tile counters (WM kernel, small layers); owned by the handle
  ~pwc_tc_conv() { if (scratch) cudaFree(scratch); }
};

namespace pwc {

__global__ void pack_conv_weight_bf16_kernel(const float* __restrict__ wt, __nv_bfloat16* __restrict__ wp,
                                             const int* __restrict__ map, int Cout, int Cin, int nchunk, int kc, int coutp,
                                             int f16) {
  // wt: [Cout][Cin][3][3] fp32 -> wp: [9][nchunk][coutp][kc] bf16 (pre-zeroed)
  const long total = (long)Cout * Cin * 9;
  for (long e = blockIdx.x * (long)blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
    const int tap = e % 9; const long r = e / 9;
    const int ci = r % Cin, co = r / Cin;
    const int row = map ? map[ci] : ci;
    const int ch = row / kc, kk = row - ch * kc;
    reinterpret_cast<uint16_t*>(wp)[(((long)tap * nchunk + ch) * coutp + co) * kc + kk] = pack1_h16(wt[e], f16);
  }
}

static int g_num_sms = 0;
static thread_local int g_half_format = 0;        // 0 = bf16, 1 = IEEE half (pwc_set_half_format)
// split-precision options of the create call in progress (pwc_tc_conv_create_split sets them around pwc_tc_conv_create)
struct SplitOpts { int in_lo_off, out_lo_off, up_lo_off, up_dup_off, products; };
static thread_local SplitOpts g_split = {0, 0, 0, 0, 3};
int half_format() { return g_half_format; }

static int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}

// Halo kernel: stride 1, dilation <= 8 (TMA elementStrides limit), one input tensor.  PWC_TC_HALO=0 forces the
// per-tap kernel (debugging / A-B comparisons); PWC_TC_HALO_MINKC (default 16) is the smallest channel chunk it takes.
static bool halo_eligible(int stride, int dil, bool has_in2, int kc, int max_dil = 8) {
  if (!env_int("PWC_TC_HALO", 1)) return false;
  return stride == 1 && dil >= 1 && dil <= max_dil && !has_in2 && (kc == 16 || kc == 32 || kc == 64) &&
         kc >= env_int("PWC_TC_HALO_MINKC", 16);
}

static int create_halo(pwc_tc_conv** handle, const void* in, long in_pitch, int in_coff, int cin_span, int kc,
                       const void* w_packed, const float* bias, int cout, int coutp,
                       void* out, long out_pitch, int out_coff, void* out2, long out2_pitch, int out2_coff, int n_split,
                       int out_f32, const float* residual, long res_pitch, int N, int H, int W, int dil, float slope) {
  if (coutp % 16 || coutp < 16 || coutp > 256 || cout > coutp || (in_pitch & 7) || (in_coff & 7) ||
      (!out_f32 && ((out_pitch & 7) || (out_coff & 7) || (out2 && ((out2_pitch & 7) || (out2_coff & 7))))))
    return (int)cudaErrorInvalidValue;
  if (g_split.in_lo_off || g_split.out_lo_off) return (int)cudaErrorInvalidValue;   // no split-precision form of this kernel
  pwc_tc_conv* h = new (std::nothrow) pwc_tc_conv();
  if (!h) return (int)cudaErrorMemoryAllocation;
  h->halo = 1;
  HaloParams& p = h->hp;
  p.N = N; p.H = H; p.W = W; p.dil = dil;
  p.cout = cout; p.coutp = coutp; p.kc = kc; p.nchunk = cdiv(cin_span, kc);
  const int Hd = cdiv(H, dil), Wd = cdiv(W, dil);
  auto ntiles = [&](int halves) { return (long)N * dil * dil * cdiv(Hd, kHaloTH) * cdiv(Wd, 8 * halves); };
  int halves = env_int("PWC_TC_HALO_HALVES", 0);
  if (halves != 1 && halves != 2) halves = (ntiles(2) >= (long)g_num_sms * 3 / 2 && 2 * coutp <= 256) ? 2 : 1;
  if (2 * halves * coutp > 512) halves = 1;
  p.halves = halves;
  const long tiles = ntiles(halves);
  int nsplit = 1;
  if (tiles < g_num_sms) {                      // small pyramid levels: also split the output channels over CTAs
    for (int cand = 2; cand <= 8; cand *= 2)
      if (coutp % (16 * cand) == 0 && coutp / cand >= 32 && tiles * nsplit < g_num_sms) nsplit = cand;
  }
  const int forced = env_int("PWC_TC_HALO_NSPLIT", 0);
  if (forced > 0 && coutp % (16 * forced) == 0) nsplit = forced;
  p.nsplit = nsplit; p.nsl = coutp / nsplit;
  p.tiles_x = cdiv(Wd, 8 * halves); p.tiles_y = cdiv(Hd, kHaloTH);
  p.tiles_per_img = dil * dil * p.tiles_x * p.tiles_y;
  p.num_items = (int)(tiles * nsplit);
  p.pitch_px = 8 * halves + 2;
  const uint32_t row_bytes = (uint32_t)kc * 2;
  p.a_bytes_tx = (uint32_t)(kHaloTH + 2) * p.pitch_px * row_bytes;
  p.a_bytes = (p.a_bytes_tx + 1023u) & ~1023u;
  p.b_bytes_tx = (uint32_t)p.nsl * row_bytes;
  p.b_bytes = (p.b_bytes_tx + 1023u) & ~1023u;
  const uint32_t budget = 216u * 1024u - 2u * p.a_bytes;
  const uint32_t all_b = 9u * p.nchunk * p.b_bytes;
  p.b_resident = (nsplit == 1 && all_b <= budget && env_int("PWC_TC_HALO_RESIDENT", 1)) ? 1u : 0u;
  uint32_t stages = budget / p.b_bytes;
  p.b_stages = stages > 12 ? 12 : (stages < 2 ? 2 : stages);
  uint32_t cols = 32; while (cols < 2u * halves * p.nsl) cols <<= 1;
  p.tmem_cols = cols;
  p.fp16 = g_half_format; p.bo_mode = env_int("PWC_TC_HALO_BO", 0);
  p.bias = bias; p.slope = slope;
  p.out = out; p.out_pitch = out_pitch; p.out_coff = out_coff;
  p.out2 = out2; p.out2_pitch = out2_pitch; p.out2_coff = out2_coff; p.n_split = out2 ? n_split : N;
  p.out_f32 = out_f32; p.residual = residual; p.res_pitch = res_pitch;
  const uint32_t nb = p.b_resident ? 9u * p.nchunk : p.b_stages;
  h->smem_bytes = 2u * p.a_bytes + (size_t)nb * p.b_bytes + 8 * (10 + 2 * p.b_stages) + 16 + 1024;
  const CUtensorMapSwizzle swz = swizzle_for_bytes(kc * 2);
  const __nv_bfloat16* a0 = reinterpret_cast<const __nv_bfloat16*>(in) + in_coff;
  int e = make_tmap_nhwc(&h->tmA0, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a0, in_pitch, cin_span, W, H, N,
                         kc, p.pitch_px * dil, (kHaloTH + 2) * dil, dil, dil, swz);
  if (!e) e = make_tmap_2d(&h->tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w_packed, (long)9 * p.nchunk * coutp, kc, kc, p.nsl, swz);
  if (e || h->smem_bytes > 227u * 1024u) { delete h; return e ? e : (int)cudaErrorInvalidValue; }
  h->tmA1 = h->tmA0;
  h->grid = p.num_items < g_num_sms ? p.num_items : g_num_sms;
  *handle = h;
  return 0;
}

// NHWC slice with W split as (W/d, d): 5-D (C, d, W/d, H, N); box (boxC, 1, boxXg, boxH, 1) = one x phase of a tile.
// For dilations above the TMA element-stride limit of 8; needs W % d == 0.
static int make_tmap_xsplit(CUtensorMap* m, const void* base, long pitch, int C, int W, int H, int N, int d,
                            int boxC, int boxXg, int boxH, CUtensorMapSwizzle swz) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return (int)cudaErrorNotSupported;
  cuuint64_t dims[5] = {(cuuint64_t)C, (cuuint64_t)d, (cuuint64_t)(W / d), (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[4] = {(cuuint64_t)pitch * 2, (cuuint64_t)pitch * 2 * d, (cuuint64_t)pitch * 2 * W,
                           (cuuint64_t)pitch * 2 * W * H};
  cuuint32_t box[5] = {(cuuint32_t)boxC, 1, (cuuint32_t)boxXg, (cuuint32_t)boxH, 1};
  cuuint32_t es[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)cudaErrorInvalidValue;
}

// Rows of one parity of an NHWC slice, W split as (W/2, 2): 5-D (C, column parity, W/2, rows, N) with a two-row row stride;
// `base` points at the first row of that parity, img_elems = elements per image.  Box (boxC, 1, boxPairs, boxH, 1), dense.
static int make_tmap_pairs_rows(CUtensorMap* m, const void* base, long pitch, int C, int W, int rows, int N, long img_elems,
                                int boxC, int boxPairs, int boxH, CUtensorMapSwizzle swz) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return (int)cudaErrorNotSupported;
  cuuint64_t dims[5] = {(cuuint64_t)C, 2, (cuuint64_t)(W / 2), (cuuint64_t)rows, (cuuint64_t)N};
  cuuint64_t strides[4] = {(cuuint64_t)pitch * 2, (cuuint64_t)pitch * 4, (cuuint64_t)pitch * 2 * W * 2, (cuuint64_t)img_elems * 2};
  cuuint32_t box[5] = {(cuuint32_t)boxC, 1, (cuuint32_t)boxPairs, (cuuint32_t)boxH, 1};
  cuuint32_t es[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)cudaErrorInvalidValue;
}

// NHWC slice with W split as (W/2, 2): 5-D (C, column parity, W/2, H, N); box (boxC, 1, boxPairs, boxH, 1) = the columns of one
// parity of a stride-2 layer's input tile, dense.  Needs W % 2 == 0.
static int make_tmap_pairs(CUtensorMap* m, const void* base, long pitch, int C, int W, int H, int N,
                           int boxC, int boxPairs, int boxH, CUtensorMapSwizzle swz) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return (int)cudaErrorNotSupported;
  cuuint64_t dims[5] = {(cuuint64_t)C, 2, (cuuint64_t)(W / 2), (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[4] = {(cuuint64_t)pitch * 2, (cuuint64_t)pitch * 4, (cuuint64_t)pitch * 2 * W,
                           (cuuint64_t)pitch * 2 * W * H};
  cuuint32_t box[5] = {(cuuint32_t)boxC, 1, (cuuint32_t)boxPairs, (cuuint32_t)boxH, 1};
  cuuint32_t es[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)cudaErrorInvalidValue;
}

static int create_wm(pwc_tc_conv** handle, const void* in, long in_pitch, int in_coff, int cin_span, int kc,
                     const void* w_packed, const float* bias, int cout, int coutp,
                     void* out, long out_pitch, int out_coff, void* out2, long out2_pitch, int out2_coff, int n_split,
                     int out_f32, const float* residual, long res_pitch, int N, int H, int W, int dil, float slope, int stride = 1) {
  if (coutp % 16 || coutp < 16 || coutp > 128 || cout > coutp || (in_pitch & 7) || (in_coff & 7))
    return (int)cudaErrorInvalidValue;
  // stride 2 (WmParams::s2): from here on H, W are the OUTPUT dims (tiles, stores); Hin, Win describe the input map
  const int s2 = stride == 2 ? 1 : 0, Hin = H, Win = W;
  if (stride != 1 && stride != 2) return (int)cudaErrorInvalidValue;
  if (s2) {
    if (dil != 1 || (W & 1) || out_f32 || (kc != 16 && kc != 32 && kc != 64) || cin_span % kc) return (int)cudaErrorInvalidValue;
    H = (H + 1) / 2; W = W / 2;
  }
  const int ts = (out_f32 >> 4) & 1;              // tap-stacked weights ([chunk][tap*cout + co][64] in the tap-0 section)
  const int tlf = (out_f32 >> 5) & 1;             // tap-lane weights (WmParams::tl): cout <= 64 real channels, coutp = 128 rows
  out_f32 &= 3;
  if (tlf && (ts || dil != 1 || cout > 64 || coutp != 128 || out_f32 || kc != 64 || out2 || s2)) return (int)cudaErrorInvalidValue;
  if (ts && (dil != 1 || 9 * cout > 128 || coutp != 128 || !out_f32 || kc != 64 || s2)) return (int)cudaErrorInvalidValue;
  pwc_tc_conv* h = new (std::nothrow) pwc_tc_conv();
  if (!h) return (int)cudaErrorMemoryAllocation;
  h->halo = 2;
  WmParams& p = h->wp;
  p.ts = ts; p.ntaps = ts ? 1 : 9; p.s_stride = 0; p.s_bytes = 0;
  p.tl = tlf ? (cout <= 32 ? 3 : 2) : 0;
  if (p.tl) p.ntaps = p.tl == 3 ? 3 : 6;
  p.tl_bc = cout <= 16 ? 16 : 32;
  p.has_skip = 0; p.out_groups = 1; p.has_kk = 0; p.kk_first_tap = 0; p.kk_first_kk = 0;
  for (int i = 0; i < 9; ++i) p.kk_mask[i] = 0;
  for (int i = 0; i < 9; ++i) p.skip[i] = 0ull;
  p.N = N; p.H = H; p.W = W; p.dil = dil;
  p.split = g_split.in_lo_off ? g_split.products : 0; p.lo_off = g_split.in_lo_off; p.out_lo_off = g_split.out_lo_off;
  p.up_lo_off = g_split.up_lo_off; p.up_dup_off = g_split.up_dup_off;
  int vmul = p.split ? p.split : 1;                 // virtual chunks per logical chunk (see WmParams::split)
  p.s2 = s2; p.x_sub = 0;
  p.nhead = 0; p.head_out = nullptr; p.partial = nullptr; p.partial_pitch = 0;
  // adjacent [hi | lo] planes that fit one swizzle row (stride-2 pyramid layers of 16 / 32 channels): one chunk of both planes,
  // weights [W_hi | W_hi] then [W_lo | 0] (the layout of TcConvParams::fused)
  p.fused = (s2 && p.split == 3 && p.lo_off == cin_span && cin_span == kc && kc <= 32 && env_int("PWC_TC_FUSED_SPLIT", 1)) ? 1 : 0;
  if (p.fused) { kc *= 2; cin_span = kc; p.split = 0; p.lo_off = 0; vmul = 2; }
  p.xdup = (s2 && (p.fused || p.split == 3) && env_int("PWC_TC_WM_XDUP", 1)) ? 1 : 0;   // (stride 1: measured no gain)
  p.cout = cout; p.coutp = coutp; p.kc = kc; p.nchunk = cdiv(cin_span, kc) * vmul;
  if ((p.split || p.out_lo_off) && (dil != 1 || (kc != 64 && !s2) || (p.lo_off & 7) || (p.out_lo_off & 7))) { delete h; return (int)cudaErrorInvalidValue; }
  // large dilations: phases in x only, the y dilation is a row shift inside a (rows + 2*dil)-row halo tile -- 16 x fewer,
  // 16 x larger items than d x d sub-lattices of a 7 x 16 image (dc_conv5: N = 56 pixels per MMA otherwise)
  const int ydense = (dil >= env_int("PWC_TC_WM_YDENSE_MINDIL", 8) && !ts) ? 1 : 0;
  p.dy = ydense ? 1 : dil; p.kyr = ydense ? dil : 1;
  p.xsplit = dil > 8 ? 1 : 0;
  if (p.xsplit && (!ydense || W % dil || out2 || out_f32)) { delete h; return (int)cudaErrorInvalidValue; }
  const int Hd = ydense ? H : cdiv(H, dil), Wd = cdiv(W, dil);
  const int nph = ydense ? dil : dil * dil;
  int max_rows = (256 / dil - 2) < 32 ? ((256 / dil - 2) & ~1) : 32;
  if (ydense) { max_rows = 50 - 2 * dil; if (max_rows > 32) max_rows = 32; max_rows &= ~1; }
  auto s2_sub = [&](int r) { return ((uint32_t)(2 * r + 1) * 9u * (uint32_t)kc * 2u + 1023u) & ~1023u; };   // one parity's sub-tile
  if (s2) while (max_rows > 2 && 4u * s2_sub(max_rows) > 116u * 1024u) max_rows -= 2;
  auto even_up = [](int v) { return (v + 1) & ~1; };
  auto ntiles = [&](int rows) { return (long)N * nph * cdiv(Hd, rows) * cdiv(Wd, 8); };
  int rows = even_up(cdiv(Hd, cdiv(Hd, max_rows)));           // fewest row tiles, least padding
  if (rows > max_rows) rows = max_rows;
  if (rows < 2) rows = 2;
  if (env_int("PWC_TC_WM_ROWMODEL", 1)) {
    // cost model from tools/micro/mma_issue.cu: one M=128,K=16 MMA takes max(68, N/2) cycles, i.e. tiles below N = 128
    // pixels are free to grow; time ~ waves x (MMAs per item x cycles per MMA + fixed per-item cost)
    const long nmma = (long)cdiv(cin_span, kc) * 36 * vmul;
    long best = -1;
    for (int r = 2; r <= max_rows; r += 2) {
      const long waves = (ntiles(r) + g_num_sms - 1) / g_num_sms;
      const long cyc = 4L * r > 68 ? 4L * r : 68;
      const long cost = waves * (nmma * cyc + 2500);
      if (best < 0 || cost < best || (cost == best && cdiv(Hd, r) * r <= cdiv(Hd, rows) * rows)) { best = cost; rows = r; }
    }
  } else {
    while (ntiles(rows) < g_num_sms && rows > 8) rows = even_up(cdiv(rows, 2)) < 8 ? 8 : even_up(cdiv(rows, 2));
  }
  // thin single-chunk layers are epilogue-bound: a smaller tile leaves room for the replicated resident weight set
  if (cdiv(cin_span, kc) * vmul == 1 && coutp <= 64 && rows > 16 && !out_f32) rows = 16;
  const int forced = env_int("PWC_TC_WM_ROWS", 0);
  if (forced >= 2 && forced <= max_rows && !(forced & 1)) rows = forced;
  if (ts) {                                       // N = whole halo tile (rows+2) x 10 <= 256 pixels
    rows = cdiv(H, cdiv(H, 23));
    while (ntiles(rows) < g_num_sms && rows > 6) rows = cdiv(rows, 2);
  }
  if (p.tl) {
    // rows: a multiple of the lane count (equal pixel blocks per epilogue phase), even, N = 8 * (rows + 2) <= 256; fewest
    // waves x (MMA time per item + fixed cost), then least padding
    long best = -1; rows = 0;
    for (int r = 2; r <= 30; r += 2) {
      if (r % p.tl) continue;
      const long waves = (ntiles(r) + g_num_sms - 1) / g_num_sms;
      const long cost = waves * ((long)cdiv(cin_span, kc) * vmul * p.ntaps * 4 * (4L * (r + 2) > 68 ? 4L * (r + 2) : 68) + 2500);
      if (best < 0 || cost < best || (cost == best && cdiv(Hd, r) * r <= cdiv(Hd, rows) * rows)) { best = cost; rows = r; }
    }
    const int forced_tl = env_int("PWC_TC_TL_ROWS", 0);
    if (forced_tl >= 2 && forced_tl <= 30 && !(forced_tl & 1) && forced_tl % p.tl == 0) rows = forced_tl;
  }
  p.rows = rows; p.npx = ts ? (((rows + 2) * 10 + 15) & ~15) : p.tl ? 8 * (rows + 2) : 8 * rows;
  p.tiles_x = cdiv(Wd, 8); p.tiles_y = cdiv(Hd, rows);
  p.tiles_per_img = nph * p.tiles_x * p.tiles_y;
  p.num_items = (int)ntiles(rows);
  const uint32_t row_bytes = (uint32_t)kc * 2;
  // two-lane tap stacking: the single-lane ky = 2 slots read halo rows up to rows + 3; the second lane multiplies them by zero
  // weights, so they must be finite: the box carries two more (real or zero-filled) rows
  const int xextra = p.tl == 2 ? 2 : 0;
  p.x_bytes_tx = (uint32_t)(rows + 2 * p.kyr + xextra) * 10u * row_bytes;
  p.x_bytes = (p.x_bytes_tx + 1023u) & ~1023u;
  if (s2) { p.x_sub = s2_sub(rows); p.x_bytes_tx = 2u * (uint32_t)(2 * rows + 1) * 9u * row_bytes; p.x_bytes = 2u * p.x_sub; }
  if (ts) {
    const uint32_t need = (uint32_t)p.npx * row_bytes;        // the MMA reads npx pixel rows (tail rows: don't-care columns)
    if (p.x_bytes < ((need + 1023u) & ~1023u)) p.x_bytes = (need + 1023u) & ~1023u;
    p.s_stride = p.npx + 4;                                   // words; == 4 (mod 32) when npx % 32 == 0, else still 16B aligned
    p.s_bytes = ((uint32_t)(9 * cout) * p.s_stride * 4u + 1023u) & ~1023u;
  }
  // small pixel tiles consume a halo tile faster than one TMA round trip: deeper pipeline there
  p.x_stages = 2;
  if (!ts && p.npx <= 128 && cdiv(cin_span, kc) * vmul >= 3) {
    p.x_stages = env_int("PWC_TC_WM_XSTAGES", 2);   // measured: 4 stages gain nothing (3374 vs 3400 pairs/s), default stays 2
    if (p.x_stages < 2) p.x_stages = 2;
    if (p.x_stages > 4) p.x_stages = 4;
    while (p.x_stages > 2 && (uint32_t)p.x_stages * p.x_bytes > 64u * 1024u) --p.x_stages;
  }
  p.cin_span = cin_span;
  p.w_bytes_tx = (uint32_t)coutp * row_bytes;
  p.tma_store = (!out_f32 && !ts && env_int("PWC_TC_WM_TMASTORE", 1)) ? 1 : 0;
  p.o_half = p.tma_store ? (((uint32_t)(rows / 2) * 8u * coutp * 2u + 1023u) & ~1023u) : 0u;
  // second staging buffer (store of one half overlaps the fill of the next) while it leaves >= 4 weight stages
  p.o_nbuf = (p.tma_store && env_int("PWC_TC_WM_OBUF2", 1) &&
              (uint32_t)p.x_stages * p.x_bytes + 2u * p.o_half + 4u * 128u * row_bytes <= 200u * 1024u) ? 2 : 1;
  p.o_bytes = p.o_half * (uint32_t)p.o_nbuf;
  // weight-tile replication on M (epilogue parallelism for thin layers): only when the whole weight set stays resident.
  // per-warp output stores (no CTA-wide barrier in the epilogue) need 32-channel groups and rows divisible by the parts;
  // their staging size depends on the replication factor, so both are chosen together.
  p.cg = cdiv(out_f32 ? cout : coutp, 32); p.rep = 1;
  const uint32_t full_tile = 128u * row_bytes;
  const bool ws_ok = p.tma_store && coutp % 32 == 0 && slope >= 0.f && slope <= 1.f && env_int("PWC_TC_WM_WSTORE", 1);
  const bool rep_ok = p.tma_store && env_int("PWC_TC_WM_REP", 1) && env_int("PWC_TC_HALO_RESIDENT", 1);
  p.wstore = 0; p.o_rs = 0;
  for (int rep = rep_ok ? 4 / p.cg : 1; rep >= 1; rep >>= 1) {
    if (rep < 1) break;
    if (ws_ok && rows % rep) continue;
    uint32_t o_bytes = p.o_bytes, o_half = p.o_half; int rs = 0;
    bool fits = (rep == 1);
    if (ws_ok) {
      const int R = rows / rep;
      for (int cap = 8; cap >= 2 && !fits; cap >>= 1) {       // smaller staging steps if that lets the replicas stay resident
        rs = R < cap ? R : cap;
        while (R % rs) --rs;
        o_half = (uint32_t)rs * 8u * 64u;                    // one staging buffer of one warp
        o_bytes = 4u * 2u * o_half;
        fits = 9u * p.nchunk * full_tile <= 216u * 1024u - (uint32_t)p.x_stages * p.x_bytes - o_bytes;
      }
      if (rep == 1) {
        rs = R < 8 ? R : 8;
        while (R % rs) --rs;
        o_half = (uint32_t)rs * 8u * 64u; o_bytes = 4u * 2u * o_half;
      }
    } else if (rep > 1) {
      fits = 9u * p.nchunk * full_tile <= 216u * 1024u - (uint32_t)p.x_stages * p.x_bytes - o_bytes;
    }
    if (!fits) continue;
    p.rep = rep;
    if (ws_ok) { p.wstore = 1; p.o_rs = rs; p.o_half = o_half; p.o_nbuf = 2; p.o_bytes = o_bytes; }
    break;
  }
  if (p.tl) {
    if (!p.tma_store) { delete h; return (int)cudaErrorInvalidValue; }
    const int bh = rows / p.tl;
    int rs = bh < 8 ? bh : 8;
    while (bh % rs) --rs;
    p.cg = cdiv(cout, 32); p.rep = 1; p.wstore = 1; p.o_rs = rs; p.o_half = (uint32_t)rs * 8u * 64u; p.o_nbuf = 2;
    p.o_bytes = 4u * 2u * p.o_half;
    p.s_bytes = ((uint32_t)p.cg * rows * 8u * 32u * 4u + 1023u) & ~1023u;        // fp32 lane sums [cg][pixel][32 ch]
  }
  p.w_bytes = p.rep > 1 ? full_tile : ((p.w_bytes_tx + 1023u) & ~1023u);
  p.w_pad = full_tile > p.w_bytes ? full_tile - p.w_bytes : 0u;
  const uint32_t budget = 216u * 1024u - (uint32_t)p.x_stages * p.x_bytes - p.w_pad - p.o_bytes - p.s_bytes;
  const uint32_t all_w = 9u * p.nchunk * p.w_bytes;
  p.w_resident = (!ts && !p.tl && all_w <= budget && env_int("PWC_TC_HALO_RESIDENT", 1)) ? 1u : 0u;
  if (!p.w_resident && p.rep > 1) { delete h; return (int)cudaErrorInvalidValue; }
  // streamed weights: 3 taps per stage when the pixel tile is small (per-stage handshakes would dominate the MMAs)
  p.tps = 1;
  if (!ts && !p.tl && !p.w_resident && p.npx <= 128 && budget / (3u * p.w_bytes) >= 3 && env_int("PWC_TC_WM_TPS", 3) == 3) p.tps = 3;
  uint32_t stages = budget / (p.w_bytes * p.tps);
  p.w_stages = stages > 12 ? 12 : (stages < 2 ? 2 : stages);
  uint32_t cols = 32; while (cols < 2u * p.npx) cols <<= 1;
  if (ts) cols = 512;
  p.tmem_cols = cols;
  p.fp16 = g_half_format; p.dbg = env_int("PWC_TC_HALO_BO", 0);
  p.bias = bias; p.slope = slope;
  p.out = out; p.out_pitch = out_pitch; p.out_coff = out_coff;
  p.out2 = out2; p.out2_pitch = out2_pitch; p.out2_coff = out2_coff; p.n_split = (out2 && out_f32 != 2) ? n_split : N;
  p.out_f32 = out_f32; p.residual = residual; p.res_pitch = res_pitch;
  const uint32_t nw = p.w_resident ? 9u * p.nchunk : p.w_stages * p.tps;
  h->smem_bytes = (uint32_t)p.x_stages * p.x_bytes + (size_t)nw * p.w_bytes + p.w_pad + p.o_bytes + p.s_bytes + 8 * (14 + 2 * p.w_stages) + 16 + 1024;
  const CUtensorMapSwizzle swz = swizzle_for_bytes(kc * 2);
  const __nv_bfloat16* a0 = reinterpret_cast<const __nv_bfloat16*>(in) + in_coff;
  int e = 0;
  if (s2) {
    e = make_tmap_pairs(&h->tmA0, a0, in_pitch, cin_span + p.lo_off, Win, Hin, N, kc, 9, 2 * rows + 1, swz);
  } else if (p.xsplit) {
    if (!p.wstore) e = (int)cudaErrorInvalidValue;
    if (!e) e = make_tmap_xsplit(&h->tmA0, a0, in_pitch, cin_span, W, H, N, dil, kc, 10, rows + 2 * dil, swz);
  } else {
    // split input: one map spans both planes (the hi plane's partial last chunk then reads real neighbouring channels
    // instead of zero fill: their weight rows are zero and all stored values are finite)
    e = make_tmap_nhwc(&h->tmA0, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a0, in_pitch, cin_span + p.lo_off, W, H, N,
                       kc, 10 * dil, ydense ? rows + 2 * dil : (rows + 2 + xextra) * dil, dil, p.dy, swz);
  }
  if (!e) e = make_tmap_2d(&h->tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w_packed, (long)9 * p.nchunk * coutp, kc, kc, coutp, swz);
  h->tmO = h->tmA0; h->tmO2 = h->tmA0;
  if (!e && p.xsplit) {
    if ((out_pitch & 7) || (out_coff & 7)) e = (int)cudaErrorInvalidValue;
    if (!e) e = make_tmap_xsplit(&h->tmO, reinterpret_cast<__nv_bfloat16*>(out) + out_coff, out_pitch, coutp, W, H, N, dil,
                                 32, 8, p.o_rs, CU_TENSOR_MAP_SWIZZLE_NONE);
  } else if (!e && p.tma_store) {
    // output slice as a 4-D tensor (CoutP, W, H, images); box = half an item; dilation = element strides; no swizzle
    if ((out_pitch & 7) || (out_coff & 7) || (out2 && ((out2_pitch & 7) || (out2_coff & 7)))) e = (int)cudaErrorInvalidValue;
    const int n1 = out2 ? p.n_split : N;
    const int bc = p.tl ? p.tl_bc : p.wstore ? 32 : coutp, br = p.wstore ? p.o_rs : rows / 2;
    if (p.tl) coutp = p.tl_bc * p.cg;              // the map covers the real channel groups only (coutp = 128 counted MMA rows)
    if (p.out_lo_off && !p.wstore) e = (int)cudaErrorInvalidValue;     // split output exists in the per-warp store path only
    if (!e) e = make_tmap_nhwc(&h->tmO, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, reinterpret_cast<__nv_bfloat16*>(out) + out_coff,
                               out_pitch, coutp + p.out_lo_off, W, H, n1, bc, 8 * dil, br * p.dy, dil, p.dy, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (!e && out2) e = make_tmap_nhwc(&h->tmO2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, reinterpret_cast<__nv_bfloat16*>(out2) + out2_coff,
                                       out2_pitch, coutp + p.out_lo_off, W, H, N - n1, bc, 8 * dil, br * p.dy, dil, p.dy,
                                       CU_TENSOR_MAP_SWIZZLE_NONE);
  }
  if (e || h->smem_bytes > 227u * 1024u || cols > 512) { delete h; return e ? e : (int)cudaErrorInvalidValue; }
  h->tmA1 = h->tmA0;
  // split-K for layers whose pixel tiles cannot fill the machine (pyramid levels 5, 6): ksplit CTAs per tile, each a
  // contiguous range of 64-channel chunks; partial sums meet in an fp32 scratch, the last CTA to arrive finishes the tile.
  // OPT-IN (PWC_TC_WM_KSPLIT=1): measured on B200 it only pays for >= 8 chunks (conv6_4 24 -> 18 us) and loses on short
  // K loops (conv5aa 16 -> 34 us: the rendezvous costs ~5 us), and the atomic summation order makes replays differ in
  // the last bit, which the parity tests do not accept.
  p.ksplit = 1; p.cps = p.nchunk; p.scratch = nullptr; p.counters = nullptr; p.ks2pass = 0;
  // Deterministic two-pass split-K (default ON, PWC_TC_WM_KSPLIT2=0 turns it off): layers whose pixel tiles fill less than half
  // of the SMs and whose K loop has >= PWC_TC_WM_KSPLIT2_MINCH (6) chunks -- levels 5-6, and the split-precision level 4-6
  // decoders whose K loops are three times as long -- run ks CTAs per tile on contiguous chunk ranges; each stores its fp32
  // partial tile to its own scratch slab and wm_splitk_reduce_kernel (a second launch) sums the slabs in order.
  const bool two_pass = env_int("PWC_TC_WM_KSPLIT2", 1) && dil == 1 && !p.xsplit && p.out_groups == 1 && !p.tl &&
                        2 * p.num_items <= g_num_sms && p.nchunk >= env_int("PWC_TC_WM_KSPLIT2_MINCH", 6);
  if (p.wstore && !ts && !p.tl && p.nchunk >= 2 && (two_pass || (!p.out_lo_off && env_int("PWC_TC_WM_KSPLIT", 0)))) {
    int ks = g_num_sms / p.num_items;
    if (ks > 8) ks = 8;
    if (two_pass && ks > p.nchunk / 3) ks = p.nchunk / 3;            // at least 3 chunks per CTA: below that the pipeline fill dominates
    if (ks > p.nchunk) ks = p.nchunk;
    const int forced_ks = env_int("PWC_TC_WM_KSPLIT_N", 0);
    if (forced_ks > 0) ks = forced_ks < p.nchunk ? forced_ks : p.nchunk;
    if (ks >= 2) {
      p.cps = cdiv(p.nchunk, ks);
      p.ksplit = cdiv(p.nchunk, p.cps);
      p.ks2pass = two_pass ? 1 : 0;
      const size_t fl = (size_t)p.num_items * p.npx * 128 * (two_pass ? p.ksplit : 1);
      const size_t bytes = fl * sizeof(float) + (size_t)p.num_items * sizeof(int);
      if (cudaMalloc(&h->scratch, bytes) != cudaSuccess || cudaMemset(h->scratch, 0, bytes) != cudaSuccess) {
        delete h; return (int)cudaErrorMemoryAllocation;
      }
      p.scratch = reinterpret_cast<float*>(h->scratch);
      p.counters = reinterpret_cast<int*>(p.scratch + fl);
    }
  }
  { const int v = p.num_items * p.ksplit; h->grid = v < g_num_sms ? v : g_num_sms; }
  *handle = h;
  return 0;
}

}  // namespace pwc

extern "C" int pwc_tc_conv_create(pwc_tc_conv** handle,
                                  const void* in, long in_pitch, int in_coff,
                                  const void* in2, long in2_pitch, int in2_coff, int n_split, int cin_span, int kc,
                                  const void* w_packed, const float* bias, int cout, int coutp,
                                  void* out, long out_pitch, int out_coff,
                                  void* out2, long out2_pitch, int out2_coff, int out_f32,
                                  const float* residual, long res_pitch,
                                  int N, int Hin, int Win, int stride, int dil, float slope) {
  using namespace pwc;
  if (!handle) return (int)cudaErrorInvalidValue;
  *handle = nullptr;
  {
    // per DEVICE: the shared-memory opt-in is a per-device function attribute and PWCDCNet.forward follows its input's device
    static int attr_dev = -1;
    int dev = 0; cudaGetDevice(&dev);
    if (attr_dev != dev) {
      cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
      if (g_num_sms <= 0) g_num_sms = 148;
      cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      cudaFuncSetAttribute(conv_tc_halo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      cudaFuncSetAttribute(conv_tc_wm_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      cudaFuncSetAttribute(conv_tc_wm_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      cudaFuncSetAttribute(conv_tc_wm_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      cudaFuncSetAttribute(conv_tc_wm_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      cudaFuncSetAttribute(conv_tc_wm_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      cudaFuncSetAttribute(conv_tc_wm_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
      attr_dev = dev;
    }
  }
  const int wm_max_dil = env_int("PWC_TC_WM_YDENSE_MINDIL", 8) <= 16 ? 16 : 8;    // y-dense tiles take dilation 16 too
  if (halo_eligible(stride, dil, in2 != nullptr, kc, wm_max_dil) && coutp <= 128 && (env_int("PWC_TC_WM", 1) || out_f32 == 2) &&
      !(out_f32 == 1 && (cin_span + kc - 1) / kc < 4)) {
    const int e = create_wm(handle, in, in_pitch, in_coff, cin_span, kc, w_packed, bias, cout, coutp, out, out_pitch, out_coff,
                            out2, out2_pitch, out2_coff, n_split, out_f32, residual, res_pitch, N, Hin, Win, dil, slope);
    if (!e || dil <= 8) return e;
    *handle = nullptr;                                     // dilation 16 outside the y-dense mode's shapes: per-tap kernel
  }
  // (measured: big stride-2 layers lose there -- four input pixels per output pixel make the halo tiles small, so Cout <= 32
  // leaves one epilogue warp working and Cout >= 64 streams the whole weight set per 96-pixel tile; the layers whose 128-pixel
  // per-tap tiles cannot fill the SMs gain from its split-K: conv5a 34 -> 21 us)
  const long pt_tiles = (long)N * (((Hin - 1) / 2 + 1 + 7) / 8) * (((Win - 1) / 2 + 1 + 15) / 16);
  const int s2_mode = env_int("PWC_TC_WM_S2", 1);
  if (stride == 2 && dil == 1 && !in2 && !out2 && coutp <= 128 && !(Win & 1) && !out_f32 &&
      (s2_mode == 2 || (s2_mode == 1 && pt_tiles < g_num_sms))) {
    // stride 2 through the halo kernel's pair view (WmParams::s2); shapes it does not take fall through to the per-tap kernel
    const int e = create_wm(handle, in, in_pitch, in_coff, cin_span, kc, w_packed, bias, cout, coutp, out, out_pitch, out_coff,
                            nullptr, 0, 0, n_split, out_f32, residual, res_pitch, N, Hin, Win, dil, slope, 2);
    if (!e) return 0;
    *handle = nullptr;
  }
  if (out_f32 == 2) return (int)cudaErrorInvalidValue;     // fused predict_flow+upfeat exists in the WM kernel only
  if (halo_eligible(stride, dil, in2 != nullptr, kc)) {
    return create_halo(handle, in, in_pitch, in_coff, cin_span, kc, w_packed, bias, cout, coutp, out, out_pitch, out_coff,
                       out2, out2_pitch, out2_coff, n_split, out_f32, residual, res_pitch, N, Hin, Win, dil, slope);
  }
  if ((kc != 16 && kc != 32 && kc != 64) || cin_span % kc || coutp % 16 || coutp < 16 || coutp > 256 || cout > coutp ||
      (in_pitch & 7) || (in_coff & 7) || (in2 && ((in2_pitch & 7) || (in2_coff & 7))) ||
      (!out_f32 && ((out_pitch & 7) || (out_coff & 7) || (out2 && ((out2_pitch & 7) || (out2_coff & 7))))) ||
      (stride != 1 && stride != 2))
    return (int)cudaErrorInvalidValue;
  pwc_tc_conv* h = new (std::nothrow) pwc_tc_conv();
  if (!h) return (int)cudaErrorMemoryAllocation;
  h->halo = 0;
  TcConvParams& p = h->p;
  p.N = N; p.Hin = Hin; p.Win = Win; p.stride = stride; p.dil = dil;
  p.Hout = (Hin - 1) / stride + 1; p.Wout = (Win - 1) / stride + 1;
  p.split = g_split.in_lo_off ? g_split.products : 0; p.lo_off = g_split.in_lo_off; p.out_lo_off = g_split.out_lo_off;
  if ((p.lo_off & 7) || (p.out_lo_off & 7) || (p.out_lo_off && out_f32)) { delete h; return (int)cudaErrorInvalidValue; }
  p.cout = cout; p.coutp = coutp; p.kc = kc; p.nchunk = (cin_span / kc) * (p.split ? p.split : 1);
  p.fused = (p.split == 3 && p.lo_off == cin_span && cin_span == kc && kc <= 32 && env_int("PWC_TC_FUSED_SPLIT", 1)) ? 1 : 0;
  if (p.fused) { p.kc = kc = 2 * kc; p.nchunk = 2; p.split = 0; }      // weights: [tap][2][coutp][2*cin] = [W_hi|W_hi], [W_lo|0]
  p.n_split = (in2 || out2) ? n_split : N;
  p.tiles_x = cdiv(p.Wout, kTcTW); p.tiles_y = cdiv(p.Hout, kTcTH); p.num_tiles = p.tiles_x * p.tiles_y * N;
  p.a_bytes = (uint32_t)kTcM * kc * 2;
  p.stage_bytes = (p.a_bytes + (uint32_t)coutp * kc * 2 * (p.fused ? 2 : 1) + 1023u) & ~1023u;
  p.w_tile = ((uint32_t)coutp * kc * 2 + 1023u) & ~1023u;
  {
    const uint32_t all_w = 9u * p.nchunk * p.w_tile, a_stage = (p.a_bytes + 1023u) & ~1023u;
    p.w_resident = (all_w + 4u * a_stage <= 200u * 1024u && env_int("PWC_TC_RESIDENT", 1)) ? 1u : 0u;
    if (p.w_resident) p.stage_bytes = a_stage;
    p.xdup = (p.w_resident && p.split == 3 && !p.fused && env_int("PWC_TC_XDUP", 1)) ? 1 : 0;
  }
  int stages = (int)((200u * 1024u - (p.w_resident ? 9u * p.nchunk * p.w_tile : 0u)) / p.stage_bytes);
  p.stages = stages > 8 ? 8 : (stages < 2 ? 2 : stages);
  uint32_t cols = 32; while (cols < 2u * coutp) cols <<= 1;
  if (cols > 512) { delete h; return (int)cudaErrorInvalidValue; }
  p.tmem_cols = cols;
  p.bias = bias; p.slope = slope;
  p.out = out; p.out_pitch = out_pitch; p.out_coff = out_coff;
  p.out2 = out2; p.out2_pitch = out2_pitch; p.out2_coff = out2_coff;
  p.out_f32 = out_f32; p.residual = residual; p.res_pitch = res_pitch;
  p.fp16 = g_half_format;
  h->smem_bytes = (size_t)p.stages * p.stage_bytes + (p.w_resident ? (size_t)9 * p.nchunk * p.w_tile : 0) + 8 * (2 * p.stages + 8) +
                  (size_t)coutp * 4 + 1024;
  const CUtensorMapSwizzle swz = swizzle_for_bytes(kc * 2);
  const __nv_bfloat16* a0 = reinterpret_cast<const __nv_bfloat16*>(in) + in_coff;
  if (p.fused) cin_span = kc;                                        // both planes are one box now
  p.pairs = (stride == 2 && dil == 1 && !in2 && !(Win & 1) && Hin >= 2 && env_int("PWC_TC_PAIRS", 1)) ? 1 : 0;
  int e = 0;
  if (p.pairs) {
    const int cs = cin_span + (p.fused ? 0 : p.lo_off);
    e = make_tmap_pairs_rows(&h->tmA0, a0, in_pitch, cs, Win, (Hin + 1) / 2, N, (long)in_pitch * Win * Hin, kc, kTcTW, kTcTH, swz);
    if (!e) e = make_tmap_pairs_rows(&h->tmA1, a0 + (long)in_pitch * Win, in_pitch, cs, Win, Hin / 2, N, (long)in_pitch * Win * Hin,
                                     kc, kTcTW, kTcTH, swz);
  } else {
    e = make_tmap_nhwc(&h->tmA0, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a0, in_pitch, cin_span + (p.fused ? 0 : p.lo_off), Win, Hin,
                       in2 ? p.n_split : N, kc, kTcTW * stride, kTcTH * stride, stride, stride, swz);
  }
  if (!e && !p.pairs) {
    if (in2) {
      const __nv_bfloat16* a1 = reinterpret_cast<const __nv_bfloat16*>(in2) + in2_coff;
      e = make_tmap_nhwc(&h->tmA1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a1, in2_pitch, cin_span + (p.fused ? 0 : p.lo_off), Win, Hin, N - p.n_split,
                         kc, kTcTW * stride, kTcTH * stride, stride, stride, swz);
    } else {
      h->tmA1 = h->tmA0;
    }
  }
  if (!e) e = make_tmap_2d(&h->tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, w_packed, (long)9 * p.nchunk * coutp, kc, kc, coutp, swz);
  if (e) { delete h; return e; }
  h->grid = p.num_tiles < g_num_sms ? p.num_tiles : g_num_sms;
  *handle = h;
  return 0;
}

// Zero-block skip list: bit ch of masks[tap] set <=> the packed weight block (tap, 64-channel chunk ch) is all zero, so
// its TMA load and MMAs are dropped.  Only for the streamed-weights tcgen05 kernel with one tap per stage.
extern "C" int pwc_tc_conv_set_skip(pwc_tc_conv* h, const unsigned long long* masks) {
  if (!h || !masks || h->halo != 2) return (int)cudaErrorInvalidValue;
  pwc::WmParams& p = h->wp;
  if (p.ts || p.tl || p.nchunk > 64) return (int)cudaErrorInvalidValue;
  if (p.w_resident) return 0;                    // whole weight set lives in shared memory: nothing to skip loading
  p.tps = 1;                                     // one tap per ring stage (uses the first w_stages buffers of the ring)
  bool any = false;
  for (int i = 0; i < 9; ++i) { p.skip[i] = masks[i]; any = any || masks[i] != 0ull; }
  p.has_skip = any ? 1 : 0;
  return 0;
}

// Single-chunk layers (Cin span <= 64): bit kk (0..3) of masks[tap] set <=> the 16-channel K step kk of that tap's weight
// tile is all zero, so its MMA is dropped (block structure of space-to-depth weights).
extern "C" int pwc_tc_conv_set_kk_skip(pwc_tc_conv* h, const int* masks) {
  if (!h || !masks || h->halo != 2) return (int)cudaErrorInvalidValue;
  pwc::WmParams& p = h->wp;
  if (p.nchunk != 1 || p.ts || p.has_skip) return (int)cudaErrorInvalidValue;
  const int nkk = (p.cin_span + 15) / 16 < 4 ? (p.cin_span + 15) / 16 : 4;
  int ft = -1, fk = -1, any = 0;
  for (int t = 0; t < 9; ++t) {
    p.kk_mask[t] = masks[t] & 15;
    any |= p.kk_mask[t];
    for (int kk = 0; kk < nkk && ft < 0; ++kk)
      if (!((p.kk_mask[t] >> kk) & 1)) { ft = t; fk = kk; }
  }
  if (ft < 0) return (int)cudaErrorInvalidValue;      // every step skipped: the accumulator would never be written
  p.kk_first_tap = ft; p.kk_first_kk = fk;
  p.has_kk = any ? 1 : 0;
  return 0;
}

// Grouped output: the CoutP output channels are g groups of CoutP/g; group i is written to channel
// out_coff + i*group_pitch of the same pixel (space-to-depth views: group = one of the g pixels of a super-pixel).
extern "C" int pwc_tc_conv_set_out_groups(pwc_tc_conv* h, int g, long group_pitch) {
  using namespace pwc;
  if (!h || h->halo != 2 || g < 2 || g > 8) return (int)cudaErrorInvalidValue;
  WmParams& p = h->wp;
  if (!p.tma_store || p.coutp % g || p.out2 || p.dil != 1 || (group_pitch & 7)) return (int)cudaErrorInvalidValue;
  if (p.ks2pass) {                               // the split-K reduce pass writes plain NHWC slices: grouped layers run unsplit
    p.ksplit = 1; p.cps = p.nchunk; p.ks2pass = 0;
    h->grid = p.num_items < g_num_sms ? p.num_items : g_num_sms;
  }
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return (int)cudaErrorNotSupported;
  const int cg = p.coutp / g;
  cuuint64_t dims[5] = {(cuuint64_t)cg, (cuuint64_t)g, (cuuint64_t)p.W, (cuuint64_t)p.H, (cuuint64_t)p.N};
  cuuint64_t strides[4] = {(cuuint64_t)group_pitch * 2, (cuuint64_t)p.out_pitch * 2, (cuuint64_t)p.out_pitch * 2 * p.W,
                           (cuuint64_t)p.out_pitch * 2 * p.W * p.H};
  cuuint32_t box[5] = {(cuuint32_t)cg, (cuuint32_t)g, 8, (cuuint32_t)(p.rows / 2), 1};
  if (p.wstore) { box[0] = 32; box[1] = 1; box[3] = (cuuint32_t)p.o_rs; }
  if (p.wstore && cg % 32) return (int)cudaErrorInvalidValue;
  cuuint32_t es[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(&h->tmO, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5,
                   reinterpret_cast<__nv_bfloat16*>(p.out) + p.out_coff, dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return (int)cudaErrorInvalidValue;
  p.out_groups = g;
  return 0;
}

extern "C" int pwc_tc_conv_launch(const pwc_tc_conv* h, void* stream) {
  if (!h) return (int)cudaErrorInvalidValue;
  if (h->halo == 2) {
    const cudaStream_t st = (cudaStream_t)stream;
#define PWC_WM_LAUNCH(MODE) \
  pwc::launch_pdl(pwc::conv_tc_wm_kernel<MODE>, dim3(h->grid), dim3(pwc::kHaloThreads), h->smem_bytes, st, h->tmA0, h->tmB, h->tmO, \
                  h->tmO2, h->wp)
    if (h->wp.s2) PWC_WM_LAUNCH(5);
    else if (h->wp.tl) PWC_WM_LAUNCH(4);
    else if (h->wp.ts) PWC_WM_LAUNCH(1);
    else if (h->wp.has_skip) PWC_WM_LAUNCH(2);
    else if (h->wp.has_kk) PWC_WM_LAUNCH(3);
    else PWC_WM_LAUNCH(0);
#undef PWC_WM_LAUNCH
    if (h->wp.ks2pass) {
      const long total = (long)h->wp.num_items * h->wp.npx * (h->wp.coutp >> 2);
      const unsigned nb = (unsigned)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
      pwc::launch_pdl(pwc::wm_splitk_reduce_kernel, dim3(nb), dim3(256), 0, st, h->wp);
    }
    return (int)cudaGetLastError();
  }
  if (h->halo) {
    pwc::launch_pdl(pwc::conv_tc_halo_kernel, dim3(h->grid), dim3(pwc::kHaloThreads), h->smem_bytes, (cudaStream_t)stream,
                    h->tmA0, h->tmB, h->hp);
    return (int)cudaGetLastError();
  }
  pwc::launch_pdl(pwc::conv_tc_kernel, dim3(h->grid), dim3(pwc::kTcThreads), h->smem_bytes, (cudaStream_t)stream, h->tmA0,
                  h->tmA1, h->tmB, h->p);
  return (int)cudaGetLastError();
}

extern "C" void pwc_tc_conv_destroy(pwc_tc_conv* h) { delete h; }

// kernels one pwc_tc_conv_launch of this handle enqueues (2 = deterministic two-pass split-K: conv kernel + reduce kernel)
// Head rows of a three-lane tap-lane layer (WmParams::nhead): weight rows 96 + ky * nhead + h of the packed [128, Cin, 3, 3] tensor
// are kernel row ky of `nhead` extra output channels; their raw sums go to head_out [N*H*W][nhead] fp32.
extern "C" int pwc_tc_conv_set_head_rows(pwc_tc_conv* h, float* head_out, int nhead) {
  if (!h || h->halo != 2 || h->wp.tl != 3 || h->wp.cg != 1 || !head_out || nhead < 1 || 3 * nhead > 32) return (int)cudaErrorInvalidValue;
  if (!h->wp.nhead) {                               // 48 pixels x 32 lanes of fp32 behind the barriers (see the kernel's head warp)
    if (h->smem_bytes + 6144 + 32 > 227u * 1024u || h->wp.rows % 6) return (int)cudaErrorInvalidValue;
    h->smem_bytes += 6144 + 32;
  }
  h->wp.nhead = nhead; h->wp.head_out = head_out;
  return 0;
}

// Tap-stacked layer (out_f32 | 16): partial [N*H*W][pitch] fp32 is added to every output channel before the activation.
extern "C" int pwc_tc_conv_set_partial(pwc_tc_conv* h, const float* partial, int pitch) {
  if (!h || h->halo != 2 || !h->wp.ts || !partial || pitch < h->wp.cout) return (int)cudaErrorInvalidValue;
  h->wp.partial = partial; h->wp.partial_pitch = pitch;
  return 0;
}

extern "C" int pwc_tc_conv_num_launches(const pwc_tc_conv* h) { return !h ? 0 : (h->halo == 2 && h->wp.ks2pass) ? 2 : 1; }

extern "C" int pwc_tc_conv_create_split(pwc_tc_conv** handle,
                                        const void* in, long in_pitch, int in_coff,
                                        const void* in2, long in2_pitch, int in2_coff, int n_split, int cin_span, int kc,
                                        const void* w_packed, const float* bias, int cout, int coutp,
                                        void* out, long out_pitch, int out_coff,
                                        void* out2, long out2_pitch, int out2_coff, int out_f32,
                                        const float* residual, long res_pitch,
                                        int N, int Hin, int Win, int stride, int dil, float slope,
                                        int in_lo_off, int out_lo_off, int up_lo_off, int up_dup_off) {
  if (out_lo_off < 0 || up_lo_off < 0 || up_dup_off < 0) return (int)cudaErrorInvalidValue;
  // in_lo_off < 0: split ACTIVATIONS only (two products per K step, x_hi*W_hi + x_lo*W_hi; lo plane at -in_lo_off)
  pwc::g_split = {in_lo_off < 0 ? -in_lo_off : in_lo_off, out_lo_off, up_lo_off, up_dup_off, in_lo_off < 0 ? 2 : 3};
  const int e = pwc_tc_conv_create(handle, in, in_pitch, in_coff, in2, in2_pitch, in2_coff, n_split, cin_span, kc, w_packed, bias,
                                   cout, coutp, out, out_pitch, out_coff, out2, out2_pitch, out2_coff, out_f32, residual,
                                   res_pitch, N, Hin, Win, stride, dil, slope);
  pwc::g_split = {0, 0, 0, 0, 3};
  return e;
}

extern "C" int pwc_set_half_format(int fmt) {
  if (fmt != 0 && fmt != 1) return (int)cudaErrorInvalidValue;
  pwc::g_half_format = fmt;
  return 0;
}
extern "C" int pwc_get_half_format(void) { return pwc::g_half_format; }

// ------------------------------------------------------------------------------------------------------------------
// Split-K launch that allocates nothing (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops).  The fp32 partial-sum scratch and
// the per-tile arrival counters of the split-K path live in ONE file-scope device array of kStaticScratchBytes; every
// GPU that loads the library holds it until exit.  In host code `g_static_scratch` is only a stand-in for the symbol:
// the device address comes from cudaGetSymbolAddress, once per device with that device current.
// ------------------------------------------------------------------------------------------------------------------
namespace pwc {
constexpr size_t kStaticScratchBytes = 16u << 20;                   // 16 MiB per GPU
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];
constexpr int kMaxDevices = 64;
static void* g_static_scratch_ptr[kMaxDevices] = {};              // device address of the array, per device
static std::mutex g_static_scratch_mu;
}  // namespace pwc

extern "C" int pwc_tc_conv_launch_static_splitk(const pwc_tc_conv* h, int ksplit, void* stream) {
  using namespace pwc;
  if (!h || h->halo != 2) return (int)cudaErrorInvalidValue;
  WmParams p = h->wp;                                               // by-value copy: the handle itself is not modified
  p.ks2pass = 0;                                                    // this entry point runs the one-pass (atomic) form
  if (!p.wstore || p.ts || p.tl || p.nchunk < 2 || ksplit < 0 || p.out_lo_off) {
    fprintf(stderr, "pwc_tc_conv_launch_static_splitk: the layer has no split-K form (needs >= 2 channel chunks, "
                    "16-bit per-warp output stores, no tap stacking)\n");
    return (int)cudaErrorInvalidValue;
  }
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
  void* base = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_static_scratch_mu);
    if (!g_static_scratch_ptr[dev]) {
      void* sym = nullptr;
      e = cudaGetSymbolAddress(&sym, g_static_scratch);             // the array's address on the CURRENT device
      if (e != cudaSuccess) return (int)e;
      g_static_scratch_ptr[dev] = sym;
    }
    base = g_static_scratch_ptr[dev];
  }
  int nsm = 0;
  cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
  if (nsm <= 0) nsm = 148;
  int ks = ksplit ? ksplit : nsm / p.num_items;                     // 0 = as many CTAs per tile as fill the machine
  if (ks > 8) ks = 8;
  if (ks > p.nchunk) ks = p.nchunk;
  if (ks < 2) ks = 2;
  p.cps = cdiv(p.nchunk, ks);
  p.ksplit = cdiv(p.nchunk, p.cps);
  const size_t fl = (size_t)p.num_items * p.npx * 128;              // fp32 partial sums [item][pixel][128 channels]
  const size_t bytes = fl * sizeof(float) + (size_t)p.num_items * sizeof(int);
  if (bytes > kStaticScratchBytes) {
    fprintf(stderr, "pwc_tc_conv_launch_static_splitk: layer needs %zu bytes of split-K scratch, the static array holds %zu\n",
            bytes, kStaticScratchBytes);
    return (int)cudaErrorInvalidValue;
  }
  p.scratch = reinterpret_cast<float*>(base);
  p.counters = reinterpret_cast<int*>(p.scratch + fl);
  const cudaStream_t st = (cudaStream_t)stream;
  // every call shares the array, which keeps the last call's contents: sums and counters must start at zero
  e = cudaMemsetAsync(base, 0, bytes, st);
  if (e != cudaSuccess) return (int)e;
  const int v = p.num_items * p.ksplit;
  const int grid = v < nsm ? v : nsm;
  cudaLaunchConfig_t cfg = {};                                      // plain stream order (no programmatic early launch)
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kHaloThreads); cfg.dynamicSmemBytes = h->smem_bytes; cfg.stream = st;
  if (p.s2) e = cudaLaunchKernelEx(&cfg, conv_tc_wm_kernel<5>, h->tmA0, h->tmB, h->tmO, h->tmO2, p);
  else if (p.has_skip) e = cudaLaunchKernelEx(&cfg, conv_tc_wm_kernel<2>, h->tmA0, h->tmB, h->tmO, h->tmO2, p);
  else if (p.has_kk) e = cudaLaunchKernelEx(&cfg, conv_tc_wm_kernel<3>, h->tmA0, h->tmB, h->tmO, h->tmO2, p);
  else e = cudaLaunchKernelEx(&cfg, conv_tc_wm_kernel<0>, h->tmA0, h->tmB, h->tmO, h->tmO2, p);
  return (int)(e != cudaSuccess ? e : cudaGetLastError());
}

extern "C" int pwc_pack_conv_weight_bf16(const float* w_torch, void* w_packed, const int* chan_map,
                                         int Cout, int Cin, int cin_span, int kc, int coutp, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  const int nchunk = (cin_span + kc - 1) / kc;     // a partial last chunk is zero padded (halo kernel: TMA zero-fills A)
  cudaError_t e = cudaMemsetAsync(w_packed, 0, 2 * (size_t)9 * nchunk * coutp * kc, s);
  if (e != cudaSuccess) return (int)e;
  pwc::pack_conv_weight_bf16_kernel<<<256, 256, 0, s>>>(w_torch, reinterpret_cast<__nv_bfloat16*>(w_packed), chan_map,
                                                        Cout, Cin, nchunk, kc, coutp, pwc::g_half_format);
  return (int)cudaGetLastError();
}

// Debug: copy the halo kernel's timeline buffer (see conv_tc_halo.cuh) to host memory; returns a cudaError.
extern "C" int pwc_debug_wm_trace(unsigned long long* host_dst, int count) {
  if (!host_dst || count < 0 || count > 3 * 64 * 4) return (int)cudaErrorInvalidValue;
  return (int)cudaMemcpyFromSymbol(host_dst, pwc::g_wm_trace, sizeof(unsigned long long) * count);
}

extern "C" int pwc_debug_halo_trace(unsigned long long* host_dst, int count) {
  if (count > 8192) count = 8192;
  return (int)cudaMemcpyFromSymbol(host_dst, pwc::g_halo_trace, sizeof(unsigned long long) * count);
}
