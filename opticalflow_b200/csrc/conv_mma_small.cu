// 3x3 convolution for the SMALL pyramid levels (5 and 6: a few thousand pixels per batch, bf16 NHWC, stride 1 or 2).
// conv() of /root/reference/models/PWCNet.py:26-30 at its call sites PWCNet.py:186-195 (conv5a..conv6b) and
// PWCNet.py:202-206,217-221 (conv6_0..4, conv5_0..4); bias + LeakyReLU fused, output written into its concat slice.
//
// These layers are latency problems, not throughput problems: 0.1-2 GFLOP each, but a K loop of 9 taps x up to 12
// 64-channel chunks.  On the tcgen05 kernels (one CTA per 128 x N tile, one serial chain of several hundred MMAs behind a
// TMA/TMEM pipeline fill) they take 10-30 us at 16-64 CTAs.  Here the work is cut three ways so that every SM gets a CTA
// and no chain is long:
//   CTA   = (image, tile of TH x 16 output pixels, slice of 8*NTL output channels)
//   warp  = (wm, wk): MTW tile rows (one 16-pixel m-tile each) x every WK-th (tap, chunk) pair of the K loop
//   MMA   = mma.sync.m16n8k16 bf16 / fp32 accumulate.  A comes from the input patch staged ONCE per CTA in shared memory
//           (pixel rows padded to an odd number of 16-byte chunks: conflict-free); the 3x3 taps and the stride are just
//           different ldmatrix row addresses into that patch (implicit GEMM without im2col).  B (weights) is used by
//           exactly one warp once, so its fragments go global -> registers directly (32-byte sectors fully used), one
//           (tap, chunk) pair ahead of the MMAs.
//   K-split partial sums meet in shared memory (fixed order: deterministic), then bias + LeakyReLU + bf16 + store.
// What bounds such a layer is the per-SM fetch rate (a CTA that owns a whole K loop pulls the whole weight slice, up to
// 1 MB, through one SM's L2 port), so the K loop is ALSO cut across a thread-block cluster: CTA `ks` of a cluster of S
// takes the 64-channel chunks ks, ks+S, ... (it stages only those channels of the patch and reads only those weight
// rows), the S partial tiles are summed through distributed shared memory in rank order after one cluster barrier.
#include <cooperative_groups.h>
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "../../include/pwc_b200.h"

namespace pwc {

struct SmallConvParams {
  const __nv_bfloat16* in; long in_pitch; int in_coff;
  const __nv_bfloat16* in2; long in2_pitch; int in2_coff; int n_split_in;   // images >= n_split_in read in2 (or null)
  int span, nchunk, cs;                       // channels read, 64-channel chunks, cs = 64 * nchunk
  const __nv_bfloat16* w; const float* bias; int cout, coutp;
  __nv_bfloat16* out; long out_pitch; int out_coff;
  __nv_bfloat16* out2; long out2_pitch; int out2_coff; int n_split;         // images >= n_split write out2 (or null)
  int N, H, W, Ho, Wo, stride; float slope;
  int wm, wk;                                 // warp grid: wm x wk = warps per CTA
  int tiles_x, tiles_y, q;                    // output-channel slices per tile
  int sp;                                     // shared-memory bytes per staged pixel: 2*cs + 16
  int ih, iw;                                 // staged patch: ((TH-1)*stride + 3) x (15*stride + 3) pixels
  int S, nloc;                                // cluster size (K split over CTAs), chunks per CTA = ceil(nchunk / S)
  int off_part;                               // byte offset of this CTA's partial tile [pixel][8*NTL] fp32
};

__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int MTW, int NTL, bool F16>
__global__ void __launch_bounds__(256, 2) conv_mma_small_kernel(const SmallConvParams p) {
  extern __shared__ __align__(128) uint8_t sm_raw[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp % p.wm, wk = warp / p.wm;
  const int TH = p.wm * MTW;
  // item -> (image, tile, channel slice); slices of one tile are adjacent so that they share the patch in L2
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int ks = p.S > 1 ? (int)cluster.block_rank() : 0;
  int item = blockIdx.x / p.S;
  const int qs = item % p.q; item /= p.q;
  const int tx = item % p.tiles_x; item /= p.tiles_x;
  const int ty = item % p.tiles_y; const int n = item / p.tiles_y;
  const int oy0 = ty * TH, ox0 = tx * 16;
  const int iy0 = oy0 * p.stride - 1, ix0 = ox0 * p.stride - 1;
  const bool second_in = p.in2 != nullptr && n >= p.n_split_in;
  const __nv_bfloat16* src = second_in ? p.in2 + ((long)(n - p.n_split_in) * p.H * p.W) * p.in2_pitch + p.in2_coff
                                       : p.in + ((long)n * p.H * p.W) * p.in_pitch + p.in_coff;
  const long spitch = second_in ? p.in2_pitch : p.in_pitch;
  const uint32_t sm_addr = ptx::smem_addr(sm_raw);

  pdl_wait();                       // the input comes from the previous kernel (PDL: see common.cuh)
  pdl_launch_dependents();

  // ---- first weight fragments (independent of the patch): pair index = wk, wk + WK, ...
  const int nloc = (p.nchunk - ks + p.S - 1) / p.S;           // chunks ks, ks + S, ... of this CTA
  const int pairs = 9 * nloc;                                // local pair pr = tap * nloc + lc
  const int co_lane = qs * (NTL * 8) + (lane >> 2);
  const __nv_bfloat16* wl = p.w + (long)co_lane * 64 + 2 * (lane & 3);
  uint32_t b0[NTL][8], b1[NTL][8], b2[NTL][8];              // three register buffers: [n-tile][kk*2 + half]
  auto load_b = [&](uint32_t (&bq)[NTL][8], int pr) {
    const int tap = pr / nloc, lc = pr - tap * nloc;
    const __nv_bfloat16* wp = wl + (long)(tap * p.nchunk + ks + lc * p.S) * p.coutp * 64;   // rows [coutp][64] of (tap, chunk)
#pragma unroll
    for (int j = 0; j < NTL; ++j) {
      const bool live = co_lane + 8 * j < p.coutp;           // last slice of a coutp that is not a multiple of 8*NTL
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        bq[j][2 * kk] = live ? __ldg(reinterpret_cast<const unsigned int*>(wp + j * 8 * 64 + kk * 16)) : 0u;
        bq[j][2 * kk + 1] = live ? __ldg(reinterpret_cast<const unsigned int*>(wp + j * 8 * 64 + kk * 16 + 8)) : 0u;
      }
    }
  };
  // the first three pairs of this warp are requested before anything else (at level 6 that is the whole K loop)
  if (wk < pairs) load_b(b0, wk);
  if (wk + p.wk < pairs) load_b(b1, wk + p.wk);
  if (wk + 2 * p.wk < pairs) load_b(b2, wk + 2 * p.wk);

  // ---- stage the input patch with cp.async (16-byte groups, zero fill outside the image / beyond the channel span):
  //      every thread has all its copies in flight at once -- a load/store loop would serialise on the L2 latency
  {
    const int ncg = nloc << 3, nvalid = (p.span + 7) >> 3;   // local 16-byte groups per staged pixel; groups with real channels
    const int total = p.ih * p.iw * ncg;
    const int nthr = blockDim.x, dq = ncg ? nthr / ncg : 0, dr = ncg ? nthr - dq * ncg : 0;
    int px = ncg ? tid / ncg : 0, cg = ncg ? tid - px * ncg : 0;
    for (int idx = tid; idx < total; idx += nthr) {
      const int r = px / p.iw, c = px - r * p.iw;
      const int iy = iy0 + r, ix = ix0 + c;
      const int gcg = ((ks + (cg >> 3) * p.S) << 3) + (cg & 7);          // group index in the tensor's channel order
      const bool ok = iy >= 0 && iy < p.H && ix >= 0 && ix < p.W && gcg < nvalid;
      const __nv_bfloat16* g = ok ? src + ((long)iy * p.W + ix) * spitch + 8 * gcg : src;
      const uint32_t d = sm_addr + (uint32_t)(px * p.sp + 16 * cg);
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(d), "l"(g), "r"(ok ? 16 : 0) : "memory");
      px += dq; cg += dr;
      if (cg >= ncg) { cg -= ncg; ++px; }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
  __syncthreads();

  // ---- K loop
  float acc[MTW][NTL][4];
#pragma unroll
  for (int i = 0; i < MTW; ++i)
#pragma unroll
    for (int j = 0; j < NTL; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;
  // ldmatrix row address of this lane: output pixel (row wm*MTW + i, column lane & 15), k half lane >> 4
  const uint32_t a_lane = sm_addr + (uint32_t)(((wm * MTW * p.stride) * p.iw + (lane & 15) * p.stride) * p.sp + (lane >> 4) * 16);
  const uint32_t a_row = (uint32_t)(p.stride * p.iw * p.sp);
  auto compute = [&](const uint32_t (&bq)[NTL][8], int pr) {
    const int tap = pr / nloc, lc = pr - tap * nloc;
    const int ky = tap / 3, kx = tap - 3 * ky;
    const uint32_t a_tap = a_lane + (uint32_t)((ky * p.iw + kx) * p.sp + lc * 128);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
      uint32_t a[MTW][4];
#pragma unroll
      for (int i = 0; i < MTW; ++i) ldsm4(a[i], a_tap + i * a_row + kk * 32);
#pragma unroll
      for (int i = 0; i < MTW; ++i)
#pragma unroll
        for (int j = 0; j < NTL; ++j) mma_h16<F16>(acc[i][j], a[i], bq[j][2 * kk], bq[j][2 * kk + 1]);
    }
  };
  for (int pr = wk; pr < pairs;) {                           // weights three (tap, chunk) pairs ahead of the MMAs
    compute(b0, pr);
    if (pr + 3 * p.wk < pairs) load_b(b0, pr + 3 * p.wk);
    pr += p.wk;
    if (pr >= pairs) break;
    compute(b1, pr);
    if (pr + 3 * p.wk < pairs) load_b(b1, pr + 3 * p.wk);
    pr += p.wk;
    if (pr >= pairs) break;
    compute(b2, pr);
    if (pr + 3 * p.wk < pairs) load_b(b2, pr + 3 * p.wk);
    pr += p.wk;
  }
  __syncthreads();                                           // patch fully consumed: its memory becomes the reduction buffer

  // ---- K-split reduction through shared memory: red[wk][pixel][8*NTL] fp32
  float* red = reinterpret_cast<float*>(sm_raw);
  constexpr int NC = 8 * NTL;
  const int npix = TH * 16;
#pragma unroll
  for (int i = 0; i < MTW; ++i)
#pragma unroll
    for (int j = 0; j < NTL; ++j)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int px = (wm * MTW + i) * 16 + (lane >> 2) + 8 * h;
        *reinterpret_cast<float2*>(red + ((long)wk * npix + px) * NC + j * 8 + 2 * (lane & 3)) =
            make_float2(acc[i][j][2 * h], acc[i][j][2 * h + 1]);
      }
  __syncthreads();
  // this CTA's partial tile: sum of its K-groups in fixed order
  float* part = reinterpret_cast<float*>(sm_raw + p.off_part);
  for (int e = tid; e < npix * (NC / 2); e += blockDim.x) {
    float s0 = 0.f, s1 = 0.f;
    for (int k = 0; k < p.wk; ++k) {
      const float2 v = *reinterpret_cast<const float2*>(red + (long)k * npix * NC + 2 * e);
      s0 += v.x; s1 += v.y;
    }
    *reinterpret_cast<float2*>(part + 2 * e) = make_float2(s0, s1);
  }
  if (p.S > 1) cluster.sync(); else __syncthreads();
  const bool second = p.out2 != nullptr && n >= p.n_split;
  __nv_bfloat16* dst = second ? p.out2 + ((long)(n - p.n_split) * p.Ho * p.Wo) * p.out2_pitch + p.out2_coff
                              : p.out + ((long)n * p.Ho * p.Wo) * p.out_pitch + p.out_coff;
  const long dpitch = second ? p.out2_pitch : p.out_pitch;
  // CTA ks finishes every S-th channel pair: partial tiles of all ranks summed in rank order (deterministic)
  for (int e = ks * blockDim.x + tid; e < npix * (NC / 2); e += p.S * blockDim.x) {
    const int px = e / (NC / 2), cp = e - px * (NC / 2);
    const int oy = oy0 + (px >> 4), ox = ox0 + (px & 15);
    const int co = qs * NC + 2 * cp;
    if (oy >= p.Ho || ox >= p.Wo || co >= p.coutp) continue;
    float s0 = 0.f, s1 = 0.f;
    if (p.S > 1) {
      for (int r = 0; r < p.S; ++r) {
        const float2 v = *reinterpret_cast<const float2*>(cluster.map_shared_rank(part, r) + 2 * e);
        s0 += v.x; s1 += v.y;
      }
    } else {
      const float2 v = *reinterpret_cast<const float2*>(part + 2 * e);
      s0 = v.x; s1 = v.y;
    }
    s0 += co < p.cout ? p.bias[co] : 0.f;
    s1 += co + 1 < p.cout ? p.bias[co + 1] : 0.f;
    *reinterpret_cast<uint32_t*>(dst + ((long)oy * p.Wo + ox) * dpitch + co) =
        pack2_h16(fmaxf(s0, s0 * p.slope), fmaxf(s1, s1 * p.slope), F16 ? 1 : 0);
  }
  if (p.S > 1) cluster.sync();                               // nobody leaves while its partial tile is still being read
}

template <int MTW, int NTL, bool F16>
static int launch_small(const SmallConvParams& p, int grid, int threads, int smem, cudaStream_t s) {
  static int attr_dev = -1;
  int dev = 0; cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t e = cudaFuncSetAttribute(conv_mma_small_kernel<MTW, NTL, F16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr_dev = dev;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  at[1].id = cudaLaunchAttributeClusterDimension;
  at[1].val.clusterDim.x = p.S; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 2;
  return (int)cudaLaunchKernelEx(&cfg, conv_mma_small_kernel<MTW, NTL, F16>, p);
}

}  // namespace pwc

extern "C" int pwc_conv3x3_mma_fwd(const void* in, long in_pitch, int in_coff,
                                   const void* in2, long in2_pitch, int in2_coff, int n_split_in, int cin_span,
                                   const void* w_packed, const float* bias, int cout, int coutp,
                                   void* out, long out_pitch, int out_coff,
                                   void* out2, long out2_pitch, int out2_coff, int n_split,
                                   int N, int H, int W, int stride, float slope, void* stream) {
  using namespace pwc;
  if (N <= 0 || H <= 0 || W <= 0 || (stride != 1 && stride != 2) || cin_span <= 0 || coutp % 16 || cout > coutp ||
      (in_pitch & 7) || (in_coff & 7) || (in2 && ((in2_pitch & 7) || (in2_coff & 7))) || (out_pitch & 1) || (out_coff & 1) ||
      (out2 && ((out2_pitch & 1) || (out2_coff & 1))) || slope < 0.f || slope > 1.f ||
      ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(in2)) & 15) || (reinterpret_cast<uintptr_t>(w_packed) & 3))
    return (int)cudaErrorInvalidValue;
  SmallConvParams p;
  p.in = reinterpret_cast<const __nv_bfloat16*>(in); p.in_pitch = in_pitch; p.in_coff = in_coff;
  p.in2 = reinterpret_cast<const __nv_bfloat16*>(in2); p.in2_pitch = in2_pitch; p.in2_coff = in2_coff; p.n_split_in = n_split_in;
  p.span = cin_span; p.nchunk = cdiv(cin_span, 64); p.cs = 64 * p.nchunk;
  p.w = reinterpret_cast<const __nv_bfloat16*>(w_packed); p.bias = bias; p.cout = cout; p.coutp = coutp;
  p.out = reinterpret_cast<__nv_bfloat16*>(out); p.out_pitch = out_pitch; p.out_coff = out_coff;
  p.out2 = reinterpret_cast<__nv_bfloat16*>(out2); p.out2_pitch = out2_pitch; p.out2_coff = out2_coff; p.n_split = n_split;
  p.N = N; p.H = H; p.W = W; p.stride = stride; p.slope = slope;
  p.Ho = (H - 1) / stride + 1; p.Wo = (W - 1) / stride + 1;
  static int num_sms = 0;
  if (!num_sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev); if (num_sms <= 0) num_sms = 148; }
  // tile height 4 or 8 rows (2 x MTW), channel slice 16 or 32: the finest cut whose grid still fits the machine once,
  // else the coarsest (fewest re-reads of patch and weights); then the K loop is split over a cluster of S CTAs while
  // two CTAs per SM can still hold the whole grid
  struct Cfg { int mtw, ntl; };
  const Cfg cfgs[2] = {{2, 2}, {4, 2}};
  static const int forced = []() { const char* e = getenv("PWC_MMA_SMALL_CFG"); return e ? atoi(e) : -1; }();
  static const int forced_s = []() { const char* e = getenv("PWC_MMA_SMALL_S"); return e ? atoi(e) : 0; }();
  auto grid_of = [&](int i) { return (long)N * cdiv(p.Ho, 2 * cfgs[i].mtw) * cdiv(p.Wo, 16) * cdiv(coutp, 8 * cfgs[i].ntl); };
  int pick = 1;
  for (int i = 0; i < 2; ++i)
    if (grid_of(i) <= num_sms) { pick = i; break; }
  if (forced >= 0 && forced < 2) pick = forced;
  const int mtw = cfgs[pick].mtw, ntl = cfgs[pick].ntl;
  int S = 1;                                            // any cluster size up to 8 (portable limit), one chunk per CTA at best
  while (S < 8 && S < p.nchunk && grid_of(pick) * (S + 1) <= 2L * num_sms) ++S;
  if (forced_s >= 1 && forced_s <= 8) S = forced_s <= p.nchunk ? forced_s : p.nchunk;
  p.S = S; p.nloc = cdiv(p.nchunk, S);
  p.sp = 128 * p.nloc + 16;
  p.wm = 2; p.wk = 4;                                   // 8 warps: 2 m-groups x 4 K-groups
  const int TH = p.wm * mtw;
  p.tiles_x = cdiv(p.Wo, 16); p.tiles_y = cdiv(p.Ho, TH); p.q = cdiv(coutp, 8 * ntl);
  p.ih = (TH - 1) * stride + 3; p.iw = 15 * stride + 3;
  const int patch = p.ih * p.iw * p.sp, red = p.wk * TH * 16 * 8 * ntl * 4;
  p.off_part = ((patch > red ? patch : red) + 127) & ~127;
  const int smem = p.off_part + TH * 16 * 8 * ntl * 4;
  if (smem > 227 * 1024) return (int)cudaErrorInvalidValue;
  const long grid = (long)N * p.tiles_y * p.tiles_x * p.q * S;
  if (grid > 0x7fffffffL) return (int)cudaErrorInvalidValue;
  cudaStream_t s = (cudaStream_t)stream;
  if (half_format()) return mtw == 2 ? launch_small<2, 2, true>(p, (int)grid, 256, smem, s) : launch_small<4, 2, true>(p, (int)grid, 256, smem, s);
  if (mtw == 2) return launch_small<2, 2, false>(p, (int)grid, 256, smem, s);
  return launch_small<4, 2, false>(p, (int)grid, 256, smem, s);
}
