// Fused bilinear backward-warp + validity mask + 81-displacement cost volume + LeakyReLU.
//
// Replaces, in ONE kernel (reference paths relative to /root/reference):
//   PWCDCNet.warp ................................ models/PWCNet.py:141-177
//   channels_first x2 + correlation_forward ...... models/correlation_package/correlation_cuda_kernel.cu:46-147
//   the three fill_(0) of the binding ............ models/correlation_package/correlation_cuda.cc:40-42
//   self.leakyRELU on the cost volume ............ models/PWCNet.py:199,214,228,242,258
//
// Decomposition ("diagonal outer product").  For one warped row r of image 2 and one column x,
//     acc[i][j] = sum_c f1[r-4+i][x][c] * w2[r][x-4+j][c],   i,j in 0..8
// is the cost-volume entry of pixel (y = r-4+i, x) at displacement (dy = 4-i, dx = j-4).  One lane owns one (r,x):
// a 9x9 outer product per channel, 81 FMA for 18 shared-memory operands -- every output is computed exactly once.
// A consumer warp is 32 consecutive x of one row r; a tile is ROWS consecutive rows r x 32 columns.
//
// NHWC kernel (the product path): persistent CTAs, warp-specialised.
//   * 4 producer warps build, per (tile, 32-channel chunk), one pipeline stage in shared memory:
//       - the f1 box (rows r0-4..r0+ROWS+3, 32 columns, 32 channels) by ONE TMA instruction (SWIZZLE_128B, zero fill
//         outside the image = the correlation's zero padding);
//       - the warped image-2 rows r0..r0+ROWS-1, columns x0-4..x0+35: bilinear gather of f2 with the mask folded into
//         the corner weights, written straight into shared memory (the warped features never exist in HBM).  The
//         sample coordinates / weights are computed once per tile into a small table.
//   * ROWS consumer warps run the outer products out of shared memory (LDS.128, XOR-swizzled 16-byte chunks =
//     the TMA 128B swizzle, conflict free) with the 81 accumulators in registers across chunks, then divide by C,
//     apply LeakyReLU and store 36-byte runs through a small per-warp staging buffer.
//   * two stages, mbarrier full/empty handshakes: the gather + TMA of the next tile overlap the FMAs of this one.
// NCHW kernel (API-compatible Correlation op, correlation_cuda.cc:10-87): one CTA per tile, all threads fill then
// compute; outputs are stored coalesced along x.
#include <stdlib.h>
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "tmap.cuh"
#include "../../include/pwc_b200.h"

namespace pwc {

constexpr int kMD = 4;
constexpr int kND = 9;                 // displacements per axis
constexpr int kTW = 32;                // columns per tile (lanes)
constexpr int kW2Cols = kTW + 2 * kMD; // 40
constexpr int kChunk = 32;             // channels per pipeline stage
constexpr int kProdWarps = 4;

struct WarpCorrParams {
  const void* f1; long f1_pitch;
  const void* f2; long f2_pitch;
  const float* flow; long flow_pitch;
  float flow_scale; int align_corners; float mask_thresh; float slope;
  void* out; long out_pitch;
  int B, C, H, W;
  int tiles_x, tiles_y, num_tiles;
  int f16;                                       // 16-bit features / output are IEEE half (pwc_set_half_format) instead of bf16
};

// swizzled float4 slot of (pixel index p within its array, 16B-chunk k)
__device__ __forceinline__ int swz(int p, int k) { return p * 8 + (k ^ (p & 7)); }

__device__ __forceinline__ void fma4(float4& a, float w, const float4& v) {
  a.x = fmaf(w, v.x, a.x); a.y = fmaf(w, v.y, a.y); a.z = fmaf(w, v.z, a.z); a.w = fmaf(w, v.w, a.w);
}

// Bilinear sample position, corner weights (nw,ne,sw,se) with the zeros-padding and the 0.9999 mask folded in
// (weight 0 <=> corner contributes nothing), and clamped corner coordinates.  PWCNet.py:159-177 + ATen GridSampler.
__device__ __forceinline__ void warp_entry(const WarpCorrParams& p, float u, float vv, int xx, int yy,
                                           float4& w, uint2& xy) {
  const int W = p.W, H = p.H;
  // flo = up_flow*scale; vgrid = grid + flo; 2*v/(W-1) - 1: one rounding per op, as the reference's tensor ops
  const float vx = __fadd_rn((float)xx, __fmul_rn(u, p.flow_scale));
  const float vy = __fadd_rn((float)yy, __fmul_rn(vv, p.flow_scale));
  const float gx = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, vx), (float)max(W - 1, 1)), 1.0f);
  const float gy = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, vy), (float)max(H - 1, 1)), 1.0f);
  float ix, iy;   // ATen grid_sampler_unnormalize (GridSampler.h:27-35); /2 == *0.5 exactly
  if (p.align_corners) {
    ix = __fmul_rn(__fmul_rn(__fadd_rn(gx, 1.f), 0.5f), (float)(W - 1));
    iy = __fmul_rn(__fmul_rn(__fadd_rn(gy, 1.f), 0.5f), (float)(H - 1));
  } else {
    ix = __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(gx, 1.f), (float)W), 1.f), 0.5f);
    iy = __fmul_rn(__fsub_rn(__fmul_rn(__fadd_rn(gy, 1.f), (float)H), 1.f), 0.5f);
  }
  const float fx0 = floorf(ix), fy0 = floorf(iy);
  const float fx1 = fx0 + 1.f, fy1 = fy0 + 1.f;
  float wnw = __fmul_rn(fx1 - ix, fy1 - iy), wne = __fmul_rn(ix - fx0, fy1 - iy);
  float wsw = __fmul_rn(fx1 - ix, iy - fy0), wse = __fmul_rn(ix - fx0, iy - fy0);
  const bool okx0 = (fx0 >= 0.f && fx0 <= (float)(W - 1)), okx1 = (fx1 >= 0.f && fx1 <= (float)(W - 1));
  const bool oky0 = (fy0 >= 0.f && fy0 <= (float)(H - 1)), oky1 = (fy1 >= 0.f && fy1 <= (float)(H - 1));
  if (!(okx0 && oky0)) wnw = 0.f;
  if (!(okx1 && oky0)) wne = 0.f;
  if (!(okx0 && oky1)) wsw = 0.f;
  if (!(okx1 && oky1)) wse = 0.f;
  float m = 0.f;                       // grid_sample(ones): nw, ne, sw, se order
  m = __fadd_rn(m, wnw); m = __fadd_rn(m, wne); m = __fadd_rn(m, wsw); m = __fadd_rn(m, wse);
  if ((m < p.mask_thresh) || !(m > 0.f)) { wnw = wne = wsw = wse = 0.f; }   // PWCNet.py:174-177 (NaN -> dropped)
  const int x0 = (int)fminf(fmaxf(fx0, 0.f), (float)(W - 1)), x1 = (int)fminf(fmaxf(fx1, 0.f), (float)(W - 1));
  const int y0 = (int)fminf(fmaxf(fy0, 0.f), (float)(H - 1)), y1 = (int)fminf(fmaxf(fy1, 0.f), (float)(H - 1));
  w = make_float4(wnw, wne, wsw, wse);
  xy = make_uint2((uint32_t)x0 | ((uint32_t)x1 << 16), (uint32_t)y0 | ((uint32_t)y1 << 16));
}

// correctly rounded acc / C from rc = RN(1/C) (Markstein): same value as the `/ nelems` of correlation_cuda_kernel.cu:143
__device__ __forceinline__ float div_by(float a, float fC, float rc) {
  const float q = a * rc;
  const float r = fmaf(-q, fC, a);
  return fmaf(r, rc, q);
}

// 9x9 outer product over the live 16-byte channel chunks of one stage
__device__ __forceinline__ void outer_products(float (&acc)[kND][kND], const float4* __restrict__ f1s,
                                               const float4* __restrict__ w2s, int f1_pix0, int w2_pix0, int nk) {
#pragma unroll 1
  for (int k = 0; k < nk; ++k) {
    float4 wv[kND];
#pragma unroll
    for (int j = 0; j < kND; ++j) wv[j] = w2s[swz(w2_pix0 + j, k)];
#pragma unroll
    for (int i = 0; i < kND; ++i) {
      const float4 a = f1s[swz(f1_pix0 + i * kTW, k)];
#pragma unroll
      for (int j = 0; j < kND; ++j) {
        float s = acc[i][j];
        s = fmaf(a.x, wv[j].x, s); s = fmaf(a.y, wv[j].y, s);
        s = fmaf(a.z, wv[j].z, s); s = fmaf(a.w, wv[j].w, s);
        acc[i][j] = s;
      }
    }
  }
}

// =====================================================================================================================
// NHWC product kernel
// =====================================================================================================================
template <int ROWS> struct Geo {
  static constexpr int kThreads = (ROWS + kProdWarps) * 32;
  static constexpr int kF1Rows = ROWS + 2 * kMD;
  static constexpr int kF1Pix = kF1Rows * kTW;
  static constexpr int kW2Pix = ROWS * kW2Cols;
  static constexpr int kF1Bytes = kF1Pix * kChunk * 4;          // multiple of 4096
  static constexpr int kW2Bytes = kW2Pix * kChunk * 4;          // multiple of 5120 * ROWS (ROWS even -> 1024-aligned)
  static constexpr int kStageBytes = kF1Bytes + kW2Bytes;
  static constexpr int kStages = 2;
  static constexpr int kStgFloats = kTW * kND;                  // per consumer warp: one output row of 9 channels
  static constexpr int kStgBytes = ROWS * kStgFloats * 4;
  static constexpr int kTabBytes = kW2Pix * 24;
  static constexpr int kSmemBytes = kStages * kStageBytes + kStgBytes + kTabBytes + 64 + 1024;
};

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

template <typename T, int ROWS, bool TMA>
__global__ void __launch_bounds__(Geo<ROWS>::kThreads, 1)
warp_corr_nhwc_kernel(const __grid_constant__ CUtensorMap tmF1, const WarpCorrParams p) {
  using G = Geo<ROWS>;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t addr0 = ptx::smem_addr(smem_raw);
  const uint32_t base = (addr0 + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (base - addr0);
  float* stg_all = reinterpret_cast<float*>(sm + G::kStages * G::kStageBytes);
  float4* tabw = reinterpret_cast<float4*>(sm + G::kStages * G::kStageBytes + G::kStgBytes);
  uint2* tabxy = reinterpret_cast<uint2*>(tabw + G::kW2Pix);
  const uint32_t bar0 = base + G::kStages * G::kStageBytes + G::kStgBytes + G::kTabBytes;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (G::kStages + s); };

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int C = p.C, H = p.H, W = p.W;
  const int nchunks = cdiv(C, kChunk);

  if (tid == 0) {
    if (TMA) ptx::prefetch_tmap(&tmF1);
    for (int s = 0; s < G::kStages; ++s) {
      ptx::mbar_init(full_bar(s), kProdWarps + (TMA ? 1 : 0));
      ptx::mbar_init(empty_bar(s), ROWS);
    }
    ptx::fence_barrier_init();
    ptx::fence_proxy_async();
  }
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();

  if (warp >= ROWS) {
    // ============================== producers ==============================
    const int pw = warp - ROWS, ptid = pw * 32 + lane;
    const T* f1 = reinterpret_cast<const T*>(p.f1);
    const T* f2 = reinterpret_cast<const T*>(p.f2);
    const int k = lane & 7, sub = lane >> 3;
    int s = 0; uint32_t ph = 0;
    for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x) {
      const int per_img = p.tiles_x * p.tiles_y;
      const int n = t / per_img, rr = t - n * per_img;
      const int x0 = (rr % p.tiles_x) * kTW, r0 = (rr / p.tiles_x) * ROWS - kMD;
      // ---- L2 prefetch of the NEXT tile's flow rows and the image-2 rows its samples most likely fall in ----
      {
        const int t2 = t + gridDim.x;
        if (t2 < p.num_tiles) {
          const int n2 = t2 / per_img, rr2 = t2 - n2 * per_img;
          const int x02 = (rr2 % p.tiles_x) * kTW, r02 = (rr2 / p.tiles_x) * ROWS - kMD;
          const int xs = max(x02 - 12, 0), xe = min(x02 + kTW + 12, W);
          if (ptid < ROWS + 8) {
            const int y = r02 - 4 + ptid;
            if (y >= 0 && y < H) {
              const uintptr_t a = reinterpret_cast<uintptr_t>(f2 + (((long)n2 * H + y) * W + xs) * p.f2_pitch);
              const uintptr_t e = reinterpret_cast<uintptr_t>(f2 + (((long)n2 * H + y) * W + xe - 1) * p.f2_pitch + C);
              const uintptr_t a16 = a & ~(uintptr_t)15;
              ptx::prefetch_l2_bulk(reinterpret_cast<const void*>(a16), (uint32_t)((e - a16 + 15) & ~(uintptr_t)15));
            }
          } else if (p.flow != nullptr && ptid >= 32 && ptid < 32 + ROWS) {
            const int y = r02 + (ptid - 32);
            if (y >= 0 && y < H) {
              const int fs = max(x02 - kMD, 0), fe = min(x02 + kTW + kMD, W);
              const uintptr_t a = reinterpret_cast<uintptr_t>(p.flow + (((long)n2 * H + y) * W + fs) * p.flow_pitch);
              const uintptr_t e = reinterpret_cast<uintptr_t>(p.flow + (((long)n2 * H + y) * W + fe - 1) * p.flow_pitch + 2);
              const uintptr_t a16 = a & ~(uintptr_t)15;
              ptx::prefetch_l2_bulk(reinterpret_cast<const void*>(a16), (uint32_t)((e - a16 + 15) & ~(uintptr_t)15));
            }
          }
        }
      }
      // ---- per-tile table of sample coordinates / masked corner weights ----
      named_bar_sync(1, kProdWarps * 32);          // everyone is done reading the previous tile's table
      for (int e = ptid; e < G::kW2Pix; e += kProdWarps * 32) {
        const int q = e % kW2Cols, wr = e / kW2Cols;
        const int yy = r0 + wr, xx = x0 - kMD + q;
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        uint2 xy = make_uint2(0u, 0u);
        if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
          if (p.flow == nullptr) {                 // level 6: no warp (PWCNet.py:198)
            w.x = 1.f;
            xy = make_uint2((uint32_t)xx | ((uint32_t)xx << 16), (uint32_t)yy | ((uint32_t)yy << 16));
          } else {
            const float2 fl = *reinterpret_cast<const float2*>(p.flow + (((long)n * H + yy) * W + xx) * p.flow_pitch);
            warp_entry(p, fl.x, fl.y, xx, yy, w, xy);
          }
        }
        tabw[e] = w; tabxy[e] = xy;
      }
      named_bar_sync(1, kProdWarps * 32);
      const T* f2n = f2 + (long)n * H * W * p.f2_pitch;
      for (int ch = 0; ch < nchunks; ++ch) {
        ptx::mbar_wait(empty_bar(s), ph ^ 1);
        float4* f1s = reinterpret_cast<float4*>(sm + s * G::kStageBytes);
        float4* w2s = reinterpret_cast<float4*>(sm + s * G::kStageBytes + G::kF1Bytes);
        if (TMA) {
          if (ptid == 0) {
            ptx::mbar_expect_tx(full_bar(s), G::kF1Bytes);
            ptx::tma_load_4d(base + s * G::kStageBytes, &tmF1, full_bar(s), ch * kChunk, x0, r0 - kMD, n);
          }
        } else {
#pragma unroll 8
          for (int it = ptid; it < G::kF1Pix * 8; it += kProdWarps * 32) {
            const int kk = it & 7, px = (it >> 3) & (kTW - 1), row = it >> 8;
            const int y = r0 - kMD + row, x = x0 + px, c = ch * kChunk + 4 * kk;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (y >= 0 && y < H && x < W && c < C) v = ld4(f1 + (((long)n * H + y) * W + x) * p.f1_pitch + c);
            f1s[swz(row * kTW + px, kk)] = v;
          }
        }
        // ---- bilinear gather: 8 lanes cover the 32 channels (128 B) of one pixel, 4 pixels per warp instruction ----
        const int c = ch * kChunk + 4 * k;
        const bool cok = c < C;
        const T* f2c = f2n + c;
        constexpr int kIters = G::kW2Pix / (kProdWarps * 4);
        constexpr int kUnroll = 5;
#pragma unroll kUnroll
        for (int itp = 0; itp < kIters; ++itp) {
          const int pix = itp * (kProdWarps * 4) + pw * 4 + sub;
          const float4 w = tabw[pix];
          const uint2 xy = tabxy[pix];
          const int xa = xy.x & 0xffff, xb = xy.x >> 16, ya = xy.y & 0xffff, yb = xy.y >> 16;
          const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
          const float4 vnw = (cok && w.x != 0.f) ? ld4(f2c + ((long)ya * W + xa) * p.f2_pitch) : z;
          const float4 vne = (cok && w.y != 0.f) ? ld4(f2c + ((long)ya * W + xb) * p.f2_pitch) : z;
          const float4 vsw = (cok && w.z != 0.f) ? ld4(f2c + ((long)yb * W + xa) * p.f2_pitch) : z;
          const float4 vse = (cok && w.w != 0.f) ? ld4(f2c + ((long)yb * W + xb) * p.f2_pitch) : z;
          float4 v = z;
          fma4(v, w.x, vnw); fma4(v, w.y, vne); fma4(v, w.z, vsw); fma4(v, w.w, vse);
          w2s[swz(pix, k)] = v;
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(full_bar(s));
        if (++s == G::kStages) { s = 0; ph ^= 1; }
      }
    }
  } else {
    // ============================== consumers ==============================
    const float fC = (float)C, rc = __frcp_rn(fC);
    T* out = reinterpret_cast<T*>(p.out);
    float* stg = stg_all + warp * G::kStgFloats;
    // epilogue store pattern: element e = lane + 32 t of a [32 px][9] row block -> (px, j)
    int eoff[kND];
#pragma unroll
    for (int tt = 0; tt < kND; ++tt) {
      const int e = lane + 32 * tt, px = e / kND, j = e - px * kND;
      eoff[tt] = px * (int)p.out_pitch + j;
    }
    int s = 0; uint32_t ph = 0;
    for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x) {
      const int per_img = p.tiles_x * p.tiles_y;
      const int n = t / per_img, rr = t - n * per_img;
      const int x0 = (rr % p.tiles_x) * kTW, r0 = (rr / p.tiles_x) * ROWS - kMD;
      const int r = r0 + warp;
      const bool row_live = (r >= 0 && r < H);     // warp-uniform: rows outside the image contribute zeros
      float acc[kND][kND];
#pragma unroll
      for (int i = 0; i < kND; ++i)
#pragma unroll
        for (int j = 0; j < kND; ++j) acc[i][j] = 0.f;
      for (int ch = 0; ch < nchunks; ++ch) {
        ptx::mbar_wait(full_bar(s), ph);
        if (row_live) {
          const float4* f1s = reinterpret_cast<const float4*>(sm + s * G::kStageBytes);
          const float4* w2s = reinterpret_cast<const float4*>(sm + s * G::kStageBytes + G::kF1Bytes);
          const int nk = min(kChunk, C - ch * kChunk + 3) >> 2;     // 16B chunks holding live channels
          outer_products(acc, f1s, w2s, warp * kTW + lane, warp * kW2Cols + lane, nk);
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(empty_bar(s));
        if (++s == G::kStages) { s = 0; ph ^= 1; }
      }
      // ---- epilogue: /C (correlation_cuda_kernel.cu:143), LeakyReLU, 36-byte runs via the per-warp staging row ----
      const int ncol9 = min(kTW, W - x0) * kND;
#pragma unroll
      for (int i = 0; i < kND; ++i) {
        const int y = r - kMD + i;
        if (y < 0 || y >= H) continue;             // warp-uniform
#pragma unroll
        for (int j = 0; j < kND; ++j) stg[lane * kND + j] = lrelu(div_by(acc[i][j], fC, rc), p.slope);
        __syncwarp();
        T* orow = out + (((long)n * H + y) * W + x0) * p.out_pitch + (kND - 1 - i) * kND;
#pragma unroll
        for (int tt = 0; tt < kND; ++tt) {
          const int e = lane + 32 * tt;
          if (e < ncol9) st1(orow + eoff[tt], stg[e]);
        }
        __syncwarp();
      }
    }
  }
}

template <typename T, int ROWS, bool TMA>
static int launch_nhwc(WarpCorrParams p, cudaStream_t s, int num_sms) {
  using G = Geo<ROWS>;
  auto kern = warp_corr_nhwc_kernel<T, ROWS, TMA>;
  static int attr_dev = -1;            // per template instance; re-applied when the current device changes
  int dev = 0; cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, G::kSmemBytes);
    if (e != cudaSuccess) return (int)e;
    attr_dev = dev;
  }
  p.tiles_x = cdiv(p.W, kTW); p.tiles_y = cdiv(p.H + 2 * kMD, ROWS); p.num_tiles = p.tiles_x * p.tiles_y * p.B;
  CUtensorMap tm;
  memset(&tm, 0, sizeof(tm));
  if (TMA) {
    int e = make_tmap_nhwc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, p.f1, p.f1_pitch, p.C, p.W, p.H, p.B,
                           kChunk, kTW, G::kF1Rows, 1, 1, CU_TENSOR_MAP_SWIZZLE_128B);
    if (e) return e;
  }
  const int per_sm = (227 * 1024) / G::kSmemBytes;     // resident CTAs the shared memory allows (registers: 168/thread)
  const int cap = num_sms * (per_sm < 1 ? 1 : (per_sm > 2 ? 2 : per_sm));
  const int grid = p.num_tiles < cap ? p.num_tiles : cap;
  return (int)launch_pdl(kern, dim3(grid), dim3(G::kThreads), G::kSmemBytes, s, tm, p);
}

template <typename T, bool TMA>
static int dispatch_rows(const WarpCorrParams& p, cudaStream_t s) {
  static int num_sms = 0;
  if (!num_sms) {
    int dev = 0; cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (num_sms <= 0) num_sms = 148;
  }
  auto tiles = [&](int rows) { return (long)cdiv(p.W, kTW) * cdiv(p.H + 2 * kMD, rows) * p.B; };
  // largest row block that still gives (nearly) every SM a tile; the small pyramid levels take the smallest
  if (tiles(8) * 10 >= (long)num_sms * 9) return launch_nhwc<T, 8, TMA>(p, s, num_sms);
  if (tiles(4) * 10 >= (long)num_sms * 9) return launch_nhwc<T, 4, TMA>(p, s, num_sms);
  return launch_nhwc<T, 2, TMA>(p, s, num_sms);
}

// =====================================================================================================================
// NCHW compat kernel (fp32): one CTA per tile of 8 warped rows x 32 columns
// =====================================================================================================================
constexpr int kRowsC = 8;
constexpr int kF1PixC = (kRowsC + 2 * kMD) * kTW;   // 512
constexpr int kW2PixC = kRowsC * kW2Cols;           // 320
constexpr int kSmemC = (kF1PixC + kW2PixC) * kChunk * 4;

__global__ void __launch_bounds__(kRowsC * 32, 1)
warp_corr_nchw_kernel(const WarpCorrParams p) {
  extern __shared__ __align__(1024) float smem[];
  float4* f1s = reinterpret_cast<float4*>(smem);                    // [kF1PixC][8 chunks]
  float4* w2s = reinterpret_cast<float4*>(smem) + kF1PixC * 8;      // [kW2PixC][8 chunks]
  float* f1f = smem;
  float* w2f = smem + kF1PixC * kChunk;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int x0 = blockIdx.x * kTW;
  const int r0 = blockIdx.y * kRowsC - kMD;
  const int n = blockIdx.z;
  const int C = p.C, H = p.H, W = p.W;
  const long cs = (long)H * W;
  const float* f1 = reinterpret_cast<const float*>(p.f1) + (long)n * C * cs;
  const float* f2 = reinterpret_cast<const float*>(p.f2) + (long)n * C * cs;

  float acc[kND][kND];
#pragma unroll
  for (int i = 0; i < kND; ++i)
#pragma unroll
    for (int j = 0; j < kND; ++j) acc[i][j] = 0.f;
  const int r = r0 + warp;
  const bool row_live = (r >= 0 && r < H);

  // this thread's warped pixels (fixed across channel chunks): entries e = tid, tid+256 of the 320-pixel w2 tile
  float4 tw[2]; uint2 txy[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int e = tid + 256 * h;
    tw[h] = make_float4(0.f, 0.f, 0.f, 0.f); txy[h] = make_uint2(0u, 0u);
    if (e < kW2PixC) {
      const int q = e % kW2Cols, wr = e / kW2Cols;
      const int yy = r0 + wr, xx = x0 - kMD + q;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
        if (p.flow == nullptr) {
          tw[h].x = 1.f;
          txy[h] = make_uint2((uint32_t)xx | ((uint32_t)xx << 16), (uint32_t)yy | ((uint32_t)yy << 16));
        } else {
          const float u = p.flow[((long)n * 2 + 0) * cs + (long)yy * W + xx];
          const float vv = p.flow[((long)n * 2 + 1) * cs + (long)yy * W + xx];
          warp_entry(p, u, vv, xx, yy, tw[h], txy[h]);
        }
      }
    }
  }

  for (int c0 = 0; c0 < C; c0 += kChunk) {
    // f1 box: coalesced along x per channel, transposed into the [pixel][channel] tile
    for (int it = tid; it < kF1PixC * kChunk; it += kRowsC * 32) {
      const int px = it & (kTW - 1), row = (it >> 5) % (kRowsC + 2 * kMD), cc = it / kF1PixC;
      const int y = r0 - kMD + row, x = x0 + px, c = c0 + cc;
      float v = 0.f;
      if (y >= 0 && y < H && x < W && c < C) v = f1[(long)c * cs + (long)y * W + x];
      const int pp = row * kTW + px;
      f1f[swz(pp, cc >> 2) * 4 + (cc & 3)] = v;
    }
    // warped image-2 rows
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int e = tid + 256 * h;
      if (e < kW2PixC) {
        const int xa = txy[h].x & 0xffff, xb = txy[h].x >> 16, ya = txy[h].y & 0xffff, yb = txy[h].y >> 16;
        const float4 w = tw[h];
        for (int cc = 0; cc < kChunk; ++cc) {
          const int c = c0 + cc;
          float v = 0.f;
          if (c < C) {
            const float* fc = f2 + (long)c * cs;
            if (w.x != 0.f) v = fmaf(w.x, fc[(long)ya * W + xa], v);
            if (w.y != 0.f) v = fmaf(w.y, fc[(long)ya * W + xb], v);
            if (w.z != 0.f) v = fmaf(w.z, fc[(long)yb * W + xa], v);
            if (w.w != 0.f) v = fmaf(w.w, fc[(long)yb * W + xb], v);
          }
          w2f[swz(e, cc >> 2) * 4 + (cc & 3)] = v;
        }
      }
    }
    __syncthreads();
    if (row_live) {
      const int nk = min(kChunk, C - c0 + 3) >> 2;
      outer_products(acc, f1s, w2s, warp * kTW + lane, warp * kW2Cols + lane, nk);
    }
    __syncthreads();
  }

  const float fC = (float)C, rc = __frcp_rn(fC);
  float* out = reinterpret_cast<float*>(p.out);
  const int x = x0 + lane;
#pragma unroll
  for (int i = 0; i < kND; ++i) {
    const int y = r - kMD + i;
    if (y < 0 || y >= H || x >= W) continue;
#pragma unroll
    for (int j = 0; j < kND; ++j) {
      const int tc = (kND - 1 - i) * kND + j;
      out[(((long)n * (kND * kND) + tc) * H + y) * W + x] = lrelu(div_by(acc[i][j], fC, rc), p.slope);
    }
  }
}

static int launch_nchw(const WarpCorrParams& p, cudaStream_t s) {
  static int attr_dev = -1;
  int dev = 0; cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t e = cudaFuncSetAttribute(warp_corr_nchw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemC);
    if (e != cudaSuccess) return (int)e;
    attr_dev = dev;
  }
  dim3 grid(cdiv(p.W, kTW), cdiv(p.H + 2 * kMD, kRowsC), p.B);
  warp_corr_nchw_kernel<<<grid, kRowsC * 32, kSmemC, s>>>(p);
  return (int)cudaGetLastError();
}

}  // namespace pwc

#include "warp_corr_tc.cuh"
#include "warp_corr_small.cuh"
#include "warp_corr_mma.cuh"

extern "C" int pwc_warp_corr_lrelu_fwd(const void* f1, long f1_pitch, const void* f2, long f2_pitch,
                                       const float* flow, long flow_pitch, float flow_scale,
                                       int align_corners, float mask_thresh, float slope,
                                       void* out, long out_pitch,
                                       int B, int C, int H, int W, int md,
                                       int layout, int dtype, void* stream) {
  using namespace pwc;
  if (md != kMD || B <= 0 || C <= 0 || H <= 0 || W <= 0 || B > 65535 || H > 65535 || W > 65535)
    return (int)cudaErrorInvalidValue;
  WarpCorrParams p{f1, f1_pitch, f2, f2_pitch, flow, flow_pitch, flow_scale, align_corners, mask_thresh, slope,
                   out, out_pitch, B, C, H, W, 0, 0, 0, (dtype == PWC_BF16 && half_format()) ? 1 : 0};
  cudaStream_t s = (cudaStream_t)stream;
  if (layout == PWC_LAYOUT_NCHW) {
    if (dtype != PWC_F32) return (int)cudaErrorInvalidValue;
    return launch_nchw(p, s);
  }
  if (layout != PWC_LAYOUT_NHWC) return (int)cudaErrorInvalidValue;
  if (C % 4 || f1_pitch % 4 || f2_pitch % 4 || (flow && (flow_pitch % 2))) return (int)cudaErrorInvalidValue;
  if (out_pitch * 32 > 0x7fffffffL) return (int)cudaErrorInvalidValue;
  if (dtype == PWC_BF16) {
    // warp-level tensor-core kernel (warp_corr_mma.cuh): PWC_CORR_MMA = 0 off | 1 wide-feature levels (C > 64) | 2 all
    static const int use_mma = []() { const char* e = getenv("PWC_CORR_MMA"); return e ? atoi(e) : 1; }();
    const bool mma_ok = use_mma && (use_mma >= 2 || C > 64) && C >= 16 && C <= 256 && f1_pitch % 8 == 0 && f2_pitch % 8 == 0 &&
                        ((reinterpret_cast<uintptr_t>(f1) | reinterpret_cast<uintptr_t>(f2)) & 15) == 0 &&
                        (reinterpret_cast<uintptr_t>(out) & 15) == 0 && out_pitch % 8 == 0 && slope >= 0.f && slope <= 1.f &&
                        get_encode_tiled() != nullptr;
    if (mma_ok) return launch_warp_corr_mma(p, s);
    // thin-feature levels (C <= 64: pyramid levels 2 and 3) run the cost volume on the tensor cores
    static const int use_tc = []() { const char* e = getenv("PWC_CORR_TC"); return e ? atoi(e) : 1; }();
    const bool tc_ok = use_tc && (C == 16 || C == 32 || C == 64) && f1_pitch % 8 == 0 && f2_pitch % 8 == 0 &&
                       ((reinterpret_cast<uintptr_t>(f1) | reinterpret_cast<uintptr_t>(f2)) & 15) == 0 &&
                       (reinterpret_cast<uintptr_t>(out) & 15) == 0 && out_pitch % 8 == 0 && slope >= 0.f && slope <= 1.f &&
                       get_encode_tiled() != nullptr;
    if (tc_ok) return launch_warp_corr_tc(p, s);
    // wide-feature (small) levels 4-6: latency-oriented kernel, one warp per row displacement
    static const int use_small = []() { const char* e = getenv("PWC_CORR_SMALL"); return e ? atoi(e) : 1; }();
    // (each CTA re-gathers the 9 warped rows it needs: only worth it while the grid fits the machine once)
    const bool small_ok = use_small && !p.f16 && C > 64 && C % 4 == 0 && C <= 200 && (long)cdiv(W, 32) * H * B <= 160 && f1_pitch % 4 == 0 && f2_pitch % 4 == 0 &&
                          ((reinterpret_cast<uintptr_t>(f1) | reinterpret_cast<uintptr_t>(f2)) & 7) == 0 &&
                          (reinterpret_cast<uintptr_t>(out) & 15) == 0 && out_pitch % 8 == 0 && slope >= 0.f &&
                          slope <= 1.f && get_encode_tiled() != nullptr;
    if (small_ok) return launch_warp_corr_small(p, s);
    return p.f16 ? dispatch_rows<__half, false>(p, s) : dispatch_rows<__nv_bfloat16, false>(p, s);
  }
  if (dtype != PWC_F32) return (int)cudaErrorInvalidValue;
  // TMA needs a 16-byte aligned base and 16-byte multiples as strides; otherwise the producers load the f1 box themselves
  const bool tma_ok = ((reinterpret_cast<uintptr_t>(f1) & 15) == 0) && get_encode_tiled() != nullptr;
  return tma_ok ? dispatch_rows<float, true>(p, s) : dispatch_rows<float, false>(p, s);
}

// API-compatible op: correlation_cuda.forward for pad=md=4, k=1, strides 1 (correlation_cuda.cc:10-87).
extern "C" int pwc_corr_fwd(const void* f1, const void* f2, void* out, int B, int C, int H, int W, int md, void* stream) {
  return pwc_warp_corr_lrelu_fwd(f1, 0, f2, 0, nullptr, 0, 1.0f, 0, 0.9999f, 1.0f, out, 0, B, C, H, W, md,
                                 PWC_LAYOUT_NCHW, PWC_F32, stream);
}

namespace pwc {

// Standalone PWCDCNet.warp (PWCNet.py:141-177) for callers of the reference's method; the forward itself never
// materialises warped features (they are gathered inside the cost-volume kernels).  One thread = one output pixel: the
// sample position / corner weights / mask once (warp_entry), then a channel loop of 4 gathers; stores coalesced along x.
__global__ void __launch_bounds__(256) warp_nchw_kernel(WarpCorrParams p, const float* __restrict__ x, float* __restrict__ out) {
  const long hw = (long)p.H * p.W, total = (long)p.B * hw;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int n = (int)(i / hw); const long r = i - (long)n * hw;
    const int yy = (int)(r / p.W), xx = (int)(r - (long)yy * p.W);
    const float u = p.flow[((long)n * 2) * hw + r], v = p.flow[((long)n * 2 + 1) * hw + r];
    float4 w; uint2 xy;
    warp_entry(p, u, v, xx, yy, w, xy);
    const int x0 = xy.x & 0xffff, x1 = xy.x >> 16, y0 = xy.y & 0xffff, y1 = xy.y >> 16;
    const long o00 = (long)y0 * p.W + x0, o01 = (long)y0 * p.W + x1, o10 = (long)y1 * p.W + x0, o11 = (long)y1 * p.W + x1;
    const float* src = x + (long)n * p.C * hw;
    float* dst = out + (long)n * p.C * hw + r;
    for (int c = 0; c < p.C; ++c, src += hw, dst += hw) {
      // ATen grid_sampler_2d: out = nw_val*nw + ne_val*ne + sw_val*sw + se_val*se, accumulated in that order; corners
      // with weight 0 (outside / masked) are not read by ATen, so a NaN there must not leak: skip them the same way
      float a = 0.f;
      if (w.x != 0.f) a = fmaf(src[o00], w.x, a);
      if (w.y != 0.f) a = fmaf(src[o01], w.y, a);
      if (w.z != 0.f) a = fmaf(src[o10], w.z, a);
      if (w.w != 0.f) a = fmaf(src[o11], w.w, a);
      *dst = a;
    }
  }
}

// Correlation backward (correlation_cuda_kernel.cu:150-334, 430-564) for the PWC-Net configuration, NCHW fp32:
//   gin1[n,c,y,x] = (1/C) sum_{dy,dx} gout[n,tc,y,x]       * f2[n,c,y+dy,x+dx]
//   gin2[n,c,y,x] = (1/C) sum_{dy,dx} gout[n,tc,y-dy,x-dx] * f1[n,c,y-dy,x-dx]          tc = (dy+4)*9 + dx+4
// (terms whose displaced pixel is outside the image are 0).  Both are "81 coefficients per pixel, independent of c, times
// a 9x9 neighbourhood of the other feature map": SIGN = +1 / -1 selects which.  One thread = one pixel; the 81
// coefficients live in registers; per channel 81 loads that are coalesced along x and served by L1.  No padded NHWC
// copies (rInput1/rInput2 of correlation_cuda.cc:95-101), no atomics: deterministic.
template <int SIGN>
__global__ void __launch_bounds__(128) corr_bwd_kernel(const float* __restrict__ feat, const float* __restrict__ gout,
                                                       float* __restrict__ gin, int B, int C, int H, int W) {
  const long hw = (long)H * W;
  const int xx = blockIdx.x * 32 + (threadIdx.x & 31), yy = blockIdx.y * 4 + (threadIdx.x >> 5), n = blockIdx.z;
  if (xx >= W || yy >= H) return;
  float g[kND * kND];
  const float* go = gout + (long)n * (kND * kND) * hw;
#pragma unroll
  for (int dy = -kMD; dy <= kMD; ++dy)
#pragma unroll
    for (int dx = -kMD; dx <= kMD; ++dx) {
      const int t = (dy + kMD) * kND + dx + kMD;
      const int sy = yy + SIGN * dy, sx = xx + SIGN * dx;          // the pixel of the other map this term multiplies
      const bool in = sy >= 0 && sy < H && sx >= 0 && sx < W;
      // gin1: coefficient at the own pixel; gin2: at the displaced pixel (= where the forward read f1)
      g[t] = in ? go[(long)t * hw + (SIGN > 0 ? (long)yy * W + xx : (long)sy * W + sx)] : 0.f;
    }
  const float fC = (float)C, rc = __frcp_rn(fC);
  const float* f = feat + (long)n * C * hw;
  float* dst = gin + (long)n * C * hw + (long)yy * W + xx;
  for (int c = 0; c < C; ++c, f += hw, dst += hw) {
    float a = 0.f;
#pragma unroll
    for (int dy = -kMD; dy <= kMD; ++dy) {
      const int sy = yy + SIGN * dy;
      if (sy < 0 || sy >= H) continue;
      const float* row = f + (long)sy * W;
#pragma unroll
      for (int dx = -kMD; dx <= kMD; ++dx) {
        const int sx = xx + SIGN * dx;
        if (sx >= 0 && sx < W) a = fmaf(g[(dy + kMD) * kND + dx + kMD], __ldg(row + sx), a);
      }
    }
    *dst = div_by(a, fC, rc);
  }
}

}  // namespace pwc

extern "C" int pwc_warp_fwd(const float* x, const float* flow, float* out, int B, int C, int H, int W,
                            float flow_scale, int align_corners, float mask_thresh, void* stream) {
  using namespace pwc;
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || H > 65535 || W > 65535 || !x || !flow || !out) return (int)cudaErrorInvalidValue;
  WarpCorrParams p{x, 0, x, 0, flow, 0, flow_scale, align_corners, mask_thresh, 1.0f, out, 0, B, C, H, W, 0, 0, 0};
  const long total = (long)B * H * W;
  const long blocks = (total + 255) / 256;
  warp_nchw_kernel<<<(unsigned)(blocks < 148L * 16 ? blocks : 148L * 16), 256, 0, (cudaStream_t)stream>>>(p, x, out);
  return (int)cudaGetLastError();
}

extern "C" int pwc_corr_bwd(const float* f1, const float* f2, const float* grad_out, float* grad_f1, float* grad_f2,
                            int B, int C, int H, int W, int md, void* stream) {
  using namespace pwc;
  if (md != kMD || B <= 0 || C <= 0 || H <= 0 || W <= 0 || B > 65535 || (H + 3) / 4 > 65535 || !f1 || !f2 || !grad_out)
    return (int)cudaErrorInvalidValue;
  const dim3 grid((W + 31) / 32, (H + 3) / 4, B);
  if (grad_f1) corr_bwd_kernel<1><<<grid, 128, 0, (cudaStream_t)stream>>>(f2, grad_out, grad_f1, B, C, H, W);
  if (grad_f2) corr_bwd_kernel<-1><<<grid, 128, 0, (cudaStream_t)stream>>>(f1, grad_out, grad_f2, B, C, H, W);
  return (int)cudaGetLastError();
}

// debug: timeline marks of the tensor-core cost-volume kernel (CTA 0), see tools/trace_corr.py
extern "C" int pwc_debug_corr_trace(unsigned long long* host_dst, int count) {
  if (!host_dst || count < 0 || count > 3 * 64 * 4) return (int)cudaErrorInvalidValue;
  return (int)cudaMemcpyFromSymbol(host_dst, pwc::g_corr_trace, sizeof(unsigned long long) * count);
}
