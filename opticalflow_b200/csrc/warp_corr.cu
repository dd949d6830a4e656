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
// a 9x9 outer product per channel, 81 FMA for 18 shared-memory operands.  A warp is 32 consecutive x of one row r,
// a CTA is ROWS consecutive rows r (ROWS warps) x 32 columns.  The warped features w2 are produced straight into
// shared memory (never written to HBM), each warped row exactly once per column strip; the halo redundancy sits on
// f1 (rows r0-4..r0+ROWS+3), which is a clean box and is fetched by ONE TMA instruction per channel chunk
// (SWIZZLE_128B, zero fill outside the image).  Channels are processed in chunks of 32 with the 81 accumulators in
// registers across chunks.  Shared-memory rows are 128 B (32 fp32 channels of one pixel); the 16-byte chunk index
// is XOR-swizzled with (pixel & 7) -- the TMA 128B swizzle -- so the LDS.128 of 8 consecutive lanes hit 8 distinct
// bank groups.  Bilinear coordinates / weights / mask are computed once per CTA into a small table.
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "tmap.cuh"
#include "../../include/pwc_b200.h"

namespace pwc {

constexpr int kMD = 4;
constexpr int kND = 9;                 // displacements per axis
constexpr int kTW = 32;                // columns per CTA (lanes)
constexpr int kW2Cols = kTW + 2 * kMD; // 40
constexpr int kChunk = 32;             // channels per pass

struct WarpCorrParams {
  const void* f1; long f1_pitch;
  const void* f2; long f2_pitch;
  const float* flow; long flow_pitch;
  float flow_scale; int align_corners; float mask_thresh; float slope;
  void* out; long out_pitch;
  int B, C, H, W;
};

template <int ROWS> struct Geo {
  static constexpr int kF1Rows = ROWS + 2 * kMD;
  static constexpr int kF1Pix = kF1Rows * kTW;
  static constexpr int kW2Pix = ROWS * kW2Cols;
  static constexpr int kThreads = ROWS * 32;
  static constexpr int kFeatBytes = (kF1Pix + kW2Pix) * kChunk * 4;
  static constexpr int kTabBytes = kW2Pix * 24;
  // staging for the NHWC epilogue: sum over output rows of 32 * (9*n_i | 1) floats  <=  32*(81*ROWS + ROWS+8)
  static constexpr int kStageFloats = 32 * (81 * ROWS + ROWS + 8);
  static constexpr int kMainBytes = (kFeatBytes > kStageFloats * 4 ? kFeatBytes : kStageFloats * 4);
  static constexpr int kSmemBytes = kMainBytes + kTabBytes + 64 + 1024;   // + barrier/rowbase + alignment slack
};

// swizzled float4 slot of (pixel index p within its array, 16B-chunk k)
__device__ __forceinline__ int swz(int p, int k) { return p * 8 + (k ^ (p & 7)); }

template <typename T, bool NCHW>
__device__ __forceinline__ float4 load_feat4(const T* base, long pitch, int n, int y, int x, int c, int C, int H, int W) {
  if (NCHW) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    const T* p = base + (((long)n * C + c) * H + y) * W + x;
    const long cs = (long)H * W;
    if (c + 0 < C) v.x = ld1(p);
    if (c + 1 < C) v.y = ld1(p + cs);
    if (c + 2 < C) v.z = ld1(p + 2 * cs);
    if (c + 3 < C) v.w = ld1(p + 3 * cs);
    return v;
  } else {
    if (c >= C) return make_float4(0.f, 0.f, 0.f, 0.f);
    return ld4(base + (((long)n * H + y) * W + x) * pitch + c);
  }
}

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
