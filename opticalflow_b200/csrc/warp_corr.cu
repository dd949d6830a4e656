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
// is the cost-volume entry of pixel (y = r-4+i, x) at displacement (dy = 4-i, dx = j-4).  One lane owns
// one (r,x): a 9x9 outer product per channel, 81 FMA for 18 shared-memory operands.  A warp is 32
// consecutive x of one row r, a CTA is 8 consecutive rows r (8 warps) x 32 columns.  The warped features
// w2 are produced straight into shared memory (never written to HBM); every warped row is computed exactly
// once per CTA column strip; the redundancy sits on f1 (rows r0-4..r0+11, a clean box).
// Channels are processed in chunks of 32 with the 81 accumulators held in registers across chunks.
// Shared-memory rows are 128 B (32 fp32 channels of one pixel); the 16-byte chunk index is XOR-swizzled
// with (pixel & 7) so that the LDS.128 of 8 consecutive lanes hit 8 distinct bank groups.
#include "common.cuh"
#include "../../include/pwc_b200.h"

namespace pwc {

constexpr int kMD = 4;
constexpr int kND = 9;                 // displacements per axis
constexpr int kRows = 8;               // warped rows (warps) per CTA
constexpr int kTW = 32;                // columns per CTA (lanes)
constexpr int kF1Rows = kRows + 2 * kMD;   // 16
constexpr int kW2Cols = kTW + 2 * kMD;     // 40
constexpr int kChunk = 32;             // channels per pass
constexpr int kThreads = kRows * 32;   // 256
constexpr int kF1Pix = kF1Rows * kTW;  // 512
constexpr int kW2Pix = kRows * kW2Cols;  // 320
constexpr int kSmemFloats = (kF1Pix + kW2Pix) * kChunk;  // 26624 floats = 106496 B

struct WarpCorrParams {
  const void* f1; long f1_pitch;
  const void* f2; long f2_pitch;
  const float* flow; long flow_pitch;
  float flow_scale; int align_corners; float mask_thresh; float slope;
  void* out; long out_pitch;
  int B, C, H, W;
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

template <typename T, bool NCHW>
__global__ void __launch_bounds__(kThreads, 2)
warp_corr_lrelu_kernel(const WarpCorrParams p) {
  extern __shared__ __align__(1024) float smem[];
  float4* f1s = reinterpret_cast<float4*>(smem);                    // [kF1Pix][8 chunks]
  float4* w2s = reinterpret_cast<float4*>(smem) + kF1Pix * 8;       // [kW2Pix][8 chunks]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int x0 = blockIdx.x * kTW;
  const int r0 = blockIdx.y * kRows - kMD;       // first warped row of this CTA (may be < 0)
  const int n = blockIdx.z;
  const int C = p.C, H = p.H, W = p.W;
  const T* f1 = reinterpret_cast<const T*>(p.f1);
  const T* f2 = reinterpret_cast<const T*>(p.f2);

  float acc[kND][kND];
#pragma unroll
  for (int i = 0; i < kND; ++i)
#pragma unroll
    for (int j = 0; j < kND; ++j) acc[i][j] = 0.f;

  const int r = r0 + warp;
  const bool row_live = (r >= 0 && r < H);   // warp-uniform: rows outside the image contribute zeros

  for (int c0 = 0; c0 < C; c0 += kChunk) {
    // ---------------- fill: f1 box (rows r0-4 .. r0+11, cols x0..x0+31) ----------------
#pragma unroll 4
    for (int it = tid; it < kF1Pix * 8; it += kThreads) {
      const int k = it & 7, px = (it >> 3) & (kTW - 1), row = it >> 8;
      const int y = r0 - kMD + row, x = x0 + px;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (y >= 0 && y < H && x < W) v = load_feat4<T, NCHW>(f1, p.f1_pitch, n, y, x, c0 + 4 * k, C, H, W);
      f1s[swz(row * kTW + px, k)] = v;
    }
    // ---------------- fill: warped image-2 rows r0..r0+7, cols x0-4..x0+35 ----------------
#pragma unroll 2
    for (int it = tid; it < kW2Pix * 8; it += kThreads) {
      const int k = it & 7, pix = it >> 3;
      const int q = pix % kW2Cols, wr = pix / kW2Cols;
      const int yy = r0 + wr, xx = x0 - kMD + q;
      const int c = c0 + 4 * k;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
        if (p.flow == nullptr) {
          v = load_feat4<T, NCHW>(f2, p.f2_pitch, n, yy, xx, c, C, H, W);
        } else {
          float u, vv;
          if (NCHW) {
            u = p.flow[(((long)n * 2 + 0) * H + yy) * W + xx];
            vv = p.flow[(((long)n * 2 + 1) * H + yy) * W + xx];
          } else {
            const float2 fl = *reinterpret_cast<const float2*>(p.flow + (((long)n * H + yy) * W + xx) * p.flow_pitch);
            u = fl.x; vv = fl.y;
          }
          // PWCNet.py:159-163 (flo = up_flow*scale; vgrid = grid + flo; 2*v/(W-1) - 1), one rounding per op
          const float vx = __fadd_rn((float)xx, __fmul_rn(u, p.flow_scale));
          const float vy = __fadd_rn((float)yy, __fmul_rn(vv, p.flow_scale));
          const float gx = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, vx), (float)max(W - 1, 1)), 1.0f);
          const float gy = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, vy), (float)max(H - 1, 1)), 1.0f);
          // ATen grid_sampler_unnormalize (GridSampler.h:27-35)
          float ix, iy;
          if (p.align_corners) {
            ix = __fmul_rn(__fdiv_rn(__fadd_rn(gx, 1.f), 2.f), (float)(W - 1));
            iy = __fmul_rn(__fdiv_rn(__fadd_rn(gy, 1.f), 2.f), (float)(H - 1));
          } else {
            ix = __fdiv_rn(__fsub_rn(__fmul_rn(__fadd_rn(gx, 1.f), (float)W), 1.f), 2.f);
            iy = __fdiv_rn(__fsub_rn(__fmul_rn(__fadd_rn(gy, 1.f), (float)H), 1.f), 2.f);
          }
          const float fx0 = floorf(ix), fy0 = floorf(iy);
          const float fx1 = fx0 + 1.f, fy1 = fy0 + 1.f;
          const float wnw = __fmul_rn(fx1 - ix, fy1 - iy), wne = __fmul_rn(ix - fx0, fy1 - iy);
          const float wsw = __fmul_rn(fx1 - ix, iy - fy0), wse = __fmul_rn(ix - fx0, iy - fy0);
          // keep the int conversion safe for wild flows
          const float cx = fminf(fmaxf(fx0, -2.f), (float)W + 1.f), cy = fminf(fmaxf(fy0, -2.f), (float)H + 1.f);
          const int ix0 = (int)cx, iy0 = (int)cy;
          const bool okx0 = (fx0 >= 0.f && fx0 <= (float)(W - 1)), okx1 = (fx1 >= 0.f && fx1 <= (float)(W - 1));
          const bool oky0 = (fy0 >= 0.f && fy0 <= (float)(H - 1)), oky1 = (fy1 >= 0.f && fy1 <= (float)(H - 1));
          float m = 0.f;                       // grid_sample(ones): nw, ne, sw, se order
          if (okx0 && oky0) m = __fadd_rn(m, wnw);
          if (okx1 && oky0) m = __fadd_rn(m, wne);
          if (okx0 && oky1) m = __fadd_rn(m, wsw);
          if (okx1 && oky1) m = __fadd_rn(m, wse);
          if (!(m < p.mask_thresh) && m > 0.f) {   // PWCNet.py:174-177
            if (okx0 && oky0) fma4(v, wnw, load_feat4<T, NCHW>(f2, p.f2_pitch, n, iy0, ix0, c, C, H, W));
            if (okx1 && oky0) fma4(v, wne, load_feat4<T, NCHW>(f2, p.f2_pitch, n, iy0, ix0 + 1, c, C, H, W));
            if (okx0 && oky1) fma4(v, wsw, load_feat4<T, NCHW>(f2, p.f2_pitch, n, iy0 + 1, ix0, c, C, H, W));
            if (okx1 && oky1) fma4(v, wse, load_feat4<T, NCHW>(f2, p.f2_pitch, n, iy0 + 1, ix0 + 1, c, C, H, W));
          }
        }
      }
      w2s[swz(pix, k)] = v;
    }
    __syncthreads();

    // ---------------- compute: 9x9 outer product per 4 channels ----------------
    if (row_live) {
      const int nk = min(kChunk, C - c0 + 3) >> 2;   // 16B chunks holding live channels
      for (int k = 0; k < nk; ++k) {
        float4 wv[kND];
#pragma unroll
        for (int j = 0; j < kND; ++j) wv[j] = w2s[swz(warp * kW2Cols + lane + j, k)];
#pragma unroll
        for (int i = 0; i < kND; ++i) {
          const float4 a = f1s[swz((warp + i) * kTW + lane, k)];
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
    __syncthreads();
  }

  // ---------------- epilogue: /C (correlation_cuda_kernel.cu:143), LeakyReLU, store ----------------
  const float fC = (float)C;
  T* out = reinterpret_cast<T*>(p.out);
  const int x = x0 + lane;
  if (NCHW) {
#pragma unroll
    for (int i = 0; i < kND; ++i) {
      const int y = r - kMD + i;
      if (y < 0 || y >= H || x >= W) continue;
#pragma unroll
      for (int j = 0; j < kND; ++j) {
        const int tc = (kND - 1 - i) * kND + j;
        st1(out + (((long)n * (kND * kND) + tc) * H + y) * W + x, lrelu(__fdiv_rn(acc[i][j], fC), p.slope));
      }
    }
  } else {
    // stage [i][px][9] per warp in shared memory (f1s/w2s are dead), then write 36-byte runs cooperatively
    float* stg = smem + warp * (kND * kTW * kND);
#pragma unroll
    for (int i = 0; i < kND; ++i)
#pragma unroll
      for (int j = 0; j < kND; ++j) stg[(i * kTW + lane) * kND + j] = lrelu(__fdiv_rn(acc[i][j], fC), p.slope);
    __syncwarp();
    const int ncol = min(kTW, W - x0);
#pragma unroll
    for (int i = 0; i < kND; ++i) {
      const int y = r - kMD + i;
      if (y < 0 || y >= H) continue;   // warp-uniform
      T* orow = out + (((long)n * H + y) * W + x0) * p.out_pitch + (kND - 1 - i) * kND;
      for (int e = lane; e < ncol * kND; e += 32) {
        const int px = e / kND, j = e - px * kND;
        st1(orow + (long)px * p.out_pitch + j, stg[i * kTW * kND + e]);
      }
    }
  }
}

template <typename T, bool NCHW>
static int launch(const WarpCorrParams& p, cudaStream_t s) {
  auto kern = warp_corr_lrelu_kernel<T, NCHW>;
  const size_t smem = (size_t)kSmemFloats * sizeof(float);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  dim3 grid(cdiv(p.W, kTW), cdiv(p.H + 2 * kMD, kRows), p.B);
  kern<<<grid, kThreads, smem, s>>>(p);
  return (int)cudaGetLastError();
}

}  // namespace pwc

extern "C" int pwc_warp_corr_lrelu_fwd(const void* f1, long f1_pitch, const void* f2, long f2_pitch,
                                       const float* flow, long flow_pitch, float flow_scale,
                                       int align_corners, float mask_thresh, float slope,
                                       void* out, long out_pitch,
                                       int B, int C, int H, int W, int md,
                                       int layout, int dtype, void* stream) {
  using namespace pwc;
  if (md != kMD || B <= 0 || C <= 0 || H <= 0 || W <= 0 || B > 65535) return (int)cudaErrorInvalidValue;
  if (layout == PWC_LAYOUT_NHWC) {
    const int a = (dtype == PWC_BF16) ? 4 : 4;   // 4-channel vector loads
    if (C % a || f1_pitch % a || f2_pitch % a) return (int)cudaErrorInvalidValue;
    if (flow && (flow_pitch % 2)) return (int)cudaErrorInvalidValue;
  } else if (dtype != PWC_F32) {
    return (int)cudaErrorInvalidValue;
  }
  WarpCorrParams p{f1, f1_pitch, f2, f2_pitch, flow, flow_pitch, flow_scale, align_corners, mask_thresh, slope,
                   out, out_pitch, B, C, H, W};
  cudaStream_t s = (cudaStream_t)stream;
  if (layout == PWC_LAYOUT_NCHW) return launch<float, true>(p, s);
  if (dtype == PWC_BF16) return launch<__nv_bfloat16, false>(p, s);
  return launch<float, false>(p, s);
}

// API-compatible op: correlation_cuda.forward for pad=md=4, k=1, strides 1 (correlation_cuda.cc:10-87).
extern "C" int pwc_corr_fwd(const void* f1, const void* f2, void* out, int B, int C, int H, int W, int md, void* stream) {
  return pwc_warp_corr_lrelu_fwd(f1, 0, f2, 0, nullptr, 0, 1.0f, 0, 0.9999f, 1.0f, out, 0, B, C, H, W, md,
                                 PWC_LAYOUT_NCHW, PWC_F32, stream);
}
