// fp32 3x3 convolution on CUDA cores (implicit GEMM, FFMA), NHWC + pitch, fused bias + LeakyReLU.
// Exact-fp32 mode of the network and the thin layers; replaces nn.Conv2d + nn.LeakyReLU of conv()
// (/root/reference/models/PWCNet.py:26-30) and predict_flow() (PWCNet.py:32-33).
//
// CTA tile: 8x16 output pixels (M=128) x BN output channels, K chunk = 16 input channels of one tap.
// 256 threads, thread tile 8 pixels x TN channels (BN = 16*TN).  A is staged k-major ([16][128+4]) so a
// thread reads its 8 pixels with two LDS.128; B ([16][BN]) is read with one LDS.(32*TN).  Register
// double-buffering overlaps the next chunk's global loads with the FMAs.
#include "common.cuh"
#include "../../include/pwc_b200.h"

namespace pwc {

constexpr int kTH = 8, kTWc = 16, kBM = 128, kKC = 16, kAP = kBM + 4;

struct ConvParams {
  const float* in; long in_pitch; int in_coff;
  const float* in2; long in2_pitch; int in2_coff; int Cin;
  const float* w; const float* bias;
  float* out; long out_pitch; int out_coff;
  float* out2; long out2_pitch; int out2_coff; int n_split;
  int Cout, CinP, CoutP, N, Hin, Win, Hout, Wout, stride, dil; float slope;
  const float* residual; long res_pitch;
};

template <int TN>
__global__ void __launch_bounds__(256, 2) conv3x3_f32_kernel(const ConvParams p) {
  constexpr int BN = 16 * TN;
  __shared__ __align__(16) float As[2][kKC][kAP];
  __shared__ __align__(16) float Bs[2][kKC][BN];

  const int tid = threadIdx.x;
  const int tilesx = cdiv(p.Wout, kTWc);
  const int tile = blockIdx.x;
  const int ty0 = (tile / tilesx) * kTH, tx0 = (tile % tilesx) * kTWc;
  const int n = blockIdx.z;
  const int co0 = blockIdx.y * BN;

  // A-load mapping: 128 px x 4 float4 (16 ch) = 512 float4, 2 per thread
  const int a_cg = tid & 3;            // which 4-channel group of the chunk
  const int a_px = tid >> 2;           // 0..63, second one +64
  // B-load mapping: 16 x BN floats = 4*BN float4
  // compute mapping
  const int tx = tid & 15, ty = tid >> 4;   // tx: channel group (TN ch), ty: pixel group (8 px = half a tile row)

  float acc[8][TN];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  const int nchunk = p.CinP / kKC;
  const int nk = 9 * nchunk;
  const bool second = (n >= p.n_split);
  const long ipitch = (second && p.in2) ? p.in2_pitch : p.in_pitch;
  const float* inb = (second && p.in2) ? p.in2 + (long)(n - p.n_split) * p.Hin * p.Win * ipitch + p.in2_coff
                                       : p.in + (long)n * p.Hin * p.Win * ipitch + p.in_coff;

  float4 ra[2];
  float4 rb[(4 * BN + 255) / 256];

  auto gload = [&](int kk) {
    const int tap = kk / nchunk, ch = kk - tap * nchunk;
    const int ky = tap / 3, kx = tap - ky * 3;
    const int c = ch * kKC + a_cg * 4;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int px = a_px + 64 * h;
      const int oy = ty0 + (px >> 4), ox = tx0 + (px & 15);
      const int iy = oy * p.stride + (ky - 1) * p.dil, ix = ox * p.stride + (kx - 1) * p.dil;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (iy >= 0 && iy < p.Hin && ix >= 0 && ix < p.Win && c < p.Cin)
        v = *reinterpret_cast<const float4*>(inb + ((long)iy * p.Win + ix) * ipitch + c);
      ra[h] = v;
    }
    const float* wb = p.w + ((long)tap * p.CinP + ch * kKC) * p.CoutP + co0;
#pragma unroll
    for (int q = 0; q < (4 * BN + 255) / 256; ++q) {
      const int e = tid + 256 * q;          // float4 index in [16][BN/4]
      if (e < 4 * BN) {
        const int kr = e / (BN / 4), cc = e - kr * (BN / 4);
        rb[q] = *reinterpret_cast<const float4*>(wb + (long)kr * p.CoutP + cc * 4);
      }
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int px = a_px + 64 * h;
      As[buf][a_cg * 4 + 0][px] = ra[h].x; As[buf][a_cg * 4 + 1][px] = ra[h].y;
      As[buf][a_cg * 4 + 2][px] = ra[h].z; As[buf][a_cg * 4 + 3][px] = ra[h].w;
    }
#pragma unroll
    for (int q = 0; q < (4 * BN + 255) / 256; ++q) {
      const int e = tid + 256 * q;
      if (e < 4 * BN) {
        const int kr = e / (BN / 4), cc = e - kr * (BN / 4);
        *reinterpret_cast<float4*>(&Bs[buf][kr][cc * 4]) = rb[q];
      }
    }
  };

  gload(0);
  sstore(0);
  __syncthreads();
  for (int kk = 0; kk < nk; ++kk) {
    const int buf = kk & 1;
    if (kk + 1 < nk) gload(kk + 1);
#pragma unroll
    for (int k = 0; k < kKC; ++k) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 8 + 4]);
      float b[TN];
      if (TN == 4) {
        const float4 bv = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
        b[0] = bv.x; b[1] = bv.y; b[2 % TN] = bv.z; b[3 % TN] = bv.w;
      } else if (TN == 2) {
        const float2 bv = *reinterpret_cast<const float2*>(&Bs[buf][k][tx * 2]);
        b[0] = bv.x; b[1 % TN] = bv.y;
      } else {
        b[0] = Bs[buf][k][tx];
      }
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kk + 1 < nk) sstore(buf ^ 1);
    __syncthreads();
  }

  // epilogue
  float* ob; long opitch; int nn = n;
  if (second && p.out2) { ob = p.out2 + p.out2_coff; opitch = p.out2_pitch; nn = n - p.n_split; }
  else { ob = p.out + p.out_coff; opitch = p.out_pitch; }
  const int oy = ty0 + (ty >> 1);
  if (oy >= p.Hout) return;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int ox = tx0 + (ty & 1) * 8 + i;
    if (ox >= p.Wout) continue;
    const long pix = ((long)nn * p.Hout + oy) * p.Wout + ox;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int co = co0 + tx * TN + j;
      if (co < p.Cout) {
        float v = lrelu(acc[i][j] + p.bias[co], p.slope);
        if (p.residual) v += p.residual[pix * p.res_pitch + co];
        ob[pix * opitch + co] = v;
      }
    }
  }
}

__global__ void pack_conv_weight_kernel(const float* __restrict__ wt, float* __restrict__ wp, const int* __restrict__ map,
                                        int Cout, int Cin, int CinP, int CoutP) {
  // wt: [Cout][Cin][3][3] -> wp: [9][CinP][CoutP] (wp pre-zeroed by the caller)
  const long total = (long)Cout * Cin * 9;
  for (long e = blockIdx.x * (long)blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
    const int tap = e % 9; const long r = e / 9;
    const int ci = r % Cin, co = r / Cin;
    const int row = map ? map[ci] : ci;
    wp[((long)tap * CinP + row) * CoutP + co] = wt[e];
  }
}

}  // namespace pwc

extern "C" int pwc_conv3x3_f32_fwd(const float* in, long in_pitch, int in_coff,
                                   const float* in2, long in2_pitch, int in2_coff, int Cin,
                                   const float* w, const float* bias,
                                   float* out, long out_pitch, int out_coff,
                                   float* out2, long out2_pitch, int out2_coff, int n_split,
                                   int Cout, int N, int Hin, int Win, int stride, int dil, float slope,
                                   const float* residual, long res_pitch, void* stream) {
  using namespace pwc;
  if ((Cin & 3) || (in_coff & 3) || (in_pitch & 3) || (in2 && ((in2_coff & 3) || (in2_pitch & 3))) || (stride != 1 && stride != 2) || N <= 0 || N > 65535)
    return (int)cudaErrorInvalidValue;
  ConvParams p;
  p.in = in; p.in_pitch = in_pitch; p.in_coff = in_coff; p.Cin = Cin;
  p.in2 = in2; p.in2_pitch = in2_pitch; p.in2_coff = in2_coff; p.w = w; p.bias = bias;
  p.out = out; p.out_pitch = out_pitch; p.out_coff = out_coff;
  p.out2 = out2; p.out2_pitch = out2_pitch; p.out2_coff = out2_coff; p.n_split = (out2 || in2) ? n_split : N;
  p.Cout = Cout; p.CinP = (Cin + 15) / 16 * 16; p.CoutP = (Cout + 63) / 64 * 64;
  p.N = N; p.Hin = Hin; p.Win = Win;
  p.Hout = (Hin - 1) / stride + 1; p.Wout = (Win - 1) / stride + 1;   // k=3, pad=dil, any dilation
  p.stride = stride; p.dil = dil; p.slope = slope; p.residual = residual; p.res_pitch = res_pitch;
  const int tiles = cdiv(p.Hout, kTH) * cdiv(p.Wout, kTWc);
  cudaStream_t s = (cudaStream_t)stream;
  if (Cout > 32) {
    conv3x3_f32_kernel<4><<<dim3(tiles, cdiv(Cout, 64), N), 256, 0, s>>>(p);
  } else if (Cout > 16) {
    conv3x3_f32_kernel<2><<<dim3(tiles, 1, N), 256, 0, s>>>(p);
  } else {
    conv3x3_f32_kernel<1><<<dim3(tiles, 1, N), 256, 0, s>>>(p);
  }
  return (int)cudaGetLastError();
}

extern "C" int pwc_pack_conv_weight_f32(const float* w_torch, float* w_packed, const int* chan_map,
                                        int Cout, int Cin, int CinP, int CoutP, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(w_packed, 0, sizeof(float) * 9 * (size_t)CinP * CoutP, s);
  if (e != cudaSuccess) return (int)e;
  pwc::pack_conv_weight_kernel<<<256, 256, 0, s>>>(w_torch, w_packed, chan_map, Cout, Cin, CinP, CoutP);
  return (int)cudaGetLastError();
}
