// bf16-mode companions of the tensor-core convolutions (all arithmetic fp32, storage bf16):
//   * conv1a: the first pyramid layer (3 -> 16, stride 2, + LeakyReLU; /root/reference/models/PWCNet.py:52,184,190)
//     fused with the input split im1 = x[:,:3], im2 = x[:,3:] (PWCNet.py:181-182): reads the driver tensor
//     [B,6,H,W] fp32 NCHW directly and writes bf16 NHWC [2B,H/2,W/2,16].  Cin = 3 is HBM-bound work for CUDA cores.
//   * the 2-channel transposed convolution deconv() (PWCNet.py:35-36; call sites 208-209,222-223,236-237,252-253)
//     with fp32 or bf16 input and output.
#include "common.cuh"
#include "../../include/pwc_b200.h"

namespace pwc {

// conv1a (3 -> 16, stride 2, PWCNet.py:52) straight from the fp32 NCHW driver tensor, fused with the im1/im2 split.
// One thread = 4 adjacent output pixels x 16 channels: every broadcast weight read (LDS.128) feeds 16 FMAs, the 9 input
// columns of a row are loaded once for the 4 pixels.
__global__ void __launch_bounds__(256, 2) conv1a_bf16_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                          const float* __restrict__ bias, __nv_bfloat16* __restrict__ out,
                                                          int B, int H, int W, float slope, int frames) {
  // frames = 0: x is the driver tensor [B,6,H,W] (image b of half h at planes 6b + 3h), 2B output images (im1 half first);
  // frames = 1: x is a frame stack [B,3,H,W], B output images (shared-pyramid video plan)
  __shared__ __align__(16) float ws[27][16];
  __shared__ float bs[16];
  pdl_wait();
  pdl_launch_dependents();
  for (int e = threadIdx.x; e < 432; e += blockDim.x) {       // torch [16][3][3][3] -> [ci*9+tap][co]
    const int co = e / 27, r = e - co * 27;
    ws[r][co] = w[e];
  }
  if (threadIdx.x < 16) bs[threadIdx.x] = bias[threadIdx.x];
  __syncthreads();
  const int Ho = H >> 1, Wo = W >> 1, Wq = (Wo + 3) >> 2;
  const long hw = (long)H * W;
  const long total = (frames ? 1L : 2L) * B * Ho * Wq;
  for (long e = blockIdx.x * (long)blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
    const int qx = e % Wq; const long t = e / Wq;
    const int oy = t % Ho; const int img = t / Ho;
    const int b = img % B, half = img / B;
    const float* src = frames ? x + (long)img * 3 * hw : x + ((long)b * 6 + half * 3) * hw;
    const int ox0 = 4 * qx, ix0 = 2 * ox0 - 1;
    float acc[4][16];
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
      for (int co = 0; co < 16; ++co) acc[p][co] = bs[co];
#pragma unroll
    for (int ci = 0; ci < 3; ++ci)
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
        const int iy = 2 * oy - 1 + ky;
        const bool yok = (iy >= 0 && iy < H);
        const float* row = src + ci * hw + (long)iy * W;
        float v[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) {
          const int ix = ix0 + i;
          v[i] = (yok && ix >= 0 && ix < W) ? __ldg(row + ix) : 0.f;
        }
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const float4* wr = reinterpret_cast<const float4*>(ws[ci * 9 + ky * 3 + kx]);
#pragma unroll
          for (int c4 = 0; c4 < 4; ++c4) {
            const float4 wv = wr[c4];
#pragma unroll
            for (int p = 0; p < 4; ++p) {
              const float a = v[2 * p + kx];
              acc[p][4 * c4 + 0] = fmaf(a, wv.x, acc[p][4 * c4 + 0]);
              acc[p][4 * c4 + 1] = fmaf(a, wv.y, acc[p][4 * c4 + 1]);
              acc[p][4 * c4 + 2] = fmaf(a, wv.z, acc[p][4 * c4 + 2]);
              acc[p][4 * c4 + 3] = fmaf(a, wv.w, acc[p][4 * c4 + 3]);
            }
          }
        }
      }
    __nv_bfloat16* orow = out + (((long)img * Ho + oy) * Wo + ox0) * 16;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
      if (ox0 + p >= Wo) break;
      uint32_t pk[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        __nv_bfloat162 h = __floats2bfloat162_rn(lrelu(acc[p][2 * i], slope), lrelu(acc[p][2 * i + 1], slope));
        pk[i] = *reinterpret_cast<uint32_t*>(&h);
      }
      uint4* o = reinterpret_cast<uint4*>(orow + p * 16);
      o[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      o[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    }
  }
}

// One warp per output pixel; lanes stride over input channels (4 at a time), 2x2 live taps, shuffle-reduce 2 sums.
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) deconv4x4s2_c2_any_kernel(const TI* __restrict__ in, long in_pitch, int in_coff, int Cin,
                                                                 const float* __restrict__ w, const float* __restrict__ bias,
                                                                 TO* __restrict__ out, long out_pitch, int out_coff,
                                                                 float* __restrict__ out_b, long out_b_pitch,
                                                                 int N, int Hin, int Win, int CinP) {
  pdl_wait();
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31;
  const long wid = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int Hout = 2 * Hin, Wout = 2 * Win;
  const long total = (long)N * Hout * Wout;
  if (wid >= total) return;
  const int ox = wid % Wout; const long t = wid / Wout;
  const int oy = t % Hout; const int n = t / Hout;
  float s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int a = 0; a < 2; ++a) {
    const int ky = ((oy + 1) & 1) + 2 * a;
    const int iy2 = oy + 1 - ky;
    const int iy = iy2 >> 1;
    if (iy2 < 0 || iy >= Hin) continue;
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int kx = ((ox + 1) & 1) + 2 * b;
      const int ix2 = ox + 1 - kx;
      const int ix = ix2 >> 1;
      if (ix2 < 0 || ix >= Win) continue;
      const TI* ip = in + (((long)n * Hin + iy) * Win + ix) * in_pitch + in_coff;
      const float* w0 = w + ((long)(ky * 4 + kx) * 2 + 0) * CinP;
      const float* w1 = w0 + CinP;
      for (int c = lane * 4; c < Cin; c += 128) {
        const float4 v = ld4(ip + c);
        const float4 a0 = *reinterpret_cast<const float4*>(w0 + c);
        const float4 a1 = *reinterpret_cast<const float4*>(w1 + c);
        s0 = fmaf(v.x, a0.x, s0); s0 = fmaf(v.y, a0.y, s0); s0 = fmaf(v.z, a0.z, s0); s0 = fmaf(v.w, a0.w, s0);
        s1 = fmaf(v.x, a1.x, s1); s1 = fmaf(v.y, a1.y, s1); s1 = fmaf(v.z, a1.z, s1); s1 = fmaf(v.w, a1.w, s1);
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s0 += __shfl_xor_sync(0xffffffffu, s0, o);
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
  }
  if (lane == 0) {
    s0 += bias[0]; s1 += bias[1];
    TO* o = out + wid * out_pitch + out_coff;
    st1(o, s0); st1(o + 1, s1);
    if (out_b) { float* ob = out_b + wid * out_b_pitch; ob[0] = s0; ob[1] = s1; }
  }
}

// Cin <= 4 (deconvL: flow -> up_flow, PWCNet.py:208,222,236,252): one THREAD per output pixel.
template <typename TI, typename TO>
__global__ void __launch_bounds__(256) deconv4x4s2_c2_small_kernel(const TI* __restrict__ in, long in_pitch, int in_coff,
                                                                   const float* __restrict__ w, const float* __restrict__ bias,
                                                                   TO* __restrict__ out, long out_pitch, int out_coff,
                                                                   float* __restrict__ out_b, long out_b_pitch,
                                                                   int N, int Hin, int Win, int CinP) {
  pdl_wait();
  pdl_launch_dependents();
  const int Hout = 2 * Hin, Wout = 2 * Win;
  const long total = (long)N * Hout * Wout;
  for (long e = blockIdx.x * (long)blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
    const int ox = e % Wout; const long t = e / Wout;
    const int oy = t % Hout; const int n = t / Hout;
    float s0 = bias[0], s1 = bias[1];
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      const int ky = ((oy + 1) & 1) + 2 * a;
      const int iy2 = oy + 1 - ky;
      const int iy = iy2 >> 1;
      if (iy2 < 0 || iy >= Hin) continue;
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int kx = ((ox + 1) & 1) + 2 * b;
        const int ix2 = ox + 1 - kx;
        const int ix = ix2 >> 1;
        if (ix2 < 0 || ix >= Win) continue;
        const float4 v = ld4(in + (((long)n * Hin + iy) * Win + ix) * in_pitch + in_coff);
        const float* w0 = w + ((long)(ky * 4 + kx) * 2 + 0) * CinP;
        const float4 a0 = *reinterpret_cast<const float4*>(w0);
        const float4 a1 = *reinterpret_cast<const float4*>(w0 + CinP);
        s0 = fmaf(v.x, a0.x, s0); s0 = fmaf(v.y, a0.y, s0); s0 = fmaf(v.z, a0.z, s0); s0 = fmaf(v.w, a0.w, s0);
        s1 = fmaf(v.x, a1.x, s1); s1 = fmaf(v.y, a1.y, s1); s1 = fmaf(v.z, a1.z, s1); s1 = fmaf(v.w, a1.w, s1);
      }
    }
    TO* o = out + e * out_pitch + out_coff;
    st1(o, s0); st1(o + 1, s1);
    if (out_b) { float* ob = out_b + e * out_b_pitch; ob[0] = s0; ob[1] = s1; }
  }
}

}  // namespace pwc

extern "C" int pwc_conv1a_bf16_fwd(const float* x, const float* w, const float* bias, void* out,
                                   int B, int H, int W, float slope, void* stream) {
  if (B <= 0 || (H & 1) || (W & 1)) return (int)cudaErrorInvalidValue;
  return (int)pwc::launch_pdl(pwc::conv1a_bf16_kernel, dim3(148 * 8), dim3(256), 0, (cudaStream_t)stream, x, w, bias,
                              reinterpret_cast<__nv_bfloat16*>(out), B, H, W, slope, 0);
}

// conv1a on a stack of N frames [N,3,H,W] fp32 -> [N,H/2,W/2,16] bf16 (video plan: every frame goes through the pyramid once)
extern "C" int pwc_conv1a_frames_bf16_fwd(const float* frames, const float* w, const float* bias, void* out,
                                          int N, int H, int W, float slope, void* stream) {
  if (N <= 0 || (H & 1) || (W & 1)) return (int)cudaErrorInvalidValue;
  return (int)pwc::launch_pdl(pwc::conv1a_bf16_kernel, dim3(148 * 8), dim3(256), 0, (cudaStream_t)stream, frames, w, bias,
                              reinterpret_cast<__nv_bfloat16*>(out), N, H, W, slope, 1);
}

extern "C" int pwc_deconv4x4s2_c2_fwd(const void* in, int in_dtype, long in_pitch, int in_coff, int Cin,
                                      const float* w, const float* bias,
                                      void* out, int out_dtype, long out_pitch, int out_coff,
                                      float* out_b, long out_b_pitch,
                                      int N, int Hin, int Win, void* stream) {
  using namespace pwc;
  if ((in_pitch & 3) || (in_coff & 3) || Cin <= 0) return (int)cudaErrorInvalidValue;
  const int CinP = (Cin + 3) / 4 * 4;
  const long warps = (long)N * 4 * Hin * Win;
  const unsigned blocks = (unsigned)((warps * 32 + 255) / 256);
  cudaStream_t s = (cudaStream_t)stream;
  if (CinP == 4) {
    const long px = (long)N * 4 * Hin * Win;
    const unsigned nb = (unsigned)((px + 255) / 256 < 148 * 16 ? (px + 255) / 256 : 148 * 16);
#define PWC_DECONV_S(TI, TO)                                                                                              \
  launch_pdl(deconv4x4s2_c2_small_kernel<TI, TO>, dim3(nb), dim3(256), 0, s, reinterpret_cast<const TI*>(in), in_pitch,    \
             in_coff, w, bias, reinterpret_cast<TO*>(out), out_pitch, out_coff, out_b, out_b_pitch, N, Hin, Win, CinP)
    if (in_dtype == PWC_F32 && out_dtype == PWC_F32) PWC_DECONV_S(float, float);
    else if (in_dtype == PWC_F32 && out_dtype == PWC_BF16) PWC_DECONV_S(float, __nv_bfloat16);
    else if (in_dtype == PWC_BF16 && out_dtype == PWC_BF16) PWC_DECONV_S(__nv_bfloat16, __nv_bfloat16);
    else if (in_dtype == PWC_BF16 && out_dtype == PWC_F32) PWC_DECONV_S(__nv_bfloat16, float);
    else return (int)cudaErrorInvalidValue;
#undef PWC_DECONV_S
    return (int)cudaGetLastError();
  }
#define PWC_DECONV(TI, TO)                                                                                             \
  launch_pdl(deconv4x4s2_c2_any_kernel<TI, TO>, dim3(blocks), dim3(256), 0, s, reinterpret_cast<const TI*>(in), in_pitch, in_coff, CinP, w, \
                                                           bias, reinterpret_cast<TO*>(out), out_pitch, out_coff, out_b, \
                                                           out_b_pitch, N, Hin, Win, CinP)
  if (in_dtype == PWC_F32 && out_dtype == PWC_F32) PWC_DECONV(float, float);
  else if (in_dtype == PWC_F32 && out_dtype == PWC_BF16) PWC_DECONV(float, __nv_bfloat16);
  else if (in_dtype == PWC_BF16 && out_dtype == PWC_BF16) PWC_DECONV(__nv_bfloat16, __nv_bfloat16);
  else if (in_dtype == PWC_BF16 && out_dtype == PWC_F32) PWC_DECONV(__nv_bfloat16, float);
  else return (int)cudaErrorInvalidValue;
#undef PWC_DECONV
  return (int)cudaGetLastError();
}
