// Thin fp32 kernels around the convolutions: 2-channel transposed conv (deconvL / upfeatL), input packing,
// flow unpacking.  Reference: deconv() /root/reference/models/PWCNet.py:35-36 and its call sites
// PWCNet.py:208-209,222-223,236-237,252-253; driver tensor formats inference.py:217-223.
#include "common.cuh"
#include "../../include/pwc_b200.h"

namespace pwc {

// One warp per output pixel; lanes stride over input channels (float4), 2x2 live taps, shuffle-reduce 2 sums.
__global__ void __launch_bounds__(256) deconv4x4s2_c2_kernel(const float* __restrict__ in, long in_pitch, int in_coff, int Cin,
                                                             const float* __restrict__ w, const float* __restrict__ bias,
                                                             float* __restrict__ out, long out_pitch, int out_coff,
                                                             float* __restrict__ out_b, long out_b_pitch,
                                                             int N, int Hin, int Win, int CinP) {
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
    const int iy2 = oy + 1 - ky;             // even by construction
    const int iy = iy2 >> 1;
    if (iy2 < 0 || iy >= Hin) continue;
#pragma unroll
    for (int b = 0; b < 2; ++b) {
      const int kx = ((ox + 1) & 1) + 2 * b;
      const int ix2 = ox + 1 - kx;
      const int ix = ix2 >> 1;
      if (ix2 < 0 || ix >= Win) continue;
      const float* ip = in + (((long)n * Hin + iy) * Win + ix) * in_pitch + in_coff;
      const float* w0 = w + ((long)(ky * 4 + kx) * 2 + 0) * CinP;
      const float* w1 = w0 + CinP;
      for (int c = lane * 4; c < Cin; c += 128) {
        const float4 v = *reinterpret_cast<const float4*>(ip + c);
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
    float* o = out + wid * out_pitch + out_coff;
    o[0] = s0; o[1] = s1;
    if (out_b) { float* ob = out_b + wid * out_b_pitch; ob[0] = s0; ob[1] = s1; }
  }
}

__global__ void pack_deconv_weight_kernel(const float* __restrict__ wt, float* __restrict__ wp, const int* __restrict__ map,
                                          int Cin, int CinP) {
  // wt: [Cin][2][4][4] -> wp: [16][2][CinP] (pre-zeroed)
  const int total = Cin * 32;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    const int tap = e & 15, co = (e >> 4) & 1, ci = e >> 5;
    const int row = map ? map[ci] : ci;
    wp[((long)tap * 2 + co) * CinP + row] = wt[e];
  }
}

__global__ void pack_input_kernel(const float* __restrict__ x, float4* __restrict__ out, int B, int H, int W) {
  const long hw = (long)H * W;
  const long total = 2L * B * hw;
  for (long e = blockIdx.x * (long)blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
    const long pix = e % hw; const int img = e / hw;          // img in [0,2B): im1 of all pairs, then im2
    const int b = img % B, half = img / B;
    const float* src = x + ((long)b * 6 + half * 3) * hw + pix;
    out[e] = make_float4(src[0], src[hw], src[2 * hw], 0.f);
  }
}

__global__ void unpack_flow_kernel(const float* __restrict__ in, long pitch, int coff, float* __restrict__ out,
                                   int B, int H, int W, float scale) {
  pdl_wait();
  pdl_launch_dependents();
  const long hw = (long)H * W;
  const long total = (long)B * hw;
  for (long e = blockIdx.x * (long)blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
    const long pix = e % hw; const int b = e / hw;
    const float* s = in + e * pitch + coff;
    out[((long)b * 2 + 0) * hw + pix] = s[0] * scale;
    out[((long)b * 2 + 1) * hw + pix] = s[1] * scale;
  }
}

}  // namespace pwc

extern "C" int pwc_deconv4x4s2_c2_f32_fwd(const float* in, long in_pitch, int in_coff, int Cin,
                                          const float* w, const float* bias,
                                          float* out, long out_pitch, int out_coff,
                                          float* out_b, long out_b_pitch,
                                          int N, int Hin, int Win, void* stream) {
  if ((in_pitch & 3) || (in_coff & 3) || Cin <= 0) return (int)cudaErrorInvalidValue;
  const int CinP = (Cin + 3) / 4 * 4;
  const long warps = (long)N * 4 * Hin * Win;
  const long blocks = (warps * 32 + 255) / 256;
  pwc::deconv4x4s2_c2_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      in, in_pitch, in_coff, CinP, w, bias, out, out_pitch, out_coff, out_b, out_b_pitch, N, Hin, Win, CinP);
  return (int)cudaGetLastError();
}

extern "C" int pwc_pack_deconv_weight_f32(const float* w_torch, float* w_packed, const int* chan_map,
                                          int Cin, int CinP, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(w_packed, 0, sizeof(float) * 32 * (size_t)CinP, s);
  if (e != cudaSuccess) return (int)e;
  pwc::pack_deconv_weight_kernel<<<64, 256, 0, s>>>(w_torch, w_packed, chan_map, Cin, CinP);
  return (int)cudaGetLastError();
}

extern "C" int pwc_pack_input_f32(const float* x, float* out, int B, int H, int W, void* stream) {
  pwc::pack_input_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(x, reinterpret_cast<float4*>(out), B, H, W);
  return (int)cudaGetLastError();
}

extern "C" int pwc_unpack_flow_f32(const float* in, long in_pitch, int in_coff, float* out,
                                   int B, int H, int W, float scale, void* stream) {
  return (int)pwc::launch_pdl(pwc::unpack_flow_kernel, dim3(148 * 4), dim3(256), 0, (cudaStream_t)stream, in, in_pitch, in_coff, out,
                              B, H, W, scale);
}

extern "C" int pwc_b200_arch(void) { return 100; }
extern "C" const char* pwc_b200_version(void) { return "pwc_b200 0.1 (sm_100a)"; }
