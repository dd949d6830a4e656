// Driver-side pre/post-processing and accuracy metrics on the device: the steps the
// reference drivers run in numpy / torch on either side of PWCDCNet.forward.  All HBM-bound elementwise / gather /
// reduction kernels: coalesced along x, grid = a multiple of the SM count, grid-stride loops.
//   pre : frame_to_tensor + pad_to_multiple_of_64 + torch.cat   pwc_extract_flow_video.py:27-42,197-205
//         (uint8 HWC -> float/255, optional BGR<->RGB, replicate pad)  script_pwc.py:61-71, inference_kitti.py:53-63,216-217
//   post: unpad + flow_resize (bilinear, align_corners=True, vector rescale) + x20
//         inference_kitti.py:66-91,219-223; script_pwc.py:72-81; pwc_extract_flow_video.py:44-47
//   metrics: EPE and Fl-all over valid pixels   inference_kitti.py:94-128; inference.py:105-159
#include "common.cuh"
#include "../../include/pwc_b200.h"

namespace pwc {

// one thread = 4 consecutive output x of one (n, image, y) row, all 3 colour planes: 12 source bytes -> 3 float4 stores
__global__ void __launch_bounds__(256)
preprocess_u8_kernel(const uint8_t* __restrict__ im1, const uint8_t* __restrict__ im2, long img_stride, long row_stride,
                     int B, int H, int W, int Hp, int Wp, int swap_rb, float* __restrict__ out, int pairs,
                     int norm, float3 mean, float3 std) {
  // pairs = 1: out [B,6,Hp,Wp] (frame 1 | frame 2 of pair n); pairs = 0: out [B,3,Hp,Wp], a plain frame stack from im1
  const int Wq = Wp >> 2;
  const long total = (long)B * (pairs ? 2 : 1) * Hp * Wq;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int xq = (int)(i % Wq); long t = i / Wq;
    const int y = (int)(t % Hp); t /= Hp;
    const int im = pairs ? (int)(t & 1) : 0; const int n = pairs ? (int)(t >> 1) : (int)t;
    const uint8_t* src = (im ? im2 : im1) + (long)n * img_stride + (long)min(y, H - 1) * row_stride;   // replicate rows
    float v[3][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int x = min(4 * xq + k, W - 1);                                                            // replicate columns
      const uint8_t* px = src + 3 * x;
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c][k] = __fdiv_rn((float)px[swap_rb ? 2 - c : c], 255.0f);        // arr.astype(float32) / 255.0
      if (norm) {      // T.Normalize after T.ToTensor (inference_kitti.py:175-179, inference.py:305-310): tensor.sub_(mean).div_(std)
        v[0][k] = __fdiv_rn(__fsub_rn(v[0][k], mean.x), std.x);
        v[1][k] = __fdiv_rn(__fsub_rn(v[1][k], mean.y), std.y);
        v[2][k] = __fdiv_rn(__fsub_rn(v[2][k], mean.z), std.z);
      }
    }
    float* dst = out + (((long)n * (pairs ? 6 : 3) + 3 * im) * Hp + y) * Wp + 4 * xq;
#pragma unroll
    for (int c = 0; c < 3; ++c)
      *reinterpret_cast<float4*>(dst + (long)c * Hp * Wp) = make_float4(v[c][0], v[c][1], v[c][2], v[c][3]);
  }
}

// ATen upsample_bilinear2d, align_corners=True (UpSample.h area_pixel_compute_scale / compute_source_index):
// scale = (in - 1) / (out - 1) (0 when out == 1), src = scale * dst, i0 = (int)src, l1 = src - i0, l0 = 1 - l1
__global__ void __launch_bounds__(256)
flow_post_kernel(const float* __restrict__ flow, int B, int h, int w, int hc, int wc, float* __restrict__ out,
                 int H, int W, float mul_u, float mul_v, int out_hwc) {
  const float sy = H > 1 ? (float)(hc - 1) / (float)(H - 1) : 0.f;
  const float sx = W > 1 ? (float)(wc - 1) / (float)(W - 1) : 0.f;
  const long total = (long)B * H * W;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int x = (int)(i % W); long t = i / W;
    const int y = (int)(t % H); const int n = (int)(t / H);
    const float fy = sy * (float)y, fx = sx * (float)x;
    const int y0 = (int)fy, x0 = (int)fx;
    const int y1 = y0 + (y0 < hc - 1 ? 1 : 0), x1 = x0 + (x0 < wc - 1 ? 1 : 0);
    const float ly1 = fy - (float)y0, ly0 = 1.f - ly1, lx1 = fx - (float)x0, lx0 = 1.f - lx1;
    float r[2];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float* p = flow + ((long)n * 2 + c) * h * w;
      const float a = p[(long)y0 * w + x0], b = p[(long)y0 * w + x1], cc = p[(long)y1 * w + x0], d = p[(long)y1 * w + x1];
      r[c] = ly0 * (lx0 * a + lx1 * b) + ly1 * (lx0 * cc + lx1 * d);
    }
    r[0] *= mul_u; r[1] *= mul_v;
    if (out_hwc) {
      *reinterpret_cast<float2*>(out + 2 * i) = make_float2(r[0], r[1]);
    } else {
      out[((long)n * 2 + 0) * H * W + (long)y * W + x] = r[0];
      out[((long)n * 2 + 1) * H * W + (long)y * W + x] = r[1];
    }
  }
}

// per-sample: acc[n] = { sum of EPE over valid pixels (fp64), #valid, #outliers } ; outlier: EPE > max(abs_thr, rel_thr*|gt|)
__global__ void __launch_bounds__(256)
flow_metrics_kernel(const float* __restrict__ pred, const float* __restrict__ gt, const uint8_t* __restrict__ valid,
                    int H, int W, int layout_hwc, float abs_thr, float rel_thr, double* __restrict__ acc) {
  const int n = blockIdx.y;
  const long hw = (long)H * W;
  double s = 0.0; unsigned int nv = 0, no = 0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < hw; i += (long)gridDim.x * blockDim.x) {
    if (valid && !valid[n * hw + i]) continue;
    float pu, pv, gu, gv;
    if (layout_hwc) {
      const float2 p = *reinterpret_cast<const float2*>(pred + 2 * (n * hw + i));
      const float2 g = *reinterpret_cast<const float2*>(gt + 2 * (n * hw + i));
      pu = p.x; pv = p.y; gu = g.x; gv = g.y;
    } else {
      pu = pred[(2 * n) * hw + i]; pv = pred[(2 * n + 1) * hw + i];
      gu = gt[(2 * n) * hw + i]; gv = gt[(2 * n + 1) * hw + i];
    }
    const float du = __fsub_rn(pu, gu), dv = __fsub_rn(pv, gv);
    const float epe = __fsqrt_rn(__fadd_rn(__fmul_rn(du, du), __fmul_rn(dv, dv)));       // numpy: sqrt(du*du + dv*dv), fp32
    const float mag = __fsqrt_rn(__fadd_rn(__fmul_rn(gu, gu), __fmul_rn(gv, gv)));
    s += (double)epe; ++nv;
    if (epe > fmaxf(abs_thr, __fmul_rn(rel_thr, mag))) ++no;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    nv += __shfl_xor_sync(0xffffffffu, nv, o);
    no += __shfl_xor_sync(0xffffffffu, no, o);
  }
  __shared__ double ss[8]; __shared__ unsigned int sv[8], so[8];
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  if (lane == 0) { ss[wp] = s; sv[wp] = nv; so[wp] = no; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < 8; ++k) { s += ss[k]; nv += sv[k]; no += so[k]; }
    atomicAdd(acc + 3 * n, s);
    atomicAdd(acc + 3 * n + 1, (double)nv);            // counts below 2^53: exact in fp64 whatever the order
    atomicAdd(acc + 3 * n + 2, (double)no);
  }
}

static int grid_for(long work_items, int threads) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const long blocks = (work_items + threads - 1) / threads;
  const long cap = (long)sms * 8;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace pwc

extern "C" int pwc_preprocess_u8_fwd(const uint8_t* im1, const uint8_t* im2, long img_stride, long row_stride,
                                     int B, int H, int W, int Hp, int Wp, int swap_rb, float* out, void* stream) {
  if (B <= 0 || H <= 0 || W <= 0 || Hp < H || Wp < W || (Wp & 3) || row_stride < 3L * W ||
      (reinterpret_cast<uintptr_t>(out) & 15))
    return (int)cudaErrorInvalidValue;
  const long work = (long)B * 2 * Hp * (Wp >> 2);
  pwc::preprocess_u8_kernel<<<pwc::grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(
      im1, im2, img_stride, row_stride, B, H, W, Hp, Wp, swap_rb, out, 1, 0, make_float3(0.f, 0.f, 0.f), make_float3(1.f, 1.f, 1.f));
  return (int)cudaGetLastError();
}

// The KITTI drivers' variant (inference_kitti.py:175-179, inference.py:305-310): ToTensor (/255) followed by
// Normalize(mean, std) per OUTPUT channel (after the optional R<->B swap), then replicate pad + cat as above.
extern "C" int pwc_preprocess_u8_norm_fwd(const uint8_t* im1, const uint8_t* im2, long img_stride, long row_stride,
                                          int B, int H, int W, int Hp, int Wp, int swap_rb, const float* mean3,
                                          const float* std3, float* out, void* stream) {
  if (B <= 0 || H <= 0 || W <= 0 || Hp < H || Wp < W || (Wp & 3) || row_stride < 3L * W || !mean3 || !std3 ||
      (reinterpret_cast<uintptr_t>(out) & 15))
    return (int)cudaErrorInvalidValue;
  const long work = (long)B * 2 * Hp * (Wp >> 2);
  pwc::preprocess_u8_kernel<<<pwc::grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(
      im1, im2, img_stride, row_stride, B, H, W, Hp, Wp, swap_rb, out, 1, 1, make_float3(mean3[0], mean3[1], mean3[2]),
      make_float3(std3[0], std3[1], std3[2]));
  return (int)cudaGetLastError();
}

extern "C" int pwc_preprocess_frames_u8_fwd(const uint8_t* frames, long img_stride, long row_stride,
                                            int N, int H, int W, int Hp, int Wp, int swap_rb, float* out, void* stream) {
  if (N <= 0 || H <= 0 || W <= 0 || Hp < H || Wp < W || (Wp & 3) || row_stride < 3L * W ||
      (reinterpret_cast<uintptr_t>(out) & 15))
    return (int)cudaErrorInvalidValue;
  const long work = (long)N * Hp * (Wp >> 2);
  pwc::preprocess_u8_kernel<<<pwc::grid_for(work, 256), 256, 0, (cudaStream_t)stream>>>(
      frames, frames, img_stride, row_stride, N, H, W, Hp, Wp, swap_rb, out, 0, 0, make_float3(0.f, 0.f, 0.f), make_float3(1.f, 1.f, 1.f));
  return (int)cudaGetLastError();
}

extern "C" int pwc_flow_postprocess_fwd(const float* flow, int B, int h, int w, int crop_h, int crop_w,
                                        float* out, int H, int W, float mult, int out_hwc, void* stream) {
  if (B <= 0 || h <= 0 || w <= 0 || crop_h <= 0 || crop_w <= 0 || crop_h > h || crop_w > w || H <= 0 || W <= 0)
    return (int)cudaErrorInvalidValue;
  // flow_rs[:,0] *= new_w / W ; flow_rs[:,1] *= new_h / H (inference_kitti.py:88-90), then the caller's x`mult`
  const float mul_u = (float)((double)W / (double)crop_w) * mult, mul_v = (float)((double)H / (double)crop_h) * mult;
  pwc::flow_post_kernel<<<pwc::grid_for((long)B * H * W, 256), 256, 0, (cudaStream_t)stream>>>(
      flow, B, h, w, crop_h, crop_w, out, H, W, mul_u, mul_v, out_hwc);
  return (int)cudaGetLastError();
}

extern "C" int pwc_flow_metrics_fwd(const float* pred, const float* gt, const uint8_t* valid, int B, int H, int W,
                                    int layout_hwc, float abs_thr, float rel_thr, double* acc, void* stream) {
  if (B <= 0 || B > 65535 || H <= 0 || W <= 0) return (int)cudaErrorInvalidValue;
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(acc, 0, sizeof(double) * 3 * (size_t)B, s);
  if (e != cudaSuccess) return (int)e;
  int gx = pwc::grid_for((long)H * W, 256 * 4);
  pwc::flow_metrics_kernel<<<dim3(gx, B), 256, 0, s>>>(pred, gt, valid, H, W, layout_hwc, abs_thr, rel_thr, acc);
  return (int)cudaGetLastError();
}
