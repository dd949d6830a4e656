// bf16-mode companions of the tensor-core convolutions (all arithmetic fp32, storage bf16):
//   * conv1a: the first pyramid layer (3 -> 16, stride 2, + LeakyReLU; /root/reference/models/PWCNet.py:52,184,190)
//     fused with the input split im1 = x[:,:3], im2 = x[:,3:] (PWCNet.py:181-182): reads the driver tensor
//     [B,6,H,W] fp32 NCHW directly and writes bf16 NHWC [2B,H/2,W/2,16].  Cin = 3 is HBM-bound work for CUDA cores.
//   * the 2-channel transposed convolution deconv() (PWCNet.py:35-36; call sites 208-209,222-223,236-237,252-253)
//     with fp32 or bf16 input and output.
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "tmap.cuh"
#include "../../include/pwc_b200.h"

namespace pwc {

// conv1a (3 -> 16, stride 2, PWCNet.py:52) straight from the fp32 NCHW driver tensor, fused with the im1/im2 split.
// One thread = 4 adjacent output pixels x 16 channels: every broadcast weight read (LDS.128) feeds 16 FMAs, the 9 input
// columns of a row are loaded once for the 4 pixels.
__global__ void __launch_bounds__(256, 2) conv1a_bf16_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                          const float* __restrict__ bias, __nv_bfloat16* __restrict__ out,
                                                          int B, int H, int W, float slope, int frames, int f16) {
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
      for (int i = 0; i < 8; ++i) pk[i] = pack2_h16(lrelu(acc[p][2 * i], slope), lrelu(acc[p][2 * i + 1], slope), f16 & 1);
      uint4* o = reinterpret_cast<uint4*>(orow + p * 16);
      o[0] = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      o[1] = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    }
  }
}

// conv1a on the tensor cores without giving up fp32 inputs: the layer is FMA-issue bound on the CUDA cores (432 FMA per
// output pixel: 78 us at batch 8, about 50% of the fp32 pipe), so the 27-tap dot products run as mma.sync.m16n8k16 (K = 27
// padded to 32) with BOTH operands split into bf16 hi + lo parts (x = hi + lo exactly to 16 mantissa bits) and the three
// products hi*hi + lo*hi + hi*lo accumulated in fp32: error ~2^-16 relative, far below the bf16 rounding of the output.
// CTA = 8 output rows x 64 pixels; the fp32 input patch (3 x 17 x 132) is staged with aligned 16-byte cp.async (zero fill
// = padding); each thread gathers its A-fragment elements from it with precomputed tap offsets (im2col in registers).
constexpr int kC1M_PW = 132, kC1M_IH = 17;
constexpr uint32_t kC1M_BUF = (3 * kC1M_IH * kC1M_PW * 4 + 127) / 128 * 128;   // TMA variant: patch buffer stride
__device__ __forceinline__ void c1m_split(float v, float w, uint32_t& hi, uint32_t& lo) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(v, w);
  const float2 hf = __bfloat1622float2(h);
  const __nv_bfloat162 l = __floats2bfloat162_rn(v - hf.x, w - hf.y);
  hi = *reinterpret_cast<const uint32_t*>(&h); lo = *reinterpret_cast<const uint32_t*>(&l);
}
// activations: hi = the top 16 bits (truncation, one PRMT for the pair), lo = the exact remainder rounded to bf16
__device__ __forceinline__ void c1m_split_trunc(float v, float w, uint32_t& hi, uint32_t& lo) {
  const uint32_t vb = __float_as_uint(v), wb = __float_as_uint(w);
  hi = __byte_perm(vb, wb, 0x7632);
  const __nv_bfloat162 l = __floats2bfloat162_rn(v - __uint_as_float(vb & 0xffff0000u), w - __uint_as_float(wb & 0xffff0000u));
  lo = *reinterpret_cast<const uint32_t*>(&l);
}
__device__ __forceinline__ void c1m_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <int OFF>
__device__ __forceinline__ float c1m_lds(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1+%2];" : "=f"(v) : "r"(addr), "n"(OFF * 4));
  return v;
}
// one 16-pixel m-tile (MT = its index in the warp's 64-pixel row): every smem offset is an immediate
template <int MT>
__device__ __forceinline__ void c1m_tile(const uint32_t (&pa)[2][4], const uint32_t (&bh)[2][2][2], const uint32_t (&bl)[2][2][2],
                                         float (&acc)[2][4]) {
#pragma unroll
  for (int s = 0; s < 2; ++s) {
    float v[2][4];                                              // [pixel row half][k index q]; row g + 8 is 16 floats further
#pragma unroll
    for (int q = 0; q < 4; ++q) { v[0][q] = c1m_lds<32 * MT>(pa[s][q]); v[1][q] = c1m_lds<32 * MT + 16>(pa[s][q]); }
    uint32_t ah[4], al[4];                                      // a0: (g, k lo pair) a1: (g+8, lo) a2: (g, k hi pair) a3: (g+8, hi)
    c1m_split_trunc(v[0][0], v[0][1], ah[0], al[0]);
    c1m_split_trunc(v[1][0], v[1][1], ah[1], al[1]);
    c1m_split_trunc(v[0][2], v[0][3], ah[2], al[2]);
    c1m_split_trunc(v[1][2], v[1][3], ah[3], al[3]);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      c1m_mma(acc[j], al, bh[s][j][0], bh[s][j][1]);            // small terms first
      c1m_mma(acc[j], ah, bl[s][j][0], bl[s][j][1]);
      c1m_mma(acc[j], ah, bh[s][j][0], bh[s][j][1]);
    }
  }
}
__global__ void __launch_bounds__(256) conv1a_mma_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                         const float* __restrict__ bias, __nv_bfloat16* __restrict__ out,
                                                         int B, int H, int W, float slope, int frames, int tiles_x, int tiles_y,
                                                         int n_tiles, int f16) {
  __shared__ __align__(16) float patch[3 * kC1M_IH * kC1M_PW];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, c = lane & 3;
  const uint32_t p_addr = (uint32_t)__cvta_generic_to_shared(patch);
  // ---- weights (independent of the previous kernel): B fragments of both n-tiles and k-steps, split hi / lo
  uint32_t bh[2][2][2], bl[2][2][2];                           // [k step][n-tile][k half]
  uint32_t pa[2][4];                                            // smem address of this thread's 8 k indices, pixel g of m-tile 0
#pragma unroll
  for (int s = 0; s < 2; ++s)
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      int k = 16 * s + 2 * c + (q & 1) + 8 * (q >> 1);
      k = k < 27 ? k : 26;                                      // K padding: its weights are 0, any tap of the window will do
      const int ci = k / 9, ky = (k - 9 * ci) / 3, kx = k - 9 * ci - 3 * ky;
      // input row 2*oy - 1 + ky -> patch row 2*warp + ky; input column 2*x - 1 + kx -> patch column 2*x + 3 + kx
      pa[s][q] = p_addr + 4u * (uint32_t)((ci * kC1M_IH + ky + 2 * warp) * kC1M_PW + kx + 3 + 2 * g);
    }
#pragma unroll
  for (int s = 0; s < 2; ++s)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int k = 16 * s + 2 * c + 8 * h, co = 8 * j + g;
        const float w0 = k < 27 ? w[co * 27 + k] : 0.f, w1 = k + 1 < 27 ? w[co * 27 + k + 1] : 0.f;
        c1m_split(w0, w1, bh[s][j][h], bl[s][j][h]);
      }
  float bia[2][2];
#pragma unroll
  for (int j = 0; j < 2; ++j) { bia[j][0] = bias[8 * j + 2 * c]; bia[j][1] = bias[8 * j + 2 * c + 1]; }
  pdl_wait();
  pdl_launch_dependents();

  const int Ho = H >> 1, Wo = W >> 1;
  const long hw = (long)H * W;
  const float inv_tx = 1.0f / (float)tiles_x, inv_ty = 1.0f / (float)tiles_y;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    // tile -> (tx, ty, img) without integer divisions (exact below 2^20 tiles, checked by the launcher)
    const int t2 = (int)(((float)tile + 0.5f) * inv_tx); const int tx = tile - t2 * tiles_x;
    const int img = (int)(((float)t2 + 0.5f) * inv_ty); const int ty = t2 - img * tiles_y;
    const float* src = frames ? x + (long)img * 3 * hw : x + ((long)(img % B) * 6 + (img / B) * 3) * hw;
    const int ox0 = tx * 64, oy0 = ty * 8;
    const int xs = 2 * ox0 - 4, ys = 2 * oy0 - 1;               // patch origin; xs is a multiple of 4 like W
    // ---- stage the patch (3 planes x 17 rows x 33 quads of 16 bytes): warp = row (8 + 8, the 17th rows go to warps 0-2),
    // lane = quad; the 33rd quads of the 51 rows are one more pass of threads 0-50.  32-bit element offsets inside the image.
    {
      const int lx = xs + 4 * lane;
      const bool xok = lx >= 0 && lx < W;
      const int ihw = (int)hw;
#pragma unroll
      for (int ci = 0; ci < 3; ++ci)
#pragma unroll
        for (int rh = 0; rh < 2; ++rh) {
          const int r = warp + 8 * rh, iy = ys + r;
          const bool ok = xok && iy >= 0 && iy < H;
          const int off = ok ? ci * ihw + iy * W + lx : 0;
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::
                       "r"(p_addr + (uint32_t)((ci * kC1M_IH + r) * kC1M_PW * 4) + 16u * lane), "l"(src + off), "r"(ok ? 16 : 0) : "memory");
        }
      if (warp < 3) {
        const int iy = ys + 16;
        const bool ok = xok && iy < H;
        const int off = ok ? warp * ihw + iy * W + lx : 0;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::
                     "r"(p_addr + (uint32_t)((warp * kC1M_IH + 16) * kC1M_PW * 4) + 16u * lane), "l"(src + off), "r"(ok ? 16 : 0) : "memory");
      }
      if (tid < 3 * kC1M_IH) {
        const int ci = tid / kC1M_IH, r = tid - ci * kC1M_IH, iy = ys + r;
        const bool ok = xs + 128 < W && iy >= 0 && iy < H;
        const int off = ok ? ci * ihw + iy * W + xs + 128 : 0;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::
                     "r"(p_addr + (uint32_t)(tid * kC1M_PW * 4) + 512u), "l"(src + off), "r"(ok ? 16 : 0) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    const int oy = oy0 + warp;
    if (oy < Ho) {
      __nv_bfloat16* orow = out + (((long)img * Ho + oy) * Wo + ox0 + g) * 16 + 2 * c;
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) {
        if (ox0 + 16 * mt >= Wo) break;
        float acc[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[j][e] = 0.f;
        if (mt == 0) c1m_tile<0>(pa, bh, bl, acc);
        else if (mt == 1) c1m_tile<1>(pa, bh, bl, acc);
        else if (mt == 2) c1m_tile<2>(pa, bh, bl, acc);
        else c1m_tile<3>(pa, bh, bl, acc);
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int h = 0; h < 2; ++h)
            if (ox0 + 16 * mt + g + 8 * h < Wo) {
              const float v0 = acc[j][2 * h] + bia[j][0], v1 = acc[j][2 * h + 1] + bia[j][1];
              *reinterpret_cast<uint32_t*>(orow + (16 * mt + 8 * h) * 16 + 8 * j) = pack2_h16(lrelu(v0, slope), lrelu(v1, slope), f16 & 1);
            }
      }
    }
    __syncthreads();                                            // the patch is re-staged by the next tile
  }
}

// The same kernel with the patch brought in by TMA: ONE 3-D box (132 floats x 17 rows x 3 planes, OOB zero fill = padding)
// per tile into one of two buffers, so the next tile loads while this one computes and no thread spends issue slots on staging.
__global__ void __launch_bounds__(256) conv1a_mma_tma_kernel(const __grid_constant__ CUtensorMap tmx, const float* __restrict__ w,
                                                         const float* __restrict__ bias, __nv_bfloat16* __restrict__ out,
                                                         int B, int H, int W, float slope, int frames, int tiles_x, int tiles_y,
                                                         int n_tiles, int f16) {
  extern __shared__ __align__(128) unsigned char c1m_smem[];   // two patch buffers of kC1M_BUF bytes | two mbarriers
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, c = lane & 3;
  const uint32_t p_addr = (uint32_t)__cvta_generic_to_shared(c1m_smem);
  const uint32_t bar0 = p_addr + 2u * kC1M_BUF;
  if (tid == 0) {
    ptx::prefetch_tmap(&tmx);
    ptx::mbar_init(bar0, 1); ptx::mbar_init(bar0 + 8, 1);
    ptx::fence_barrier_init();
  }
  // ---- weights (independent of the previous kernel): B fragments of both n-tiles and k-steps, split hi / lo
  uint32_t bh[2][2][2], bl[2][2][2];                           // [k step][n-tile][k half]
  uint32_t pa[2][4];                                            // smem address of this thread's 8 k indices, pixel g of m-tile 0
#pragma unroll
  for (int s = 0; s < 2; ++s)
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      int k = 16 * s + 2 * c + (q & 1) + 8 * (q >> 1);
      k = k < 27 ? k : 26;                                      // K padding: its weights are 0, any tap of the window will do
      const int ci = k / 9, ky = (k - 9 * ci) / 3, kx = k - 9 * ci - 3 * ky;
      // input row 2*oy - 1 + ky -> patch row 2*warp + ky; input column 2*x - 1 + kx -> patch column 2*x + 3 + kx
      pa[s][q] = p_addr + 4u * (uint32_t)((ci * kC1M_IH + ky + 2 * warp) * kC1M_PW + kx + 3 + 2 * g);
    }
#pragma unroll
  for (int s = 0; s < 2; ++s)
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int k = 16 * s + 2 * c + 8 * h, co = 8 * j + g;
        const float w0 = k < 27 ? w[co * 27 + k] : 0.f, w1 = k + 1 < 27 ? w[co * 27 + k + 1] : 0.f;
        c1m_split(w0, w1, bh[s][j][h], bl[s][j][h]);
      }
  float bia[2][2];
#pragma unroll
  for (int j = 0; j < 2; ++j) { bia[j][0] = bias[8 * j + 2 * c]; bia[j][1] = bias[8 * j + 2 * c + 1]; }
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();

  const int Ho = H >> 1, Wo = W >> 1;
  const float inv_tx = 1.0f / (float)tiles_x, inv_ty = 1.0f / (float)tiles_y;
  auto decode = [&](int tile, int& ox0, int& oy0, int& img) {
    // tile -> (tx, ty, img) without integer divisions (exact below 2^20 tiles, checked by the launcher)
    const int t2 = (int)(((float)tile + 0.5f) * inv_tx); const int tx = tile - t2 * tiles_x;
    img = (int)(((float)t2 + 0.5f) * inv_ty); const int ty = t2 - img * tiles_y;
    ox0 = tx * 64; oy0 = ty * 8;
  };
  auto issue = [&](int tile, int buf) {                        // one thread
    int ox0, oy0, img;
    decode(tile, ox0, oy0, img);
    const int plane0 = frames ? img * 3 : (img % B) * 6 + (img / B) * 3;
    ptx::mbar_expect_tx(bar0 + 8u * buf, (uint32_t)(3 * kC1M_IH * kC1M_PW * 4));
    ptx::tma_load_3d(p_addr + (uint32_t)buf * kC1M_BUF, &tmx, bar0 + 8u * buf, 2 * ox0 - 4, 2 * oy0 - 1, plane0);
  };
  int tile = blockIdx.x;
  if (tid == 0 && tile < n_tiles) issue(tile, 0);
  uint32_t ph0 = 0, ph1 = 0, pdelta = kC1M_BUF;
  for (int it = 0; tile < n_tiles; tile += gridDim.x, ++it) {
    const int b = it & 1;
    if (tid == 0 && tile + (int)gridDim.x < n_tiles) issue(tile + gridDim.x, b ^ 1);   // released by the sync that ended the last tile
    if (b == 0) { ptx::mbar_wait(bar0, ph0); ph0 ^= 1; } else { ptx::mbar_wait(bar0 + 8, ph1); ph1 ^= 1; }
    int ox0, oy0, img;
    decode(tile, ox0, oy0, img);

    const int oy = oy0 + warp;
    if (oy < Ho) {
      const int opp = (f16 & 2) ? 32 : 16;                       // bit 1: split output, pixel = [hi 16 | lo 16]
      __nv_bfloat16* orow = out + (((long)img * Ho + oy) * Wo + ox0 + g) * opp + 2 * c;
#pragma unroll
      for (int mt = 0; mt < 4; ++mt) {
        if (ox0 + 16 * mt >= Wo) break;
        float acc[2][4];
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[j][e] = 0.f;
        if (mt == 0) c1m_tile<0>(pa, bh, bl, acc);
        else if (mt == 1) c1m_tile<1>(pa, bh, bl, acc);
        else if (mt == 2) c1m_tile<2>(pa, bh, bl, acc);
        else c1m_tile<3>(pa, bh, bl, acc);
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
          for (int h = 0; h < 2; ++h)
            if (ox0 + 16 * mt + g + 8 * h < Wo) {
              const float v0 = acc[j][2 * h] + bia[j][0], v1 = acc[j][2 * h + 1] + bia[j][1];
              const float r0 = lrelu(v0, slope), r1 = lrelu(v1, slope);
              const uint32_t hi = pack2_h16(r0, r1, f16 & 1);
              *reinterpret_cast<uint32_t*>(orow + (16 * mt + 8 * h) * opp + 8 * j) = hi;
              if (f16 & 2) {
                const float2 hf = unpack2_h16(hi, f16 & 1);
                *reinterpret_cast<uint32_t*>(orow + (16 * mt + 8 * h) * opp + 16 + 8 * j) = pack2_h16(r0 - hf.x, r1 - hf.y, f16 & 1);
              }
            }
      }
    }
    __syncthreads();                                            // this buffer may be refilled
#pragma unroll
    for (int s = 0; s < 2; ++s)
#pragma unroll
      for (int q = 0; q < 4; ++q) pa[s][q] += pdelta;
    pdelta = 0u - pdelta;
  }
}


static int launch_conv1a(const float* x, const float* w, const float* bias, void* out, int B, int H, int W, float slope,
                         int frames, cudaStream_t s, int split = 0) {
  const int fmt = half_format() | (split ? 2 : 0);
  static const int use_mma = []() { const char* e = getenv("PWC_CONV1A_MMA"); return e ? atoi(e) : 1; }();
  if (use_mma && (W & 3) == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    const int tiles_x = cdiv(W >> 1, 64), tiles_y = cdiv(H >> 1, 8);
    const long nt = (long)(frames ? B : 2 * B) * tiles_x * tiles_y;
    if (nt < (1L << 20) && 3L * H * W < 0x7fffffffL) {
      static const int ctas = []() { const char* e = getenv("PWC_CONV1A_CTAS"); return e ? atoi(e) : 148 * 4; }();
      static const int use_tma = []() { const char* e = getenv("PWC_CONV1A_TMA"); return e ? atoi(e) : 1; }();
      EncodeTiledFn enc = use_tma ? get_encode_tiled() : nullptr;
      if (enc) {
        const size_t smem = 2 * kC1M_BUF + 16;
        static int attr_dev = -1;                                 // the opt-in is a per-device attribute
        int dev = 0; cudaGetDevice(&dev);
        if (attr_dev != dev) {
          const cudaError_t attr = cudaFuncSetAttribute(conv1a_mma_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                        (int)(2 * kC1M_BUF + 16));
          if (attr != cudaSuccess) return (int)attr;
          attr_dev = dev;
        }
        CUtensorMap tm;
        cuuint64_t dims[3] = {(cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)(frames ? 3 * B : 6 * B)};
        cuuint64_t strides[2] = {(cuuint64_t)W * 4, (cuuint64_t)W * H * 4};
        cuuint32_t box[3] = {(cuuint32_t)kC1M_PW, (cuuint32_t)kC1M_IH, 3}, es[3] = {1, 1, 1};
        if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(x), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
          return (int)cudaErrorInvalidValue;
        return (int)launch_pdl(conv1a_mma_tma_kernel, dim3((unsigned)(nt < ctas ? nt : ctas)), dim3(256), smem, s, tm, w, bias,
                               reinterpret_cast<__nv_bfloat16*>(out), B, H, W, slope, frames, tiles_x, tiles_y, (int)nt, fmt);
      }
      if (split) return (int)cudaErrorNotSupported;
      return (int)launch_pdl(conv1a_mma_kernel, dim3((unsigned)(nt < ctas ? nt : ctas)), dim3(256), 0, s, x, w, bias,
                             reinterpret_cast<__nv_bfloat16*>(out), B, H, W, slope, frames, tiles_x, tiles_y, (int)nt, half_format());
    }
  }
  if (split) return (int)cudaErrorNotSupported;              // the split store exists in the TMA kernel only
  return (int)launch_pdl(conv1a_bf16_kernel, dim3(148 * 8), dim3(256), 0, s, x, w, bias, reinterpret_cast<__nv_bfloat16*>(out),
                         B, H, W, slope, frames, half_format());
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
                                                                   int N, int Hin, int Win, int CinP, int lo_off, int dup_off) {
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
    if (lo_off) { st1(o + lo_off, s0 - ld1(o)); st1(o + lo_off + 1, s1 - ld1(o + 1)); }     // exact remainders of the roundings
    if (dup_off) { st1(o + dup_off, s0); st1(o + dup_off + 1, s1); }
    if (out_b) { float* ob = out_b + e * out_b_pitch; ob[0] = s0; ob[1] = s1; }
  }
}

}  // namespace pwc

// conv1a with a split-precision output (conv_mode 'mixed'): out [2B | N, H/2, W/2, 32] = [hi 16 | lo 16] per pixel
extern "C" int pwc_conv1a_split_fwd(const float* x, const float* w, const float* bias, void* out,
                                    int B, int H, int W, float slope, int frames, void* stream) {
  if (B <= 0 || (H & 1) || (W & 3)) return (int)cudaErrorInvalidValue;
  return pwc::launch_conv1a(x, w, bias, out, B, H, W, slope, frames ? 1 : 0, (cudaStream_t)stream, 1);
}

extern "C" int pwc_conv1a_bf16_fwd(const float* x, const float* w, const float* bias, void* out,
                                   int B, int H, int W, float slope, void* stream) {
  if (B <= 0 || (H & 1) || (W & 1)) return (int)cudaErrorInvalidValue;
  return pwc::launch_conv1a(x, w, bias, out, B, H, W, slope, 0, (cudaStream_t)stream);
}

// conv1a on a stack of N frames [N,3,H,W] fp32 -> [N,H/2,W/2,16] bf16 (video plan: every frame goes through the pyramid once)
extern "C" int pwc_conv1a_frames_bf16_fwd(const float* frames, const float* w, const float* bias, void* out,
                                          int N, int H, int W, float slope, void* stream) {
  if (N <= 0 || (H & 1) || (W & 1)) return (int)cudaErrorInvalidValue;
  return pwc::launch_conv1a(frames, w, bias, out, N, H, W, slope, 1, (cudaStream_t)stream);
}

static int deconv_launch(const void* in, int in_dtype, long in_pitch, int in_coff, int Cin,
                         const float* w, const float* bias,
                         void* out, int out_dtype, long out_pitch, int out_coff,
                         float* out_b, long out_b_pitch,
                         int N, int Hin, int Win, void* stream, int lo_off, int dup_off) {
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
             in_coff, w, bias, reinterpret_cast<TO*>(out), out_pitch, out_coff, out_b, out_b_pitch, N, Hin, Win, CinP, lo_off, dup_off)
    const bool f16 = half_format() != 0;             // PWC_BF16 = "the 16-bit storage format": IEEE half after pwc_set_half_format(1)
    if (in_dtype == PWC_F32 && out_dtype == PWC_F32) PWC_DECONV_S(float, float);
    else if (f16 && in_dtype == PWC_F32 && out_dtype == PWC_BF16) PWC_DECONV_S(float, __half);
    else if (f16 && in_dtype == PWC_BF16 && out_dtype == PWC_BF16) PWC_DECONV_S(__half, __half);
    else if (f16 && in_dtype == PWC_BF16 && out_dtype == PWC_F32) PWC_DECONV_S(__half, float);
    else if (in_dtype == PWC_F32 && out_dtype == PWC_BF16) PWC_DECONV_S(float, __nv_bfloat16);
    else if (in_dtype == PWC_BF16 && out_dtype == PWC_BF16) PWC_DECONV_S(__nv_bfloat16, __nv_bfloat16);
    else if (in_dtype == PWC_BF16 && out_dtype == PWC_F32) PWC_DECONV_S(__nv_bfloat16, float);
    else return (int)cudaErrorInvalidValue;
#undef PWC_DECONV_S
    return (int)cudaGetLastError();
  }
  if (lo_off || dup_off) return (int)cudaErrorNotSupported;
#define PWC_DECONV(TI, TO)                                                                                             \
  launch_pdl(deconv4x4s2_c2_any_kernel<TI, TO>, dim3(blocks), dim3(256), 0, s, reinterpret_cast<const TI*>(in), in_pitch, in_coff, CinP, w, \
                                                           bias, reinterpret_cast<TO*>(out), out_pitch, out_coff, out_b, \
                                                           out_b_pitch, N, Hin, Win, CinP)
  const bool f16 = half_format() != 0;
  if (in_dtype == PWC_F32 && out_dtype == PWC_F32) PWC_DECONV(float, float);
  else if (f16 && in_dtype == PWC_F32 && out_dtype == PWC_BF16) PWC_DECONV(float, __half);
  else if (f16 && in_dtype == PWC_BF16 && out_dtype == PWC_BF16) PWC_DECONV(__half, __half);
  else if (f16 && in_dtype == PWC_BF16 && out_dtype == PWC_F32) PWC_DECONV(__half, float);
  else if (in_dtype == PWC_F32 && out_dtype == PWC_BF16) PWC_DECONV(float, __nv_bfloat16);
  else if (in_dtype == PWC_BF16 && out_dtype == PWC_BF16) PWC_DECONV(__nv_bfloat16, __nv_bfloat16);
  else if (in_dtype == PWC_BF16 && out_dtype == PWC_F32) PWC_DECONV(__nv_bfloat16, float);
  else return (int)cudaErrorInvalidValue;
#undef PWC_DECONV
  return (int)cudaGetLastError();
}

extern "C" int pwc_deconv4x4s2_c2_fwd(const void* in, int in_dtype, long in_pitch, int in_coff, int Cin,
                                      const float* w, const float* bias,
                                      void* out, int out_dtype, long out_pitch, int out_coff,
                                      float* out_b, long out_b_pitch,
                                      int N, int Hin, int Win, void* stream) {
  return deconv_launch(in, in_dtype, in_pitch, in_coff, Cin, w, bias, out, out_dtype, out_pitch, out_coff, out_b, out_b_pitch,
                       N, Hin, Win, stream, 0, 0);
}

// deconvL (2 -> 2 channels, PWCNet.py:208...) with a split-precision result in a 16-bit concat slice: channels out_coff + {0,1}
// take hi = RN16(v), out_coff + lo_off + {0,1} the exact remainders RN16(v - hi), and (dup_off != 0) out_coff + dup_off + {0,1}
// a second copy of hi -- the K-padding form of the split for single-precision consumers.  Cin <= 4 only.
extern "C" int pwc_deconv4x4s2_c2_split_fwd(const void* in, int in_dtype, long in_pitch, int in_coff, int Cin,
                                            const float* w, const float* bias,
                                            void* out, int out_dtype, long out_pitch, int out_coff,
                                            float* out_b, long out_b_pitch,
                                            int N, int Hin, int Win, int lo_off, int dup_off, void* stream) {
  if (lo_off < 0 || dup_off < 0 || out_dtype != PWC_BF16) return (int)cudaErrorInvalidValue;
  return deconv_launch(in, in_dtype, in_pitch, in_coff, Cin, w, bias, out, out_dtype, out_pitch, out_coff, out_b, out_b_pitch,
                       N, Hin, Win, stream, lo_off, dup_off);
}

namespace pwc {
// channel slice copy between two NHWC+pitch 16-bit tensors, 16 bytes per thread (C % 8 == 0): HBM-bound plumbing kernel
__global__ void __launch_bounds__(256) copy_channels_h16_kernel(const uint16_t* __restrict__ src, long sp, const uint16_t* dummy,
                                                                uint16_t* __restrict__ dst, long dp, int cvec, long npix) {
  pdl_wait();
  pdl_launch_dependents();
  const long total = npix * cvec;
  for (long e = (long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long)gridDim.x * blockDim.x) {
    const long px = e / cvec; const int v = (int)(e - px * cvec);
    *reinterpret_cast<uint4*>(dst + px * dp + 8 * v) = *reinterpret_cast<const uint4*>(src + px * sp + 8 * v);
  }
}
}  // namespace pwc

// dst[px, dst_coff + c] = src[px, src_coff + c], c < C, for npix pixels of two NHWC+pitch 16-bit tensors (conv_mode 'mixed':
// the hi plane of the image-1 pyramid features -> the c1 slot of the decoder concat buffer, PWCNet.py:215,229,243,259).
extern "C" int pwc_copy_channels_h16(const void* src, long src_pitch, int src_coff, void* dst, long dst_pitch, int dst_coff,
                                     int C, long npix, void* stream) {
  using namespace pwc;
  if (C <= 0 || (C & 7) || (src_pitch & 7) || (dst_pitch & 7) || (src_coff & 7) || (dst_coff & 7) || npix <= 0 ||
      ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15))
    return (int)cudaErrorInvalidValue;
  const long total = npix * (C / 8);
  const unsigned nb = (unsigned)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  return (int)launch_pdl(copy_channels_h16_kernel, dim3(nb), dim3(256), 0, (cudaStream_t)stream,
                         reinterpret_cast<const uint16_t*>(src) + src_coff, src_pitch, (const uint16_t*)nullptr,
                         reinterpret_cast<uint16_t*>(dst) + dst_coff, dst_pitch, C / 8, npix);
}
