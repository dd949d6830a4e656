// 3x3 convolution for the THIN feature-pyramid layers (Cin, Cout in {16, 32, 64}; stride 1 or 2; bf16 NHWC):
// conv1aa, conv1b, conv2a, conv2aa, conv2b, conv3a, conv3aa, conv3b of /root/reference/models/PWCNet.py:52-60
// (call sites PWCNet.py:184-195), bias + LeakyReLU fused, second output / second input for the im1 | im2 halves.
//
// These layers are HBM bound (0.3-2.4 kFLOP per byte moved... i.e. 18-37 us of traffic at batch 8 against 5-15 us of tensor
// time): what matters is that every byte is moved once, coalesced, with the copies of the next tile in flight while the
// current one is multiplied -- not the MMA rate.  On the tcgen05 path their 128 x 256 tiles, TMEM round trip and
// per-item pipeline fill cost 2-4x the traffic time.  Here:
//   * persistent CTAs (grid = SMs x resident CTAs), the whole 3x3 x Cin x Cout weight set staged ONCE per CTA in shared
//     memory (rows padded to an odd number of 16-byte chunks: conflict-free ldmatrix);
//   * tile = 8 output rows x 32 or 64 pixels; its input patch (10 x 34|66, or 17 x 65 at stride 2) is double buffered
//     and fetched with cp.async (zero fill = the convolution's padding) one tile ahead, addresses from a per-pixel table;
//   * warp = one output row = two 16-pixel m-tiles; taps, stride and channel blocks are ldmatrix row addresses into
//     the patch (implicit GEMM, no im2col); mma.sync.m16n8k16 bf16 with fp32 accumulators in registers;
//   * epilogue: bias + LeakyReLU + bf16 in registers, a per-warp shared-memory transpose, 16-byte coalesced stores
//     straight into the concat slice (or the second tensor for the im2 half of the batch).
#include "common.cuh"
#include "ptx_sm100.cuh"
#include "tmap.cuh"
#include "../../include/pwc_b200.h"

namespace pwc {

struct ThinConvParams {
  const __nv_bfloat16* in; long in_pitch; int in_coff;
  const __nv_bfloat16* in2; long in2_pitch; int in2_coff; int n_split_in;
  const __nv_bfloat16* w; const float* bias; int cin, cout;
  __nv_bfloat16* out; long out_pitch; int out_coff;
  __nv_bfloat16* out2; long out2_pitch; int out2_coff; int n_split;
  int N, H, W, Ho, Wo, stride; float slope;
  int tiles_x, tiles_y, ntiles;
  int ih, iw, pp, wp;                          // patch rows / columns, bytes per patch pixel, bytes per weight row
  int off_patch, patch_bytes, off_ostg, off_tab;   // shared-memory layout (weights at 0)
  int mp;                                      // 32-pixel m-tile pairs per tile row: tile = 8 rows x 32*mp pixels
  float* flow_out; const float* residual; long res_pitch;   // flow head: channels 0,1 + residual -> fp32 NCHW [N,2,Ho,Wo]
};

__device__ __forceinline__ void thin_ldsm4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void thin_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void thin_cp16(uint32_t dst, const void* src, bool ok) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(ok ? 16 : 0) : "memory");
}

// NT = output-channel n-tiles (Cout = 8 NT), KS = 16-channel k steps per tap (Cin = 16 KS)
// SPLIT (conv_mode 'mixed', fp32-grade pyramid): every tensor is two 16-bit planes of the same pixel row, hi = RN16(v) in
// channels [0, C) and lo = RN16(v - hi) in [C, 2C); weight rows hold [W_hi | W_lo].  Per K step three products meet in the
// fp32 accumulator: x_hi*W_hi + x_hi*W_lo + x_lo*W_hi (the dropped x_lo*W_lo is 2^-22 relative).  KS counts LOGICAL steps.
template <int NT, int KS, bool F16, bool SPLIT = false>
__global__ void __launch_bounds__(256) conv_thin_kernel(const ThinConvParams p) {
  extern __shared__ __align__(128) uint8_t thin_smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t sm0 = ptx::smem_addr(thin_smem);
  const uint32_t w_addr = sm0, patch_addr = sm0 + p.off_patch;
  constexpr int NCH = (SPLIT ? 4 : 2) * KS;                     // 16-byte chunks per pixel / per weight row
  constexpr int NP = SPLIT ? 2 * NT : NT;                       // 16-byte chunks per output pixel
  constexpr int OP = NP * 16 + 16;                              // bytes per staged output pixel (padded)
  uint8_t* ostg = thin_smem + p.off_ostg + warp * 16 * OP;

  // ---- weights: [9 * Cout rows][Cin] bf16, once per CTA (independent of the previous kernel: before pdl_wait)
  for (int e = tid; e < 9 * NT * 8 * NCH; e += 256) {
    const int row = e / NCH, ch = e - row * NCH;
    thin_cp16(w_addr + row * p.wp + ch * 16, p.w + (long)row * p.cin + ch * 8, true);
  }
  float bias2[NT][2];
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const int c0 = j * 8 + 2 * (lane & 3);
    bias2[j][0] = c0 < p.cout ? p.bias[c0] : 0.f;
    bias2[j][1] = c0 + 1 < p.cout ? p.bias[c0 + 1] : 0.f;
  }
  pdl_wait();
  pdl_launch_dependents();

  // The copies a thread issues for a patch are the same for every tile up to the tile origin: element e = tid + 256 k is
  // (patch pixel e / NCH, 16-byte chunk e % NCH).  The pixel's shared-memory offset, its offset from the tile origin and its
  // (row, column) are computed ONCE (table in shared memory), so that the per-tile staging loop is a table read, a bounds
  // test and one address add per copy.
  uint2* tab = reinterpret_cast<uint2*>(thin_smem + p.off_tab);
  const int nel = p.ih * p.iw * NCH;
  for (int px = tid; px < p.ih * p.iw; px += 256) {
    const int r = px / p.iw, c = px - r * p.iw;
    tab[px] = make_uint2((uint32_t)(r * p.W + c), ((uint32_t)r << 16) | (uint32_t)c);
  }
  __syncthreads();
  // tile walk without divisions: (n, ty, tx) advance by the decomposition of gridDim.x
  const int per_img = p.tiles_x * p.tiles_y;
  const int g_n = gridDim.x / per_img, g_r = gridDim.x - g_n * per_img;
  const int g_y = g_r / p.tiles_x, g_x = g_r - g_y * p.tiles_x;
  auto advance = [&](int& n, int& ty, int& tx) {
    tx += g_x; ty += g_y; n += g_n;
    if (tx >= p.tiles_x) { tx -= p.tiles_x; ++ty; }
    if (ty >= p.tiles_y) { ty -= p.tiles_y; ++n; }
  };

  auto stage = [&](int n, int ty, int tx, int buf) {
    const bool sec = p.in2 != nullptr && n >= p.n_split_in;
    const int sp = (int)(sec ? p.in2_pitch : p.in_pitch);
    const int iy0 = ty * 8 * p.stride - 1, ix0 = tx * 32 * p.mp * p.stride - 1;
    const __nv_bfloat16* src = sec ? p.in2 + ((long)(n - p.n_split_in) * p.H * p.W) * p.in2_pitch + p.in2_coff
                                   : p.in + ((long)n * p.H * p.W) * p.in_pitch + p.in_coff;
    const __nv_bfloat16* org = src + ((long)iy0 * p.W + ix0) * sp;       // may point before the tensor: only offset from
    const uint32_t dst0 = patch_addr + buf * p.patch_bytes;
#pragma unroll 2
    for (int e = tid; e < nel; e += 256) {
      const int px = e / NCH;                                            // NCH is a power of two
      const uint2 t = tab[px];
      const int ch = e & (NCH - 1);
      const bool ok = (unsigned)(iy0 + (int)(t.y >> 16)) < (unsigned)p.H && (unsigned)(ix0 + (int)(t.y & 0xffffu)) < (unsigned)p.W;
      thin_cp16(dst0 + (uint32_t)(px * p.pp) + ch * 16, ok ? (const void*)(org + (int)t.x * sp + ch * 8) : (const void*)src, ok);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  int tile = blockIdx.x;
  int cn = tile / per_img, cty = (tile - cn * per_img) / p.tiles_x, ctx = tile - cn * per_img - cty * p.tiles_x;   // current tile
  if (tile < p.ntiles) stage(cn, cty, ctx, 0);                   // (commits the weight copies too)
  const uint32_t a_lane = (uint32_t)(((warp * p.stride) * p.iw + (lane & 15) * p.stride) * p.pp + (lane >> 4) * 16);
  const uint32_t b_lane = w_addr + (uint32_t)(((lane & 7) + ((lane >> 4) << 3)) * p.wp + ((lane >> 3) & 1) * 16);
  uint32_t a_off[9];                                             // tap offsets inside the patch, fixed for the kernel
#pragma unroll
  for (int tap = 0; tap < 9; ++tap) a_off[tap] = (uint32_t)(((tap / 3) * p.iw + (tap % 3)) * p.pp);
  const uint32_t a_m1 = (uint32_t)(16 * p.stride * p.pp), b_tap = (uint32_t)(NT * 8 * p.wp);
  for (int it = 0; tile < p.ntiles; tile += gridDim.x, ++it) {
    const int buf = it & 1;
    const int next = tile + gridDim.x;
    int nn = cn, nty = cty, ntx = ctx;
    advance(nn, nty, ntx);
    if (next < p.ntiles) {
      stage(nn, nty, ntx, buf ^ 1);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();                                             // patch of this tile (and the weights) visible to all warps

    const int tx = ctx, ty = cty, n = cn;
    const int oy = ty * 8 + warp;
    for (int pr = 0; pr < p.mp && oy < p.Ho; ++pr) {             // 32 pixels (two m-tiles) at a time
      float acc[2][NT][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[m][j][e] = 0.f;
      const uint32_t a_base = patch_addr + buf * p.patch_bytes + a_lane + (uint32_t)pr * 2u * a_m1;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const uint32_t a_tap = a_base + a_off[tap];
        const uint32_t bt = b_lane + tap * b_tap;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          uint32_t a[2][4];
          thin_ldsm4(a[0], a_tap + ks * 32);
          thin_ldsm4(a[1], a_tap + a_m1 + ks * 32);
          uint32_t al[2][4];
          if (SPLIT) {
            thin_ldsm4(al[0], a_tap + (KS + ks) * 32);
            thin_ldsm4(al[1], a_tap + a_m1 + (KS + ks) * 32);
          }
#pragma unroll
          for (int j = 0; j < NT; j += 2) {
            uint32_t b[4];
            thin_ldsm4(b, bt + (uint32_t)(j * 8 * p.wp) + ks * 32);
            mma_h16<F16>(acc[0][j], a[0], b[0], b[1]); mma_h16<F16>(acc[0][j + 1], a[0], b[2], b[3]);
            mma_h16<F16>(acc[1][j], a[1], b[0], b[1]); mma_h16<F16>(acc[1][j + 1], a[1], b[2], b[3]);
            if (SPLIT) {
              mma_h16<F16>(acc[0][j], al[0], b[0], b[1]); mma_h16<F16>(acc[0][j + 1], al[0], b[2], b[3]);
              mma_h16<F16>(acc[1][j], al[1], b[0], b[1]); mma_h16<F16>(acc[1][j + 1], al[1], b[2], b[3]);
              thin_ldsm4(b, bt + (uint32_t)(j * 8 * p.wp) + (KS + ks) * 32);
              mma_h16<F16>(acc[0][j], a[0], b[0], b[1]); mma_h16<F16>(acc[0][j + 1], a[0], b[2], b[3]);
              mma_h16<F16>(acc[1][j], a[1], b[0], b[1]); mma_h16<F16>(acc[1][j + 1], a[1], b[2], b[3]);
            }
          }
        }
      }
      // ---- epilogue: registers -> per-warp staging [16 px][Cout] -> 16-byte coalesced stores
      const bool second = p.out2 != nullptr && n >= p.n_split;
      __nv_bfloat16* dst = second ? p.out2 + ((long)(n - p.n_split) * p.Ho * p.Wo) * p.out2_pitch + p.out2_coff
                                  : p.out + ((long)n * p.Ho * p.Wo) * p.out_pitch + p.out_coff;
      const long dp = second ? p.out2_pitch : p.out_pitch;
      if (p.flow_out != nullptr) {
        // flow head (dc_conv7 + the residual add of PWCNet.py:268): output channels 0,1 live in the lanes with lane % 4 == 0
        // of n-tile 0; they leave as fp32 straight into the NCHW result, no activation
        if ((lane & 3) == 0) {
#pragma unroll
          for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int ox = (tx * p.mp + pr) * 32 + m * 16 + (lane >> 2) + 8 * h;
              if (ox < p.Wo) {
                const long pix = ((long)n * p.Ho + oy) * p.Wo + ox;
                const float2 r = *reinterpret_cast<const float2*>(p.residual + pix * p.res_pitch);
                float* o = p.flow_out + ((long)n * 2 * p.Ho + oy) * p.Wo + ox;
                o[0] = acc[m][0][2 * h] + bias2[0][0] + r.x;
                o[(long)p.Ho * p.Wo] = acc[m][0][2 * h + 1] + bias2[0][1] + r.y;
              }
            }
        }
        continue;
      }
#pragma unroll
      for (int m = 0; m < 2; ++m) {
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const float v0 = acc[m][j][2 * h] + bias2[j][0], v1 = acc[m][j][2 * h + 1] + bias2[j][1];
            const float r0 = fmaxf(v0, v0 * p.slope), r1 = fmaxf(v1, v1 * p.slope);
            const uint32_t hi = pack2_h16(r0, r1, F16 ? 1 : 0);
            *reinterpret_cast<uint32_t*>(ostg + ((lane >> 2) + 8 * h) * OP + j * 16 + 4 * (lane & 3)) = hi;
            if (SPLIT) {
              const float2 hf = unpack2_h16(hi, F16 ? 1 : 0);
              *reinterpret_cast<uint32_t*>(ostg + ((lane >> 2) + 8 * h) * OP + (NT + j) * 16 + 4 * (lane & 3)) =
                  pack2_h16(r0 - hf.x, r1 - hf.y, F16 ? 1 : 0);
            }
          }
        __syncwarp();
        const int ox0 = (tx * p.mp + pr) * 32 + m * 16;
#pragma unroll
        for (int c = lane; c < 16 * NP; c += 32) {
          const int px = c / NP, part = c - px * NP;
          if (ox0 + px < p.Wo) {
            const uint4 v = *reinterpret_cast<const uint4*>(ostg + px * OP + part * 16);
            *reinterpret_cast<uint4*>(dst + ((long)oy * p.Wo + ox0 + px) * dp + part * 8) = v;
          }
        }
        __syncwarp();
      }
    }
    __syncthreads();                                             // every warp is done with this buffer before it is refilled
    cn = nn; cty = nty; ctx = ntx;
  }
}

template <int NT, int KS, bool F16, bool SPLIT = false>
static int launch_thin(ThinConvParams& p, cudaStream_t s) {
  static int attr_dev = -1, num_sms = 0, max_ctas = 1;
  int dev = 0; cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t e = cudaFuncSetAttribute(conv_thin_kernel<NT, KS, F16, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return (int)e;
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (num_sms <= 0) num_sms = 148;
    attr_dev = dev;
  }
  const int wbytes = 9 * NT * 8 * p.wp;
  int smem = 0;
  static const int mp_env = []() { const char* e = getenv("PWC_THIN_MP"); return e ? atoi(e) : 0; }();
  for (p.mp = (mp_env == 1 || p.Wo <= 32) ? 1 : 2; p.mp >= 1; --p.mp) {   // 64-pixel tiles when two patch buffers fit
    p.tiles_x = cdiv(p.Wo, 32 * p.mp);
    p.iw = (32 * p.mp - 1) * p.stride + 3;
    p.off_patch = (wbytes + 127) & ~127;
    p.patch_bytes = (p.ih * p.iw * p.pp + 127) & ~127;
    p.off_ostg = p.off_patch + 2 * p.patch_bytes;
    p.off_tab = p.off_ostg + 8 * 16 * ((SPLIT ? 2 : 1) * NT * 16 + 16);
    smem = p.off_tab + p.ih * p.iw * 8;
    if (smem <= 113 * 1024 || (p.mp == 1 && smem <= 227 * 1024)) break;   // keep two CTAs per SM if the wide tile allows it
  }
  if (p.mp < 1) return (int)cudaErrorInvalidValue;
  const long nt = (long)p.N * p.tiles_x * p.tiles_y;
  if (nt > 0x7fffffffL) return (int)cudaErrorInvalidValue;
  p.ntiles = (int)nt;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&max_ctas, conv_thin_kernel<NT, KS, F16, SPLIT>, 256, smem) != cudaSuccess || max_ctas < 1)
    max_ctas = 1;
  const long slots = (long)num_sms * max_ctas;
  const int grid = (int)(p.ntiles < slots ? p.ntiles : slots);
  return (int)launch_pdl(conv_thin_kernel<NT, KS, F16, SPLIT>, dim3(grid), dim3(256), smem, s, p);
}

}  // namespace pwc

namespace pwc {

// ---- 16 -> 16 channels, stride 1 (conv1aa, conv1b: half-resolution layers, 32-byte pixels; HBM bound: 64 B of traffic per
// output pixel against 9.2 kFLOP).  The generic kernel above spends ~200 warp instructions per 16-pixel m-tile on this shape
// (weights re-read from shared memory for every m-tile, transpose epilogue); here ALL weight fragments of the layer live in
// 36 registers, a tap is one ldmatrix.x4 whose address is a per-lane base + compile-time offset, and the epilogue stores
// bf16 pairs straight from the accumulators (the two n-tiles complete every 32-byte pixel).  Persistent CTAs; tile = 8 rows
// x 64 pixels, warp = row; the 10 x 66-pixel patch is staged with cp.async (zero fill = padding), its two 16-byte halves
// swapped on every other group of 4 pixels so that the 8 row addresses of an ldmatrix phase hit 8 different bank groups.
constexpr int kC16PW = 66, kC16IH = 10, kC16Patch = kC16IH * kC16PW * 32;

template <int OFF>
__device__ __forceinline__ void c16_ldsm4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4+%5];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr), "n"(OFF));
}
// two m-tiles (MT, MT + 1) per call: four independent accumulator chains hide the HMMA / ldmatrix latency
template <int MT, int KY, bool F16>
__device__ __forceinline__ void c16_taps(const uint32_t (&abase)[3], const uint32_t (&bw)[9][2][2], float (&acc)[2][2][4]) {
#pragma unroll
  for (int kx = 0; kx < 3; ++kx) {
    uint32_t a0[4], a1[4];
    c16_ldsm4<(KY * kC16PW + 16 * MT) * 32>(a0, abase[kx]);
    c16_ldsm4<(KY * kC16PW + 16 * MT + 16) * 32>(a1, abase[kx]);
    mma_h16<F16>(acc[0][0], a0, bw[KY * 3 + kx][0][0], bw[KY * 3 + kx][0][1]);
    mma_h16<F16>(acc[0][1], a0, bw[KY * 3 + kx][1][0], bw[KY * 3 + kx][1][1]);
    mma_h16<F16>(acc[1][0], a1, bw[KY * 3 + kx][0][0], bw[KY * 3 + kx][0][1]);
    mma_h16<F16>(acc[1][1], a1, bw[KY * 3 + kx][1][0], bw[KY * 3 + kx][1][1]);
  }
}
template <int MT, bool F16>
__device__ __forceinline__ void c16_tile(const uint32_t (&abase)[3], const uint32_t (&bw)[9][2][2], float (&acc)[2][2][4]) {
  c16_taps<MT, 0, F16>(abase, bw, acc);
  c16_taps<MT, 1, F16>(abase, bw, acc);
  c16_taps<MT, 2, F16>(abase, bw, acc);
}

template <bool F16>
__global__ void __launch_bounds__(256, 3) conv_c16_kernel(const ThinConvParams p, int n_tiles) {
  __shared__ __align__(128) unsigned char patch[kC16Patch];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, c = lane & 3;
  const uint32_t p_addr = (uint32_t)__cvta_generic_to_shared(patch);
  // weights [9][16 cout][16 cin]: B fragment of (tap, n-tile j) = rows 8j+g, k pairs {2c,2c+1} and {2c+8,2c+9}
  uint32_t bw[9][2][2];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const uint32_t* wr = reinterpret_cast<const uint32_t*>(p.w + (t * 16 + 8 * j + g) * 16);
      bw[t][j][0] = __ldg(wr + c); bw[t][j][1] = __ldg(wr + 4 + c);
    }
  float bia[2][2];
#pragma unroll
  for (int j = 0; j < 2; ++j) { bia[j][0] = p.bias[8 * j + 2 * c]; bia[j][1] = p.bias[8 * j + 2 * c + 1]; }
  // ldmatrix row of this lane: matrix m = lane / 8 -> pixels 8*(m & 1) + (lane & 7), channel half m >> 1
  uint32_t abase[3];
  {
    const int px = (lane & 7) + 8 * ((lane >> 3) & 1), half = lane >> 4;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const int q = px + kx;                                     // patch column of m-tile 0 (16 * MT more: same swap phase)
      abase[kx] = p_addr + (uint32_t)((warp * kC16PW + q) * 32 + 16 * (half ^ ((q >> 2) & 1)));
    }
  }
  pdl_wait();
  pdl_launch_dependents();

  const int H = p.H, W = p.W;
  const float inv_tx = 1.0f / (float)p.tiles_x, inv_ty = 1.0f / (float)p.tiles_y;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int t2 = (int)(((float)tile + 0.5f) * inv_tx); const int tx = tile - t2 * p.tiles_x;
    const int img = (int)(((float)t2 + 0.5f) * inv_ty); const int ty = t2 - img * p.tiles_y;
    const int ox0 = tx * 64, oy0 = ty * 8;
    const bool sec = p.in2 != nullptr && img >= p.n_split_in;
    const __nv_bfloat16* src = sec ? p.in2 + ((long)(img - p.n_split_in) * H * W) * p.in2_pitch + p.in2_coff
                                   : p.in + ((long)img * H * W) * p.in_pitch + p.in_coff;
    const int pitch = (int)(sec ? p.in2_pitch : p.in_pitch);
    // ---- stage: 10 rows x 66 pixels x 2 halves of 16 bytes = 40 warp items of 32 chunks (row, 16-pixel quarter) + the last
    // two pixels of the 10 rows (40 chunks: items 40, 41); item = warp + 8 i
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const int it = warp + 8 * i;
      if (it >= 42) break;
      int r, q;
      if (it < 40) { r = it >> 2; q = (lane >> 1) + 16 * (it & 3); }
      else { const int idx = (it - 40) * 32 + lane; r = idx >> 2; q = 64 + ((idx >> 1) & 1); }
      const int h = lane & 1;
      const int iy = oy0 - 1 + r, ix = ox0 - 1 + q;
      const bool ok = r < kC16IH && iy >= 0 && iy < H && ix >= 0 && ix < W;
      const int off = ok ? (iy * W + ix) * pitch + 8 * h : 0;
      if (r < kC16IH)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::
                     "r"(p_addr + (uint32_t)((r * kC16PW + q) * 32 + 16 * (h ^ ((q >> 2) & 1)))), "l"(src + off), "r"(ok ? 16 : 0) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();

    const int oy = oy0 + warp;
    if (oy < H) {
      const bool second = p.out2 != nullptr && img >= p.n_split;
      __nv_bfloat16* orow = second ? p.out2 + (((long)(img - p.n_split) * H + oy) * W + ox0 + g) * p.out2_pitch + p.out2_coff + 2 * c
                                   : p.out + (((long)img * H + oy) * W + ox0 + g) * p.out_pitch + p.out_coff + 2 * c;
      const long opitch = second ? p.out2_pitch : p.out_pitch;
#pragma unroll
      for (int mp = 0; mp < 2; ++mp) {                          // pairs of 16-pixel m-tiles
        if (ox0 + 32 * mp >= W) break;
        float acc[2][2][4];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[m][j][e] = 0.f;
        if (mp == 0) c16_tile<0, F16>(abase, bw, acc);
        else c16_tile<2, F16>(abase, bw, acc);
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int h = 0; h < 2; ++h)
              if (ox0 + 32 * mp + 16 * m + g + 8 * h < W) {
                const float v0 = acc[m][j][2 * h] + bia[j][0], v1 = acc[m][j][2 * h + 1] + bia[j][1];
                *reinterpret_cast<uint32_t*>(orow + (32 * mp + 16 * m + 8 * h) * opitch + 8 * j) =
                    pack2_h16(lrelu(v0, p.slope), lrelu(v1, p.slope), F16 ? 1 : 0);
              }
      }
    }
    __syncthreads();                                            // the patch is re-staged by the next tile
  }
}

// ---- the same layer with the patch brought in by TMA (one box per tile, SWIZZLE_32B = the XOR pattern above, OOB zero fill
// = padding) into two buffers: the next tile loads while this one computes, and the 1320 cp.async + address arithmetic per
// tile disappear from the LSU / issue slots.  Patch rows are 72 pixels apart (2304 B: keeps the swizzle phase of a pixel
// independent of its row, so tap offsets stay immediates).
constexpr int kC16TPW = 72, kC16TPatch = kC16IH * kC16TPW * 32;

template <int MT, int KY, bool F16>
__device__ __forceinline__ void c16t_taps(const uint32_t (&ab)[3], const uint32_t (&bw)[9][2][2], float (&acc)[2][2][4]) {
#pragma unroll
  for (int kx = 0; kx < 3; ++kx) {
    uint32_t a0[4], a1[4];
    c16_ldsm4<(KY * kC16TPW + 16 * MT) * 32>(a0, ab[kx]);
    c16_ldsm4<(KY * kC16TPW + 16 * MT + 16) * 32>(a1, ab[kx]);
    mma_h16<F16>(acc[0][0], a0, bw[KY * 3 + kx][0][0], bw[KY * 3 + kx][0][1]);
    mma_h16<F16>(acc[0][1], a0, bw[KY * 3 + kx][1][0], bw[KY * 3 + kx][1][1]);
    mma_h16<F16>(acc[1][0], a1, bw[KY * 3 + kx][0][0], bw[KY * 3 + kx][0][1]);
    mma_h16<F16>(acc[1][1], a1, bw[KY * 3 + kx][1][0], bw[KY * 3 + kx][1][1]);
  }
}
template <int MT, bool F16>
__device__ __forceinline__ void c16t_tile(const uint32_t (&ab)[3], const uint32_t (&bw)[9][2][2], float (&acc)[2][2][4]) {
  c16t_taps<MT, 0, F16>(ab, bw, acc);
  c16t_taps<MT, 1, F16>(ab, bw, acc);
  c16t_taps<MT, 2, F16>(ab, bw, acc);
}

template <bool F16>
__global__ void __launch_bounds__(256, 3)
conv_c16_tma_kernel(const __grid_constant__ CUtensorMap tm1, const __grid_constant__ CUtensorMap tm2, const ThinConvParams p,
                    int n_tiles) {
  __shared__ __align__(1024) unsigned char patch[2 * kC16TPatch];
  __shared__ __align__(8) unsigned long long bars[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, c = lane & 3;
  const uint32_t p_addr = (uint32_t)__cvta_generic_to_shared(patch);
  const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(bars);
  if (tid == 0) {
    ptx::prefetch_tmap(&tm1);
    ptx::mbar_init(bar0, 1); ptx::mbar_init(bar0 + 8, 1);
    ptx::fence_barrier_init();
  }
  uint32_t bw[9][2][2];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const uint32_t* wr = reinterpret_cast<const uint32_t*>(p.w + (t * 16 + 8 * j + g) * 16);
      bw[t][j][0] = __ldg(wr + c); bw[t][j][1] = __ldg(wr + 4 + c);
    }
  float bia[2][2];
#pragma unroll
  for (int j = 0; j < 2; ++j) { bia[j][0] = p.bias[8 * j + 2 * c]; bia[j][1] = p.bias[8 * j + 2 * c + 1]; }
  uint32_t abase[3];
  {
    const int px = (lane & 7) + 8 * ((lane >> 3) & 1), half = lane >> 4;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const int q = px + kx;
      abase[kx] = p_addr + (uint32_t)((warp * kC16TPW + q) * 32 + 16 * (half ^ ((q >> 2) & 1)));
    }
  }
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();

  const int H = p.H, W = p.W;
  const float inv_tx = 1.0f / (float)p.tiles_x, inv_ty = 1.0f / (float)p.tiles_y;
  auto decode = [&](int tile, int& ox0, int& oy0, int& img) {
    const int t2 = (int)(((float)tile + 0.5f) * inv_tx); const int tx = tile - t2 * p.tiles_x;
    img = (int)(((float)t2 + 0.5f) * inv_ty); const int ty = t2 - img * p.tiles_y;
    ox0 = tx * 64; oy0 = ty * 8;
  };
  auto issue = [&](int tile, int buf) {                        // one thread
    int ox0, oy0, img;
    decode(tile, ox0, oy0, img);
    const bool sec = p.in2 != nullptr && img >= p.n_split_in;
    ptx::mbar_expect_tx(bar0 + 8u * buf, (uint32_t)kC16TPatch);
    ptx::tma_load_4d(p_addr + (uint32_t)buf * kC16TPatch, sec ? &tm2 : &tm1, bar0 + 8u * buf, 0, ox0 - 1, oy0 - 1,
                     sec ? img - p.n_split_in : img);
  };
  int tile = blockIdx.x;
  if (tid == 0 && tile < n_tiles) issue(tile, 0);
  uint32_t ph0 = 0, ph1 = 0;
  for (int it = 0; tile < n_tiles; tile += gridDim.x, ++it) {
    const int b = it & 1;
    if (tid == 0 && tile + (int)gridDim.x < n_tiles) issue(tile + gridDim.x, b ^ 1);   // that buffer was released by the last sync
    if (b == 0) { ptx::mbar_wait(bar0, ph0); ph0 ^= 1; } else { ptx::mbar_wait(bar0 + 8, ph1); ph1 ^= 1; }
    int ox0, oy0, img;
    decode(tile, ox0, oy0, img);
    const uint32_t boff = (uint32_t)b * kC16TPatch;
    const uint32_t ab[3] = {abase[0] + boff, abase[1] + boff, abase[2] + boff};
    const int oy = oy0 + warp;
    if (oy < H) {
      const bool second = p.out2 != nullptr && img >= p.n_split;
      __nv_bfloat16* orow = second ? p.out2 + (((long)(img - p.n_split) * H + oy) * W + ox0 + g) * p.out2_pitch + p.out2_coff + 2 * c
                                   : p.out + (((long)img * H + oy) * W + ox0 + g) * p.out_pitch + p.out_coff + 2 * c;
      const long opitch = second ? p.out2_pitch : p.out_pitch;
#pragma unroll
      for (int mp = 0; mp < 2; ++mp) {
        if (ox0 + 32 * mp >= W) break;
        float acc[2][2][4];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[m][j][e] = 0.f;
        if (mp == 0) c16t_tile<0, F16>(ab, bw, acc);
        else c16t_tile<2, F16>(ab, bw, acc);
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int h = 0; h < 2; ++h)
              if (ox0 + 32 * mp + 16 * m + g + 8 * h < W) {
                const float v0 = acc[m][j][2 * h] + bia[j][0], v1 = acc[m][j][2 * h + 1] + bia[j][1];
                *reinterpret_cast<uint32_t*>(orow + (32 * mp + 16 * m + 8 * h) * opitch + 8 * j) =
                    pack2_h16(lrelu(v0, p.slope), lrelu(v1, p.slope), F16 ? 1 : 0);
              }
      }
    }
    __syncthreads();                                            // buffer b may be refilled
  }
}

static int launch_c16_tma(ThinConvParams& p, cudaStream_t s) {
  p.tiles_x = cdiv(p.W, 64); p.tiles_y = cdiv(p.H, 8);
  const long nt = (long)p.N * p.tiles_x * p.tiles_y;
  const int n1 = p.in2 ? p.n_split_in : p.N;
  CUtensorMap tm1, tm2;
  int e = make_tmap_nhwc(&tm1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p.in + p.in_coff, p.in_pitch, 16, p.W, p.H, n1 > 0 ? n1 : 1, 16,
                         kC16TPW, kC16IH, 1, 1, CU_TENSOR_MAP_SWIZZLE_32B);
  if (e) return e;
  tm2 = tm1;
  if (p.in2 && p.N > p.n_split_in) {
    e = make_tmap_nhwc(&tm2, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p.in2 + p.in2_coff, p.in2_pitch, 16, p.W, p.H, p.N - p.n_split_in,
                       16, kC16TPW, kC16IH, 1, 1, CU_TENSOR_MAP_SWIZZLE_32B);
    if (e) return e;
  }
  static const int ctas = []() { const char* e = getenv("PWC_THIN_C16_CTAS"); return e ? atoi(e) : 148 * 3; }();
  if (half_format())
    return (int)launch_pdl(conv_c16_tma_kernel<true>, dim3((unsigned)(nt < ctas ? nt : ctas)), dim3(256), 0, s, tm1, tm2, p, (int)nt);
  return (int)launch_pdl(conv_c16_tma_kernel<false>, dim3((unsigned)(nt < ctas ? nt : ctas)), dim3(256), 0, s, tm1, tm2, p, (int)nt);
}

// ---- 16 -> 16 channels, stride 1, SPLIT precision (conv1aa, conv1b of the 'mixed' plan): the TMA kernel above with [hi | lo] pixels.
// Pixels are 64 bytes (16 hi + 16 lo channels), the patch comes in SWIZZLE_64B (the 16-byte chunk index is XORed with bits 7-8 of
// the address = (pixel >> 1) & 3: conflict-free ldmatrix; 72-pixel rows = 9 x 512 B keep that phase independent of the row and of
// the 16-pixel m-tile, so taps stay immediates), BOTH weight planes live in registers (72), and a tap issues its three products
// x_hi*W_hi, x_hi*W_lo, x_lo*W_hi on four accumulator chains; the epilogue stores hi and the rounding residual lo straight from
// the accumulators.  The generic kernel spends a quarter of its instructions staging the patch with cp.async and re-reads the
// weights from shared memory for every m-tile.
constexpr int kC16SPW = 72, kC16SPatch = kC16IH * kC16SPW * 64;

template <int MT, int KY, bool F16>
__device__ __forceinline__ void c16s_taps(const uint32_t (&ah)[3], const uint32_t (&al)[3], const uint32_t (&bh)[9][2][2],
                                          const uint32_t (&bl)[9][2][2], float (&acc)[2][2][4]) {
#pragma unroll
  for (int kx = 0; kx < 3; ++kx) {
    constexpr int O0 = (KY * kC16SPW + 16 * MT) * 64, O1 = O0 + 16 * 64;
    const int t = KY * 3 + kx;
    uint32_t a0[4], a1[4];
    c16_ldsm4<O0>(a0, ah[kx]);
    c16_ldsm4<O1>(a1, ah[kx]);
    mma_h16<F16>(acc[0][0], a0, bh[t][0][0], bh[t][0][1]); mma_h16<F16>(acc[0][1], a0, bh[t][1][0], bh[t][1][1]);
    mma_h16<F16>(acc[1][0], a1, bh[t][0][0], bh[t][0][1]); mma_h16<F16>(acc[1][1], a1, bh[t][1][0], bh[t][1][1]);
    mma_h16<F16>(acc[0][0], a0, bl[t][0][0], bl[t][0][1]); mma_h16<F16>(acc[0][1], a0, bl[t][1][0], bl[t][1][1]);
    mma_h16<F16>(acc[1][0], a1, bl[t][0][0], bl[t][0][1]); mma_h16<F16>(acc[1][1], a1, bl[t][1][0], bl[t][1][1]);
    c16_ldsm4<O0>(a0, al[kx]);
    c16_ldsm4<O1>(a1, al[kx]);
    mma_h16<F16>(acc[0][0], a0, bh[t][0][0], bh[t][0][1]); mma_h16<F16>(acc[0][1], a0, bh[t][1][0], bh[t][1][1]);
    mma_h16<F16>(acc[1][0], a1, bh[t][0][0], bh[t][0][1]); mma_h16<F16>(acc[1][1], a1, bh[t][1][0], bh[t][1][1]);
  }
}

template <bool F16>
__global__ void __launch_bounds__(256, 2)
conv_c16_split_tma_kernel(const __grid_constant__ CUtensorMap tm1, const ThinConvParams p, int n_tiles) {
  extern __shared__ __align__(1024) uint8_t c16s_smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, c = lane & 3;
  const uint32_t p_addr = (ptx::smem_addr(c16s_smem) + 1023u) & ~1023u;
  const uint32_t bar0 = p_addr + 2u * kC16SPatch;
  if (tid == 0) {
    ptx::prefetch_tmap(&tm1);
    ptx::mbar_init(bar0, 1); ptx::mbar_init(bar0 + 8, 1);
    ptx::fence_barrier_init();
  }
  // weight rows [9 * 16][W_hi(16) | W_lo(16)]: B fragments of both planes
  uint32_t bh[9][2][2], bl[9][2][2];
#pragma unroll
  for (int t = 0; t < 9; ++t)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const uint32_t* wr = reinterpret_cast<const uint32_t*>(p.w + (t * 16 + 8 * j + g) * 32);
      bh[t][j][0] = __ldg(wr + c); bh[t][j][1] = __ldg(wr + 4 + c);
      bl[t][j][0] = __ldg(wr + 8 + c); bl[t][j][1] = __ldg(wr + 12 + c);
    }
  float bia[2][2];
#pragma unroll
  for (int j = 0; j < 2; ++j) { bia[j][0] = p.bias[8 * j + 2 * c]; bia[j][1] = p.bias[8 * j + 2 * c + 1]; }
  uint32_t ah[3], al[3];
  {
    const int px = (lane & 7) + 8 * ((lane >> 3) & 1), half = lane >> 4;
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const int q = px + kx, sw = (q >> 1) & 3;
      ah[kx] = p_addr + (uint32_t)((warp * kC16SPW + q) * 64 + 16 * (half ^ sw));
      al[kx] = p_addr + (uint32_t)((warp * kC16SPW + q) * 64 + 16 * ((half + 2) ^ sw));
    }
  }
  __syncthreads();
  pdl_wait();
  pdl_launch_dependents();

  const int H = p.H, W = p.W;
  const float inv_tx = 1.0f / (float)p.tiles_x, inv_ty = 1.0f / (float)p.tiles_y;
  auto decode = [&](int tile, int& ox0, int& oy0, int& img) {
    const int t2 = (int)(((float)tile + 0.5f) * inv_tx); const int tx = tile - t2 * p.tiles_x;
    img = (int)(((float)t2 + 0.5f) * inv_ty); const int ty = t2 - img * p.tiles_y;
    ox0 = tx * 64; oy0 = ty * 8;
  };
  auto issue = [&](int tile, int buf) {                        // one thread
    int ox0, oy0, img;
    decode(tile, ox0, oy0, img);
    ptx::mbar_expect_tx(bar0 + 8u * buf, (uint32_t)kC16SPatch);
    ptx::tma_load_4d(p_addr + (uint32_t)buf * kC16SPatch, &tm1, bar0 + 8u * buf, 0, ox0 - 1, oy0 - 1, img);
  };
  int tile = blockIdx.x;
  if (tid == 0 && tile < n_tiles) issue(tile, 0);
  uint32_t ph0 = 0, ph1 = 0;
  for (int it = 0; tile < n_tiles; tile += gridDim.x, ++it) {
    const int b = it & 1;
    if (tid == 0 && tile + (int)gridDim.x < n_tiles) issue(tile + gridDim.x, b ^ 1);   // that buffer was released by the last sync
    if (b == 0) { ptx::mbar_wait(bar0, ph0); ph0 ^= 1; } else { ptx::mbar_wait(bar0 + 8, ph1); ph1 ^= 1; }
    int ox0, oy0, img;
    decode(tile, ox0, oy0, img);
    const uint32_t boff = (uint32_t)b * kC16SPatch;
    const uint32_t abh[3] = {ah[0] + boff, ah[1] + boff, ah[2] + boff};
    const uint32_t abl[3] = {al[0] + boff, al[1] + boff, al[2] + boff};
    const int oy = oy0 + warp;
    if (oy < H) {
      __nv_bfloat16* orow = p.out + (((long)img * H + oy) * W + ox0 + g) * p.out_pitch + p.out_coff + 2 * c;
      const long opitch = p.out_pitch;
#pragma unroll
      for (int mp = 0; mp < 2; ++mp) {
        if (ox0 + 32 * mp >= W) break;
        float acc[2][2][4];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[m][j][e] = 0.f;
        if (mp == 0) {
          c16s_taps<0, 0, F16>(abh, abl, bh, bl, acc); c16s_taps<0, 1, F16>(abh, abl, bh, bl, acc); c16s_taps<0, 2, F16>(abh, abl, bh, bl, acc);
        } else {
          c16s_taps<2, 0, F16>(abh, abl, bh, bl, acc); c16s_taps<2, 1, F16>(abh, abl, bh, bl, acc); c16s_taps<2, 2, F16>(abh, abl, bh, bl, acc);
        }
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int h = 0; h < 2; ++h)
              if (ox0 + 32 * mp + 16 * m + g + 8 * h < W) {
                const float v0 = acc[m][j][2 * h] + bia[j][0], v1 = acc[m][j][2 * h + 1] + bia[j][1];
                const float r0 = fmaxf(v0, v0 * p.slope), r1 = fmaxf(v1, v1 * p.slope);      // 0 <= slope <= 1
                const uint32_t hi = pack2_h16(r0, r1, F16 ? 1 : 0);
                const float2 hb = unpack2_h16(hi, F16 ? 1 : 0);
                uint32_t* o = reinterpret_cast<uint32_t*>(orow + (32 * mp + 16 * m + 8 * h) * opitch + 8 * j);
                o[0] = hi;
                o[8] = pack2_h16(r0 - hb.x, r1 - hb.y, F16 ? 1 : 0);                        // lo plane: 16 channels further
              }
      }
    }
    __syncthreads();                                            // buffer b may be refilled
  }
}

static int launch_c16_split_tma(ThinConvParams& p, cudaStream_t s) {
  p.tiles_x = cdiv(p.W, 64); p.tiles_y = cdiv(p.H, 8);
  const long nt = (long)p.N * p.tiles_x * p.tiles_y;
  CUtensorMap tm1;
  int e = make_tmap_nhwc(&tm1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, p.in + p.in_coff, p.in_pitch, 32, p.W, p.H, p.N, 32,
                         kC16SPW, kC16IH, 1, 1, CU_TENSOR_MAP_SWIZZLE_64B);
  if (e) return e;
  const int smem = 2 * kC16SPatch + 16 + 1024;
  static int attr_dev = -1;
  int dev = 0; cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t ce = cudaFuncSetAttribute(conv_c16_split_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (ce == cudaSuccess) ce = cudaFuncSetAttribute(conv_c16_split_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (ce != cudaSuccess) return (int)ce;
    attr_dev = dev;
  }
  static const int ctas = []() { const char* e = getenv("PWC_THIN_C16S_CTAS"); return e ? atoi(e) : 148 * 2; }();
  if (half_format())
    return (int)launch_pdl(conv_c16_split_tma_kernel<true>, dim3((unsigned)(nt < ctas ? nt : ctas)), dim3(256), smem, s, tm1, p, (int)nt);
  return (int)launch_pdl(conv_c16_split_tma_kernel<false>, dim3((unsigned)(nt < ctas ? nt : ctas)), dim3(256), smem, s, tm1, p, (int)nt);
}

static int launch_c16(ThinConvParams& p, cudaStream_t s) {
  p.tiles_x = cdiv(p.W, 64); p.tiles_y = cdiv(p.H, 8);
  const long nt = (long)p.N * p.tiles_x * p.tiles_y;
  static const int ctas = []() { const char* e = getenv("PWC_THIN_C16_CTAS"); return e ? atoi(e) : 148 * 3; }();
  if (half_format())
    return (int)launch_pdl(conv_c16_kernel<true>, dim3((unsigned)(nt < ctas ? nt : ctas)), dim3(256), 0, s, p, (int)nt);
  return (int)launch_pdl(conv_c16_kernel<false>, dim3((unsigned)(nt < ctas ? nt : ctas)), dim3(256), 0, s, p, (int)nt);
}

}  // namespace pwc

static int thin_dispatch(pwc::ThinConvParams& p, int nt8, int ks, cudaStream_t s) {
  using namespace pwc;
#define PWC_THIN(NTv, KSv) \
  if (nt8 == NTv && ks == KSv) return half_format() ? launch_thin<NTv, KSv, true>(p, s) : launch_thin<NTv, KSv, false>(p, s)
  PWC_THIN(2, 1); PWC_THIN(4, 1); PWC_THIN(8, 1);
  PWC_THIN(2, 2); PWC_THIN(4, 2); PWC_THIN(8, 2);
  PWC_THIN(2, 4); PWC_THIN(4, 4); PWC_THIN(8, 4);
#undef PWC_THIN
  return (int)cudaErrorInvalidValue;
}

// Flow head on the thin-layer kernel: 3x3 conv Cin -> 2 (weights zero-padded to 16 output rows), + bias + residual,
// fp32 NCHW result.  dc_conv7 and `flow2 + dc_conv7(x)` of PWCNet.py:132,268.
extern "C" int pwc_conv3x3_flow_head_fwd(const void* in, long in_pitch, int in_coff, int cin,
                                         const void* w_packed16, const float* bias2, const float* residual, long res_pitch,
                                         float* out_nchw, int N, int H, int W, void* stream) {
  using namespace pwc;
  if (N <= 0 || H <= 0 || W <= 0 || !(cin == 16 || cin == 32 || cin == 64) || (in_pitch & 7) || (in_coff & 7) || !residual ||
      (res_pitch & 1) || ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(w_packed16)) & 15) ||
      (reinterpret_cast<uintptr_t>(residual) & 7))
    return (int)cudaErrorInvalidValue;
  ThinConvParams p;
  p.in = reinterpret_cast<const __nv_bfloat16*>(in); p.in_pitch = in_pitch; p.in_coff = in_coff;
  p.in2 = nullptr; p.in2_pitch = 0; p.in2_coff = 0; p.n_split_in = N;
  p.w = reinterpret_cast<const __nv_bfloat16*>(w_packed16); p.bias = bias2; p.cin = cin; p.cout = 2;
  p.out = nullptr; p.out_pitch = 0; p.out_coff = 0; p.out2 = nullptr; p.out2_pitch = 0; p.out2_coff = 0; p.n_split = N;
  p.N = N; p.H = H; p.W = W; p.stride = 1; p.slope = 1.f; p.Ho = H; p.Wo = W;
  p.tiles_y = cdiv(H, 8); p.ih = 10; p.pp = 2 * cin + 16; p.wp = 2 * cin + 16;
  p.flow_out = out_nchw; p.residual = residual; p.res_pitch = res_pitch;
  return thin_dispatch(p, 2, cin / 16, (cudaStream_t)stream);
}

extern "C" int pwc_conv3x3_thin_fwd(const void* in, long in_pitch, int in_coff,
                                    const void* in2, long in2_pitch, int in2_coff, int n_split_in, int cin,
                                    const void* w_packed, const float* bias, int cout,
                                    void* out, long out_pitch, int out_coff,
                                    void* out2, long out2_pitch, int out2_coff, int n_split,
                                    int N, int H, int W, int stride, float slope, void* stream) {
  using namespace pwc;
  const bool cin_ok = cin == 16 || cin == 32 || cin == 64, cout_ok = cout == 16 || cout == 32 || cout == 64;
  if (N <= 0 || H <= 0 || W <= 0 || (stride != 1 && stride != 2) || !cin_ok || !cout_ok ||
      (in_pitch & 7) || (in_coff & 7) || (in2 && ((in2_pitch & 7) || (in2_coff & 7))) ||
      (out_pitch & 7) || (out_coff & 7) || (out2 && ((out2_pitch & 7) || (out2_coff & 7))) || slope < 0.f || slope > 1.f ||
      ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(in2) | reinterpret_cast<uintptr_t>(out) |
        reinterpret_cast<uintptr_t>(out2) | reinterpret_cast<uintptr_t>(w_packed)) & 15))
    return (int)cudaErrorInvalidValue;
  ThinConvParams p;
  p.in = reinterpret_cast<const __nv_bfloat16*>(in); p.in_pitch = in_pitch; p.in_coff = in_coff;
  p.in2 = reinterpret_cast<const __nv_bfloat16*>(in2); p.in2_pitch = in2_pitch; p.in2_coff = in2_coff; p.n_split_in = n_split_in;
  p.w = reinterpret_cast<const __nv_bfloat16*>(w_packed); p.bias = bias; p.cin = cin; p.cout = cout;
  p.out = reinterpret_cast<__nv_bfloat16*>(out); p.out_pitch = out_pitch; p.out_coff = out_coff;
  p.out2 = reinterpret_cast<__nv_bfloat16*>(out2); p.out2_pitch = out2_pitch; p.out2_coff = out2_coff; p.n_split = n_split;
  p.N = N; p.H = H; p.W = W; p.stride = stride; p.slope = slope;
  p.Ho = (H - 1) / stride + 1; p.Wo = (W - 1) / stride + 1;
  p.tiles_y = cdiv(p.Ho, 8);
  p.ih = 7 * stride + 3;
  p.pp = 2 * cin + 16; p.wp = 2 * cin + 16;
  p.flow_out = nullptr; p.residual = nullptr; p.res_pitch = 0;
  static const int use_c16 = []() { const char* e = getenv("PWC_THIN_C16"); return e ? atoi(e) : 1; }();
  const long max_pitch = in_pitch > in2_pitch ? in_pitch : in2_pitch;
  if (use_c16 && cin == 16 && cout == 16 && stride == 1 && (long)N * cdiv(H, 8) * cdiv(W, 64) < (1L << 20) &&
      (long)H * W * max_pitch < 0x7fffffffL)
  {
    static const int use_tma = []() { const char* e = getenv("PWC_THIN_C16_TMA"); return e ? atoi(e) : 1; }();
    return use_tma ? launch_c16_tma(p, (cudaStream_t)stream) : launch_c16(p, (cudaStream_t)stream);
  }
  return thin_dispatch(p, cout / 8, cin / 16, (cudaStream_t)stream);
}

// Split-precision form of the thin layers (conv_mode 'mixed': fp32-grade pyramid levels 1 and 2, PWCNet.py:52-56,184-187):
// in / out are [hi | lo] plane pairs -- 2*cin / 2*cout 16-bit channels per pixel at in_coff / out_coff --, w_packed holds rows
// [9][cout][W_hi(cin) | W_lo(cin)] (pwc_pack_conv_weight_bf16 with Cin = span = kc = 2*cin).  cin, cout in {16, 32}, stride 1.
extern "C" int pwc_conv3x3_thin_split_strided_fwd(const void* in, long in_pitch, int in_coff, int cin,
                                                  const void* w_packed, const float* bias, int cout,
                                                  void* out, long out_pitch, int out_coff,
                                                  int N, int H, int W, int stride, float slope, void* stream) {
  using namespace pwc;
  if (N <= 0 || H <= 0 || W <= 0 || (stride != 1 && stride != 2) ||
      !(cin == 16 || cin == 32) || !(cout == 16 || cout == 32) || (in_pitch & 7) || (in_coff & 7) ||
      (out_pitch & 7) || (out_coff & 7) || slope < 0.f || slope > 1.f || in_pitch < 2 * cin || out_pitch < 2 * cout ||
      ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(w_packed)) & 15))
    return (int)cudaErrorInvalidValue;
  ThinConvParams p;
  p.in = reinterpret_cast<const __nv_bfloat16*>(in); p.in_pitch = in_pitch; p.in_coff = in_coff;
  p.in2 = nullptr; p.in2_pitch = 0; p.in2_coff = 0; p.n_split_in = N;
  p.w = reinterpret_cast<const __nv_bfloat16*>(w_packed); p.bias = bias; p.cin = 2 * cin; p.cout = cout;
  p.out = reinterpret_cast<__nv_bfloat16*>(out); p.out_pitch = out_pitch; p.out_coff = out_coff;
  p.out2 = nullptr; p.out2_pitch = 0; p.out2_coff = 0; p.n_split = N;
  p.N = N; p.H = H; p.W = W; p.stride = stride; p.slope = slope;
  p.Ho = (H - 1) / stride + 1; p.Wo = (W - 1) / stride + 1;
  p.tiles_y = cdiv(p.Ho, 8); p.ih = 7 * stride + 3;
  p.pp = 4 * cin + 16; p.wp = 4 * cin + 16;
  p.flow_out = nullptr; p.residual = nullptr; p.res_pitch = 0;
  cudaStream_t s = (cudaStream_t)stream;
  const bool f16 = half_format() != 0;
  static const int use_c16s = []() { const char* e = getenv("PWC_THIN_C16_SPLIT"); return e ? atoi(e) : 1; }();
  if (use_c16s && cin == 16 && cout == 16 && stride == 1 && (long)N * cdiv(H, 8) * cdiv(W, 64) < (1L << 20))
    return launch_c16_split_tma(p, s);
#define PWC_THIN_SPLIT(NTv, KSv) \
  if (cout == 8 * NTv && cin == 16 * KSv) return f16 ? launch_thin<NTv, KSv, true, true>(p, s) : launch_thin<NTv, KSv, false, true>(p, s)
  PWC_THIN_SPLIT(2, 1); PWC_THIN_SPLIT(4, 1); PWC_THIN_SPLIT(2, 2); PWC_THIN_SPLIT(4, 2);
#undef PWC_THIN_SPLIT
  return (int)cudaErrorInvalidValue;
}

extern "C" int pwc_conv3x3_thin_split_fwd(const void* in, long in_pitch, int in_coff, int cin,
                                          const void* w_packed, const float* bias, int cout,
                                          void* out, long out_pitch, int out_coff,
                                          int N, int H, int W, float slope, void* stream) {
  return pwc_conv3x3_thin_split_strided_fwd(in, in_pitch, in_coff, cin, w_packed, bias, cout, out, out_pitch, out_coff, N, H, W, 1,
                                            slope, stream);
}
