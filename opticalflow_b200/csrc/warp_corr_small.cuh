// Fused warp + cost volume for the SMALL pyramid levels (bf16 features, C = 96 / 128 / 196; a few thousand pixels per
// batch): the work is microseconds, so the kernel is laid out for latency, not throughput -- one CTA per (image, row,
// 32-column segment) and one warp per row displacement, all 9 warps independent between the two CTA barriers:
//   warp d : gathers warped image-2 row y+d-4 (40 columns x C channels, mask folded into the corner weights) into its
//            own shared-memory row buffer, then lane x accumulates the 9 column displacements over all channels
//            (f1 row and warped row as bf16 pairs, odd word stride -> conflict-free), /C, LeakyReLU, bf16 -> staging;
//   CTA    : one TMA tensor store of [32 px][80 ch] into the concat slice + channel 80 from a side array.
// Included by warp_corr.cu (shares WarpCorrParams / warp_entry).
#pragma once

namespace pwc {

constexpr int kSmWarps = 9;
constexpr int kSmThreads = kSmWarps * 32;
constexpr int kSmCols = 40;

struct WarpCorrSmallParams {
  WarpCorrParams g;
  int stride_w;          // words (bf16 pairs) per pixel row in shared memory: C/2 rounded up to odd
  int dbg;
};

__host__ __device__ inline int sm_small_bytes(int stride_w) {
  // staging [32][80] bf16 (TMA source, first: 128-byte aligned) | ch80 [32] u16 | f1 [32][stride] | w2 [9][40][stride]
  // | table [9][40] (float4 + uint2)
  return 32 * 160 + 64 + (32 + kSmWarps * kSmCols) * stride_w * 4 + kSmWarps * kSmCols * 24;
}

__global__ void __launch_bounds__(kSmThreads)
warp_corr_small_kernel(const __grid_constant__ CUtensorMap tmOut, const WarpCorrSmallParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const WarpCorrParams& g = p.g;
  const int C = g.C, H = g.H, W = g.W, S = p.stride_w;
  uint16_t* stg = reinterpret_cast<uint16_t*>(smem_raw);
  uint16_t* ch80 = reinterpret_cast<uint16_t*>(smem_raw + 32 * 160);
  uint32_t* f1s = reinterpret_cast<uint32_t*>(smem_raw + 32 * 160 + 64);
  uint32_t* w2all = f1s + 32 * S;
  float4* tabw_all = reinterpret_cast<float4*>(w2all + kSmWarps * kSmCols * S);
  uint2* tabxy_all = reinterpret_cast<uint2*>(tabw_all + kSmWarps * kSmCols);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int x0 = blockIdx.x * 32, y = blockIdx.y, n = blockIdx.z;
  const __nv_bfloat16* f1 = reinterpret_cast<const __nv_bfloat16*>(g.f1);
  const __nv_bfloat16* f2n = reinterpret_cast<const __nv_bfloat16*>(g.f2) + (long)n * H * W * g.f2_pitch;
  const int ng = C >> 2;                                  // 8-byte channel groups per pixel

  pdl_wait();
  pdl_launch_dependents();
  const bool tr = p.dbg && blockIdx.x == 0 && blockIdx.y == 3 && blockIdx.z == 0 && tid == 4 * 32;
  tc_mark(tr, 0, 0, 0);
  // ---- image-1 row segment -> shared memory (all warps; coalesced 8-byte loads) ----
  for (int it = tid; it < 32 * ng; it += kSmThreads) {
    const int px = it / ng, gq = it - px * ng;
    uint2 v = make_uint2(0u, 0u);
    if (x0 + px < W) v = *reinterpret_cast<const uint2*>(f1 + (((long)n * H + y) * W + x0 + px) * g.f1_pitch + 4 * gq);
    f1s[px * S + 2 * gq] = v.x; f1s[px * S + 2 * gq + 1] = v.y;
  }

  // ---- warp d: warped row yy = y + d - 4 ----
  const int yy = y + warp - 4;
  const bool row_live = (yy >= 0 && yy < H);              // warp-uniform
  uint32_t* w2s = w2all + warp * kSmCols * S;
  float4* tw = tabw_all + warp * kSmCols;
  uint2* txy = tabxy_all + warp * kSmCols;
  if (row_live) {
    for (int e = lane; e < kSmCols; e += 32) {
      const int xx = x0 - 4 + e;
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      uint2 xy = make_uint2(0u, 0u);
      if (xx >= 0 && xx < W) {
        if (g.flow == nullptr) {
          w.x = 1.f;
          xy = make_uint2((uint32_t)xx | ((uint32_t)xx << 16), (uint32_t)yy | ((uint32_t)yy << 16));
        } else {
          const float2 fl = *reinterpret_cast<const float2*>(g.flow + (((long)n * H + yy) * W + xx) * g.flow_pitch);
          warp_entry(g, fl.x, fl.y, xx, yy, w, xy);
        }
      }
      tw[e] = w; txy[e] = xy;
    }
    __syncwarp();
    tc_mark(tr, 0, 0, 1);
    // tasks = (pixel, 8-byte channel group), lanes on adjacent groups; 8 tasks per lane per batch with all 32 corner
    // loads issued before the first blend (shared-memory stores would otherwise serialise the loop on load latency)
    const int ntask = kSmCols * ng;
    for (int t0 = 0; t0 < ntask; t0 += 8 * 32) {
      uint2 cv[8][4];
      float4 wv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int task = t0 + u * 32 + lane;
        const uint2 z = make_uint2(0u, 0u);
        cv[u][0] = cv[u][1] = cv[u][2] = cv[u][3] = z;
        wv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (task < ntask) {
          const int px = task / ng, gq = task - px * ng;
          const float4 w = tw[px];
          const uint2 xy = txy[px];
          wv[u] = w;
          const int xa = xy.x & 0xffff, xb = xy.x >> 16, ya = xy.y & 0xffff, yb = xy.y >> 16;
          const __nv_bfloat16* fc = f2n + 4 * gq;
          if (w.x != 0.f) cv[u][0] = *reinterpret_cast<const uint2*>(fc + ((long)ya * W + xa) * g.f2_pitch);
          if (w.y != 0.f) cv[u][1] = *reinterpret_cast<const uint2*>(fc + ((long)ya * W + xb) * g.f2_pitch);
          if (w.z != 0.f) cv[u][2] = *reinterpret_cast<const uint2*>(fc + ((long)yb * W + xa) * g.f2_pitch);
          if (w.w != 0.f) cv[u][3] = *reinterpret_cast<const uint2*>(fc + ((long)yb * W + xb) * g.f2_pitch);
        }
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int task = t0 + u * 32 + lane;
        if (task < ntask) {
          const int px = task / ng, gq = task - px * ng;
          const float4 w = wv[u];
          const uint32_t av[2] = {cv[u][0].x, cv[u][0].y}, bv[2] = {cv[u][1].x, cv[u][1].y};
          const uint32_t c2v[2] = {cv[u][2].x, cv[u][2].y}, dv[2] = {cv[u][3].x, cv[u][3].y};
          uint32_t o[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            float lo = 0.f, hi = 0.f;
            lo = fmaf(w.x, __uint_as_float(av[e] << 16), lo); hi = fmaf(w.x, __uint_as_float(av[e] & 0xffff0000u), hi);
            lo = fmaf(w.y, __uint_as_float(bv[e] << 16), lo); hi = fmaf(w.y, __uint_as_float(bv[e] & 0xffff0000u), hi);
            lo = fmaf(w.z, __uint_as_float(c2v[e] << 16), lo); hi = fmaf(w.z, __uint_as_float(c2v[e] & 0xffff0000u), hi);
            lo = fmaf(w.w, __uint_as_float(dv[e] << 16), lo); hi = fmaf(w.w, __uint_as_float(dv[e] & 0xffff0000u), hi);
            const __nv_bfloat162 pr = __floats2bfloat162_rn(lo, hi);
            o[e] = *reinterpret_cast<const uint32_t*>(&pr);
          }
          w2s[px * S + 2 * gq] = o[0]; w2s[px * S + 2 * gq + 1] = o[1];
        }
      }
    }
  }
  tc_mark(tr, 0, 0, 2);
  __syncthreads();                                        // f1 row complete (and this warp's w2 row, trivially)
  tc_mark(tr, 0, 0, 3);

  // ---- lane x: 9 column displacements over all channels ----
  float acc[9];
#pragma unroll
  for (int j = 0; j < 9; ++j) acc[j] = 0.f;
  if (row_live) {
    const uint32_t* fa = f1s + lane * S;
    const uint32_t* wb = w2s + lane * S;
#pragma unroll 2
    for (int c2 = 0; c2 < (C >> 1); ++c2) {
      const uint32_t a = fa[c2];
      const float ax = __uint_as_float(a << 16), ay = __uint_as_float(a & 0xffff0000u);
#pragma unroll
      for (int j = 0; j < 9; ++j) {
        const uint32_t b = wb[j * S + c2];
        acc[j] = fmaf(ax, __uint_as_float(b << 16), acc[j]);
        acc[j] = fmaf(ay, __uint_as_float(b & 0xffff0000u), acc[j]);
      }
    }
  }
  tc_mark(tr, 0, 1, 0);
  const float rc = __frcp_rn((float)C);
#pragma unroll
  for (int j = 0; j < 9; ++j) {
    const float q1 = acc[j] * rc;
    const __nv_bfloat16 hv = __float2bfloat16_rn(fmaxf(q1, q1 * g.slope));
    const uint16_t bits = *reinterpret_cast<const uint16_t*>(&hv);
    const int ch = warp * 9 + j;
    if (ch < 80) stg[lane * 80 + ch] = bits; else ch80[lane] = bits;
  }
  ptx::fence_proxy_async();
  __syncthreads();
  if (tid == 0) {
    ptx::tma_store_4d(&tmOut, ptx::smem_addr(stg), 0, x0, y, n);
    ptx::bulk_commit();
  }
  if (tid < 32 && x0 + tid < W)
    reinterpret_cast<uint16_t*>(g.out)[(((long)n * H + y) * W + x0 + tid) * g.out_pitch + 80] = ch80[tid];
  tc_mark(tr, 0, 1, 1);
  if (tid == 0) ptx::bulk_wait0();
  tc_mark(p.dbg && blockIdx.x == 0 && blockIdx.y == 3 && blockIdx.z == 0 && tid == 0, 0, 1, 2);
}

static int launch_warp_corr_small(const WarpCorrParams& g, cudaStream_t s) {
  WarpCorrSmallParams p;
  p.g = g;
  p.stride_w = (g.C / 2) | 1;
  { const char* e = getenv("PWC_CORR_TRACE"); p.dbg = e ? atoi(e) : 0; }
  const int smem = sm_small_bytes(p.stride_w);
  static int attr_dev = -1;
  int dev = 0; cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t e = cudaFuncSetAttribute(warp_corr_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) return (int)e;
    attr_dev = dev;
  }
  if (smem > 200 * 1024) return (int)cudaErrorInvalidValue;
  CUtensorMap tmo;
  int e = make_tmap_nhwc(&tmo, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, g.out, g.out_pitch, 80, g.W, g.H, g.B,
                         80, 32, 1, 1, 1, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (e) return e;
  dim3 grid(cdiv(g.W, 32), g.H, g.B);
  return (int)launch_pdl(warp_corr_small_kernel, grid, dim3(kSmThreads), smem, s, tmo, p);
}

}  // namespace pwc
