// Fused warp + cost volume, bf16 NHWC features of any width (C = 32 ... 196), warp-level tensor-core version:
// every warp gathers, multiplies and writes out -- no role specialisation, latency is hidden by 2 CTAs (18 warps) per SM.
//
// A CTA owns a strip of XS = 16*NMI output columns of one image and slides down a range of rows in blocks of R rows.
// Shared memory holds a ring of R+8 WARPED image-2 rows (XS+8 pixels each; bilinear gather with the 0.9999 mask folded
// into the corner weights -- the warped features never exist in HBM; each row is gathered once per strip range and
// reused by the 9 row displacements x R rows that need it) and the R image-1 rows of the block.  Pixel rows are padded
// to an odd number of 16-byte chunks, so ldmatrix reads and the gather's 16-byte stores are bank-conflict free.
// The 81 dot products of a pixel are the near-diagonal band of a GEMM (f1 row) x (warped row)^T:
//     D[x][x'] = sum_c f1[y][x0+x][c] * w2[y+dy][x0-4+x'][c],     out channel 9*(dy+4) + (dx+4) = D[x][x + dx + 4]
// One warp task = (row, dy): per 16-pixel m-tile only the 3 n-tiles (24 columns) that hold its band are multiplied
// (mma.sync.m16n8k16 bf16, fp32 accumulate).  In the mma accumulator layout every thread knows its (row, column) at
// compile time up to the lane index, so the band elements are written straight from the accumulator registers to the
// staging tile [pixel][80 ch] -- no transposition, no shuffles.  /C, LeakyReLU and the bf16 rounding are applied on
// the way; each output row leaves as one TMA tensor store into the concat slice (channel 80 by plain stores).
// Included by warp_corr.cu (shares WarpCorrParams / warp_entry).
#pragma once

namespace pwc {

constexpr int kMmaThreadsMax = 576;

struct WarpCorrMmaParams {
  WarpCorrParams g;
  int R;                 // output rows per block
  int nk;                // K = 16 steps: ceil(C / 16)
  int pitch;             // shared-memory bytes per pixel: 32*nk + 16
  int strips, ranges, blocks, num_items;
  int off_f1, off_stg, off_c80;     // byte offsets in shared memory (ring at 0; the table aliases the staging tile)
};

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2(uint32_t (&r)[2], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// bilinear blend of four bf16 pairs (one 32-bit word per corner): fp32 FMAs in nw, ne, sw, se order, one rounding
template <bool F16>
__device__ __forceinline__ uint32_t blend_pair(const float4& w, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  float lo = 0.f, hi = 0.f;
  const float2 fa = unpack2_h16(a, F16), fb = unpack2_h16(b, F16), fc = unpack2_h16(c, F16), fd = unpack2_h16(d, F16);
  lo = fmaf(w.x, fa.x, lo); hi = fmaf(w.x, fa.y, hi);
  lo = fmaf(w.y, fb.x, lo); hi = fmaf(w.y, fb.y, hi);
  lo = fmaf(w.z, fc.x, lo); hi = fmaf(w.z, fc.y, hi);
  lo = fmaf(w.w, fd.x, lo); hi = fmaf(w.w, fd.y, hi);
  return pack2_h16(lo, hi, F16);
}

// 16 bytes (8 channels) at p, or its first `nvalid` channels (4) zero-extended
__device__ __forceinline__ uint4 ld_cg(const __nv_bfloat16* p, int nvalid) {
  if (nvalid >= 8) return *reinterpret_cast<const uint4*>(p);
  const uint2 h = *reinterpret_cast<const uint2*>(p);
  return make_uint4(h.x, h.y, 0u, 0u);
}

template <int NMI, bool F16>
__global__ void __launch_bounds__(kMmaThreadsMax, 1)
warp_corr_mma_kernel(const __grid_constant__ CUtensorMap tmOut, const WarpCorrMmaParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr int XS = 16 * NMI, WX = XS + 8, NT = 2 * NMI + 1;      // strip width, window width, n-tiles of the window
  constexpr int kU = 4;                                             // gather tasks in flight per thread
  const WarpCorrParams& g = p.g;
  const int C = g.C, H = g.H, W = g.W, R = p.R, RR = R + 8;
  const int pitch = p.pitch, ncg = 2 * p.nk;
  uint8_t* ring = smem_raw;
  uint8_t* f1s = smem_raw + p.off_f1;
  uint16_t* stg = reinterpret_cast<uint16_t*>(smem_raw + p.off_stg);
  uint16_t* c80 = reinterpret_cast<uint16_t*>(smem_raw + p.off_c80);
  // sample table (live between the TMA-store read of the staging tile, which it aliases, and phase 3):
  // corner weights | element offsets of the 4 corners inside image n of f2 | ring byte offset of the pixel (+1: any weight != 0)
  float4* tabw = reinterpret_cast<float4*>(smem_raw + p.off_stg);
  uint4* taboff = reinterpret_cast<uint4*>(tabw + RR * WX);
  uint32_t* tabdst = reinterpret_cast<uint32_t*>(taboff + RR * WX);
  const uint32_t ring_addr = ptx::smem_addr(ring), f1s_addr = ptx::smem_addr(f1s), stg_addr = ptx::smem_addr(stg);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthr = blockDim.x, nwarps = nthr >> 5;
  const __nv_bfloat16* f1 = reinterpret_cast<const __nv_bfloat16*>(g.f1);
  const __nv_bfloat16* f2 = reinterpret_cast<const __nv_bfloat16*>(g.f2);
  uint16_t* outp = reinterpret_cast<uint16_t*>(g.out);
  const float rc = __frcp_rn((float)C);
  const int dq = nthr / ncg, dr = nthr - dq * ncg;                  // task index step nthr as (pixel, channel group)

  pdl_wait();                       // features / flow come from the previous kernels (PDL: see common.cuh)
  pdl_launch_dependents();

  for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
    const int per_img = p.strips * p.ranges;
    const int n = item / per_img;
    const int rem_i = item - n * per_img;
    const int rg = rem_i / p.strips;
    const int x0 = (rem_i - rg * p.strips) * XS;
    const int qb = p.blocks / p.ranges, rb = p.blocks - qb * p.ranges;
    const int b0 = rg * qb + min(rg, rb), T = qb + (rg < rb ? 1 : 0);
    const __nv_bfloat16* f2n = f2 + (long)n * H * W * g.f2_pitch;

    for (int t = 0; t < T; ++t) {
      const int y0 = (b0 + t) * R;
      const int ynew0 = (t == 0) ? y0 - 4 : y0 + 4;                 // first warped row to gather, and how many
      const int nnew = (t == 0) ? RR : R;

      // ---------------- phase 1: sample table of the new warped rows ----------------
      if (tid < R) ptx::bulk_wait_read0();                          // previous block's TMA stores have read the staging tile
      __syncthreads();
      for (int e = tid; e < nnew * WX; e += nthr) {
        const int i = e / WX, qx = e - i * WX;
        const int yy = ynew0 + i, xx = x0 - 4 + qx;
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        uint2 xy = make_uint2(0u, 0u);
        if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
          if (g.flow == nullptr) {
            w.x = 1.f;
            xy = make_uint2((uint32_t)xx | ((uint32_t)xx << 16), (uint32_t)yy | ((uint32_t)yy << 16));
          } else {
            const float2 fl = *reinterpret_cast<const float2*>(g.flow + (((long)n * H + yy) * W + xx) * g.flow_pitch);
            warp_entry(g, fl.x, fl.y, xx, yy, w, xy);
          }
        }
        const int slot = (yy + 4) % RR;                             // yy >= -4
        const uint32_t xa = xy.x & 0xffff, xb = xy.x >> 16, ya = xy.y & 0xffff, yb = xy.y >> 16;
        const uint32_t fp = (uint32_t)g.f2_pitch, ra = ya * (uint32_t)W, rb = yb * (uint32_t)W;
        const bool live = (w.x != 0.f) | (w.y != 0.f) | (w.z != 0.f) | (w.w != 0.f);
        tabw[e] = w;
        taboff[e] = make_uint4((ra + xa) * fp, (ra + xb) * fp, (rb + xa) * fp, (rb + xb) * fp);
        tabdst[e] = (uint32_t)((slot * WX + qx) * pitch) | (live ? 1u : 0u);
      }
      __syncthreads();

      // ---------------- phase 2: image-1 rows of the block + bilinear gather of the new warped rows ----------------
      {
        const int nf1 = R * XS * ncg;
        const __nv_bfloat16* f1n = f1 + (long)n * H * W * g.f1_pitch;
        int pix = tid / ncg, cg = tid - pix * ncg;
        for (int it = tid; it < nf1; it += nthr) {
          const int r = pix / XS, px = pix - r * XS;                // XS is a power of two
          const int y = y0 + r, x = x0 + px, c = cg * 8;
          uint4 v = make_uint4(0u, 0u, 0u, 0u);
          if (y < H && x < W && c < C) v = ld_cg(f1n + (uint32_t)(y * W + x) * (uint32_t)g.f1_pitch + c, C - c);
          *reinterpret_cast<uint4*>(f1s + pix * pitch + cg * 16) = v;
          pix += dq; cg += dr;
          if (cg >= ncg) { cg -= ncg; ++pix; }
        }
      }
      {
        const int ntask = nnew * WX * ncg;
        int pix = tid / ncg, cg = tid - pix * ncg;
        for (int base = tid; base < ntask; base += kU * nthr) {
          uint4 cv[kU][4];
          uint32_t dstu[kU];
          int pixu[kU];
#pragma unroll
          for (int u = 0; u < kU; ++u) {
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);
            cv[u][0] = cv[u][1] = cv[u][2] = cv[u][3] = z;
            pixu[u] = pix;
            dstu[u] = 0xffffffffu;
            if (base + u * nthr < ntask) {
              const uint32_t dst = tabdst[pix];
              dstu[u] = (dst & ~1u) + cg * 16;
              const int c = cg * 8, nv = C - c;
              if ((dst & 1u) && nv > 0) {
                const uint4 o = taboff[pix];
                const __nv_bfloat16* fc = f2n + c;
                if (nv >= 8) {
                  cv[u][0] = *reinterpret_cast<const uint4*>(fc + o.x);
                  cv[u][1] = *reinterpret_cast<const uint4*>(fc + o.y);
                  cv[u][2] = *reinterpret_cast<const uint4*>(fc + o.z);
                  cv[u][3] = *reinterpret_cast<const uint4*>(fc + o.w);
                } else {
                  cv[u][0] = ld_cg(fc + o.x, 4); cv[u][1] = ld_cg(fc + o.y, 4);
                  cv[u][2] = ld_cg(fc + o.z, 4); cv[u][3] = ld_cg(fc + o.w, 4);
                }
              }
            }
            pix += dq; cg += dr;
            if (cg >= ncg) { cg -= ncg; ++pix; }
          }
#pragma unroll
          for (int u = 0; u < kU; ++u) {
            if (dstu[u] != 0xffffffffu) {
              const float4 w = tabw[pixu[u]];
              uint4 o;
              o.x = blend_pair<F16>(w, cv[u][0].x, cv[u][1].x, cv[u][2].x, cv[u][3].x);
              o.y = blend_pair<F16>(w, cv[u][0].y, cv[u][1].y, cv[u][2].y, cv[u][3].y);
              o.z = blend_pair<F16>(w, cv[u][0].z, cv[u][1].z, cv[u][2].z, cv[u][3].z);
              o.w = blend_pair<F16>(w, cv[u][0].w, cv[u][1].w, cv[u][2].w, cv[u][3].w);
              *reinterpret_cast<uint4*>(ring + dstu[u]) = o;
            }
          }
        }
      }
      __syncthreads();

      // ---------------- phase 3: band GEMM per (row, dy) + write-out to the staging tile ----------------
      const int cr = 2 * (lane & 3) - (lane >> 2);                   // (column - row) of accumulator register 0 in its 16x8 tile
      const uint32_t aLane = f1s_addr + (uint32_t)((lane & 15) * pitch + (lane >> 4) * 16);          // ldmatrix row addresses
      const uint32_t bLane = ring_addr + (uint32_t)(((lane & 7) + ((lane >> 4) << 3)) * pitch + ((lane >> 3) & 1) * 16);
      for (int task = warp; task < R * 9; task += nwarps) {
        const int r = task / 9, dy = task - 9 * r;
        const int y = y0 + r;
        if (y >= H) continue;
        const int slot = (y + dy) % RR;                              // warped row y + dy - 4
        const uint32_t aBase = aLane + (uint32_t)(r * XS * pitch);
        const uint32_t bBase = bLane + (uint32_t)(slot * WX * pitch);
        float acc[NMI][3][4];
#pragma unroll
        for (int mi = 0; mi < NMI; ++mi)
#pragma unroll
          for (int tt = 0; tt < 3; ++tt)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[mi][tt][e] = 0.f;
#pragma unroll 2
        for (int k = 0; k < p.nk; ++k) {
          uint32_t a[NMI][4], b[NT + 1][2];
#pragma unroll
          for (int mi = 0; mi < NMI; ++mi) ldsm_x4(a[mi], aBase + (uint32_t)(mi * 16 * pitch + k * 32));
#pragma unroll
          for (int j = 0; j + 1 < NT; j += 2) {
            uint32_t q4[4];
            ldsm_x4(q4, bBase + (uint32_t)(j * 8 * pitch + k * 32));
            b[j][0] = q4[0]; b[j][1] = q4[1]; b[j + 1][0] = q4[2]; b[j + 1][1] = q4[3];
          }
          ldsm_x2(b[NT - 1], bBase + (uint32_t)((NT - 1) * 8 * pitch + k * 32));
#pragma unroll
          for (int mi = 0; mi < NMI; ++mi)
#pragma unroll
            for (int tt = 0; tt < 3; ++tt) mma_h16<F16>(acc[mi][tt], a[mi], b[2 * mi + tt][0], b[2 * mi + tt][1]);
        }
        // accumulator (mi, tt, e): row = 16 mi + lane/4 + 8 (e>>1), column = 8 (2 mi + tt) + 2 (lane%4) + (e&1);
        // j = column - row = 8 (tt - (e>>1)) + (e&1) + cr is the x displacement index when 0 <= j <= 8: per thread the
        // four predicates are lane constants and every staging address is one base + a compile-time offset
        const uint32_t sbase = stg_addr + (uint32_t)(((r * XS + (lane >> 2)) * 80 + dy * 9 + cr) * 2);
        if (dy < 8) {
#pragma unroll
          for (int mi = 0; mi < NMI; ++mi)
#pragma unroll
            for (int tt = 0; tt < 3; ++tt)
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int d8 = tt - (e >> 1);                         // compile time: only 0 and 1 can hold band elements
                if (d8 < 0 || d8 > 1) continue;
                const float q = acc[mi][tt][e] * rc;
                const uint16_t hv = pack1_h16(fmaxf(q, q * g.slope), F16);
                const bool on = d8 == 0 ? (cr >= -(e & 1)) : (cr <= -(e & 1));
                if (on) asm volatile("st.shared.b16 [%0], %1;" :: "r"(sbase + (uint32_t)(((16 * mi + 8 * (e >> 1)) * 80 + 8 * d8 + (e & 1)) * 2)),
                                     "h"(hv) : "memory");
              }
        } else {                                                      // dy = 8: its j = 8 is channel 80 -> side array
#pragma unroll
          for (int mi = 0; mi < NMI; ++mi)
#pragma unroll
            for (int tt = 0; tt < 3; ++tt)
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const int d8 = tt - (e >> 1);
                if (d8 < 0 || d8 > 1) continue;
                const int j = 8 * d8 + (e & 1) + cr;
                if ((unsigned)j <= 8u) {
                  const float q = acc[mi][tt][e] * rc;
                  const uint16_t hv = pack1_h16(fmaxf(q, q * g.slope), F16);
                  const int px = r * XS + 16 * mi + 8 * (e >> 1) + (lane >> 2);
                  if (j < 8) stg[px * 80 + 72 + j] = hv;
                  else c80[px] = hv;
                }
              }
        }
      }
      ptx::fence_proxy_async();                                       // staging writes -> visible to the TMA store
      __syncthreads();
      if (tid < R && y0 + tid < H) {
        ptx::tma_store_4d(&tmOut, stg_addr + (uint32_t)(tid * XS * 160), 0, x0, y0 + tid, n);
        ptx::bulk_commit();
      }
      for (int i = tid; i < R * XS; i += nthr) {
        const int r = i / XS, px = i - r * XS;
        if (y0 + r < H && x0 + px < W) outp[(((long)n * H + y0 + r) * W + x0 + px) * g.out_pitch + 80] = c80[i];
      }
    }
  }
  if (tid < p.R) ptx::bulk_wait0();
}

static int launch_warp_corr_mma(const WarpCorrParams& g, cudaStream_t s) {
  static int attr_dev = -1, num_sms = 0;
  int dev = 0; cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t e = cudaFuncSetAttribute(warp_corr_mma_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(warp_corr_mma_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(warp_corr_mma_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(warp_corr_mma_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return (int)e;
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (num_sms <= 0) num_sms = 148;
    attr_dev = dev;
  }
  WarpCorrMmaParams p;
  p.g = g;
  // 16-pixel strips for 16-pixel-wide levels and when 32-pixel strips cannot give half of the SMs a CTA
  const int nmi = (g.W <= 16 || (long)g.B * cdiv(g.W, 32) * cdiv(g.H, 4) * 2 <= num_sms) ? 1 : 2, XS = 16 * nmi, WX = XS + 8;
  p.nk = cdiv(g.C, 16);
  p.pitch = 32 * p.nk + 16;
  static const int r_env = []() { const char* e = getenv("PWC_CORR_MMA_R"); return e ? atoi(e) : 0; }();
  // rows per block: 4; 2 while blocks of 4 rows leave more than half of the SMs without an item (levels 5 and 6 at batch 8:
  // 12.9 -> 9.9 us and 13.3 -> 10.4 us; level 4 with 112 items is faster at 4: 13.7 vs 17.5 us)
  int R = r_env > 0 ? r_env : ((long)g.B * cdiv(g.W, XS) * cdiv(g.H, 4) * 2 <= num_sms ? 2 : 4);
  int smem = 0;
  for (;; R >>= 1) {
    const int ring = (R + 8) * WX * p.pitch, f1b = R * XS * p.pitch;
    const int stgb = R * XS * 160, tabb = (R + 8) * WX * 36;
    p.off_f1 = ring;
    p.off_stg = (ring + f1b + 127) & ~127;
    p.off_c80 = p.off_stg + (stgb > tabb ? stgb : tabb);
    smem = p.off_c80 + R * XS * 2;
    if (smem <= 227 * 1024) break;
    if (R == 1) return (int)cudaErrorInvalidValue;
  }
  p.R = R;
  const int cps = (smem + 1024) * 2 <= 228 * 1024 ? 2 : 1;          // CTAs per SM
  static const int thr_env = []() { const char* e = getenv("PWC_CORR_MMA_THREADS"); return e ? atoi(e) : 0; }();
  const int threads = thr_env > 0 ? thr_env : (cps == 2 ? 288 : 576);
  p.strips = cdiv(g.W, XS);
  p.blocks = cdiv(g.H, R);
  // row ranges per strip: the split that minimises (items per SM) x (rows gathered + rows multiplied) per item
  const long slots = (long)num_sms * cps;
  long best = -1; int best_k = 1;
  for (int k = 1; k <= p.blocks; ++k) {
    const long items = (long)g.B * p.strips * k;
    const long cost = (long)cdiv((int)items, (int)slots) * (2L * cdiv(p.blocks, k) * R + 8);
    if (best < 0 || cost < best) { best = cost; best_k = k; }
  }
  p.ranges = best_k;
  p.num_items = g.B * p.strips * p.ranges;
  CUtensorMap tmo;                                        // cost-volume channels 0..79 of the output slice
  int e = make_tmap_nhwc(&tmo, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, g.out, g.out_pitch, 80, g.W, g.H, g.B,
                         80, XS, 1, 1, 1, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (e) return e;
  const int grid = (int)(p.num_items < slots ? p.num_items : slots);
  if (g.f16) {
    if (nmi == 1) return (int)launch_pdl(warp_corr_mma_kernel<1, true>, dim3(grid), dim3(threads), smem, s, tmo, p);
    return (int)launch_pdl(warp_corr_mma_kernel<2, true>, dim3(grid), dim3(threads), smem, s, tmo, p);
  }
  if (nmi == 1) return (int)launch_pdl(warp_corr_mma_kernel<1, false>, dim3(grid), dim3(threads), smem, s, tmo, p);
  return (int)launch_pdl(warp_corr_mma_kernel<2, false>, dim3(grid), dim3(threads), smem, s, tmo, p);
}

}  // namespace pwc
