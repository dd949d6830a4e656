// Fused warp + cost volume on the tensor cores (bf16 features, C <= 64): the 81 displacement dot products of a pixel
// are the near-diagonal band of a GEMM between image-1 rows and WARPED image-2 rows,
//     D_dy[(r,x), (r,x')] = sum_c f1[y0+r][x0+x][c] * w2[y0+r+dy][x0-4+x'][c],   band: x' - x = dx+4 in 0..8.
// Tile = 4 rows x 32 columns of output pixels = the M = 128 operand (one TMA box, SWIZZLE_128B, zero fill = padding);
// for each of the 9 row displacements the N = 160 operand is 4 warped rows x 40 columns, a window of a shared-memory
// RING of warped rows that the producer warps fill by bilinear gather (mask folded into the corner weights) -- every
// warped row is gathered once per (32-column strip, row range) and reused by the 9 tiles that need it; warped features
// never exist in HBM.  Only 9 of 160 columns per row are wanted (5.6% MMA utilisation), which still costs less tensor
// time than the HBM traffic of the tile; what the tensor cores remove is the 81*C FMA/pixel of the CUDA-core kernel.
//   warp 0      : TMA loader of the image-1 tiles (2 stages)
//   warp 1      : MMA issuer; accumulators triple buffered in TMEM (3 x 160 columns), one buffer per dy
//   warps 2..5  : epilogue.  Warp q owns tile row q (TMEM lanes 32q..32q+31) and its 40-column window; the band is
//                 extracted by a skewed round trip through a per-warp scratch (thread L writes its 40 values, reads back
//                 [L+j], j = 0..8: bank-conflict free both ways), then /C, LeakyReLU, bf16 and staged [pixel][81];
//                 after the 9th dy the warp writes 162-byte runs coalesced into the concat slice.
//   warps 6..13 : producers (gather).  Ring = 4 groups of 4 warped rows (+3 mirror rows so a 4-row window never wraps).
// Included by warp_corr.cu (shares WarpCorrParams / warp_entry / div_by).
#pragma once

namespace pwc {

constexpr int kTcR = 4;                    // rows per tile
constexpr int kTcBW = 40;                  // warped pixels per ring row (32 + 2*4)
constexpr int kTcN = kTcR * kTcBW;         // 160 warped pixels per ring group
constexpr int kTcPassRows = 6;             // one MMA pass = a 6-row window = 3 row displacements for all 4 tile rows
constexpr int kTcNP = kTcPassRows * kTcBW; // 240 = MMA N
constexpr int kTcMirror = 5;               // ring rows 0..4 repeated behind row 15 so a 6-row window never wraps
constexpr int kTcRingRows = 16;
constexpr int kTcRowBytes = kTcBW * 128;   // 5120: one ring row, 128-byte (64 x bf16) pixel rows
constexpr int kTcRingBytes = (kTcRingRows + kTcMirror) * kTcRowBytes;
constexpr int kTcABytes = 128 * 128;
constexpr int kTcScrBytes = 0;
constexpr int kTcStgStride = 80;           // bf16 per staged pixel: dense 160-byte rows = the TMA store box
constexpr int kTcStgBytes = 4 * 32 * kTcStgStride * 2;
constexpr int kTcTabBytes = 2 * kTcN * 24;          // one coordinate/weight table per producer team
constexpr int kTcProdWarps = 8;
constexpr int kTcThreads = (6 + kTcProdWarps) * 32;
constexpr int kTcBarBytes = 8 * (2 + 2 + 4 + 4 + 3 + 3) + 16;
constexpr int kTcSmemBytes = kTcRingBytes + 2 * kTcABytes + kTcScrBytes + kTcStgBytes + kTcTabBytes + kTcBarBytes + 1024;

struct WarpCorrTcParams {
  WarpCorrParams g;
  int strips, ranges, tile_rows, num_items;     // tile_rows = ceil(H/4); an item = (image, strip, row range)
  int nk;                                       // K = 16 steps holding real channels (C/16)
  int dbg;                                      // PWC_CORR_TRACE=1: CTA 0 records clock64() marks (tools/trace_corr.py)
};

// debug timeline of CTA 0: [role 0 producer | 1 mma | 2 epilogue][index][4 marks]
__device__ unsigned long long g_corr_trace[3 * 64 * 4];
__device__ __forceinline__ void tc_mark(bool on, int role, int idx, int k) {
  if (on && idx < 64) g_corr_trace[(role * 64 + idx) * 4 + k] = clock64();
}

// item -> image n, strip x0, first tile row, tile count
__device__ __forceinline__ void tc_decode(const WarpCorrTcParams& p, int item, int& n, int& x0, int& ty0, int& T) {
  const int per_img = p.strips * p.ranges;
  n = item / per_img;
  const int r = item - n * per_img;
  const int rg = r / p.strips;
  x0 = (r - rg * p.strips) * 32;
  const int q = p.tile_rows / p.ranges, rem = p.tile_rows - q * p.ranges;
  ty0 = rg * q + min(rg, rem);
  T = q + (rg < rem ? 1 : 0);
}

// LeakyReLU (0 <= slope <= 1) of two packed 16-bit values: max(x, x*s1 + x*s2) with s1 + s2 = slope to ~2^-20
template <bool F16>
__device__ __forceinline__ uint32_t tc_lrelu2(uint32_t x, uint32_t s1, uint32_t s2) {
  if (F16) {
    const __half2 v = *reinterpret_cast<const __half2*>(&x);
    const __half2 y = __hmax2(v, __hfma2(v, *reinterpret_cast<const __half2*>(&s1), __hmul2(v, *reinterpret_cast<const __half2*>(&s2))));
    return *reinterpret_cast<const uint32_t*>(&y);
  }
  const __nv_bfloat162 v = *reinterpret_cast<const __nv_bfloat162*>(&x);
  const __nv_bfloat162 y = __hmax2(v, __hfma2(v, *reinterpret_cast<const __nv_bfloat162*>(&s1),
                                            __hmul2(v, *reinterpret_cast<const __nv_bfloat162*>(&s2))));
  return *reinterpret_cast<const uint32_t*>(&y);
}

template <bool F16>
__global__ void __launch_bounds__(kTcThreads, 1)
warp_corr_tc_kernel(const __grid_constant__ CUtensorMap tmF1, const __grid_constant__ CUtensorMap tmOut,
                    const WarpCorrTcParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t addr0 = ptx::smem_addr(smem_raw);
  const uint32_t base = (addr0 + 1023u) & ~1023u;
  uint8_t* sm = smem_raw + (base - addr0);
  const uint32_t ring_base = base;
  const uint32_t a_base = ring_base + kTcRingBytes;
  uint8_t* ring = sm;
  uint16_t* stg_all = reinterpret_cast<uint16_t*>(sm + kTcRingBytes + 2 * kTcABytes + kTcScrBytes);
  float4* tabw = reinterpret_cast<float4*>(sm + kTcRingBytes + 2 * kTcABytes + kTcScrBytes + kTcStgBytes);
  uint2* tabxy = reinterpret_cast<uint2*>(tabw + 2 * kTcN);
  const uint32_t bar0 = a_base + 2 * kTcABytes + kTcScrBytes + kTcStgBytes + kTcTabBytes;
  auto a_full = [&](int i) { return bar0 + 8u * i; };
  auto a_empty = [&](int i) { return bar0 + 8u * (2 + i); };
  auto b_full = [&](int i) { return bar0 + 8u * (4 + i); };
  auto b_empty = [&](int i) { return bar0 + 8u * (8 + i); };
  auto d_full = [&](int i) { return bar0 + 8u * (12 + i); };
  auto d_empty = [&](int i) { return bar0 + 8u * (15 + i); };
  const uint32_t tmem_slot = bar0 + 8u * 18;
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(sm + (tmem_slot - base));

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const WarpCorrParams& g = p.g;
  const int C = g.C, H = g.H, W = g.W;

  if (tid == 0) {
    ptx::prefetch_tmap(&tmF1);
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(a_full(i), 1); ptx::mbar_init(a_empty(i), 1); }
    for (int i = 0; i < 4; ++i) { ptx::mbar_init(b_full(i), kTcProdWarps / 2); ptx::mbar_init(b_empty(i), 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(d_full(i), 1); ptx::mbar_init(d_empty(i), 4); }
    ptx::fence_barrier_init();
    ptx::fence_proxy_async();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, 512);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  pdl_wait();                     // features / flow come from the previous kernels (PDL: see common.cuh)
  pdl_launch_dependents();

  if (warp == 0) {
    // ============================== image-1 tile loader ==============================
    int as = 0; uint32_t aph = 0;
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
      int n, x0, ty0, T;
      tc_decode(p, item, n, x0, ty0, T);
      for (int t = 0; t < T; ++t) {
        ptx::mbar_wait_sleep(a_empty(as), aph ^ 1, 200);
        if (ptx::elect_one()) {
          ptx::mbar_expect_tx(a_full(as), kTcABytes);
          ptx::tma_load_4d(a_base + as * kTcABytes, &tmF1, a_full(as), 0, x0, (ty0 + t) * kTcR, n);
        }
        __syncwarp();
        if (++as == 2) { as = 0; aph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer ==============================
    const uint32_t idesc = F16 ? (ptx::make_idesc_bf16(128, kTcNP) & ~((7u << 7) | (7u << 10))) : ptx::make_idesc_bf16(128, kTcNP);
    int as = 0; uint32_t aph = 0;
    int db = 0; uint32_t dph = 0;
    int gcount = 0, waited = -1;           // global group counter / highest group whose b_full has been observed
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
      int n, x0, ty0, T;
      tc_decode(p, item, n, x0, ty0, T);
      for (int t = 0; t < T; ++t) {
        const int G = gcount + t;
        const bool tr = p.dbg && blockIdx.x == 0 && lane == 0;
        tc_mark(tr, 1, G, 0);
        ptx::mbar_wait(a_full(as), aph);
        tc_mark(tr, 1, G, 1);
        const uint64_t da = ptx::make_kmajor_desc(a_base + as * kTcABytes, 1024, 2u);
        for (int ps = 0; ps < 3; ++ps) {
          const int dyb = 3 * ps;                        // window rows 4G+dyb .. 4G+dyb+5: displacements dyb..dyb+2
          const int need = G + (ps == 0 ? 1 : 2);
          while (waited < need) { ++waited; ptx::mbar_wait(b_full(waited & 3), (uint32_t)(waited >> 2) & 1u); }
          const bool tr2 = tr && t == 2;
          tc_mark(tr2, 1, 20 + ps, 0);
          ptx::mbar_wait(d_empty(db), dph ^ 1);
          ptx::tc_fence_after();
          tc_mark(tr2, 1, 20 + ps, 1);
          if (ps == 0) tc_mark(tr, 1, G, 2);
          if (ps == 2) tc_mark(tr, 1, G, 3);
          if (ptx::elect_one()) {
            const uint32_t srow = (uint32_t)((4 * G + dyb) & 15);
            const uint64_t dbd = ptx::make_kmajor_desc(ring_base + srow * kTcRowBytes, 1024, 2u);
            const uint32_t d_tmem = tmem_base + (uint32_t)(db * 256);
            ptx::umma_bf16(d_tmem, da, dbd, idesc, 0u);
            if (p.nk > 1) ptx::umma_bf16(d_tmem, da + 2, dbd + 2, idesc, 1u);
            if (p.nk > 2) ptx::umma_bf16(d_tmem, da + 4, dbd + 4, idesc, 1u);
            if (p.nk > 3) ptx::umma_bf16(d_tmem, da + 6, dbd + 6, idesc, 1u);
            ptx::umma_commit(d_full(db));
            if (ps == 2) {
              ptx::umma_commit(a_empty(as));
              ptx::umma_commit(b_empty(G & 3));
              if (t == T - 1) { ptx::umma_commit(b_empty((G + 1) & 3)); ptx::umma_commit(b_empty((G + 2) & 3)); }
            }
          }
          __syncwarp();
          tc_mark(tr2, 1, 20 + ps, 2);
          if (++db == 2) { db = 0; dph ^= 1; }
        }
        if (++as == 2) { as = 0; aph ^= 1; }
      }
      gcount += T + 2;
    }
  } else if (warp < 6) {
    // ============================== epilogue ==============================
    const int q = warp & 3;                               // TMEM lane quarter = tile row
    uint16_t* stg = stg_all + q * 32 * kTcStgStride;
    const uint32_t stg_addr = a_base + 2 * kTcABytes + kTcScrBytes + (uint32_t)(q * 32 * kTcStgStride * 2);
    const bool b16 = lane & 16, b8 = lane & 8, b4 = lane & 4, b2 = lane & 2, b1 = lane & 1;
    // byte-permute selectors of the last shifter level (lane parity): even / odd channel position, carry, merge with the carry
    const uint32_t selE = b1 ? 0x5432u : 0x3210u, selO = b1 ? 0x7654u : 0x5432u;
    const uint32_t selC = b1 ? 0x4432u : 0x4410u, selM = b1 ? 0x7610u : 0x5410u;
    // LeakyReLU slope as hi + lo of the storage format, both halves of a packed register
    const uint32_t sl1 = pack2_h16(g.slope, g.slope, F16);
    const float sl1f = unpack2_h16(sl1, F16).x;
    const uint32_t sl2 = pack2_h16(g.slope - sl1f, g.slope - sl1f, F16);
    int db = 0; uint32_t dph = 0;
    int tcount = 0;
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
      int n, x0, ty0, T;
      tc_decode(p, item, n, x0, ty0, T);
      for (int t = 0; t < T; ++t) {
        const int y = (ty0 + t) * kTcR + q;
        const bool tr = p.dbg && blockIdx.x == 0 && warp == 2 && lane == 0;
        const bool tr2 = tr && tcount == 2;
        // The epilogue works on PACKED 16-bit pairs: the 40 accumulator columns of a (row, dy) are rounded to the storage
        // format first (the 1/C scale is folded into the gather weights, exact for C = 16 / 32 / 64), then the band [lane+j],
        // j = 0..8, is extracted by a 4-level barrel shifter on 20 packed registers (31 selects) plus one byte permute per
        // output word (the lane's parity and the parity of the channel position 9*dy pick the halves) -- 36 instead of 71
        // selects per 9 outputs -- and LeakyReLU runs on packed words (HMUL2 + HFMA2 + HMNMX2 per two channels, the slope
        // as a two-term sum so that the product is correctly rounded).  ow[] = the pixel's 80 channels, 40 words.
        uint32_t ow[40];
        uint32_t carry = 0;                               // a 9-element group ends on an odd position every other dy
        uint16_t ch80 = 0;
        uint16_t* srow = stg + lane * kTcStgStride;       // this lane's pixel: [80] halfwords, 160-byte rows
#pragma unroll
        for (int ps = 0; ps < 3; ++ps) {
          if (ps == 0) tc_mark(tr, 2, tcount, 0);
          ptx::mbar_wait(d_full(db), dph);
          ptx::tc_fence_after();
          if (ps == 0) tc_mark(tr, 2, tcount, 1);
          if (ps == 2) tc_mark(tr, 2, tcount, 2);
#pragma unroll
          for (int d = 0; d < 3; ++d) {
            tc_mark(tr2, 2, 20 + 3 * ps + d, 0);
            // tile row q at displacement dy = 3*ps + d reads window row q + d, columns lane .. lane+8 of its 40
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(db * 256 + (q + d) * kTcBW);
            uint32_t P[20];
            {
              uint32_t v0[16], v1[16], v2[8];
              ptx::tmem_ld16(taddr, v0);
              ptx::tmem_ld16(taddr + 16, v1);
              ptx::tmem_ld8(taddr + 32, v2);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                P[i] = pack2_h16(__uint_as_float(v0[2 * i]), __uint_as_float(v0[2 * i + 1]), F16);
                P[8 + i] = pack2_h16(__uint_as_float(v1[2 * i]), __uint_as_float(v1[2 * i + 1]), F16);
              }
#pragma unroll
              for (int i = 0; i < 4; ++i) P[16 + i] = pack2_h16(__uint_as_float(v2[2 * i]), __uint_as_float(v2[2 * i + 1]), F16);
            }
            tc_mark(tr2, 2, 20 + 3 * ps + d, 1);
            if (d == 2) {                                 // accumulator buffer fully read: hand it back to the MMA warp
              ptx::tc_fence_before();
              __syncwarp();
              if (lane == 0) ptx::mbar_arrive(d_empty(db));
            }
            uint32_t s1[12], s2[8], s3[6], s4[5];
#pragma unroll
            for (int i = 0; i < 12; ++i) s1[i] = b16 ? P[i + 8] : P[i];
#pragma unroll
            for (int i = 0; i < 8; ++i) s2[i] = b8 ? s1[i + 4] : s1[i];
#pragma unroll
            for (int i = 0; i < 6; ++i) s3[i] = b4 ? s2[i + 2] : s2[i];
#pragma unroll
            for (int i = 0; i < 5; ++i) s4[i] = b2 ? s3[i + 1] : s3[i];
            const int dy = 3 * ps + d, pos = 9 * dy;      // compile time
            if ((pos & 1) == 0) {
#pragma unroll
              for (int k = 0; k < 4; ++k) ow[pos / 2 + k] = tc_lrelu2<F16>(__byte_perm(s4[k], s4[k + 1], selE), sl1, sl2);
              carry = __byte_perm(s4[4], 0u, selC);       // element 8 in the low half
            } else {
              ow[(pos - 1) / 2] = tc_lrelu2<F16>(__byte_perm(carry, s4[0], selM), sl1, sl2);
#pragma unroll
              for (int k = 1; k < 5; ++k) ow[(pos - 1) / 2 + k] = tc_lrelu2<F16>(__byte_perm(s4[k - 1], s4[k], selO), sl1, sl2);
            }
            tc_mark(tr2, 2, 20 + 3 * ps + d, 3);
          }
          if (++db == 2) { db = 0; dph ^= 1; }
        }
        ch80 = (uint16_t)(tc_lrelu2<F16>(carry, sl1, sl2) & 0xffffu);      // channel 80 goes out from the register
        // the previous tile's TMA store has read the staging rows: now this lane's row may be rewritten (conflict-free STS.128)
        if (lane == 0) ptx::bulk_wait_read0();
        __syncwarp();
#pragma unroll
        for (int k = 0; k < 10; ++k)
          *reinterpret_cast<uint4*>(srow + 8 * k) = make_uint4(ow[4 * k], ow[4 * k + 1], ow[4 * k + 2], ow[4 * k + 3]);
        // ---- write-out: the warp's 32 pixels x 80 channels leave as ONE TMA tensor store (box 160 B x 32 px, clipped
        //      at the image edge) into the concat slice; channel 80 is stored from the register ----
        tc_mark(tr2, 2, 40, 0);
        ptx::fence_proxy_async();                         // STS.128 of every lane -> visible to the async proxy
        __syncwarp();
        if (ptx::elect_one()) {
          ptx::tma_store_4d(&tmOut, stg_addr, 0, x0, y, n);
          ptx::bulk_commit();
        }
        __syncwarp();
        if (y < H && x0 + lane < W)
          reinterpret_cast<uint16_t*>(g.out)[(((long)n * H + y) * W + x0 + lane) * g.out_pitch + 80] = ch80;
        tc_mark(tr2, 2, 40, 1);
        tc_mark(p.dbg && blockIdx.x == 0 && warp == 2 && lane == 0, 2, tcount, 3);
        ++tcount;
      }
    }
    ptx::bulk_wait0();
  } else {
    // ============================== producers: warped image-2 rows into the ring ==============================
    // two teams of 4 warps take alternate groups, so the latency chain of one group (flow -> table -> corner loads
    // -> blend -> stores) overlaps the other team's
    const int team = (warp - 6) >> 2;
    const int ptid = tid - 6 * 32 - team * 128;
    constexpr int kP = 128;
    constexpr int kBatch = 5;                             // tasks per thread per batch (C = 32: one batch per group)
    float4* tw = tabw + team * kTcN;
    uint2* txy = tabxy + team * kTcN;
    const __nv_bfloat16* f2 = reinterpret_cast<const __nv_bfloat16*>(g.f2);
    const int sh = (C == 64) ? 3 : (C == 32) ? 2 : 1;     // 16-byte channel groups per pixel = 1 << sh
    const float rcC = 1.0f / (float)C;
    const int ntask = kTcN << sh;
    int gcount = 0;
    for (int item = blockIdx.x; item < p.num_items; item += gridDim.x) {
      int n, x0, ty0, T;
      tc_decode(p, item, n, x0, ty0, T);
      const __nv_bfloat16* f2n = f2 + (long)n * H * W * g.f2_pitch;
      for (int gi = 0; gi < T + 2; ++gi) {
        const int G = gcount + gi, slot = G & 3;
        if ((G & 1) != team) continue;
        const int yb = ty0 * kTcR - 4 + 4 * gi;            // first warped row of the group
        const bool tr = p.dbg && blockIdx.x == 0 && ptid == 0;
        tc_mark(tr, 0, G, 0);
        named_bar_sync(1 + team, kP);                      // previous group's table fully consumed
        for (int e = ptid; e < kTcN; e += kP) {
          const int i = e / kTcBW, qx = e - i * kTcBW;
          const int yy = yb + i, xx = x0 - 4 + qx;
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
          // the cost volume's 1/C (correlation_cuda_kernel.cu:143) rides on the warped operand: C is 16, 32 or 64 here, so the
          // scale is a power of two and commutes with every rounding of the blend and of the dot product
          w.x *= rcC; w.y *= rcC; w.z *= rcC; w.w *= rcC;
          tw[e] = w; txy[e] = xy;
        }
        named_bar_sync(1 + team, kP);
        tc_mark(tr, 0, G, 1);
        for (int b0 = 0; b0 < ntask; b0 += kBatch * kP) {
          // corner loads are issued BEFORE waiting for the ring slot: only the stores need it
          uint4 cv[kBatch][4];
#pragma unroll
          for (int u = 0; u < kBatch; ++u) {
            const int task = b0 + ptid + u * kP;
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);
            cv[u][0] = cv[u][1] = cv[u][2] = cv[u][3] = z;
            if (task < ntask) {
              const int pix = task >> sh, cg = task & ((1 << sh) - 1);
              const float4 w = tw[pix];
              const uint2 xy = txy[pix];
              const int xa = xy.x & 0xffff, xb = xy.x >> 16, ya = xy.y & 0xffff, yb2 = xy.y >> 16;
              const __nv_bfloat16* fc = f2n + cg * 8;
              if (w.x != 0.f) cv[u][0] = *reinterpret_cast<const uint4*>(fc + ((long)ya * W + xa) * g.f2_pitch);
              if (w.y != 0.f) cv[u][1] = *reinterpret_cast<const uint4*>(fc + ((long)ya * W + xb) * g.f2_pitch);
              if (w.z != 0.f) cv[u][2] = *reinterpret_cast<const uint4*>(fc + ((long)yb2 * W + xa) * g.f2_pitch);
              if (w.w != 0.f) cv[u][3] = *reinterpret_cast<const uint4*>(fc + ((long)yb2 * W + xb) * g.f2_pitch);
            }
          }
          if (b0 == 0) {
            ptx::mbar_wait_sleep(b_empty(slot), ((uint32_t)(G >> 2) & 1u) ^ 1u, 100);
            tc_mark(tr, 0, G, 2);
          }
#pragma unroll
          for (int u = 0; u < kBatch; ++u) {
            const int task = b0 + ptid + u * kP;
            if (task < ntask) {
              const int pix = task >> sh, cg = task & ((1 << sh) - 1);
              const float4 w = tw[pix];
              const uint32_t a[4] = {cv[u][0].x, cv[u][0].y, cv[u][0].z, cv[u][0].w};
              const uint32_t b[4] = {cv[u][1].x, cv[u][1].y, cv[u][1].z, cv[u][1].w};
              const uint32_t c[4] = {cv[u][2].x, cv[u][2].y, cv[u][2].z, cv[u][2].w};
              const uint32_t d[4] = {cv[u][3].x, cv[u][3].y, cv[u][3].z, cv[u][3].w};
              uint32_t o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                // bf16 pair -> two fp32 (low half = even channel); same FMA order as the fp32 kernel's gather
                float lo = 0.f, hi = 0.f;
                const float2 fa = unpack2_h16(a[e], F16), fb = unpack2_h16(b[e], F16), fc2 = unpack2_h16(c[e], F16),
                             fd = unpack2_h16(d[e], F16);
                lo = fmaf(w.x, fa.x, lo); hi = fmaf(w.x, fa.y, hi);
                lo = fmaf(w.y, fb.x, lo); hi = fmaf(w.y, fb.y, hi);
                lo = fmaf(w.z, fc2.x, lo); hi = fmaf(w.z, fc2.y, hi);
                lo = fmaf(w.w, fd.x, lo); hi = fmaf(w.w, fd.y, hi);
                o[e] = pack2_h16(lo, hi, F16);
              }
              const int i = pix / kTcBW, qx = pix - i * kTcBW;
              const uint32_t off = (uint32_t)((slot * 4 + i) * kTcRowBytes + qx * 128 + ((cg ^ (qx & 7)) << 4));
              *reinterpret_cast<uint4*>(ring + off) = make_uint4(o[0], o[1], o[2], o[3]);
              if (slot == 0 || (slot == 1 && i == 0))        // mirror of ring rows 0..4 behind row 15
                *reinterpret_cast<uint4*>(ring + off + kTcRingRows * kTcRowBytes) = make_uint4(o[0], o[1], o[2], o[3]);
            }
          }
        }
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(b_full(slot));
        tc_mark(tr, 0, G, 3);
      }
      gcount += T + 2;
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, 512);
}

static int launch_warp_corr_tc(const WarpCorrParams& g, cudaStream_t s) {
  static int attr_dev = -1, num_sms = 0;
  int dev = 0; cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t e = cudaFuncSetAttribute(warp_corr_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(warp_corr_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes);
    if (e != cudaSuccess) return (int)e;
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (num_sms <= 0) num_sms = 148;
    attr_dev = dev;
  }
  WarpCorrTcParams p;
  p.g = g;
  p.strips = cdiv(g.W, 32);
  p.tile_rows = cdiv(g.H, kTcR);
  p.nk = g.C / 16;
  { const char* e = getenv("PWC_CORR_TRACE"); p.dbg = e ? atoi(e) : 0; }
  // row ranges per strip: fewest waves x (tiles + 2 warm-up groups) per item
  long best = -1; int best_k = 1;
  for (int k = 1; k <= p.tile_rows; ++k) {
    const long items = (long)g.B * p.strips * k;
    const long cost = (long)cdiv((int)items, num_sms) * (cdiv(p.tile_rows, k) + 2);
    if (best < 0 || cost < best) { best = cost; best_k = k; }
  }
  p.ranges = best_k;
  p.num_items = g.B * p.strips * p.ranges;
  CUtensorMap tm;
  int e = make_tmap_nhwc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, g.f1, g.f1_pitch, g.C, g.W, g.H, g.B,
                         64, 32, kTcR, 1, 1, CU_TENSOR_MAP_SWIZZLE_128B);
  if (e) return e;
  CUtensorMap tmo;                                      // cost-volume channels 0..79 of the output slice
  e = make_tmap_nhwc(&tmo, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, g.out, g.out_pitch, 80, g.W, g.H, g.B,
                     80, 32, 1, 1, 1, CU_TENSOR_MAP_SWIZZLE_NONE);
  if (e) return e;
  const int grid = p.num_items < num_sms ? p.num_items : num_sms;
  if (g.f16) return (int)launch_pdl(warp_corr_tc_kernel<true>, dim3(grid), dim3(kTcThreads), kTcSmemBytes, s, tm, tmo, p);
  return (int)launch_pdl(warp_corr_tc_kernel<false>, dim3(grid), dim3(kTcThreads), kTcSmemBytes, s, tm, tmo, p);
}

}  // namespace pwc
