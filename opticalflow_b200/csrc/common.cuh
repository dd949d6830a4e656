// Shared device helpers for libpwc_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#ifndef PWC_ARCH_OK
#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libpwc_b200 is written for sm_100a (B200) only"
#endif
#endif

#include <stdlib.h>
#include <utility>

namespace pwc {

// ---- programmatic dependent launch (PDL, on by default, PWC_PDL=0 turns it off; +1.2% on the batch-8 forward: 3914 vs 3867
// pairs/s on the same box, since the pyramid kernels load their weights before the wait): every kernel of the forward can be launched with the programmatic-serialization
// attribute, so its CTAs may be scheduled while the previous kernel drains; it runs its prologue (barrier init, TMEM
// allocation, tensor-map prefetch, weight loads -- nothing the previous kernel produced) and then blocks in pdl_wait()
// until the previous grid has completed and flushed.  No kernel touches activations before pdl_wait().
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline int pdl_enabled() {
  static const int v = []() { const char* e = getenv("PWC_PDL"); return e ? atoi(e) : 1; }();
  return v;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

__host__ __device__ __forceinline__ int cdiv(int a, int b) { return (a + b - 1) / b; }

__device__ __forceinline__ float lrelu(float v, float slope) { return v > 0.f ? v : v * slope; }

// ---- 4-channel vector load/store with dtype conversion (fp32 math everywhere) ----
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 ld4(const __nv_bfloat16* p) {
  uint2 r = *reinterpret_cast<const uint2*>(p);
  __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&r.x);
  __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&r.y);
  float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
  return make_float4(fa.x, fa.y, fb.x, fb.y);
}
__device__ __forceinline__ float ld1(const float* p) { return *p; }
__device__ __forceinline__ float ld1(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void st1(float* p, float v) { *p = v; }
__device__ __forceinline__ void st1(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

__device__ __forceinline__ float4 ld4(const __half* p) {
  uint2 r = *reinterpret_cast<const uint2*>(p);
  float2 fa = __half22float2(*reinterpret_cast<__half2*>(&r.x)), fb = __half22float2(*reinterpret_cast<__half2*>(&r.y));
  return make_float4(fa.x, fa.y, fb.x, fb.y);
}
__device__ __forceinline__ float ld1(const __half* p) { return __half2float(*p); }
__device__ __forceinline__ void st1(__half* p, float v) { *p = __float2half_rn(v); }

// ---- the 16-bit storage format of the tensor-core conv modes is a run-time choice: bf16 (conv_mode 'bf16', f16 == 0) or
// IEEE half (conv_mode 'f16': 3 more mantissa bits, ~7x lower end-to-end flow error at the same speed).
// tcgen05 kind::f16 and mma.sync take either; these helpers are the only places where the two differ.
__device__ __forceinline__ uint32_t pack2_h16(float a, float b, int f16) {
  if (f16) { __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint16_t pack1_h16(float a, int f16) {
  if (f16) { __half h = __float2half_rn(a); return *reinterpret_cast<uint16_t*>(&h); }
  __nv_bfloat16 h = __float2bfloat16_rn(a);
  return *reinterpret_cast<uint16_t*>(&h);
}
__device__ __forceinline__ float2 unpack2_h16(uint32_t v, int f16) {
  if (f16) return __half22float2(*reinterpret_cast<__half2*>(&v));
  return make_float2(__uint_as_float(v << 16), __uint_as_float(v & 0xffff0000u));
}
__device__ __forceinline__ float unpack1_h16(uint16_t v, int f16) {
  if (f16) return __half2float(*reinterpret_cast<__half*>(&v));
  return __uint_as_float((uint32_t)v << 16);
}
template <bool F16>
__device__ __forceinline__ void mma_h16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if (F16)
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  else
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// host side: the format the NEXT create / launch call of this thread uses (pwc_set_half_format, include/pwc_b200.h)
int half_format();

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

}  // namespace pwc
