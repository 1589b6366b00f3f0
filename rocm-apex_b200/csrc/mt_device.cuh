// mt_device.cuh -- device side of the multi-tensor launch: descriptor-table view, tile walk,
// vectorised streaming loop and deterministic CTA reductions.  sm_100a only.
//
// Replaces (and is deliberately not shaped like) the reference harness
// csrc/multi_tensor_apply.cuh:19-42: there, <=110 tensor addresses and <=320 (tensor, chunk)
// records ride in a ~3 KB by-value kernel argument and a list needs many launches.  Here the
// tensor list lives in a device-resident table (uploaded once, reused while pointers are
// unchanged), ONE launch covers any number of tensors, and every CTA finds its tile with a
// binary search over a per-tensor tile prefix.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <type_traits>
#include <tuple>

namespace b200mt {

// ------------------------------------------------------------------------------------------
// Descriptor table as the kernels see it (all pointers are device pointers into one blob).
// ------------------------------------------------------------------------------------------
struct TableView {
  const unsigned long long* ptrs;   // [depth][ntensors] base addresses, list-major
  const long long*          numel;  // [ntensors]
  const int*                tile_prefix;  // [ntensors+1] exclusive prefix of tiles per tensor
  int       ntensors;
  int       ntiles;
  int       nonempty;     // tensors with at least one element
  int       multichunk;   // tensors longer than one chunk
  int       tile_elems;   // elements per scheduling tile
  int       ppc;          // tiles per full chunk (1 for element-wise ops)
  long long chunk_elems;  // chunk length in elements (== tile_elems when ppc == 1)
};

struct TileLoc {
  int       tensor;
  int       chunk;   // chunk index inside the tensor (== tile index inside tensor when ppc==1)
  long long off;     // element offset of the tile inside the tensor
  int       len;     // elements in the tile
};

#ifndef B200MT_WARP_SEARCH
#define B200MT_WARP_SEARCH 1   // 0: plain binary search (log2 T dependent loads), 1: 32-way warp search (log33 T rounds)
#endif

// Last tensor whose first tile is <= tile (upper_bound(tile_prefix, tile) - 1).  Empty tensors have zero tiles
// and can never be selected.  Must be called by all 32 lanes of a warp with the same arguments.
//
// The look-up is the prologue of every CTA: nothing of the tile can be requested before it is known, and a binary
// search is a chain of log2(T) DEPENDENT loads (8 for the ResNet-50 list, 14 for 10,000 tensors), each an L1/L2 round
// trip during which the CTA's slot moves no data.  Here the 32 lanes probe 32 evenly spaced entries at once and a
// ballot picks the sub-range: ceil(log33 T) rounds (2 up to 1,089 tensors, 3 up to 35,937).
__device__ __forceinline__ int find_tensor(const int* __restrict__ prefix, int ntensors, int tile) {
  int lo = 0, hi = ntensors;  // invariant: prefix[lo] <= tile < prefix[hi]
#if B200MT_WARP_SEARCH
  const int lane = threadIdx.x & 31;
  while (hi - lo > 1) {
    const long long span = hi - lo;
    const int idx = lo + (int)(((lane + 1) * span) / 33);          // in [lo, hi - 1], non-decreasing in lane
    const bool le = __ldg(prefix + idx) <= tile;                    // true for a leading run of lanes (prefix is sorted)
    const int c = __popc(__ballot_sync(0xffffffffu, le));
    const int nlo = c ? lo + (int)((c * span) / 33) : lo;           // probe of the last lane that said "<="
    const int nhi = (c < 32) ? lo + (int)(((c + 1) * span) / 33) : hi;   // probe of the first lane that said ">"
    lo = nlo; hi = nhi;
  }
#else
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (__ldg(prefix + mid) <= tile) lo = mid; else hi = mid;
  }
#endif
  return lo;
}

__device__ __forceinline__ TileLoc locate_tile(const TableView& tv, int tile) {
  const int lo = find_tensor(tv.tile_prefix, tv.ntensors, tile);
  TileLoc L;
  L.tensor = lo;
  const int local = tile - __ldg(tv.tile_prefix + lo);
  const long long n = __ldg(tv.numel + lo);
  const int chunk = local / tv.ppc;
  const int piece = local - chunk * tv.ppc;
  L.chunk = chunk;
  L.off = (long long)chunk * tv.chunk_elems + (long long)piece * tv.tile_elems;
  long long chunk_end = (long long)(chunk + 1) * tv.chunk_elems;
  if (chunk_end > n) chunk_end = n;
  long long len = chunk_end - L.off;
  if (len > tv.tile_elems) len = tv.tile_elems;
  L.len = (int)len;
  return L;
}

// ------------------------------------------------------------------------------------------
// Element types: 4-element ("quad") vector access + scalar access, everything widened to fp32.
// Streaming cache policy: loads and stores do not allocate in L1 (no reuse on this path).
// ------------------------------------------------------------------------------------------
#ifndef B200MT_CACHE_POLICY
#define B200MT_CACHE_POLICY 1   // 0: default caching, 1: L1::no_allocate, 2: .cs (evict-first)
#endif

__device__ __forceinline__ float4 ld_f4(const float4* p) {
  float4 r;
#if B200MT_CACHE_POLICY == 1
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
#elif B200MT_CACHE_POLICY == 2
  r = __ldcs(p);
#else
  r = *p;
#endif
  return r;
}
__device__ __forceinline__ void st_f4(float4* p, float4 v) {
#if B200MT_CACHE_POLICY == 1
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
#elif B200MT_CACHE_POLICY == 2
  __stcs(p, v);
#else
  *p = v;
#endif
}
__device__ __forceinline__ uint2 ld_u2(const uint2* p) {
  uint2 r;
#if B200MT_CACHE_POLICY == 1
  asm volatile("ld.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
               : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
#elif B200MT_CACHE_POLICY == 2
  r = __ldcs(p);
#else
  r = *p;
#endif
  return r;
}
__device__ __forceinline__ void st_u2(uint2* p, uint2 v) {
#if B200MT_CACHE_POLICY == 1
  asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};"
               :: "l"(p), "r"(v.x), "r"(v.y) : "memory");
#elif B200MT_CACHE_POLICY == 2
  __stcs(p, v);
#else
  *p = v;
#endif
}

// 256-bit global accesses (sm_100+: LDG.E.256 / STG.E.256), 32-byte aligned
__device__ __forceinline__ void ld_f8(const void* p, float (&r)[8]) {
  asm volatile("ld.global.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7])
               : "l"(p) : "memory");
}
__device__ __forceinline__ void st_f8(void* p, const float (&r)[8]) {
  asm volatile("st.global.L1::no_allocate.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "l"(p), "f"(r[0]), "f"(r[1]), "f"(r[2]), "f"(r[3]), "f"(r[4]), "f"(r[5]), "f"(r[6]), "f"(r[7])
               : "memory");
}
// ... with an L2 eviction priority: POL 1 = evict_first for data this kernel is the last to touch for a while, 2 = evict_last
// for data the next kernel of the stream reads back.  PTX takes the qualifier itself on 256-bit accesses only; 128-bit
// accesses carry it as a cache-hint policy operand (ends up in the access descriptor, two uniform registers).
template <int POL> __device__ __forceinline__ unsigned long long l2_policy_word() {
  unsigned long long pol;
  if constexpr (POL == 1) asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  else asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
template <int POL> __device__ __forceinline__ void ld_f4_l2(const void* p, float (&r)[4]) {
  asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
               : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]) : "l"(p), "l"(l2_policy_word<POL>()) : "memory");
}
template <int POL> __device__ __forceinline__ void st_f4_l2(void* p, const float (&r)[4]) {
  asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;"
               :: "l"(p), "f"(r[0]), "f"(r[1]), "f"(r[2]), "f"(r[3]), "l"(l2_policy_word<POL>()) : "memory");
}
template <int POL> __device__ __forceinline__ void ld_f8_l2(const void* p, float (&r)[8]) {
  if constexpr (POL == 1)
    asm volatile("ld.global.L1::no_allocate.L2::evict_first.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7])
                 : "l"(p) : "memory");
  else if constexpr (POL == 2)
    asm volatile("ld.global.L1::no_allocate.L2::evict_last.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7])
                 : "l"(p) : "memory");
  else
    ld_f8(p, r);
}
template <int POL> __device__ __forceinline__ void st_f8_l2(void* p, const float (&r)[8]) {
  if constexpr (POL == 1)
    asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "f"(r[0]), "f"(r[1]), "f"(r[2]), "f"(r[3]), "f"(r[4]), "f"(r[5]), "f"(r[6]), "f"(r[7])
                 : "memory");
  else if constexpr (POL == 2)
    asm volatile("st.global.L1::no_allocate.L2::evict_last.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "f"(r[0]), "f"(r[1]), "f"(r[2]), "f"(r[3]), "f"(r[4]), "f"(r[5]), "f"(r[6]), "f"(r[7])
                 : "memory");
  else
    st_f8(p, r);
}
__device__ __forceinline__ uint4 ld_u4(const void* p) {
  uint4 r;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_u4(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void bf16x2_to_f(unsigned v, float& lo, float& hi) {
  lo = __uint_as_float(v << 16); hi = __uint_as_float(v & 0xffff0000u);
}
__device__ __forceinline__ unsigned f_to_bf16x2(float lo, float hi) {
  __nv_bfloat162 a = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<unsigned*>(&a);
}
__device__ __forceinline__ void f16x2_to_f(unsigned v, float& lo, float& hi) {
  float2 a = __half22float2(*reinterpret_cast<__half2*>(&v)); lo = a.x; hi = a.y;
}
__device__ __forceinline__ unsigned f_to_f16x2(float lo, float hi) {
  __half2 a = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<unsigned*>(&a);
}

// Elem<T>: SIZE bytes per element; VECTORISED types are accessed as vectors of V = 4 or 8 elements
// (V*SIZE bytes, aligned to that), the others element by element.
template <typename T> struct Elem;

template <> struct Elem<float> {
  static constexpr int SIZE = 4, QUAD_ALIGN = 16;
  static __device__ __forceinline__ void ld4(const void* p, float (&r)[4]) {
    float4 v = ld_f4(reinterpret_cast<const float4*>(p));
    r[0] = v.x; r[1] = v.y; r[2] = v.z; r[3] = v.w;
  }
  static __device__ __forceinline__ void st4(void* p, const float (&r)[4]) {
    st_f4(reinterpret_cast<float4*>(p), make_float4(r[0], r[1], r[2], r[3]));
  }
  static __device__ __forceinline__ void ld8(const void* p, float (&r)[8]) { ld_f8(p, r); }
  static __device__ __forceinline__ void st8(void* p, const float (&r)[8]) { st_f8(p, r); }
  static __device__ __forceinline__ float ld1(const void* p) { return *reinterpret_cast<const float*>(p); }
  static __device__ __forceinline__ void st1(void* p, float v) { *reinterpret_cast<float*>(p) = v; }
  static __device__ __forceinline__ float round_trip(float v) { return v; }
};

template <> struct Elem<__half> {
  static constexpr int SIZE = 2, QUAD_ALIGN = 8;
  static __device__ __forceinline__ void ld4(const void* p, float (&r)[4]) {
    uint2 v = ld_u2(reinterpret_cast<const uint2*>(p));
    float2 a = __half22float2(*reinterpret_cast<__half2*>(&v.x));
    float2 b = __half22float2(*reinterpret_cast<__half2*>(&v.y));
    r[0] = a.x; r[1] = a.y; r[2] = b.x; r[3] = b.y;
  }
  static __device__ __forceinline__ void st4(void* p, const float (&r)[4]) {
    __half2 a = __floats2half2_rn(r[0], r[1]);
    __half2 b = __floats2half2_rn(r[2], r[3]);
    uint2 v;
    v.x = *reinterpret_cast<unsigned*>(&a);
    v.y = *reinterpret_cast<unsigned*>(&b);
    st_u2(reinterpret_cast<uint2*>(p), v);
  }
  static __device__ __forceinline__ void ld8(const void* p, float (&r)[8]) {
    uint4 v = ld_u4(p);
    f16x2_to_f(v.x, r[0], r[1]); f16x2_to_f(v.y, r[2], r[3]); f16x2_to_f(v.z, r[4], r[5]); f16x2_to_f(v.w, r[6], r[7]);
  }
  static __device__ __forceinline__ void st8(void* p, const float (&r)[8]) {
    st_u4(p, make_uint4(f_to_f16x2(r[0], r[1]), f_to_f16x2(r[2], r[3]), f_to_f16x2(r[4], r[5]), f_to_f16x2(r[6], r[7])));
  }
  static __device__ __forceinline__ float ld1(const void* p) { return __half2float(*reinterpret_cast<const __half*>(p)); }
  static __device__ __forceinline__ void st1(void* p, float v) { *reinterpret_cast<__half*>(p) = __float2half_rn(v); }
  static __device__ __forceinline__ float round_trip(float v) { return __half2float(__float2half_rn(v)); }
};

template <> struct Elem<__nv_bfloat16> {
  static constexpr int SIZE = 2, QUAD_ALIGN = 8;
  static __device__ __forceinline__ void ld4(const void* p, float (&r)[4]) {
    uint2 v = ld_u2(reinterpret_cast<const uint2*>(p));
    // bf16 -> fp32 is a 16-bit left shift
    r[0] = __uint_as_float(v.x << 16); r[1] = __uint_as_float(v.x & 0xffff0000u);
    r[2] = __uint_as_float(v.y << 16); r[3] = __uint_as_float(v.y & 0xffff0000u);
  }
  static __device__ __forceinline__ void st4(void* p, const float (&r)[4]) {
    __nv_bfloat162 a = __floats2bfloat162_rn(r[0], r[1]);
    __nv_bfloat162 b = __floats2bfloat162_rn(r[2], r[3]);
    uint2 v;
    v.x = *reinterpret_cast<unsigned*>(&a);
    v.y = *reinterpret_cast<unsigned*>(&b);
    st_u2(reinterpret_cast<uint2*>(p), v);
  }
  static __device__ __forceinline__ void ld8(const void* p, float (&r)[8]) {
    uint4 v = ld_u4(p);
    bf16x2_to_f(v.x, r[0], r[1]); bf16x2_to_f(v.y, r[2], r[3]); bf16x2_to_f(v.z, r[4], r[5]); bf16x2_to_f(v.w, r[6], r[7]);
  }
  static __device__ __forceinline__ void st8(void* p, const float (&r)[8]) {
    st_u4(p, make_uint4(f_to_bf16x2(r[0], r[1]), f_to_bf16x2(r[2], r[3]), f_to_bf16x2(r[4], r[5]), f_to_bf16x2(r[6], r[7])));
  }
  static __device__ __forceinline__ float ld1(const void* p) { return __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(p)); }
  static __device__ __forceinline__ void st1(void* p, float v) { *reinterpret_cast<__nv_bfloat16*>(p) = __float2bfloat16_rn(v); }
  static __device__ __forceinline__ float round_trip(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
};

// fp64 lists (multi_tensor_adam accepts them): arithmetic stays fp32 as in the reference
// (MATH_T = float, csrc/multi_tensor_adam.cu:19), so a quad is four 8-byte accesses.
template <> struct Elem<double> {
  static constexpr int SIZE = 8, QUAD_ALIGN = 8;
  static __device__ __forceinline__ void ld4(const void* p, float (&r)[4]) {
    const double* d = reinterpret_cast<const double*>(p);
#pragma unroll
    for (int i = 0; i < 4; i++) r[i] = (float)d[i];
  }
  static __device__ __forceinline__ void st4(void* p, const float (&r)[4]) {
    double* d = reinterpret_cast<double*>(p);
#pragma unroll
    for (int i = 0; i < 4; i++) d[i] = (double)r[i];
  }
  static __device__ __forceinline__ void ld8(const void* p, float (&r)[8]) {
    const double* d = reinterpret_cast<const double*>(p);
#pragma unroll
    for (int i = 0; i < 8; i++) r[i] = (float)d[i];
  }
  static __device__ __forceinline__ void st8(void* p, const float (&r)[8]) {
    double* d = reinterpret_cast<double*>(p);
#pragma unroll
    for (int i = 0; i < 8; i++) d[i] = (double)r[i];
  }
  static __device__ __forceinline__ float ld1(const void* p) { return (float)*reinterpret_cast<const double*>(p); }
  static __device__ __forceinline__ void st1(void* p, float v) { *reinterpret_cast<double*>(p) = (double)v; }
  static __device__ __forceinline__ float round_trip(float v) { return v; }
};

// e5m2 bytes (B200MT_U8 lists of maybe_cast / strided_check_finite): the upper byte of an fp16.
// Packing follows the reference exactly (apex/contrib/csrc/optimizers/fused_adam_cuda_kernel.cu:395-463):
// add 1/8 of the value's power of two (round half up in magnitude), round to fp16, keep the high byte.
struct e5m2 {};

__device__ __forceinline__ float e5m2_to_f(unsigned byte) {
  return __half2float(__ushort_as_half((unsigned short)(byte << 8)));
}
__device__ __forceinline__ unsigned f_to_e5m2(float v) {
  const float pow2 = __uint_as_float(__float_as_uint(v) & 0xFF800000u);
  return (unsigned)(__half_as_ushort(__float2half_rn(v + pow2 / 8.0f)) >> 8);
}

template <> struct Elem<e5m2> {
  static constexpr int SIZE = 1, QUAD_ALIGN = 4;
  static __device__ __forceinline__ void ld4(const void* p, float (&r)[4]) {
    const unsigned v = *reinterpret_cast<const unsigned*>(p);
#pragma unroll
    for (int i = 0; i < 4; i++) r[i] = e5m2_to_f((v >> (8 * i)) & 0xffu);
  }
  static __device__ __forceinline__ void st4(void* p, const float (&r)[4]) {
    unsigned v = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) v |= f_to_e5m2(r[i]) << (8 * i);
    *reinterpret_cast<unsigned*>(p) = v;
  }
  static __device__ __forceinline__ void ld8(const void* p, float (&r)[8]) {
    const uint2 v = ld_u2(reinterpret_cast<const uint2*>(p));
#pragma unroll
    for (int i = 0; i < 4; i++) { r[i] = e5m2_to_f((v.x >> (8 * i)) & 0xffu); r[4 + i] = e5m2_to_f((v.y >> (8 * i)) & 0xffu); }
  }
  static __device__ __forceinline__ void st8(void* p, const float (&r)[8]) {
    uint2 v = make_uint2(0u, 0u);
#pragma unroll
    for (int i = 0; i < 4; i++) { v.x |= f_to_e5m2(r[i]) << (8 * i); v.y |= f_to_e5m2(r[4 + i]) << (8 * i); }
    st_u2(reinterpret_cast<uint2*>(p), v);
  }
  static __device__ __forceinline__ float ld1(const void* p) { return e5m2_to_f(*reinterpret_cast<const unsigned char*>(p)); }
  static __device__ __forceinline__ void st1(void* p, float v) { *reinterpret_cast<unsigned char*>(p) = (unsigned char)f_to_e5m2(v); }
  static __device__ __forceinline__ float round_trip(float v) { return e5m2_to_f(f_to_e5m2(v)); }
};

// Multicast (NVLS) views of 16-bit lists, for the fused compute+collective kernels.  The pointer is
// an address inside a CUDA multicast object bound to the same symmetric buffer on every GPU of the
// NVSwitch domain:
//   ld4 = multimem.ld_reduce.add : the switch fetches the 8 bytes from EVERY GPU and returns their sum
//         (accumulated in fp32) -- the reduce-scatter happens inside the gradient load;
//   st4 = multimem.st            : one store, replicated by the switch into every GPU's buffer --
//         the all-gather happens inside the parameter write.
// Only whole, 8-byte aligned quads exist in this address space (the host entry point enforces it),
// so the scalar accessors must never run.
struct mc_bf16 {};
struct mc_f16 {};

template <> struct Elem<mc_bf16> {
  static constexpr int SIZE = 2, QUAD_ALIGN = 8;
  static __device__ __forceinline__ void ld4(const void* p, float (&r)[4]) {
    unsigned x, y;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v2.bf16x2 {%0,%1}, [%2];"
                 : "=r"(x), "=r"(y) : "l"(p) : "memory");
    r[0] = __uint_as_float(x << 16); r[1] = __uint_as_float(x & 0xffff0000u);
    r[2] = __uint_as_float(y << 16); r[3] = __uint_as_float(y & 0xffff0000u);
  }
  static __device__ __forceinline__ void st4(void* p, const float (&r)[4]) {
    __nv_bfloat162 a = __floats2bfloat162_rn(r[0], r[1]);
    __nv_bfloat162 b = __floats2bfloat162_rn(r[2], r[3]);
    asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1,%2};"
                 :: "l"(p), "r"(*reinterpret_cast<unsigned*>(&a)), "r"(*reinterpret_cast<unsigned*>(&b)) : "memory");
  }
  static __device__ __forceinline__ void ld8(const void* p, float (&r)[8]) {
    unsigned x, y, z, w;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "l"(p) : "memory");
    bf16x2_to_f(x, r[0], r[1]); bf16x2_to_f(y, r[2], r[3]); bf16x2_to_f(z, r[4], r[5]); bf16x2_to_f(w, r[6], r[7]);
  }
  static __device__ __forceinline__ void st8(void* p, const float (&r)[8]) {
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(f_to_bf16x2(r[0], r[1])), "r"(f_to_bf16x2(r[2], r[3])), "r"(f_to_bf16x2(r[4], r[5])),
                    "r"(f_to_bf16x2(r[6], r[7])) : "memory");
  }
  static __device__ __forceinline__ float ld1(const void*) { __trap(); return 0.f; }
  static __device__ __forceinline__ void st1(void*, float) { __trap(); }
  static __device__ __forceinline__ float round_trip(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
};

template <> struct Elem<mc_f16> {
  static constexpr int SIZE = 2, QUAD_ALIGN = 8;
  static __device__ __forceinline__ void ld4(const void* p, float (&r)[4]) {
    unsigned x, y;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v2.f16x2 {%0,%1}, [%2];"
                 : "=r"(x), "=r"(y) : "l"(p) : "memory");
    float2 a = __half22float2(*reinterpret_cast<__half2*>(&x));
    float2 b = __half22float2(*reinterpret_cast<__half2*>(&y));
    r[0] = a.x; r[1] = a.y; r[2] = b.x; r[3] = b.y;
  }
  static __device__ __forceinline__ void st4(void* p, const float (&r)[4]) {
    __half2 a = __floats2half2_rn(r[0], r[1]);
    __half2 b = __floats2half2_rn(r[2], r[3]);
    asm volatile("multimem.st.relaxed.sys.global.v2.f16x2 [%0], {%1,%2};"
                 :: "l"(p), "r"(*reinterpret_cast<unsigned*>(&a)), "r"(*reinterpret_cast<unsigned*>(&b)) : "memory");
  }
  static __device__ __forceinline__ void ld8(const void* p, float (&r)[8]) {
    unsigned x, y, z, w;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "l"(p) : "memory");
    f16x2_to_f(x, r[0], r[1]); f16x2_to_f(y, r[2], r[3]); f16x2_to_f(z, r[4], r[5]); f16x2_to_f(w, r[6], r[7]);
  }
  static __device__ __forceinline__ void st8(void* p, const float (&r)[8]) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f16x2 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(f_to_f16x2(r[0], r[1])), "r"(f_to_f16x2(r[2], r[3])), "r"(f_to_f16x2(r[4], r[5])),
                    "r"(f_to_f16x2(r[6], r[7])) : "memory");
  }
  static __device__ __forceinline__ float ld1(const void*) { __trap(); return 0.f; }
  static __device__ __forceinline__ void st1(void*, float) { __trap(); }
  static __device__ __forceinline__ float round_trip(float v) { return __half2float(__float2half_rn(v)); }
};

// Two-GPU peer view of a 16-bit list (b200mt_dist_adam_p2p): the pointer addresses THIS GPU's copy of a buffer that has a
// twin at a fixed byte distance in the peer GPU's memory (both mapped into this process: symmetric memory).
//   ld = this GPU's value + the peer's value, summed in fp32 and rounded once to the list's type -- the reduce-scatter of
//        two ranks done by the load (what NCCL would hand over, without the trip through a reduced buffer);
//   st = the value stored here AND in the peer's twin -- the all-gather of two ranks done by the store.
// With two ranks this moves 2 bytes per element per direction for each collective; the multicast path moves
// 2*(1 + 1/W) = 3: the switch pulls the requester's own half through the links as well.
// The distances live in constant memory, written on the launching stream before the launch (b200mt_dist_adam_p2p).
struct pair_bf16 {};
__constant__ long long c_pair_delta[2];   // [0]: loads (gradients), [1]: stores (parameter copy); bytes to the peer's twin

__device__ __forceinline__ uint4 ld_u4_sys(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_u4_sys(void* p, uint4 v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <> struct Elem<pair_bf16> {
  static constexpr int SIZE = 2, QUAD_ALIGN = 8;
  static __device__ __forceinline__ void ld8(const void* p, float (&r)[8]) {
    const uint4 a = ld_u4(p);
    const uint4 b = ld_u4_sys(reinterpret_cast<const char*>(p) + c_pair_delta[0]);
    float x[8], y[8];
    bf16x2_to_f(a.x, x[0], x[1]); bf16x2_to_f(a.y, x[2], x[3]); bf16x2_to_f(a.z, x[4], x[5]); bf16x2_to_f(a.w, x[6], x[7]);
    bf16x2_to_f(b.x, y[0], y[1]); bf16x2_to_f(b.y, y[2], y[3]); bf16x2_to_f(b.z, y[4], y[5]); bf16x2_to_f(b.w, y[6], y[7]);
#pragma unroll
    for (int i = 0; i < 8; i++) r[i] = __bfloat162float(__float2bfloat16_rn(x[i] + y[i]));
  }
  static __device__ __forceinline__ void st8(void* p, const float (&r)[8]) {
    const uint4 v = make_uint4(f_to_bf16x2(r[0], r[1]), f_to_bf16x2(r[2], r[3]), f_to_bf16x2(r[4], r[5]), f_to_bf16x2(r[6], r[7]));
    st_u4(p, v);
    st_u4_sys(reinterpret_cast<char*>(p) + c_pair_delta[1], v);
  }
  // only whole 16-byte vectors exist on this path (the host entry point enforces it)
  static __device__ __forceinline__ void ld4(const void*, float (&)[4]) { __trap(); }
  static __device__ __forceinline__ void st4(void*, const float (&)[4]) { __trap(); }
  static __device__ __forceinline__ float ld1(const void*) { __trap(); return 0.f; }
  static __device__ __forceinline__ void st1(void*, float) { __trap(); }
  static __device__ __forceinline__ float round_trip(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
};

// Store-only multicast view of an e5m2 byte list (the compressed all-gather of the sharded Adam):
// 4 or 8 packed bytes per multimem.st, replicated by the switch.
struct mc_e5m2 {};

template <> struct Elem<mc_e5m2> {
  static constexpr int SIZE = 1, QUAD_ALIGN = 4;
  static __device__ __forceinline__ void ld4(const void*, float (&)[4]) { __trap(); }
  static __device__ __forceinline__ void ld8(const void*, float (&)[8]) { __trap(); }
  static __device__ __forceinline__ void st4(void* p, const float (&r)[4]) {
    unsigned v = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) v |= f_to_e5m2(r[i]) << (8 * i);
    asm volatile("multimem.st.relaxed.sys.global.b32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
  }
  static __device__ __forceinline__ void st8(void* p, const float (&r)[8]) {
    unsigned lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) { lo |= f_to_e5m2(r[i]) << (8 * i); hi |= f_to_e5m2(r[4 + i]) << (8 * i); }
    const unsigned long long v = ((unsigned long long)hi << 32) | lo;
    asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
  }
  static __device__ __forceinline__ float ld1(const void*) { __trap(); return 0.f; }
  static __device__ __forceinline__ void st1(void*, float) { __trap(); }
  static __device__ __forceinline__ float round_trip(float v) { return e5m2_to_f(f_to_e5m2(v)); }
};

// vector access of V = 4 or 8 elements
// POL: L2 eviction priority (see ld_f8_l2); honoured for 256-bit fp32 vectors, ignored elsewhere
template <typename T, int V, int POL = 0> __device__ __forceinline__ void ldv(const void* p, float (&r)[V]) {
  static_assert(V == 4 || V == 8, "vector width");
  if constexpr (V == 8 && POL != 0 && std::is_same<T, float>::value) ld_f8_l2<POL>(p, r);
  else if constexpr (V == 4 && POL != 0 && std::is_same<T, float>::value) ld_f4_l2<POL>(p, r);
  else if constexpr (V == 4) Elem<T>::ld4(p, r); else Elem<T>::ld8(p, r);
}
template <typename T, int V, int POL = 0> __device__ __forceinline__ void stv(void* p, const float (&r)[V]) {
  if constexpr (V == 8 && POL != 0 && std::is_same<T, float>::value) st_f8_l2<POL>(p, r);
  else if constexpr (V == 4 && POL != 0 && std::is_same<T, float>::value) st_f4_l2<POL>(p, r);
  else if constexpr (V == 4) Elem<T>::st4(p, r); else Elem<T>::st8(p, r);
}
// per-list policy out of an op's two-bits-per-list word (list k in bits 2k, 2k+1)
constexpr int l2_policy(unsigned word, int k) { return (int)((word >> (2 * k)) & 3u); }
// required alignment, in bytes, of a V-element vector of T (element-wise types: just the element)
template <typename T, int V> constexpr int vec_align() {
  return (Elem<T>::QUAD_ALIGN > Elem<T>::SIZE) ? Elem<T>::SIZE * V : Elem<T>::SIZE;
}

// compile-time loop
template <int N, int I = 0, class F>
__device__ __forceinline__ void static_for(F&& f) {
  if constexpr (I < N) {
    f(std::integral_constant<int, I>{});
    static_for<N, I + 1>(f);
  }
}

template <typename... Ts> struct TypeList {
  static constexpr int size = sizeof...(Ts);
  template <int K> using at = typename std::tuple_element<K, std::tuple<Ts...>>::type;
};

// ------------------------------------------------------------------------------------------
// Deterministic CTA-wide sum: xor-shuffle inside each warp, one shared-memory hop, then the
// first warp reduces the per-warp values in a fixed order.  Result valid in thread 0.
// All threads of the CTA must call it.
// ------------------------------------------------------------------------------------------
template <int THREADS>
__device__ __forceinline__ float cta_sum(float v, float* smem /* >= THREADS/32 floats */) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = THREADS / 32;
  __syncthreads();  // protect smem reuse across consecutive calls
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  float r = 0.f;
  if (warp == 0) {
    r = (lane < NW) ? smem[lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
  }
  return r;
}

// Same shape, max instead of sum (values must be >= 0: the L-infinity norm).  fmaxf drops NaN, as the
// reference's MaxNormFunctor does (csrc/multi_tensor_l2norm_kernel.cu:132-155).
template <int THREADS>
__device__ __forceinline__ float cta_max(float v, float* smem) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = THREADS / 32;
  __syncthreads();
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  float r = 0.f;
  if (warp == 0) {
    r = (lane < NW) ? smem[lane] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) r = fmaxf(r, __shfl_xor_sync(0xffffffffu, r, o));
  }
  return r;
}

// ------------------------------------------------------------------------------------------
// The streaming kernel.  Op supplies:
//   using Types = TypeList<...>;            element type of every list
//   static constexpr int THREADS, UNROLL, MIN_CTAS, VEC;
//   static constexpr unsigned STORES;       bit k set: list k is written back
//   struct State;                           per-thread registers
//   bool  begin(State&) const;              false: whole kernel is a no-op (noop gating)
//   unsigned load_mask() const;             bit k set: list k is read (others arrive as 0.f)
//   unsigned store_mask() const;            run-time subset of STORES (default: all of STORES)
//   void  tile_begin(State&, const TileLoc&) const;
//   void  elem(State&, float (&x)[NL]) const;      the per-element math, fp32 registers
//   void  tile_end(State&, int tile, const TileLoc&, float* smem) const;   (CTA-uniform)
//   void  end(State&) const;
// ------------------------------------------------------------------------------------------
template <class Op>
__global__ void __launch_bounds__(Op::THREADS, Op::MIN_CTAS)
mt_stream_kernel(const TableView tv, const Op op) {
  using Types = typename Op::Types;
  constexpr int NL = Types::size;
  constexpr int THREADS = Op::THREADS;
  constexpr int U = Op::UNROLL;
  constexpr int V = Op::VEC;   // elements per vector access: 4 (128-bit fp32) or 8 (256-bit fp32, 128-bit 16-bit types)
  __shared__ float red_smem[THREADS / 32];

  // Programmatic dependent launch (no-ops unless the launch asked for it): let the next kernel of the
  // stream become resident as soon as every CTA of this grid has started, and do this CTA's tile lookup
  // -- which reads only the descriptor table, never data of the previous kernel -- before waiting for
  // the previous kernel to finish and flush.
  asm volatile("griddepcontrol.launch_dependents;");
  TileLoc L0 = {};
  if ((int)blockIdx.x < tv.ntiles) L0 = locate_tile(tv, blockIdx.x);
  asm volatile("griddepcontrol.wait;" ::: "memory");

  typename Op::State st;
  if (!op.begin(st)) return;
  const unsigned lmask = op.load_mask();
  const unsigned smask = op.store_mask();
  const int tid = threadIdx.x;

  for (int tile = blockIdx.x; tile < tv.ntiles; tile += gridDim.x) {
    const TileLoc L = (tile == (int)blockIdx.x) ? L0 : locate_tile(tv, tile);
    op.tile_begin(st, L);

    // base address of this tile in every list + common head peel to reach vector alignment
    char* base[NL];
    int head = -1;
    bool vec_ok = true;
    static_for<NL>([&](auto k) {
      using T = typename Types::template at<k.value>;
      unsigned long long a = __ldg(tv.ptrs + (size_t)k.value * tv.ntensors + L.tensor);
      a += (unsigned long long)L.off * Elem<T>::SIZE;
      base[k.value] = reinterpret_cast<char*>(a);
      constexpr int A = vec_align<T, V>();
      constexpr int P = A / Elem<T>::SIZE;  // alignment period in elements
      if constexpr (P > 1) {
        const int mis = (int)(a % A) / Elem<T>::SIZE;  // elements past a boundary
        const int hk = (P - mis) % P;
        static_assert(P == V, "all vectorised element types have a V-element period");
        if (head < 0) head = hk; else vec_ok = vec_ok && (head == hk);
      }
    });
    if (head < 0) head = 0;
    if (!vec_ok || head > L.len) head = L.len;  // phases disagree: whole tile on the scalar path
    const int nq = (L.len - head) / V;
    const int tail0 = head + nq * V;

    // ---- vector body: UNROLL vectors per thread in flight, all loads issued before any math
    for (int q0 = 0; q0 < nq; q0 += THREADS * U) {
      float x[U][NL][V];
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int q = q0 + tid + u * THREADS;
        if (q < nq) {
          static_for<NL>([&](auto k) {
            using T = typename Types::template at<k.value>;
            if (lmask & (1u << k.value)) {
              ldv<T, V, l2_policy(Op::L2_LOADS, k.value)>(base[k.value] + ((size_t)head + (size_t)q * V) * Elem<T>::SIZE, x[u][k.value]);
            } else {
#pragma unroll
              for (int i = 0; i < V; i++) x[u][k.value][i] = 0.f;
            }
          });
        }
      }
#pragma unroll
      for (int u = 0; u < U; u++) {
        const int q = q0 + tid + u * THREADS;
        if (q < nq) {
#pragma unroll
          for (int i = 0; i < V; i++) {
            float e[NL];
#pragma unroll
            for (int k = 0; k < NL; k++) e[k] = x[u][k][i];
            op.elem(st, e);
#pragma unroll
            for (int k = 0; k < NL; k++) x[u][k][i] = e[k];
          }
          static_for<NL>([&](auto k) {
            using T = typename Types::template at<k.value>;
            if constexpr ((Op::STORES >> k.value) & 1u) {
              if (smask & (1u << k.value))
                stv<T, V, l2_policy(Op::L2_STORES, k.value)>(base[k.value] + ((size_t)head + (size_t)q * V) * Elem<T>::SIZE, x[u][k.value]);
            }
          });
        }
      }
    }

    // ---- scalar head/tail (< 2*V elements when vectorised, the whole tile otherwise)
    const int nscalar = head + (L.len - tail0);
    for (int s = tid; s < nscalar; s += THREADS) {
      const int i = (s < head) ? s : (tail0 + (s - head));
      float e[NL];
      static_for<NL>([&](auto k) {
        using T = typename Types::template at<k.value>;
        e[k.value] = (lmask & (1u << k.value)) ? Elem<T>::ld1(base[k.value] + (size_t)i * Elem<T>::SIZE) : 0.f;
      });
      op.elem(st, e);
      static_for<NL>([&](auto k) {
        using T = typename Types::template at<k.value>;
        if constexpr ((Op::STORES >> k.value) & 1u) {
          if (smask & (1u << k.value))
            Elem<T>::st1(base[k.value] + (size_t)i * Elem<T>::SIZE, e[k.value]);
        }
      });
    }

    op.tile_end(st, tile, L, red_smem);
  }
  op.end(st);
}

// 1 - beta^step evaluated in double and narrowed, as the reference's device-side
// `1 - std::pow(beta, step)` does (std::pow(float, int) promotes to double;
// csrc/multi_tensor_lamb_mp.cu:66-70, multi_tensor_distopt_adam_kernel.cu:59-63).  The value is tiny
// for beta2 at small steps, so fp32 powf would lose ~1e-4 relative.  Lane 0 computes, warp broadcasts.
__device__ __forceinline__ float bias_correction_dev(float beta, int step) {
  float r = 0.f;
  if ((threadIdx.x & 31) == 0) r = (float)(1.0 - pow((double)beta, (double)step));
  return __shfl_sync(0xffffffffu, r, 0);
}

// Division / square root of the update rules.  ncu on the mixed-precision Adam and LAMB kernels showed ~120
// executed instructions per element with IEEE `div.rn` / `sqrt.rn` (each expands to a Newton sequence plus a
// slow path for special values), enough to make those kernels issue-bound rather than HBM-bound.  The approximate
// forms (MUFU.RCP / MUFU.SQRT, <= 2 ulp) are what the reference itself computes with: it is built --use_fast_math
// (setup.py:210).  B200MT_APPROX_MATH=0 restores the IEEE forms.
#ifndef B200MT_APPROX_MATH
#define B200MT_APPROX_MATH 1
#endif
__device__ __forceinline__ float fast_div(float a, float b) {
#if B200MT_APPROX_MATH
  return __fdividef(a, b);
#else
  return a / b;
#endif
}
__device__ __forceinline__ float fast_sqrt(float a) {
#if B200MT_APPROX_MATH
  float r;
  asm("sqrt.approx.f32 %0, %1;" : "=f"(r) : "f"(a));
  return r;
#else
  return sqrtf(a);
#endif
}
// x / c for a divisor c that is constant over a tile: multiply by the reciprocal taken once
__device__ __forceinline__ float div_const(float x, float c, float rc) {
#if B200MT_APPROX_MATH
  (void)c;
  return x * rc;
#else
  (void)rc;
  return x / c;
#endif
}

// amp_C.multi_tensor_adam's per-element rule (csrc/multi_tensor_adam.cu:94-112), shared by the streaming AdamOp
// and the bulk-copy kernel of ops_adam_tma.cu so that both give the same bits.
__device__ __forceinline__ void adam_rule(float gin, float& p, float& m, float& v, float beta1, float beta2,
                                          float bc1, float bc2, float eps, float lr, float decay, float grad_scale,
                                          int mode) {
  float g = gin * grad_scale;
  const float rbc1 = 1.f / bc1, rbc2 = 1.f / bc2;   // loop-invariant: hoisted by the compiler
  if (mode == 0) {  // L2: decay folded into the gradient
    g = g + decay * p;
    m = beta1 * m + (1.f - beta1) * g;
    v = beta2 * v + (1.f - beta2) * g * g;
    const float mu = div_const(m, bc1, rbc1), vu = div_const(v, bc2, rbc2);
    const float denom = fast_sqrt(vu) + eps;
    p = p - lr * fast_div(mu, denom);
  } else {          // AdamW: decoupled decay
    m = beta1 * m + (1.f - beta1) * g;
    v = beta2 * v + (1.f - beta2) * g * g;
    const float mu = div_const(m, bc1, rbc1), vu = div_const(v, bc2, rbc2);
    const float denom = fast_sqrt(vu) + eps;
    p = p - lr * (fast_div(mu, denom) + decay * p);
  }
}

// CTAs per SM the light streaming ops (scale, axpby, cast, the norm passes: ~20 live registers) are compiled for.
// ncu on the ResNet-50 unscale showed 44.7 % warps active with the register cap of 4 CTAs/SM of round 1
// (profiles/adam_scale_r01_raw.csv).  Swept per op on B200 with the L2 flushed (profiles/sweep_light_r02.txt):
//   scale / axpby   8 CTAs/SM (32 registers): 1000x64K scale 112.5 -> 88.1 us, axpby 149.5 -> 129.1 us
//   fp32 norms      6 CTAs/SM (42 registers; 8 spills the 16 accumulating registers): 1000x64K l2norm 72.7 -> 60.5 us
//   16-bit norms    4 CTAs/SM (one 128-bit load per thread in flight is enough; more CTAs only add tile prologues)
// B200MT_LIGHT_MIN_CTAS, when defined, overrides all of them (the sweep builds).
#ifdef B200MT_LIGHT_MIN_CTAS
#define B200MT_SCALE_MIN_CTAS B200MT_LIGHT_MIN_CTAS
#define B200MT_NORM32_MIN_CTAS B200MT_LIGHT_MIN_CTAS
#define B200MT_NORM16_MIN_CTAS B200MT_LIGHT_MIN_CTAS
#endif
#ifndef B200MT_SCALE_MIN_CTAS
#define B200MT_SCALE_MIN_CTAS 8
#endif
#ifndef B200MT_NORM32_MIN_CTAS
#define B200MT_NORM32_MIN_CTAS 6
#endif
#ifndef B200MT_NORM16_MIN_CTAS
#define B200MT_NORM16_MIN_CTAS 4
#endif

// Default no-op hooks an Op can inherit.
struct OpBase {
  struct State {};
  static constexpr int THREADS = 256;
  static constexpr int UNROLL = 4;
  static constexpr int MIN_CTAS = 4;
  static constexpr int VEC = 4;
  static constexpr unsigned L2_LOADS = 0, L2_STORES = 0;   // two bits per list: 0 none, 1 evict_first, 2 evict_last (ldv / stv)
  __device__ __forceinline__ unsigned store_mask() const { return 0xffffffffu; }
  template <class S> __device__ __forceinline__ bool begin(S&) const { return true; }
  template <class S> __device__ __forceinline__ void tile_begin(S&, const TileLoc&) const {}
  template <class S> __device__ __forceinline__ void tile_end(S&, int, const TileLoc&, float*) const {}
  template <class S> __device__ __forceinline__ void end(S&) const {}
};

}  // namespace b200mt
