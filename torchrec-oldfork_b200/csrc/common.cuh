// Shared device helpers for the B200 sparse-path kernels (sm_100a only).
#pragma once
#include <type_traits>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200_sparse.h"

#define B200_LAUNCH_CHECK()                 \
  do {                                      \
    cudaError_t e__ = cudaGetLastError();   \
    if (e__ != cudaSuccess) return (int)e__; \
  } while (0)

#define B200_CUDA_TRY(expr)                 \
  do {                                      \
    cudaError_t e__ = (expr);               \
    if (e__ != cudaSuccess) return (int)e__; \
  } while (0)

namespace b200 {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;
// SM count of the current device, queried once per device (B200: 148 = 2 dies x 74 SMs).  Host code only.
inline int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}
#define kNumSMs (::b200::num_sms())

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---- 128/64-bit global accesses with cache hints ------------------------------------
// Embedding rows are touched once per kernel: stream them past L1.
__device__ __forceinline__ uint4 ldg_stream_16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ldg_stream_8(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
               : "=r"(r.x), "=r"(r.y)
               : "l"(p));
  return r;
}
// Plain (coherent) loads: rows that this kernel also writes (weights in the update).
__device__ __forceinline__ uint4 ld_16(const void* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ uint2 ld_8(const void* p) { return *reinterpret_cast<const uint2*>(p); }
__device__ __forceinline__ void st_stream_16(void* p, uint4 v) {
  asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void st_stream_8(void* p, uint2 v) {
  asm volatile("st.global.cs.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

// ---- element-type traits: a "vec" is EPV consecutive row elements held as floats ------
template <typename W>
struct ElemCvt;
template <>
struct ElemCvt<float> {
  static __device__ __forceinline__ void unpack2(uint32_t, float&, float&) {}
};
template <>
struct ElemCvt<__half> {
  static __device__ __forceinline__ void unpack2(uint32_t u, float& a, float& b) {
    __half2 h = *reinterpret_cast<__half2*>(&u);
    float2 f = __half22float2(h);
    a = f.x;
    b = f.y;
  }
  static __device__ __forceinline__ uint32_t pack2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};
template <>
struct ElemCvt<__nv_bfloat16> {
  static __device__ __forceinline__ void unpack2(uint32_t u, float& a, float& b) {
    // bf16 -> fp32 is a 16-bit shift
    a = __uint_as_float(u << 16);
    b = __uint_as_float(u & 0xffff0000u);
  }
  static __device__ __forceinline__ uint32_t pack2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
};

// Load EPV elements of type W starting at p (16-byte or 8-byte aligned as EPV*sizeof(W)).
template <typename W, int EPV, bool STREAM>
__device__ __forceinline__ void load_vec(const W* p, float (&f)[EPV]) {
  if constexpr (sizeof(W) == 4) {
    static_assert(EPV == 4, "fp32 rows use 4-element vectors");
    uint4 r = STREAM ? ldg_stream_16(p) : ld_16(p);
    f[0] = __uint_as_float(r.x);
    f[1] = __uint_as_float(r.y);
    f[2] = __uint_as_float(r.z);
    f[3] = __uint_as_float(r.w);
  } else if constexpr (EPV == 8) {
    uint4 r = STREAM ? ldg_stream_16(p) : ld_16(p);
    ElemCvt<W>::unpack2(r.x, f[0], f[1]);
    ElemCvt<W>::unpack2(r.y, f[2], f[3]);
    ElemCvt<W>::unpack2(r.z, f[4], f[5]);
    ElemCvt<W>::unpack2(r.w, f[6], f[7]);
  } else {
    static_assert(EPV == 4, "16-bit rows use 8- or 4-element vectors");
    uint2 r = STREAM ? ldg_stream_8(p) : ld_8(p);
    ElemCvt<W>::unpack2(r.x, f[0], f[1]);
    ElemCvt<W>::unpack2(r.y, f[2], f[3]);
  }
}

// Store EPV floats as type O at p (round-to-nearest-even for 16-bit types).
template <typename O, int EPV, bool STREAM>
__device__ __forceinline__ void store_vec(O* p, const float (&f)[EPV]) {
  if constexpr (sizeof(O) == 4) {
#pragma unroll
    for (int i = 0; i < EPV; i += 4) {
      uint4 v = make_uint4(__float_as_uint(f[i]), __float_as_uint(f[i + 1]),
                           __float_as_uint(f[i + 2]), __float_as_uint(f[i + 3]));
      if (STREAM)
        st_stream_16(p + i, v);
      else
        *reinterpret_cast<uint4*>(p + i) = v;
    }
  } else if constexpr (EPV == 8) {
    uint4 v = make_uint4(ElemCvt<O>::pack2(f[0], f[1]), ElemCvt<O>::pack2(f[2], f[3]),
                         ElemCvt<O>::pack2(f[4], f[5]), ElemCvt<O>::pack2(f[6], f[7]));
    if (STREAM)
      st_stream_16(p, v);
    else
      *reinterpret_cast<uint4*>(p) = v;
  } else {
    uint2 v = make_uint2(ElemCvt<O>::pack2(f[0], f[1]), ElemCvt<O>::pack2(f[2], f[3]));
    if (STREAM)
      st_stream_8(p, v);
    else
      *reinterpret_cast<uint2*>(p) = v;
  }
}

// ---- stochastic rounding of fp32 -> fp16 / bf16 (reference: stochastic_rounding_vector,
// fbgemm_gpu/include/fbgemm_gpu/fbgemm_cuda_utils.cuh, Philox4x32-10 seeded from the CUDA generator).
// Random bits are added below the last mantissa bit the 16-bit type keeps (13 for fp16, 16 for bf16)
// and the remainder is cleared, so the value handed to the round-to-nearest pack is already
// representable: P(round up) = discarded fraction / ulp.  Exact for normal fp16 values (like the
// reference, fp16 subnormals fall back to round-to-nearest).
struct SrArgs {
  uint64_t seed;
  uint64_t offset;
  const void* base;  // first element of the weight buffer: vector index = (p - base) / EPV
  int32_t on;
};

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t h0 = __umulhi(0xD2511F53u, c.x), l0 = 0xD2511F53u * c.x;
    const uint32_t h1 = __umulhi(0xCD9E8D57u, c.z), l1 = 0xCD9E8D57u * c.z;
    c = make_uint4(h1 ^ c.y ^ k.x, l1, h0 ^ c.w ^ k.y, l0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

template <typename W>
__device__ __forceinline__ float sr_bits(float x, uint32_t r16) {
  if constexpr (sizeof(W) == 2) {
    constexpr bool kBf16 = std::is_same<W, __nv_bfloat16>::value;
    constexpr uint32_t kMask = kBf16 ? 0xffffu : 0x1fffu;
    const uint32_t u = __float_as_uint(x);
    if ((u & 0x7f800000u) == 0x7f800000u) return x;  // inf / nan pass through
    return __uint_as_float((u + (r16 & kMask)) & ~kMask);
  } else {
    return x;
  }
}

// Perturbs the EPV fp32 values that are about to be stored at p (type W) so that the
// round-to-nearest store rounds them stochastically.  One Philox call per vector.
template <typename W, int EPV>
__device__ __forceinline__ void sr_prepare(const SrArgs& sr, const W* p, float (&w)[EPV]) {
  if constexpr (sizeof(W) == 2) {
    if (sr.on) {
      const uint64_t vec = (uint64_t)(p - reinterpret_cast<const W*>(sr.base)) / EPV;
      const uint4 r = philox4x32_10(make_uint4((uint32_t)vec, (uint32_t)(vec >> 32), (uint32_t)sr.offset,
                                               (uint32_t)(sr.offset >> 32)),
                                    make_uint2((uint32_t)sr.seed, (uint32_t)(sr.seed >> 32)));
      const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
      if constexpr (EPV == 8) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          w[2 * i] = sr_bits<W>(w[2 * i], rr[i] & 0xffffu);
          w[2 * i + 1] = sr_bits<W>(w[2 * i + 1], rr[i] >> 16);
        }
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) w[i] = sr_bits<W>(w[i], rr[i] >> 16);
      }
    }
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// Index / offset loads with a runtime width (int32 or int64); the branch is warp-uniform.
__device__ __forceinline__ int64_t load_index(const void* base, int64_t i, bool is64) {
  return is64 ? reinterpret_cast<const int64_t*>(base)[i]
              : (int64_t) reinterpret_cast<const int32_t*>(base)[i];
}

}  // namespace b200
