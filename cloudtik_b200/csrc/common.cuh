// Shared device/host helpers for the b200dlrm kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200dlrm.h"

namespace b200 {

// ---------------------------------------------------------------- host-side status
void set_error(const char* fmt, ...);

#define B200_REQUIRE(cond, code, ...)  \
  do {                                 \
    if (!(cond)) {                     \
      ::b200::set_error(__VA_ARGS__);  \
      return (code);                   \
    }                                  \
  } while (0)

#define B200_CHECK_LAUNCH(name)                                                   \
  do {                                                                            \
    cudaError_t e__ = cudaGetLastError();                                         \
    if (e__ != cudaSuccess) {                                                     \
      ::b200::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__));  \
      return B200_ELAUNCH;                                                        \
    }                                                                             \
  } while (0)

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static inline int dtype_size(int dt) { return dt == B200_BF16 ? 2 : (dt == B200_I8 ? 1 : 4); }

// ---------------------------------------------------------------- device helpers
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// 128-bit accesses.  Loads are volatile asm so that a batch of them stays in program order, and
// keep() pins a loaded value as "used" by an (empty) volatile asm: issuing every load of a batch
// and only then keep()-ing them forces all the loads in flight together.  Left to itself the
// compiler interleaves load / use pairs to save registers, which serialises the memory latency
// these HBM-bound kernels depend on hiding.
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {  // read-only data, no L1 allocation
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ld_nc_v2(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_v4(const void* p) {  // coherent load for read-modify-write rows
  uint4 r;
  asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_cg_v4(const void* p) {  // straight from L2: data another SM has just written
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void keep(uint4& v) { asm volatile("" : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w)); }
__device__ __forceinline__ void keep(uint2& v) { asm volatile("" : "+r"(v.x), "+r"(v.y)); }
__device__ __forceinline__ void st_cs_v4(void* p, uint4 v) { __stcs(reinterpret_cast<uint4*>(p), v); }
__device__ __forceinline__ void st_cs_v2(void* p, uint2 v) { __stcs(reinterpret_cast<uint2*>(p), v); }
__device__ __forceinline__ void st_v4(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }

// fire-and-forget vector reductions performed by the L2 atomic unit (no value returns to the SM)
__device__ __forceinline__ void red_add_v4_f32(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red_add_v4_bf16x2(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("red.global.add.noftz.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d)
               : "memory");
}

__device__ __forceinline__ void red_add_v2_bf16x2(void* p, uint32_t a, uint32_t b) {
  asm volatile("red.global.add.noftz.v2.bf16x2 [%0], {%1,%2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}

// ---- mbarrier + 1-D bulk async copy (TMA engine; SASS: UBLKCP / SYNCS)
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// global -> shared bulk copy; src, dst 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(smem_u32(dst_smem)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}

// ---- dtype helpers: a 16-byte vector of WT holds VE elements
template <typename T>
struct Vec16;
template <>
struct Vec16<float> {
  static constexpr int VE = 4;
  __device__ static __forceinline__ void accum(float* acc, uint4 v) {
    acc[0] += __uint_as_float(v.x);
    acc[1] += __uint_as_float(v.y);
    acc[2] += __uint_as_float(v.z);
    acc[3] += __uint_as_float(v.w);
  }
  __device__ static __forceinline__ void unpack(float* out, uint4 v) {
    out[0] = __uint_as_float(v.x);
    out[1] = __uint_as_float(v.y);
    out[2] = __uint_as_float(v.z);
    out[3] = __uint_as_float(v.w);
  }
};
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
template <>
struct Vec16<__nv_bfloat16> {
  static constexpr int VE = 8;
  __device__ static __forceinline__ void accum(float* acc, uint4 v) {
    acc[0] += bf16lo(v.x);
    acc[1] += bf16hi(v.x);
    acc[2] += bf16lo(v.y);
    acc[3] += bf16hi(v.y);
    acc[4] += bf16lo(v.z);
    acc[5] += bf16hi(v.z);
    acc[6] += bf16lo(v.w);
    acc[7] += bf16hi(v.w);
  }
  __device__ static __forceinline__ void unpack(float* out, uint4 v) {
    out[0] = bf16lo(v.x);
    out[1] = bf16hi(v.x);
    out[2] = bf16lo(v.y);
    out[3] = bf16hi(v.y);
    out[4] = bf16lo(v.z);
    out[5] = bf16hi(v.z);
    out[6] = bf16lo(v.w);
    out[7] = bf16hi(v.w);
  }
};

template <>
struct Vec16<int8_t> {  // quantised rows (forward only): 16 values per vector, the row scale is applied by the caller
  static constexpr int VE = 16;
  __device__ static __forceinline__ void unpack(float* out, uint4 v) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int h = 0; h < 4; ++h) {
      out[4 * h + 0] = (float)(int8_t)(w[h] & 0xffu);
      out[4 * h + 1] = (float)(int8_t)((w[h] >> 8) & 0xffu);
      out[4 * h + 2] = (float)(int8_t)((w[h] >> 16) & 0xffu);
      out[4 * h + 3] = (float)(int8_t)(w[h] >> 24);
    }
  }
  __device__ static __forceinline__ void accum(float* acc, uint4 v) {
    float f[16];
    unpack(f, v);
#pragma unroll
    for (int x = 0; x < 16; ++x) acc[x] += f[x];
  }
};

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);  // .x = a (low half)
  return *reinterpret_cast<uint32_t*>(&h);
}

// store N consecutive fp32 values as OT starting at p (p suitably aligned)
template <typename OT, int N, bool STREAM>
__device__ __forceinline__ void store_vec(OT* p, const float* v);

template <>
__device__ __forceinline__ void store_vec<float, 4, true>(float* p, const float* v) {
  st_cs_v4(p, make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]),
                         __float_as_uint(v[3])));
}
template <>
__device__ __forceinline__ void store_vec<float, 8, true>(float* p, const float* v) {
  store_vec<float, 4, true>(p, v);
  store_vec<float, 4, true>(p + 4, v + 4);
}
template <>
__device__ __forceinline__ void store_vec<__nv_bfloat16, 4, true>(__nv_bfloat16* p, const float* v) {
  st_cs_v2(p, make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3])));
}
template <>
__device__ __forceinline__ void store_vec<__nv_bfloat16, 8, true>(__nv_bfloat16* p, const float* v) {
  st_cs_v4(p, make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                         pack_bf16x2(v[6], v[7])));
}
template <>
__device__ __forceinline__ void store_vec<float, 16, true>(float* p, const float* v) {
  store_vec<float, 8, true>(p, v);
  store_vec<float, 8, true>(p + 8, v + 8);
}
template <>
__device__ __forceinline__ void store_vec<__nv_bfloat16, 16, true>(__nv_bfloat16* p, const float* v) {
  store_vec<__nv_bfloat16, 8, true>(p, v);
  store_vec<__nv_bfloat16, 8, true>(p + 8, v + 8);
}
template <>
__device__ __forceinline__ void store_vec<float, 4, false>(float* p, const float* v) {
  st_v4(p, make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]),
                      __float_as_uint(v[3])));
}
template <>
__device__ __forceinline__ void store_vec<__nv_bfloat16, 8, false>(__nv_bfloat16* p, const float* v) {
  st_v4(p, make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                      pack_bf16x2(v[6], v[7])));
}

// add N consecutive fp32 values into OT memory at p with fire-and-forget vector reductions (p 8/16-byte aligned)
template <typename OT, int N>
__device__ __forceinline__ void red_vec(OT* p, const float* v) {
  if constexpr (sizeof(OT) == 4) {
#pragma unroll
    for (int h = 0; h < N / 4; ++h)
      red_add_v4_f32(reinterpret_cast<float*>(p) + 4 * h, v[4 * h], v[4 * h + 1], v[4 * h + 2], v[4 * h + 3]);
  } else if constexpr (N == 4) {
    red_add_v2_bf16x2(p, pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]));
  } else {
#pragma unroll
    for (int h = 0; h < N / 8; ++h)
      red_add_v4_bf16x2(p + 8 * h, pack_bf16x2(v[8 * h], v[8 * h + 1]), pack_bf16x2(v[8 * h + 2], v[8 * h + 3]),
                        pack_bf16x2(v[8 * h + 4], v[8 * h + 5]), pack_bf16x2(v[8 * h + 6], v[8 * h + 7]));
  }
}

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <>
__device__ __forceinline__ float to_f32<int8_t>(int8_t v) { return (float)v; }
template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

#endif  // __CUDACC__
}  // namespace b200
