// Pairwise dot-product interaction in fp32 at the HBM roofline (CUDA cores, register-tiled).
//
// Replaces DLRM_Net.interact_features and its autograd (reference P/dlrm_s_pytorch.py:309-334) for the
// reference's DEFAULT precision.  fp32 parity (1e-5) rules out single-pass TF32 tensor-core math, and
// the op is HBM-bound anyway (3 GFLOP against 0.5 GB at B = 32768), so the kernel is built around the
// memory system: ONE WARP PER SAMPLE, the sample's whole [F, D] tile held in registers (lane l owns
// D/32 consecutive columns of every feature: 128-bit coalesced row loads, F of them in flight per
// lane), no shared memory, no block barrier.
//   forward : per group of 32 pairs each lane forms its partial dot products, then a 31-shuffle
//             transpose-reduce leaves pair (32 g + lane) complete on lane `lane` -> coalesced store.
//   backward: the sample's dR pair values are loaded coalesced (one per lane per group of 32) and
//             broadcast by shuffle; every pair value s updates dT_i += s T_j and dT_j += s T_i in
//             registers; dT rows leave as 128-bit coalesced stores.
// F is a template parameter (pair indices must be compile-time for the register tile); the Criteo
// shape F = 27 is instantiated, other shapes use interact_simt.cu.
#include <algorithm>
#include <utility>

#include "common.cuh"

namespace b200 {

constexpr int IF_THREADS = 128;

template <int VPL>
__device__ __forceinline__ void ld_row(const float* p, float* v) {
  if constexpr (VPL == 4) {
    const uint4 q = ld_nc_v4(p);
    v[0] = __uint_as_float(q.x), v[1] = __uint_as_float(q.y), v[2] = __uint_as_float(q.z), v[3] = __uint_as_float(q.w);
  } else if constexpr (VPL == 2) {
    const uint2 q = ld_nc_v2(p);
    v[0] = __uint_as_float(q.x), v[1] = __uint_as_float(q.y);
  } else {
    v[0] = __ldg(p);
  }
}

template <int VPL>
__device__ __forceinline__ void st_row(float* p, const float* v) {
  if constexpr (VPL == 4) {
    st_cs_v4(p, make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3])));
  } else if constexpr (VPL == 2) {
    st_cs_v2(p, make_uint2(__float_as_uint(v[0]), __float_as_uint(v[1])));
  } else {
    *p = v[0];
  }
}

// 32 values per lane -> lane L ends with the sum over all lanes of value L (31 shuffles)
__device__ __forceinline__ float transpose_reduce32(float (&p)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int k = 0; k < off; ++k) {
      const float a = p[k], c = p[k + off];
      const float send = up ? a : c;
      const float keep = up ? c : a;
      p[k] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return p[0];
}

// strict-lower-triangle pair p -> (i, j), row-major (p = i(i-1)/2 + j, j < i), at compile time
constexpr int pair_i_c(int p) {
  int i = 1;
  while ((i + 1) * i / 2 <= p) ++i;
  return i;
}
constexpr int pair_j_c(int p) { return p - pair_i_c(p) * (pair_i_c(p) - 1) / 2; }

template <int F, int VPL, int P>
__device__ __forceinline__ float pair_dot(const float (&t)[F][VPL]) {
  constexpr int NP = F * (F - 1) / 2;
  if constexpr (P >= NP) {
    return 0.f;
  } else {
    constexpr int I = pair_i_c(P), J = pair_j_c(P);
    float s = 0.f;
#pragma unroll
    for (int v = 0; v < VPL; ++v) s = fmaf(t[I][v], t[J][v], s);
    return s;
  }
}

// one group of 32 pairs: partial dots (register indices are template constants), transpose-reduce, store
template <int F, int VPL, int G, int... K>
__device__ __forceinline__ void fwd_group(const float (&t)[F][VPL], float* Rpairs, int lane,
                                          std::integer_sequence<int, K...>) {
  constexpr int NP = F * (F - 1) / 2;
  float p[32];
  ((p[K] = pair_dot<F, VPL, G * 32 + K>(t)), ...);
  const float tot = transpose_reduce32(p, lane);
  if (G * 32 + lane < NP) Rpairs[G * 32 + lane] = tot;
}

template <int F, int VPL, int... G>
__device__ __forceinline__ void fwd_groups(const float (&t)[F][VPL], float* Rpairs, int lane,
                                           std::integer_sequence<int, G...>) {
  (fwd_group<F, VPL, G>(t, Rpairs, lane, std::make_integer_sequence<int, 32>{}), ...);
}

template <int F, int VPL>
__global__ void __launch_bounds__(IF_THREADS, VPL >= 2 ? 3 : 5)
    interact_fwd_f32_kernel(const float* __restrict__ T, float* __restrict__ R, int64_t B, int64_t ldr, bool r_vec) {
  constexpr int D = 32 * VPL;
  constexpr int NP = F * (F - 1) / 2;
  constexpr int W = D + NP;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * IF_THREADS + threadIdx.x) >> 5;
  const int64_t nwarp = ((int64_t)gridDim.x * IF_THREADS) >> 5;
  for (int64_t b = warp; b < B; b += nwarp) {
    const float* Tb = T + b * (int64_t)F * D + lane * VPL;
    float t[F][VPL];
#pragma unroll
    for (int f = 0; f < F; ++f) ld_row<VPL>(Tb + f * D, t[f]);
    float* Rb = R + b * ldr;
    if (r_vec) {  // x pass-through
      st_row<VPL>(Rb + lane * VPL, t[0]);
    } else {  // R rows not 16-byte aligned (unpadded ldr = D + 351): scalar stores, still one 128 B line per 32 lanes
#pragma unroll
      for (int v = 0; v < VPL; ++v) Rb[lane * VPL + v] = t[0][v];
    }
    fwd_groups<F, VPL>(t, Rb + D, lane, std::make_integer_sequence<int, (NP + 31) / 32>{});
    for (int c = W + lane; c < ldr; c += 32) Rb[c] = 0.f;  // caller-requested padding columns
  }
}

template <int F, int VPL>
__global__ void __launch_bounds__(IF_THREADS)
    interact_bwd_f32_kernel(const float* __restrict__ T, const float* __restrict__ dR, float* __restrict__ dT,
                            int64_t B, int64_t ldr, bool r_vec) {
  constexpr int D = 32 * VPL;
  constexpr int NP = F * (F - 1) / 2;
  constexpr int NGRP = (NP + 31) / 32;
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * IF_THREADS + threadIdx.x) >> 5;
  const int64_t nwarp = ((int64_t)gridDim.x * IF_THREADS) >> 5;
  for (int64_t b = warp; b < B; b += nwarp) {
    const float* Tb = T + b * (int64_t)F * D + lane * VPL;
    const float* dRb = dR + b * ldr;
    float t[F][VPL], acc[F][VPL], sv[NGRP];
#pragma unroll
    for (int f = 0; f < F; ++f) ld_row<VPL>(Tb + f * D, t[f]);
#pragma unroll
    for (int g = 0; g < NGRP; ++g) sv[g] = (g * 32 + lane < NP) ? __ldg(dRb + D + g * 32 + lane) : 0.f;
    if (r_vec) {  // gradient of the x pass-through
      ld_row<VPL>(dRb + lane * VPL, acc[0]);
    } else {
#pragma unroll
      for (int v = 0; v < VPL; ++v) acc[0][v] = __ldg(dRb + lane * VPL + v);
    }
#pragma unroll
    for (int f = 1; f < F; ++f)
#pragma unroll
      for (int v = 0; v < VPL; ++v) acc[f][v] = 0.f;
#pragma unroll
    for (int i = 1; i < F; ++i) {
#pragma unroll
      for (int j = 0; j < i; ++j) {
        const int pidx = i * (i - 1) / 2 + j;
        const float s = __shfl_sync(0xffffffffu, sv[pidx >> 5], pidx & 31);
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
          acc[i][v] = fmaf(s, t[j][v], acc[i][v]);
          acc[j][v] = fmaf(s, t[i][v], acc[j][v]);
        }
      }
    }
    float* dTb = dT + b * (int64_t)F * D + lane * VPL;
#pragma unroll
    for (int f = 0; f < F; ++f) st_row<VPL>(dTb + f * D, acc[f]);
  }
}

bool interact_f32_supported(int64_t B, int64_t F, int64_t D, int dtype, const void* T, const void* R, int64_t ldr) {
  if (dtype != B200_F32 || F != 27 || !(D == 32 || D == 64 || D == 128)) return false;
  const int vb = (int)(D / 32) * 4;  // bytes per lane per row of T / dT
  return (reinterpret_cast<uintptr_t>(T) % vb) == 0 && (reinterpret_cast<uintptr_t>(R) % 4) == 0;
}

// rows of R / dR can be read and written with D/32-float vectors?
static bool rows_vectorisable(const void* R, int64_t D, int64_t ldr) {
  const int vb = (int)(D / 32) * 4;
  return (reinterpret_cast<uintptr_t>(R) % vb) == 0 && (ldr * 4) % vb == 0;
}

template <int F, int VPL>
static int launch_fwd(const float* T, float* R, int64_t B, int64_t ldr, cudaStream_t st) {
  static const int per_sm = []() {
    int n = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, interact_fwd_f32_kernel<F, VPL>, IF_THREADS, 0);
    return n;
  }();
  const int64_t grid = std::min<int64_t>(ceil_div(B, IF_THREADS / 32), (int64_t)148 * std::max(per_sm, 1));
  interact_fwd_f32_kernel<F, VPL><<<(unsigned)grid, IF_THREADS, 0, st>>>(T, R, B, ldr, rows_vectorisable(R, 32 * VPL, ldr));
  B200_CHECK_LAUNCH("interact_fwd_f32");
  return B200_OK;
}

template <int F, int VPL>
static int launch_bwd(const float* T, const float* dR, float* dT, int64_t B, int64_t ldr, cudaStream_t st) {
  static const int per_sm = []() {
    int n = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, interact_bwd_f32_kernel<F, VPL>, IF_THREADS, 0);
    return n;
  }();
  const int64_t grid = std::min<int64_t>(ceil_div(B, IF_THREADS / 32), (int64_t)148 * std::max(per_sm, 1));
  interact_bwd_f32_kernel<F, VPL><<<(unsigned)grid, IF_THREADS, 0, st>>>(T, dR, dT, B, ldr, rows_vectorisable(dR, 32 * VPL, ldr));
  B200_CHECK_LAUNCH("interact_bwd_f32");
  return B200_OK;
}

int interact_fwd_f32(const void* T, void* R, int64_t B, int64_t F, int64_t D, int64_t ldr, cudaStream_t st) {
  const float* t = static_cast<const float*>(T);
  float* r = static_cast<float*>(R);
  if (D == 128) return launch_fwd<27, 4>(t, r, B, ldr, st);
  if (D == 64) return launch_fwd<27, 2>(t, r, B, ldr, st);
  return launch_fwd<27, 1>(t, r, B, ldr, st);
}

int interact_bwd_f32(const void* T, const void* dR, void* dT, int64_t B, int64_t F, int64_t D, int64_t ldr,
                     cudaStream_t st) {
  const float* t = static_cast<const float*>(T);
  const float* g = static_cast<const float*>(dR);
  float* o = static_cast<float*>(dT);
  if (D == 128) return launch_bwd<27, 4>(t, g, o, B, ldr, st);
  if (D == 64) return launch_bwd<27, 2>(t, g, o, B, ldr, st);
  return launch_bwd<27, 1>(t, g, o, B, ldr, st);
}

}  // namespace b200
