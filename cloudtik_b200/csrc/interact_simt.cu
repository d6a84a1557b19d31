// Pairwise dot-product interaction, CUDA-core version (any F, any D).
//
// Replaces DLRM_Net.interact_features (reference P/dlrm_s_pytorch.py:309-334).  This is the
// general-shape kernel: it serves shapes the tcgen05 kernel (interact_tc.cu) does not tile
// (F > 32, D not a multiple of 16/32, tiny test shapes).  One CTA per sample, T[b] staged in
// shared memory as fp32; fp32 accumulation throughout.
#include <algorithm>

#include "common.cuh"

namespace b200 {

constexpr int INT_THREADS = 128;

// strict-lower-triangle pair p -> (i, j), row-major: p = i(i-1)/2 + j, j < i
__device__ __forceinline__ void pair_ij(int p, int& i, int& j) {
  i = (int)((1.f + sqrtf(1.f + 8.f * (float)p)) * 0.5f);
  while (i * (i - 1) / 2 > p) --i;
  while ((i + 1) * i / 2 <= p) ++i;
  j = p - i * (i - 1) / 2;
}

template <typename ET>
__global__ void __launch_bounds__(INT_THREADS)
    interact_fwd_simt_kernel(const ET* __restrict__ T, ET* __restrict__ R, int64_t B, int F, int D, int64_t ldr,
                             int self) {
  extern __shared__ float sT[];  // [F][D+1]
  const int ld = D + 1;
  // self = 1 (arch_interaction_itself, T/dlrm_s_pytorch.py:428): pairs (i, j <= i) instead of (i, j < i)
  const int npairs = F * (F - 1) / 2 + (self ? F : 0);
  const int rw = D + npairs;
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    const ET* Tb = T + b * (int64_t)F * D;
    for (int i = threadIdx.x; i < F * D; i += INT_THREADS) sT[(i / D) * ld + (i % D)] = to_f32<ET>(Tb[i]);
    __syncthreads();
    ET* Rb = R + b * ldr;
    for (int d = threadIdx.x; d < D; d += INT_THREADS) Rb[d] = Tb[d];
    for (int c = rw + threadIdx.x; c < ldr; c += INT_THREADS) Rb[c] = from_f32<ET>(0.f);
    for (int p = threadIdx.x; p < npairs; p += INT_THREADS) {
      int i, j;
      pair_ij(p, i, j);
      i -= self;  // with the diagonal: strict lower triangle of an (F+1) x (F+1) matrix, row shifted by one
      const float* a = sT + i * ld;
      const float* c = sT + j * ld;
      float acc = 0.f;
      for (int d = 0; d < D; ++d) acc = fmaf(a[d], c[d], acc);
      Rb[D + p] = from_f32<ET>(acc);
    }
    __syncthreads();
  }
}

template <typename ET>
__global__ void __launch_bounds__(INT_THREADS)
    interact_bwd_simt_kernel(const ET* __restrict__ T, const ET* __restrict__ dR, ET* __restrict__ dT,
                             int64_t B, int F, int D, int64_t ldr, int self) {
  extern __shared__ float sm[];  // sT [F][D] then sS [F][F]
  float* sT = sm;
  float* sS = sm + F * D;
  const int npairs = F * (F - 1) / 2 + (self ? F : 0);
  for (int64_t b = blockIdx.x; b < B; b += gridDim.x) {
    const ET* Tb = T + b * (int64_t)F * D;
    const ET* dRb = dR + b * ldr;
    for (int i = threadIdx.x; i < F * D; i += INT_THREADS) sT[i] = to_f32<ET>(Tb[i]);
    if (!self)
      for (int i = threadIdx.x; i < F; i += INT_THREADS) sS[i * F + i] = 0.f;
    for (int p = threadIdx.x; p < npairs; p += INT_THREADS) {
      int i, j;
      pair_ij(p, i, j);
      i -= self;
      const float v = to_f32<ET>(dRb[D + p]);
      if (i == j) {
        sS[i * F + i] = 2.f * v;  // S = dZ + dZ^t
      } else {
        sS[i * F + j] = v;
        sS[j * F + i] = v;
      }
    }
    __syncthreads();
    ET* dTb = dT + b * (int64_t)F * D;
    for (int d = threadIdx.x; d < D; d += INT_THREADS) {
      for (int f = 0; f < F; ++f) {
        float acc = (f == 0) ? to_f32<ET>(dRb[d]) : 0.f;
        for (int g = 0; g < F; ++g) acc = fmaf(sS[f * F + g], sT[g * D + d], acc);
        dTb[f * D + d] = from_f32<ET>(acc);
      }
    }
    __syncthreads();
  }
}

int interact_fwd_simt(const void* T, void* R, int64_t B, int64_t F, int64_t D, int64_t ldr, int dtype,
                      cudaStream_t st, int self) {
  const size_t smem = (size_t)F * (D + 1) * 4;
  B200_REQUIRE(smem <= 200 * 1024, B200_EUNSUPPORTED, "interact_fwd: F*D too large for shared memory");
  const int grid = (int)std::min<int64_t>(B, 148 * 16);
  if (dtype == B200_F32) {
    cudaFuncSetAttribute(interact_fwd_simt_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    interact_fwd_simt_kernel<float><<<grid, INT_THREADS, smem, st>>>((const float*)T, (float*)R, B, (int)F, (int)D, ldr, self);
  } else {
    cudaFuncSetAttribute(interact_fwd_simt_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    interact_fwd_simt_kernel<__nv_bfloat16>
        <<<grid, INT_THREADS, smem, st>>>((const __nv_bfloat16*)T, (__nv_bfloat16*)R, B, (int)F, (int)D, ldr, self);
  }
  B200_CHECK_LAUNCH("interact_fwd_simt");
  return B200_OK;
}

int interact_bwd_simt(const void* T, const void* dR, void* dT, int64_t B, int64_t F, int64_t D, int64_t ldr,
                      int dtype, cudaStream_t st, int self) {
  const size_t smem = ((size_t)F * D + (size_t)F * F) * 4;
  B200_REQUIRE(smem <= 200 * 1024, B200_EUNSUPPORTED, "interact_bwd: F*D too large for shared memory");
  const int grid = (int)std::min<int64_t>(B, 148 * 16);
  if (dtype == B200_F32) {
    cudaFuncSetAttribute(interact_bwd_simt_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    interact_bwd_simt_kernel<float>
        <<<grid, INT_THREADS, smem, st>>>((const float*)T, (const float*)dR, (float*)dT, B, (int)F, (int)D, ldr, self);
  } else {
    cudaFuncSetAttribute(interact_bwd_simt_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    interact_bwd_simt_kernel<__nv_bfloat16><<<grid, INT_THREADS, smem, st>>>(
        (const __nv_bfloat16*)T, (const __nv_bfloat16*)dR, (__nv_bfloat16*)dT, B, (int)F, (int)D, ldr, self);
  }
  B200_CHECK_LAUNCH("interact_bwd_simt");
  return B200_OK;
}

}  // namespace b200
