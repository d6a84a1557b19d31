// ReLU backward fused with the bias gradient of the Linear layer in front of it (bf16, sm_100a).
//
// The dense MLPs stay on torch / cuBLAS (reference P/dlrm_s_pytorch.py:172-211 builds Linear+ReLU
// stacks; its IPEX fast path fuses them, :926-953).  Around the backward GEMMs autograd launches, per
// layer, a threshold_backward pass over the [B, N] output gradient AND a column reduction of the same
// matrix for the bias gradient -- two trips through HBM plus (in the bf16 fast path) a skinny GEMM with
// a split-K reduction.  This kernel makes ONE trip: it reads dY and the saved activation Y, writes
// dZ = dY * (Y > 0) and accumulates the fp32 column sums of dZ on the way.
//   grid = (column tiles of 256, row chunks of 32..128 rows, sized so that even a narrow layer fills the
//   GPU); thread = 8 consecutive columns (one 16-byte vector) x every 8th row of the chunk, 4 row vectors
//   of each operand in flight;
//   per-chunk sums go to a [chunks, N] fp32 scratch and a second tiny kernel adds the chunks in a fixed order
//   (deterministic: no atomics).
#include <algorithm>

#include "common.cuh"

namespace b200 {

constexpr int RB_THREADS = 256;
constexpr int RB_ROWS = 128;   // most rows per CTA
constexpr int RB_ROWS_MIN = 32;  // fewest (= one pass of the 8 row lanes x 4 vectors in flight)
constexpr int RB_CV = 32;      // column vectors (8 bf16 each) per CTA = 256 columns
constexpr int RB_RL = RB_THREADS / RB_CV;  // row lanes per CTA

__global__ void __launch_bounds__(RB_THREADS)
    relu_bwd_bias_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y, uint4* __restrict__ dz,
                         float* __restrict__ partial, int64_t B, int64_t NV /* N / 8 */, int rows_per_cta) {
  __shared__ float s_sum[RB_RL][RB_CV * 8 + 8];
  const int tx = threadIdx.x % RB_CV, ty = threadIdx.x / RB_CV;
  const int64_t cv = (int64_t)blockIdx.x * RB_CV + tx;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
  const int64_t r1 = min(r0 + rows_per_cta, B);
  float acc[8];
#pragma unroll
  for (int x = 0; x < 8; ++x) acc[x] = 0.f;
  if (cv < NV) {
    for (int64_t r = r0 + ty; r < r1; r += 4 * RB_RL) {
      uint4 g[4], a[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t rr = r + u * RB_RL;
        const bool ok = rr < r1;
        g[u] = ok ? ld_nc_v4(dy + rr * NV + cv) : make_uint4(0, 0, 0, 0);
        a[u] = ok ? ld_nc_v4(y + rr * NV + cv) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int64_t rr = r + u * RB_RL;
        if (rr >= r1) continue;
        uint32_t gw[4] = {g[u].x, g[u].y, g[u].z, g[u].w};
        const uint32_t aw[4] = {a[u].x, a[u].y, a[u].z, a[u].w};
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          // bf16 pairs: keep a gradient where the activation is > 0 (sign clear and non-zero; NaN counts as kept
          // exactly like torch's `self <= threshold ? 0 : grad`)
          const uint32_t lo_a = aw[h] & 0xffffu, hi_a = aw[h] >> 16;
          const bool keep_lo = !(lo_a == 0u || lo_a == 0x8000u || ((lo_a & 0x8000u) && (lo_a & 0x7fffu) <= 0x7f80u));
          const bool keep_hi = !(hi_a == 0u || hi_a == 0x8000u || ((hi_a & 0x8000u) && (hi_a & 0x7fffu) <= 0x7f80u));
          uint32_t w = gw[h];
          if (!keep_lo) w &= 0xffff0000u;
          if (!keep_hi) w &= 0x0000ffffu;
          gw[h] = w;
          acc[2 * h] += __uint_as_float(w << 16);
          acc[2 * h + 1] += __uint_as_float(w & 0xffff0000u);
        }
        st_v4(dz + rr * NV + cv, make_uint4(gw[0], gw[1], gw[2], gw[3]));
      }
    }
  }
#pragma unroll
  for (int x = 0; x < 8; ++x) s_sum[ty][tx * 8 + x] = acc[x];
  __syncthreads();
  // RB_CV * 8 = 256 columns, one per thread: add the row lanes in order
  const int c = threadIdx.x;
  const int64_t col = (int64_t)blockIdx.x * RB_CV * 8 + c;
  if (col < NV * 8) {
    float s = 0.f;
#pragma unroll
    for (int l = 0; l < RB_RL; ++l) s += s_sum[l][c];
    partial[(int64_t)blockIdx.y * NV * 8 + col] = s;
  }
}

// second stage: db[col] = sum over the chunks.  CTA = 32 columns x BS_LANES chunk lanes: every lane adds chunks lane,
// lane + BS_LANES, ... (8 loads in flight), then the lanes are added in order through shared memory (deterministic).
// 32 lanes: up to 1024 chunks of a narrow layer are 4 round trips to L2 (the grid is only N / 32 blocks, so the
// kernel is as long as its dependent loads: with 8 lanes it took 11 us of a 1.26 ms step, seven times per step).
constexpr int BS_LANES = 32;
__global__ void __launch_bounds__(32 * BS_LANES)
    bias_chunks_sum_kernel(const float* __restrict__ partial, float* __restrict__ db, int64_t chunks, int64_t N) {
  __shared__ float s[BS_LANES][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t col = (int64_t)blockIdx.x * 32 + tx;
  float acc = 0.f;
  if (col < N) {
    for (int64_t c = ty; c < chunks; c += 8 * BS_LANES) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = (c + BS_LANES * u < chunks) ? partial[(c + BS_LANES * u) * N + col] : 0.f;
#pragma unroll
      for (int u = 0; u < 8; ++u) acc += v[u];
    }
  }
  s[ty][tx] = acc;
  __syncthreads();
  if (ty == 0 && col < N) {
    float t = 0.f;
#pragma unroll
    for (int l = 0; l < BS_LANES; ++l) t += s[l][tx];
    db[col] = t;
  }
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_relu_bwd_bias_workspace_bytes(int64_t B, int64_t N) {
  if (B <= 0 || N <= 0) return 0;
  return (size_t)ceil_div(B, RB_ROWS_MIN) * (size_t)N * sizeof(float);  // the most chunks a launch may use
}

extern "C" int b200_relu_bwd_bias(const void* d_dy, const void* d_y, void* d_dz, float* d_db, int64_t B, int64_t N,
                                  int dtype, void* ws, size_t ws_bytes, b200_stream_t stream) {
  B200_REQUIRE(B >= 0 && N >= 0, B200_EINVAL, "relu_bwd_bias: negative size");
  B200_REQUIRE(dtype == B200_BF16, B200_EUNSUPPORTED, "relu_bwd_bias: bf16 only");
  B200_REQUIRE(N % 8 == 0, B200_EUNSUPPORTED, "relu_bwd_bias: N must be a multiple of 8");
  if (N == 0) return B200_OK;
  B200_REQUIRE(d_db, B200_EINVAL, "relu_bwd_bias: null pointer");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (B == 0) {
    cudaMemsetAsync(d_db, 0, (size_t)N * sizeof(float), st);
    return B200_OK;
  }
  B200_REQUIRE(d_dy && d_y && d_dz && ws, B200_EINVAL, "relu_bwd_bias: null pointer");
  B200_REQUIRE(((reinterpret_cast<uintptr_t>(d_dy) | reinterpret_cast<uintptr_t>(d_y) |
                 reinterpret_cast<uintptr_t>(d_dz)) & 15) == 0,
               B200_EINVAL, "relu_bwd_bias: operands must be 16-byte aligned");
  B200_REQUIRE(ws_bytes >= b200_relu_bwd_bias_workspace_bytes(B, N), B200_EWORKSPACE,
               "relu_bwd_bias: workspace too small");
  // rows per CTA: as many as keep >= 8 CTAs per SM in the grid (a 256-column layer has ONE column tile),
  // a multiple of 32, within [32, 128]; more only if the grid's y dimension would overflow
  const int64_t NV = N / 8, col_tiles = ceil_div(NV, RB_CV);
  int64_t rows = B / ceil_div((int64_t)148 * 8, col_tiles) / RB_ROWS_MIN * RB_ROWS_MIN;
  rows = std::max<int64_t>(RB_ROWS_MIN, std::min<int64_t>(RB_ROWS, rows));
  if (ceil_div(B, rows) > 65535) rows = ceil_div(ceil_div(B, (int64_t)65535), RB_ROWS_MIN) * RB_ROWS_MIN;
  const int64_t chunks = ceil_div(B, rows);
  dim3 grid((unsigned)ceil_div(NV, RB_CV), (unsigned)chunks);
  relu_bwd_bias_kernel<<<grid, RB_THREADS, 0, st>>>(static_cast<const uint4*>(d_dy), static_cast<const uint4*>(d_y),
                                                    static_cast<uint4*>(d_dz), static_cast<float*>(ws), B, NV, (int)rows);
  bias_chunks_sum_kernel<<<(unsigned)ceil_div(N, 32), 32 * BS_LANES, 0, st>>>(static_cast<const float*>(ws), d_db, chunks, N);
  B200_CHECK_LAUNCH("relu_bwd_bias");
  return B200_OK;
}
