// Row-wise sharded tables: keep, of every bag, the indices whose row this rank owns (sm_100a).
//
// BASELINE configs[4] (not in the reference, which only shards table-wise): rank r of W owns the rows
// {i : i mod W == r} of a hot table and stores them densely as local row i / W.  Every rank sees the bags of the
// GLOBAL batch and filters them down to its rows.  The output has a FIXED capacity and its fill level stays on the
// device (out_offsets[n_bags]), so the step around it has no data-dependent shape and no host synchronisation:
// it can be captured in a CUDA graph.  Three launches: per-bag counts + per-block sums, scan of the block sums,
// in-block scan + compaction.
#include "common.cuh"

namespace b200 {

constexpr int RW_THREADS = 256;  // bags per block

__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_sums, int& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += u;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  int before = 0;
  total = 0;
#pragma unroll
  for (int w = 0; w < RW_THREADS / 32; ++w) {
    const int c = warp_sums[w];
    before += (w < warp) ? c : 0;
    total += c;
  }
  return before + incl - v;
}

__global__ void __launch_bounds__(RW_THREADS)
    rowwise_count_kernel(const int64_t* __restrict__ indices, const int64_t* __restrict__ offsets, int64_t n_bags,
                         int64_t world, int64_t rank, int32_t* __restrict__ cnt, int64_t* __restrict__ blk_sums) {
  __shared__ int warp_sums[RW_THREADS / 32];
  const int64_t g = (int64_t)blockIdx.x * RW_THREADS + threadIdx.x;
  int c = 0;
  if (g < n_bags) {
    const int64_t s = offsets[g], e = offsets[g + 1];
    for (int64_t j = s; j < e; ++j) c += (indices[j] % world == rank) ? 1 : 0;
    cnt[g] = c;
  }
  int total;
  block_exclusive_scan(c, warp_sums, total);
  if (threadIdx.x == 0) blk_sums[blockIdx.x] = total;
}

// single block: exclusive scan of the block sums (n_blk <= a few thousand), total behind the last entry
__global__ void __launch_bounds__(1024) rowwise_scan_kernel(int64_t* __restrict__ blk_sums, int64_t n_blk) {
  __shared__ int64_t wsum[32];
  __shared__ int64_t carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < n_blk; base += 1024) {
    const int64_t i = base + threadIdx.x;
    const int64_t v = i < n_blk ? blk_sums[i] : 0;
    int64_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int64_t u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    int64_t before = carry;
    for (int w = 0; w < warp; ++w) before += wsum[w];
    if (i < n_blk) blk_sums[i] = before + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) blk_sums[n_blk] = carry;
}

__global__ void __launch_bounds__(RW_THREADS)
    rowwise_compact_kernel(const int64_t* __restrict__ indices, const int64_t* __restrict__ offsets, int64_t n_bags,
                           int64_t world, int64_t rank, const int32_t* __restrict__ cnt,
                           const int64_t* __restrict__ blk_sums, int64_t n_blk, int64_t* __restrict__ out_indices,
                           int64_t capacity, int64_t* __restrict__ out_offsets, int32_t* error) {
  __shared__ int warp_sums[RW_THREADS / 32];
  const int64_t g = (int64_t)blockIdx.x * RW_THREADS + threadIdx.x;
  const int c = g < n_bags ? cnt[g] : 0;
  int total;
  const int64_t pos = blk_sums[blockIdx.x] + block_exclusive_scan(c, warp_sums, total);
  if (g < n_bags) {
    out_offsets[g] = min(pos, capacity);
    const int64_t s = offsets[g], e = offsets[g + 1];
    int64_t o = pos;
    for (int64_t j = s; j < e; ++j) {
      const int64_t idx = indices[j];
      if (idx % world == rank) {
        if (o < capacity) out_indices[o] = idx / world;
        ++o;
      }
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const int64_t all = blk_sums[n_blk];
    out_offsets[n_bags] = min(all, capacity);
    if (all > capacity && error) *error = 2;  // more owned indices than the fixed capacity: bags were truncated
  }
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_rowwise_filter_workspace_bytes(int64_t n_bags) {
  const int64_t nb = n_bags < 0 ? 0 : n_bags;
  return align_up((size_t)nb * 4, 256) + align_up((size_t)(ceil_div(nb, RW_THREADS) + 1) * 8, 256);
}

extern "C" int b200_rowwise_filter(const int64_t* d_indices, const int64_t* d_offsets, int64_t n_bags, int64_t world,
                                   int64_t rank, int64_t* d_out_indices, int64_t out_capacity,
                                   int64_t* d_out_offsets, int32_t* d_error, void* ws, size_t ws_bytes,
                                   b200_stream_t stream) {
  B200_REQUIRE(n_bags >= 0 && world >= 1 && rank >= 0 && rank < world && out_capacity >= 0, B200_EINVAL,
               "rowwise_filter: bad argument");
  B200_REQUIRE(d_out_offsets, B200_EINVAL, "rowwise_filter: null output");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (n_bags == 0) {
    cudaMemsetAsync(d_out_offsets, 0, sizeof(int64_t), st);
    return B200_OK;
  }
  B200_REQUIRE(d_indices && d_offsets && (d_out_indices || out_capacity == 0) && ws, B200_EINVAL,
               "rowwise_filter: null pointer");
  B200_REQUIRE(ws_bytes >= b200_rowwise_filter_workspace_bytes(n_bags), B200_EWORKSPACE,
               "rowwise_filter: workspace too small");
  const int64_t n_blk = ceil_div(n_bags, RW_THREADS);
  B200_REQUIRE(n_blk < ((int64_t)1 << 31), B200_EUNSUPPORTED, "rowwise_filter: too many bags");
  int32_t* cnt = static_cast<int32_t*>(ws);
  int64_t* blk = reinterpret_cast<int64_t*>(static_cast<char*>(ws) + align_up((size_t)n_bags * 4, 256));
  rowwise_count_kernel<<<(unsigned)n_blk, RW_THREADS, 0, st>>>(d_indices, d_offsets, n_bags, world, rank, cnt, blk);
  rowwise_scan_kernel<<<1, 1024, 0, st>>>(blk, n_blk);
  rowwise_compact_kernel<<<(unsigned)n_blk, RW_THREADS, 0, st>>>(d_indices, d_offsets, n_bags, world, rank, cnt, blk,
                                                                 n_blk, d_out_indices, out_capacity, d_out_offsets,
                                                                 d_error);
  B200_CHECK_LAUNCH("rowwise_filter");
  return B200_OK;
}
