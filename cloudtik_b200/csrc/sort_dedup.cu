// Stable LSD radix sort of (key32, val32) pairs + run-length de-duplication (sm_100a).
//
// The index stage of the sparse embedding backward: equals torch.sort(stable=True) /
// torch.unique(return_counts=True) on the global row ids, which is what the reference's
// `param.add_(sparse_grad)` (optimizer.step, P/dlrm_s_pytorch.py:1059) relies on.
//
// One sweep over the keys takes the digit histograms of ALL passes (they do not depend on the order of the keys);
// after that every 8-bit digit pass is ONE kernel: a CTA takes the next tile of 4096 keys (ticket counter: tiles
// start in position order), ranks its keys (warp match-any, stable, no atomics on the data path), publishes the
// tile's 256 digit counts and finds how many keys of each digit the earlier tiles hold by decoupled look-back over
// the published counts (thread d walks back for digit d), then scatters.  2 + passes launches instead of 3 per pass.
// Pure integer work, HBM/L2-bound; tiles are sized so that the whole (key,val) working set of a DLRM batch stays in
// the 126 MB L2 between passes.
#include <stdlib.h>

#include <algorithm>
#include <mutex>

#include "common.cuh"
#include "sort_dedup.cuh"

namespace b200 {

// ------------------------------------------------------------------ histogram
__global__ void __launch_bounds__(SORT_THREADS) radix_hist_kernel(const uint32_t* __restrict__ keys,
                                                                  int64_t n, int shift,
                                                                  uint32_t* __restrict__ counts,
                                                                  int nblk) {
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * SORT_CHUNK;
#pragma unroll 4
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int64_t j = base + i * SORT_THREADS + threadIdx.x;
    if (j < n) atomicAdd(&h[(keys[j] >> shift) & 255u], 1u);
  }
  __syncthreads();
  counts[(int64_t)threadIdx.x * nblk + blockIdx.x] = h[threadIdx.x];
}

// ------------------------------------------------------------------ single-block exclusive scan
// In-place exclusive scan of `total` uint32 values; optionally writes the grand total.
__global__ void __launch_bounds__(1024) scan_exclusive_kernel(uint32_t* __restrict__ data,
                                                              int64_t total,
                                                              int32_t* __restrict__ grand_total) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  const int64_t per = (total + 1023) / 1024;
  const int64_t lo = min(total, (int64_t)tid * per), hi = min(total, lo + per);
  uint32_t sum = 0;
  for (int64_t i = lo; i < hi; ++i) sum += data[i];
  // block exclusive scan of per-thread sums
  uint32_t incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = warp_sums[lane];
    uint32_t wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += v;
    }
    warp_sums[lane] = wi - w;  // exclusive
    if (lane == 31 && grand_total) *grand_total = (int32_t)wi;
  }
  __syncthreads();
  uint32_t run = warp_sums[warp] + incl - sum;
  for (int64_t i = lo; i < hi; ++i) {
    const uint32_t v = data[i];
    data[i] = run;
    run += v;
  }
}

// ------------------------------------------------------------------ per-digit row scan
// grid = 256 blocks: block d turns row d of the digit-major table counts[256][nblk] into exclusive
// prefixes (in place) and writes the row total; the scatter kernel scans the 256 totals itself.
__global__ void __launch_bounds__(256) radix_rowscan_kernel(uint32_t* __restrict__ counts, int nblk,
                                                            uint32_t* __restrict__ totals) {
  __shared__ uint32_t wsum[8];
  __shared__ uint32_t carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t* row = counts + (int64_t)blockIdx.x * nblk;
  if (tid == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nblk; base += 256) {
    const int i = base + tid;
    const uint32_t v = i < nblk ? row[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    uint32_t before = carry;
#pragma unroll
    for (int w = 0; w < 8; ++w) before += (w < warp) ? wsum[w] : 0u;
    if (i < nblk) row[i] = before + incl - v;
    __syncthreads();
    if (tid == 255) carry = before + incl;
    __syncthreads();
  }
  if (tid == 0) totals[blockIdx.x] = carry;
}

// ------------------------------------------------------------------ stable scatter
__global__ void __launch_bounds__(SORT_THREADS)
    radix_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                         uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int64_t n,
                         int shift, const uint32_t* __restrict__ counts, int nblk,
                         const uint32_t* __restrict__ totals) {
  constexpr int NW = SORT_THREADS / 32;
  constexpr int WCHUNK = SORT_CHUNK / NW;  // keys per warp, contiguous
  __shared__ uint32_t whist[NW][256];
  __shared__ uint32_t dsum[NW];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < NW * 256; i += SORT_THREADS) (&whist[0][0])[i] = 0;
  // global start of digit `tid` = exclusive scan of the 256 digit totals (every block redoes it)
  uint32_t dbase;
  {
    const uint32_t v = totals[tid];
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    if (lane == 31) dsum[warp] = incl;
    __syncthreads();
    uint32_t before = 0;
#pragma unroll
    for (int w = 0; w < NW; ++w) before += (w < warp) ? dsum[w] : 0u;
    dbase = before + incl - v;
  }
  __syncthreads();

  const int64_t wbase = (int64_t)blockIdx.x * SORT_CHUNK + (int64_t)warp * WCHUNK;
  uint32_t key[SORT_ITEMS], val[SORT_ITEMS], rank[SORT_ITEMS];
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int64_t j = wbase + i * 32 + lane;
    const bool ok = j < n;
    key[i] = ok ? keys_in[j] : 0u;
    val[i] = ok ? vals_in[j] : 0u;
  }
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int64_t j = wbase + i * 32 + lane;
    const bool ok = j < n;
    const uint32_t d = ok ? ((key[i] >> shift) & 255u) : 0xffffffffu;
    const uint32_t m = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(m) - 1;
    uint32_t pre = 0;
    if (ok && lane == leader) {
      pre = whist[warp][d];
      whist[warp][d] = pre + __popc(m);
    }
    pre = __shfl_sync(0xffffffffu, pre, leader);
    rank[i] = pre + __popc(m & lt);
    __syncwarp();
  }
  __syncthreads();
  {  // thread `tid` owns digit `tid`: turn per-warp counts into global start positions
    uint32_t run = dbase + counts[(int64_t)tid * nblk + blockIdx.x];
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const uint32_t c = whist[w][tid];
      whist[w][tid] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int64_t j = wbase + i * 32 + lane;
    if (j < n) {
      const uint32_t pos = whist[warp][(key[i] >> shift) & 255u] + rank[i];
      keys_out[pos] = key[i];
      vals_out[pos] = val[i];
    }
  }
}

// ------------------------------------------------------------------ one-sweep passes
constexpr int MAX_PASSES = 4;
constexpr uint32_t LB_AGG = 1u << 30, LB_INC = 2u << 30, LB_MASK = (1u << 30) - 1u;

// digit histograms of every pass in one sweep: hist_all[pass][256] (zeroed by the caller)
__global__ void __launch_bounds__(256) radix_hist_all_kernel(const uint32_t* __restrict__ keys, int64_t n, int passes,
                                                             uint32_t* __restrict__ hist_all) {
  __shared__ uint32_t h[MAX_PASSES][256];
  for (int i = threadIdx.x; i < MAX_PASSES * 256; i += 256) (&h[0][0])[i] = 0;
  __syncthreads();
  for (int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x; j < n; j += (int64_t)gridDim.x * 256) {
    const uint32_t k = keys[j];
#pragma unroll
    for (int q = 0; q < MAX_PASSES; ++q)
      if (q < passes) atomicAdd(&h[q][(k >> (8 * q)) & 255u], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * 256; i += 256) {
    const uint32_t c = (&h[0][0])[i];
    if (c) atomicAdd(&hist_all[i], c);
  }
}

// One digit pass.  status[tile][256]: bits 31:30 = 0 nothing yet / LB_AGG tile count / LB_INC inclusive prefix over
// all tiles up to this one, bits 29:0 the count.  Zeroed by the caller, like *ticket.
__global__ void __launch_bounds__(SORT_THREADS)
    radix_onesweep_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                          uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int64_t n, int shift,
                          const uint32_t* __restrict__ hist, uint32_t* __restrict__ status,
                          uint32_t* __restrict__ ticket) {
  constexpr int NW = SORT_THREADS / 32;
  constexpr int WCHUNK = SORT_CHUNK / NW;  // keys per warp, contiguous
  __shared__ uint32_t whist[NW][256];
  __shared__ uint32_t dsum[NW];
  __shared__ uint32_t s_tile;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);  // tiles are handed out in the order CTAs start running
  for (int i = tid; i < NW * 256; i += SORT_THREADS) (&whist[0][0])[i] = 0;
  // global start of digit `tid` = exclusive scan of the 256 digit totals of this pass
  uint32_t dbase;
  {
    const uint32_t v = hist[tid];
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += u;
    }
    if (lane == 31) dsum[warp] = incl;
    __syncthreads();
    uint32_t before = 0;
#pragma unroll
    for (int w = 0; w < NW; ++w) before += (w < warp) ? dsum[w] : 0u;
    dbase = before + incl - v;
  }
  __syncthreads();
  const uint32_t tile = s_tile;

  const int64_t wbase = (int64_t)tile * SORT_CHUNK + (int64_t)warp * WCHUNK;
  uint32_t key[SORT_ITEMS], val[SORT_ITEMS], rank[SORT_ITEMS];
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int64_t j = wbase + i * 32 + lane;
    const bool ok = j < n;
    key[i] = ok ? keys_in[j] : 0u;
    val[i] = ok ? vals_in[j] : 0u;
  }
  const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int64_t j = wbase + i * 32 + lane;
    const bool ok = j < n;
    const uint32_t d = ok ? ((key[i] >> shift) & 255u) : 0xffffffffu;
    const uint32_t m = __match_any_sync(0xffffffffu, d);
    const int leader = __ffs(m) - 1;
    uint32_t pre = 0;
    if (ok && lane == leader) {
      pre = whist[warp][d];
      whist[warp][d] = pre + __popc(m);
    }
    pre = __shfl_sync(0xffffffffu, pre, leader);
    rank[i] = pre + __popc(m & lt);
    __syncwarp();
  }
  __syncthreads();
  {  // thread `tid` owns digit `tid`: tile count, look-back, then per-warp start positions
    uint32_t cnt = 0;
#pragma unroll
    for (int w = 0; w < NW; ++w) cnt += whist[w][tid];
    volatile uint32_t* st = status + (size_t)tile * 256 + tid;
    uint32_t excl = 0;
    if (tile == 0) {
      *st = LB_INC | cnt;
    } else {
      *st = LB_AGG | cnt;
      for (int64_t t = (int64_t)tile - 1; t >= 0;) {
        const uint32_t v = *reinterpret_cast<volatile uint32_t*>(status + (size_t)t * 256 + tid);
        if ((v >> 30) == 0) continue;  // not published yet: poll again
        excl += v & LB_MASK;
        if (v & LB_INC) break;
        --t;
      }
      *st = LB_INC | (excl + cnt);
    }
    uint32_t run = dbase + excl;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const uint32_t c = whist[w][tid];
      whist[w][tid] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < SORT_ITEMS; ++i) {
    const int64_t j = wbase + i * 32 + lane;
    if (j < n) {
      const uint32_t pos = whist[warp][(key[i] >> shift) & 255u] + rank[i];
      keys_out[pos] = key[i];
      vals_out[pos] = val[i];
    }
  }
}

// ------------------------------------------------------------------ run-length dedup
__device__ __forceinline__ bool is_head(const uint32_t* __restrict__ k, int64_t j) {
  return j == 0 || k[j] != k[j - 1];
}

__global__ void __launch_bounds__(DEDUP_THREADS) head_count_kernel(const uint32_t* __restrict__ skeys,
                                                                   int64_t n,
                                                                   uint32_t* __restrict__ blk_counts) {
  __shared__ uint32_t wsum[DEDUP_THREADS / 32];
  const int64_t base = (int64_t)blockIdx.x * DEDUP_CHUNK;
  uint32_t c = 0;
#pragma unroll
  for (int i = 0; i < DEDUP_ITEMS; ++i) {
    const int64_t j = base + i * DEDUP_THREADS + threadIdx.x;
    if (j < n) c += is_head(skeys, j) ? 1u : 0u;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t s = 0;
    for (int w = 0; w < DEDUP_THREADS / 32; ++w) s += wsum[w];
    blk_counts[blockIdx.x] = s;
  }
}

__global__ void __launch_bounds__(DEDUP_THREADS)
    compact_kernel(const uint32_t* __restrict__ skeys, int64_t n, const uint32_t* __restrict__ blk_offsets,
                   const int32_t* __restrict__ num_unique, uint32_t* __restrict__ unique,
                   uint32_t* __restrict__ seg, uint32_t* __restrict__ win_run,
                   const uint32_t* __restrict__ svals, uint32_t* __restrict__ first_val) {
  __shared__ uint32_t wsum[DEDUP_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t base = (int64_t)blockIdx.x * DEDUP_CHUNK;
  uint32_t run = blk_offsets[blockIdx.x];
  if (blockIdx.x == 0 && tid == 0) seg[*num_unique] = (uint32_t)n;
#pragma unroll 1
  for (int i = 0; i < DEDUP_ITEMS; ++i) {
    const int64_t j = base + i * DEDUP_THREADS + tid;
    const bool h = j < n && is_head(skeys, j);
    const uint32_t b = __ballot_sync(0xffffffffu, h);
    if (lane == 0) wsum[warp] = __popc(b);
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < DEDUP_THREADS / 32; ++w) {
      const uint32_t c = wsum[w];
      before += (w < warp) ? c : 0u;
      total += c;
    }
    if (h) {
      const uint32_t pos = run + before + __popc(b & ((1u << lane) - 1u));
      unique[pos] = skeys[j];
      seg[pos] = (uint32_t)j;
      if (first_val != nullptr) first_val[pos] = svals[j];  // payload of the run's first element
    }
    if (win_run != nullptr && j < n && (tid % HOT_CH) == 0)  // run containing this window start
      win_run[j / HOT_CH] = run + before + __popc(b & (0xffffffffu >> (31 - lane))) - 1u;
    run += total;
    __syncthreads();
  }
}

__global__ void empty_result_kernel(uint32_t* seg, int32_t* num_unique) {
  seg[0] = 0;
  *num_unique = 0;
}

// ------------------------------------------------------------------ host drivers
size_t sort_ws_bytes(int64_t n) {
  const int64_t nblk = ceil_div(n > 0 ? n : 1, SORT_CHUNK);
  const int64_t nblk2 = ceil_div(n > 0 ? n : 1, DEDUP_CHUNK);
  size_t b = 0;
  b += 4 * align_up((size_t)n * 4, 256);              // keys A/B, vals A/B
  b += align_up((size_t)256 * nblk * 4, 256);         // (kept: the three-kernel pass of B200_SORT_ONESWEEP=0)
  b += align_up((size_t)(nblk2 + 1) * 4, 256);        // dedup block counts
  b += 1024;                                          // 256 digit totals
  b += 256;
  // one-sweep state, zeroed by ONE memset per sort: tickets [4] | hist_all [4][256] | status [4][nblk][256]
  b += 256 + (size_t)MAX_PASSES * 1024 + (size_t)MAX_PASSES * nblk * 1024;
  return b;
}

int sort_pairs(const uint32_t* keys, const uint32_t* vals, int64_t n, uint64_t key_bound,
               uint32_t* out_keys, uint32_t* out_vals, void* ws, cudaStream_t st) {
  const int nblk = (int)ceil_div(n, SORT_CHUNK);
  char* w = static_cast<char*>(ws);
  const size_t stride = align_up((size_t)n * 4, 256);
  uint32_t* kbuf[2] = {reinterpret_cast<uint32_t*>(w), reinterpret_cast<uint32_t*>(w + stride)};
  uint32_t* vbuf[2] = {reinterpret_cast<uint32_t*>(w + 2 * stride),
                       reinterpret_cast<uint32_t*>(w + 3 * stride)};
  uint32_t* counts = reinterpret_cast<uint32_t*>(w + 4 * stride);
  // 256 digit totals: own 1 KB slot behind the dedup block counts (see sort_ws_bytes)
  uint32_t* totals = reinterpret_cast<uint32_t*>(w + 4 * stride + align_up((size_t)256 * nblk * 4, 256) +
                                                 align_up((size_t)(ceil_div(n, DEDUP_CHUNK) + 1) * 4, 256));
  int bits = 1;
  while (bits < 32 && (key_bound - 1) >> bits) ++bits;
  const int passes = (bits + 7) / 8;
  const uint32_t* ksrc = keys;
  const uint32_t* vsrc = vals;
  // Measured on B200 (tools/sort_ab.sh, 8-bit digits, 4 passes): 852 k keys 0.130 ms one-sweep vs 0.140 ms with three
  // kernels per pass; 6.8 M keys 0.631 vs 0.593 ms (about 46 G keys/s per pass either way: the ranking and the 4-byte
  // scatter bound a pass, not the launches; the look-back traffic costs a little at large n).  So the one-sweep
  // passes serve the step-sized sorts, the three-kernel pass the large ones.  B200_SORT_ONESWEEP=0 / 1 forces one.
  static const int onesweep_env = []() {
    const char* e = getenv("B200_SORT_ONESWEEP");
    return e == nullptr ? -1 : atoi(e);
  }();
  const bool onesweep = onesweep_env < 0 ? n <= ((int64_t)2 << 20) : onesweep_env != 0;
  if (onesweep) {
    char* os = w + 4 * stride + align_up((size_t)256 * nblk * 4, 256) +
               align_up((size_t)(ceil_div(n, DEDUP_CHUNK) + 1) * 4, 256) + 1024 + 256;
    uint32_t* tickets = reinterpret_cast<uint32_t*>(os);
    uint32_t* hist_all = reinterpret_cast<uint32_t*>(os + 256);
    uint32_t* status = reinterpret_cast<uint32_t*>(os + 256 + MAX_PASSES * 1024);
    cudaMemsetAsync(os, 0, 256 + (size_t)MAX_PASSES * 1024 + (size_t)passes * nblk * 1024, st);
    const int hgrid = (int)std::min<int64_t>(ceil_div(n, 256 * 8), 148 * 8);
    radix_hist_all_kernel<<<hgrid, 256, 0, st>>>(keys, n, passes, hist_all);
    for (int pass = 0; pass < passes; ++pass) {
      uint32_t* kdst = (pass == passes - 1) ? out_keys : kbuf[pass & 1];
      uint32_t* vdst = (pass == passes - 1) ? out_vals : vbuf[pass & 1];
      radix_onesweep_kernel<<<nblk, SORT_THREADS, 0, st>>>(ksrc, vsrc, kdst, vdst, n, pass * 8, hist_all + pass * 256,
                                                           status + (size_t)pass * nblk * 256, tickets + pass);
      ksrc = kdst;
      vsrc = vdst;
    }
    B200_CHECK_LAUNCH("radix sort (one-sweep)");
    return B200_OK;
  }
  for (int pass = 0; pass < passes; ++pass) {
    uint32_t* kdst = (pass == passes - 1) ? out_keys : kbuf[pass & 1];
    uint32_t* vdst = (pass == passes - 1) ? out_vals : vbuf[pass & 1];
    const int shift = pass * 8;
    radix_hist_kernel<<<nblk, SORT_THREADS, 0, st>>>(ksrc, n, shift, counts, nblk);
    radix_rowscan_kernel<<<256, 256, 0, st>>>(counts, nblk, totals);
    radix_scatter_kernel<<<nblk, SORT_THREADS, 0, st>>>(ksrc, vsrc, kdst, vdst, n, shift, counts, nblk, totals);
    ksrc = kdst;
    vsrc = vdst;
  }
  B200_CHECK_LAUNCH("radix sort");
  return B200_OK;
}

int dedup_runs(const uint32_t* skeys, int64_t n, uint32_t* unique, uint32_t* seg, int32_t* num_unique,
               void* ws, cudaStream_t st, uint32_t* win_run, const uint32_t* svals, uint32_t* first_val) {
  char* w = static_cast<char*>(ws);
  const size_t stride = align_up((size_t)n * 4, 256);
  const int64_t nblk = ceil_div(n, SORT_CHUNK);
  uint32_t* blk = reinterpret_cast<uint32_t*>(w + 4 * stride + align_up((size_t)256 * nblk * 4, 256));
  const int nblk2 = (int)ceil_div(n, DEDUP_CHUNK);
  head_count_kernel<<<nblk2, DEDUP_THREADS, 0, st>>>(skeys, n, blk);
  scan_exclusive_kernel<<<1, 1024, 0, st>>>(blk, nblk2, num_unique);
  compact_kernel<<<nblk2, DEDUP_THREADS, 0, st>>>(skeys, n, blk, num_unique, unique, seg, win_run, svals, first_val);
  B200_CHECK_LAUNCH("dedup");
  return B200_OK;
}

int empty_result(uint32_t* seg, int32_t* num_unique, cudaStream_t st) {
  empty_result_kernel<<<1, 1, 0, st>>>(seg, num_unique);
  B200_CHECK_LAUNCH("empty_result");
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_sort_workspace_bytes(int64_t n) { return sort_ws_bytes(n < 0 ? 0 : n); }

extern "C" int b200_sort_dedup(const uint32_t* d_keys, const uint32_t* d_vals, int64_t n,
                               uint64_t key_bound, uint32_t* d_sorted_keys, uint32_t* d_sorted_vals,
                               uint32_t* d_unique, uint32_t* d_seg_offsets, int32_t* d_num_unique,
                               void* ws, size_t ws_bytes, b200_stream_t stream) {
  B200_REQUIRE(n >= 0, B200_EINVAL, "sort_dedup: negative n");
  B200_REQUIRE(n < ((int64_t)1 << 31), B200_EUNSUPPORTED, "sort_dedup: n >= 2^31");
  B200_REQUIRE(key_bound >= 1 && key_bound <= ((uint64_t)1 << 32), B200_EUNSUPPORTED,
               "sort_dedup: key_bound must be in [1, 2^32]");
  B200_REQUIRE(d_seg_offsets && d_num_unique, B200_EINVAL, "sort_dedup: null output");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (n == 0) return empty_result(d_seg_offsets, d_num_unique, st);
  B200_REQUIRE(d_keys && d_vals && d_sorted_keys && d_sorted_vals && d_unique && ws, B200_EINVAL,
               "sort_dedup: null pointer");
  B200_REQUIRE(ws_bytes >= sort_ws_bytes(n), B200_EWORKSPACE, "sort_dedup: workspace too small");
  int rc = sort_pairs(d_keys, d_vals, n, key_bound, d_sorted_keys, d_sorted_vals, ws, st);
  if (rc) return rc;
  return dedup_runs(d_sorted_keys, n, d_unique, d_seg_offsets, d_num_unique, ws, st);
}

namespace b200 {
// ------------------------------------------------------------------ entry point that allocates nothing
//
// Who this is for: callers that are handed device pointers but have no way to obtain device
// memory of their own.  Typical cases: a trainer whose framework has already claimed all of HBM
// for its caching allocator (a DLRM with 96 GB of tables leaves a bare cudaMalloc nothing to
// find, and the framework's allocator is not reachable from a plain C caller); a latency-critical
// loop that must not stall (cudaMalloc / cudaFree synchronise the whole device); and thin
// language bindings (cgo, JNI, ctypes callbacks) that cannot own the lifetime of a device buffer.
//
// The sort / dedup scratch therefore lives in the library image itself: a file-scope device array
// of fixed size, present once on every GPU that loads libb200dlrm.so and held until the process
// exits.  Consequences the caller must accept:
//   * n is limited to B200_SORT_NOALLOC_MAX_KEYS keys (larger n is refused with B200_EUNSUPPORTED);
//   * all calls on a device share that one array, so two calls that may overlap on the GPU
//     (different streams, different host threads) race -- serialise them;
//   * the array keeps whatever the previous call left in it.  Nothing in it has to start at zero:
//     every word a call reads (digit histograms, dedup block counts, ping-pong key/value buffers)
//     is written earlier in the same call, so no clearing pass is needed.
constexpr int64_t NOALLOC_MAX_KEYS = B200_SORT_NOALLOC_MAX_KEYS;
constexpr size_t cx_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
constexpr size_t cx_ceil_div(size_t a, size_t b) { return (a + b - 1) / b; }
// same carve-up as sort_ws_bytes(), evaluated at compile time for the maximum n
constexpr size_t NOALLOC_SCRATCH_BYTES =
    4 * cx_align_up((size_t)NOALLOC_MAX_KEYS * 4, 256) +
    cx_align_up((size_t)256 * cx_ceil_div(NOALLOC_MAX_KEYS, SORT_CHUNK) * 4, 256) +
    cx_align_up((cx_ceil_div(NOALLOC_MAX_KEYS, DEDUP_CHUNK) + 1) * 4, 256) + 1024 + 256 +
    256 + (size_t)MAX_PASSES * 1024 + (size_t)MAX_PASSES * cx_ceil_div(NOALLOC_MAX_KEYS, SORT_CHUNK) * 1024;

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is only a handle for the symbol, not a device address: the
// address differs per device and has to be asked for with that device current.
static std::mutex g_noalloc_mu;
static void* g_noalloc_dev_ptr[64] = {};

static int noalloc_scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  B200_REQUIRE(e == cudaSuccess && dev >= 0 && dev < 64, B200_ELAUNCH, "sort_dedup_noalloc: cudaGetDevice failed: %s",
               cudaGetErrorString(e));
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_dev_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    B200_REQUIRE(e == cudaSuccess && ptr != nullptr, B200_ELAUNCH,
                 "sort_dedup_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    g_noalloc_dev_ptr[dev] = ptr;
  }
  *out = g_noalloc_dev_ptr[dev];
  return B200_OK;
}

}  // namespace b200

extern "C" int b200_sort_dedup_noalloc(const uint32_t* d_keys, const uint32_t* d_vals, int64_t n,
                                       uint64_t key_bound, uint32_t* d_sorted_keys, uint32_t* d_sorted_vals,
                                       uint32_t* d_unique, uint32_t* d_seg_offsets, int32_t* d_num_unique,
                                       b200_stream_t stream) {
  B200_REQUIRE(n >= 0, B200_EINVAL, "sort_dedup_noalloc: negative n");
  B200_REQUIRE(n <= NOALLOC_MAX_KEYS, B200_EUNSUPPORTED,
               "sort_dedup_noalloc: n = %lld exceeds the built-in scratch (max %lld keys); use b200_sort_dedup "
               "with a caller-provided workspace",
               (long long)n, (long long)NOALLOC_MAX_KEYS);
  B200_REQUIRE(key_bound >= 1 && key_bound <= ((uint64_t)1 << 32), B200_EUNSUPPORTED,
               "sort_dedup_noalloc: key_bound must be in [1, 2^32]");
  B200_REQUIRE(d_seg_offsets && d_num_unique, B200_EINVAL, "sort_dedup_noalloc: null output");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (n == 0) return empty_result(d_seg_offsets, d_num_unique, st);
  B200_REQUIRE(d_keys && d_vals && d_sorted_keys && d_sorted_vals && d_unique, B200_EINVAL,
               "sort_dedup_noalloc: null pointer");
  void* scratch = nullptr;
  int rc = noalloc_scratch_for_current_device(&scratch);
  if (rc) return rc;
  rc = sort_pairs(d_keys, d_vals, n, key_bound, d_sorted_keys, d_sorted_vals, scratch, st);
  if (rc) return rc;
  return dedup_runs(d_sorted_keys, n, d_unique, d_seg_offsets, d_num_unique, scratch, st);
}
