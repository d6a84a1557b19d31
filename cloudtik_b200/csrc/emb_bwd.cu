// EmbeddingBag backward fused with the SGD / Adagrad row update (sm_100a).
//
// Replaces autograd's sparse EmbeddingBag backward + torch.optim.SGD.step on the 26 sparse
// gradients (reference P/dlrm_s_pytorch.py:1055,1059) -- see include/b200dlrm.h.
//
//   1. build_keys : key[j] = global row id (table-local index + row base), val[j] = global
//                   bag id t*B+b   (coalesced; bag of j found from the CSR offsets)
//   2. sort_pairs / dedup_runs (sort_dedup.cu): stable radix sort, unique rows, run starts
//   3. update     : the sorted positions are cut into windows of HOT_CH; one sub-warp group per
//                   window sums the gradient rows run by run (fp32, in bag order) and gives every
//                   row ONE read-modify-write (weight row, Adagrad state row, bf16 trail); runs that
//                   cross windows are joined in shared memory, runs that cross CTAs by the CTA that
//                   draws the last ticket on the run's counter (same kernel, no second launch).
//                   The [nnz, D] COO gradient is never materialised.
#include <stdlib.h>

#include <algorithm>

#include "common.cuh"
#include "sort_dedup.cuh"

namespace b200 {

constexpr int KEYS_THREADS = 256;
constexpr int UPD_THREADS = 256;

// Runs longer than HOT_CH positions ("hot" rows: tiny tables, Zipf heads) are never summed by one
// group serially: every window pre-reduces the (at most two) runs that cross its borders into fp32
// pieces and the pieces of a row are added in window order.  Deterministic, no data-path atomics
// (the fp32 SGD variant issues ONE vector reduction per 16 bytes of a row, from one group).
constexpr int PERSIST_DEFAULT = 0;  // window_update_kernel launch mode when B200_UPD_PERSIST is unset (see launch_update)
constexpr int OPT_STORE = 2;  // internal "optimizer": write the per-row gradient sums to values[run index]
constexpr int MIN_NG = UPD_THREADS / 32;  // fewest windows per block (one warp per window)


struct BwdLayout {
  uint32_t *keys, *vals, *skeys, *svals, *unique, *seg, *win_run, *first_val;
  int32_t* num_unique;
  int64_t* n_valid;  // sorted positions that hold real keys (= nnz, or fewer after b200_emb_bwd_prepare_partial)
  int32_t* blk_cnt;  // [nblk]: tickets of the cross-block join, per head block (zero between calls)
  float* partial;   // [2 * nblk][D], nblk = blocks of the window kernel (<= ceil(nwin / MIN_NG))
  void* sort_ws;
  size_t total;
};

static BwdLayout bwd_layout(void* ws, int64_t n, int64_t D) {
  BwdLayout L;
  char* w = static_cast<char*>(ws);
  const size_t s = align_up((size_t)n * 4, 256);
  size_t o = 0;
  L.keys = reinterpret_cast<uint32_t*>(w + o), o += s;
  L.vals = reinterpret_cast<uint32_t*>(w + o), o += s;
  L.skeys = reinterpret_cast<uint32_t*>(w + o), o += s;
  L.svals = reinterpret_cast<uint32_t*>(w + o), o += s;
  L.unique = reinterpret_cast<uint32_t*>(w + o), o += s;
  L.first_val = reinterpret_cast<uint32_t*>(w + o), o += s;
  L.seg = reinterpret_cast<uint32_t*>(w + o), o += align_up((size_t)(n + 1) * 4, 256);
  L.num_unique = reinterpret_cast<int32_t*>(w + o);
  L.n_valid = reinterpret_cast<int64_t*>(w + o + 64);
  o += 256;
  L.blk_cnt = reinterpret_cast<int32_t*>(w + o), o += align_up((size_t)(ceil_div(ceil_div(n, HOT_CH), MIN_NG) + 3) * 4, 256);  // + chunk scheduler (2 ints)
  L.win_run = reinterpret_cast<uint32_t*>(w + o), o += align_up((size_t)(ceil_div(n, HOT_CH) + 1) * 4, 256);
  L.partial = reinterpret_cast<float*>(w + o),
  o += align_up((size_t)(2 * ceil_div(ceil_div(n, HOT_CH), MIN_NG)) * D * 4, 256);
  L.sort_ws = w + o, o += sort_ws_bytes(n);
  L.total = o;
  return L;
}

// ------------------------------------------------------------------ 1. keys
// grid = (ceil(B/256), T); one warp per 32 consecutive bags of one table.
// Positions [*d_valid, n) of a partially filled key array (b200_emb_bwd_prepare_partial) get the sentinel key
// `sentinel` (one past the last real row id): they sort behind every real key and the update kernels stop at
// *n_valid.  Also publishes n_valid in the workspace.
__global__ void __launch_bounds__(256)
    keys_tail_kernel(uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, const int64_t* __restrict__ d_valid,
                     int64_t n, uint32_t sentinel, int64_t* __restrict__ n_valid_out, int32_t* __restrict__ blk_cnt,
                     int64_t n_blk_cnt) {
  const int64_t nv = d_valid ? min(max(*d_valid, (int64_t)0), n) : n;
  if (blockIdx.x == 0 && threadIdx.x == 0) *n_valid_out = nv;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_blk_cnt; j += (int64_t)gridDim.x * blockDim.x)
    blk_cnt[j] = 0;  // tickets of the cross-block join (the join also leaves them at zero)
  for (int64_t j = nv + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (int64_t)gridDim.x * blockDim.x) {
    keys[j] = sentinel;
    vals[j] = 0u;
  }
}

__global__ void __launch_bounds__(KEYS_THREADS)
    build_keys_kernel(const int64_t* __restrict__ indices, const int64_t* __restrict__ offsets,
                      const int64_t* __restrict__ rows, const int64_t* __restrict__ row_base, int64_t B,
                      uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, int32_t* error) {
  const int lane = threadIdx.x & 31;
  const int t = blockIdx.y;
  const int64_t b = (int64_t)blockIdx.x * KEYS_THREADS + threadIdx.x;
  const int64_t b_warp = b - lane;
  if (b_warp >= B) return;
  const int64_t g = (int64_t)t * B + b;
  const bool valid = b < B;
  const int64_t s = valid ? offsets[g] : 0;
  const int64_t e = valid ? offsets[g + 1] : 0;
  const uint64_t nrows = (uint64_t)rows[t];
  const int64_t base = row_base[t];
  const bool all_one = __all_sync(0xffffffffu, !valid || (e - s) == 1);
  if (all_one) {
    if (valid) {
      int64_t idx = indices[s];
      if ((uint64_t)idx >= nrows) {
        if (error) *error = 1;
        idx = 0;
      }
      keys[s] = (uint32_t)(idx + base);
      vals[s] = (uint32_t)g;
    }
    return;
  }
  for (int q = 0; q < 32; ++q) {
    const int64_t sq = __shfl_sync(0xffffffffu, s, q);
    const int64_t eq = __shfl_sync(0xffffffffu, e, q);
    const uint32_t gq = (uint32_t)(g - lane + q);
    for (int64_t j = sq + lane; j < eq; j += 32) {
      int64_t idx = indices[j];
      if ((uint64_t)idx >= nrows) {
        if (error) *error = 1;
        idx = 0;
      }
      keys[j] = (uint32_t)(idx + base);
      vals[j] = gq;
    }
  }
}

// ------------------------------------------------------------------ 3. update
struct UpdParams {
  void* const* weights;
  float* const* state;
  // bf16 tables with a 16-bit "trail": weight = upper half, trail = lower half of the fp32 master value
  // (the reference's IPEX split-SGD storage, P/dlrm_s_pytorch.py:920-933); same [rows, D] layout
  uint16_t* const* trail;
  const int64_t* row_base;
  const uint32_t *unique, *seg, *svals;
  const int32_t* num_unique;
  const void* grad;
  int64_t grad_stride_table, grad_stride_bag;
  const float* d_lr;
  float lr, eps;
  uint32_t B;
  int D, V;
  const uint32_t* skeys;
  const uint32_t* win_run;
  const uint32_t* first_val;
  float* values;  // OPT_STORE: [n, D] fp32 coalesced gradient rows, row u = sum of run u
  // OPT_STORE with by_key: values is a DENSE [total_rows, D] gradient, the sum of run u goes to row unique[u]
  // (rows no bag touched are left as the caller initialised them)
  int by_key;
  // sharded gradients (fused mirror all-to-all over peer memory): the gradient row of bag b of local
  // table t lives on rank r = b / lN at grad_ptrs[r] + ((b - r*lN) * F + feat0 + t) * D
  const void* const* grad_ptrs;
  int64_t lN, F, feat0;
  int T;
  float* partial;
  int64_t n;            // capacity of the sorted arrays (sizes the grid)
  const int64_t* d_n;   // real keys among them, read at kernel time (the sort puts them first)
  int32_t* blk_cnt;
  // persistent launch: the grid is one wave of blocks, each claims further 256-key chunks from sched[0]
  // (sched[1] counts finished blocks; the last one leaves both at zero)
  int32_t* sched;
  int persist;
  int kflush;  // persistent launch: chunks with border runs a block collects before it takes their tickets (<= 16)
};

// load VE consecutive gradient elements (GT) as fp32
template <typename GT, int VE>
__device__ __forceinline__ void load_grad(const GT* p, float* out);
template <>
__device__ __forceinline__ void load_grad<float, 4>(const float* p, float* out) {
  Vec16<float>::unpack(out, ld_nc_v4(p));
}
template <>
__device__ __forceinline__ void load_grad<float, 8>(const float* p, float* out) {
  Vec16<float>::unpack(out, ld_nc_v4(p));
  Vec16<float>::unpack(out + 4, ld_nc_v4(p + 4));
}
template <>
__device__ __forceinline__ void load_grad<__nv_bfloat16, 8>(const __nv_bfloat16* p, float* out) {
  Vec16<__nv_bfloat16>::unpack(out, ld_nc_v4(p));
}
template <>
__device__ __forceinline__ void load_grad<__nv_bfloat16, 4>(const __nv_bfloat16* p, float* out) {
  const uint2 v = ld_nc_v2(p);
  out[0] = bf16lo(v.x);
  out[1] = bf16hi(v.x);
  out[2] = bf16lo(v.y);
  out[3] = bf16hi(v.y);
}

// raw (still packed) gradient vector of VE elements: kept packed while in flight, unpacked when added
template <typename GT, int VE>
struct GradVec;
template <>
struct GradVec<float, 4> {
  uint4 v;
  __device__ __forceinline__ void load(const float* p) { v = ld_nc_v4(p); }
  __device__ __forceinline__ void zero() { v = make_uint4(0, 0, 0, 0); }
  __device__ __forceinline__ void pin() { keep(v); }
  __device__ __forceinline__ void add_to(float* a) const { Vec16<float>::accum(a, v); }
};
template <>
struct GradVec<float, 8> {
  uint4 v0, v1;
  __device__ __forceinline__ void load(const float* p) { v0 = ld_nc_v4(p); v1 = ld_nc_v4(p + 4); }
  __device__ __forceinline__ void zero() { v0 = v1 = make_uint4(0, 0, 0, 0); }
  __device__ __forceinline__ void pin() { keep(v0); keep(v1); }
  __device__ __forceinline__ void add_to(float* a) const { Vec16<float>::accum(a, v0); Vec16<float>::accum(a + 4, v1); }
};
template <>
struct GradVec<__nv_bfloat16, 8> {
  uint4 v;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { v = ld_nc_v4(p); }
  __device__ __forceinline__ void zero() { v = make_uint4(0, 0, 0, 0); }
  __device__ __forceinline__ void pin() { keep(v); }
  __device__ __forceinline__ void add_to(float* a) const { Vec16<__nv_bfloat16>::accum(a, v); }
};
template <>
struct GradVec<__nv_bfloat16, 4> {
  uint2 v;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { v = ld_nc_v2(p); }
  __device__ __forceinline__ void zero() { v = make_uint2(0, 0); }
  __device__ __forceinline__ void pin() { keep(v); }
  __device__ __forceinline__ void add_to(float* a) const {
    a[0] += bf16lo(v.x);
    a[1] += bf16hi(v.x);
    a[2] += bf16lo(v.y);
    a[3] += bf16hi(v.y);
  }
};

// ---- 3a. window kernel: uniform work per group, independent of the duplicate structure ----
// One group of G lanes owns a window of HOT_CH consecutive SORTED positions; a block owns NG
// consecutive windows.  A group stages its window's keys / gradient-row addresses in shared
// memory (coalesced), then streams the gradient rows in batches (BS rows in flight per lane,
// plus the weight rows of the runs that end in the batch) and sums run by run, in position
// (= bag) order:
//   * a run that lies entirely inside the window is finished right there: ONE read-modify-write
//     of its weight row (and Adagrad state row);
//   * a run that crosses a window border leaves an fp32 piece in SHARED memory: slot 0 if it
//     began at or before the window's first position, slot 1 if it began inside.  After a block
//     barrier the group holding the head of such a run adds the pieces of the following windows
//     in window order and finishes the row -- no trip through global memory, no extra launch;
//   * only a run that crosses a BLOCK border leaves a partial row in global memory (same slot
//     rule, per block); the cross-block join at the end of the kernel adds those in block order and updates the row.
// Every weight row is written by exactly one group: deterministic, no atomics.
template <int OPT, typename WT>
__device__ __forceinline__ void apply_row(char* wp, float* sp, const float* acc, const float* w_in, float lr,
                                          float eps, char* tp = nullptr, uint4 tv = make_uint4(0, 0, 0, 0),
                                          const uint4* sv = nullptr) {
  constexpr int VE = Vec16<WT>::VE;
  float w[VE];
#pragma unroll
  for (int x = 0; x < VE; ++x) w[x] = w_in[x];
  if constexpr (sizeof(WT) == 2) {
    if (tp != nullptr) {  // exact fp32 master = (bf16 bits << 16) | trail bits
      const uint32_t t4[4] = {tv.x, tv.y, tv.z, tv.w};
#pragma unroll
      for (int x = 0; x < VE; ++x)
        w[x] = __uint_as_float(__float_as_uint(w[x]) | ((x & 1) ? (t4[x >> 1] >> 16) : (t4[x >> 1] & 0xffffu)));
    }
  }
  if (OPT == OPT_STORE) {
#pragma unroll
    for (int x = 0; x < VE; ++x) w[x] = acc[x];
  } else if (OPT == B200_OPT_ADAGRAD) {
#pragma unroll
    for (int h = 0; h < VE / 4; ++h) {
      float st[4];
      Vec16<float>::unpack(st, sv != nullptr ? sv[h] : ld_v4(sp + 4 * h));  // sv: state vectors fetched with the batch
#pragma unroll
      for (int x = 0; x < 4; ++x) {
        const float g = acc[4 * h + x];
        st[x] += g * g;
        w[4 * h + x] -= lr * g / (sqrtf(st[x]) + eps);
      }
      store_vec<float, 4, false>(sp + 4 * h, st);
    }
  } else {
#pragma unroll
    for (int x = 0; x < VE; ++x) w[x] -= lr * acc[x];
  }
  if constexpr (sizeof(WT) == 2) {
    if (tp != nullptr) {  // split back: upper half -> bf16 table (truncation keeps hi|lo exact), lower -> trail
      uint32_t hi[4], lo[4];
#pragma unroll
      for (int x = 0; x < 4; ++x) {
        const uint32_t a = __float_as_uint(w[2 * x]), b = __float_as_uint(w[2 * x + 1]);
        hi[x] = (a >> 16) | (b & 0xffff0000u);
        lo[x] = (a & 0xffffu) | (b << 16);
      }
      st_v4(wp, make_uint4(hi[0], hi[1], hi[2], hi[3]));
      st_v4(tp, make_uint4(lo[0], lo[1], lo[2], lo[3]));
      return;
    }
  }
  store_vec<WT, VE, false>(reinterpret_cast<WT*>(wp), w);
}

// one update of run u's row with the gradient sum `acc` (lane l of the group holds vectors l, l + G, ...)
template <typename WT, int G, int NV, int OPT, bool RED>
__device__ __forceinline__ void finish_run(const UpdParams& p, int64_t u, float (*acc)[Vec16<WT>::VE], int lane) {
  constexpr int VE = Vec16<WT>::VE;
  const float lr = p.d_lr ? *p.d_lr : p.lr;
  char* wp;
  float* sp = nullptr;
  char* tp = nullptr;
  if (OPT == OPT_STORE) {
    wp = reinterpret_cast<char*>(p.values) + ((size_t)(p.by_key ? (int64_t)p.unique[u] : u) * p.D + lane * VE) * sizeof(WT);
  } else {
    const uint32_t tt = p.first_val[u] / p.B;
    const int64_t row = (int64_t)p.unique[u] - p.row_base[tt];
    wp = static_cast<char*>(p.weights[tt]) + ((size_t)row * p.D + lane * VE) * sizeof(WT);
    if (OPT == B200_OPT_ADAGRAD) sp = p.state[tt] + (size_t)row * p.D + lane * VE;
    if constexpr (sizeof(WT) == 2) {
      if (p.trail != nullptr) tp = wp + (reinterpret_cast<char*>(p.trail[tt]) - static_cast<char*>(p.weights[tt]));
    }
  }
#pragma unroll
  for (int kk = 0; kk < NV; ++kk)
    if (lane + kk * G < p.V) {
      if constexpr (RED) {
        red_add_v4_f32(reinterpret_cast<float*>(wp + (size_t)kk * G * 16), -lr * acc[kk][0], -lr * acc[kk][1],
                       -lr * acc[kk][2], -lr * acc[kk][3]);
      } else {
        float wf[VE] = {};
        uint4 tv = make_uint4(0, 0, 0, 0);
        if constexpr (OPT != OPT_STORE) {
          Vec16<WT>::unpack(wf, ld_v4(wp + (size_t)kk * G * 16));
          if (tp) tv = ld_v4(tp + (size_t)kk * G * 16);
        }
        apply_row<OPT, WT>(wp + (size_t)kk * G * 16, sp ? sp + kk * G * VE : nullptr, acc[kk], wf, lr, p.eps,
                           tp ? tp + (size_t)kk * G * 16 : nullptr, tv);
      }
    }
}

constexpr int UPD_TMAX = 256;  // tables whose base pointers / row bases are cached in shared memory

// RED (fp32 SGD only): the row update W += -lr * sum is linear, so it is issued as ONE vector
// reduction per 16 bytes (REDG.ADD.F32x4, executed by the L2 atomic unit) instead of a load +
// store: no weight row ever travels to the SM, no registers wait for it.  Each row still gets
// exactly one update per step from exactly one group, so the result stays deterministic.
template <typename WT, typename GT, int G, int NV, int OPT, bool RED, bool TRAIL>
__global__ void __launch_bounds__(UPD_THREADS, (sizeof(GT) == 2 && !TRAIL && OPT != B200_OPT_ADAGRAD) ? 3 : 2) window_update_kernel(const UpdParams p) {
  static_assert(!RED || (OPT == B200_OPT_SGD && sizeof(WT) == 4), "RED path is fp32 SGD only");
  static_assert(!TRAIL || sizeof(WT) == 2, "a trail belongs to bf16 tables");
  constexpr int VE = Vec16<WT>::VE;
  constexpr int NG = UPD_THREADS / G;
  constexpr int C = HOT_CH;
  constexpr int BS0 = (NV * VE >= 32) ? 1 : (NV * VE >= 16 ? 2 : (NV * VE >= 8 ? 4 : 8));
  // gradient rows in flight per lane; Adagrad also keeps the state rows of the batch in flight (below), so it takes
  // half the batch to stay inside the register budget
  constexpr int BS = RED ? 2 * BS0 : ((OPT == B200_OPT_ADAGRAD && BS0 > 1) ? BS0 / 2 : BS0);
  __shared__ int64_t s_base[UPD_TMAX];
  __shared__ char* s_wptr[UPD_TMAX];
  __shared__ float* s_sptr[UPD_TMAX];
  __shared__ int64_t s_tdelta[UPD_TMAX];  // trail base - weight base (bytes), 0 = no trail
  __shared__ uint32_t s_key[NG][C + 2];  // [0] = key before the window, [C+1] = key after it
  __shared__ int64_t s_goff[NG][C];      // address of the gradient row of each position
  __shared__ uint16_t s_tab[NG][C];
  extern __shared__ __align__(16) float s_part[];  // [NG][2][D]: pieces of the runs that cross window borders
  const int gid = threadIdx.x / G, lane = threadIdx.x % G;
  const bool cached = p.T <= UPD_TMAX && OPT != OPT_STORE;
  if (cached) {
    for (int i = threadIdx.x; i < p.T; i += UPD_THREADS) {
      s_base[i] = p.row_base[i];
      s_wptr[i] = static_cast<char*>(p.weights[i]);
      s_sptr[i] = (OPT == B200_OPT_ADAGRAD) ? p.state[i] : nullptr;
      s_tdelta[i] = (sizeof(WT) == 2 && p.trail != nullptr)
                        ? reinterpret_cast<char*>(p.trail[i]) - static_cast<char*>(p.weights[i]) : 0;
    }
  }
  constexpr bool has_trail = TRAIL;  // compile-time: plain bf16 tables carry no trail registers
  const int64_t n_eff = p.d_n ? min(*p.d_n, p.n) : p.n;
  const float lr = p.d_lr ? *p.d_lr : p.lr;
  // A block works on CHUNKS of NG windows (BC keys): the one of its index and, in a persistent launch, further ones
  // claimed from a device counter (the claim for the next chunk is issued at the top of the current one).  What the
  // cross-block join needs about a chunk's border runs is recorded per chunk and the join itself (fence, tickets,
  // finishing) is taken once per KMAX chunks: its fence has to drain every update the block still has in flight, which
  // costs ~2 us of a block's ~15 us when it is paid per chunk.
  constexpr int64_t BC = (int64_t)NG * C;
  constexpr int KMAX = 16;
  __shared__ int64_t s_xs[KMAX][6];  // per recorded chunk: {run, start, end} of the entering and of the leaving run
  __shared__ int64_t s_job[2 * KMAX][3];  // {run index, first position of the run, last block of the run}
  __shared__ int s_njobs;
  __shared__ int64_t s_claim;
  int64_t blk = blockIdx.x;
  bool more = true;
  while (more) {
  int nrec = 0;  // chunks of this round with a run across a border (block-uniform)
  for (;;) {
  if (blk * BC >= n_eff) {  // behind the last real key (uniform)
    more = false;
    break;
  }
  int64_t nxt = 0;
  if (p.persist && threadIdx.x == 0) nxt = (int64_t)gridDim.x + atomicAdd(&p.sched[0], 1);
  const int64_t w = blk * NG + gid;
  const int64_t a = w * C;
  const int cnt = a < n_eff ? (int)min((int64_t)C, n_eff - a) : 0;
  for (int i = lane; i < cnt; i += G) {
    const uint32_t g = p.svals[a + i];
    const uint32_t tt = g / p.B;
    s_key[gid][i + 1] = p.skeys[a + i];
    s_tab[gid][i] = (uint16_t)tt;
    const int64_t bb = g - tt * p.B;
    const GT* gaddr;
    if (p.grad_ptrs != nullptr) {
      const int64_t r = bb / p.lN;
      gaddr = static_cast<const GT*>(p.grad_ptrs[r]) + ((bb - r * p.lN) * p.F + p.feat0 + tt) * p.D;
    } else {
      gaddr = static_cast<const GT*>(p.grad) + (int64_t)tt * p.grad_stride_table + bb * p.grad_stride_bag;
    }
    s_goff[gid][i] = reinterpret_cast<int64_t>(gaddr);  // absolute address of the gradient row
  }
  if (lane == 0 && cnt > 0) {
    // sentinels: a key that differs from its neighbour when there is no neighbour
    s_key[gid][0] = a > 0 ? p.skeys[a - 1] : ~p.skeys[a];
    s_key[gid][cnt + 1] = a + cnt < n_eff ? p.skeys[a + cnt] : ~p.skeys[a + cnt - 1];
  }
  __syncthreads();
  // (a group whose window lies beyond the end has cnt == 0: it skips the work but keeps the barriers)
  // Does a run cross this block's left / right border?  Known now from the staged keys (block-uniform); thread 0
  // fetches what the cross-block join at the end needs about those runs right away, so that those dependent loads
  // are long done when it gets there.
  const int64_t x_first = blk * BC;
  const int x_ngv = (int)min((int64_t)NG, (n_eff - x_first + C - 1) / C);
  const int x_cntl = (int)min((int64_t)C, n_eff - (x_first + (int64_t)(x_ngv - 1) * C));
  const bool x_enters = s_key[0][0] == s_key[0][1];
  const bool x_leaves = s_key[x_ngv - 1][x_cntl] == s_key[x_ngv - 1][x_cntl + 1];
  if (threadIdx.x == 0 && (x_enters || x_leaves)) {
    int64_t* s_x = s_xs[nrec];
    int64_t uE = -1, uL = -2;
    if (x_enters) {
      uE = p.win_run[blk * NG];
      s_x[1] = p.seg[uE], s_x[2] = p.seg[uE + 1];
    }
    if (x_leaves) {
      uL = p.win_run[(blk + 1) * NG];
      s_x[4] = p.seg[uL], s_x[5] = p.seg[uL + 1];
    }
    s_x[0] = uE, s_x[3] = uL;
  }

  float acc[NV][VE];
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int x = 0; x < VE; ++x) acc[k][x] = 0.f;
  // is the run being summed a continuation from the previous window, and where is its head?
  const bool cont0 = cnt > 0 && s_key[gid][0] == s_key[gid][1];
  bool cont = cont0;
  int cur_head = cont ? -1 : 0;
  const int64_t uidx0 = (OPT == OPT_STORE && cnt > 0) ? (int64_t)p.win_run[w] : 0;
  int64_t uidx = uidx0;  // index of the run being summed
  // Load/store variants: the group that holds the head of a run crossing its right border will finish that row
  // after the block barrier -- fetch the weight (and trail) row NOW so that the read overlaps the window loop.
  [[maybe_unused]] uint4 pre_w[NV], pre_t[NV];
  [[maybe_unused]] bool pre_ok = false;
  if constexpr (!RED && OPT != OPT_STORE) {
    const bool own_tail = cnt > 0 && s_key[gid][cnt] == s_key[gid][cnt + 1] &&
                          !(cont0 && s_key[gid][1] == s_key[gid][cnt]);
    if (own_tail) {
      pre_ok = true;
      const int tt = s_tab[gid][cnt - 1];
      const int64_t row = (int64_t)s_key[gid][cnt] - (cached ? s_base[tt] : p.row_base[tt]);
      const char* wq = (cached ? s_wptr[tt] : static_cast<char*>(p.weights[tt])) + ((size_t)row * p.D + lane * VE) * sizeof(WT);
      const int64_t td = !has_trail ? 0 : (cached ? s_tdelta[tt] : reinterpret_cast<char*>(p.trail[tt]) - static_cast<char*>(p.weights[tt]));
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const bool in = lane + k * G < p.V;
        pre_w[k] = in ? ld_v4(wq + (size_t)k * G * 16) : make_uint4(0, 0, 0, 0);
        pre_t[k] = (in && has_trail) ? ld_v4(wq + td + (size_t)k * G * 16) : make_uint4(0, 0, 0, 0);
      }
    }
  }

  for (int c0 = 0; c0 < cnt; c0 += BS) {
    GradVec<GT, VE> g[BS][NV];
    uint4 wv[RED ? 1 : BS][NV];
    [[maybe_unused]] uint4 tv[TRAIL ? BS : 1][NV];  // trail halves of the rows in wv
    // Adagrad: the fp32 state vectors of the rows in wv, loaded with the batch -- fetched inside the update they cost
    // one exposed memory latency per finished run
    [[maybe_unused]] uint4 sv[(OPT == B200_OPT_ADAGRAD) ? BS : 1][NV][VE / 4];
    [[maybe_unused]] int64_t tdel[TRAIL ? BS : 1];
    char* wp[RED ? 1 : BS];
    float* sp[RED ? 1 : BS];
    bool ends[BS], closed[BS];
#pragma unroll
    for (int c = 0; c < BS; ++c) {
      const int pos = c0 + c;
      const bool ok = pos < cnt;
      const uint32_t key = ok ? s_key[gid][pos + 1] : 0u;
      if (ok && pos > 0 && s_key[gid][pos] != key) {  // a run begins here, inside the window
        cont = false;
        cur_head = pos;
        if (OPT == OPT_STORE) ++uidx;
      }
      ends[c] = ok && s_key[gid][pos + 2] != key;
      closed[c] = ends[c] && !cont;  // began and ends inside this window: finished here
      const GT* gp = reinterpret_cast<const GT*>(ok ? s_goff[gid][pos] : s_goff[gid][0]) + lane * VE;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        if (ok && lane + k * G < p.V) {
          g[c][k].load(gp + k * G * VE);
        } else {
          g[c][k].zero();
        }
      }
      if constexpr (!RED) {
        wp[c] = nullptr;
        sp[c] = nullptr;
#pragma unroll
        for (int k = 0; k < NV; ++k) wv[c][k] = make_uint4(0, 0, 0, 0);
        if (closed[c]) {
          if (OPT == OPT_STORE) {
            wp[c] = reinterpret_cast<char*>(p.values) + ((size_t)(p.by_key ? (int64_t)key : uidx) * p.D + lane * VE) * sizeof(WT);
          } else {
            const int tt = s_tab[gid][pos];
            const int64_t row = (int64_t)key - (cached ? s_base[tt] : p.row_base[tt]);
            wp[c] = (cached ? s_wptr[tt] : static_cast<char*>(p.weights[tt])) + ((size_t)row * p.D + lane * VE) * sizeof(WT);
            if (OPT == B200_OPT_ADAGRAD) sp[c] = (cached ? s_sptr[tt] : p.state[tt]) + (size_t)row * p.D + lane * VE;
#pragma unroll
            for (int k = 0; k < NV; ++k)
              if (lane + k * G < p.V) wv[c][k] = ld_v4(wp[c] + (size_t)k * G * 16);
            if constexpr (OPT == B200_OPT_ADAGRAD) {
#pragma unroll
              for (int k = 0; k < NV; ++k)
#pragma unroll
                for (int h = 0; h < VE / 4; ++h)
                  sv[c][k][h] = (lane + k * G < p.V) ? ld_v4(sp[c] + k * G * VE + 4 * h) : make_uint4(0, 0, 0, 0);
            }
            if constexpr (TRAIL) {
              tdel[c] = cached ? s_tdelta[tt] : reinterpret_cast<char*>(p.trail[tt]) - static_cast<char*>(p.weights[tt]);
#pragma unroll
              for (int k = 0; k < NV; ++k)
                tv[c][k] = (lane + k * G < p.V) ? ld_v4(wp[c] + tdel[c] + (size_t)k * G * 16) : make_uint4(0, 0, 0, 0);
            }
          }
        }
      }
    }
    // every load of the batch is issued by now; pin them so they stay in flight together
#pragma unroll
    for (int c = 0; c < BS; ++c)
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        g[c][k].pin();
        if constexpr (!RED) keep(wv[c][k]);
      }
#pragma unroll
    for (int c = 0; c < BS; ++c) {
#pragma unroll
      for (int k = 0; k < NV; ++k) g[c][k].add_to(acc[k]);
      if (ends[c]) {
        if (closed[c]) {
#pragma unroll
          for (int k = 0; k < NV; ++k) {
            if (lane + k * G < p.V) {
              if constexpr (RED) {
                const int pos = c0 + c;
                const int tt = s_tab[gid][pos];
                const int64_t row = (int64_t)s_key[gid][pos + 1] - (cached ? s_base[tt] : p.row_base[tt]);
                float* dstw = reinterpret_cast<float*>(cached ? s_wptr[tt] : static_cast<char*>(p.weights[tt])) +
                              (size_t)row * p.D + lane * VE + k * G * VE;
                red_add_v4_f32(dstw, -lr * acc[k][0], -lr * acc[k][1], -lr * acc[k][2], -lr * acc[k][3]);
              } else {
                float wf[VE];
                Vec16<WT>::unpack(wf, wv[c][k]);
                const uint4* svp = nullptr;
                if constexpr (OPT == B200_OPT_ADAGRAD) svp = &sv[c][k][0];
                if constexpr (TRAIL) {
                  apply_row<OPT, WT>(wp[c] + (size_t)k * G * 16, sp[c] ? sp[c] + k * G * VE : nullptr, acc[k], wf, lr,
                                     p.eps, wp[c] + tdel[c] + (size_t)k * G * 16, tv[c][k], svp);
                } else {
                  apply_row<OPT, WT>(wp[c] + (size_t)k * G * 16, sp[c] ? sp[c] + k * G * VE : nullptr, acc[k], wf, lr,
                                     p.eps, nullptr, make_uint4(0, 0, 0, 0), svp);
                }
              }
            }
          }
        } else {  // continuation from the previous window that ends here: piece in slot 0
          float* dst = s_part + (size_t)(2 * gid) * p.D + lane * VE;
#pragma unroll
          for (int k = 0; k < NV; ++k)
            if (lane + k * G < p.V) {
#pragma unroll
              for (int x = 0; x < VE; ++x) dst[k * G * VE + x] = acc[k][x];
            }
        }
#pragma unroll
        for (int k = 0; k < NV; ++k)
#pragma unroll
          for (int x = 0; x < VE; ++x) acc[k][x] = 0.f;
      }
    }
  }
  // a run still open at the right border: piece in slot 0 if it began at or before the first position
  const bool open_r = cnt > 0 && s_key[gid][cnt] == s_key[gid][cnt + 1];
  const int tail_slot = cur_head <= 0 ? 0 : 1;
  if (open_r) {
    float* dst = s_part + (size_t)(2 * gid + tail_slot) * p.D + lane * VE;
#pragma unroll
    for (int k = 0; k < NV; ++k)
      if (lane + k * G < p.V) {
#pragma unroll
        for (int x = 0; x < VE; ++x) dst[k * G * VE + x] = acc[k][x];
      }
  }
  __syncthreads();
  if (cnt > 0) {
  // ---- in-block join.  The group holding the HEAD of a run that crosses its right border adds the
  // slot-0 pieces of the following windows (window order) and finishes the run; group 0 does the same
  // for the run that enters the block from the left.  What crosses a block border goes to p.partial.
  const int64_t blk_first = blk * BC;
  const int ngv = (int)min((int64_t)NG, (n_eff - blk_first + C - 1) / C);  // windows of this block with data
  const int cnt_last = (int)min((int64_t)C, n_eff - (blk_first + (int64_t)(ngv - 1) * C));
  const uint32_t key_after_block = s_key[ngv - 1][cnt_last + 1];
  // run entering from the previous window covers this whole window and goes on: some earlier group owns it
  const bool pass_through = cont0 && open_r && s_key[gid][1] == s_key[gid][cnt];
  auto finish = [&](int slot, uint32_t key, bool from_left, bool open, int pos, int64_t u, bool use_pre) {
    const float* src = s_part + (size_t)(2 * gid + slot) * p.D + lane * VE;
#pragma unroll
    for (int k = 0; k < NV; ++k)
#pragma unroll
      for (int x = 0; x < VE; ++x) acc[k][x] = (lane + k * G < p.V) ? src[k * G * VE + x] : 0.f;
    int j = gid + 1;
    if (open) {
      for (; j < ngv && s_key[j][1] == key; ++j) {
        const float* q = s_part + (size_t)(2 * j) * p.D + lane * VE;
#pragma unroll
        for (int k = 0; k < NV; ++k)
          if (lane + k * G < p.V) {
#pragma unroll
            for (int x = 0; x < VE; ++x) acc[k][x] += q[k * G * VE + x];
          }
      }
    }
    const bool to_right = open && j == ngv && key_after_block == key;
    if (from_left || to_right) {  // crosses a block border: per-block partial row, finished by the cross-block join below
      const int bslot = (from_left || (gid == 0 && slot == 0)) ? 0 : 1;
      float* dst = p.partial + (size_t)(2 * blk + bslot) * p.D + lane * VE;
#pragma unroll
      for (int k = 0; k < NV; ++k)
        if (lane + k * G < p.V) {
#pragma unroll
          for (int h = 0; h < VE / 4; ++h) store_vec<float, 4, false>(dst + k * G * VE + 4 * h, acc[k] + 4 * h);
        }
      return;
    }
    char* wp;
    float* sp = nullptr;
    char* tp = nullptr;
    if (OPT == OPT_STORE) {
      wp = reinterpret_cast<char*>(p.values) + ((size_t)(p.by_key ? (int64_t)key : u) * p.D + lane * VE) * sizeof(WT);
    } else {
      const int tt = s_tab[gid][pos];
      const int64_t row = (int64_t)key - (cached ? s_base[tt] : p.row_base[tt]);
      wp = (cached ? s_wptr[tt] : static_cast<char*>(p.weights[tt])) + ((size_t)row * p.D + lane * VE) * sizeof(WT);
      if (OPT == B200_OPT_ADAGRAD) sp = (cached ? s_sptr[tt] : p.state[tt]) + (size_t)row * p.D + lane * VE;
      if (has_trail) tp = wp + (cached ? s_tdelta[tt] : reinterpret_cast<char*>(p.trail[tt]) - static_cast<char*>(p.weights[tt]));
    }
#pragma unroll
    for (int k = 0; k < NV; ++k)
      if (lane + k * G < p.V) {
        if constexpr (RED) {
          red_add_v4_f32(reinterpret_cast<float*>(wp + (size_t)k * G * 16), -lr * acc[k][0], -lr * acc[k][1],
                         -lr * acc[k][2], -lr * acc[k][3]);
        } else {
          float wf[VE] = {};
          uint4 tv = make_uint4(0, 0, 0, 0);
          if constexpr (OPT != OPT_STORE) {
            if (use_pre) {
              Vec16<WT>::unpack(wf, pre_w[k]);
              tv = pre_t[k];
            } else {
              Vec16<WT>::unpack(wf, ld_v4(wp + (size_t)k * G * 16));
              if (tp) tv = ld_v4(tp + (size_t)k * G * 16);
            }
          }
          apply_row<OPT, WT>(wp + (size_t)k * G * 16, sp ? sp + k * G * VE : nullptr, acc[k], wf, lr, p.eps,
                             tp ? tp + (size_t)k * G * 16 : nullptr, tv);
        }
      }
  };
  if (gid == 0 && cont0)  // the run that enters the block from the left (its head is in an earlier block)
    finish(0, s_key[0][1], true, pass_through, 0, uidx0, false);
  if (open_r && !pass_through)  // the run whose head is in this window and that crosses its right border
    finish(tail_slot, s_key[gid][cnt], false, true, cnt - 1, uidx, pre_ok);
  }

  if (x_enters || x_leaves) ++nrec;
  if (!p.persist) {
    more = false;
    break;
  }
  if (threadIdx.x == 0) s_claim = nxt;
  __syncthreads();  // (also: every group is done with this chunk's shared-memory pieces)
  blk = s_claim;
  if (nrec >= min(p.kflush, KMAX)) break;
  }  // chunks of one round

  // ---- cross-block join, inside this kernel.  A run that crosses chunk borders has left one partial row per chunk
  // it touches (above).  Every such chunk takes a ticket on the counter of the run's HEAD chunk once its partial row
  // is visible device-wide; whoever draws the last ticket adds the partial rows of all the run's chunks in chunk
  // order (whoever does it, the order is the same: deterministic) with all its groups -- strided rows, PB in
  // flight, combined in shared memory -- and gives the row its one update.  No second kernel, no launch gap.
  if (nrec == 0) continue;  // (block-uniform) no run crossed a border of this round's chunks
  __threadfence();
  if (threadIdx.x == 0) s_njobs = 0;
  __syncthreads();
  if (threadIdx.x < 2 * nrec) {  // one thread per (chunk, border): the atomics' round trips overlap
    const int64_t* x = s_xs[threadIdx.x >> 1];
    const bool left = (threadIdx.x & 1) == 0;
    auto ticket = [&](int64_t u, int64_t s0, int64_t e0) {
      const int64_t h = s0 / BC, w1 = (min(e0, n_eff) - 1) / BC;
      const int need = (int)(w1 - h + 1);
      if (atomicAdd(&p.blk_cnt[h], 1) + 1 == need) {
        p.blk_cnt[h] = 0;  // every contributor has drawn: ready for the next call on this workspace
        __threadfence();
        const int j = atomicAdd(&s_njobs, 1);
        s_job[j][0] = u, s_job[j][1] = s0, s_job[j][2] = w1;
      }
    };
    if (left && x[0] >= 0) ticket(x[0], x[1], x[2]);
    if (!left && x[3] >= 0 && x[3] != x[0]) ticket(x[3], x[4], x[5]);
  }
  __syncthreads();
  const int njobs = s_njobs;
  constexpr int PB = (NV * VE >= 32) ? 2 : (NV * VE >= 16 ? 4 : 8);
  // Runs over at most NG chunks (nearly all of them): one GROUP per job, the jobs of a block side by side and without
  // block barriers.  The rows are added in chunk order, which is also the order of the cooperative path below for
  // such a run (one row per group, groups combined in order): the result does not depend on the path.
  for (int jb = gid; jb < njobs; jb += NG) {
    const int64_t u = s_job[jb][0], s0 = s_job[jb][1], w1 = s_job[jb][2];
    const int64_t w0 = s0 / BC;
    if (w1 - w0 >= NG) continue;
    float sum[NV][VE];
#pragma unroll
    for (int kk = 0; kk < NV; ++kk)
#pragma unroll
      for (int x = 0; x < VE; ++x) sum[kk][x] = 0.f;
    for (int64_t ib = w0; ib <= w1; ib += PB) {
      uint4 f[PB][NV][VE / 4];
#pragma unroll
      for (int c = 0; c < PB; ++c) {
        const int64_t i = ib + c;
        const bool ok = i <= w1;
        const float* src = p.partial + (size_t)(2 * (ok ? i : w0) + ((ok && s0 > i * BC) ? 1 : 0)) * p.D + lane * VE;
#pragma unroll
        for (int kk = 0; kk < NV; ++kk)
#pragma unroll
          for (int h = 0; h < VE / 4; ++h)
            f[c][kk][h] = (ok && lane + kk * G < p.V) ? ld_cg_v4(src + kk * G * VE + 4 * h) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int c = 0; c < PB; ++c)
#pragma unroll
        for (int kk = 0; kk < NV; ++kk)
#pragma unroll
          for (int h = 0; h < VE / 4; ++h) keep(f[c][kk][h]);
#pragma unroll
      for (int c = 0; c < PB; ++c)
#pragma unroll
        for (int kk = 0; kk < NV; ++kk)
#pragma unroll
          for (int h = 0; h < VE / 4; ++h) Vec16<float>::accum(sum[kk] + 4 * h, f[c][kk][h]);
    }
    finish_run<WT, G, NV, OPT, RED>(p, u, sum, lane);
  }
  // longer runs (hot rows of tiny tables): all groups of the block on one job
  for (int jb = 0; jb < njobs; ++jb) {
    const int64_t u = s_job[jb][0], s0 = s_job[jb][1], w1 = s_job[jb][2];
    const int64_t w0 = s0 / BC;
    if (w1 - w0 < NG) continue;  // (block-uniform) done above
    float sum[NV][VE];
#pragma unroll
    for (int kk = 0; kk < NV; ++kk)
#pragma unroll
      for (int x = 0; x < VE; ++x) sum[kk][x] = 0.f;
    for (int64_t ib = w0 + gid; ib <= w1; ib += (int64_t)NG * PB) {
      uint4 f[PB][NV][VE / 4];
#pragma unroll
      for (int c = 0; c < PB; ++c) {
        const int64_t i = ib + (int64_t)c * NG;
        const bool ok = i <= w1;
        // slot rule: the block holding the run's head wrote slot 1 (slot 0 if the run starts at its first position)
        const float* src = p.partial + (size_t)(2 * (ok ? i : w0) + ((ok && s0 > i * BC) ? 1 : 0)) * p.D + lane * VE;
#pragma unroll
        for (int kk = 0; kk < NV; ++kk)
#pragma unroll
          for (int h = 0; h < VE / 4; ++h)
            f[c][kk][h] = (ok && lane + kk * G < p.V) ? ld_cg_v4(src + kk * G * VE + 4 * h) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int c = 0; c < PB; ++c)
#pragma unroll
        for (int kk = 0; kk < NV; ++kk)
#pragma unroll
          for (int h = 0; h < VE / 4; ++h) keep(f[c][kk][h]);
#pragma unroll
      for (int c = 0; c < PB; ++c)
#pragma unroll
        for (int kk = 0; kk < NV; ++kk)
#pragma unroll
          for (int h = 0; h < VE / 4; ++h) Vec16<float>::accum(sum[kk] + 4 * h, f[c][kk][h]);
    }
    float* mine = s_part + (size_t)gid * p.D + lane * VE;  // (the in-block pieces are consumed: reuse the buffer)
#pragma unroll
    for (int kk = 0; kk < NV; ++kk)
      if (lane + kk * G < p.V) {
#pragma unroll
        for (int x = 0; x < VE; ++x) mine[kk * G * VE + x] = sum[kk][x];
      }
    __syncthreads();
    if (gid == 0) {
      for (int g = 1; g < NG; ++g) {
        const float* q = s_part + (size_t)g * p.D + lane * VE;
#pragma unroll
        for (int kk = 0; kk < NV; ++kk)
          if (lane + kk * G < p.V) {
#pragma unroll
            for (int x = 0; x < VE; ++x) sum[kk][x] += q[kk * G * VE + x];
          }
      }
      finish_run<WT, G, NV, OPT, RED>(p, u, sum, lane);
    }
    __syncthreads();
  }
  }  // rounds
  if (p.persist && threadIdx.x == 0) {
    // the last block to finish leaves the scheduler at zero for the next launch on this workspace
    if (atomicAdd(&p.sched[1], 1) + 1 == (int)gridDim.x) p.sched[0] = 0, p.sched[1] = 0;
  }
}

// scalar fallback (rows not a whole number of 16-byte vectors): thread per (unique row, d)
template <typename WT, typename GT, int OPT>
__global__ void emb_update_scalar_kernel(const UpdParams p, int64_t n_max) {
  const int64_t nU = *p.num_unique;
  const float lr = p.d_lr ? *p.d_lr : p.lr;
  const int64_t total = nU * p.D;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t u = i / p.D;
    const int d = (int)(i % p.D);
    const uint32_t s = p.seg[u], e = p.seg[u + 1];
    if (p.d_n && (int64_t)s >= *p.d_n) continue;  // the run of sentinel keys behind the real ones
    const uint32_t t = p.svals[s] / p.B;
    const int64_t row = (int64_t)p.unique[u] - p.row_base[t];
    float acc = 0.f;
    for (uint32_t j = s; j < e; ++j) {
      const uint32_t g = p.svals[j];
      const uint32_t b = g - t * p.B;
      acc += to_f32<GT>(static_cast<const GT*>(p.grad)[(int64_t)t * p.grad_stride_table +
                                                      (int64_t)b * p.grad_stride_bag + d]);
    }
    if (OPT == OPT_STORE) {
      p.values[(p.by_key ? (int64_t)p.unique[u] : u) * p.D + d] = acc;
      continue;
    }
    WT* w = static_cast<WT*>(p.weights[t]) + row * p.D + d;
    float wv = to_f32<WT>(*w);
    uint16_t* tr = (sizeof(WT) == 2 && p.trail != nullptr) ? p.trail[t] + row * p.D + d : nullptr;
    if (tr) wv = __uint_as_float(__float_as_uint(wv) | *tr);
    if (OPT == B200_OPT_ADAGRAD) {
      float* sp = p.state[t] + row * p.D + d;
      const float st = *sp + acc * acc;
      *sp = st;
      wv -= lr * acc / (sqrtf(st) + p.eps);
    } else {
      wv -= lr * acc;
    }
    if (tr) {
      const uint32_t bits = __float_as_uint(wv);
      *reinterpret_cast<uint16_t*>(w) = (uint16_t)(bits >> 16);
      *tr = (uint16_t)(bits & 0xffffu);
    } else {
      *w = from_f32<WT>(wv);
    }
  }
}

__global__ void unique_to_i64_kernel(const uint32_t* __restrict__ unique, const int32_t* __restrict__ num_unique,
                                     int64_t* __restrict__ out, int64_t n) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n && i < *num_unique) out[i] = (int64_t)unique[i];
}

template <typename WT, typename GT, int OPT>
static int launch_update(const UpdParams& p, int64_t n, bool vec_ok, cudaStream_t st) {
  if (!vec_ok) {
    const int blocks = (int)std::min<int64_t>(ceil_div(n * p.D, 256), 148 * 16);
    emb_update_scalar_kernel<WT, GT, OPT><<<blocks, 256, 0, st>>>(p, n);
    B200_CHECK_LAUNCH("emb_update_scalar");
    return B200_OK;
  }
  const int V = p.V;
  const int64_t nwin = ceil_div(n, HOT_CH);
  constexpr bool CAN_RED = OPT == B200_OPT_SGD && sizeof(WT) == 4;
  static const bool use_red = []() {
    const char* e = getenv("B200_SGD_RED");
    return e == nullptr || e[0] != '0';
  }();
  // B200_UPD_PERSIST: 0 = one chunk per block, 1 = always one wave of chunk-claiming blocks, unset / "auto" = from two
  // waves of chunks on; B200_UPD_GRID overrides the size of that wave (tests)
  static const int persist_mode = []() {
    const char* e = getenv("B200_UPD_PERSIST");
    return e == nullptr ? PERSIST_DEFAULT : (e[0] == '0' ? 0 : (e[0] == '1' ? 1 : 2));
  }();
  static const int persist_kflush = []() {
    const char* e = getenv("B200_UPD_KFLUSH");
    const int v = e == nullptr ? 16 : atoi(e);
    return v < 1 ? 1 : (v > 16 ? 16 : v);
  }();
  static const int64_t persist_grid = []() {
    const char* e = getenv("B200_UPD_GRID");
    return e == nullptr ? (int64_t)0 : (int64_t)atoll(e);
  }();
#define B200_UPD_LAUNCH(G, NV, RED)                                                                               \
  if constexpr (sizeof(WT) == 2) {                                                                                \
    if (p.trail != nullptr) {                                                                                     \
      B200_UPD_LAUNCH_T(G, NV, RED, true)                                                                         \
    } else {                                                                                                      \
      B200_UPD_LAUNCH_T(G, NV, RED, false)                                                                        \
    }                                                                                                             \
  } else {                                                                                                        \
    B200_UPD_LAUNCH_T(G, NV, RED, false)                                                                          \
  }
#define B200_UPD_LAUNCH_T(G, NV, RED, TRAIL)                                                                               \
  {                                                                                                               \
    constexpr int NG_ = UPD_THREADS / G;                                                                          \
    const size_t part_bytes = (size_t)NG_ * 2 * p.D * sizeof(float);                                              \
    auto kern = window_update_kernel<WT, GT, G, NV, OPT, RED, TRAIL>;                                                  \
    if (part_bytes > 48 * 1024) { /* per device, so set on every such launch (cheap) rather than once per process */ \
      const cudaError_t attr_rc =                                                                                 \
          cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)part_bytes);               \
      B200_REQUIRE(attr_rc == cudaSuccess && part_bytes <= 128 * 1024, B200_EUNSUPPORTED,                         \
                   "emb_update: row too long for the in-block join");                                             \
    }                                                                                                             \
    const int64_t nblk = ceil_div(nwin, NG_);                                                                     \
    UpdParams q = p;                                                                                              \
    int64_t grid = nblk;                                                                                          \
    q.persist = 0;                                                                                                \
    q.kflush = persist_kflush;                                                                                    \
    if (persist_mode != 0 && p.sched != nullptr) {                                                                \
      int occ = 0, dev = 0, sms = 0;                                                                              \
      cudaGetDevice(&dev);                                                                                        \
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);                                          \
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, UPD_THREADS, part_bytes);                         \
      const int64_t wave = persist_grid > 0 ? persist_grid : (int64_t)occ * sms;                                  \
      /* one wave of blocks claiming chunks pays when there are several waves of chunks */                        \
      if (wave > 0 && (persist_mode == 1 || nblk >= 2 * wave)) {                                                  \
        q.persist = 1;                                                                                            \
        grid = std::min(nblk, wave);                                                                              \
      }                                                                                                           \
    }                                                                                                             \
    kern<<<(unsigned)grid, UPD_THREADS, part_bytes, st>>>(q);                                                     \
  }
#define B200_UPD_CASE(G, NV)                 \
  if constexpr (CAN_RED) {                   \
    if (use_red) {                           \
      B200_UPD_LAUNCH(G, NV, true);          \
    } else {                                 \
      B200_UPD_LAUNCH(G, NV, false);         \
    }                                        \
  } else {                                   \
    B200_UPD_LAUNCH(G, NV, false);           \
  }
  if (V <= 4) { B200_UPD_CASE(4, 1); }
  else if (V <= 8) { B200_UPD_CASE(8, 1); }
  else if (V <= 16) { B200_UPD_CASE(16, 1); }
  else if (V <= 32) { B200_UPD_CASE(32, 1); }
  else if (V <= 64) { B200_UPD_CASE(32, 2); }
  else { B200_UPD_CASE(32, 4); }
#undef B200_UPD_CASE
#undef B200_UPD_LAUNCH
#undef B200_UPD_LAUNCH_T
  B200_CHECK_LAUNCH("emb_update");
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_emb_bwd_workspace_bytes(int64_t nnz, int64_t D) {
  return bwd_layout(nullptr, nnz < 0 ? 0 : nnz, D < 1 ? 1 : D).total;
}

static int check_bwd_common(int64_t T, int64_t B, int64_t D, int64_t nnz, int64_t total_rows) {
  B200_REQUIRE(T >= 0 && B >= 0 && D > 0 && nnz >= 0, B200_EINVAL, "emb_bwd: negative size");
  B200_REQUIRE(total_rows > 0 && (uint64_t)total_rows <= ((uint64_t)1 << 32), B200_EUNSUPPORTED,
               "emb_bwd: total rows over all tables must be <= 2^32 (32-bit sort keys)");
  B200_REQUIRE(T * B < ((int64_t)1 << 32) && nnz < ((int64_t)1 << 31) && T <= 65535,
               B200_EUNSUPPORTED, "emb_bwd: T*B must be < 2^32 and nnz < 2^31");
  return B200_OK;
}

static int emb_bwd_prepare_impl(const int64_t* d_rows, const int64_t* d_row_base, const int64_t* d_indices,
                                const int64_t* d_offsets, int64_t nnz, const int64_t* d_nnz_valid, int64_t T,
                                int64_t B, int64_t D, int64_t total_rows, void* ws, size_t ws_bytes,
                                int32_t* d_error, b200_stream_t stream) {
  // a partially filled index array needs one more key value (the sentinel) behind the real row ids
  int rc = check_bwd_common(T, B, D, nnz, total_rows + (d_nnz_valid ? 1 : 0));
  if (rc) return rc;
  if (T == 0 || B == 0 || nnz == 0) return B200_OK;
  B200_REQUIRE(d_rows && d_row_base && d_indices && d_offsets && ws, B200_EINVAL,
               "emb_bwd_prepare: null pointer");
  B200_REQUIRE(ws_bytes >= b200_emb_bwd_workspace_bytes(nnz, D), B200_EWORKSPACE,
               "emb_bwd: workspace too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  BwdLayout L = bwd_layout(ws, nnz, D);
  const int64_t n_blk_cnt = ceil_div(ceil_div(nnz, HOT_CH), MIN_NG) + 3;  // tickets + the chunk scheduler's two counters
  keys_tail_kernel<<<(unsigned)std::min<int64_t>(d_nnz_valid ? 256 : ceil_div(n_blk_cnt, 256), 256), 256, 0, st>>>(
      L.keys, L.vals, d_nnz_valid, nnz, (uint32_t)total_rows, L.n_valid, L.blk_cnt, n_blk_cnt);
  dim3 kgrid((unsigned)ceil_div(B, KEYS_THREADS), (unsigned)T);
  build_keys_kernel<<<kgrid, KEYS_THREADS, 0, st>>>(d_indices, d_offsets, d_rows, d_row_base, B, L.keys,
                                                    L.vals, d_error);
  B200_CHECK_LAUNCH("build_keys");
  rc = sort_pairs(L.keys, L.vals, nnz, (uint64_t)total_rows + (d_nnz_valid ? 1 : 0), L.skeys, L.svals, L.sort_ws, st);
  if (rc) return rc;
  return dedup_runs(L.skeys, nnz, L.unique, L.seg, L.num_unique, L.sort_ws, st, L.win_run, L.svals, L.first_val);
}

extern "C" int b200_emb_bwd_prepare(const int64_t* d_rows, const int64_t* d_row_base,
                                    const int64_t* d_indices, const int64_t* d_offsets, int64_t nnz,
                                    int64_t T, int64_t B, int64_t D, int64_t total_rows, void* ws, size_t ws_bytes,
                                    int32_t* d_error, b200_stream_t stream) {
  return emb_bwd_prepare_impl(d_rows, d_row_base, d_indices, d_offsets, nnz, nullptr, T, B, D, total_rows, ws,
                              ws_bytes, d_error, stream);
}

extern "C" int b200_emb_bwd_prepare_partial(const int64_t* d_rows, const int64_t* d_row_base,
                                            const int64_t* d_indices, const int64_t* d_offsets, int64_t nnz_capacity,
                                            const int64_t* d_nnz_valid, int64_t T, int64_t B, int64_t D,
                                            int64_t total_rows, void* ws, size_t ws_bytes, int32_t* d_error,
                                            b200_stream_t stream) {
  B200_REQUIRE(d_nnz_valid, B200_EINVAL, "emb_bwd_prepare_partial: null count");
  return emb_bwd_prepare_impl(d_rows, d_row_base, d_indices, d_offsets, nnz_capacity, d_nnz_valid, T, B, D,
                              total_rows, ws, ws_bytes, d_error, stream);
}

static int emb_bwd_apply_impl(void* const* d_weights, float* const* d_state, uint16_t* const* d_trail,
                              const int64_t* d_row_base,
                              int64_t nnz, const void* d_grad, int64_t T, int64_t B, int64_t D,
                              int64_t grad_stride_table, int64_t grad_stride_bag, int w_dtype, int grad_dtype,
                              int optimizer, float lr, const float* d_lr, float eps, void* ws, size_t ws_bytes,
                              b200_stream_t stream, const void* const* d_grad_ptrs, int64_t lN, int64_t F,
                              int64_t feat0) {
  B200_REQUIRE(T >= 0 && B >= 0 && D > 0 && nnz >= 0, B200_EINVAL, "emb_bwd: negative size");
  if (T == 0 || B == 0 || nnz == 0) return B200_OK;
  B200_REQUIRE(d_weights && d_row_base && (d_grad || d_grad_ptrs) && ws, B200_EINVAL, "emb_bwd_apply: null pointer");
  B200_REQUIRE(!d_grad_ptrs || (lN > 0 && B % lN == 0 && F > 0), B200_EINVAL,
               "emb_bwd_apply_sharded: B must be a multiple of lN");
  B200_REQUIRE(optimizer == B200_OPT_SGD || (optimizer == B200_OPT_ADAGRAD && d_state), B200_EINVAL,
               "emb_bwd: bad optimizer / missing state");
  B200_REQUIRE((w_dtype == B200_F32 || w_dtype == B200_BF16) &&
                   (grad_dtype == B200_F32 || grad_dtype == B200_BF16),
               B200_EINVAL, "emb_bwd: bad dtype");
  B200_REQUIRE(ws_bytes >= b200_emb_bwd_workspace_bytes(nnz, D), B200_EWORKSPACE,
               "emb_bwd: workspace too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  BwdLayout L = bwd_layout(ws, nnz, D);
  UpdParams p;
  p.weights = d_weights;
  p.state = d_state;
  p.trail = (w_dtype == B200_BF16) ? d_trail : nullptr;
  p.row_base = d_row_base;
  p.unique = L.unique;
  p.seg = L.seg;
  p.svals = L.svals;
  p.num_unique = L.num_unique;
  p.grad = d_grad;
  p.grad_stride_table = grad_stride_table;
  p.grad_stride_bag = grad_stride_bag;
  p.d_lr = d_lr;
  p.lr = lr;
  p.eps = eps;
  p.B = (uint32_t)B;
  p.D = (int)D;
  p.skeys = L.skeys;
  p.win_run = L.win_run;
  p.first_val = L.first_val;
  p.T = (int)T;
  p.partial = L.partial;
  p.n = nnz;
  p.d_n = L.n_valid;
  p.blk_cnt = L.blk_cnt;
  p.sched = L.blk_cnt + ceil_div(ceil_div(nnz, HOT_CH), MIN_NG) + 1;
  p.persist = 0;
  p.kflush = 16;
  p.values = nullptr;
  p.by_key = 0;
  p.grad_ptrs = d_grad_ptrs;
  p.lN = lN;
  p.F = F;
  p.feat0 = feat0;
  const int wsz = dtype_size(w_dtype), gsz = dtype_size(grad_dtype);
  p.V = (int)(D * wsz / 16);
  const int ve = 16 / wsz;  // elements per weight vector
  const bool vec_ok = (D * wsz) % 16 == 0 && p.V <= 128 &&
                      (reinterpret_cast<uintptr_t>(d_grad) % 16) == 0 &&
                      (grad_stride_table * gsz) % 16 == 0 && (grad_stride_bag * gsz) % 16 == 0 &&
                      (ve * gsz) % 8 == 0 && (!d_grad_ptrs || (D * gsz) % 16 == 0);
  B200_REQUIRE(vec_ok || !d_grad_ptrs, B200_EUNSUPPORTED,
               "emb_bwd_apply_sharded: row size must be a multiple of 16 bytes");
#define B200_UPD_DISPATCH(OPT)                                                                  \
  (w_dtype == B200_F32                                                                          \
       ? (grad_dtype == B200_F32 ? launch_update<float, float, OPT>(p, nnz, vec_ok, st)         \
                                 : launch_update<float, __nv_bfloat16, OPT>(p, nnz, vec_ok, st)) \
       : (grad_dtype == B200_F32                                                                \
              ? launch_update<__nv_bfloat16, float, OPT>(p, nnz, vec_ok, st)                     \
              : launch_update<__nv_bfloat16, __nv_bfloat16, OPT>(p, nnz, vec_ok, st)))
  if (optimizer == B200_OPT_ADAGRAD) return B200_UPD_DISPATCH(B200_OPT_ADAGRAD);
  return B200_UPD_DISPATCH(B200_OPT_SGD);
#undef B200_UPD_DISPATCH
}

extern "C" int b200_emb_bwd_apply(void* const* d_weights, float* const* d_state, uint16_t* const* d_trail,
                                  const int64_t* d_row_base, int64_t nnz, const void* d_grad, int64_t T,
                                  int64_t B, int64_t D, int64_t grad_stride_table,
                                  int64_t grad_stride_bag, int w_dtype, int grad_dtype, int optimizer,
                                  float lr, const float* d_lr, float eps, void* ws, size_t ws_bytes,
                                  b200_stream_t stream) {
  B200_REQUIRE(d_grad != nullptr || nnz == 0, B200_EINVAL, "emb_bwd_apply: null gradient");
  return emb_bwd_apply_impl(d_weights, d_state, d_trail, d_row_base, nnz, d_grad, T, B, D, grad_stride_table,
                            grad_stride_bag, w_dtype, grad_dtype, optimizer, lr, d_lr, eps, ws, ws_bytes, stream,
                            nullptr, 0, 0, 0);
}

extern "C" int b200_emb_bwd_apply_sharded(void* const* d_weights, float* const* d_state,
                                          uint16_t* const* d_trail, const int64_t* d_row_base, int64_t nnz,
                                          const void* const* d_grad_ptrs, int64_t T, int64_t B, int64_t D,
                                          int64_t lN, int64_t F, int64_t feat0, int w_dtype, int grad_dtype,
                                          int optimizer, float lr, const float* d_lr, float eps, void* ws,
                                          size_t ws_bytes, b200_stream_t stream) {
  B200_REQUIRE(d_grad_ptrs != nullptr, B200_EINVAL, "emb_bwd_apply_sharded: null pointer table");
  return emb_bwd_apply_impl(d_weights, d_state, d_trail, d_row_base, nnz, nullptr, T, B, D, 0, 0, w_dtype,
                            grad_dtype, optimizer, lr, d_lr, eps, ws, ws_bytes, stream, d_grad_ptrs, lN, F, feat0);
}

extern "C" int b200_emb_bwd_update(void* const* d_weights, float* const* d_state, uint16_t* const* d_trail,
                                   const int64_t* d_rows, const int64_t* d_row_base,
                                   const int64_t* d_indices, const int64_t* d_offsets, int64_t nnz,
                                   const void* d_grad, int64_t T, int64_t B, int64_t D,
                                   int64_t grad_stride_table, int64_t grad_stride_bag, int w_dtype,
                                   int grad_dtype, int optimizer, float lr, const float* d_lr,
                                   float eps, int64_t total_rows, void* ws, size_t ws_bytes,
                                   int32_t* d_error, b200_stream_t stream) {
  int rc = b200_emb_bwd_prepare(d_rows, d_row_base, d_indices, d_offsets, nnz, T, B, D, total_rows, ws,
                                ws_bytes, d_error, stream);
  if (rc) return rc;
  return b200_emb_bwd_apply(d_weights, d_state, d_trail, d_row_base, nnz, d_grad, T, B, D, grad_stride_table,
                            grad_stride_bag, w_dtype, grad_dtype, optimizer, lr, d_lr, eps, ws, ws_bytes,
                            stream);
}

// segment sums of the gradient rows into `d_values`: row u of a compact [U, D] array, or (by_key) row unique[u]
// of a dense [total_rows, D] array.  The workspace must hold the result of b200_emb_bwd_prepare.
static int store_run_sums(const int64_t* d_row_base, int64_t nnz, const void* d_grad, int64_t T, int64_t B, int64_t D,
                          int64_t grad_stride_table, int64_t grad_stride_bag, int grad_dtype, float* d_values,
                          int by_key, void* ws, cudaStream_t st) {
  BwdLayout L = bwd_layout(ws, nnz, D);
  UpdParams p{};
  p.row_base = d_row_base;
  p.unique = L.unique;
  p.seg = L.seg;
  p.svals = L.svals;
  p.num_unique = L.num_unique;
  p.grad = d_grad;
  p.grad_stride_table = grad_stride_table;
  p.grad_stride_bag = grad_stride_bag;
  p.B = (uint32_t)B;
  p.D = (int)D;
  p.V = (int)(D * 4 / 16);
  p.skeys = L.skeys;
  p.win_run = L.win_run;
  p.first_val = L.first_val;
  p.partial = L.partial;
  p.values = d_values;
  p.by_key = by_key;
  p.n = nnz;
  p.d_n = L.n_valid;
  p.blk_cnt = L.blk_cnt;
  p.sched = L.blk_cnt + ceil_div(ceil_div(nnz, HOT_CH), MIN_NG) + 1;
  p.persist = 0;
  p.kflush = 16;
  p.T = (int)T;
  const int gsz = dtype_size(grad_dtype);
  const bool vec_ok = (D * 4) % 16 == 0 && p.V <= 128 && (reinterpret_cast<uintptr_t>(d_grad) % 16) == 0 &&
                      (reinterpret_cast<uintptr_t>(d_values) % 16) == 0 && (grad_stride_table * gsz) % 16 == 0 &&
                      (grad_stride_bag * gsz) % 16 == 0;
  return grad_dtype == B200_F32 ? launch_update<float, float, OPT_STORE>(p, nnz, vec_ok, st)
                                : launch_update<float, __nv_bfloat16, OPT_STORE>(p, nnz, vec_ok, st);
}

extern "C" int b200_emb_bwd_dense_grad(const int64_t* d_row_base, int64_t nnz, const void* d_grad, int64_t T,
                                       int64_t B, int64_t D, int64_t grad_stride_table, int64_t grad_stride_bag,
                                       int grad_dtype, float* d_dense_grad, void* ws, size_t ws_bytes,
                                       b200_stream_t stream) {
  B200_REQUIRE(T >= 0 && B >= 0 && D > 0 && nnz >= 0, B200_EINVAL, "emb_bwd_dense_grad: negative size");
  if (T == 0 || B == 0 || nnz == 0) return B200_OK;
  B200_REQUIRE(d_row_base && d_grad && d_dense_grad && ws, B200_EINVAL, "emb_bwd_dense_grad: null pointer");
  B200_REQUIRE(grad_dtype == B200_F32 || grad_dtype == B200_BF16, B200_EINVAL, "emb_bwd_dense_grad: bad dtype");
  B200_REQUIRE(ws_bytes >= b200_emb_bwd_workspace_bytes(nnz, D), B200_EWORKSPACE,
               "emb_bwd_dense_grad: workspace too small");
  return store_run_sums(d_row_base, nnz, d_grad, T, B, D, grad_stride_table, grad_stride_bag, grad_dtype,
                        d_dense_grad, 1, ws, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int b200_emb_bwd_coalesce(const int64_t* d_rows, const int64_t* d_row_base, const int64_t* d_indices,
                                     const int64_t* d_offsets, int64_t nnz, const void* d_grad, int64_t T,
                                     int64_t B, int64_t D, int64_t grad_stride_table, int64_t grad_stride_bag,
                                     int grad_dtype, int64_t total_rows, int64_t* d_unique_rows, float* d_values,
                                     int32_t* d_num_unique, void* ws, size_t ws_bytes, int32_t* d_error,
                                     b200_stream_t stream) {
  B200_REQUIRE(d_num_unique, B200_EINVAL, "emb_bwd_coalesce: null output");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (T == 0 || B == 0 || nnz == 0) {
    cudaMemsetAsync(d_num_unique, 0, sizeof(int32_t), st);
    return B200_OK;
  }
  B200_REQUIRE(d_unique_rows && d_values && d_grad, B200_EINVAL, "emb_bwd_coalesce: null pointer");
  B200_REQUIRE(grad_dtype == B200_F32 || grad_dtype == B200_BF16, B200_EINVAL, "emb_bwd_coalesce: bad dtype");
  int rc = b200_emb_bwd_prepare(d_rows, d_row_base, d_indices, d_offsets, nnz, T, B, D, total_rows, ws, ws_bytes,
                                d_error, stream);
  if (rc) return rc;
  rc = store_run_sums(d_row_base, nnz, d_grad, T, B, D, grad_stride_table, grad_stride_bag, grad_dtype, d_values, 0,
                      ws, st);
  if (rc) return rc;
  BwdLayout L = bwd_layout(ws, nnz, D);
  unique_to_i64_kernel<<<(unsigned)ceil_div(nnz, 256), 256, 0, st>>>(L.unique, L.num_unique, d_unique_rows, nnz);
  cudaMemcpyAsync(d_num_unique, L.num_unique, sizeof(int32_t), cudaMemcpyDeviceToDevice, st);
  B200_CHECK_LAUNCH("emb_bwd_coalesce");
  return B200_OK;
}

// ---- the three names SURVEY.md section 8b lists for this boundary, as thin forms of the entry points above
extern "C" size_t b200_workspace_bytes(int64_t nnz, int64_t D) { return b200_emb_bwd_workspace_bytes(nnz, D); }

extern "C" int b200_emb_bwd_sgd(void* const* d_weights, uint16_t* const* d_trail, const int64_t* d_rows,
                                const int64_t* d_row_base, const int64_t* d_indices, const int64_t* d_offsets,
                                int64_t nnz, const void* d_grad, int64_t T, int64_t B, int64_t D,
                                int64_t grad_stride_table, int64_t grad_stride_bag, int w_dtype, int grad_dtype,
                                float lr, const float* d_lr, int64_t total_rows, void* ws, size_t ws_bytes,
                                int32_t* d_error, b200_stream_t stream) {
  return b200_emb_bwd_update(d_weights, nullptr, d_trail, d_rows, d_row_base, d_indices, d_offsets, nnz, d_grad, T, B,
                             D, grad_stride_table, grad_stride_bag, w_dtype, grad_dtype, B200_OPT_SGD, lr, d_lr, 0.f,
                             total_rows, ws, ws_bytes, d_error, stream);
}

extern "C" int b200_emb_bwd_adagrad(void* const* d_weights, float* const* d_state, uint16_t* const* d_trail,
                                    const int64_t* d_rows, const int64_t* d_row_base, const int64_t* d_indices,
                                    const int64_t* d_offsets, int64_t nnz, const void* d_grad, int64_t T, int64_t B,
                                    int64_t D, int64_t grad_stride_table, int64_t grad_stride_bag, int w_dtype,
                                    int grad_dtype, float lr, const float* d_lr, float eps, int64_t total_rows,
                                    void* ws, size_t ws_bytes, int32_t* d_error, b200_stream_t stream) {
  return b200_emb_bwd_update(d_weights, d_state, d_trail, d_rows, d_row_base, d_indices, d_offsets, nnz, d_grad, T, B,
                             D, grad_stride_table, grad_stride_bag, w_dtype, grad_dtype, B200_OPT_ADAGRAD, lr, d_lr,
                             eps, total_rows, ws, ws_bytes, d_error, stream);
}
