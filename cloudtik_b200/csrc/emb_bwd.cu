// EmbeddingBag backward fused with the SGD / Adagrad row update (sm_100a).
//
// Replaces autograd's sparse EmbeddingBag backward + torch.optim.SGD.step on the 26 sparse
// gradients (reference P/dlrm_s_pytorch.py:1055,1059) -- see include/b200dlrm.h.
//
//   1. build_keys : key[j] = global row id (table-local index + row base), val[j] = global
//                   bag id t*B+b   (coalesced; bag of j found from the CSR offsets)
//   2. sort_pairs / dedup_runs (sort_dedup.cu): stable radix sort, unique rows, run starts
//   3. update     : one sub-warp group per unique row: sum grad rows of the run (fp32, in
//                   bag order), then ONE read-modify-write of the weight row (and of the
//                   Adagrad state row).  The [nnz, D] COO gradient is never materialised.
#include <algorithm>

#include "common.cuh"
#include "sort_dedup.cuh"

namespace b200 {

constexpr int KEYS_THREADS = 256;
constexpr int UPD_THREADS = 256;

// Runs longer than HOT_CH positions ("hot" rows: tiny tables, Zipf heads) are not summed by one
// group serially: the sorted array is cut into windows of HOT_CH positions, every window
// pre-reduces the (at most two) hot runs that cross it into fp32 partial rows, and the update
// kernel adds the partials of a hot row in window order.  Deterministic, no atomics.
struct BwdLayout {
  uint32_t *keys, *vals, *skeys, *svals, *unique, *seg, *win_run;
  int32_t* num_unique;
  float* partial;  // [2 * ceil(n / HOT_CH)][D]
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
  L.seg = reinterpret_cast<uint32_t*>(w + o), o += align_up((size_t)(n + 1) * 4, 256);
  L.num_unique = reinterpret_cast<int32_t*>(w + o), o += 256;
  L.win_run = reinterpret_cast<uint32_t*>(w + o), o += align_up((size_t)(ceil_div(n, HOT_CH) + 1) * 4, 256);
  L.partial = reinterpret_cast<float*>(w + o), o += align_up((size_t)(2 * ceil_div(n, HOT_CH)) * D * 4, 256);
  L.sort_ws = w + o, o += sort_ws_bytes(n);
  L.total = o;
  return L;
}

// ------------------------------------------------------------------ 1. keys
// grid = (ceil(B/256), T); one warp per 32 consecutive bags of one table.
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
  float* partial;
  int64_t n;
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
  uint2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  out[0] = bf16lo(v.x);
  out[1] = bf16hi(v.x);
  out[2] = bf16lo(v.y);
  out[3] = bf16hi(v.y);
}

// One group per window of HOT_CH sorted positions: partial sums of the hot runs crossing it.
// slot 0 = the run that contains the window's first position, slot 1 = a hot run that starts
// inside the window (a window cannot hold a third: hot runs are longer than a window).
// The run containing each window start comes from the dedup stage (win_run), so no search.
template <typename GT, int VE, int G, int NV>
__global__ void __launch_bounds__(UPD_THREADS) hot_window_kernel(const UpdParams p) {
  constexpr int NG = UPD_THREADS / G;
  constexpr int CH = 16;  // gradient rows in flight per lane
  __shared__ uint32_t s_bag[NG][HOT_CH];
  const int gid = threadIdx.x / G, lane = threadIdx.x % G;
  const int64_t w = (int64_t)blockIdx.x * NG + gid;
  const int64_t a = w * HOT_CH;
  if (a >= p.n) return;
  const int64_t b = min(a + (int64_t)HOT_CH, p.n);
  const bool cont_before = a > 0 && p.skeys[a - 1] == p.skeys[a];
  const bool cont_after = b < p.n && p.skeys[b - 1] == p.skeys[b];
  if (!cont_before && !cont_after) return;  // no run longer than a window can touch this one
  const uint32_t gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << ((threadIdx.x & 31) & ~(G - 1)));

  auto reduce_range = [&](int64_t lo, int64_t hi, int slot) {
    const int cnt = (int)(hi - lo);
    for (int i = lane; i < cnt; i += G) s_bag[gid][i] = p.svals[lo + i];  // coalesced
    __syncwarp(gmask);
    const uint32_t tt = s_bag[gid][0] / p.B;
    const GT* gt = static_cast<const GT*>(p.grad) + (int64_t)tt * p.grad_stride_table - (int64_t)(tt * p.B) * p.grad_stride_bag;
    float acc[NV][VE];
#pragma unroll
    for (int k = 0; k < NV; ++k)
#pragma unroll
      for (int x = 0; x < VE; ++x) acc[k][x] = 0.f;
    for (int j = 0; j < cnt; j += CH) {
      float g[CH][NV][VE];
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const bool ok = j + c < cnt;
        const GT* gp = gt + (int64_t)(ok ? s_bag[gid][j + c] : tt * p.B) * p.grad_stride_bag;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          const int vi = lane + k * G;
          if (ok && vi < p.V) {
            load_grad<GT, VE>(gp + vi * VE, g[c][k]);
          } else {
#pragma unroll
            for (int x = 0; x < VE; ++x) g[c][k][x] = 0.f;
          }
        }
      }
#pragma unroll
      for (int c = 0; c < CH; ++c)
#pragma unroll
        for (int k = 0; k < NV; ++k)
#pragma unroll
          for (int x = 0; x < VE; ++x) acc[k][x] += g[c][k][x];
    }
    float* dst = p.partial + (size_t)(2 * w + slot) * p.D;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int vi = lane + k * G;
      if (vi < p.V) {
#pragma unroll
        for (int h = 0; h < VE / 4; ++h) store_vec<float, 4, false>(dst + vi * VE + 4 * h, acc[k] + 4 * h);
      }
    }
    __syncwarp(gmask);
  };
  const int64_t ua = p.win_run[w];
  const int64_t sa = p.seg[ua], ea = p.seg[ua + 1];
  if (ea - sa > HOT_CH) reduce_range(a, min(b, ea), 0);
  if (ea < b) {
    const int64_t ub = (b < p.n) ? (int64_t)p.win_run[w + 1] - (p.skeys[b] != p.skeys[b - 1] ? 1 : 0)
                                 : (int64_t)(*p.num_unique) - 1;
    const int64_t sb = p.seg[ub], eb = p.seg[ub + 1];
    if (eb - sb > HOT_CH) reduce_range(sb, b, 1);
  }
}

template <typename WT, typename GT, int G, int NV, int OPT>
__global__ void __launch_bounds__(UPD_THREADS) emb_update_kernel(const UpdParams p) {
  constexpr int VE = Vec16<WT>::VE;
  constexpr int NG = UPD_THREADS / G;
  constexpr int R = 4;  // unique rows in flight per group
  const int gid = threadIdx.x / G, lane = threadIdx.x % G;
  const int64_t nU = *p.num_unique;
  const int64_t u0 = (int64_t)blockIdx.x * (NG * R);
  if (u0 >= nU) return;
  const float lr = p.d_lr ? *p.d_lr : p.lr;

  uint32_t s[R], e[R], t[R];
  char* wrow[R];
  float* srow[R];
  bool valid[R];
  bool single = true;
#pragma unroll
  for (int q = 0; q < R; ++q) {
    const int64_t u = u0 + q * NG + gid;
    valid[q] = u < nU;
    s[q] = valid[q] ? p.seg[u] : 0;
    e[q] = valid[q] ? p.seg[u + 1] : 0;
    const uint32_t key = valid[q] ? p.unique[u] : 0;
    const uint32_t g0 = valid[q] ? p.svals[s[q]] : 0;
    t[q] = g0 / p.B;
    const int64_t row = valid[q] ? (int64_t)key - p.row_base[t[q]] : 0;
    wrow[q] = valid[q] ? static_cast<char*>(p.weights[t[q]]) + (size_t)row * p.D * sizeof(WT) : nullptr;
    if (OPT == B200_OPT_ADAGRAD) srow[q] = valid[q] ? p.state[t[q]] + (size_t)row * p.D : nullptr;
    single = single && (e[q] - s[q] <= 1);
  }

  float acc[R][NV][VE];
#pragma unroll
  for (int q = 0; q < R; ++q)
#pragma unroll
    for (int k = 0; k < NV; ++k)
#pragma unroll
      for (int x = 0; x < VE; ++x) acc[q][k][x] = 0.f;

  auto grad_row = [&](uint32_t g, uint32_t tt) -> const GT* {
    const uint32_t b = g - tt * p.B;
    return static_cast<const GT*>(p.grad) + (int64_t)tt * p.grad_stride_table + (int64_t)b * p.grad_stride_bag;
  };

  if (single) {
    // every run has one element (the common one-hot, mostly-unique case): all loads first
#pragma unroll
    for (int q = 0; q < R; ++q) {
      if (valid[q]) {
        const GT* gp = grad_row(p.svals[s[q]], t[q]);
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          const int vi = lane + k * G;
          if (vi < p.V) load_grad<GT, VE>(gp + vi * VE, acc[q][k]);
        }
      }
    }
  } else {
    // hot rows first: add their per-window partial sums (hot_window_kernel) in window order
    uint32_t j[R], je[R];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      j[q] = s[q];
      je[q] = valid[q] ? e[q] : s[q];
      if (valid[q] && e[q] - s[q] > (uint32_t)HOT_CH) {
        je[q] = s[q];  // nothing left for the walk below
        const uint32_t w0 = s[q] / HOT_CH, w1 = (e[q] - 1) / HOT_CH;
        for (uint32_t w = w0; w <= w1; ++w) {
          const int slot = (s[q] > w * HOT_CH) ? 1 : 0;
          const float* src = p.partial + (size_t)(2 * w + slot) * p.D;
#pragma unroll
          for (int k = 0; k < NV; ++k) {
            const int vi = lane + k * G;
            if (vi < p.V) {
#pragma unroll
              for (int h = 0; h < VE / 4; ++h) {
                float f[4];
                Vec16<float>::unpack(f, ld_v4(src + vi * VE + 4 * h));
#pragma unroll
                for (int x = 0; x < 4; ++x) acc[q][k][4 * h + x] += f[x];
              }
            }
          }
        }
      }
    }
    // the four runs advance together: R independent (bag id -> gradient row) chains in flight,
    // each run still summed in sorted (= bag) order
    while (true) {
      bool act[R];
      bool any = false;
#pragma unroll
      for (int q = 0; q < R; ++q) {
        act[q] = j[q] < je[q];
        any = any || act[q];
      }
      if (!any) break;
      uint32_t g[R];
#pragma unroll
      for (int q = 0; q < R; ++q) g[q] = act[q] ? p.svals[j[q]] : 0u;
      float tmp[R][NV][VE];
#pragma unroll
      for (int q = 0; q < R; ++q) {
        const GT* gp = grad_row(g[q], t[q]);
#pragma unroll
        for (int k = 0; k < NV; ++k) {
          const int vi = lane + k * G;
          if (act[q] && vi < p.V) {
            load_grad<GT, VE>(gp + vi * VE, tmp[q][k]);
          } else {
#pragma unroll
            for (int x = 0; x < VE; ++x) tmp[q][k][x] = 0.f;
          }
        }
      }
#pragma unroll
      for (int q = 0; q < R; ++q) {
#pragma unroll
        for (int k = 0; k < NV; ++k)
#pragma unroll
          for (int x = 0; x < VE; ++x) acc[q][k][x] += tmp[q][k][x];
        j[q] += act[q] ? 1u : 0u;
      }
    }
  }

  // one read-modify-write per unique row
#pragma unroll
  for (int q = 0; q < R; ++q) {
    if (!valid[q]) continue;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int vi = lane + k * G;
      if (vi >= p.V) continue;
      float w[VE];
      char* wp = wrow[q] + (size_t)vi * 16;
      Vec16<WT>::unpack(w, ld_v4(wp));
      if (OPT == B200_OPT_ADAGRAD) {
        float* sp = srow[q] + vi * VE;
#pragma unroll
        for (int h = 0; h < VE / 4; ++h) {
          float st[4];
          Vec16<float>::unpack(st, ld_v4(sp + 4 * h));
#pragma unroll
          for (int x = 0; x < 4; ++x) {
            const float g = acc[q][k][4 * h + x];
            st[x] += g * g;
            w[4 * h + x] -= lr * g / (sqrtf(st[x]) + p.eps);
          }
          store_vec<float, 4, false>(sp + 4 * h, st);
        }
      } else {
#pragma unroll
        for (int x = 0; x < VE; ++x) w[x] -= lr * acc[q][k][x];
      }
      store_vec<WT, VE, false>(reinterpret_cast<WT*>(wp), w);
    }
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
    const uint32_t t = p.svals[s] / p.B;
    const int64_t row = (int64_t)p.unique[u] - p.row_base[t];
    float acc = 0.f;
    for (uint32_t j = s; j < e; ++j) {
      const uint32_t g = p.svals[j];
      const uint32_t b = g - t * p.B;
      acc += to_f32<GT>(static_cast<const GT*>(p.grad)[(int64_t)t * p.grad_stride_table +
                                                      (int64_t)b * p.grad_stride_bag + d]);
    }
    WT* w = static_cast<WT*>(p.weights[t]) + row * p.D + d;
    float wv = to_f32<WT>(*w);
    if (OPT == B200_OPT_ADAGRAD) {
      float* sp = p.state[t] + row * p.D + d;
      const float st = *sp + acc * acc;
      *sp = st;
      wv -= lr * acc / (sqrtf(st) + p.eps);
    } else {
      wv -= lr * acc;
    }
    *w = from_f32<WT>(wv);
  }
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
  constexpr int VE = Vec16<WT>::VE;
  const int64_t nwin = ceil_div(n, HOT_CH);
#define B200_UPD_CASE(G, NV)                                                                       \
  hot_window_kernel<GT, VE, G, NV><<<(unsigned)ceil_div(nwin, UPD_THREADS / G), UPD_THREADS, 0, st>>>(p); \
  emb_update_kernel<WT, GT, G, NV, OPT>                                                            \
      <<<(unsigned)ceil_div(n, (UPD_THREADS / G) * 4), UPD_THREADS, 0, st>>>(p)
  if (V <= 1) { B200_UPD_CASE(1, 1); }
  else if (V <= 2) { B200_UPD_CASE(2, 1); }
  else if (V <= 4) { B200_UPD_CASE(4, 1); }
  else if (V <= 8) { B200_UPD_CASE(8, 1); }
  else if (V <= 16) { B200_UPD_CASE(16, 1); }
  else if (V <= 32) { B200_UPD_CASE(32, 1); }
  else if (V <= 64) { B200_UPD_CASE(32, 2); }
  else { B200_UPD_CASE(32, 4); }
#undef B200_UPD_CASE
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

extern "C" int b200_emb_bwd_prepare(const int64_t* d_rows, const int64_t* d_row_base,
                                    const int64_t* d_indices, const int64_t* d_offsets, int64_t nnz,
                                    int64_t T, int64_t B, int64_t D, int64_t total_rows, void* ws, size_t ws_bytes,
                                    int32_t* d_error, b200_stream_t stream) {
  int rc = check_bwd_common(T, B, D, nnz, total_rows);
  if (rc) return rc;
  if (T == 0 || B == 0 || nnz == 0) return B200_OK;
  B200_REQUIRE(d_rows && d_row_base && d_indices && d_offsets && ws, B200_EINVAL,
               "emb_bwd_prepare: null pointer");
  B200_REQUIRE(ws_bytes >= b200_emb_bwd_workspace_bytes(nnz, D), B200_EWORKSPACE,
               "emb_bwd: workspace too small");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  BwdLayout L = bwd_layout(ws, nnz, D);
  dim3 kgrid((unsigned)ceil_div(B, KEYS_THREADS), (unsigned)T);
  build_keys_kernel<<<kgrid, KEYS_THREADS, 0, st>>>(d_indices, d_offsets, d_rows, d_row_base, B, L.keys,
                                                    L.vals, d_error);
  B200_CHECK_LAUNCH("build_keys");
  rc = sort_pairs(L.keys, L.vals, nnz, (uint64_t)total_rows, L.skeys, L.svals, L.sort_ws, st);
  if (rc) return rc;
  return dedup_runs(L.skeys, nnz, L.unique, L.seg, L.num_unique, L.sort_ws, st, L.win_run);
}

extern "C" int b200_emb_bwd_apply(void* const* d_weights, float* const* d_state,
                                  const int64_t* d_row_base, int64_t nnz, const void* d_grad, int64_t T,
                                  int64_t B, int64_t D, int64_t grad_stride_table,
                                  int64_t grad_stride_bag, int w_dtype, int grad_dtype, int optimizer,
                                  float lr, const float* d_lr, float eps, void* ws, size_t ws_bytes,
                                  b200_stream_t stream) {
  B200_REQUIRE(T >= 0 && B >= 0 && D > 0 && nnz >= 0, B200_EINVAL, "emb_bwd: negative size");
  if (T == 0 || B == 0 || nnz == 0) return B200_OK;
  B200_REQUIRE(d_weights && d_row_base && d_grad && ws, B200_EINVAL, "emb_bwd_apply: null pointer");
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
  p.partial = L.partial;
  p.n = nnz;
  const int wsz = dtype_size(w_dtype), gsz = dtype_size(grad_dtype);
  p.V = (int)(D * wsz / 16);
  const int ve = 16 / wsz;  // elements per weight vector
  const bool vec_ok = (D * wsz) % 16 == 0 && p.V <= 128 &&
                      (reinterpret_cast<uintptr_t>(d_grad) % 16) == 0 &&
                      (grad_stride_table * gsz) % 16 == 0 && (grad_stride_bag * gsz) % 16 == 0 &&
                      (ve * gsz) % 8 == 0;
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

extern "C" int b200_emb_bwd_update(void* const* d_weights, float* const* d_state,
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
  return b200_emb_bwd_apply(d_weights, d_state, d_row_base, nnz, d_grad, T, B, D, grad_stride_table,
                            grad_stride_bag, w_dtype, grad_dtype, optimizer, lr, d_lr, eps, ws, ws_bytes,
                            stream);
}
