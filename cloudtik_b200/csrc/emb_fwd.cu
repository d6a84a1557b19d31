// Sum-pooled EmbeddingBag forward, all tables in one launch (sm_100a).
//
// Replaces the 26-iteration Python loop over nn.EmbeddingBag in DLRM_Net.apply_emb
// (reference P/dlrm_s_pytorch.py:290-305) -- see include/b200dlrm.h.
//
// Mapping: grid = (tiles of `tile` consecutive bags, T tables).  A CTA stages its tile of
// CSR offsets and (when it fits) the tile's window of indices into shared memory with 1-D
// bulk async copies (TMA engine, mbarrier completion), then sub-warp groups of G lanes own
// one bag at a time: lane l of a group moves the 16-byte vectors l, l+G, ... of each row
// (G*16 bytes coalesced per row).  The staged tile is first compacted to 32-bit offsets / row
// ids (bounds check done once there).  One-hot tiles take a copy-through path with four bags in
// flight per group; pooled bags accumulate in fp32 in index order, four rows in flight.
#include <stdlib.h>

#include <algorithm>

#include "common.cuh"

namespace b200 {

constexpr int FWD_THREADS = 256;
constexpr int FWD_MAX_TILE = 128;  // bags per CTA (upper bound; host picks <= this)
constexpr int FWD_IDX_CAP = 2048;  // staged index window, int64 entries (16 KB)

struct EmbFwdParams {
  const void* const* weights;
  const int64_t* rows;
  const int64_t* indices;
  const int64_t* offsets;
  void* out;
  int64_t nnz;
  int64_t B;
  int64_t out_stride_table;
  int64_t out_stride_bag;
  int32_t* error;
  int D;
  int V;     // 16-byte vectors per row
  int tile;  // bags per CTA
  // sharded output (fused pooled-embedding all-to-all over peer memory): bag b of local table t is
  // written to rank r = b / lN at out_ptrs[r] + ((b - r*lN) * F + feat0 + t) * D.  lN % tile == 0.
  void* const* out_ptrs;
  int64_t lN;
  int64_t F;
  int64_t feat0;
  const float* const* scales;  // int8 tables: per-table pointer to the fp32 row scales (else unused)
  // accumulate != 0: ADD the pooled row into the output (vector reductions; the caller zeroed it) instead of
  // storing it, and skip bags that hold no index -- row-wise sharded tables, where every rank contributes the
  // partial sum of the rows it owns to the same slot of the destination rank's interaction input
  int accumulate;
  // sharded output: bag tiles are visited in an order rotated by this many tiles, so that the ranks of a job do not
  // all write to the same destination rank at the same time (rank r starts with destination r + 1)
  int tile_rot;
};

// acc += row (fp32 / bf16 tables) or acc += scale * row (int8 tables)
template <typename WT>
__device__ __forceinline__ void accum_row(float* acc, uint4 v, float s) {
  if constexpr (sizeof(WT) == 1) {
    float f[16];
    Vec16<WT>::unpack(f, v);
#pragma unroll
    for (int x = 0; x < 16; ++x) acc[x] = fmaf(s, f[x], acc[x]);
  } else {
    Vec16<WT>::accum(acc, v);
  }
}

template <typename WT, typename OT, int G, int NV>
__global__ void __launch_bounds__(FWD_THREADS) emb_fwd_kernel(const EmbFwdParams p) {
  constexpr int VE = Vec16<WT>::VE;
  constexpr int NG = FWD_THREADS / G;
  __shared__ __align__(16) int64_t s_off[FWD_MAX_TILE + 2];
  __shared__ __align__(16) int64_t s_idx[FWD_IDX_CAP + 2];
  __shared__ __align__(8) uint64_t s_bar;

  const int tid = threadIdx.x;
  const int t = blockIdx.y;
  const int64_t b0 = (int64_t)((blockIdx.x + (unsigned)p.tile_rot) % gridDim.x) * p.tile;
  const int nb = (int)min((int64_t)p.tile, p.B - b0);
  const int64_t* offp = p.offsets + (int64_t)t * p.B + b0;
  const int n_off = nb + 1;

  // ---- stage the offsets tile (bulk copy when 16-byte aligned)
  const bool off_al = (reinterpret_cast<uintptr_t>(offp) & 15) == 0;
  if (tid == 0) {
    mbar_init(&s_bar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (off_al) {
    const int nbulk = n_off & ~1;
    if (tid == 0) {
      mbar_expect_tx(&s_bar, nbulk * 8);
      bulk_g2s(s_off, offp, nbulk * 8, &s_bar);
    }
    if (tid == 32 && (n_off & 1)) s_off[n_off - 1] = offp[n_off - 1];
    mbar_wait(&s_bar, 0);
  } else {
    for (int i = tid; i < n_off; i += FWD_THREADS) s_off[i] = offp[i];
  }
  __syncthreads();

  // ---- stage the index window of this tile
  const int64_t i0 = s_off[0], i1 = s_off[nb];
  const int64_t a0 = i0 & ~(int64_t)1;
  const int64_t a1 = min((i1 + 1) & ~(int64_t)1, p.nnz);
  const bool idx_al = (reinterpret_cast<uintptr_t>(p.indices) & 15) == 0;
  const bool staged = idx_al && i1 > i0 && (a1 - a0) <= FWD_IDX_CAP && i0 >= 0 && i1 <= p.nnz;
  if (staged) {
    const int cnt = (int)(a1 - a0);
    const int nbulk = cnt & ~1;
    if (tid == 0 && nbulk > 0) {
      mbar_expect_tx(&s_bar, nbulk * 8);
      bulk_g2s(s_idx, p.indices + a0, nbulk * 8, &s_bar);
    }
    if (tid == 32 && (cnt & 1)) s_idx[cnt - 1] = p.indices[a0 + cnt - 1];
    if (nbulk > 0) mbar_wait(&s_bar, off_al ? 1 : 0);
    __syncthreads();
  }

  const int gid = tid / G;
  const int lane = tid % G;
  const int64_t nrows = p.rows[t];
  constexpr bool Q8 = sizeof(WT) == 1;
  const float* S = Q8 ? p.scales[t] : nullptr;  // row scales of an int8 table
  const char* W = static_cast<const char*>(p.weights[t]);
  const size_t row_bytes = (size_t)p.D * sizeof(WT);
  OT* out_t;
  if (p.out_ptrs != nullptr) {  // this tile's bags all belong to one destination rank
    const int64_t r = b0 / p.lN;
    out_t = static_cast<OT*>(p.out_ptrs[r]) + ((b0 - r * p.lN) * p.F + p.feat0 + t) * p.D;
  } else {
    out_t = static_cast<OT*>(p.out) + (int64_t)t * p.out_stride_table + b0 * p.out_stride_bag;
  }

  // ---- compact the staged tile to 32 bits: bag offsets relative to i0, row ids (bounds-checked
  //      once here, 0xffffffff = out of range) -- keeps 64-bit arithmetic out of the gather loop
  __shared__ int s_rel[FWD_MAX_TILE + 1];
  __shared__ uint32_t s_row[FWD_IDX_CAP + 2];
  const bool staged32 = staged && nrows < 0xffffffffll;
  bool all_one = true;
  if (staged32) {
    for (int i = tid; i <= nb; i += FWD_THREADS) s_rel[i] = (int)(s_off[i] - i0);
    for (int i = tid; i < nb; i += FWD_THREADS) all_one = all_one && (s_off[i + 1] - s_off[i] == 1);
    const int cnt = (int)(i1 - i0), shift = (int)(i0 - a0);
    for (int k = tid; k < cnt; k += FWD_THREADS) {
      const int64_t idx = s_idx[k + shift];
      const bool ok = (uint64_t)idx < (uint64_t)nrows;
      if (!ok && p.error) *p.error = 1;
      s_row[k] = ok ? (uint32_t)idx : 0xffffffffu;
    }
  }
  const bool fast = __syncthreads_and(staged32 && all_one) != 0;

  const char* Wl = W + (size_t)lane * 16;
  OT* outl = out_t + lane * VE;

  if (fast) {
    // every bag of the tile holds exactly one index (Criteo one-hot): bag k reads row s_row[k]
    for (int r = gid; r < nb; r += 4 * NG) {
      uint32_t row[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) row[u] = (r + u * NG < nb) ? s_row[r + u * NG] : 0xffffffffu;
      uint4 v[4][NV];
      float sc[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const char* src = Wl + (size_t)row[u] * row_bytes;
        sc[u] = (Q8 && row[u] != 0xffffffffu) ? __ldg(S + row[u]) : 1.f;
#pragma unroll
        for (int k = 0; k < NV; ++k)
          v[u][k] = (row[u] != 0xffffffffu && lane + k * G < p.V) ? ld_nc_v4(src + (size_t)k * G * 16)
                                                                 : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (r + u * NG < nb) {
          OT* dst = outl + (int64_t)(r + u * NG) * p.out_stride_bag;
#pragma unroll
          for (int k = 0; k < NV; ++k) {
            if (lane + k * G < p.V) {
              if (p.accumulate) {
                if (row[u] != 0xffffffffu) {
                  float f[VE];
                  Vec16<WT>::unpack(f, v[u][k]);
                  if constexpr (Q8) {
#pragma unroll
                    for (int x = 0; x < VE; ++x) f[x] *= sc[u];
                  }
                  red_vec<OT, VE>(dst + k * G * VE, f);
                }
              } else if constexpr (sizeof(WT) == sizeof(OT)) {
                st_cs_v4(dst + k * G * VE, v[u][k]);
              } else {
                float f[VE];
                Vec16<WT>::unpack(f, v[u][k]);
                if constexpr (Q8) {
#pragma unroll
                  for (int x = 0; x < VE; ++x) f[x] *= sc[u];
                }
                store_vec<OT, VE, true>(dst + k * G * VE, f);
              }
            }
          }
        }
      }
    }
    return;
  }

  if (staged32) {
    // pooled / ragged bags, indices staged: fp32 accumulation in index order, 4 rows in flight
    for (int bag = gid; bag < nb; bag += NG) {
      const int sb = s_rel[bag], eb = s_rel[bag + 1];
      float acc[NV][VE];
#pragma unroll
      for (int k = 0; k < NV; ++k)
#pragma unroll
        for (int x = 0; x < VE; ++x) acc[k][x] = 0.f;
      for (int j = sb; j < eb; j += 4) {
        uint4 v[4][NV];
        float sc[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint32_t row = (j + q < eb) ? s_row[j + q] : 0xffffffffu;
          const char* src = Wl + (size_t)row * row_bytes;
          sc[q] = (Q8 && row != 0xffffffffu) ? __ldg(S + row) : 0.f;
#pragma unroll
          for (int k = 0; k < NV; ++k)
            v[q][k] = (row != 0xffffffffu && lane + k * G < p.V) ? ld_nc_v4(src + (size_t)k * G * 16)
                                                               : make_uint4(0, 0, 0, 0);
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
          for (int k = 0; k < NV; ++k) accum_row<WT>(acc[k], v[q][k], sc[q]);
      }
      OT* dst = outl + (int64_t)bag * p.out_stride_bag;
      if (p.accumulate) {
        if (eb > sb) {
#pragma unroll
          for (int k = 0; k < NV; ++k)
            if (lane + k * G < p.V) red_vec<OT, VE>(dst + k * G * VE, acc[k]);
        }
        continue;
      }
#pragma unroll
      for (int k = 0; k < NV; ++k)
        if (lane + k * G < p.V) store_vec<OT, VE, true>(dst + k * G * VE, acc[k]);
    }
    return;
  }

  // tile's index window does not fit the staging buffer (very long bags) or is misaligned:
  // indices come straight from global memory (one broadcast load per row)
  for (int bag = gid; bag < nb; bag += NG) {
    const int64_t sb = s_off[bag], eb = s_off[bag + 1];
    float acc[NV][VE];
#pragma unroll
    for (int k = 0; k < NV; ++k)
#pragma unroll
      for (int x = 0; x < VE; ++x) acc[k][x] = 0.f;
    for (int64_t j = sb; j < eb; j += 4) {
      uint4 v[4][NV];
      float sc[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        bool ok = false;
        int64_t idx = 0;
        if (j + q < eb) {
          idx = __ldg(p.indices + j + q);
          ok = (uint64_t)idx < (uint64_t)nrows;
          if (!ok && p.error) *p.error = 1;
        }
        sc[q] = (Q8 && ok) ? __ldg(S + idx) : 0.f;
        const char* src = Wl + (ok ? (size_t)idx : 0) * row_bytes;
#pragma unroll
        for (int k = 0; k < NV; ++k)
          v[q][k] = (ok && lane + k * G < p.V) ? ld_nc_v4(src + (size_t)k * G * 16) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int k = 0; k < NV; ++k) accum_row<WT>(acc[k], v[q][k], sc[q]);
    }
    OT* dst = outl + (int64_t)bag * p.out_stride_bag;
    if (p.accumulate) {
      if (eb > sb) {
#pragma unroll
        for (int k = 0; k < NV; ++k)
          if (lane + k * G < p.V) red_vec<OT, VE>(dst + k * G * VE, acc[k]);
      }
      continue;
    }
#pragma unroll
    for (int k = 0; k < NV; ++k)
      if (lane + k * G < p.V) store_vec<OT, VE, true>(dst + k * G * VE, acc[k]);
  }
}

// ------------------------------------------------------------------ persistent, software-pipelined variant
// grid = a few CTAs per SM; every CTA walks over (table, bag tile) work items w = blockIdx.x, + gridDim.x, ...
// While the sub-warp groups gather tile i, the TMA engine is already fetching what the next tiles need:
//   offsets of tile i+3  -> one of three offset slots   (issued by thread 0 at the top of iteration i)
//   indices of tile i+2  -> one of two index slots      (their window is known once offsets(i+2) have landed)
// and the 64 -> 32 bit compaction (bounds check, offsets relative to the window) of tile i+1 is done by all
// threads in the shadow of tile i's FIRST batch of row loads: issue the loads, compact, only then consume them.
// One __syncthreads per tile.  The one-shot kernel above spends about a third of a CTA's life in the chain
// offsets -> wait -> indices -> wait -> compaction before its first row load; here that chain is off the path.
constexpr int FWD_OSLOTS = 3;
constexpr int FWD_OFF_PAD = FWD_MAX_TILE + 4;

struct FwdTileInfo {
  int64_t i0;   // first index position of the tile's window
  int staged;   // window staged in s_idx (fits, in range)
  int shift;    // i0 - (16-byte aligned start of the bulk copy)
};

template <typename WT, typename OT, int G, int NV>
__global__ void __launch_bounds__(FWD_THREADS, 4)
    emb_fwd_persist_kernel(const EmbFwdParams p, int n_items, int tiles_per_table, int cap) {
  constexpr int VE = Vec16<WT>::VE;
  constexpr int NG = FWD_THREADS / G;
  constexpr bool Q8 = sizeof(WT) == 1;
  constexpr int UF = NV == 1 ? 8 : 4;  // rows in flight per group on the one-hot path
  extern __shared__ __align__(16) unsigned char fwd_smem[];
  // carve-up (all 16-byte aligned): s_off[3][OFF_PAD] i64 | s_idx[2][cap+2] i64 | s_off64[2][MAX_TILE+2] i64 |
  //                                 s_row[2][cap+2] u32 | s_rel[2][MAX_TILE+4] i32
  int64_t* s_off = reinterpret_cast<int64_t*>(fwd_smem);
  int64_t* s_idx = s_off + FWD_OSLOTS * FWD_OFF_PAD;
  int64_t* s_off64 = s_idx + 2 * (cap + 2);
  uint32_t* s_row = reinterpret_cast<uint32_t*>(s_off64 + 2 * (FWD_MAX_TILE + 2));
  int* s_rel = reinterpret_cast<int*>(s_row + 2 * (cap + 2));
  __shared__ __align__(8) uint64_t bar_off[FWD_OSLOTS];
  __shared__ __align__(8) uint64_t bar_idx[2];
  __shared__ FwdTileInfo d_info[FWD_OSLOTS];
  __shared__ int d_osh[FWD_OSLOTS];     // first offset of the tile sits at s_off[slot][d_osh]
  __shared__ int d_notone[FWD_OSLOTS];  // some bag of the tile does not hold exactly one index

  const int tid = threadIdx.x, gid = tid / G, lane = tid % G;
  const int n_mine = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // blockIdx.x < n_items
  const int64_t n_off_total = (int64_t)(n_items / tiles_per_table) * p.B + 1;  // entries of the CSR array

  auto item = [&](int li, int& t, int64_t& b0, int& nb) {
    const int w = (int)blockIdx.x + li * (int)gridDim.x;
    t = w / tiles_per_table;
    b0 = (int64_t)((w - t * tiles_per_table + p.tile_rot) % tiles_per_table) * p.tile;
    nb = (int)min((int64_t)p.tile, p.B - b0);
  };
  // ---- thread 0 only: start the bulk copy of tile li's offsets
  auto issue_off = [&](int li) {
    int t, nb;
    int64_t b0;
    item(li, t, b0, nb);
    const int slot = li % FWD_OSLOTS;
    int64_t* dst = s_off + slot * FWD_OFF_PAD;
    const int64_t first = (int64_t)t * p.B + b0;
    const int64_t a = first & ~(int64_t)1;
    const int sh = (int)(first - a);
    int n = nb + 1 + sh;
    if ((n & 1) && a + n + 1 <= n_off_total) ++n;  // round the copy up to 16 bytes when the array allows it
    const int nbulk = n & ~1;
    if (n & 1) dst[n - 1] = p.offsets[a + n - 1];  // very last tile of the array only
    d_osh[slot] = sh;
    if (nbulk > 0) {
      mbar_expect_tx(&bar_off[slot], nbulk * 8);
      bulk_g2s(dst, p.offsets + a, nbulk * 8, &bar_off[slot]);
    } else {
      mbar_arrive(&bar_off[slot]);
    }
  };
  // ---- thread 0 only: offsets of tile li have landed -> start the bulk copy of its index window
  auto issue_idx = [&](int li) {
    int t, nb;
    int64_t b0;
    item(li, t, b0, nb);
    const int slot = li % FWD_OSLOTS, islot = li & 1;
    mbar_wait(&bar_off[slot], (uint32_t)(li / FWD_OSLOTS) & 1u);
    const int64_t* o = s_off + slot * FWD_OFF_PAD + d_osh[slot];
    const int64_t i0 = o[0], i1 = o[nb];
    const int64_t a0 = i0 & ~(int64_t)1;
    FwdTileInfo inf;
    inf.i0 = i0;
    inf.shift = (int)(i0 - a0);
    int64_t n = i1 - a0;
    inf.staged = (i1 > i0 && n <= cap && i0 >= 0 && i1 <= p.nnz) ? 1 : 0;
    d_info[slot] = inf;
    if (inf.staged) {
      if ((n & 1) && a0 + n + 1 <= p.nnz) ++n;
      const int nbulk = (int)n & ~1;
      int64_t* dst = s_idx + islot * (cap + 2);
      if (n & 1) dst[n - 1] = p.indices[a0 + n - 1];
      if (nbulk > 0) {
        mbar_expect_tx(&bar_idx[islot], nbulk * 8);
        bulk_g2s(dst, p.indices + a0, nbulk * 8, &bar_idx[islot]);
        return;
      }
    }
    mbar_arrive(&bar_idx[islot]);
  };
  // ---- all threads: 64 -> 32 bit compaction of tile li (its index copy was issued an iteration ago)
  auto compact = [&](int li) {
    int t, nb;
    int64_t b0;
    item(li, t, b0, nb);
    const int slot = li % FWD_OSLOTS, islot = li & 1, buf = li & 1;
    mbar_wait(&bar_idx[islot], (uint32_t)(li / 2) & 1u);
    const FwdTileInfo inf = d_info[slot];
    const int64_t* o = s_off + slot * FWD_OFF_PAD + d_osh[slot];
    const int64_t nrows = p.rows[t];
    if (inf.staged && nrows < 0xffffffffll) {
      int* rel = s_rel + buf * (FWD_MAX_TILE + 4);
      bool one = true;
      for (int i = tid; i <= nb; i += FWD_THREADS) {
        rel[i] = (int)(o[i] - inf.i0);
        if (i < nb) one = one && (o[i + 1] - o[i] == 1);
      }
      if (!one) d_notone[slot] = 1;
      const int cnt = (int)(o[nb] - inf.i0);
      const int64_t* src = s_idx + islot * (cap + 2) + inf.shift;
      uint32_t* row = s_row + buf * (cap + 2);
      for (int k = tid; k < cnt; k += FWD_THREADS) {
        const int64_t idx = src[k];
        const bool ok = (uint64_t)idx < (uint64_t)nrows;
        if (!ok && p.error) *p.error = 1;
        row[k] = ok ? (uint32_t)idx : 0xffffffffu;
      }
    } else {  // window too long for the staging buffer: the gather reads the indices from global memory
      int64_t* o64 = s_off64 + buf * (FWD_MAX_TILE + 2);
      for (int i = tid; i <= nb; i += FWD_THREADS) o64[i] = o[i];
      if (tid == 0) d_notone[slot] = 2;  // marks "unstaged"
    }
  };

  if (tid == 0) {
    for (int s_ = 0; s_ < FWD_OSLOTS; ++s_) {
      mbar_init(&bar_off[s_], 1);
      d_notone[s_] = 0;
    }
    mbar_init(&bar_idx[0], 1);
    mbar_init(&bar_idx[1], 1);
    fence_mbar_init();
    for (int li = 0; li < FWD_OSLOTS && li < n_mine; ++li) issue_off(li);
    for (int li = 0; li < 2 && li < n_mine; ++li) issue_idx(li);
  }
  __syncthreads();
  compact(0);
  __syncthreads();

  const size_t row_bytes = (size_t)p.D * sizeof(WT);
  for (int li = 0; li < n_mine; ++li) {
    if (tid == 0) {
      // slot li % 3 held tile li's offsets (compacted one iteration ago): reuse it for tile li + 3
      if (li + FWD_OSLOTS < n_mine) issue_off(li + FWD_OSLOTS);
      d_notone[(li + 2) % FWD_OSLOTS] = 0;
      if (li >= 0 && li + 2 < n_mine) issue_idx(li + 2);
    }
    int t, nb;
    int64_t b0;
    item(li, t, b0, nb);
    const int slot = li % FWD_OSLOTS, buf = li & 1;
    const int mode = d_notone[slot];  // 0: every bag holds one index, 1: ragged / pooled, 2: unstaged
    const int64_t nrows = p.rows[t];
    const float* S = Q8 ? p.scales[t] : nullptr;
    const char* Wl = static_cast<const char*>(p.weights[t]) + (size_t)lane * 16;
    OT* out_t;
    if (p.out_ptrs != nullptr) {  // this tile's bags all belong to one destination rank
      const int64_t r = b0 / p.lN;
      out_t = static_cast<OT*>(p.out_ptrs[r]) + ((b0 - r * p.lN) * p.F + p.feat0 + t) * p.D;
    } else {
      out_t = static_cast<OT*>(p.out) + (int64_t)t * p.out_stride_table + b0 * p.out_stride_bag;
    }
    OT* outl = out_t + lane * VE;
    const uint32_t* row_s = s_row + buf * (cap + 2);
    const int* rel = s_rel + buf * (FWD_MAX_TILE + 4);
    bool next_done = li + 1 >= n_mine;
    // compaction of the NEXT tile, run once by every thread in the shadow of this tile's first batch of loads
    auto mid = [&]() {
      if (!next_done) {
        compact(li + 1);
        next_done = true;
      }
    };

    if (mode == 0) {
      // every bag of the tile holds exactly one index (Criteo one-hot): bag k reads row s_row[k]
      for (int r = gid; r < nb; r += UF * NG) {
        uint32_t row[UF];
#pragma unroll
        for (int u = 0; u < UF; ++u) row[u] = (r + u * NG < nb) ? row_s[r + u * NG] : 0xffffffffu;
        uint4 v[UF][NV];
        float sc[UF];
#pragma unroll
        for (int u = 0; u < UF; ++u) {
          const char* src = Wl + (size_t)row[u] * row_bytes;
          sc[u] = (Q8 && row[u] != 0xffffffffu) ? __ldg(S + row[u]) : 1.f;
#pragma unroll
          for (int k = 0; k < NV; ++k)
            v[u][k] = (row[u] != 0xffffffffu && lane + k * G < p.V) ? ld_nc_v4(src + (size_t)k * G * 16)
                                                                   : make_uint4(0, 0, 0, 0);
        }
        mid();
#pragma unroll
        for (int u = 0; u < UF; ++u) {
          if (r + u * NG < nb) {
            OT* dst = outl + (int64_t)(r + u * NG) * p.out_stride_bag;
#pragma unroll
            for (int k = 0; k < NV; ++k) {
              if (lane + k * G < p.V) {
                if (p.accumulate) {
                  if (row[u] != 0xffffffffu) {
                    float f[VE];
                    Vec16<WT>::unpack(f, v[u][k]);
                    if constexpr (Q8) {
#pragma unroll
                      for (int x = 0; x < VE; ++x) f[x] *= sc[u];
                    }
                    red_vec<OT, VE>(dst + k * G * VE, f);
                  }
                } else if constexpr (sizeof(WT) == sizeof(OT)) {
                  st_cs_v4(dst + k * G * VE, v[u][k]);
                } else {
                  float f[VE];
                  Vec16<WT>::unpack(f, v[u][k]);
                  if constexpr (Q8) {
#pragma unroll
                    for (int x = 0; x < VE; ++x) f[x] *= sc[u];
                  }
                  store_vec<OT, VE, true>(dst + k * G * VE, f);
                }
              }
            }
          }
        }
      }
    } else if (mode == 1) {
      // pooled / ragged bags, indices staged: fp32 accumulation in index order, 4 rows in flight
      for (int bag = gid; bag < nb; bag += NG) {
        const int sb = rel[bag], eb = rel[bag + 1];
        float acc[NV][VE];
#pragma unroll
        for (int k = 0; k < NV; ++k)
#pragma unroll
          for (int x = 0; x < VE; ++x) acc[k][x] = 0.f;
        for (int j = sb; j < eb; j += 4) {
          uint4 v[4][NV];
          float sc[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const uint32_t row = (j + q < eb) ? row_s[j + q] : 0xffffffffu;
            const char* src = Wl + (size_t)row * row_bytes;
            sc[q] = (Q8 && row != 0xffffffffu) ? __ldg(S + row) : 0.f;
#pragma unroll
            for (int k = 0; k < NV; ++k)
              v[q][k] = (row != 0xffffffffu && lane + k * G < p.V) ? ld_nc_v4(src + (size_t)k * G * 16)
                                                                 : make_uint4(0, 0, 0, 0);
          }
          mid();
#pragma unroll
          for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int k = 0; k < NV; ++k) accum_row<WT>(acc[k], v[q][k], sc[q]);
        }
        OT* dst = outl + (int64_t)bag * p.out_stride_bag;
        if (p.accumulate) {
          if (eb > sb) {
#pragma unroll
            for (int k = 0; k < NV; ++k)
              if (lane + k * G < p.V) red_vec<OT, VE>(dst + k * G * VE, acc[k]);
          }
          continue;
        }
#pragma unroll
        for (int k = 0; k < NV; ++k)
          if (lane + k * G < p.V) store_vec<OT, VE, true>(dst + k * G * VE, acc[k]);
      }
    } else {
      // index window does not fit the staging buffer (very long bags): indices straight from global memory
      const int64_t* o64 = s_off64 + buf * (FWD_MAX_TILE + 2);
      for (int bag = gid; bag < nb; bag += NG) {
        const int64_t sb = o64[bag], eb = o64[bag + 1];
        float acc[NV][VE];
#pragma unroll
        for (int k = 0; k < NV; ++k)
#pragma unroll
          for (int x = 0; x < VE; ++x) acc[k][x] = 0.f;
        for (int64_t j = sb; j < eb; j += 4) {
          uint4 v[4][NV];
          float sc[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            bool ok = false;
            int64_t idx = 0;
            if (j + q < eb && j + q >= 0 && j + q < p.nnz) {
              idx = __ldg(p.indices + j + q);
              ok = (uint64_t)idx < (uint64_t)nrows;
              if (!ok && p.error) *p.error = 1;
            }
            sc[q] = (Q8 && ok) ? __ldg(S + idx) : 0.f;
            const char* src = Wl + (ok ? (size_t)idx : 0) * row_bytes;
#pragma unroll
            for (int k = 0; k < NV; ++k)
              v[q][k] = (ok && lane + k * G < p.V) ? ld_nc_v4(src + (size_t)k * G * 16) : make_uint4(0, 0, 0, 0);
          }
          mid();
#pragma unroll
          for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int k = 0; k < NV; ++k) accum_row<WT>(acc[k], v[q][k], sc[q]);
        }
        OT* dst = outl + (int64_t)bag * p.out_stride_bag;
        if (p.accumulate) {
          if (eb > sb) {
#pragma unroll
            for (int k = 0; k < NV; ++k)
              if (lane + k * G < p.V) red_vec<OT, VE>(dst + k * G * VE, acc[k]);
          }
          continue;
        }
#pragma unroll
        for (int k = 0; k < NV; ++k)
          if (lane + k * G < p.V) store_vec<OT, VE, true>(dst + k * G * VE, acc[k]);
      }
    }
    mid();  // groups without a bag in this tile (and empty bags) still owe their share of the compaction
    __syncthreads();
  }
}

// Fallback for rows that are not a whole number of 16-byte vectors (tiny D in tests).
template <typename WT, typename OT>
__global__ void emb_fwd_scalar_kernel(const EmbFwdParams p, int64_t T) {
  const int64_t n = T * p.B * p.D;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int d = (int)(i % p.D);
    const int64_t g = i / p.D;
    const int64_t t = g / p.B, b = g % p.B;
    const WT* W = static_cast<const WT*>(p.weights[t]);
    const int64_t nrows = p.rows[t];
    float acc = 0.f;
    for (int64_t j = p.offsets[g]; j < p.offsets[g + 1]; ++j) {
      const int64_t idx = p.indices[j];
      if ((uint64_t)idx < (uint64_t)nrows) {
        if constexpr (sizeof(WT) == 1) {
          acc = fmaf(p.scales[t][idx], to_f32<WT>(W[idx * p.D + d]), acc);
        } else {
          acc += to_f32<WT>(W[idx * p.D + d]);
        }
      } else if (p.error) {
        *p.error = 1;
      }
    }
    static_cast<OT*>(p.out)[t * p.out_stride_table + b * p.out_stride_bag + d] = from_f32<OT>(acc);
  }
}

static int fwd_sm_count() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

template <typename WT, typename OT, int G, int NV>
static int launch_fwd_persist(const EmbFwdParams& p, int64_t T, int cap, int ctas_per_sm, cudaStream_t st) {
  const int64_t tiles = ceil_div(p.B, p.tile);
  const int64_t n_items = tiles * T;
  const size_t smem = (size_t)FWD_OSLOTS * FWD_OFF_PAD * 8 + 2 * (size_t)(cap + 2) * 8 + 2 * (FWD_MAX_TILE + 2) * 8 +
                      2 * (size_t)(cap + 2) * 4 + 2 * (FWD_MAX_TILE + 4) * 4;
  auto kern = emb_fwd_persist_kernel<WT, OT, G, NV>;
  if (smem > 48 * 1024) {
    const cudaError_t rc = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    B200_REQUIRE(rc == cudaSuccess, B200_ELAUNCH, "emb_fwd: cannot reserve %zu bytes of shared memory", smem);
  }
  const int64_t grid = std::min<int64_t>(n_items, (int64_t)fwd_sm_count() * ctas_per_sm);
  kern<<<(unsigned)grid, FWD_THREADS, smem, st>>>(p, (int)n_items, (int)tiles, cap);
  B200_CHECK_LAUNCH("emb_fwd_persist");
  return B200_OK;
}

template <typename WT, typename OT>
static int launch_fwd(const EmbFwdParams& p, int64_t T, cudaStream_t st) {
  const int V = p.V;
  // Which launch shape?  Measured on B200 (tools/fwd_ab.sh): one-shot CTAs at 6 per SM gather 4.67 TB/s on the
  // config-3 tables and 5.85 TB/s on uniform ones, the persistent pipelined kernel at 4 per SM 4.51 / 5.41 (it
  // hides the staging chain but holds fewer rows in flight and balances its static work split worse across the two
  // dies).  So the local forward stays one-shot; the peer-memory forward, which NVLink bounds at a fraction of
  // that rate, runs persistent with few CTAs per SM and leaves the other slots to the kernels it runs next to.
  // B200_FWD_PERSIST=0 / 1 forces one shape for A/B runs.
  static const int persist = []() {
    const char* e = getenv("B200_FWD_PERSIST");
    return e == nullptr ? -1 : atoi(e);
  }();
  static const int env_ctas = []() {
    const char* e = getenv("B200_FWD_CTAS");
    return e == nullptr ? 0 : atoi(e);
  }();
  const int64_t n_items = ceil_div(p.B, p.tile) * T;
  const bool aligned = (reinterpret_cast<uintptr_t>(p.offsets) & 15) == 0 &&
                       (p.nnz == 0 || (reinterpret_cast<uintptr_t>(p.indices) & 15) == 0);
  // (from 4 ranks on: at 2 ranks half of the rows stay local and the one-shot kernel is faster, 104 vs 124 us)
  const bool want_persist = persist < 0 ? (p.out_ptrs != nullptr && p.lN > 0 && p.B / p.lN >= 4) : persist != 0;
  if (want_persist && aligned && n_items < ((int64_t)1 << 30) && p.nnz > 0) {
    // staging capacity: the average window of a tile with 2x head-room for ragged bags (a longer window falls
    // back to unstaged reads for that tile), at least 256 entries
    const int64_t avg_pool = std::max<int64_t>(1, (p.nnz + T * p.B - 1) / (T * p.B));
    int cap = (int)std::min<int64_t>(FWD_IDX_CAP, std::max<int64_t>(256, 2 * avg_pool * p.tile));
    cap = (cap + 1) & ~1;
    // resident CTAs per SM: enough row loads in flight for HBM when the output is local; the peer-memory variant is
    // bound by NVLink and keeps half of the slots free for the kernels it runs next to
    int ctas = env_ctas > 0 ? env_ctas : (p.out_ptrs != nullptr ? 3 : 4);
#define B200_FWDP_CASE(G, NV) return launch_fwd_persist<WT, OT, G, NV>(p, T, cap, ctas, st)
    if (V <= 1) B200_FWDP_CASE(1, 1);
    else if (V <= 2) B200_FWDP_CASE(2, 1);
    else if (V <= 4) B200_FWDP_CASE(4, 1);
    else if (V <= 8) B200_FWDP_CASE(8, 1);
    else if (V <= 16) B200_FWDP_CASE(16, 1);
    else if (V <= 32) B200_FWDP_CASE(32, 1);
    else if (V <= 64) B200_FWDP_CASE(32, 2);
    else B200_FWDP_CASE(32, 4);
#undef B200_FWDP_CASE
  }
  dim3 grid((unsigned)ceil_div(p.B, p.tile), (unsigned)T);
#define B200_FWD_CASE(G, NV) emb_fwd_kernel<WT, OT, G, NV><<<grid, FWD_THREADS, 0, st>>>(p)
  if (V <= 1) B200_FWD_CASE(1, 1);
  else if (V <= 2) B200_FWD_CASE(2, 1);
  else if (V <= 4) B200_FWD_CASE(4, 1);
  else if (V <= 8) B200_FWD_CASE(8, 1);
  else if (V <= 16) B200_FWD_CASE(16, 1);
  else if (V <= 32) B200_FWD_CASE(32, 1);
  else if (V <= 64) B200_FWD_CASE(32, 2);
  else B200_FWD_CASE(32, 4);
#undef B200_FWD_CASE
  B200_CHECK_LAUNCH("emb_fwd");
  return B200_OK;
}

template <typename WT, typename OT>
static int launch_fwd_scalar(const EmbFwdParams& p, int64_t T, cudaStream_t st) {
  const int64_t n = T * p.B * p.D;
  const int blocks = (int)std::min<int64_t>(ceil_div(n, 256), 148 * 16);
  emb_fwd_scalar_kernel<WT, OT><<<blocks, 256, 0, st>>>(p, T);
  B200_CHECK_LAUNCH("emb_fwd_scalar");
  return B200_OK;
}

}  // namespace b200

using namespace b200;

static int emb_fwd_impl(const void* const* d_weights, const int64_t* d_rows, const int64_t* d_indices,
                        const int64_t* d_offsets, void* d_out, int64_t nnz, int64_t T, int64_t B, int64_t D,
                        int64_t out_stride_table, int64_t out_stride_bag, int w_dtype, int out_dtype,
                        int32_t* d_error, b200_stream_t stream, void* const* d_out_ptrs, int64_t lN, int64_t F,
                        int64_t feat0, const float* const* d_scales = nullptr, int accumulate = 0,
                        int64_t first_dest = 0) {
  B200_REQUIRE(T >= 0 && B >= 0 && D > 0 && nnz >= 0, B200_EINVAL, "emb_fwd: negative size");
  if (T == 0 || B == 0) return B200_OK;
  B200_REQUIRE(d_weights && d_rows && d_offsets && (d_out || d_out_ptrs) && (d_indices || nnz == 0), B200_EINVAL,
               "emb_fwd: null pointer");
  B200_REQUIRE(T <= 65535, B200_EUNSUPPORTED, "emb_fwd: more than 65535 tables");
  B200_REQUIRE((w_dtype == B200_F32 || w_dtype == B200_BF16 || (w_dtype == B200_I8 && d_scales)) &&
                   (out_dtype == B200_F32 || out_dtype == B200_BF16),
               B200_EINVAL, "emb_fwd: bad dtype");
  EmbFwdParams p;
  p.accumulate = accumulate;
  p.scales = d_scales;
  p.weights = d_weights;
  p.rows = d_rows;
  p.indices = d_indices;
  p.offsets = d_offsets;
  p.out = d_out;
  p.nnz = nnz;
  p.B = B;
  p.out_stride_table = out_stride_table;
  p.out_stride_bag = out_stride_bag;
  p.error = d_error;
  p.D = (int)D;
  p.out_ptrs = d_out_ptrs;
  p.lN = lN;
  p.F = F;
  p.feat0 = feat0;
  const int ws = dtype_size(w_dtype), os = dtype_size(out_dtype);
  const int64_t row_bytes = D * ws;
  p.V = (int)(row_bytes / 16);
  // bags per CTA: keep the tile's index window inside the staged capacity on average
  const int64_t avg_pool = (nnz + T * B - 1) / (T * B);
  int tile = FWD_MAX_TILE;
  while (tile > 8 && (int64_t)tile * (avg_pool > 0 ? avg_pool : 1) > FWD_IDX_CAP - 2) tile >>= 1;
  if (d_out_ptrs) {
    B200_REQUIRE(lN > 0 && B % lN == 0 && F > 0, B200_EINVAL, "emb_fwd_sharded: B must be a multiple of lN");
    while (tile > 1 && lN % tile != 0) tile >>= 1;  // a tile never straddles two destination ranks
    p.out_stride_bag = F * D;
  }
  p.tile = tile;
  p.tile_rot = 0;
  if (d_out_ptrs && first_dest > 0) {
    B200_REQUIRE(first_dest < B / lN, B200_EINVAL, "emb_fwd_sharded: first destination rank out of range");
    p.tile_rot = (int)(first_dest * (lN / tile));  // lN % tile == 0: a whole number of tiles per destination
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool vec_ok = row_bytes % 16 == 0 && p.V <= 128 &&
                      (d_out_ptrs ? (D * os) % 16 == 0
                                  : ((reinterpret_cast<uintptr_t>(d_out) & 15) == 0 &&
                                     (out_stride_table * os) % 16 == 0 && (out_stride_bag * os) % 16 == 0));
  B200_REQUIRE(vec_ok || !d_out_ptrs, B200_EUNSUPPORTED, "emb_fwd_sharded: row size must be a multiple of 16 bytes");
  B200_REQUIRE(!accumulate || (vec_ok && ((16 / ws) * os) % 8 == 0), B200_EUNSUPPORTED,
               "emb_fwd (accumulate): row size must be a multiple of 16 bytes");
#define B200_DISPATCH(FN)                                                          \
  (w_dtype == B200_F32                                                             \
       ? (out_dtype == B200_F32 ? FN<float, float>(p, T, st)                       \
                                : FN<float, __nv_bfloat16>(p, T, st))              \
       : (out_dtype == B200_F32 ? FN<__nv_bfloat16, float>(p, T, st)               \
                                : FN<__nv_bfloat16, __nv_bfloat16>(p, T, st)))
  if (w_dtype == B200_I8) {
    if (vec_ok) return out_dtype == B200_F32 ? launch_fwd<int8_t, float>(p, T, st) : launch_fwd<int8_t, __nv_bfloat16>(p, T, st);
    return out_dtype == B200_F32 ? launch_fwd_scalar<int8_t, float>(p, T, st)
                                 : launch_fwd_scalar<int8_t, __nv_bfloat16>(p, T, st);
  }
  if (vec_ok) return B200_DISPATCH(launch_fwd);
  return B200_DISPATCH(launch_fwd_scalar);
#undef B200_DISPATCH
}

extern "C" int b200_emb_fwd(const void* const* d_weights, const int64_t* d_rows,
                            const int64_t* d_indices, const int64_t* d_offsets, void* d_out,
                            int64_t nnz, int64_t T, int64_t B, int64_t D, int64_t out_stride_table,
                            int64_t out_stride_bag, int w_dtype, int out_dtype, int32_t* d_error,
                            b200_stream_t stream) {
  return emb_fwd_impl(d_weights, d_rows, d_indices, d_offsets, d_out, nnz, T, B, D, out_stride_table,
                      out_stride_bag, w_dtype, out_dtype, d_error, stream, nullptr, 0, 0, 0);
}

extern "C" int b200_emb_fwd_q8(const void* const* d_weights, const float* const* d_scales, const int64_t* d_rows,
                               const int64_t* d_indices, const int64_t* d_offsets, void* d_out, int64_t nnz,
                               int64_t T, int64_t B, int64_t D, int64_t out_stride_table, int64_t out_stride_bag,
                               int out_dtype, int32_t* d_error, b200_stream_t stream) {
  B200_REQUIRE(d_scales || T == 0 || B == 0, B200_EINVAL, "emb_fwd_q8: null scale table");
  return emb_fwd_impl(d_weights, d_rows, d_indices, d_offsets, d_out, nnz, T, B, D, out_stride_table,
                      out_stride_bag, B200_I8, out_dtype, d_error, stream, nullptr, 0, 0, 0, d_scales);
}

extern "C" int b200_emb_fwd_sharded_from(const void* const* d_weights, const int64_t* d_rows,
                                         const int64_t* d_indices, const int64_t* d_offsets,
                                         void* const* d_out_ptrs, int64_t nnz, int64_t T, int64_t B, int64_t D,
                                         int64_t lN, int64_t F, int64_t feat0, int w_dtype, int out_dtype,
                                         int accumulate, int64_t first_dest_rank, int32_t* d_error,
                                         b200_stream_t stream) {
  B200_REQUIRE(d_out_ptrs, B200_EINVAL, "emb_fwd_sharded_from: null pointer table");
  return emb_fwd_impl(d_weights, d_rows, d_indices, d_offsets, nullptr, nnz, T, B, D, 0, 0, w_dtype, out_dtype,
                      d_error, stream, d_out_ptrs, lN, F, feat0, nullptr, accumulate ? 1 : 0, first_dest_rank);
}

extern "C" int b200_emb_fwd_sharded_add(const void* const* d_weights, const int64_t* d_rows,
                                        const int64_t* d_indices, const int64_t* d_offsets,
                                        void* const* d_out_ptrs, int64_t nnz, int64_t T, int64_t B, int64_t D,
                                        int64_t lN, int64_t F, int64_t feat0, int w_dtype, int out_dtype,
                                        int32_t* d_error, b200_stream_t stream) {
  B200_REQUIRE(d_out_ptrs, B200_EINVAL, "emb_fwd_sharded_add: null pointer table");
  return emb_fwd_impl(d_weights, d_rows, d_indices, d_offsets, nullptr, nnz, T, B, D, 0, 0, w_dtype, out_dtype,
                      d_error, stream, d_out_ptrs, lN, F, feat0, nullptr, 1);
}

extern "C" int b200_emb_fwd_sharded(const void* const* d_weights, const int64_t* d_rows,
                                    const int64_t* d_indices, const int64_t* d_offsets,
                                    void* const* d_out_ptrs, int64_t nnz, int64_t T, int64_t B, int64_t D,
                                    int64_t lN, int64_t F, int64_t feat0, int w_dtype, int out_dtype,
                                    int32_t* d_error, b200_stream_t stream) {
  B200_REQUIRE(d_out_ptrs, B200_EINVAL, "emb_fwd_sharded: null pointer table");
  return emb_fwd_impl(d_weights, d_rows, d_indices, d_offsets, nullptr, nnz, T, B, D, 0, 0, w_dtype, out_dtype,
                      d_error, stream, d_out_ptrs, lN, F, feat0);
}

// ------------------------------------------------------------------ linearisation of per-table bags
// grid = (chunks, T): table t copies its indices to their place in the merged array (and, shifted by the table's
// row base, into the global-row-id array) and its B start offsets, shifted by the indices before it.
namespace b200 {
__global__ void __launch_bounds__(256)
    linearize_kernel(const int64_t* const* __restrict__ idx_ptrs, const int64_t* const* __restrict__ off_ptrs,
                     const int64_t* __restrict__ nnz_prefix, const int64_t* __restrict__ row_base, int64_t B, int T,
                     int64_t* __restrict__ indices, int64_t* __restrict__ offsets, int64_t* __restrict__ iwro) {
  const int t = blockIdx.y;
  const int64_t p0 = nnz_prefix[t], n = nnz_prefix[t + 1] - p0;
  const int64_t* src = idx_ptrs[t];
  const int64_t base = row_base ? row_base[t] : 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int64_t v = src[i];
    indices[p0 + i] = v;
    if (iwro) iwro[p0 + i] = v + base;
  }
  const int64_t* so = off_ptrs[t];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < B; i += stride) offsets[(int64_t)t * B + i] = so[i] + p0;
  if (t == T - 1 && blockIdx.x == 0 && threadIdx.x == 0) offsets[(int64_t)T * B] = nnz_prefix[T];
}
}  // namespace b200

extern "C" int b200_linearize(const int64_t* const* d_idx_ptrs, const int64_t* const* d_off_ptrs,
                              const int64_t* d_nnz_prefix, const int64_t* d_row_base, int64_t T, int64_t B,
                              int64_t max_nnz, int64_t* d_indices, int64_t* d_offsets, int64_t* d_iwro,
                              b200_stream_t stream) {
  B200_REQUIRE(T >= 0 && B >= 0 && max_nnz >= 0, B200_EINVAL, "linearize: negative size");
  if (T == 0) return B200_OK;
  B200_REQUIRE(T <= 65535, B200_EUNSUPPORTED, "linearize: more than 65535 tables");
  B200_REQUIRE(d_idx_ptrs && d_off_ptrs && d_nnz_prefix && d_indices && d_offsets, B200_EINVAL,
               "linearize: null pointer");
  const int64_t work = std::max<int64_t>(std::max(max_nnz, B), 1);
  const unsigned gx = (unsigned)std::min<int64_t>(ceil_div(work, 256 * 4), 1024);
  b200::linearize_kernel<<<dim3(gx, (unsigned)T), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      d_idx_ptrs, d_off_ptrs, d_nnz_prefix, d_row_base, B, (int)T, d_indices, d_offsets, d_iwro);
  B200_CHECK_LAUNCH("linearize");
  return B200_OK;
}
