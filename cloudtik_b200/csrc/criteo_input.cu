// Criteo binary-record decode + feature transform on the device (sm_100a).
//
// Replaces, for the training loop's input side, CriteoBinDataset.__getitem__ and _transform_features
// (reference T/data_loader_terabyte.py:259-271 and :83-102): the host hands over the RAW int32
// records of a batch (160 bytes per sample: label, 13 dense counts, 26 categorical ids -- one
// pinned-memory copy, a third of the bytes of the transformed tensors) and ONE kernel writes every
// tensor the step consumes, in the layouts the other kernels read:
//   dense   fp32 [rows, dense_stride]  = log(float(x) + 1), columns >= n_dense zero (GEMM K padding)
//   label   fp32 [rows]
//   indices int64 [n_tab, N]           = id mod max_ind_range (floor-mod), TRANSPOSED (table-major)
//   offsets int64 [n_tab * N + 1]      = arange (one id per bag), the merged-embedding CSR array
// A CTA takes CR_TILE consecutive records = one contiguous 10 KB span: a single 1-D bulk async
// copy (TMA) into shared memory, then coalesced stores (512 B runs per table for the ids).
#include <math.h>

#include "common.cuh"

namespace b200 {

constexpr int CR_TILE = 64;      // records per CTA
constexpr int CR_THREADS = 256;
constexpr int CR_MAX_FEA = 64;   // int32 fields per record (Criteo: 40)

struct CriteoParams {
  const int32_t* rec;
  int64_t N;
  int tot_fea, n_dense;
  int64_t max_ind_range;
  int tab_begin, n_tab;
  int64_t row_begin, n_rows;
  float* dense;
  int64_t dense_stride;
  float* label;
  int64_t* indices;
  int64_t* offsets;
};

__global__ void __launch_bounds__(CR_THREADS) criteo_transform_kernel(const CriteoParams p) {
  __shared__ __align__(128) int32_t s_rec[CR_TILE * CR_MAX_FEA];
  __shared__ __align__(8) uint64_t s_bar;
  const int tid = threadIdx.x;
  const int64_t s0 = (int64_t)blockIdx.x * CR_TILE;
  const int ns = (int)min((int64_t)CR_TILE, p.N - s0);
  const int words = ns * p.tot_fea;
  const int32_t* src = p.rec + s0 * p.tot_fea;
  const uint32_t bytes = (uint32_t)words * 4u;
  const bool bulk = ((reinterpret_cast<uintptr_t>(src) | bytes) & 15u) == 0;
  if (bulk) {
    if (tid == 0) {
      mbar_init(&s_bar, 1);
      fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
      mbar_expect_tx(&s_bar, bytes);
      bulk_g2s(s_rec, src, bytes, &s_bar);
    }
    mbar_wait(&s_bar, 0);
  } else {
    for (int i = tid; i < words; i += CR_THREADS) s_rec[i] = src[i];
    __syncthreads();
  }

  // categorical ids: [n_tab, N] table-major, 64 consecutive int64 per (CTA, table) + the arange offsets
  for (int i = tid; i < p.n_tab * CR_TILE; i += CR_THREADS) {
    const int tl = i / CR_TILE, j = i % CR_TILE;
    if (j < ns) {
      int64_t v = s_rec[j * p.tot_fea + 1 + p.n_dense + p.tab_begin + tl];
      if (p.max_ind_range > 0) {
        v %= p.max_ind_range;
        if (v < 0) v += p.max_ind_range;  // floor-mod: the sign of the divisor, like torch / numpy
      }
      const int64_t o = (int64_t)tl * p.N + s0 + j;
      p.indices[o] = v;
      p.offsets[o] = o;
    }
  }
  if (blockIdx.x == 0 && tid == 0 && p.n_tab > 0) p.offsets[(int64_t)p.n_tab * p.N] = (int64_t)p.n_tab * p.N;

  // dense features and labels of the samples in [row_begin, row_begin + n_rows)
  const int ds = (int)p.dense_stride;
  for (int i = tid; i < ns * ds; i += CR_THREADS) {
    const int j = i / ds, c = i % ds;
    const int64_t r = s0 + j - p.row_begin;
    if (r >= 0 && r < p.n_rows)
      p.dense[r * p.dense_stride + c] = c < p.n_dense ? logf((float)s_rec[j * p.tot_fea + 1 + c] + 1.0f) : 0.f;
  }
  for (int j = tid; j < ns; j += CR_THREADS) {
    const int64_t r = s0 + j - p.row_begin;
    if (r >= 0 && r < p.n_rows) p.label[r] = (float)s_rec[j * p.tot_fea];
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_criteo_transform(const int32_t* d_records, int64_t n_samples, int tot_fea, int n_dense,
                                     int64_t max_ind_range, int tab_begin, int tab_count, int64_t row_begin,
                                     int64_t row_count, float* d_dense, int64_t dense_stride, float* d_label,
                                     int64_t* d_indices, int64_t* d_offsets, b200_stream_t stream) {
  B200_REQUIRE(n_samples >= 0 && tot_fea >= 1 && tot_fea <= CR_MAX_FEA && n_dense >= 0 && n_dense < tot_fea,
               B200_EINVAL, "criteo_transform: bad record shape (at most 64 int32 fields)");
  B200_REQUIRE(tab_begin >= 0 && tab_count >= 0 && 1 + n_dense + tab_begin + tab_count <= tot_fea, B200_EINVAL,
               "criteo_transform: table range outside the record");
  B200_REQUIRE(row_begin >= 0 && row_count >= 0 && row_begin + row_count <= n_samples, B200_EINVAL,
               "criteo_transform: sample range outside the batch");
  B200_REQUIRE(dense_stride >= n_dense && dense_stride <= 4096, B200_EINVAL, "criteo_transform: bad dense stride");
  B200_REQUIRE(max_ind_range >= 0, B200_EINVAL, "criteo_transform: negative max_ind_range");
  if (n_samples == 0) return B200_OK;
  B200_REQUIRE(d_records && (row_count == 0 || (d_dense && d_label)) && (tab_count == 0 || (d_indices && d_offsets)),
               B200_EINVAL, "criteo_transform: null pointer");
  CriteoParams p;
  p.rec = d_records;
  p.N = n_samples;
  p.tot_fea = tot_fea;
  p.n_dense = n_dense;
  p.max_ind_range = max_ind_range;
  p.tab_begin = tab_begin;
  p.n_tab = tab_count;
  p.row_begin = row_begin;
  p.n_rows = row_count;
  p.dense = d_dense;
  p.dense_stride = dense_stride;
  p.label = d_label;
  p.indices = d_indices;
  p.offsets = d_offsets;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  criteo_transform_kernel<<<(unsigned)ceil_div(n_samples, CR_TILE), CR_THREADS, 0, st>>>(p);
  B200_CHECK_LAUNCH("criteo_transform");
  return B200_OK;
}
