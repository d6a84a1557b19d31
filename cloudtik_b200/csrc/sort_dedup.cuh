// Internal interface of the sort / dedup stage (shared by sort_dedup.cu and emb_bwd.cu).
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int SORT_THREADS = 256;  // must be 256: one thread per 8-bit digit
constexpr int SORT_ITEMS = 16;
constexpr int SORT_CHUNK = SORT_THREADS * SORT_ITEMS;  // keys per block
constexpr int DEDUP_THREADS = 256;
constexpr int DEDUP_ITEMS = 8;
constexpr int DEDUP_CHUNK = DEDUP_THREADS * DEDUP_ITEMS;
// window of sorted positions one group of emb_bwd.cu's update kernel owns (divides DEDUP_THREADS)
constexpr int HOT_CH = 32;

// workspace: [keysA | keysB | valsA | valsB | digit histograms | dedup block counts]
size_t sort_ws_bytes(int64_t n);

// stable LSD radix sort of (key, val) pairs, keys < key_bound <= 2^32; input untouched,
// out_* must not alias the input.  n > 0.
int sort_pairs(const uint32_t* keys, const uint32_t* vals, int64_t n, uint64_t key_bound,
               uint32_t* out_keys, uint32_t* out_vals, void* ws, cudaStream_t st);

// run-length de-duplication of sorted keys: unique[u], seg[u] (start of run u), seg[U] = n,
// *num_unique = U.  n > 0.  Uses the tail of the same workspace.
// win_run (optional, may be null): win_run[w] = index of the run that contains sorted position
// w * HOT_CH, for every window start below n.
// svals / first_val (optional): first_val[u] = svals[seg[u]], the payload of run u's first element.
int dedup_runs(const uint32_t* skeys, int64_t n, uint32_t* unique, uint32_t* seg, int32_t* num_unique,
               void* ws, cudaStream_t st, uint32_t* win_run = nullptr, const uint32_t* svals = nullptr,
               uint32_t* first_val = nullptr);

int empty_result(uint32_t* seg, int32_t* num_unique, cudaStream_t st);

}  // namespace b200
