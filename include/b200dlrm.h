/*
 * b200dlrm.h -- C ABI of the B200-native DLRM training hot path.
 *
 * Drop-in boundary: these entry points are what a binding for the
 * reference's DLRM step would call in place of the ATen / IPEX operators it uses today.
 * Reference paths: P/ = applications/ai/quickstart/models/recommendation/pytorch/dlrm/product,
 * T/ = .../dlrm/training/bfloat16.
 *
 * Conventions
 *   - plain C: raw DEVICE pointers, int64_t sizes, dtype enum, a CUDA stream handle.
 *   - every function returns an int status (0 = ok); no exception crosses the ABI.
 *     b200_last_error() returns a thread-local message for the last non-zero status.
 *   - the caller owns every buffer, including workspace (sizes from *_workspace_bytes).
 *   - calls are stream-ordered, never synchronise, never allocate: they may be captured
 *     in a CUDA graph.  The library keeps no global device state (b200_sort_dedup_noalloc's
 *     built-in scratch excepted).
 *   - element strides are in ELEMENTS of the buffer's dtype.
 */
#ifndef B200DLRM_H_
#define B200DLRM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* b200_stream_t; /* == cudaStream_t */

enum { B200_F32 = 0, B200_BF16 = 1, B200_I8 = 2 /* tables of b200_emb_fwd_q8 only */ };

enum {
  B200_OK = 0,
  B200_EINVAL = 1,       /* bad argument (null pointer, negative size, misaligned buffer) */
  B200_EUNSUPPORTED = 2, /* valid but not implemented shape / dtype */
  B200_EWORKSPACE = 3,   /* workspace too small */
  B200_ELAUNCH = 4       /* CUDA launch error */
};

/* optimiser selector for b200_emb_bwd_update */
enum { B200_OPT_SGD = 0, B200_OPT_ADAGRAD = 1 };

const char* b200_last_error(void);
int b200_version(void);

/*
 * Sum-pooled EmbeddingBag forward over ALL tables in one launch.
 * Replaces the per-table loop `V = E(sparse_index_group_batch, sparse_offset_group_batch)` of
 * DLRM_Net.apply_emb (P/dlrm_s_pytorch.py:290-305, T/dlrm_s_pytorch.py:383-405) and the forward
 * half of ipex MergedEmbeddingBagWithSGD (call site P/dlrm_s_pytorch.py:288-289).
 *
 *   d_weights   device array [T] of table base pointers, table t is [rows_t, D] row-major
 *   d_rows      device array [T] of row counts (for the optional bounds check)
 *   d_indices   int64 [nnz]  table-local row ids, tables concatenated in order
 *   d_offsets   int64 [T*B+1] CSR row pointers over all T*B bags (bag g = t*B + b)
 *   d_out       out[t*out_stride_table + b*out_stride_bag + d]; an empty bag gives zeros
 *   d_error     optional int32[1]; set to 1 if an index is out of range (the row is then
 *               skipped). May be NULL.
 */
int b200_emb_fwd(const void* const* d_weights, const int64_t* d_rows, const int64_t* d_indices,
                 const int64_t* d_offsets, void* d_out, int64_t nnz, int64_t T, int64_t B, int64_t D,
                 int64_t out_stride_table, int64_t out_stride_bag, int w_dtype, int out_dtype,
                 int32_t* d_error, b200_stream_t stream);

/*
 * Stable sort + run-length de-duplication of global row ids: the index stage of the sparse
 * backward.  Equals torch.sort(stable=True) / torch.unique(return_counts=True) /
 * sparse.coalesce().indices() which the reference reaches through
 * `optimizer.step()` -> `param.add_(sparse_grad)` (P/dlrm_s_pytorch.py:1059).
 *
 *   d_keys        uint32 [n] keys (global row ids < key_bound)
 *   d_vals        uint32 [n] payload carried through the sort (position or bag id)
 *   outputs       d_sorted_keys[n], d_sorted_vals[n], d_unique[n], d_seg_offsets[n+1]
 *                 (first *d_num_unique(+1) entries valid), d_num_unique int32[1]
 *   ws            workspace of b200_sort_workspace_bytes(n) bytes
 */
size_t b200_sort_workspace_bytes(int64_t n);
int b200_sort_dedup(const uint32_t* d_keys, const uint32_t* d_vals, int64_t n, uint64_t key_bound,
                    uint32_t* d_sorted_keys, uint32_t* d_sorted_vals, uint32_t* d_unique,
                    uint32_t* d_seg_offsets, int32_t* d_num_unique, void* ws, size_t ws_bytes,
                    b200_stream_t stream);

/*
 * Same result as b200_sort_dedup, for callers that cannot provide a workspace because they cannot
 * allocate device memory at all (all of HBM already belongs to their framework's allocator, an
 * allocation would stall a latency-critical loop, or the caller is a thin language binding that
 * cannot own a device buffer).  No counterpart in the reference: this entry point exists only for
 * such callers.
 *
 * The scratch is a fixed array inside the library image (about 4.5 MB on every GPU that loads the
 * library, kept until the process exits); the call allocates nothing, neither on the device nor
 * through any library.  Limits and rules:
 *   - n <= B200_SORT_NOALLOC_MAX_KEYS, otherwise B200_EUNSUPPORTED and nothing is launched;
 *   - the first call on each device looks up the array's address on that device (the device the
 *     stream belongs to must be current); make that first call before any stream capture begins;
 *   - every call on a device uses the same array: calls that can overlap on the GPU race, so
 *     issue them on one stream or order them yourself.  The array keeps the previous call's
 *     contents; nothing in it needs to be zero on entry.
 */
#define B200_SORT_NOALLOC_MAX_KEYS 262144
int b200_sort_dedup_noalloc(const uint32_t* d_keys, const uint32_t* d_vals, int64_t n,
                            uint64_t key_bound, uint32_t* d_sorted_keys, uint32_t* d_sorted_vals,
                            uint32_t* d_unique, uint32_t* d_seg_offsets, int32_t* d_num_unique,
                            b200_stream_t stream);

/*
 * EmbeddingBag backward fused with the optimiser row update; never materialises the
 * [nnz, D] COO gradient.  Replaces autograd's _embedding_bag_sparse_backward (triggered by
 * E.backward(), P/dlrm_s_pytorch.py:1055) + torch.optim.SGD.step on the sparse grads
 * (P/dlrm_s_pytorch.py:805,1059; lr/world for sharded tables T/dlrm_s_pytorch.py:1020-1031)
 * and the backward half of ipex MergedEmbeddingBagWithSGD (P/dlrm_s_pytorch.py:801,1027).
 * Adagrad follows torch.optim.Adagrad's sparse branch (no call site in the reference).
 *
 *   d_row_base  int64 [T+1] exclusive prefix sum of table row counts
 *   d_grad      grad of the pooled output, addressed like d_out of b200_emb_fwd
 *   d_state     Adagrad: device array [T] of fp32 state_sum tables; SGD: NULL
 *   d_trail     bf16 tables only, optional: device array [T] of uint16 [rows_t, D] tables holding the
 *               LOWER 16 bits of the fp32 master weights (the bf16 table holds the upper 16): the
 *               update is then exact in fp32 while the forward gathers half the bytes -- the
 *               split-SGD storage of the reference's to_bfloat16_train() (P/dlrm_s_pytorch.py:920-933).
 *               NULL: plain bf16 rows, rounded to nearest after the update.
 *   lr / d_lr   learning rate by value, or (if d_lr != NULL) read from device memory at
 *               kernel time so a captured graph follows the LR schedule
 *   ws          b200_emb_bwd_workspace_bytes(nnz, D) bytes
 */
size_t b200_emb_bwd_workspace_bytes(int64_t nnz, int64_t D);
int b200_emb_bwd_update(void* const* d_weights, float* const* d_state, uint16_t* const* d_trail,
                        const int64_t* d_rows,
                        const int64_t* d_row_base, const int64_t* d_indices,
                        const int64_t* d_offsets, int64_t nnz, const void* d_grad, int64_t T,
                        int64_t B, int64_t D, int64_t grad_stride_table, int64_t grad_stride_bag,
                        int w_dtype, int grad_dtype, int optimizer, float lr, const float* d_lr,
                        float eps, int64_t total_rows, void* ws, size_t ws_bytes, int32_t* d_error,
                        b200_stream_t stream);

/*
 * The two halves of b200_emb_bwd_update, callable separately (same workspace in between):
 *   prepare: keys + stable sort + dedup.  Depends only on the step's indices, so a caller may
 *            issue it on a side stream while the forward / dense backward runs.
 *   apply:   segment-reduce the gradient rows of each unique row and update it.
 */
int b200_emb_bwd_prepare(const int64_t* d_rows, const int64_t* d_row_base, const int64_t* d_indices,
                         const int64_t* d_offsets, int64_t nnz, int64_t T, int64_t B, int64_t D,
                         int64_t total_rows, void* ws, size_t ws_bytes, int32_t* d_error,
                         b200_stream_t stream);
int b200_emb_bwd_apply(void* const* d_weights, float* const* d_state, uint16_t* const* d_trail,
                       const int64_t* d_row_base,
                       int64_t nnz, const void* d_grad, int64_t T, int64_t B, int64_t D,
                       int64_t grad_stride_table, int64_t grad_stride_bag, int w_dtype, int grad_dtype,
                       int optimizer, float lr, const float* d_lr, float eps, void* ws, size_t ws_bytes,
                       b200_stream_t stream);

/*
 * Table-wise sharded variants: the pooled-embedding all-to-all of the reference's distributed_forward
 * (ext_dist.alltoall, P/dlrm_s_pytorch.py:351,354 -> P/extend_distributed.py:271-333) fused into the
 * two embedding kernels over NVLink peer memory (one process per GPU, every rank maps every rank's
 * buffers, e.g. torch symmetric memory / CUDA IPC):
 *   fwd:   bag b (index in the GLOBAL batch) of local table t is written straight to its destination
 *          rank r = b / lN:  d_out_ptrs[r] + ((b - r*lN) * F + feat0 + t) * D, i.e. into that rank's
 *          interaction input T[lN, F, D] (feat0 = 1 + global id of this rank's first table);
 *   apply: the gradient row of that bag is read from d_grad_ptrs[r] at the same offset of rank r's
 *          dT[lN, F, D] (the mirror all-to-all of P/extend_distributed.py:323-333).
 * The caller provides the cross-rank ordering: a barrier after every rank's fwd kernel before T is
 * read, and one after every rank's interaction backward before apply reads the gradients.
 * B (global batch) must be a multiple of lN; rows must be a multiple of 16 bytes.
 */
int b200_emb_fwd_sharded(const void* const* d_weights, const int64_t* d_rows, const int64_t* d_indices,
                         const int64_t* d_offsets, void* const* d_out_ptrs, int64_t nnz, int64_t T,
                         int64_t B, int64_t D, int64_t lN, int64_t F, int64_t feat0, int w_dtype,
                         int out_dtype, int32_t* d_error, b200_stream_t stream);
/* b200_emb_fwd_sharded (accumulate = 0) / b200_emb_fwd_sharded_add (accumulate = 1) with the order of the bag tiles
 * rotated so that the kernel starts with the bags of rank `first_dest_rank` and goes round from there.  Every rank
 * of a job passes (its own rank + 1) mod world: at any moment the ranks then write to DIFFERENT destinations, the
 * schedule of an all-to-all, instead of all converging on rank 0's NVLink ingress first, then on rank 1's, ... */
int b200_emb_fwd_sharded_from(const void* const* d_weights, const int64_t* d_rows, const int64_t* d_indices,
                              const int64_t* d_offsets, void* const* d_out_ptrs, int64_t nnz, int64_t T, int64_t B,
                              int64_t D, int64_t lN, int64_t F, int64_t feat0, int w_dtype, int out_dtype,
                              int accumulate, int64_t first_dest_rank, int32_t* d_error, b200_stream_t stream);
int b200_emb_bwd_apply_sharded(void* const* d_weights, float* const* d_state, uint16_t* const* d_trail,
                               const int64_t* d_row_base,
                               int64_t nnz, const void* const* d_grad_ptrs, int64_t T, int64_t B,
                               int64_t D, int64_t lN, int64_t F, int64_t feat0, int w_dtype,
                               int grad_dtype, int optimizer, float lr, const float* d_lr, float eps,
                               void* ws, size_t ws_bytes, b200_stream_t stream);

/*
 * Row-wise sharded hot tables (BASELINE configs[4]; not in the reference, whose distributed_forward shards table-wise
 * only, T/dlrm_s_pytorch.py:483-522).  Rank r of W owns the rows {i : i mod W == r} of a table, stored densely as
 * local row i / W.
 *   b200_rowwise_filter          bags over global row ids (CSR d_offsets[n_bags+1]) -> bags over the rows this rank
 *                                owns: d_out_indices (local ids, FIXED capacity) and d_out_offsets[n_bags+1], whose
 *                                last entry is the number of indices kept.  The fill level never reaches the host,
 *                                so the step has no data-dependent shape (graph-capturable).  If more indices are
 *                                owned than out_capacity, the overflow is dropped and *d_error is set to 2.
 *   b200_emb_fwd_sharded_add     b200_emb_fwd_sharded, but the pooled partial sum is ADDED (red.global.add over
 *                                peer memory) to the destination slot, which the destination rank zeroed; bags
 *                                without an owned index send nothing.  W ranks together build the pooled row: the
 *                                reduce-scatter of the row-wise scheme fused into the lookup.
 *   b200_emb_bwd_prepare_partial b200_emb_bwd_prepare for an index array filled up to *d_nnz_valid (device-side,
 *                                = d_out_offsets[n_bags] of the filter): the unused tail sorts behind the real keys
 *                                and the following b200_emb_bwd_apply[_sharded] stops at the fill level.
 * The backward needs no extra entry point: b200_emb_bwd_apply_sharded already gathers the gradient rows of every
 * rank's samples (the all-gather of the row-wise scheme).
 */
size_t b200_rowwise_filter_workspace_bytes(int64_t n_bags);
int b200_rowwise_filter(const int64_t* d_indices, const int64_t* d_offsets, int64_t n_bags, int64_t world,
                        int64_t rank, int64_t* d_out_indices, int64_t out_capacity, int64_t* d_out_offsets,
                        int32_t* d_error, void* ws, size_t ws_bytes, b200_stream_t stream);
int b200_emb_fwd_sharded_add(const void* const* d_weights, const int64_t* d_rows, const int64_t* d_indices,
                             const int64_t* d_offsets, void* const* d_out_ptrs, int64_t nnz, int64_t T, int64_t B,
                             int64_t D, int64_t lN, int64_t F, int64_t feat0, int w_dtype, int out_dtype,
                             int32_t* d_error, b200_stream_t stream);
int b200_emb_bwd_prepare_partial(const int64_t* d_rows, const int64_t* d_row_base, const int64_t* d_indices,
                                 const int64_t* d_offsets, int64_t nnz_capacity, const int64_t* d_nnz_valid,
                                 int64_t T, int64_t B, int64_t D, int64_t total_rows, void* ws, size_t ws_bytes,
                                 int32_t* d_error, b200_stream_t stream);

/*
 * EmbeddingBag backward as a COALESCED sparse gradient (no optimiser fused): for callers that keep
 * the per-table `nn.EmbeddingBag(sparse=True)` seam of DLRM_Net.apply_emb (P/dlrm_s_pytorch.py:299-305)
 * and let torch.optim.SGD / Adagrad consume the sparse gradient (P/dlrm_s_pytorch.py:805,1059).
 * Equals `weight.grad.coalesce()` of the reference: d_unique_rows[u] = u-th distinct global row id
 * (ascending), d_values[u, :] = sum of the gradient rows of all bags that looked it up (fp32),
 * *d_num_unique = number of distinct rows.  Outputs must hold nnz rows.  Same workspace as
 * b200_emb_bwd_update.
 */
int b200_emb_bwd_coalesce(const int64_t* d_rows, const int64_t* d_row_base, const int64_t* d_indices,
                          const int64_t* d_offsets, int64_t nnz, const void* d_grad, int64_t T, int64_t B,
                          int64_t D, int64_t grad_stride_table, int64_t grad_stride_bag, int grad_dtype,
                          int64_t total_rows, int64_t* d_unique_rows, float* d_values,
                          int32_t* d_num_unique, void* ws, size_t ws_bytes, int32_t* d_error,
                          b200_stream_t stream);

/*
 * EmbeddingBag backward as a DENSE gradient, for the small tables the reference replicates on every rank instead
 * of sharding them: `nn.EmbeddingBag(n, m, mode="sum", sparse=False)` when n < --sparse-dense-boundary
 * (T/dlrm_s_pytorch.py:266-269), wrapped in DDP so that their dense gradients are all-reduced with the MLPs'
 * (T/dlrm_s_pytorch.py:988-997).  Second half of a two-call sequence: after b200_emb_bwd_prepare on the same
 * workspace, every distinct row gets the sum of the gradient rows of the bags that looked it up:
 *   d_dense_grad[row_base[t] + r, :] = sum_{bags b of table t containing r (with multiplicity)} grad[t, b, :]
 * fp32, rows no bag touched are NOT written (the caller zeroes the buffer).  Hot rows (a 3-row table receives
 * B / 3 gradients per row) go through the same windowed reduction as the fused update: deterministic, no atomics.
 */
int b200_emb_bwd_dense_grad(const int64_t* d_row_base, int64_t nnz, const void* d_grad, int64_t T, int64_t B,
                            int64_t D, int64_t grad_stride_table, int64_t grad_stride_bag, int grad_dtype,
                            float* d_dense_grad, void* ws, size_t ws_bytes, b200_stream_t stream);

/*
 * Pairwise dot-product feature interaction, forward and backward.
 * Replaces DLRM_Net.interact_features (P/dlrm_s_pytorch.py:309-334, T/dlrm_s_pytorch.py:407-444;
 * cat + bmm + tril gather + cat) and ipex.nn.functional.interaction(*T) (P/dlrm_s_pytorch.py:312).
 *
 *   d_T   [B, F, D] contiguous: feature 0 is the bottom-MLP output x, 1..F-1 the pooled rows
 *   d_R   [B, ldr], ldr >= W = D + F(F-1)/2 (row stride in elements):
 *         R[b, :W] = [x_b, <T_i,T_j> for i in 1..F-1 for j in 0..i-1]; R[b, W:ldr] = 0 (padding a
 *         caller may ask for so that the next GEMM sees a K that is a multiple of 8)
 *   bwd:  d_dT[B, F, D] = (dZ + dZ^t) T with dZ scattered from d_dR[:, D:W], plus d_dR[:, :D]
 *         added into feature 0.  d_dR has the same row stride ldr.
 * bf16 with F <= 32 and D in {128, 256} runs on tcgen05 tensor cores (TMA-fed, TMEM
 * accumulators); other shapes / fp32 run a CUDA-core kernel.
 */
int b200_interact_fwd(const void* d_T, void* d_R, int64_t B, int64_t F, int64_t D, int64_t ldr,
                      int dtype, b200_stream_t stream);
int b200_interact_bwd(const void* d_T, const void* d_dR, void* d_dT, int64_t B, int64_t F,
                      int64_t D, int64_t ldr, int dtype, b200_stream_t stream);

/*
 * The interaction with options.  flags = 0 is b200_interact_fwd / _bwd.  B200_INTERACT_ITSELF adds the diagonal
 * (a feature's dot product with itself): pairs (i, j <= i) in row-major order, W = D + F(F+1)/2 -- the training
 * variant's --arch-interaction-itself (T/dlrm_s_pytorch.py:428: `offset = 1 if self.arch_interaction_itself`);
 * in backward a diagonal entry contributes 2 * dZ[i,i] * T[i].  This option runs the general CUDA-core kernel.
 * (--arch-interaction-op=cat, T/dlrm_s_pytorch.py:434-436, needs no kernel: it is a view of the pooled buffer.)
 */
enum { B200_INTERACT_ITSELF = 1 };
int b200_interact_fwd_ex(const void* d_T, void* d_R, int64_t B, int64_t F, int64_t D, int64_t ldr, int dtype,
                         int flags, b200_stream_t stream);
int b200_interact_bwd_ex(const void* d_T, const void* d_dR, void* d_dT, int64_t B, int64_t F, int64_t D,
                         int64_t ldr, int dtype, int flags, b200_stream_t stream);

/*
 * fp32 interaction forward on the tensor cores: 3xTF32 (every operand split into two TF32 numbers, three
 * tcgen05.mma kind::tf32 products, fp32 accumulators in TMEM), same result layout as b200_interact_fwd with
 * dtype B200_F32 and the same 1e-5 bound (reference default precision, P/dlrm_s_pytorch.py:315-333).
 * F <= 32, D in {32, 64, 128}.  b200_interact_fwd dispatches fp32 to the CUDA-core kernel by default (faster on
 * B200: this variant is bound by shared-memory operand bandwidth); B200_INTERACT_TF32=1 in the environment makes it
 * take this one.
 */
int b200_interact_fwd_tf32(const void* d_T, void* d_R, int64_t B, int64_t F, int64_t D, int64_t ldr,
                           b200_stream_t stream);

/*
 * The names SURVEY.md section 8b lists for the fused backward, as thin forms of b200_emb_bwd_update with the
 * optimiser fixed (SGD: torch.optim.SGD on the sparse gradients, P/dlrm_s_pytorch.py:805,1059; Adagrad: torch's
 * sparse Adagrad branch, no call site in the reference), and of b200_emb_bwd_workspace_bytes.
 */
size_t b200_workspace_bytes(int64_t nnz, int64_t D);
int b200_emb_bwd_sgd(void* const* d_weights, uint16_t* const* d_trail, const int64_t* d_rows,
                     const int64_t* d_row_base, const int64_t* d_indices, const int64_t* d_offsets, int64_t nnz,
                     const void* d_grad, int64_t T, int64_t B, int64_t D, int64_t grad_stride_table,
                     int64_t grad_stride_bag, int w_dtype, int grad_dtype, float lr, const float* d_lr,
                     int64_t total_rows, void* ws, size_t ws_bytes, int32_t* d_error, b200_stream_t stream);
int b200_emb_bwd_adagrad(void* const* d_weights, float* const* d_state, uint16_t* const* d_trail,
                         const int64_t* d_rows, const int64_t* d_row_base, const int64_t* d_indices,
                         const int64_t* d_offsets, int64_t nnz, const void* d_grad, int64_t T, int64_t B, int64_t D,
                         int64_t grad_stride_table, int64_t grad_stride_bag, int w_dtype, int grad_dtype, float lr,
                         const float* d_lr, float eps, int64_t total_rows, void* ws, size_t ws_bytes,
                         int32_t* d_error, b200_stream_t stream);

/*
 * Per-table bags -> the merged layout of the calls above, in one launch.  Replaces
 * MergedEmbeddingBagWithSGD.linearize_indices_and_offsets (call sites P/dlrm_s_pytorch.py:908-915,1008-1014):
 *   d_idx_ptrs / d_off_ptrs  device arrays [T] of pointers to table t's int64 indices [nnz_t] / start offsets [B]
 *   d_nnz_prefix             int64 [T+1] exclusive prefix sum of nnz_t (host-known: the sizes of the index arrays)
 *   d_row_base               int64 [T+1] or NULL (then d_iwro must be NULL too)
 *   outputs                  d_indices [sum nnz] = concatenation; d_offsets [T*B+1] = start offsets shifted by the
 *                            indices of the tables before, then the total; d_iwro [sum nnz] or NULL = global row ids
 *   max_nnz                  largest nnz_t (sizes the grid)
 */
int b200_linearize(const int64_t* const* d_idx_ptrs, const int64_t* const* d_off_ptrs, const int64_t* d_nnz_prefix,
                   const int64_t* d_row_base, int64_t T, int64_t B, int64_t max_nnz, int64_t* d_indices,
                   int64_t* d_offsets, int64_t* d_iwro, b200_stream_t stream);

/* ---------------------------------------------------------------------------
 * Input side: Criteo binary records -> the tensors the step consumes.
 *
 * Replaces CriteoBinDataset.__getitem__ (T/data_loader_terabyte.py:259-271: np.frombuffer of
 * int32 records, view(-1, tot_fea), column split) and _transform_features (:83-102:
 * ids % max_ind_range, log(float(x) + 1), ids transposed to table-major, offsets = arange per
 * table, label as float column) with one kernel over the RAW records already copied to the device.
 *
 *   d_records  int32 [n_samples, tot_fea]: label, n_dense counts, then the categorical ids
 *   tab_begin / tab_count    categorical features to emit (a rank's table slice), for ALL samples
 *   row_begin / row_count    samples whose dense features and labels to emit (a rank's local batch)
 *   d_dense    fp32 [row_count, dense_stride], columns n_dense..dense_stride-1 are written as 0
 *   d_label    fp32 [row_count]
 *   d_indices  int64 [tab_count, n_samples]  (= lS_i of the slice, or the merged-embedding index array)
 *   d_offsets  int64 [tab_count * n_samples + 1] = 0, 1, 2, ... (one id per bag; lS_o rows are
 *              d_offsets[t * n_samples ...] - t * n_samples, i.e. the same arange the reference builds)
 * max_ind_range == 0 leaves the ids as they are; otherwise floor-mod (never negative), like torch.
 * tot_fea <= 64.
 */
int b200_criteo_transform(const int32_t* d_records, int64_t n_samples, int tot_fea, int n_dense,
                          int64_t max_ind_range, int tab_begin, int tab_count, int64_t row_begin,
                          int64_t row_count, float* d_dense, int64_t dense_stride, float* d_label,
                          int64_t* d_indices, int64_t* d_offsets, b200_stream_t stream);

/* ---------------------------------------------------------------------------
 * Inference with int8 tables.  The reference's --int8 path quantises every
 * EmbeddingBag weight with a symmetric int8 observer (P/dlrm_s_pytorch.py:428-437,868-883: QConfig(weight =
 * PerChannelMinMaxObserver(qint8, per_channel_symmetric)), calibration file P/int8_configure.json); here a
 * table is int8 [rows, D] plus one fp32 scale per ROW (zero point 0), and the pooled forward dequantises on
 * the fly:  out[t, b, :] = sum over the bag of  d_scales[t][r] * float(d_weights[t][r, :]).
 * Same arguments as b200_emb_fwd otherwise (d_scales: T device pointers to fp32 [rows_t]); forward only.
 */
int b200_emb_fwd_q8(const void* const* d_weights, const float* const* d_scales, const int64_t* d_rows,
                    const int64_t* d_indices, const int64_t* d_offsets, void* d_out, int64_t nnz, int64_t T,
                    int64_t B, int64_t D, int64_t out_stride_table, int64_t out_stride_bag, int out_dtype,
                    int32_t* d_error, b200_stream_t stream);

/* ---------------------------------------------------------------------------
 * Dense-MLP backward glue.  The MLP GEMMs stay on torch / cuBLAS (north_star); what autograd launches
 * AROUND them per (Linear, ReLU) pair -- threshold_backward over the [B, N] output gradient plus a column
 * reduction of the same matrix for the bias gradient (reference MLP build P/dlrm_s_pytorch.py:172-211;
 * its IPEX path fuses Linear+ReLU, :926-953) -- is one pass here:
 *   d_dz[b, n] = d_y[b, n] > 0 ? d_dy[b, n] : 0        (bf16, [B, N] contiguous, N % 8 == 0)
 *   d_db[n]    = sum_b d_dz[b, n]                       (fp32, summed in a fixed order: deterministic)
 * ws: b200_relu_bwd_bias_workspace_bytes(B, N) bytes of scratch.
 */
size_t b200_relu_bwd_bias_workspace_bytes(int64_t B, int64_t N);
int b200_relu_bwd_bias(const void* d_dy, const void* d_y, void* d_dz, float* d_db, int64_t B, int64_t N,
                       int dtype, void* ws, size_t ws_bytes, b200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200DLRM_H_ */
