/*
 * b200_sparse.h — C ABI of the B200-native sparse hot path (libb200sparse.so).
 *
 * Every entry point takes raw DEVICE pointers, sizes and a cudaStream_t (passed as
 * void*), launches on that stream without any device-wide synchronisation, and returns
 * 0 on success or a cudaError_t / negative B200_E* code.  No torch types cross this
 * boundary; the Python host layer (torchrec-oldfork_b200/fbgemm_gpu) binds it with
 * ctypes and registers the torch.ops.fbgemm.* operators on top of it.
 *
 * Each function cites the reference interface it replaces.  Paths are relative to the
 * reference checkout; "fbgemm_gpu/" abbreviates third_party/fbgemm/fbgemm_gpu/.
 */
#ifndef B200_SPARSE_H_
#define B200_SPARSE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* dtype codes shared by all entry points */
enum {
  B200_I32 = 0,
  B200_I64 = 1,
  B200_F32 = 2,
  B200_F16 = 3,
  B200_BF16 = 4,
  B200_U8 = 5
};

/* error codes below zero are ours; positive values are cudaError_t */
enum {
  B200_OK = 0,
  B200_EINVAL = -1,      /* bad argument combination */
  B200_EUNSUPPORTED = -2, /* valid in the reference, not built here */
  B200_ETOOBIG = -3       /* the problem needs more scratch than a fixed-size entry point owns */
};

enum { B200_POOL_SUM = 0, B200_POOL_MEAN = 1 };
enum { B200_OPT_SGD = 0, B200_OPT_ROWWISE_ADAGRAD = 1 };
enum { B200_BOUNDS_FATAL = 0, B200_BOUNDS_WARNING = 1, B200_BOUNDS_IGNORE = 2 };
enum { B200_CUMSUM_COMPLETE = 0, B200_CUMSUM_INCLUSIVE = 1, B200_CUMSUM_EXCLUSIVE = 2 };

/* Human-readable text for a return code of any function below. */
const char* b200_error_string(int code);
/* Library/ABI version, bumped when a signature changes. */
int b200_abi_version(void);
/* Bytes of the small fixed scratch area (`ws`) the scan-based ops need. */
size_t b200_scan_workspace_bytes(void);

/* ---------------------------------------------------------------- jagged integer ops */

/* torch.ops.fbgemm.asynchronous_{complete,inclusive,exclusive}_cumsum
 * (fbgemm_gpu/src/sparse_ops.cu:189,226,263; CPU twin sparse_ops_cpu.cpp:1028-1090).
 * complete: out has n+1 entries, out[0]=0.  dtype is B200_I32 or B200_I64. */
int b200_cumsum(const void* in, void* out, int64_t n, int dtype, int mode, void* ws,
                void* stream);

/* torch.ops.fbgemm.permute_2D_sparse_data (fbgemm_gpu/src/sparse_ops.cu:309-486; CPU
 * sparse_ops_cpu.cpp:512) and, with B == 1, permute_1D_sparse_data (sparse_ops.cu:490-665).
 * lengths is [T, B]; permute is int32[Tp] with entries in [0, T) (duplicates allowed).
 * Step 1 writes out_lengths [Tp, B] and the two exclusive scans (in_offsets [T*B+1],
 * out_offsets [Tp*B+1], both int64 scratch supplied by the caller).  Step 2 copies the
 * value (and optional weight) segments.  out_offsets[Tp*B] is the permuted total.  */
int b200_permute_lengths(const int32_t* permute, const void* lengths, int lengths_dtype,
                         int32_t T, int32_t Tp, int32_t B, void* out_lengths,
                         int64_t* in_offsets, int64_t* out_offsets, void* ws, void* stream);
int b200_permute_data(const int32_t* permute, int32_t Tp, int32_t B,
                      const int64_t* in_offsets, const int64_t* out_offsets,
                      const void* values, void* out_values, int value_bytes,
                      const void* weights, void* out_weights, int weight_bytes,
                      int64_t out_capacity, void* stream);

/* permute_2D_sparse_data in ONE launch for callers that know the permuted total (`out_capacity`;
 * torchrec passes it: jagged_tensor.py:946 `permuted_lengths_sum`, dist_data.py:257).  A feature
 * row's B bags are one contiguous value range, so only T per-feature totals are computed (into
 * `ws`), never per-bag offsets.  T, Tp <= 1024, else B200_EUNSUPPORTED (use the two-step pair). */
int b200_permute_2d_fused(const int32_t* permute, const void* lengths, int lengths_dtype,
                          int32_t T, int32_t Tp, int32_t B, void* out_lengths,
                          const void* values, void* out_values, int value_bytes,
                          const void* weights, void* out_weights, int weight_bytes,
                          int64_t out_capacity, void* ws, void* stream);

/* torch.ops.fbgemm.expand_into_jagged_permute (fbgemm_gpu/src/sparse_ops.cu:671-730). */
int b200_expand_into_jagged_permute(const int32_t* permute, const void* input_offsets,
                                    const void* output_offsets, int dtype, int64_t permute_size,
                                    void* output_permute, int64_t output_size, void* stream);

/* torch.ops.fbgemm.offsets_range (fbgemm_gpu/src/sparse_ops.cu:62-187). */
int b200_offsets_range(const void* offsets, int dtype, int64_t n_offsets, void* out,
                       int64_t range_size, void* stream);

/* torch.ops.fbgemm.block_bucketize_sparse_features (fbgemm_gpu/src/sparse_ops.cu:736-1228;
 * unsigned arithmetic of kernel1/kernel2 :759-761,806-812).  lengths [T*B] (offset dtype),
 * indices [N] (index dtype), block_sizes [T] (index dtype).  new_lengths is [W*T*B].
 * Scratch: offsets [T*B+1] and new_offsets [W*T*B+1], both int64.  Optional outputs may
 * be NULL.  Output order inside a bucket keeps the in-bag order of the input. */
int b200_block_bucketize(const void* lengths, int offset_dtype, const void* indices,
                         int index_dtype, const void* block_sizes, int32_t T, int32_t B,
                         int32_t W, int64_t N, const float* weights, void* new_lengths,
                         void* new_indices, float* new_weights, void* new_pos,
                         void* unbucketize_permute, int64_t* offsets_scratch,
                         int64_t* new_offsets_scratch, void* ws, void* stream);

/* torch.ops.fbgemm.bounds_check_indices (fbgemm_gpu/codegen/embedding_bounds_check.cu:12-110).
 * In-place on indices / offsets.  warning is int64[1] (count of violations). */
int b200_bounds_check(const int64_t* rows_per_table, void* indices, void* offsets, int dtype,
                      int32_t T, int32_t B, int64_t num_indices, int mode, int64_t* warning,
                      void* ws, void* stream);

/* ---------------------------------------------------------------- pooled lookup (TBE) */

/* split_embedding_codegen_forward_{un,}weighted_cuda
 * (fbgemm_gpu/codegen/embedding_forward_split_template.cu:43-474).
 * weights: flat table-concatenated buffer; weights_offsets int64[T] (element offsets);
 * D_offsets int32[T+1]; offsets [T*B+1] table-major; output [B, total_D] row-major.
 * per_sample_weights may be NULL.  All D must be multiples of 4. */
int b200_tbe_forward(const void* weights, int weight_dtype, const int64_t* weights_offsets,
                     const int32_t* D_offsets, int32_t T, int32_t B, int32_t total_D,
                     int32_t max_D, int32_t all_D_mult8, const void* indices,
                     const void* offsets, int index_dtype, const float* per_sample_weights,
                     int pooling_mode, void* output, int output_dtype, void* stream);

/* transpose_embedding_input (fbgemm_gpu/src/split_embeddings_utils.cu:63-306: linearize_index_kernel,
 * radix_sort_pairs :276-306): linearise (table,row) keys and sort them stably - hand-written segmented LSD
 * radix sort (csrc/tbe_sort.cu), one segment per table, 9-bit digits, passes = ceil(log2(rows_t) / 9) per
 * table.  sorted_keys holds uint32 keys when key_bits <= 32, else uint64.  sorted_payload holds
 * b_t = t*B+b, or the original position p when payload_is_position != 0 (weighted lookups); bag_of_pos
 * (int32[N], may be NULL unless payload_is_position) receives b_t per position.  Equal keys keep position
 * order (== a stable argsort, bit for bit).  max_table_bits: ceil(log2(rows)) of the largest table when the
 * caller knows it (saves launching passes that no table needs), 0 when unknown; it must be 0 unless
 * hash_size_cumsum is non-decreasing (features grouped by table).  sort_ws: b200_tbe_sort_workspace_bytes(N, key_bits, T) bytes. */
size_t b200_tbe_sort_workspace_bytes(int64_t N, int key_bits, int32_t T);
int b200_tbe_sort_indices(const void* indices, const void* offsets, int index_dtype,
                          const int64_t* hash_size_cumsum, int32_t T, int32_t B, int64_t N,
                          int key_bits, int max_table_bits, int payload_is_position, void* sorted_keys,
                          int32_t* sorted_payload, int32_t* bag_of_pos, void* sort_ws,
                          size_t sort_ws_bytes, void* stream);

/* split_embedding_backward_codegen_{sgd,rowwise_adagrad}_{un,}weighted_exact_cuda
 * (fbgemm_gpu/codegen/embedding_backward_split_template.cu:81-1138, optimizer bodies
 * embedding_backward_code_generator.py:376-401,629-632).  Consumes the sorted keys,
 * reduces grad_output rows per unique (table,row) in sorted order (deterministic) and
 * applies the optimizer in place.  partial_scratch: b200_tbe_backward_scratch_bytes(). */
size_t b200_tbe_backward_scratch_bytes(int64_t N, int32_t max_D);
int b200_tbe_backward(const float* grad_output, int64_t grad_stride, const void* sorted_keys,
                      const int32_t* sorted_payload, const int32_t* bag_of_pos, int64_t N,
                      int key_bits, void* weights, int weight_dtype,
                      const int64_t* weights_offsets, const int32_t* D_offsets,
                      const int64_t* hash_size_cumsum, int32_t T, int32_t B, int32_t max_D,
                      int32_t all_D_mult8, const void* offsets, int index_dtype,
                      const float* per_sample_weights, int pooling_mode, int optimizer,
                      float* momentum1, const int64_t* momentum1_offsets, float learning_rate,
                      float eps, void* partial_scratch, void* stream);

/* b200_tbe_backward without a scratch argument, for callers that cannot allocate (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The scratch is a file-scope `static __device__` array of
 * b200_tbe_backward_static_max_scratch_bytes() = 16 MiB inside the library (every GPU that loads the
 * library holds it until the process exits); its device address is looked up once per device with
 * cudaGetSymbolAddress, with the CALLER'S CURRENT DEVICE (which must be the device of `stream` and of
 * every pointer).  The call makes no cudaMalloc / cudaFree / allocating library call.  Problems with
 * b200_tbe_backward_scratch_bytes(N, max_D) above the limit return B200_ETOOBIG (cfg2 - 1,703,936 ids,
 * dim 128 - needs 13.8 MiB and fits).  All calls on one GPU share the array: nothing in it has to start
 * at zero (each call overwrites what it reads), but two calls that overlap on the GPU (different
 * streams, or graph branches) race - serialise them or give one of them b200_tbe_backward with its
 * own scratch.  Arguments and results are otherwise those of b200_tbe_backward. */
size_t b200_tbe_backward_static_max_scratch_bytes(void);
int b200_tbe_backward_static(const float* grad_output, int64_t grad_stride, const void* sorted_keys,
                             const int32_t* sorted_payload, const int32_t* bag_of_pos, int64_t N,
                             int key_bits, void* weights, int weight_dtype,
                             const int64_t* weights_offsets, const int32_t* D_offsets,
                             const int64_t* hash_size_cumsum, int32_t T, int32_t B, int32_t max_D,
                             int32_t all_D_mult8, const void* offsets, int index_dtype,
                             const float* per_sample_weights, int pooling_mode, int optimizer,
                             float* momentum1, const int64_t* momentum1_offsets,
                             float learning_rate, float eps, void* stream);

/* b200_tbe_backward with stochastic rounding of the updated fp16 / bf16 rows (the reference's
 * `stochastic_rounding` optimizer argument, default true: embedding_backward_split_template.cu:367-
 * 380, stochastic_rounding_vector in fbgemm_cuda_utils.cuh).  (sr_seed, sr_offset) is the Philox
 * state of the step - the caller advances the offset every call, as at::PhiloxCudaState does;
 * the random stream is keyed by the element's position in `weights`, so a run is reproducible.
 * fp32 tables ignore it. */
int b200_tbe_backward_sr(const float* grad_output, int64_t grad_stride, const void* sorted_keys,
                         const int32_t* sorted_payload, const int32_t* bag_of_pos, int64_t N,
                         int key_bits, void* weights, int weight_dtype,
                         const int64_t* weights_offsets, const int32_t* D_offsets,
                         const int64_t* hash_size_cumsum, int32_t T, int32_t B, int32_t max_D,
                         int32_t all_D_mult8, const void* offsets, int index_dtype,
                         const float* per_sample_weights, int pooling_mode, int optimizer,
                         float* momentum1, const int64_t* momentum1_offsets, float learning_rate,
                         float eps, void* partial_scratch, uint64_t sr_seed, uint64_t sr_offset,
                         void* stream);

/* split_embedding_codegen_grad_indice_weights_cuda
 * (fbgemm_gpu/codegen/embedding_backward_split_indice_weights_template.cu). */
int b200_tbe_grad_indice_weights(const float* grad_output, int64_t grad_stride,
                                 const void* weights, int weight_dtype,
                                 const int64_t* weights_offsets, const int32_t* D_offsets,
                                 int32_t T, int32_t B, const void* indices, const void* offsets,
                                 int index_dtype, float* grad_indice_weights, void* stream);

/* ---------------------------------------------------------------- int8 inference path */

/* torch.ops.fbgemm.FloatToFusedNBitRowwiseQuantizedSBHalf(w, 8)
 * (fbgemm_gpu/src/quantize_ops.cu `_float_to_fusednbitrowwise_cuda_kernel`, CPU
 * quantize_ops_cpu.cpp): rows of ncols uint8 followed by fp16 scale, fp16 bias. */
int b200_quantize_rowwise_8bit_sbhalf(const float* in, int64_t nrows, int32_t ncols,
                                      uint8_t* out, void* stream);

/* int_nbit_split_embedding_codegen_lookup_function, INT8 rows only
 * (fbgemm_gpu/codegen/embedding_forward_quantized_split_template.cu:154-450).
 * Row layout: [fp16 scale, fp16 bias, D x uint8], padded to row_alignment bytes;
 * weights_offsets are BYTE offsets.  idx == -1 contributes zeros.  Output fp16/fp32. */
int b200_tbe_int8_forward(const uint8_t* weights, const int64_t* weights_offsets,
                          const int32_t* D_offsets, int32_t T, int32_t B, int32_t total_D,
                          int32_t max_D, int32_t row_alignment, const void* indices,
                          const void* offsets, int index_dtype,
                          const float* per_sample_weights, int pooling_mode, void* output,
                          int output_dtype, void* stream);

/* ---------------------------------------------------------------- DLRM dot interaction */

/* InteractionArch.forward (torchrec/models/dlrm.py:193-219):
 * dense [B, D], sparse [B, F, D] -> out [B, D + (F+1)F/2] = [dense | triu(X X^T, 1)],
 * X = [dense; sparse]; row-major upper triangle, fp32. */
/* out rows start every out_stride floats (>= the row width); columns past the row width are
 * written as zeros, so with out_stride = round_up(width, 4) the consumer may run an aligned
 * out_stride-wide GEMM over the buffer (the 479-wide row of Criteo DLRM otherwise forces
 * cuBLAS onto its unaligned kernels). */
int b200_interaction_forward(const float* dense, const float* sparse, int32_t B, int32_t F,
                             int32_t D, float* out, int64_t out_stride, void* stream);

/* The same forward with the sample loads done by the tensor memory accelerator: per 32-column stage one lane
 * issues two cp.async.bulk.tensor copies (dense row; the sample's F sparse rows) into a 128-byte-swizzled
 * shared-memory tile, completion on an mbarrier; MMA fragment loads follow the swizzle (bank-conflict free
 * without row padding).  Bit-identical to b200_interaction_forward.  B200_EUNSUPPORTED unless F + 1 <= 32,
 * F >= 1, D % 32 == 0 and 16-byte aligned inputs.  Not the default: 258.6 vs 249.4 us at the Criteo shape
 * (B200_IX_TMA=1 routes b200_interaction_forward here). */
int b200_interaction_forward_tma(const float* dense, const float* sparse, int32_t B, int32_t F, int32_t D,
                                 float* out, int64_t out_stride, void* stream);
/* Backward of the same (autograd of the bmm + triu gather in torchrec/models/dlrm.py:205-219):
 * grad_out [B, D + (F+1)F/2] at a pitch of grad_stride floats
 * -> grad_dense [B, D], grad_sparse [B,F,D]. */
int b200_interaction_backward(const float* dense, const float* sparse, const float* grad_out,
                              int64_t grad_stride, int32_t B, int32_t F, int32_t D,
                              float* grad_dense, float* grad_sparse, void* stream);

/* b200_interaction_backward whose sparse gradient leaves feature by feature: the row of feature f
 * and sample s is stored at (float*)feat_base[f] + s * feat_pitch[f] (device arrays [F]; addresses
 * 16-byte aligned, pitches in elements).  With peer-mapped addresses this is the backward
 * all-to-all of table-wise sharded embeddings (comm_ops.py:585-640: recat copy + all_to_all_single)
 * fused into the kernel that produces the gradient.  B200_EUNSUPPORTED when F + 1 > 32 or D % 32. */
int b200_interaction_backward_scatter(const float* dense, const float* sparse, const float* grad_out,
                                      int64_t grad_stride, int32_t B, int32_t F, int32_t D,
                                      float* grad_dense, const int64_t* feat_base,
                                      const int64_t* feat_pitch, void* stream);

/* ---------------------------------------------------------------- sharded data movement */

/* Receive-side layout change of PooledEmbeddingsAllToAll
 * (torchrec/distributed/comm_ops.py:555-560 split+view+cat, and its backward :418-428):
 * src is W contiguous blocks, block r = [B_local, dims[r]]; dst is [B_local, sum dims]
 * (forward=1) or the inverse with an optional scale (forward=0; the 1/W of :527-528). */
int b200_pooled_a2a_recat(const float* src, float* dst, const int32_t* dim_cumsum /*[W+1]*/,
                          int32_t W, int32_t B_local, int forward, float scale, void* stream);

/* Sequence (non-pooled, PoolingMode.NONE) embeddings - split_embedding_nobag_codegen_forward /
 * backward kernels of the reference (codegen/embedding_forward_split_template.cu `nobag` variants).
 * out[p] = indices[p] + hash_size_cumsum[t(p)], t(p) = the feature whose id range
 * [offsets[t*B], offsets[(t+1)*B]) holds position p.  With all tables of one dim D laid back to
 * back this is the row number in ONE virtual table, and the sequence lookup / fused backward are
 * b200_tbe_forward / b200_tbe_sort_indices / b200_tbe_backward with T = 1, B = N, L = 1. */
int b200_linearize_seq_indices(const void* indices, const void* offsets, int index_dtype,
                               const int64_t* hash_size_cumsum /*[T+1]*/, int32_t T, int32_t B,
                               int64_t N, int64_t* out, void* stream);

/* ---------------------------------------------------------------- dense-layer glue
 *
 * Backward of the ReLU + bias of one Linear+ReLU layer (torchrec/modules/mlp.py Perceptron :19-79;
 * PyTorch: threshold_backward + sum(dim=0)) in ONE pass over the gradient:
 *   grad_pre[b, n] = grad_out[b, n] * (out[b, n] > 0),  grad_bias[n] = sum_b grad_pre[b, n].
 * grad_out / out are [B, N] fp32 with row strides in elements (multiples of 4); grad_pre is dense
 * [B, N]; N % 4 == 0, N <= 1024.  Deterministic (per-CTA partial sums added in a fixed order in
 * `scratch`, b200_relu_bias_grad_scratch_bytes(N) bytes).  The layer's GEMMs stay on cuBLAS. */
size_t b200_relu_bias_grad_scratch_bytes(int32_t N);
int b200_relu_bias_grad(const float* grad_out, int64_t grad_out_stride, const float* out,
                        int64_t out_stride, int64_t B, int32_t N, float* grad_pre,
                        float* grad_bias, float* scratch, void* stream);

/* ---------------------------------------------------------------- fused lookup + exchange (NVLink peer memory)
 *
 * One process per GPU; every rank maps its peers' buffers (CUDA IPC / symmetric memory) and passes
 * device arrays of W base pointers.  These entry points replace lookup -> alltoall_pooled
 * (comm_ops.py:203-560) / reduce_scatter_pooled (:382-890) and their backward collectives with
 * direct peer stores / loads inside the lookup kernels; the caller brackets them with cross-rank
 * barriers (all ranks' stores done before anyone reads; all reads done before buffers are reused). */

/* b200_tbe_forward whose pooled row of global sample b is written to
 * dst_bases[b / B_dst] + (b % B_dst) * dst_pitch + dst_col[t]  (elements of the output dtype).
 * skip_empty != 0: empty bags store nothing (used for the partial sums of row-wise sharding, where a bag is
 * empty on every rank that owns none of its rows).  flag_bases != NULL: every non-empty bag also sets the
 * byte flag_bases[b / B_dst][(b % B_dst) * flag_pitch + flag_col[t]] for b200_peer_stage_reduce_flagged. */
int b200_tbe_forward_peer(const void* weights, int weight_dtype, const int64_t* weights_offsets,
                          const int32_t* D_offsets, int32_t T, int32_t B, int32_t total_D,
                          int32_t max_D, int32_t all_D_mult8, const void* indices,
                          const void* offsets, int index_dtype, const float* per_sample_weights,
                          int pooling_mode, void* const* dst_bases /*device [n_dst]*/,
                          int32_t n_dst, int32_t B_dst, int64_t dst_pitch,
                          const int32_t* dst_col /*device [T]*/, int output_dtype,
                          int32_t skip_empty, uint8_t* const* flag_bases /*device [n_dst], may be NULL*/,
                          int64_t flag_pitch, const int32_t* flag_col /*device [T]*/, void* stream);

/* b200_tbe_backward whose gradient row of global sample b is read from
 * grad_bases[b / B_src] + (b % B_src) * grad_stride + src_col[t], scaled by grad_scale
 * (the 1/W of comm_ops.py:527-528, 883).  per_sample_weights / bag_of_pos as in b200_tbe_backward (weighted
 * features); stochastic_rounding, sr_seed, sr_offset as in b200_tbe_backward_sr (fp16 / bf16 tables). */
int b200_tbe_backward_peer(const float* const* grad_bases /*device [n_src]*/, int32_t n_src,
                           int32_t B_src, int64_t grad_stride, const int32_t* src_col /*device [T]*/,
                           float grad_scale, const void* sorted_keys,
                           const int32_t* sorted_payload, const int32_t* bag_of_pos, int64_t N,
                           int key_bits, void* weights, int weight_dtype,
                           const int64_t* weights_offsets, const int32_t* D_offsets,
                           const int64_t* hash_size_cumsum, int32_t T, int32_t B, int32_t max_D,
                           int32_t all_D_mult8, const void* offsets, int index_dtype,
                           const float* per_sample_weights, int pooling_mode, int optimizer,
                           float* momentum1, const int64_t* momentum1_offsets,
                           float learning_rate, float eps, void* partial_scratch,
                           int32_t stochastic_rounding, uint64_t sr_seed, uint64_t sr_offset, void* stream);

/* Row-wise sharding, the reduction half of reduce_scatter_pooled (torchrec/distributed/comm_ops.py:382,848-890;
 * PooledEmbeddingsReduceScatter, dist_data.py:745): out[b, out_col4[c] : +4] = sum_r stage[r][b][4c : 4c+4] (rank order) for every
 * float4 column c of the stage [W][B_local][Dw] filled by the peers' b200_tbe_forward_peer;
 * out_col4 (device int32 [Dw/4], multiples of 4) places each feature anywhere in the output row.
 * clear_stage != 0: every non-zero slot read is zeroed again, ready for a skip_empty forward. */
int b200_peer_stage_reduce(float* stage, int32_t W, int32_t B_local, int32_t Dw, float* out,
                           int64_t out_pitch, const int32_t* out_col4, int32_t clear_stage,
                           void* stream);

/* Stream-ordered cross-rank barrier over peer memory (what brackets the peer stores / loads above; the
 * reference's collectives synchronise inside NCCL, comm_ops.py:490,592,866,922).  pads: device array of W
 * pointers to the ranks' signal pads (uint32[W] each, zero-initialised symmetric memory); epoch_ctr: a
 * device-local uint32 counter owned by this barrier instance (zero-initialised, bumped by the kernel).
 * Every rank must call it the same number of times.  One 32-thread kernel; graph-capturable. */
int b200_peer_barrier(uint32_t* const* pads, int32_t rank, int32_t W, uint32_t* epoch_ctr, void* stream);

/* The same reduction driven by presence flags: flags is [B_local][F][W] bytes (set by the peers'
 * b200_tbe_forward_peer with flag_pitch = F*W, flag_col[t] = f*W + rank); feature f owns the staged float4
 * columns [feat_col4[f], feat_col4[f+1]).  Only flagged slots are read (rank order, deterministic), consumed
 * flags are cleared, the staging data is left as is.  Replaces the dense pass of
 * reduce_scatter_pooled (comm_ops.py:382) when most (sample, feature, rank) slots are empty.  W <= 32. */
int b200_peer_stage_reduce_flagged(const float* stage, uint8_t* flags, int32_t W, int32_t B_local,
                                   int32_t Dw, int32_t F, const int32_t* feat_col4, float* out,
                                   int64_t out_pitch, const int32_t* out_col4, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_SPARSE_H_ */
