// C ABI of librdp_kernels.so: every hand-written sm_100a kernel of raydp_b200.
//
// Conventions: all pointers are device pointers unless a comment says "host"; `stream` is a
// cudaStream_t passed as void*; every function returns 0 on success or a negative code, with a
// message available from rdp_last_error().  The reference (pang-wu/raydp) ships no native code
// at all, so each group cites the *implicit library op* of the reference's training loop that
// the kernel replaces.
#pragma once
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

// ---- common --------------------------------------------------------------------------------
const char* rdp_last_error(void);
int rdp_cuda_init(int device);   // cudaSetDevice + capability check; returns the SM count
int rdp_num_sms(void);

// ---- K6: tcgen05/TMEM GEMMs (cuBLAS behind nn.Linear; reference torch/estimator.py:276-286) --
// C[M,N] = A[M,K] . B[N,K]^T, bf16 operands, fp32 accumulate.  Epilogue: +bias[N], ReLU,
// ReLU-mask by aux[M,N] (aux > 0), bf16 or fp32 store, per-column statistics into stats[2,N]
// (mode 1: sum(v), sum(v^2); mode 2: sum(v), sum(v*aux)); stats must be zeroed by the caller.
int rdp_gemm_tn(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int out_f32,
                int M, int N, int K, const float* bias, int relu, const void* aux, int ld_aux,
                int aux_relu_mask, int stats_mode, float* stats, void* stream);
// dgrad fused with the BatchNorm + ReLU backward of the layer it feeds (Linear -> ReLU -> BN chains):
// out[M,N] (bf16) = (aux > 0) ? alpha[n] * (a . b^T)[m,n] + beta[n] * aux[m,n] + delta[n] : 0, and
// colsum[n] += sum_m out[m,n] (zero it first; = the bias gradient of that layer).  aux is the layer's ReLU
// output, alpha/beta/delta come from rdp_bn_bwd_coef.  N % 64 == 0, N <= 256.
int rdp_gemm_tn_bnbwd(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int M, int N, int K,
                      const void* aux, int ld_aux, const float* alpha, const float* beta, const float* delta,
                      float* colsum, void* stream);
// out[Mw,Nw] (fp32, += with atomics; zero it first) = sum_r g[r,:Mw]^T x[r,:Nw].
int rdp_gemm_wgrad(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows,
                   int Mw, int Nw, void* stream);
// CTA-pair (cta_group::2, 256 x 256 cluster tile) kernels for large problems.  rdp_gemm_tn and
// rdp_gemm_wgrad dispatch to them on their own (plain bf16 output, N % 256 == 0, >= 37 tiles); these
// entry points force the path.  rdp_gemm2_tn: out (bf16) = relu?(a . b^T + bias), N % 256 == 0.
// rdp_gemm_set_dynamic_tiles(1): the pair kernels take their tiles by hardware work stealing (cluster
// launch control) instead of a static persistent loop; use it when GEMMs share the SMs with an
// overlapped all-reduce kernel (process-wide switch, default 0; env RAYDP_B200_DYNAMIC_TILES=1).
int rdp_gemm_set_dynamic_tiles(int on);
// mh: 0 = choose, 1 = 256-row cluster tiles (TMEM double-buffered), 2 = 512-row cluster tiles.
int rdp_gemm2_tn(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int M, int N, int K,
                 const float* bias, int relu, int mh, void* stream);
int rdp_gemm2_wgrad(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows,
                    int Mw, int Nw, int mh, void* stream);
// out[M, N] (bf16) = a[M, K] . b[K, N] with b row-major (MN-major B operand): dgrad dX = dY . W straight
// from the row-major weight, no transposed copy.  N % 256 == 0.
int rdp_gemm2_nn(const void* a, int lda, const void* b, int ldb, void* out, int ldc, int M, int N, int K, int mh,
                 void* stream);

// Allocation-free variant of rdp_gemm_tn's statistics path.  It has no counterpart in the
// reference: it exists for callers that must not allocate device memory (a plugin inside a host
// whose device memory belongs to that host's own pool; a latency-critical loop that cannot take
// cudaMalloc's implicit synchronisation).  out = relu?(a . b^T + bias) as in rdp_gemm_tn, plus the
// per-column batch mean and biased variance of `out`.  The [2, N] accumulator is a __device__
// array inside the library (8 KB per GPU, held until exit), so N <= 1024 or the call is rejected.
// The first call on a device resolves the array's device address: make it before any stream
// capture.  All calls share the array (cleared on entry, on `stream`): calls that overlap on the
// GPU race.  No cudaMalloc / cudaFree / allocating library call is made.
int rdp_gemm_tn_bnstats_noalloc(const void* a, int lda, const void* b, int ldb, void* out, int ldc,
                                int out_f32, int M, int N, int K, const float* bias, int relu,
                                float* mean_out, float* var_out, void* stream);

// ---- K7: BatchNorm1d folded into the neighbouring GEMMs (cuDNN/ATen batch_norm) -------------
int rdp_bn_fold(const float* stats, float count, const float* gamma, const float* beta,
                float* running_mean, float* running_var, float momentum, float eps, int use_running,
                float* a, float* c, float* mean, float* rstd, const float* w, const float* b, int N, int K,
                void* w_folded, int ldw, void* w_folded_t, int ldwt, float* b_folded, void* stream);
int rdp_cast_weight(const float* w, int N, int K, void* out, int ldo, void* out_t, int ldt, void* stream);
// The same for up to 16 layers in one launch; bias[l] (fp32, N[l]) is copied to bias_out[l] (padded operand) too.
// out_t / ldt / bias / bias_out may be null.
int rdp_cast_weight_many(const float* const* w, const int* N, const int* K, void* const* out, const int* ldo,
                         void* const* out_t, const int* ldt, const float* const* bias, float* const* bias_out,
                         int n_layers, void* stream);
int rdp_bn_relu_bwd(const void* dy, const void* z, void* dpre, int B, int N, int ld, const float* stats2,
                    const float* mean, const float* rstd, const float* a, float* dgamma, float* dbeta,
                    float* dbias, int has_bn, void* stream);
int rdp_grad_fixup(float* dw, const float* db, const float* a, const float* c, int N, int K, void* stream);
// BatchNorm backward without a pass over the batch: from the raw weight gradient dw[N,K] (= dpre^T r) and
// bias gradient db[N] of Linear l and its bf16 transposed weight w_t[K,N], derive for BN l-1 (count rows):
// dgamma, dbeta and the per-column constants alpha/beta/delta of rdp_gemm_tn_bnbwd; dw is rewritten in
// place into the gradient of the unfolded weight (a[j] * dw[i,j] + c[j] * db[i], as rdp_grad_fixup does).
int rdp_bn_bwd_coef(float* dw, int ldw, const float* db, const void* w_t, int ldt, int N, int K, const float* a,
                    const float* c, const float* mean, const float* rstd, float count, float* dgamma,
                    float* dbeta, float* alpha, float* beta, float* delta, void* stream);

// ---- K8: last Linear + loss + dloss + metric sums (ATen losses, torchmetrics, loss.item()) --
// loss_kind: 0 MSE, 1 SmoothL1(beta=1), 2 BCEWithLogits.  accum[8] += {loss, |err|, err^2, n,
// correct, 0, 0, 0}.  grad_scale = 1/B trains; 0 evaluates (no backward outputs written).
int rdp_mlp_head(const void* z, int B, int K, int ld, const float* a, const float* c, const float* w,
                 const float* bias, const float* target, int loss_kind, float grad_scale, float* pred,
                 float* accum, void* dy, float* stats2, float* dw, float* db, void* stream);

// ---- fused optimizers over a flat fp32 parameter buffer (torch.optim.Adam / SGD) ------------
// lr_ptr / step_ptr: optional device scalars read by the kernel instead of the by-value lr / step
// (lets a captured CUDA graph follow lr schedules and the Adam bias correction).
int rdp_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                  float eps, float weight_decay, int step, float grad_scale, int adamw,
                  const float* lr_ptr, const float* step_ptr, void* shadow, void* stream);
int rdp_sgd_step(float* p, const float* g, float* buf, int64_t n, float lr, float momentum, float dampening,
                 float weight_decay, int nesterov, int first_step, float grad_scale,
                 const float* lr_ptr, const float* step_ptr, void* shadow, void* stream);
// (`shadow`: optional bf16 copy of the updated parameters, written by the same kernel so the GEMM
//  operands never need a separate fp32 -> bf16 cast pass; NULL to skip.)
// Generic Linear(+ReLU) backward helper: g_out = (act > 0) ? g : 0 (skipped when act == NULL) and
// dbias[N] += column sums of the result, one pass over the [M, N] bf16 matrices.
int rdp_relu_bwd_bias(const void* g, const void* act, void* g_out, float* dbias, int M, int N, int ld, void* stream);
int rdp_add_scalar(float* x, float v, void* stream);
// host_pinned[0..n) = src[0..n) by a device store into page-locked host memory (n <= 32): the per-step
// result read-back without a device->host memcpy in the copy queue (see profiles/ingest_timeline.md).
int rdp_publish_host(const float* src, float* host_pinned, int n, void* stream);
// host (pinned or pageable) -> device async copy on `stream`
int rdp_memcpy_h2d_async(void* dst, const void* src_host, uint64_t bytes, void* stream);
int rdp_memcpy_d2h_async(void* dst_host, const void* src, uint64_t bytes, void* stream);

// ---- K5: ingestion (pandas/numpy `to_torch` on the CPU; reference torch/estimator.py:227-232)
// cols / dtypes / table_rows / cross_* are HOST arrays (copied into the launch parameters).
// dtype codes: 0 f32, 1 f64, 2 i32, 3 i64, 4 i16, 5 i8.
int rdp_ingest_pack(const void* const* cols, const int* dtypes, int n_cols, int64_t rows, int64_t src_offset,
                    void* out, int ld, int out_f32, const float* shift, const float* scale,
                    const void* label, int label_dtype, float* label_out, void* stream);
int rdp_ingest_indices(const void* const* cols, const int* dtypes, int n_cols, int64_t rows, int64_t src_offset,
                       const int* table_rows, int n_cross, const int* cross_a, const int* cross_b,
                       const int* cross_rows, int* out, void* stream);
// v2: `cols[s]` is the value buffer of SOURCE column s holding `widths[s]` consecutive features per row
// (1 = scalar column; w = fixed-width vector column, child values row-major [rows, w]); `valid[s]` is the
// column's Arrow validity bitmap (device copy; NULL = no nulls) whose bit `vbit0[s] + row` belongs to `row`
// of the staged range.  A null feature / label becomes `null_value` (pass NAN for the reference's
// pandas semantics, 0 for fill-with-zero); a null index becomes row 0.  ld >= sum(widths).
int rdp_ingest_pack_v2(const void* const* cols, const void* const* valid, const int* vbit0, const int* widths,
                       const int* dtypes, int n_src, int64_t rows, int64_t src_offset, void* out, int ld,
                       int out_f32, const float* shift, const float* scale, float null_value, const void* label,
                       const void* label_valid, int label_vbit0, int label_dtype, float* label_out, void* stream);
int rdp_ingest_indices_v2(const void* const* cols, const void* const* valid, const int* vbit0, const int* dtypes,
                          int n_cols, int64_t rows, int64_t src_offset, const int* table_rows, int n_cross,
                          const int* cross_a, const int* cross_b, const int* cross_rows, int* out, void* stream);
// All host->device copies of one chunk in one call: dst[i] (device address), src_host[i] (pinned host
// address), bytes[i]; one cudaMemcpyAsync per entry on `stream`.
int rdp_ingest_stage(const uint64_t* dst, const uint64_t* src_host, const uint64_t* bytes, int n, void* stream);

// ---- K1 / K2: peer-memory gradient all-reduce fused with the optimizer ---------------------------
// (NCCL all-reduce issued by DDP + torch.optim step; reference torch/estimator.py:243, :284-286)
// Symmetric region of one rank, in floats: send[0] | send[1] | recv[0] | recv[1] | flags(64 B);
// allocate 4 * slot_floats * 4 + 256 bytes with rdp_comm_alloc, exchange rdp_comm_ipc_handle blobs
// (64 bytes) between the ranks' processes and map them with rdp_comm_ipc_open.
int rdp_comm_alloc(uint64_t bytes, void** out);
int rdp_comm_free(void* ptr);
int rdp_comm_ipc_handle(void* ptr, void* handle64);
int rdp_comm_ipc_open(const void* handle64, void** out);
int rdp_comm_ipc_close(void* ptr);
int rdp_comm_can_access_peer(int dev, int peer);
struct rdp_opt_args {
  int kind;                      // 0: out = scale * sum ; 1: Adam / AdamW ; 2: SGD
  float* p; float* m; float* v;  // parameters and optimizer state (fp32, n floats)
  float lr, beta1, beta2, eps, weight_decay, momentum, dampening;
  int adamw, nesterov;
  const float* lr_ptr; const float* step_ptr;   // optional device scalars (override lr / step)
  void* shadow;                  // optional bf16 copy of the updated parameters (n elements), or NULL
};
// bases[q]: rank q's symmetric region as mapped in this process (host array of `world` pointers).
// n_local_ranks = 1 (normal: ins/outs/epochs/counters/opts have one entry) or = world (single-GPU
// emulation of all ranks in one cooperative launch).  epochs[l] -> one uint32, counters[l] -> four
// uint32, both zero-initialised device memory owned by the caller.  algo: 0 auto, 1 one-shot, 2 two-shot,
// 3 two-shot on send[0]/recv[0] only (registered gradient buffers: `ins[l]` may BE send[0], no copy-in).
int rdp_allreduce(void* const* bases, uint64_t slot_floats, int rank, int world, int n_local_ranks,
                  const float* const* ins, float* const* outs, uint32_t* const* epochs,
                  uint32_t* const* counters, const struct rdp_opt_args* opts, int64_t n, float scale,
                  int algo, void* stream);

// bf16 on the wire, fp32 accumulate: two-shot all-reduce for link-bound gradient buckets.  Same region and
// arguments as rdp_allreduce (the send / recv slots are used as bf16); n % 8 == 0; ins[] may not alias the region.
int rdp_allreduce_bf16(void* const* bases, uint64_t slot_floats, int rank, int world, int n_local_ranks,
                       const float* const* ins, float* const* outs, uint32_t* const* epochs,
                       uint32_t* const* counters, const rdp_opt_args* opts, int64_t n, float scale, void* stream);

// One-round ("LL") all-reduce for latency-bound sizes: every rank pushes {value, epoch} packets into slot `rank`
// of every peer's receive area and then only polls its own memory -- no signal round.  Region per rank:
// 2 (parity) x 8 (source slots) x 2 * slot_floats floats, zero-initialised.  Same argument conventions as
// rdp_allreduce; sums in rank order (bit-identical replicas); optimizer epilogue optional.
int rdp_allreduce_ll(void* const* bases, uint64_t slot_floats, int rank, int world, int n_local_ranks,
                     const float* const* ins, float* const* outs, uint32_t* const* epochs,
                     uint32_t* const* counters, const struct rdp_opt_args* opts, int64_t n, float scale,
                     void* stream);
// Bounded peer waits for every kernel above and below: after `seconds` (0 = wait forever, the default) a spin
// on a peer's flag gives up, writes a non-zero code into *host_status (page-locked host memory; may be NULL)
// and all later waits of the process return at once, so the kernels terminate and the host can raise.
int rdp_comm_set_timeout(double seconds, uint32_t* host_status);

// ---- S2: symmetric memory and multicast objects from the driver's virtual-memory API ---------------------
// (cuMemCreate + POSIX file descriptors + cuMulticast*: no CUDA IPC, no dependency on another library's
// allocator.)  rdp_symm_supported: 0 = unsupported, 1 = symmetric memory, 2 = symmetric memory + multicast.
// Sizes must be multiples of rdp_symm_granularity(device, multicast, n_devices).  The fd of
// rdp_symm_export_fd / rdp_mc_export_fd goes to the peers over a unix socket (SCM_RIGHTS); the receiver
// imports it, which maps the peer's memory for its own device, and closes the fd.  Multicast: rank 0 creates
// the object, every rank imports it, calls rdp_mc_add_device, then (after all have) rdp_mc_bind with its
// own allocation and rdp_mc_map.  rdp_symm_free unmaps and releases without any cooperation of peers.
int rdp_symm_supported(int device);
int rdp_symm_granularity(int device, int multicast, int n_devices, uint64_t* out);
int rdp_symm_alloc(uint64_t bytes, int device, void** handle_out, void** ptr_out);
int rdp_symm_export_fd(void* handle, int* fd_out);
int rdp_symm_import_fd(int fd, uint64_t bytes, int device, void** handle_out, void** ptr_out);
int rdp_symm_free(void* handle);
int rdp_mc_create(uint64_t bytes, int n_devices, void** handle_out);
int rdp_mc_import_fd(int fd, uint64_t bytes, void** handle_out);
int rdp_mc_add_device(void* mc, int device);
int rdp_mc_bind(void* mc, void* symm);
int rdp_mc_map(void* mc, int device, void** ptr_out);
int rdp_close_fd(int fd);

// ---- K3: NVLS all-reduce (multimem.ld_reduce + multimem.st through the NVSwitch) ------------------
// mc_ptr: multicast mapping of all ranks' buffers; local_ptr: this rank's unicast mapping;
// flag_ptrs[q]: 256 zero-initialised bytes inside rank q's buffer (unicast, peer-mapped).
// In place: every rank's buffer ends up holding scale * sum.  opt (optional) applies Adam/SGD from
// the reduced buffer in the same kernel; out (optional) receives a copy of the reduced gradient.
int rdp_allreduce_nvls(void* mc_ptr, void* local_ptr, void* const* flag_ptrs, int rank, int world,
                       uint32_t* epoch, uint32_t* counters, const struct rdp_opt_args* opt, float* out,
                       int64_t n, float scale, int overlap_shape, void* stream);
// bf16 wire through the switch (multimem.ld_reduce ... .acc::f32.v4.bf16x2): the multicast buffer holds n bf16
// values, `in` is the fp32 gradient, the average goes to the optimizer epilogue and / or `out` (fp32).  n % 8 == 0.
int rdp_allreduce_nvls_bf16(void* mc_ptr, void* local_ptr, void* const* flag_ptrs, int rank, int world,
                            uint32_t* epoch, uint32_t* counters, const rdp_opt_args* opt, const float* in,
                            float* out, int64_t n, float scale, int overlap_shape, void* stream);
// overlap_shape != 0: launch 256-thread, <= 40-register blocks (one per SM) that fit beside a resident
// tcgen05 GEMM CTA, for all-reduces meant to run concurrently with backward GEMMs.

// All-to-all over peer memory: chunk q of `in` ([world][floats_per_pair]) lands in slot `rank` of rank q's
// receive buffer (recv_bases[q], [world][floats_per_pair]); flag_ptrs[q] as for rdp_allreduce_nvls.
// parity_stride_floats != 0: every rank has TWO receive buffers that far apart, used alternately (call parity =
// (epoch + 1) & 1), and the call costs one flag round; 0: one buffer and a second ("consumed") round.
// counters: 16 zero-initialised words.
int rdp_alltoall(const void* in, void* const* recv_bases, void* const* flag_ptrs, int rank, int world,
                 int64_t floats_per_pair, int64_t parity_stride_floats, uint32_t* epoch, uint32_t* counters,
                 void* stream);

// One flag round between all ranks on `stream`; the flag area (>= 256 bytes, zero-initialised) sits at
// bases[q] + flag_offset_bytes in every rank's symmetric region; `epoch` is a local device uint32.
int rdp_comm_barrier(void* const* bases, uint64_t flag_offset_bytes, int rank, int world, uint32_t* epoch,
                     void* stream);

// ---- K4: embedding lookup fused with the model-parallel all-to-all (nn.Embedding + [absent] a2a) ---
// tables[t]: local fp32 [rows, D]; slots[t]: feature row (1..S-1) in the slab; cols[t]: column in the
// index matrix.  idx[s] / feats[s]: source rank s' int32 index matrix [batch[s], idx_ld] and bf16 slab
// [batch[s], S, D] (peer mappings for remote ranks).  Host arrays; copied into the launch parameters.
int rdp_embedding_fwd(float* const* tables, const int* slots, const int* cols, int n_tables, const int* const* idx,
                      void* const* feats, const int* batch, int n_src, int idx_ld, int S, int D, void* stream);
// Pull gradient rows from every rank's dfeats slab and apply table[row] += -lr * grad_scale * g.
// lr_ptr (optional device scalar) overrides lr, so a captured CUDA graph follows lr schedules.
int rdp_embedding_bwd_sgd(float* const* tables, const int* slots, const int* cols, int n_tables,
                          const int* const* idx, void* const* dfeats, const int* batch, int n_src, int idx_ld,
                          int S, int D, float lr, float grad_scale, const float* lr_ptr, void* stream);
// K4 v2 (csrc/kernels/embedding_bag.cu): pooled bags, vector index dispatch, uneven batches, deterministic
// duplicate combine and row-wise optimizers.  idx_t[s]: rank s' TRANSPOSED int32 index matrix [n_cols, ldb]
// (ldb % 4 == 0, negative index = padding); lens[t]: bag length of table t (null = 1); batch[s]: capacity of rank
// s' batch, batch_ptr[s] (optional, may be a peer pointer): its actual batch this step; mean: 0 sum, 1 mean.
int rdp_embedding_bag_fwd(float* const* tables, const int* slots, const int* cols, const int* lens, int n_tables,
                          const int* const* idx_t, void* const* feats, const int* batch, const int* const* batch_ptr,
                          int n_src, int ldb, int S, int D, int mean, void* stream);
// Backward + update.  A claim pass gives every touched row of every local table a slot (flag[t]: [rows_t] int32,
// zero between steps; list: [acc_rows] (table, row) int pairs; count: one int); the combine pass adds every
// gradient row into acc ([acc_rows, D] 64-bit fixed point, zero between steps; acc_rows >= indices of a step =
// sum over sources and tables of batch * bag length); the update pass touches every listed row once.
// kind 0: SGD; 1: row-wise Adagrad, state[t] [rows_t]; 2: lazy Adam,
// state[t] = m then v ([2, rows_t, D], state_stride[t] = rows_t * D), bias correction from *step_ptr.
int rdp_embedding_bag_bwd(float* const* tables, const int* slots, const int* cols, const int* lens, int n_tables,
                          const int* const* idx_t, void* const* dfeats, const int* batch, const int* const* batch_ptr,
                          int n_src, int ldb, int S, int D, int mean, float grad_scale,
                          unsigned long long* acc, long long acc_rows, int* const* flag, int* list, int* count,
                          float* const* state, const int64_t* state_stride, int kind, float lr, float eps, float beta1,
                          float beta2, float weight_decay, const float* lr_ptr, const float* step_ptr, void* stream);
// Deterministic reductions (bit-reproducible training).  rdp_gemm_wgrad_det: out[Mw, Nw] = G^T X (out contiguous,
// ldo == Nw; OVERWRITES out): batch splits store into ws ([splits, Mw, Nw] fp32, splits limited by
// ws_floats) and are added in split order.  rdp_colsum_ordered: out[i][0..n[i]) = column sums of the bf16
// matrices src[i] ([rows, ld[i]]) by a fixed summation tree (bias gradients); up to 16 matrices per launch.
int rdp_gemm_wgrad_det(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows, int Mw,
                       int Nw, float* ws, long long ws_floats, void* stream);
// Two-phase form: rdp_gemm_wgrad_partial stores the split tiles of one layer (*splits_out = splits; 1 means the
// tile went straight into out), rdp_wgrad_reduce_many adds the splits of up to 16 layers in one launch.
int rdp_gemm_wgrad_partial(const void* g, int ldg, const void* x, int ldx, float* out, int ldo, int rows, int Mw,
                           int Nw, float* ws, long long ws_floats, int* splits_out, void* stream);
int rdp_wgrad_reduce_many(const float* const* ws, const long long* elems, const int* splits, float* const* outs,
                          int n, void* stream);
// ws / counters: scratch for the two-level tree (rows are cut into up to 32 chunks; ws holds chunks x groups x 128
// floats with groups = sum ceil(n[i] / 128); the first `groups` words of counters are zero on entry and on exit);
// without scratch one CTA per 64 columns walks all rows.
int rdp_colsum_ordered(const void* const* src, const int* ld, const int* n, float* const* out, int n_seg, int rows,
                       float* ws, long long ws_floats, unsigned int* counters, int n_counters, void* stream);
// ---- histogram gradient boosting (XGBoostEstimator, tree_method "gpu_hist") -------------------------------------
// binned_t: uint16 bin ids, FEATURE-major [F, n]; slot[r]: histogram slot of row r's node in this level (-1: skip);
// hist: fp64 [n_slots, 2 (g, h), F, n_bins], zeroed by the caller.
int rdp_gbt_hist(const void* binned_t, const float* g, const float* h, const int* slot, double* hist, long long n,
                 int F, int n_bins, int n_slots, void* stream);
// After the splits of a level: node_of[r] moves to the child (rows whose bin == miss_bin follow default_left), and
// slot[r] = in_sample[r] ? next_slot[node] : -1.  feat[node] < 0 means the node was not split.
int rdp_gbt_partition(const void* binned_t, int* node_of, int* slot, const void* in_sample, const int* feat,
                      const int* thr_bin, const int* left, const int* right, const void* default_left,
                      const int* next_slot, long long n, int miss_bin, void* stream);
// ---- K9: DLRM dot interaction (torch.bmm + tril gather + cat + pad) ---------------------------------
// feats [B, S, D] bf16 -> out [B, ld_out] bf16 = [row 0 (D) | S(S-1)/2 pair dots, i > j row-major | 0 pad].
int rdp_dot_interaction_fwd(const void* feats, void* out, int B, int S, int D, int ld_out, void* stream);
int rdp_dot_interaction_bwd(const void* feats, const void* dout, void* dfeats, int B, int S, int D, int ld_out,
                            void* stream);
// BCEWithLogits(mean) on column 0 of a bf16 logit matrix; writes dlogits[:, 0] (other width_d-1 columns 0);
// accum[0] += loss sum, [3] += count, [4] += correct, [5] += batch-mean loss.  dlogits may be null (eval);
// dbias (optional, +=) receives sum(dlogits[:, 0]), the bias gradient of the last Linear.
int rdp_bce_logits(const void* logits, int ld, const float* target, void* dlogits, int ld_d, int width_d,
                   float* accum, int B, float grad_scale, float* dbias, void* stream);

#ifdef __cplusplus
}
#endif
