/* libtfop_b200 — C ABI of the B200 (sm_100a) replica-runtime kernels.
 *
 * Every function returns 0 on success; on failure it returns non-zero and
 * tfop_last_error() gives a thread-local description. Streams are passed as
 * cudaStream_t (a CUstream handle; torch.cuda.current_stream().cuda_stream
 * works). dtype tags: 0 = fp32, 1 = bf16.
 *
 * The reference (ryantd/tf-operator) ships no native code at all (SURVEY.md
 * 2.8c); each entry point below cites the reference call site whose path it
 * replaces.
 */
#ifndef TFOP_B200_H_
#define TFOP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __CUDA_RUNTIME_H__
typedef struct CUstream_st* cudaStream_t;
#endif

typedef struct tfop_comm tfop_comm;

enum { TFOP_DT_F32 = 0, TFOP_DT_BF16 = 1 };

const char* tfop_last_error(void);
int tfop_version(void);

/* ---- symmetric peer memory (csrc/symm.cu) -------------------------------
 * Replaces the DNS-name + port addressing of pkg/controller.v1/pytorch/
 * pytorch.go:13-68 with CUDA IPC handles exchanged at rendezvous. */
int tfop_comm_handle_bytes(void);
int tfop_comm_create(int rank, int world, size_t data_bytes, tfop_comm** out);
int tfop_comm_get_handle(tfop_comm* c, void* out_bytes);
int tfop_comm_open_peers(tfop_comm* c, const void* all_handles);
int tfop_comm_create_emulated(int world, size_t data_bytes, tfop_comm** out);
void* tfop_comm_buffer(tfop_comm* c, int peer);
size_t tfop_comm_buffer_bytes(tfop_comm* c);
int tfop_comm_rank(tfop_comm* c);
int tfop_comm_world(tfop_comm* c);
int tfop_comm_check_error(tfop_comm* c, int* flag);
/* Spin limit of every cross-rank flag barrier, in ms of a 2 GHz SM clock (default 120,000). When
 * it expires the kernel ABANDONS the collective (reduces and stores nothing) and sets the
 * communicator's error word; the replica is expected to exit with a retryable code so the
 * operator re-creates it (restartPolicy ExitCode, pkg/controller.v1/tensorflow/pod.go:140-159). */
int tfop_comm_set_timeout_ms(long long ms);
/* Device address of the error word (uint32, 0 = healthy) for asynchronous polling. */
void* tfop_comm_error_ptr(tfop_comm* c);
/* tfop_comm_destroy refuses (returns an error, frees nothing) while peers may still read or
 * write this rank's buffers: run a cross-rank exit barrier after the last collective, call
 * tfop_comm_quiesce, then destroy. Single-rank and emulated communicators need no quiesce. */
int tfop_comm_quiesce(tfop_comm* c);
int tfop_comm_destroy(tfop_comm* c);

/* ---- K1/K6/K8: fused P2P all-reduce (csrc/allreduce.cu) ------------------
 * Replaces the NCCL all-reduce of examples/v1/distribution_strategy/keras-API/
 * multi_worker_strategy-with-keras.py:76-77 and the rabit histogram
 * all-reduce of examples/xgboost/xgboostjob.yaml:19-25. algo: 0 auto,
 * 1 one-shot, 2 two-shot. */
int tfop_allreduce(tfop_comm* c, const void* in, int in_dtype, void* out, int out_dtype,
                   int wire_dtype, size_t n, size_t sym_off, float scale, int algo,
                   int max_blocks, cudaStream_t stream);
int tfop_allreduce_emulated(tfop_comm** comms, int world, const void* const* ins, int in_dtype,
                            void* const* outs, int out_dtype, int wire_dtype, size_t n,
                            size_t sym_off, float scale, int algo, int max_blocks,
                            cudaStream_t stream);
int tfop_allreduce_set_oneshot_max_bytes(size_t bytes);

/* ---- K2: NVLS (in-switch) all-reduce (csrc/nvls.cu) ------------------------
 * Same reference call site as tfop_allreduce. Set-up: rank 0 tfop_nvls_create -> send the fd
 * to the other ranks (SCM_RIGHTS) -> they tfop_nvls_import; all: tfop_nvls_add_device,
 * barrier, tfop_nvls_bind, barrier. Probe with tfop_nvls_supported first. */
int tfop_nvls_supported(int* out);
int tfop_nvls_round_size(int world, size_t bytes, size_t* out);
int tfop_nvls_create(tfop_comm* c, size_t bytes, int* out_fd);
int tfop_nvls_import(tfop_comm* c, size_t bytes, int fd);
int tfop_nvls_add_device(tfop_comm* c);
int tfop_nvls_bind(tfop_comm* c);
void* tfop_nvls_local_ptr(tfop_comm* c);
size_t tfop_nvls_bytes(tfop_comm* c);
int tfop_allreduce_nvls(tfop_comm* c, const void* in, int in_dtype, void* out, int out_dtype,
                        int wire_dtype, size_t n, size_t sym_off, float scale, int max_blocks,
                        cudaStream_t stream);
int tfop_nvls_destroy(tfop_comm* c);

/* ---- fused flat-buffer optimizers (csrc/optim.cu) ------------------------
 * The optimizer step that follows the gradient exchange in
 * examples/v1/dist-mnist/dist_mnist.py:194-208. hp_dev: device float[8] =
 * {lr, momentum, weight_decay, grad_scale, beta1, beta2, eps, step}. */
int tfop_sgd_momentum(float* master, float* mom, const void* grad, int grad_dtype,
                      void* weights_bf16, size_t n, size_t n_decay, const float* hp_dev,
                      int nesterov, cudaStream_t stream);
int tfop_adamw(float* master, float* m1, float* m2, const void* grad, int grad_dtype,
               void* weights_bf16, size_t n, size_t n_decay, const float* hp_dev,
               cudaStream_t stream);
int tfop_zero(void* p, size_t bytes, cudaStream_t stream);

/* ---- multi-tensor gradient pack (csrc/pack.cu) ---------------------------
 * The "pack gradients into few tensors" step before the collective in
 * keras-API/multi_worker_strategy-with-keras.py:76-77. nseg <= 128. */
typedef struct tfop_pack_seg {
    const void* src;
    unsigned long long dst_off; /* element offset in dst, multiple of 8 */
    unsigned long long n;       /* elements */
} tfop_pack_seg;
int tfop_multi_pack_bf16(const tfop_pack_seg* segs, int nseg, int src_dtype, void* dst,
                         float scale, cudaStream_t stream);

/* ---- input pipeline tail (csrc/preproc.cu) -------------------------------
 * Per-step feed of examples/v1/dist-mnist/dist_mnist.py:275-281. */
int tfop_u8_to_bf16_normalize(const void* in_u8, void* out_bf16, size_t n, int channels,
                              const float* mean, const float* stdv, cudaStream_t stream);

/* uint8 [N,H,W,3] -> bf16 [N,H/2+3,W/2+3,16]: space-to-depth layout under which the
 * ResNet 7x7/2 stem conv is a 4x4/1 conv over 16 channels (no reference counterpart). */
int tfop_stem_space_to_depth(const void* in_u8, void* out_bf16, int N, int H, int W,
                             const float* mean, const float* stdv, cudaStream_t stream);

/* ---- fused BatchNorm / ReLU / residual, NHWC bf16 (csrc/bn.cu) -----------
 * No reference counterpart (the reference ships no kernels); compute side of
 * the PyTorchJob ResNet-50 config. `workspace` = tfop_bn_workspace_bytes()
 * bytes of zero-initialised device memory, reusable across layers on one
 * stream. */
size_t tfop_bn_workspace_bytes(void);
int tfop_bn_fwd_stats(const void* x, size_t M, int C, const void* gamma, const void* beta,
                      int param_dtype, float* running_mean, float* running_var, float momentum,
                      float eps, float* scale_shift, float* save_mean, float* save_invstd,
                      void* workspace, cudaStream_t stream);
/* Allocation-free variant of tfop_bn_fwd_stats. It has no counterpart in the reference.
 * For callers that must not allocate device memory (the GPU is owned by the host
 * application's pool allocator, or cudaMalloc's device synchronisation is unacceptable):
 * the scratch is a built-in 33,024-byte __device__ array, so C <= 4096 channels (larger C is
 * an error). One array per device, shared by every call and cleared at the start of each:
 * calls that overlap on the GPU race, keep them on one stream. The first call on a device
 * resolves the array's address; make it before any stream capture. */
int tfop_bn_fwd_stats_noalloc(const void* x, size_t M, int C, const void* gamma,
                              const void* beta, int param_dtype, float* running_mean,
                              float* running_var, float momentum, float eps,
                              float* scale_shift, float* save_mean, float* save_invstd,
                              cudaStream_t stream);
int tfop_bn_fwd_apply(const void* x, const void* res, void* y, void* mask,
                      const float* scale_shift, size_t M, int C, int relu, cudaStream_t stream);
int tfop_bn_bwd_reduce(const void* dy, const void* x, const void* mask, const float* save_mean,
                       const float* save_invstd, size_t M, int C, void* dgamma, void* dbeta,
                       int param_dtype, float* coef, void* workspace, cudaStream_t stream);
int tfop_bn_bwd_apply(const void* dy, const void* x, const void* mask, const float* save_mean,
                      const float* save_invstd, const void* gamma, int param_dtype,
                      const float* coef, void* dx, void* dres, size_t M, int C,
                      cudaStream_t stream);

/* Stem tail fused (csrc/bn.cu): y = maxpool3x3/2/pad1(relu(bn(x))) without materialising the BN
 * output; forward takes scale_shift from tfop_bn_fwd_stats, backward gathers the pooled gradient
 * through the argmax codes inside both BN backward passes. NHWC bf16, C % 64 == 0. No reference
 * counterpart (the reference ships no kernels). */
int tfop_stem_bn_relu_pool_fwd(const void* x, const float* scale_shift, void* y, void* idx, int N,
                               int H, int W, int C, cudaStream_t stream);
int tfop_stem_bn_pool_bwd_reduce(const void* dpool, const void* idx, const void* x,
                                 const float* save_mean, const float* save_invstd, int N, int H,
                                 int W, int C, void* dgamma, void* dbeta, int param_dtype,
                                 float* coef, void* workspace, cudaStream_t stream);
int tfop_stem_bn_pool_bwd_apply(const void* dpool, const void* idx, const void* x,
                                const float* save_mean, const float* save_invstd,
                                const void* gamma, int param_dtype, const float* coef, void* dx,
                                int N, int H, int W, int C, cudaStream_t stream);

/* ---- MaxPool 3x3/2 pad 1, NHWC bf16 (csrc/pool.cu) -----------------------
 * No reference counterpart; ResNet stem pool. idx: 1 byte per output element. */
int tfop_maxpool3x3s2_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C,
                          cudaStream_t stream);
int tfop_maxpool3x3s2_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C,
                          cudaStream_t stream);

/* ---- K4/K5: parameter-server embedding over peer memory (csrc/ps.cu) ------
 * The variable pull / gradient push + optimizer-on-PS of
 * examples/v1/dist-mnist/dist_mnist.py:161-208. Row r lives on shard r % num_shards at local
 * row r / num_shards; weight/accum are the shard bases as mapped in the calling process. */
typedef struct tfop_ps_table {
    float* weight[8];
    float* accum[8];
    int num_shards;
    int dim;                  /* multiple of 4 */
    long long rows_per_shard;
} tfop_ps_table;
int tfop_ps_gather(const tfop_ps_table* table, const void* ids_i64, size_t n_ids, void* out,
                   int out_dtype, int max_blocks, cudaStream_t stream);
int tfop_ps_push_adagrad(const tfop_ps_table* table, const void* uniq_ids_i64,
                         const void* seg_off_i64, const void* perm_i64, size_t n_uniq,
                         const void* grad, int grad_dtype, float lr, float eps, float grad_scale,
                         int max_blocks, cudaStream_t stream);
/* The same push with duplicate ids merged on the device (hash insert, compact, sum, apply:
 * four kernels, no sort, no host round trip, no allocation). ids as the batch names them
 * (negative = padding). The six scratch arrays are sized by tfop_ps_dedup_workspace; before the
 * first use fill keys with 0xFF and the rest with zero - every call leaves them in that state. */
int tfop_ps_set_push_mode(int fetch);   /* 1: fetch-and-add accumulators; 0 (default): load + RED */
int tfop_ps_dedup_workspace(size_t n_ids, int dim, size_t* sizes6);
int tfop_ps_push_adagrad_dedup(const tfop_ps_table* table, const void* ids_i64, size_t n_ids,
                               const void* grad, int grad_dtype, float lr, float eps,
                               float grad_scale, void* ws_keys, void* ws_slot_of, void* ws_dense,
                               void* ws_uniq, void* ws_count, void* ws_sums, int max_blocks,
                               cudaStream_t stream);

/* ---- gradient histogram build for boosted trees (csrc/hist.cu) --------------
 * The per-worker step in front of the rabit histogram all-reduce of
 * examples/xgboost/xgboostjob.yaml:19-25. bins: uint8 [rows, F]; out: fp32 [F][256][2],
 * zeroed by the caller; row_idx: int32 rows of the node, or NULL for all rows. */
int tfop_hist_build(const void* bins_u8, const float* grad, const float* hess,
                    const int* row_idx, size_t n_rows, int F, float* out, cudaStream_t stream);

/* ---- K9: fused transformer-block elementwise ops (csrc/layernorm.cu) ------
 * No reference counterpart (the reference ships no model code); the op list is BASELINE.json
 * config 3. All tensors bf16, 16-byte aligned, H % 8 == 0.
 * tfop_ln_fwd: y = LayerNorm(h + bias + res) * gamma + beta in one pass; s_out (the bf16 LN
 *   input), mean, rstd (fp32 [rows]) are saved for backward; bias / res / s_out may be NULL.
 * tfop_ln_bwd: one pass giving dx and, accumulated with fp32 RED into caller-zeroed [H]
 *   buffers, dgamma, dbeta and dbias = colsum(dx) (the bias gradient of the producing linear
 *   layer; may be NULL). H <= 1536 for both.
 * tfop_bias_gelu_fwd: y = gelu_erf(h + bias).
 * tfop_bias_act_bwd: act=1: dh = dy * gelu'(h + bias), dbias += colsum(dh); act=0: dbias +=
 *   colsum(dy) only. */
int tfop_ln_fwd(const void* h, const void* bias, const void* res, const void* gamma,
                const void* beta, void* y, void* s_out, float* mean, float* rstd, int rows, int H,
                float eps, cudaStream_t stream);
int tfop_ln_bwd(const void* dy, const void* s, const float* mean, const float* rstd,
                const void* gamma, void* dx, float* dgamma, float* dbeta, float* dbias, int rows,
                int H, cudaStream_t stream);
/* Test hook: 1 routes every LayerNorm backward through the warp-per-row kernel (the default for
 * H % 256 == 0 is the tiled kernel: H/8 threads per row, 4 rows per thread and trip). */
int tfop_ln_force_warp_rows(int on);
int tfop_ln_set_bwd_rows(int rows);   /* rows per thread and trip of the tiled backward: 2 (default) or 4 */
int tfop_bias_gelu_fwd(const void* h, const void* bias, void* y, int rows, int H,
                       cudaStream_t stream);
int tfop_bias_act_bwd(const void* dy, const void* h, const void* bias, void* dh, float* dbias,
                      int rows, int H, int act, cudaStream_t stream);

/* ---- K10: softmax cross-entropy with ignore_index over bf16 logits (csrc/xent.cu) ----
 * Mean over labelled rows (int64 labels). Forward reads only labelled rows and saves one fp32
 * log-sum-exp per row; backward writes bf16 dlogits (zeros for unlabelled rows). inv_n and
 * grad_out are device scalars. No reference counterpart (BASELINE.json config 3). */
int tfop_xent_fwd(const void* logits, long ld, const void* labels, int rows, int V,
                  long long ignore_index, const float* inv_n, float* row_loss, float* lse,
                  cudaStream_t stream);
int tfop_xent_bwd(const void* logits, long ld, const void* labels, int rows, int V,
                  long long ignore_index, const float* inv_n, const float* lse,
                  const float* grad_out, void* dlogits, long ld_out, cudaStream_t stream);

/* ---- K11: ResNet stem as an implicit GEMM on tcgen05 (csrc/conv_stem.cu) ----
 * y[N,OH,OW,64] = conv4x4/1(x[N,OH+3,OW+3,16], w[64,4,4,16]), bf16 NHWC / KRSC; x comes from
 * tfop_stem_space_to_depth. The im2col operand is read straight from x through a 5-D TMA map
 * with overlapping windows; the 32 KB weight stays in shared memory. OH % 8 == 0, OW % 16 == 0.
 * No reference counterpart (BASELINE.json config 2 names the model). */
/* dw[64,4,4,16] fp32, zeroed by the caller: weight gradient of the stem (split-K over all CTAs,
 * both operands MN-major straight from dy and the overlapping-window map of x, RED at the end). */
int tfop_conv_stem_s2d_wgrad(const void* x, const void* dy, float* dw, int n, int oh, int ow,
                             cudaStream_t stream);
int tfop_conv_stem_set_mode(int mode); /* test hook: 0 = 4-D window map (SW128), 1 = 5-D (SW32) */
int tfop_conv_stem_s2d_fwd(const void* x, const void* w, void* y, int n, int oh, int ow,
                           cudaStream_t stream);

/* bf16 transpose out[C,R] = in[R,C]^T (csrc/transpose.cu); feeds the K-major operands of the
 * weight-gradient GEMM. No reference counterpart. */
int tfop_transpose_bf16(const void* in, void* out, int R, int C, int ld_in, int ld_out,
                        cudaStream_t stream);

/* ---- K7: tcgen05 / TMEM / TMA bf16 GEMM (csrc/gemm_tcgen05.cu) -----------
 * C[M,N] = epi(A[M,K] * B[N,K]^T + bias); the dense layers of
 * examples/v1/dist-mnist/dist_mnist.py:188-192. epilogue: 0 none, 1 bias,
 * 2 bias+GeLU(erf), 3 bias+ReLU, 4 bias+GeLU(tanh). tile_n: 0 auto/128/256. */
int tfop_gemm_bf16_tn(const void* A, const void* B, void* C, const void* bias, int bias_dtype,
                      int M, int N, int K, int lda, int ldb, int ldc, int out_dtype,
                      int epilogue, int tile_n, cudaStream_t stream);
/* General form: a_mn / b_mn = 1 means that operand is stored transposed ([K, M] resp. [K, N]
 * row-major, "MN-major") and is fed to the tensor core as is - dgrad (x @ W, b_mn) and wgrad
 * (dY^T @ X, a_mn + b_mn) of a linear layer need no transposed copies. */
int tfop_gemm_bf16(const void* A, const void* B, void* C, const void* bias, int bias_dtype,
                   int M, int N, int K, int lda, int ldb, int ldc, int out_dtype, int epilogue,
                   int tile_n, int a_mn, int b_mn, cudaStream_t stream);
/* -1 auto (default: pairs whenever there are >= 2 m-blocks), 0 never, 1 always: run the GEMM
 * CTAs as cta_group::2 pairs (2-CTA clusters; one tcgen05.mma covers 256 rows, each SM stages
 * only half of the B tile). */
int tfop_gemm_set_cluster_mode(int mode);
/* Split-K (several CTAs per output tile on different K ranges, partial tiles added into a
 * zeroed C by TMA reduce-add stores): 0 = heuristic (used when the tiles alone fill < 75 % of a
 * wave, no epilogue), -1 = never, n > 0 = force n splits. */
int tfop_gemm_set_splits(int s);
/* Benchmark hook: band height of the tile order, in scheduling units (0 = built-in default). */
int tfop_gemm_set_group_m(int g);
/* Test hook: 1 forces the per-thread direct C stores even where the TMA-store epilogue
 * (32x128-byte swizzled boxes, used when ldc*elt and C are 16-byte multiples) applies. */
int tfop_gemm_force_direct_store(int on);

/* ---- fused GEMM -> reduce-scatter -> all-gather (csrc/gemm_allreduce.cu) ---------
 * Gradient GEMM + NCCL all-reduce of keras-API/multi_worker_strategy-with-keras.py:76-77,
 * 117-120 as ONE kernel over peer memory: out[M,N] (bf16, in every rank's symmetric buffer at
 * out_off, row pitch ldo) = scale * sum_r A_r[M,K] * B_r[N,K]^T. The workspace at ws_off
 * (tfop_gemm_allreduce_workspace_bytes, zero-initialised once) holds partial-tile slots and
 * counters. */
size_t tfop_gemm_allreduce_workspace_bytes(int M, int N, int world);
int tfop_gemm_allreduce_set_timeout(long long cycles);
int tfop_gemm_allreduce_bf16_tn(tfop_comm* c, const void* A, const void* B, int M, int N, int K,
                                int lda, int ldb, int ldo, float scale, size_t ws_off,
                                size_t out_off, int max_ctas, cudaStream_t stream);
/* NVLS flavour: out region in the multicast-bound buffer, partials stay local (stored by the
 * cta_group::2 GEMM's TMA epilogue), reducer warps in the same kernel let the NVSwitch reduce and
 * broadcast the tiles as they complete. a_mn / b_mn as in tfop_gemm_bf16 ((1,1) = dY^T X from the
 * stored activations). ws_off: tfop_gemm_allreduce_nvls_workspace_bytes(M, N, world) zeroed bytes
 * of counters in the IPC buffer, tied to one (M, N, world) and one tile width. N, ldo: multiples
 * of 8. _tn = both operands K-major. Hooks: tile width 0 (heuristic) / 128 / 256; legacy = 1 runs
 * the first-generation kernel (one CTA per tile, three stages, K-major operands only). */
size_t tfop_gemm_allreduce_nvls_workspace_bytes(int M, int N, int world);
int tfop_gemm_allreduce_nvls_bf16(tfop_comm* c, const void* A, const void* B, int M, int N, int K,
                                  int lda, int ldb, int ldo, int a_mn, int b_mn, float scale,
                                  size_t ws_off, size_t out_off, int max_ctas, cudaStream_t stream);
int tfop_gemm_allreduce_nvls_bf16_tn(tfop_comm* c, const void* A, const void* B, int M, int N,
                                     int K, int lda, int ldb, int ldo, float scale, size_t ws_off,
                                     size_t out_off, int max_ctas, cudaStream_t stream);
int tfop_gemm_allreduce_set_tile_n(int tile_n);
int tfop_gemm_allreduce_set_legacy(int on);
int tfop_gemm_allreduce_emulated(tfop_comm** comms, int world, const void* const* As,
                                 const void* const* Bs, int M, int N, int K, int lda, int ldb,
                                 int ldo, float scale, size_t ws_off, size_t out_off,
                                 cudaStream_t stream);

/* ---- NHWC bf16 convolutions on tcgen05 implicit GEMMs (csrc/conv_tc.cu) --------------------
 * The Conv2D layers of keras-API/multi_worker_strategy-with-keras.py:42-60 (cuDNN in the
 * reference stack). x[N,H,W,Cin], w[Cout,k,k,Cin] (KRSC), y[N,OH,OW,Cout]; k in {1,3}, stride in
 * {1,2}, padding k/2, channel counts multiples of 64. `stats` (may be NULL): [2,Cout] fp32 to
 * which the per-channel sum and sum of squares of y are ADDED (BatchNorm statistics without a
 * pass over y, see tfop_bn_finalize). tfop_conv_dgrad with accumulate != 0 ADDS the gradient into
 * dx (the shortcut branch's gradient) by TMA reduce-add. tfop_conv_wgrad adds fp32 partial tiles
 * into dw[Cout,k,k,Cin], which the caller zeroes. */
int tfop_conv_fwd(const void* x, const void* w, void* y, float* stats, int N, int H, int W,
                  int Cin, int Cout, int ksize, int stride, cudaStream_t stream);
int tfop_conv_dgrad(const void* dy, const void* w, void* dx, int N, int H, int W, int Cin,
                    int Cout, int ksize, int stride, int accumulate, cudaStream_t stream);
int tfop_conv_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin,
                    int Cout, int ksize, int stride, cudaStream_t stream);
/* Test hook: -1 auto, 0 never pair CTAs (cta_group::2), 1 pair wherever the tile allows. */
int tfop_conv_set_cluster_mode(int mode);
/* BatchNorm statistics from sums[2][C] a convolution epilogue accumulated over M rows: same
 * outputs as tfop_bn_fwd_stats; clears `sums` for the next step. */
int tfop_bn_finalize(float* sums, size_t M, int C, const void* gamma, const void* beta,
                     int param_dtype, float* running_mean, float* running_var, float momentum,
                     float eps, float* scale_shift, float* save_mean, float* save_invstd,
                     cudaStream_t stream);

/* ---- multi-head self-attention, head dimension 64, no mask (csrc/attention.cu) --------------
 * Worker-side math of the BERT-large TFJob (the reference's trainers compute in
 * examples/v1/dist-mnist/dist_mnist.py:188-192). qkv[B*S, 3H] = [Q | K | V] packed as the
 * projection GEMM writes it, H = heads * 64, S a multiple of 128; out[B*S, H];
 * lse[B, heads, S] fp32 (base-2 log-sum-exp). tfop_attn_bwd writes dqkv[B*S, 3H] in the layout
 * of qkv; delta[B, heads, S] and dq32[B*S, H] (fp32) are caller-provided scratch. */
int tfop_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int heads, int head_dim,
                  cudaStream_t stream);
int tfop_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse,
                  void* dqkv, float* delta, float* dq32, int B, int S, int heads, int head_dim,
                  cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* TFOP_B200_H_ */
