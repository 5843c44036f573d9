/* dmp_b200.h -- C ABI of libdmp_b200.so (sm_100a kernels) and libdmp_b200_rt.so (host runtime).
 *
 * Conventions
 *   - every kernel entry point returns 0 on success, a CUDA error code or -1 on failure; dmp_last_error()
 *     then holds a message (thread local);
 *   - pointers are device pointers unless stated; `stream` is a cudaStream_t; work is enqueued, never synchronised;
 *   - activations are channels-last: x[M, C] with C contiguous (M = N*H*W); dtype codes: 0 = fp32, 1 = bf16, 2 = fp64;
 *   - "R/" cites /root/reference/code/distributed_training/, "T/" cites the installed torch sources.
 */
#ifndef DMP_B200_H
#define DMP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* dmp_stream_t; /* cudaStream_t */

/* ------------------------------------------------------------------ library state */
const char* dmp_last_error(void);
unsigned long long dmp_launch_count(void); /* kernels launched by this library since the last reset */
void dmp_reset_launch_count(void);
int dmp_version(void);
int dmp_device_cc(void);                                              /* e.g. 100 for sm_100 */
int dmp_l2_flush(void* buf, long long bytes, dmp_stream_t stream);    /* benchmark helper */
/* Stream ordering for the native Reducer launch path (csrc/core.cu): comm waits for cur and for the side streams that are part
 * of cur's capture (or for all of them outside a capture); cur waits for comm.  Counterpart of the event record / wait pairs
 * c10d's Reducer issues around each bucket's all-reduce (T/nn/parallel/distributed.py:1183-1304). */
int dmp_stream_fork(dmp_stream_t cur, dmp_stream_t comm, const dmp_stream_t* side, int n_side);
int dmp_stream_join(dmp_stream_t cur, dmp_stream_t comm);

/* ------------------------------------------------------------------ tcgen05 GEMM (csrc/kernels/gemm_tcgen05.cu)
 * bf16 x bf16 -> fp32 accumulate in TMEM.  layout 0: A[M,K] B[N,K] (1x1 conv / linear forward, R/model/mobilenetv2.py:17,21,55);
 * layout 1: A[M,K] B[K,N] (data gradient); layout 2: A[K,M] B[K,N] (weight gradient, split-K).
 * out_mode 0: bf16 store, 1: fp32 store, 2: fp32 atomic accumulate.  bias: optional fp32[N]. */
int dmp_gemm_bf16(int layout, const void* A, long long lda, const void* B, long long ldb, void* D, long long ldd, int M,
                  int N, int K, const float* bias, int out_mode, int splits, dmp_stream_t stream);
/* NT GEMM with bf16 output that also accumulates per-column {sum, sum of squares} of the rounded outputs into stats[2N]
 * (double): the 1x1 convolution hands its BatchNorm the batch statistics, see dmp_bn_apply_sums. */
int dmp_gemm_bf16_colstats(const void* A, long long lda, const void* B, long long ldb, void* D, long long ldd, int M, int N,
                           int K, double* stats, dmp_stream_t stream);
/* NN data-gradient GEMM of a 1x1 convolution whose input was act(BatchNorm(y_raw)) (R/model/mobilenetv2.py:33-34): stores the
 * ReLU-masked gradient G and accumulates {sum G, sum G*xhat} per column into stats[2N] -- the BatchNorm backward's reductions
 * come out of the producer of its incoming gradient; dmp_bn_bwd_dx_sums finishes the job in one elementwise pass. */
int dmp_gemm_bf16_dgrad_bnbwd(const void* A, long long lda, const void* B, long long ldb, void* G, long long ldd, int M, int N, int K,
                              const void* y_raw, long long ld_y, const float* mean, const float* invstd, const float* gamma,
                              const float* beta, int relu, double* stats, dmp_stream_t stream);
/* nh x nb independent GEMMs of one shape in one persistent launch; operand z = (zh, zb) lives at base + zh*stride_h + zb*stride_b
 * (elements).  The six matrix products of softmax attention, walking the fused [B,S,3,heads,d] projection in place.
 * NOT IN REFERENCE (BASELINE.json config 4: BERT-base pipeline). */
int dmp_gemm_bf16_batched(int layout, const void* A, long long lda, long long a_sh, long long a_sb, const void* B, long long ldb,
                          long long b_sh, long long b_sb, void* D, long long ldd, long long d_sh, long long d_sb, int M, int N, int K,
                          int nh, int nb, dmp_stream_t stream);
int dmp_gemm_ref_f32out(int layout, const void* A, long long lda, const void* B, long long ldb, float* D, long long ldd,
                        int M, int N, int K, const float* bias, int accumulate, dmp_stream_t stream); /* CUDA-core oracle */

/* ------------------------------------------------------------------ implicit-GEMM dense convolution (csrc/kernels/conv_igemm_tcgen05.cu)
 * 'same' stride-1 KxK convolution on channels-last bf16: 4-D TMA boxes with zero-filled halo feed tcgen05.mma, no im2col buffer.
 * NOT IN REFERENCE beyond its Cin=3 stem (R/model/mobilenetv2.py:52); serves the ResNet/VGG configs.  wp = weight permuted to
 * [Cout][kh][kw][Cin] (bf16).  The stride-1 data gradient is dmp_conv_igemm_fwd on dy with the flipped, transposed filter.
 * dmp_conv_igemm_wgrad accumulates into dwp (fp32, [Cout][kh][kw][Cin]) with split-K atomics. */
int dmp_conv_igemm_fwd(const void* x, const void* wp, const float* bias, void* y, int N, int H, int W, int Cin, int Cout, int KH,
                       int KW, int pad, dmp_stream_t stream);
int dmp_conv_igemm_wgrad(const void* dy, const void* x, float* dwp, int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad,
                         dmp_stream_t stream);
/* Strided (1 or 2) KxK forward with TMA element strides; stride-2 data gradient as four parity-class implicit GEMMs over dy (wd =
 * weight permuted to [Cin][kh][kw][Cout]); weight gradient with strided x boxes, optionally (kcrs) accumulated straight into the
 * [Cout][Cin][KH][KW] parameter gradient.  ResNet-18/50 down-sampling blocks. */
int dmp_conv_igemm_fwd_strided(const void* x, const void* wp, const float* bias, void* y, int N, int H, int W, int Cin, int Cout, int KH,
                               int KW, int pad, int stride, dmp_stream_t stream);
int dmp_conv_igemm_dgrad_s2(const void* dy, const void* wd, void* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad,
                            dmp_stream_t stream);
int dmp_conv_igemm_wgrad_strided(const void* dy, const void* x, float* dw, int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad,
                                 int stride, int kcrs, dmp_stream_t stream);

/* ------------------------------------------------------------------ optimizer (csrc/kernels/sgd.cu; R/data_parallel.py:90-91,111)
 * hp_dev = device {lr, momentum, weight_decay, grad_scale}. */
int dmp_sgd_flat(float* master, void* shadow_bf16, void* grad, int grad_dtype, float* momentum, long long n,
                 const float* hp_dev, int nesterov, int zero_grad, dmp_stream_t stream);
int dmp_cast_flat(const void* src, void* dst, long long n, int direction, dmp_stream_t stream);

/* ------------------------------------------------------------------ loss (csrc/kernels/xent.cu; R/utils.py:58-65,215-229)
 * rows: scratch of 3*B floats.  out: {mean loss, top-1 hits, top-5 hits}. */
int dmp_xent_fwd_bwd(const void* logits, const long long* targets, void* dlogits, float* rows, float* out, int B, int C,
                     int dtype, float grad_scale, dmp_stream_t stream);

/* ------------------------------------------------------------------ BatchNorm (csrc/kernels/bn.cu; R/model/mobilenetv2.py:18-22,53,56)
 * accum: caller-provided scratch of 2*C doubles (zeroed by the call). */
int dmp_bn_stats(const void* x, int dtype, long long M, int C, double* accum, dmp_stream_t stream);
int dmp_bn_finalize(const double* accum, double count, const float* gamma, const float* beta, float* running_mean,
                    float* running_var, float momentum, float eps, float* mean, float* invstd, float* scale, float* shift,
                    int C, dmp_stream_t stream);
int dmp_bn_apply(const void* x, int dtype, const float* scale, const float* shift, const void* residual, void* y,
                 long long M, int C, int relu, dmp_stream_t stream);
int dmp_bn_bwd_reduce(const void* dy, const void* y, const void* x, int dtype, const float* mean, const float* invstd,
                      long long M, int C, int relu, double* accum, dmp_stream_t stream);
int dmp_bn_bwd_dx(const void* dy, const void* y, const void* x, int dtype, const float* mean, const float* invstd,
                  const float* gamma, const double* accum_global, double count, void* dx, void* dres, long long M, int C,
                  int relu, dmp_stream_t stream);
int dmp_bn_bwd_params(const double* accum_local, float* dgamma, float* dbeta, int C, int accumulate, dmp_stream_t stream);
/* Single-launch (cooperative grid) train-mode BatchNorm: statistics -> grid barrier -> normalise(+residual,+ReLU) with running
 * statistics / batch counter update, and reductions -> grid barrier -> dx(+dres) with dgamma/dbeta.  ws: per-layer workspace of
 * dmp_bn_fused_ws_bytes(C) bytes, zeroed once (double-buffered accumulators + device-side parity, graph-replay safe).
 * Return 1 = shape not launchable cooperatively, use the split entry points above.  (T/nn/functional.py batch_norm semantics) */
long long dmp_bn_fused_ws_bytes(int C);
/* BatchNorm forward from sums a *_colstats producer accumulated: sums = {double[2C]; int ticket} of dmp_bn_sums_bytes(C),
 * zeroed once by the caller and re-zeroed by the kernel itself after every use. */
long long dmp_bn_sums_bytes(int C);
int dmp_bn_apply_sums(const void* x, int dtype, const void* residual, void* y, long long M, int C, const float* gamma,
                      const float* beta, float* running_mean, float* running_var, long long* num_batches_tracked,
                      float momentum, float eps, float* mean, float* invstd, double* sums, int relu, dmp_stream_t stream);
int dmp_bn_fwd_fused(const void* x, int dtype, const void* residual, void* y, long long M, int C, const float* gamma,
                     const float* beta, float* running_mean, float* running_var, long long* num_batches_tracked, float momentum,
                     float eps, float* mean, float* invstd, double* ws, int relu, dmp_stream_t stream);
int dmp_bn_bwd_fused(const void* dy, const void* y, const void* x, int dtype, const float* mean, const float* invstd,
                     const float* gamma, const float* beta, void* dx, void* dres, float* dgamma, float* dbeta,
                     int accumulate_params, long long M, int C, int relu, double* ws, dmp_stream_t stream);
/* Statistics + finalize for callers that may not allocate device memory (their process gave all of HBM to its own arena, or
 * bans allocator calls in its steady-state loop): same results as dmp_bn_stats + dmp_bn_finalize, but the double[2*C]
 * accumulator is a fixed 64 KiB array inside the library instead of a caller buffer.  No counterpart in the reference (it has
 * no native code and allocates freely through torch).  C <= dmp_bn_noalloc_max_channels() (4096), larger C is refused.
 * One array per GPU shared by all calls: each call clears what it uses; calls that can overlap on a GPU race.  The first call
 * on a device resolves the array's address (cudaGetSymbolAddress) -- make it at start-up, before any stream capture.
 * Calls no cudaMalloc/cudaFree. */
/* SyncBatchNorm, one cooperative launch per direction (csrc/comm/syncbn.cu): local sums -> grid barrier -> flag barrier with the
 * peers + peer loads of every rank's partial sums -> grid barrier -> normalise / dx.  ws_ptrs[W] = every rank's peer-mapped
 * workspace of dmp_syncbn_ws_bytes(C) (zeroed once), state_ptrs[W] = communicator states.  Return 1 = not launchable
 * cooperatively (use dmp_bn_stats + dmp_allreduce_fused + dmp_bn_finalize + dmp_bn_apply).  (R/Readme.md:151; torch
 * nn/modules/_functions.py:39-179 is the behaviour matched.) */
long long dmp_syncbn_ws_bytes(int C);
int dmp_syncbn_fwd(const void* x, const void* residual, void* y, long long M, int C, const float* gamma, const float* beta,
                   float* running_mean, float* running_var, long long* num_batches_tracked, float momentum, float eps, float* mean,
                   float* invstd, void* const* ws_ptrs, void* const* state_ptrs, int rank, int W, double* total, int relu,
                   dmp_stream_t stream);
int dmp_syncbn_bwd(const void* dy, const void* y, const void* x, const float* mean, const float* invstd, const float* gamma, void* dx,
                   void* dres, float* dgamma, float* dbeta, int accumulate_params, long long M, int C, int relu, void* const* ws_ptrs,
                   void* const* state_ptrs, int rank, int W, double* total, dmp_stream_t stream);
/* BatchNorm backward from reductions that already exist (sums = {sum g, sum g*xhat} + ticket, g already ReLU-masked): one
 * elementwise pass, dx may alias g; block 0 writes dgamma/dbeta, the last block clears the sums. */
int dmp_bn_bwd_dx_sums(const void* g, const void* x, int dtype, const float* mean, const float* invstd, const float* gamma, void* dx,
                       float* dgamma, float* dbeta, int accumulate_params, long long M, int C, double* sums, dmp_stream_t stream);
int dmp_bn_stats_finalize_noalloc(const void* x, int dtype, long long M, int C, const float* gamma, const float* beta,
                                  float* running_mean, float* running_var, float momentum, float eps, float* mean,
                                  float* invstd, float* scale, float* shift, dmp_stream_t stream);
int dmp_bn_noalloc_max_channels(void);

/* ------------------------------------------------------------------ depthwise 3x3 (csrc/kernels/dwconv.cu; R/model/mobilenetv2.py:19,33) */
int dmp_dwconv3x3_fwd(const void* x, const void* w, int w_dtype, void* y, int N, int H, int W, int C, int stride,
                      dmp_stream_t stream);
int dmp_dwconv3x3_fwd_colstats(const void* x, const void* w, int w_dtype, void* y, int N, int H, int W, int C, int stride,
                               double* stats, dmp_stream_t stream);
int dmp_dwconv3x3_dgrad(const void* dy, const void* w, int w_dtype, void* dx, int N, int H, int W, int C, int stride,
                        dmp_stream_t stream);
/* expand -> bn1 -> relu -> depthwise -> bn2 chain (R/model/mobilenetv2.py:32-33) without ever writing relu(bn1(.)) to memory:
 * the forward normalises the RAW expand output while loading it (coefficients from the producer's batch sums, which it also
 * publishes and clears), the weight gradient re-creates the activation the same way, and the data gradient applies the ReLU
 * mask and accumulates bn1's backward reductions {sum g, sum g*xhat}.  bf16 weights [C,1,3,3], bf16 NHWC activations. */
int dmp_dwconv3x3_fwd_bnin_colstats(const void* x_raw, const void* w_bf16, void* y, int N, int H, int W, int C, int stride,
                                    double* in_sums, const float* gamma, const float* beta, float eps, float momentum,
                                    float* running_mean, float* running_var, long long* num_batches_tracked, float* mean_out,
                                    float* invstd_out, double* out_stats, dmp_stream_t stream);
int dmp_dwconv3x3_dgrad_bnbwd(const void* dy, const void* w_bf16, void* g, int N, int H, int W, int C, int stride, const void* y_raw,
                              const float* mean, const float* invstd, const float* gamma, const float* beta, double* bwd_sums,
                              dmp_stream_t stream);
int dmp_dwconv3x3_wgrad_bnin(const void* dy, const void* x_raw, float* dw, int N, int H, int W, int C, int stride, int accumulate,
                             const float* mean, const float* invstd, const float* gamma, const float* beta, dmp_stream_t stream);
int dmp_dwconv3x3_wgrad(const void* dy, const void* x, float* dw, int N, int H, int W, int C, int stride, int accumulate,
                        dmp_stream_t stream);

/* ------------------------------------------------------------------ pooling and transformer row-wise kernels (csrc/kernels/elementwise.cu)
 * Max-pool for the ResNet / VGG family the reference lists (R/data_parallel.py:58-73); LayerNorm, GELU, softmax for the BERT-base
 * pipeline config (NOT IN REFERENCE).  bf16 activations, fp32 statistics; idx = winning tap per output element (uint8). */
int dmp_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride, int pad, dmp_stream_t stream);
int dmp_maxpool_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int k, int stride, int pad, dmp_stream_t stream);
int dmp_layernorm_fwd(const void* x, const void* res, const float* gamma, const float* beta, void* y, void* z, float* mean, float* rstd,
                      long long rows, int H, float eps, dmp_stream_t stream);
int dmp_layernorm_bwd(const void* dy, const void* z, const float* gamma, const float* mean, const float* rstd, void* dz, float* dgamma,
                      float* dbeta, long long rows, int H, dmp_stream_t stream);
int dmp_gelu_fwd(const void* x, void* y, long long n, dmp_stream_t stream);
int dmp_gelu_bwd(const void* dy, const void* x, void* dx, long long n, dmp_stream_t stream);
int dmp_softmax_fwd(const void* s, void* p, long long rows, int L, float scale, dmp_stream_t stream);
int dmp_softmax_bwd(const void* dp, const void* p, void* ds, long long rows, int L, float scale, dmp_stream_t stream);

/* ------------------------------------------------------------------ small ops (csrc/kernels/misc.cu; R/model/mobilenetv2.py:52,57,69-75) */
int dmp_stem_conv_fwd(const void* x, int x_layout, const void* w, int w_dtype, void* y, int N, int Cin, int H, int W,
                      int Cout, int KH, int KW, int stride, int pad, dmp_stream_t stream);
int dmp_stem_conv_wgrad(const void* dy, const void* x, int x_layout, float* dw, int N, int Cin, int H, int W, int Cout,
                        int KH, int KW, int stride, int pad, int accumulate, dmp_stream_t stream);
/* Large stems (7x7) as GEMMs on the tensor cores: bf16 im2col rows [pixel][Kp] (Kp % 8 == 0, zero beyond Cin*KH*KW) for
 * dmp_gemm_bf16 (forward: layout 0 with the filter packed by dmp_pack_rows_bf16; weight gradient: layout 2, fp32 accumulate).
 * Same op as dmp_stem_conv_*: R/data_parallel.py:58-73 lists ResNet, whose first layer is the 7x7 stride-2 convolution. */
int dmp_stem_im2col(const void* x, int x_layout, void* cols, int N, int Cin, int H, int W, int KH, int KW, int stride, int pad,
                    int Kp, dmp_stream_t stream);
int dmp_pack_rows_bf16(const void* src, int src_dtype, void* dst, int rows, int cols, int dst_pitch, dmp_stream_t stream);
/* w[Cout][Cin][KH][KW] (fp32 0 / bf16 1) -> bf16 operand of the implicit-GEMM convolutions: mode 0 [Cout][KH][KW][Cin],
 * 1 [Cin][KH][KW][Cout], 2 [Cin][flipped KH][flipped KW][Cout] (data gradients).  The dense 3x3 convolutions of the ResNet / VGG
 * models the reference lists (R/data_parallel.py:58-73). */
int dmp_pack_filter(const void* w, int w_dtype, void* out, int Cout, int Cin, int KH, int KW, int mode, dmp_stream_t stream);
int dmp_avgpool_fwd(const void* x, void* y, int N, int HW, int C, int relu, dmp_stream_t stream);
int dmp_avgpool_bwd(const void* dy, const void* x, void* dx, int N, int HW, int C, int relu, dmp_stream_t stream);
int dmp_linear_small_fwd(const void* x, const void* w, int w_dtype, const float* bias, float* y, int B, int N, int K,
                         dmp_stream_t stream);
int dmp_linear_small_bwd(const float* dy, const void* x, const void* w, int w_dtype, void* dx, float* dw, float* db, int B,
                         int N, int K, int accumulate, dmp_stream_t stream);
int dmp_add_act(const void* a, const void* b, void* y, long long n, int relu, dmp_stream_t stream);
int dmp_relu_bwd(const void* dy, const void* y, void* dx, long long n, dmp_stream_t stream);

/* ------------------------------------------------------------------ peer-memory collectives (csrc/comm/collectives.cu)
 * Pointer tables (`void* const*`) are HOST arrays of device addresses, one per rank, valid on the calling rank.
 * state[r]: rank r's communicator state of dmp_comm_state_bytes() bytes, zero-initialised once.
 * E3 DDP bucket all-reduce (T/nn/parallel/distributed.py:1255-1280 Reducer; default_hooks.py:26-32), D2 broadcast_coalesced
 * (T/nn/parallel/comm.py:50-67), D4 gather (:258), D6 reduce_add_coalesced (:124-168). */
long long dmp_comm_state_bytes(void);
int dmp_comm_max_ranks(void);
int dmp_comm_max_blocks(void);
int dmp_allreduce_fused(int algo, int dtype, void* const* bufs, void* const* state, int rank, int W, long long n,
                        float scale, void* out, float* master, float* momentum, void* shadow, const float* hp_dev,
                        int nesterov, void* const* param_peers, void* const* shadow_peers, int zero_after, int max_blocks,
                        dmp_stream_t stream);
int dmp_broadcast(const void* src, void* const* dst, int n_dst, void* mc_dst, void* const* dst_bf16, long long bytes,
                  void* const* state, int rank, int W, int is_src, int max_blocks, dmp_stream_t stream);
int dmp_gather_peer(void* const* src, const long long* bytes, int n_src, void* dst, int max_blocks, dmp_stream_t stream);
int dmp_copy_peer(const void* src, void* dst, long long bytes, int max_blocks, dmp_stream_t stream);
int dmp_reduce_add_peer(int dtype, void* const* src, int W, long long n, float scale, float* out_f32, float* master,
                        float* momentum, void* shadow, const float* hp_dev, int nesterov, int max_blocks,
                        dmp_stream_t stream);
/* Same, optionally zeroing every source after it was read (persistent DataParallel replicas accumulate into their arenas). */
int dmp_reduce_add_peer_z(int dtype, void* const* src, int W, long long n, float scale, float* out_f32, float* master, float* momentum,
                          void* shadow, const float* hp_dev, int nesterov, int zero_src, int max_blocks, dmp_stream_t stream);
int dmp_comm_status(void* state_local, dmp_stream_t stream); /* 0 healthy, 1+peer: barrier with that peer timed out */

/* NVLS multicast objects for one process driving several GPUs (csrc/comm/multicast.cu): per-device allocations bound to one
 * multicast address; a multimem.st to out_mc (dmp_broadcast's mc_dst) lands in every device's copy (out_unicast[i]). */
int dmp_mc_supported(const int* devs, int ndev);
int dmp_mc_create(const int* devs, int ndev, long long bytes, void** out_handle, void** out_unicast, void** out_mc,
                  long long* out_size);
int dmp_mc_destroy(void* handle);
/* Cross-process multicast object (one process per GPU).  Rank 0 creates it and hands *out_fd (a POSIX file descriptor) to the
 * other ranks (SCM_RIGHTS over a Unix socket), which import it; then every rank adds its device, ALL ranks synchronise, and
 * every rank binds + maps: *out_unicast is the rank's own copy, *out_mc the multicast address for multimem.* instructions. */
int dmp_mc_xp_create(int dev, int ndev_total, long long bytes, void** out_handle, int* out_fd, long long* out_size);
int dmp_mc_xp_import(int dev, int fd, int ndev_total, long long bytes, void** out_handle, long long* out_size);
int dmp_mc_xp_add_device(void* handle);
int dmp_mc_xp_bind_map(void* handle, void** out_unicast, void** out_mc);
int dmp_mc_xp_destroy(void* handle);
/* In-place fp32 all-reduce through the NVSwitch: multimem.ld_reduce of the caller's 1/W slice, scale, multimem.st. */
int dmp_allreduce_nvls(void* mc, void* const* state, int rank, int W, long long n, float scale, int max_blocks,
                       dmp_stream_t stream);
/* All-reduce + SGD through the switch: multimem.ld_reduce of the owned slice, momentum update, multimem.st of fp32 master, bf16
 * shadow and zeroed gradients into every copy.  (4 + 10)/W bytes per element on each GPU's link instead of 10 (W-1)/W. */
int dmp_allreduce_nvls_sgd(void* mc_grad, void* mc_master, void* mc_shadow, const float* master, float* momentum, const float* hp_dev,
                           int nesterov, void* const* state, int rank, int W, long long n, float scale, int max_blocks,
                           dmp_stream_t stream);

/* peer-memory plumbing */
int dmp_symm_malloc(long long bytes, void** out);
int dmp_symm_free(void* p);
int dmp_ipc_export(void* p, void* handle64);
int dmp_ipc_import(const void* handle64, void** out);
int dmp_ipc_close(void* p);
int dmp_ipc_handle_bytes(void);
int dmp_enable_peer_access(int peer);

/* ------------------------------------------------------------------ host runtime (libdmp_b200_rt.so; csrc/runtime/)
 * Reducer bookkeeping (T/nn/parallel/distributed.py:1183-1304; studied at /root/reference/Readme.md:145-157) and the pipeline
 * partitioner / schedules replacing R/model_parallel.py:102-103,129,143-144 and R/utils.py:34-210. */
int dmp_rt_version(void);
int dmp_bucket_assign(int n, const int64_t* nbytes, const int32_t* keys, const int64_t* limits, int n_limits,
                      int32_t* out_bucket_of_param);
void* dmp_reducer_create(int n_params, int n_buckets, const int32_t* bucket_of_param);
void dmp_reducer_destroy(void* h);
void dmp_reducer_prepare(void* h);
int dmp_reducer_mark_ready(void* h, int p, int unused, int32_t* out, int cap);
int dmp_reducer_finalize(void* h, int32_t* out_missing, int cap);
int dmp_reducer_num_launched(void* h);
long long dmp_reducer_iterations(void* h);
int dmp_reducer_ready_order(void* h, int32_t* out, int cap);
void dmp_reducer_unused_flags(void* h, int32_t* out);
/* Native launch path: per-bucket launch plans (struct DmpBucketPlan / DmpLaunchFns of csrc/runtime/reducer.cpp, mirrored by
 * ctypes structures in parallel/reducer.py).  dmp_reducer_grad_ready is the whole autograd hook: mark, decide, order the
 * communication stream, call the collective's entry point -- c10d's Reducer::autograd_hook -> mark_variable_ready ->
 * all_reduce_bucket, which the reference README dissects (/root/reference/Readme.md:145-157). */
struct DmpLaunchFns { /* entry points of libdmp_b200.so (or stand-ins with the same signatures) */
    int (*fork)(dmp_stream_t cur, dmp_stream_t comm, const dmp_stream_t* side, int n_side);
    int (*join)(dmp_stream_t cur, dmp_stream_t comm);
    int (*allreduce_fused)(int algo, int dtype, void* bufs, void* state, int rank, int world, long long n, float scale, void* out,
                           void* master, void* momentum, void* shadow, void* hp, int nesterov, void* master_ptrs, void* shadow_ptrs,
                           int zero_after, int max_blocks, dmp_stream_t stream);
    int (*allreduce_nvls)(void* mc, void* state, int rank, int world, long long n, float scale, int max_blocks, dmp_stream_t stream);
    int (*allreduce_nvls_sgd)(void* mc_grad, void* mc_master, void* mc_shadow, void* master, void* momentum, void* hp, int nesterov,
                              void* state, int rank, int world, long long n, float scale, int max_blocks, dmp_stream_t stream);
    int (*sgd_flat)(void* master, void* shadow, void* grad, int grad_dtype, void* momentum, long long n, void* hp, int nesterov,
                    int zero_grad, dmp_stream_t stream);
};
struct DmpBucketPlan { /* kind: 0 caller launches, 1 peer all-reduce (+SGD), 2 switch-side all-reduce, 3 switch-side all-reduce + SGD in
                          one kernel, 4 flat SGD only (one rank), 5 nothing, 6 switch-side all-reduce then flat SGD */
    int kind, algo, dtype, rank, world, nesterov, zero_after, max_blocks;
    long long numel;
    float scale;
    void *bufs, *state, *out, *master, *momentum, *shadow, *hp, *master_ptrs, *shadow_ptrs;
    void *mc_grad, *mc_master, *mc_shadow;
    void* grad;
};
int dmp_reducer_set_plans(void* h, const void* fns, const void* plans, int n_plans, dmp_stream_t comm, const dmp_stream_t* side,
                          int n_side);
void dmp_reducer_clear_plans(void* h);
int dmp_reducer_grad_ready(void* h, int p, int unused, dmp_stream_t cur, int32_t* out_caller, int cap);
int dmp_reducer_finish(void* h, dmp_stream_t cur, int32_t* out_missing, int cap);
long long dmp_reducer_native_launches(void* h);
int dmp_reducer_plan_bytes(void);
int dmp_reducer_fns_bytes(void);
int dmp_partition_balanced(int n, const double* cost, int W, int32_t* out_bounds);
int dmp_schedule_generate(int kind, int W, int s, int M, int32_t* out, int cap);
double dmp_schedule_simulate(int kind, int W, int M, double tf, double tb, int32_t* out_peak_inflight);

#ifdef __cplusplus
}
#endif
#endif /* DMP_B200_H */
