/* C ABI of libb200ddp_kernels.so -- the sm_100a kernel library of distributeddataparallel_b200.
 *
 * Conventions
 *   - every function launches asynchronously on `stream` and returns a cudaError_t as int
 *     (0 = launched); nothing here allocates device memory: callers pass outputs and scratch;
 *   - activations are NHWC ([N,H,W,C], C contiguous), filters KRSC ([K,R,S,C]), fp32 unless
 *     a suffix says otherwise;
 *   - "REF" cites /root/reference/dpp.py, "SURVEY Kn" the kernel inventory in SURVEY.md §2.5.
 */
#ifndef B200DDP_H_
#define B200DDP_H_

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- programmatic dependent launch ------------------------------------------------------
 * The kernels of the training step are launched with cudaLaunchAttributeProgrammaticStreamSerialization:
 * a kernel's prologue may overlap the tail of its predecessor in the stream, and it executes
 * griddepcontrol.wait before its first global-memory access.  Default on (env B200DDP_PDL=0 turns it off). */
void b200_set_pdl(int enabled);

/* ---- input pipeline --------------------------------------------------------------------
 * uint8 NHWC image batch -> fp32 NHWC, (x/255 - mean)/std: the ToTensor+Normalize transform
 * of REF:32, done on the device after a 4x smaller H2D copy. */
int b200_preprocess_u8(const uint8_t* in, float* out, size_t n, float mean, float stdv, cudaStream_t stream);

/* ---- convolution as implicit GEMM (cuDNN on the reference path, REF:50-52).
 * `workspace`/`counters` are split-K scratch: partial tiles and per-tile arrival tickets
 * (counters must be zero on entry and are left zero).  `stats` (optional, [2*K], pre-zeroed
 * or accumulating) receives per-output-channel sum and sum of squares for BatchNorm.
 * A Linear layer is the R=S=H=W=1 case. */
int b200_conv2d_fprop(const float* x, const float* w, const float* bias, float* y, float* stats,
                      int Nb, int H, int W, int C, int K, int R, int S, int stride, int pad,
                      float* workspace, size_t workspace_bytes, unsigned* counters, int max_counters,
                      cudaStream_t stream);
int b200_conv2d_dgrad(const float* dy, const float* w, float* dx,
                      int Nb, int H, int W, int C, int K, int R, int S, int stride, int pad,
                      float* workspace, size_t workspace_bytes, unsigned* counters, int max_counters,
                      cudaStream_t stream);
/* dw may point straight into a gradient-bucket slot (removes the reducer copy-in). */
int b200_conv2d_wgrad(const float* dy, const float* x, float* dw, int accumulate,
                      int Nb, int H, int W, int C, int K, int R, int S, int stride, int pad,
                      float* workspace, size_t workspace_bytes, unsigned* counters, int max_counters,
                      cudaStream_t stream);

/* Forward convolution that allocates nothing and needs no caller scratch.  No counterpart in the
 * reference (REF has no native code); it exists for hosts that cannot call cudaMalloc where they
 * need the op: code running while a sibling stream is being graph-captured, plug-ins inside a
 * runtime that owns the device memory, loops that have frozen their allocations.  Same
 * arithmetic and arguments as b200_conv2d_fprop minus the scratch parameters; split-K scratch is
 * a 4 MiB + 4 KiB array inside the library image (held by every GPU that loads the library,
 * until exit).
 *   Limit: the output (N*P*Q x K) may span at most 1024 tiles of 64x64; larger problems return
 *          cudaErrorInvalidValue and b200_conv2d_fprop_noalloc_error() explains why.
 *   First call on a device resolves the array's device address (cudaGetSymbolAddress) -- make
 *          that first call with the device current and before any stream capture begins.
 *   All calls share the one array: calls that can overlap on the GPU (different streams, no
 *          ordering between them) race on it.  Tickets are cleared at the start of every call. */
int b200_conv2d_fprop_noalloc(const float* x, const float* w, const float* bias, float* y, float* stats,
                              int Nb, int H, int W, int C, int K, int R, int S, int stride, int pad,
                              cudaStream_t stream);
const char* b200_conv2d_fprop_noalloc_error(void);
size_t b200_conv2d_fprop_noalloc_scratch_bytes(void);

/* ---- BatchNorm2d (+residual add, +ReLU) training kernels ------------------- */
int b200_bn_stats(const float* x, float* sums, int M, int C, cudaStream_t stream);
int b200_bn_fwd_apply(const float* x, const float* sums, const float* gamma, const float* beta,
                      const float* residual, float* out, float* running_mean, float* running_var,
                      long long* num_batches_tracked, float* save_mean, float* save_invstd,
                      float* zero_after, int zero_count, int M, int C, float momentum, float eps,
                      int relu, int training, cudaStream_t stream);
int b200_bn_bwd_reduce(const float* dy, const float* out, const float* x, const float* save_mean,
                       const float* save_invstd, float* sums, int M, int C, int relu, cudaStream_t stream);
int b200_bn_bwd_apply(const float* dy, const float* out, const float* x, const float* gamma,
                      const float* save_mean, const float* save_invstd, const float* sums, float* dx,
                      float* dres, float* dgamma, float* dbeta, int accumulate_param_grads,
                      float* zero_after, int zero_count, int M, int C, int relu, cudaStream_t stream);

/* Cluster BatchNorm (one launch, one 8-CTA cluster, tensor register-resident, fixed-order reduction through
 * distributed shared memory: bit-reproducible).  For M*C <= 131072 and C/4 dividing 512; otherwise returns
 * cudaErrorNotSupported and the caller uses the streaming kernels above.  The forward computes exact two-pass
 * statistics itself (no `sums` input): it is the variance-robust / deterministic alternative to statistics
 * accumulated by the convolution epilogue. */
int b200_bn_bwd_cluster(const float* dy, const float* out, const float* x, const float* gamma,
                        const float* save_mean, const float* save_invstd, float* dx, float* dres, float* dgamma,
                        float* dbeta, int accumulate_param_grads, float* zero_after, int zero_count, int M, int C,
                        int relu, cudaStream_t stream);
int b200_bn_fwd_cluster(const float* x, const float* gamma, const float* beta, const float* residual, float* out,
                        float* running_mean, float* running_var, long long* num_batches_tracked, float* save_mean,
                        float* save_invstd, float* zero_after, int zero_count, int M, int C, float momentum,
                        float eps, int relu, cudaStream_t stream);

/* reduce + apply in one launch (blocks meet on `tickets`, 2 zero-initialised words left at zero) */
int b200_bn_bwd_fused(const float* dy, const float* out, const float* x, const float* gamma,
                      const float* save_mean, const float* save_invstd, float* sums, float* dx, float* dres,
                      float* dgamma, float* dbeta, int accumulate_param_grads, float* zero_after,
                      int zero_count, unsigned* tickets, int M, int C, int relu, cudaStream_t stream);

/* ---- pooling --------------------------------------------------------------- */
int b200_maxpool3x3s2_fwd(const float* x, float* y, uint8_t* idx, int Nb, int H, int W, int C, cudaStream_t stream);
int b200_maxpool3x3s2_bwd(const float* dy, const uint8_t* idx, float* dx, int Nb, int H, int W, int C,
                          cudaStream_t stream);
int b200_avgpool_fwd(const float* x, float* y, int Nb, int HW, int C, cudaStream_t stream);
int b200_avgpool_bwd(const float* dy, float* dx, int Nb, int HW, int C, cudaStream_t stream);

/* ---- loss: mean softmax cross-entropy, forward and logits gradient in one launch
 * (nn.CrossEntropyLoss at REF:40). `row_loss_scratch` needs B floats. */
int b200_softmax_xent(const float* logits, const long long* target, float* loss, float* dlogits,
                      float* row_loss_scratch, int B, int V, cudaStream_t stream);

/* ---- small elementwise helpers ----------------------------------------------------------------- */
int b200_scale(const float* a, float* out, size_t n, const float* scale_dev, float scale_host, cudaStream_t stream);
int b200_add(const float* a, const float* b, float* out, size_t n, cudaStream_t stream);
int b200_colsum(const float* a, float* out, int M, int N, int accumulate, cudaStream_t stream);
int b200_colsum_bf16(const void* a, void* out, int M, int N, int accumulate, cudaStream_t stream);
/* one-launch bias gradient: acc (fp32 [>= N]) and ticket (1 word) are zero on entry and left zero on exit */
int b200_colsum_bf16_fused(const void* a, void* out, float* acc, unsigned* ticket, int M, int N, cudaStream_t stream);
int b200_colsum_bf16_acc(const void* a, float* acc, int M, int N, cudaStream_t stream);   /* tall matrices: fp32 atomics */

/* ---- fused multi-tensor optimizers (optim.SGD at REF:41,53) -------------------------
 * recs: device array of {float* p, g, s0, s1; int64 n; int64 dtype (0 fp32, 1 bf16 p/g)}; chunks: device array of
 * {int tensor, int chunk}; hyper: device floats (layout in csrc/kernels/optim.cu). */
int b200_optim_chunk_elems(void);
int b200_optim_rec_bytes(void);
int b200_fused_sgd(const void* recs, const void* chunks, int num_chunks, const float* hyper, cudaStream_t stream);
int b200_fused_adam(const void* recs, const void* chunks, int num_chunks, float* hyper,
                    float* const* step_ptrs, int num_tensors, cudaStream_t stream);

/* ---- tcgen05 / TMEM tensor-core path (csrc/kernels/gemm_tcgen05.cu) ------------
 * Dense GEMM with fp32 output; inputs fp32 (TF32 math) or bf16.  mode 0: A[M,K] B[N,K]^T,
 * mode 1: A[M,K] B[K,N], mode 2: A[K,M]^T B[K,N].  Dimensions that are contiguous in memory
 * must be multiples of 4 (fp32) / 8 (bf16).  `error` (optional device word) is set if an
 * internal wait times out.  The conv entry points return cudaErrorNotSupported (801) for
 * shapes that need the FFMA fallback (channels not a multiple of 4). */
/* b200_gemm_tc_ex with a fused epilogue on the persistent 2-CTA kernel (bf16 in/out): epi 1 = GELU forward
 * (aux_out <- pre-activation, out <- gelu), 2 = GELU backward (out = result * gelu'(aux)), 3 = residual add
 * (out = result + aux).  801 when the shape is not eligible: run the unfused kernels instead. */
int b200_gemm_tc_epi(const void* A, const void* B, void* out, const void* bias, const void* aux, void* aux_out, int epi,
                     int M, int N, int K, int mode, int flags, const long long* desc, float* workspace,
                     size_t workspace_bytes, unsigned* counters, int max_counters, unsigned* error, cudaStream_t stream);
int b200_gemm_tc(const void* A, const void* B, float* out, const float* bias, int M, int N, int K, int mode,
                 int bf16, int accumulate, float* workspace, size_t workspace_bytes, unsigned* counters,
                 int max_counters, unsigned* error, cudaStream_t stream);
/* general form: explicit leading dimensions, batching (attention) and bf16 output; see the .cu */
int b200_gemm_tc_ex(const void* A, const void* B, void* out, const void* bias, int M, int N, int K, int mode,
                    int flags, const long long* desc, float* workspace, size_t workspace_bytes,
                    unsigned* counters, int max_counters, unsigned* error, cudaStream_t stream);
/* weight-gradient GEMM fused with the all-reduce (average) of its output over NVLink peers: the
 * epilogue pushes partial tiles to their owner rank and finished tiles into every rank's copy
 * (SURVEY W12).  `peer` = host copy of the PeerReduce struct in csrc/kernels/gemm_tcgen05.cu. */
int b200_gemm_tc_peer_bn(int N);
int b200_peer_reduce_bytes(void);
int b200_gemm_tc_peer(const void* A, const void* B, const void* bias, int M, int N, int K, int mode, int flags,
                      const long long* desc, const void* peer, unsigned* error, cudaStream_t stream);
void b200_gemm_tc_use_tma(int enabled);   /* dense operands via TMA (default) or cp.async gathers */
void b200_gemm_tc_use_s2_dgrad(int enabled);    /* stride-2 dgrad of >= 16384-pixel inputs as four stride-1 parity problems (default on) */
void b200_gemm_tc_use_filter3d(int enabled);    /* stride-1 dgrad fetches the KRSC filter through a 3-D TMA map (default on) */
void b200_gemm_tc_use_mn32_tma(int enabled);    /* fp32 MN-major operands fetched by TMA (default on) */
void b200_gemm_tc_cluster_reduce(int enabled);  /* split-K: splits of a tile form a cluster, reduce via DSMEM (for the calls that follow) */
void b200_gemm_tc_split_cap(int cap);            /* upper bound on split-K for the calls that follow (default 32) */
void b200_gemm_tc_use_im2col_tma(int enabled);  /* im2col-mode TMA for convolution activations (default on) */
void b200_gemm_tc_use_2cta(int mode);     /* cta_group::2 cluster kernels: 0 off, 1 one tile per cluster, 2 persistent (default) */
int b200_conv2d_fprop_tc(const float* x, const float* w, float* y, float* stats, int Nb, int H, int W, int C,
                         int K, int R, int S, int stride, int pad, float* workspace, size_t workspace_bytes,
                         unsigned* counters, int max_counters, unsigned* error, cudaStream_t stream);
int b200_conv2d_dgrad_tc(const float* dy, const float* w, float* dx, int Nb, int H, int W, int C, int K,
                         int R, int S, int stride, int pad, float* workspace, size_t workspace_bytes,
                         unsigned* counters, int max_counters, unsigned* error, cudaStream_t stream);
int b200_conv2d_wgrad_tc(const float* dy, const float* x, float* dw, int accumulate, int Nb, int H, int W,
                         int C, int K, int R, int S, int stride, int pad, float* workspace,
                         size_t workspace_bytes, unsigned* counters, int max_counters, unsigned* error,
                         cudaStream_t stream);

/* ---- transformer blocks (GPT-2 configuration): fp32 math on fp32 / bf16 storage -------- */
int b200_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                       int M, int D, float eps, int is_bf16, cudaStream_t st);
/* bf16, D in {768, 1024, 1280}: dx (+ dres: gradient of the residual branch), dgamma, dbeta in one launch; acc
 * (fp32 [2*D]) and ticket (1 word) zero on entry, left zero on exit; 801 for other widths */
int b200_layernorm_bwd_fused(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                             const void* dres, void* dx, void* dgamma, void* dbeta, float* acc, unsigned* ticket, int M,
                             int D, cudaStream_t st);
int b200_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                       void* dx, float* dgamma_acc, float* dbeta_acc, int M, int D, int is_bf16, cudaStream_t st);
int b200_cast_from_f32(const float* in, void* out, size_t n, int accumulate, int is_bf16, cudaStream_t st);
int b200_gelu_fwd(const void* x, void* y, size_t n, int is_bf16, cudaStream_t st);
int b200_gelu_bwd(const void* x, const void* dy, void* dx, size_t n, int is_bf16, cudaStream_t st);
int b200_causal_softmax_fwd(void* s, int T_len, size_t rows, float scale, int is_bf16, cudaStream_t st);
int b200_causal_softmax_bwd(const void* p, void* dp, int T_len, size_t rows, float scale, int is_bf16, cudaStream_t st);
/* Fused causal self-attention, head dimension 64, bf16 (csrc/kernels/attention_tcgen05.cu): reads q|k|v in place
 * from qkv[B*T, 3*H*64]; out[B*T, H*64]; lse[B*H, T] (log2 domain) feeds the backward.  The backward needs two
 * fp32 scratch arrays: delta[B*H*T] and dq32[B*T*H*64].  801 = TMA descriptor unavailable (use the GEMM path).
 * No counterpart on the reference's path (dpp.py trains a CNN); stock PyTorch uses a flash-attention library. */
int b200_attention_fwd(const void* qkv, void* out, float* lse, int B, int H, int T, unsigned* error, cudaStream_t st);
int b200_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                       float* dq32, void* dqkv, int B, int H, int T, unsigned* error, cudaStream_t st);
int b200_embedding_fwd(const long long* idx, const void* wte, const void* wpe, void* out, size_t rows, int T_len,
                       int D, int is_bf16, cudaStream_t st);
int b200_embedding_bwd(const long long* idx, const void* dout, float* dwte_acc, float* dwpe_acc, size_t rows,
                       int T_len, int D, int is_bf16, cudaStream_t st);
int b200_scale_inplace(void* x, size_t n, const float* scale_dev, int is_bf16, cudaStream_t st);
int b200_xent_rows(const void* logits, const long long* target, void* dlogits, float* row_loss, float* loss,
                   int B, int V, int ld, int is_bf16, cudaStream_t st);

/* ---- peer-to-peer collectives over NVLink-mapped symmetric arenas (SURVEY K2-K6, W2-W7;
 * ProcessGroupNCCL on the reference path, REF:21).  `ctx` points to a HOST copy of the CommCtx
 * struct (csrc/kernels/comm.cu): arena and signal-pad base of every rank as mapped in this
 * process, local epoch counters, error word, rank/world/channel.  Objects are addressed by
 * byte offset inside the arena; all collectives are in place.
 * dtype 0 = fp32, 1 = bf16 (fp32 accumulate); algo 1 = one-shot, 2 = two-shot, 3 = NVLS (multimem through the
 * arena's multicast mapping).  Every in-kernel wait is bounded by the context's timeout_ns (0 = 10 minutes): a
 * collective whose peer does not arrive writes the context's error word, skips its peer traffic and NaN-fills its
 * share of the local result. */
int b200_comm_ctx_bytes(void);
int b200_comm_max_blocks(void);
size_t b200_comm_signal_bytes(void);
size_t b200_comm_epoch_bytes(void);
size_t b200_allreduce_oneshot_max_bytes(void);
int b200_allreduce(const void* ctx, size_t offset, size_t nbytes, int dtype, float scale, int algo,
                   int max_blocks, cudaStream_t stream);
/* small messages (<= b200_allreduce_push_max_bytes()): one flag exchange instead of two barriers; `stage_off` is the
 * arena offset of a per-channel staging object of b200_allreduce_push_stage_bytes() bytes */
size_t b200_allreduce_push_max_bytes(void);
size_t b200_allreduce_push_stage_bytes(void);
int b200_allreduce_push(const void* ctx, size_t offset, size_t nbytes, int dtype, float scale, size_t stage_off,
                        cudaStream_t stream);
/* fp32 data with a 16-bit wire format (wire 1 = bf16, 2 = fp16) and fp32 accumulation: the fused form of the
 * bf16/fp16 compress comm hooks.  nbytes % 32 == 0; staging object of nbytes/2 bytes at `stage_off`. */
int b200_allreduce_wire16(const void* ctx, size_t offset, size_t nbytes, int wire, float scale, size_t stage_off,
                          int max_blocks, cudaStream_t stream);
int b200_broadcast(const void* ctx, size_t offset, size_t nbytes, int root, int max_blocks, cudaStream_t stream);
int b200_barrier(const void* ctx, cudaStream_t stream);
/* device-timed logger: write %globaltimer (ns) to *dst on `stream`; the collective kernels stamp their own start / end
 * into the context's `stamps` pair when it is set */
int b200_stamp(unsigned long long* dst, cudaStream_t stream);
/* set-up helpers (these DO allocate / map memory; call once, outside any capture) */
int b200_arena_alloc(size_t bytes, void** out);
int b200_arena_free(void* p);
int b200_ipc_get_handle(void* p, unsigned char out[64]);
int b200_ipc_open_handle(const unsigned char in[64], void** out);
int b200_ipc_close_handle(void* p);

#ifdef __cplusplus
}
#endif
#endif /* B200DDP_H_ */
