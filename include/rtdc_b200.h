/* C ABI of librtdc_b200.so — the hand-written sm_100a kernels and native runtime pieces of
 * ray_torch_distributed_checkpoint_b200. Python binds these with ctypes (ops/_lib.py); any other
 * host language can do the same. Every function returns 0 on success, non-zero on failure, in which
 * case rtdc_last_error() holds a message for the calling thread. All kernels are enqueued on the
 * stream passed in (a cudaStream_t / CUstream as void*); nothing synchronises the device unless
 * its comment says so.
 *
 * "ref:" comments cite the reference (outerbounds/ray-torch-distributed-checkpoint) line whose
 * behaviour the entry point covers; the reference itself ships no native code.
 */
#ifndef RTDC_B200_H
#define RTDC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* rtdc_last_error(void);
/* Library/ABI version and the architecture the kernels were compiled for ("sm_100a"). */
int rtdc_version(void);
const char* rtdc_arch(void);

/* ------------------------------------------------------------------------------------------
 * GEMM on the 5th-generation tensor cores (tcgen05.mma kind::tf32, accumulators in TMEM, operands
 * staged by TMA with 128-byte swizzle), fp32 in / fp32 out.
 *   D[M,N] = epilogue( sum_k A(m,k) * B(n,k) )
 * precision = 3: "3xTF32" — each fp32 operand tile is split in shared memory into a tf32 head and
 *   a tf32 tail and three MMAs (tail*head + head*tail + head*head) accumulate in fp32, which
 *   recovers fp32-GEMM accuracy (~2^-22 relative per product) on the tensor cores.
 * precision = 1: single tf32 MMA (10-bit mantissa operands).
 * Covers ref: my_ray_module.py:99-106 (Linear+ReLU+Dropout forward, K1-K3), :159 (backward
 * dgrad/wgrad, K6), :279-282 (inference forward, K12).
 */
typedef struct rtdc_gemm_desc {
  /* operands */
  const float* a; int lda; int a_mn; /* a_mn=0: A is [M,K] row-major (K-major); 1: A is [K,M] row-major */
  const float* b; int ldb; int b_mn; /* b_mn=0: B is [N,K] row-major (K-major); 1: B is [K,N] row-major */
  int M, N, K;
  /* output: out[m*ldo + n]; columns N..n_store-1 are written as 0 (padding kept clean) */
  float* out; int ldo; int n_store;
  /* epilogue, applied in this order */
  const float* bias;      /* [N] added per column, or NULL */
  int relu;               /* max(x,0) */
  float drop_p;           /* inverted dropout with probability p (0 disables) */
  uint64_t seed;          /* Philox key */
  const uint32_t* step_ptr; /* device pointer to a step counter mixed into the Philox counter, or NULL */
  uint32_t step_imm;      /* used when step_ptr is NULL */
  uint32_t salt;          /* distinguishes layers sharing a seed */
  const float* mask_src; int ldm; float mask_scale; /* x *= (mask_src[m*ldm+n] > 0) ? mask_scale : 0, or NULL */
  /* execution */
  int precision;          /* 3 or 1 */
  int bn;                 /* N tile (0 = choose) */
  int splits;             /* split-K factor (0 = choose). >1 needs scratch */
  float* scratch; size_t scratch_bytes; /* split-K partial tiles, see rtdc_gemm_scratch_bytes */
  /* optional fused column sums of the output (bias gradient of the layer that consumes it): the epilogue writes, for
   * every 32-row group g of the output, sum_rows out[rows of g, n] to colsum_partial[g*ld_colsum + n]
   * (4*ceil(M/128) groups; rows past M contribute nothing). Sum those rows (e.g. rtdc_colsum) to get the total. */
  float* colsum_partial; int ld_colsum;
  /* optional exact-operand / affine-input folding (3xTF32 only). a_exact / b_exact != 0 promise that every element of
   * that operand has its low 13 mantissa bits clear (e.g. small integers such as raw uint8 pixels widened to fp32), so
   * the operand needs no low-order correction pass: the kernel skips its split and one of the three MMAs per K step
   * with no loss of accuracy. A broken promise silently degrades that operand to tf32 precision.
   * The accumulator is then mapped as  acc*alpha (alpha == 0 means 1)  + bias[n] + row_bias[m]*row_bias_scale
   * before relu/dropout/mask, which lets a caller fold an input normalisation x = s*raw + t into the layer:
   *   forward  W x + b          = s*(W raw) + (b + t*rowsum(W))          -> alpha = s, bias = rtdc_fold_input_bias(...)
   *   wgrad    dY^T x           = s*(dY^T raw) + t*colsum(dY) 1^T        -> alpha = s, row_bias = db, row_bias_scale = t */
  int a_exact, b_exact;
  float alpha;
  const float* row_bias; float row_bias_scale;
  /* optional bit-packed activation mask. mask_bits_out: the epilogue also writes, for every output row, one uint32 per 32
   * columns with bit q = (out[m, 32w + q] > 0) to mask_bits_out[m*ld_mask_out + w] (needs no K split). mask_bits_in: the
   * same layout consumed as the backward mask instead of mask_src (x *= bit ? mask_scale : 0): a dgrad then reads
   * 1/32 of the bytes of the activation it masks with. Leading dimensions in 32-bit words, >= ceil(N/32). */
  uint32_t* mask_bits_out; int ld_mask_out;
  const uint32_t* mask_bits_in; int ld_mask_in;
} rtdc_gemm_desc;

/* bf16 GEMM (tcgen05.mma kind::f16, fp32 accumulation in TMEM, TMA-fed, persistent): D[M,N] = act(alpha * sum_k A(m,k) B(n,k) + bias[n]).
 * a_mn / b_mn as in rtdc_gemm_desc (0: [rows, K] row-major; 1: [K, rows] row-major), so forward (x W^T), dgrad (dy W) and wgrad
 * (dy^T x) of a Linear layer need no transposed copies. Pointers 16-byte aligned, leading dimensions multiples of 8 elements.
 * out is bf16 (out_fp32 = 0) or fp32 [M, ldo]. splits > 1 (0 = choose; only for fp32 output without bias / activation) splits K over
 * CTAs that add into the zeroed output with fp32 atomics (wgrad: K = the batch's tokens). act: 0 none, 1 relu, 2 gelu(tanh), 3 gelu(erf).
 * The reference has no bf16 path (it trains its MLP in fp32, no autocast); this serves the bf16 GPT-2 / ResNet-50 configurations. */
typedef struct rtdc_gemm_bf16_desc {
  const void* a; int lda; int a_mn;
  const void* b; int ldb; int b_mn;
  int M, N, K;
  void* out; int ldo; int out_fp32;
  const void* bias; int bias_fp32;   /* [N] bf16 (or fp32), or NULL */
  int act;
  float alpha;                       /* 0 means 1 */
  int splits;
} rtdc_gemm_bf16_desc;
int rtdc_gemm_bf16(const rtdc_gemm_bf16_desc* d, void* stream);

/* Softmax cross-entropy over a wide vocabulary fused with its backward (the LM head of the GPT-2 configuration): per row of the
 * bf16 logits [rows, ld] (first `vocab` columns valid; 16-byte aligned, ld a multiple of 8) writes row_loss[row] =
 * logsumexp(row) - row[target] (natural log, fp32) and overwrites the row IN PLACE with (softmax - onehot) * *grad_scale in
 * bf16 — the A operand of the head's dgrad / wgrad GEMMs. Rows whose target is ignore_index (or outside [0, vocab)) get a zero
 * loss and a zero gradient. grad_scale is a DEVICE scalar (1 / number of counted rows for a mean reduction), so a caller
 * needs no host synchronisation to form it. One pass over HBM: a block stages its row in shared memory (rows up to 110 KB;
 * wider ones are re-read through L2). No reference counterpart (my_ray_module.py:141 is a 10-class loss). */
int rtdc_lm_head_ce_bf16(void* logits, long long ld, const int64_t* targets, int rows, int vocab, long long ignore_index,
                         const float* grad_scale, float* row_loss, void* stream);

/* bf16 LayerNorm over the last dimension (C a multiple of 8, <= 1024; rows 16-byte aligned), fp32 math, one warp per row.
 * Forward: y = (x - mean) * rstd * gamma + beta (beta may be NULL); mean / rstd [rows] fp32 are saved for the backward pass.
 * Backward: dx, and in the same pass over dy and x the weight / bias gradients (dgamma, dbeta bf16 [C]; dbeta may be NULL)
 * through per-block partial sums in `scratch` (rtdc_layernorm_bwd_scratch_bytes; no initialisation needed) that a second
 * small kernel adds in block order. No reference counterpart (the reference MLP has no normalisation layer). */
int rtdc_layernorm_fwd_bf16(const void* x, long long ldx, const void* gamma, const void* beta, float eps, int rows, int C,
                            void* y, long long ldy, float* mean, float* rstd, void* stream);
size_t rtdc_layernorm_bwd_scratch_bytes(int rows, int C);
int rtdc_layernorm_bwd_bf16(const void* dy, long long lddy, const void* x, long long ldx, const void* gamma, const float* mean,
                            const float* rstd, int rows, int C, void* dx, long long lddx, void* dgamma, void* dbeta,
                            void* scratch, size_t scratch_bytes, void* stream);
/* out[c] = sum over rows of x[r, c] for a bf16 matrix (cols and the row pitch multiples of 8): the bias gradient of a bf16
 * Linear layer. out is bf16 (out_fp32 == 0) or fp32; scratch: rtdc_colsum_bf16_scratch_bytes bytes, no initialisation. */
size_t rtdc_colsum_bf16_scratch_bytes(int rows, int cols);
int rtdc_colsum_bf16(const void* x, long long ldx, int rows, int cols, void* out, int out_fp32, void* scratch,
                     size_t scratch_bytes, void* stream);

/* out[n] = bias[n] (0 when NULL) + t * sum_k w[n*ldw + k], n < rows: the effective bias of a Linear layer whose input
 * normalisation x = s*raw + t has been folded into the GEMM (see rtdc_gemm_desc.alpha). One warp per row. */
int rtdc_fold_input_bias(const float* w, int ldw, int rows, int cols, const float* bias, float t, float* out, void* stream);

/* Bytes of scratch rtdc_gemm needs for this descriptor (0 when it will not split K): room for the
 * per-split partial tiles, which a second kernel sums in split order. Contents need no initialisation.
 * Calls that share a scratch buffer must not overlap on the GPU. */
size_t rtdc_gemm_scratch_bytes(const rtdc_gemm_desc* d);
int rtdc_gemm(const rtdc_gemm_desc* d, void* stream);
/* Tile/split plan the launcher would use: out[0]=bn, out[1]=stages, out[2]=splits, out[3]=grid.x, out[4]=smem bytes */
int rtdc_gemm_plan(const rtdc_gemm_desc* d, int* out5);

/* ------------------------------------------------------------------------------------------
 * Fused softmax-cross-entropy (K4/K5). logits [rows, ldz] (first `classes` columns valid),
 * labels int64 [rows]. Writes the mean loss (sum/rows, ref: nn.CrossEntropyLoss() default,
 * my_ray_module.py:141,156), optionally dlogits [rows, ldz] = (softmax - onehot) * grad_scale,
 * multiplied by (logit > 0) when relu_mask != 0 (the reference's trailing ReLU, :106), padding
 * columns zeroed; optionally counts argmax==label into correct (ref: :170); optionally writes the
 * column sums of dlogits into dbias_out[classes] (the bias gradient of the layer that produced the logits).
 * scratch: rtdc_softmax_ce_scratch_bytes(rows) bytes, zero before first use, left zero after.
 */
size_t rtdc_softmax_ce_scratch_bytes(int rows);
int rtdc_softmax_ce(const float* logits, int ldz, const int64_t* labels, int rows, int classes,
                    float* loss_out, float* dlogits, float grad_scale, int relu_mask,
                    unsigned long long* correct_accum, float* loss_sum_accum, float* dbias_out,
                    void* scratch, size_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused training "head": last Linear (K -> classes, classes = 10 or 16, K <= 512) + trailing ReLU +
 * softmax cross-entropy + that layer's whole backward, in one pass over the hidden activations h
 * [rows, ldh] (exact fp32 FFMA): writes loss (mean), dw [classes, K], db [classes] and
 * dh[rows, lddh] = (dlogits . w) * (h > 0) * mask_scale, with dlogits = (softmax - onehot) * grad_scale
 * masked by the logit ReLU. Replaces a skinny forward GEMM, the loss kernel, a skinny wgrad (+ split-K
 * reduce), the bias column sum and a skinny dgrad.
 * ref: my_ray_module.py:105-106 (Linear(512,10)+ReLU), :156 (loss), :159 (backward).
 * scratch: rtdc_mlp_head_scratch_bytes bytes (per-block partials; no initialisation needed). */
size_t rtdc_mlp_head_scratch_bytes(int rows, int K, int classes);
int rtdc_mlp_head_train(const float* h, int ldh, const float* w, const float* bias, const int64_t* labels, int rows,
                        int K, int classes, float grad_scale, float mask_scale, float* dh, int lddh, float* dw,
                        float* db, float* loss, float* db_below /* [K] column sums of dh, or NULL */,
                        void* scratch, size_t scratch_bytes, void* stream);
/* The same in two launches: part 1 = the pass over h (dh is final, dw/db/loss/db_below are still per-block partials in
 * scratch), part 2 = the reduction of those partials (same arguments, same scratch, ordered after part 1); part 0 = both.
 * A caller can run part 2 on a side stream next to the kernels that only need dh. */
int rtdc_mlp_head_train_part(const float* h, int ldh, const float* w, const float* bias, const int64_t* labels, int rows,
                             int K, int classes, float grad_scale, float mask_scale, float* dh, int lddh, float* dw,
                             float* db, float* loss, float* db_below, void* scratch, size_t scratch_bytes, void* stream, int part);

/* ------------------------------------------------------------------------------------------
 * Column sums of a [rows, cols] fp32 matrix (bias gradients, K6): out[c] = sum_r x[r*ld + c].
 * Deterministic two-stage reduction; scratch: rtdc_colsum_scratch_bytes(rows, cols).
 */
size_t rtdc_colsum_scratch_bytes(int rows, int cols);
int rtdc_colsum(const float* x, int ld, int rows, int cols, float* out, void* scratch, size_t scratch_bytes,
                void* stream);

/* Column sums without any device allocation, for callers that cannot call cudaMalloc (a server whose
 * memory arena already owns all free HBM, a latency-critical loop that must avoid the device-wide
 * synchronisation of cudaMalloc/cudaFree, an embedder without a caching allocator). Same result as
 * rtdc_colsum, but the scratch (partial sums + arrival counters) is a fixed array inside the library
 * image: RTDC_COLSUM_NOALLOC_SCRATCH_BYTES of device memory on every GPU that loads the library, held
 * until the process exits.
 *   - cols must be <= RTDC_COLSUM_NOALLOC_MAX_COLS; wider inputs are rejected with an error message
 *     (use rtdc_colsum with your own scratch for those). rows is unlimited.
 *   - The first call on each device looks the array's address up (cudaGetSymbolAddress, with that
 *     device current) and remembers it; make that first call before starting any stream capture.
 *   - The scratch used is the CURRENT device's: call it with the device that owns `stream` (and x / out) current.
 *   - No cudaMalloc / cudaFree / allocating library routine is called, on the first call or later.
 *   - All calls on a device share the one array. Each call clears the counters it depends on before
 *     launching, so a call never inherits state from the previous one; but two calls that can run on
 *     the GPU at the same time (different streams without an ordering between them) race on it and
 *     give wrong sums. Order them on one stream or with events.
 * This entry point has no counterpart in the reference (which ships no native code). */
#define RTDC_COLSUM_NOALLOC_MAX_COLS 1024
#define RTDC_COLSUM_NOALLOC_SCRATCH_BYTES (4096 + 64 * (RTDC_COLSUM_NOALLOC_MAX_COLS / 128) * 128 * 4)
int rtdc_colsum_noalloc(const float* x, int ld, int rows, int cols, float* out, void* stream);

/* uint8 pixels -> normalised fp32 ((x/255 - mean)/std; ref: my_ray_module.py:38 ToTensor+Normalize). */
int rtdc_u8_normalize(const uint8_t* src, float* dst, size_t n, float mean, float std, void* stream);
/* uint8 -> fp32 x*scale + shift (one fma). scale = 1, shift = 0 widens raw pixels exactly, the form the exact-operand
 * GEMM path (rtdc_gemm_desc.a_exact / b_exact) consumes. */
int rtdc_u8_affine(const uint8_t* src, float* dst, size_t n, float scale, float shift, void* stream);
/* Same with the grid capped at `blocks` thread blocks (0 = no cap): a conversion that runs on a side stream next to a
 * tensor-bound kernel is paced down so that it does not fight that kernel for issue slots and L2 bandwidth. */
int rtdc_u8_affine_paced(const uint8_t* src, float* dst, size_t n, float scale, float shift, int blocks, void* stream);

/* Input batches out of a device-resident uint8 dataset (K9, resident mode): out[r,:] = src[idx[r],:] for r < rows
 * (row_bytes a multiple of 16, pointers 16-byte aligned), y_out[r] = (int64) labels[idx[r]] when labels is given.
 * ref: what DataLoader(shuffle=True) + default_collate do on the host (my_ray_module.py:74), done at HBM speed. */
int rtdc_u8_gather_rows(const uint8_t* src, const int32_t* idx, int rows, int row_bytes, uint8_t* out,
                        const uint8_t* labels_or_null, int64_t* y_out_or_null, void* stream);
/* Input batches by the copy engines (K9, streaming mode): n pieces of a PINNED HOST buffer, src_host_base + src_offsets[i],
 * sizes[i] bytes each (host arrays), are copied back to back into dst_dev with one cudaMemcpyAsync per piece on `stream`. */
int rtdc_h2d_pieces(void* dst_dev, const void* src_host_base, const int64_t* src_offsets, const int64_t* sizes, int n, void* stream);

/* ReLU(+inverted dropout) backward with padding in one pass: dz[r,c] = dy[r,c] * (y[r,c] > 0 ? scale : 0) for c < cols (y = the
 * layer's output; y NULL copies dy), dz[r,c] = 0 for cols <= c < cols_out. Used for the last layer of a fused chain, whose
 * gradient does not come out of a dgrad epilogue: cols_out rounds the row pitch up to what TMA needs (16 bytes).
 * ref: autograd's threshold_backward + dropout backward behind my_ray_module.py:159 (K5/K6). */
int rtdc_act_backward(const float* dy, int lddy, const float* y_or_null, int ldy, int rows, int cols, int cols_out, float scale,
                      float* dz, int lddz, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused optimizer steps over a flat parameter arena (K7). hyper points to device memory so a
 * captured CUDA graph sees updated values.
 * SGD (ref: torch.optim.SGD(lr, momentum=0.9), my_ray_module.py:142,160):
 *   g' = g*grad_scale + wd*p ; m = mu*m + (1-damp)*g' (m = g' on the first step) ; p -= lr*(nesterov ? g'+mu*m : m)
 */
typedef struct rtdc_sgd_hyper { float lr, momentum, dampening, weight_decay, grad_scale; int nesterov; int first_step; int pad; } rtdc_sgd_hyper;
int rtdc_sgd_step(float* p, const float* g, float* m, size_t n, const rtdc_sgd_hyper* hyper_dev, void* stream);
/* Same, and the block that finishes last also does what rtdc_counter_advance does (first_step = 0, step_counter[0] += 1):
 * one launch less per training step. hyper_dev->pad is the arrival counter: keep it 0 between calls. */
int rtdc_sgd_step_advance(float* p, const float* g, float* m, size_t n, rtdc_sgd_hyper* hyper_dev, uint32_t* step_counter_or_null,
                          void* stream);
/* Same again, and that last block also stores *loss_dev into loss_log_host[step_counter[0] % loss_log_slots] before it bumps the
 * counter. loss_log_host is PINNED HOST memory (mapped into the device address space, e.g. cudaHostAlloc / a torch pinned
 * tensor; loss_log_slots a power of two): every step's loss crosses PCIe by itself, so a host can follow the training loss
 * without synchronising with the step stream or adding a copy to a captured graph. */
int rtdc_sgd_step_advance_log(float* p, const float* g, float* m, size_t n, rtdc_sgd_hyper* hyper_dev, uint32_t* step_counter_or_null,
                              const float* loss_dev_or_null, float* loss_log_host_or_null, uint32_t loss_log_slots, void* stream);
/* Adam / AdamW (decoupled != 0): step counter lives in hyper_dev->step and is advanced by rtdc_adam_advance. */
typedef struct rtdc_adam_hyper { float lr, beta1, beta2, eps, weight_decay, grad_scale; int decoupled; int step; } rtdc_adam_hyper;
int rtdc_adam_step(float* p, const float* g, float* m, float* v, size_t n, const rtdc_adam_hyper* hyper_dev, void* stream);
/* bf16 parameters: the update runs on the fp32 `master` copy (and fp32 m, v); p_bf16 is rewritten from the master every step,
 * g_bf16 is the bf16 gradient arena. Same hyper-parameters and step counter. */
int rtdc_adam_step_bf16(void* p_bf16, const void* g_bf16, float* master, float* m, float* v, size_t n,
                        const rtdc_adam_hyper* hyper_dev, void* stream);
int rtdc_adam_advance(rtdc_adam_hyper* hyper_dev, void* stream);
/* step_counter[0] += 1 on the stream (dropout step counter, first_step flag clearing). */
int rtdc_counter_advance(uint32_t* counter_dev, rtdc_sgd_hyper* sgd_hyper_or_null, void* stream);

/* ------------------------------------------------------------------------------------------
 * Symmetric memory + collectives over NVLink 5 / NVSwitch (C1-C5). One process per GPU; each rank
 * calls rtdc_symm_create with the same `bytes` and the same rendezvous prefix (a path prefix for
 * unix sockets, e.g. "/tmp/rtdc_<job>"); the call exchanges POSIX fd handles of cuMemCreate
 * allocations and maps every peer. Then, after any host-level barrier between the ranks, each calls
 * rtdc_symm_bind_multicast (NVLS; silently stays off where unsupported).
 * ref: replaces the NCCL process group that Ray Train / torch DDP set up implicitly
 * (my_ray_module.py:135,244-250).
 */
#define RTDC_SYMM_MAX_WORLD 8
#define RTDC_SYMM_MAX_BLOCKS 160
typedef struct rtdc_symm_info_t {
  int rank, world, device;
  size_t bytes;               /* usable bytes per rank */
  void* local_ptr;            /* this rank's buffer */
  void* multicast_ptr;        /* NVLS address covering all ranks' buffers, or NULL */
  void* peer_ptrs[RTDC_SYMM_MAX_WORLD]; /* every rank's buffer as mapped here */
  void** peer_table_dev;      /* device array of the same pointers */
  void** pad_table_dev;       /* device array of per-rank signal pads */
  unsigned int* epochs_dev;   /* per-block barrier epochs (local) */
} rtdc_symm_info_t;
int rtdc_symm_create(int rank, int world, int device, size_t bytes, const char* rendezvous_prefix,
                     int enable_multicast, double timeout_s, void** out_handle);
int rtdc_symm_bind_multicast(void* handle);
int rtdc_symm_info(void* handle, rtdc_symm_info_t* out);
int rtdc_symm_destroy(void* handle);

/* In-place all-reduce (sum, then * scale) of `count` elements at byte `offset` of every rank's
 * symmetric buffer. Every rank must issue the same sequence of collective calls on a handle.
 *   algo: 0 = choose by size, 1 = one-shot (out-of-place into `out_local`, which must not alias the
 *         symmetric buffer), 2 = two-shot (reduce-scatter + all-gather over P2P loads/stores),
 *         3 = NVLS (multimem.ld_reduce + multimem.st through the switch; needs multicast_ptr).
 *   dtype: 0 = fp32, 1 = bf16 (accumulated in fp32).
 * The sum is formed in rank order 0..world-1 by a single owner per element (algo 1: by everyone in
 * the same order), so all ranks end with bit-identical values.
 * ref: the DDP gradient bucket all-reduce with its 1/world pre-scale (C3; my_ray_module.py:159). */
int rtdc_symm_all_reduce(void* handle, size_t offset, size_t count, int dtype, float scale, int algo,
                         void* out_local, int max_blocks, void* stream);
/* The distributed optimizer step as ONE kernel (replaces rtdc_symm_all_reduce + rtdc_sgd_step_advance): gradient, parameter
 * and momentum arenas of `count` fp32 elements live at byte offsets g/p/m_offset of every rank's symmetric buffer. Rank r owns
 * the r-th of `world` equal slices (in 16-byte vectors): it sums that slice of all ranks' gradients (algo 3: multimem.ld_reduce
 * through the switch; algo 2: peer loads added in rank order), scales by 1/world, applies the SGD update of rtdc_sgd_hyper to
 * its slice of parameters and momentum, and writes the new parameters into EVERY rank's parameter arena (multimem.st / peer
 * stores). Per GPU that is count/world in and count/world out instead of an all-reduce's 2x, and 1/world of the optimizer
 * math. Momentum is only maintained by the owner of a slice: rtdc_symm_gather_slices(m_offset, count) completes a rank's copy
 * (checkpoints). Counter / first-step / loss-log handling as rtdc_sgd_step_advance_log. algo 0 = NVLS when bound and
 * world > 2, else P2P. Every rank must call it in the same order as its other collectives on this handle.
 * ref: DDP all-reduce (my_ray_module.py:159) + optimizer.step() (:160). */
int rtdc_symm_sgd_step(void* handle, size_t g_offset, size_t p_offset, size_t m_offset, size_t count, int algo,
                       rtdc_sgd_hyper* hyper_dev, uint32_t* step_counter_or_null, const float* loss_dev_or_null,
                       float* loss_log_host_or_null, uint32_t loss_log_slots, int max_blocks, void* stream);
/* Copy, for every other rank q, slice q of the `count`-element arena at `offset` from rank q's buffer into this rank's (no
 * barrier, not a collective: the caller knows the owners are idle, see above). */
int rtdc_symm_gather_slices(void* handle, size_t offset, size_t count, void* stream);
/* Copy `bytes` at `offset` of rank `root`'s symmetric buffer into dst_local on every rank (C2/C4). */
int rtdc_symm_broadcast(void* handle, size_t offset, size_t bytes, int root, void* dst_local, void* stream);
/* Device-side barrier across ranks on the stream (C5). */
int rtdc_symm_barrier(void* handle, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RTDC_B200_H */
