/* pipe_b200 native kernels — C ABI.
 *
 * One shared library (pipe_b200/_native/libpipe_b200.so, sm_100a only).  Every
 * entry point takes raw device pointers and a cudaStream_t, launches on that
 * stream and returns 0 on success, -1 on failure (pb2_last_error() says why).
 * Nothing here allocates device memory unless stated: scratch is passed in by
 * the caller.  The Python package binds these with ctypes (pipe_b200/ops/_lib.py);
 * other runtimes can link the library directly.
 *
 * Dropout everywhere is counter-based: a stateless generator (two SplitMix64
 * finalisers) keyed by `seed`, with the element's flat index in the output tensor
 * as counter and `rng_stream` separating call sites, so a backward kernel (or a
 * checkpoint recompute) given the same (seed, rng_stream) regenerates the
 * identical mask.
 */
#ifndef PIPE_B200_H_
#define PIPE_B200_H_

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ------------------------------------------------------------ */
const char* pb2_last_error(void);          /* message of the last failure on this thread */
long long pb2_launch_count(void);          /* kernels launched by this library so far */
int pb2_version(void);

/* ---- GEMM: D[M,N] = A[M,K] * B[N,K]^T, bf16 in, fp32 accumulate (tcgen05 + TMEM + TMA) ----
 * Reference counterpart: the cuBLAS calls behind nn.Linear / nn.MultiheadAttention
 * (main.py:45,148); SURVEY.md K2, K4, K6, K7. */
#define PB2_GEMM_COUNTER_BYTES (64u * 1024u)
#define PB2_GEMM_WORKSPACE_BYTES (96u * 1024u * 1024u + PB2_GEMM_COUNTER_BYTES)

typedef struct pb2_gemm_desc {
    const void* a;        /* bf16 [M, K], row stride lda (elements, multiple of 8): rides the UMMA M axis */
    const void* b;        /* bf16 [N, K], row stride ldb: rides the UMMA N axis */
    int M, N, K;
    long long lda, ldb;
    void* out;            /* bf16 or fp32 */
    long long ld_out;     /* row stride of the output tensor */
    int out_f32;          /* 1: fp32 output, 0: bf16 */
    int transpose_out;    /* 1: element (m, n) is stored at out[n * ld_out + m] (swap-AB) */
    const float* bias;    /* optional fp32 bias */
    int bias_along_m;     /* bias index is m (1) or n (0) */
    int relu;             /* apply max(0, x) after bias */
    float drop_p;         /* dropout probability applied after bias/ReLU (0 = off) */
    unsigned long long seed, rng_stream;
    const void* mask_src; /* optional bf16, output-shaped: out *= (mask_src > 0) * mask_scale  (ReLU+dropout backward) */
    float mask_scale;
    const void* addend;   /* optional bf16, output-shaped: out += addend (residual gradient) */
    int accumulate;       /* fp32 output only: out += result */
    float alpha;          /* scales the accumulator first (0 means 1) */
    float* sumsq;         /* optional device scalar: += sum of squares of the stored values (grad-norm fusion) */
    int block_n;          /* N tile, multiple of 16 in [16,256]; 0 = choose */
    int splits;           /* split-K factor; 0 = choose */
    int stages;           /* TMA ring depth; 0 = choose */
    int reduce_mode;      /* split-K reduction: 0 = choose, 1 = global scratch workspace, 2 = cluster / DSMEM */
    int a_static;         /* 1: A is a weight that the kernel launched just before this one on the stream does
                             not write, so its first tiles may be fetched while that kernel is still running
                             (programmatic dependent launch); 0: wait for the predecessor before any load */
    int persistent;       /* tile scheduling when there is no split-K: 0 = choose (persistent CTAs with a
                             double-buffered TMEM accumulator when there are more tiles than SMs),
                             1 = always persistent, 2 = one CTA per tile */
    int a_mn_major;       /* 1: `a` points to the TRANSPOSE, a bf16 [K, M] matrix with M contiguous and row stride lda — the
                             tensor core reads it as an MN-major operand.  This is how a dgrad GEMM consumes the forward
                             weight [out, in] as it is (M = in, K = out) instead of a transposed copy. */
} pb2_gemm_desc;

size_t pb2_gemm_workspace_bytes(void);
/* `workspace` (device, zero-initialised once, reusable across launches on one stream) holds the
 * split-K partial tiles and per-tile arrival counters; may be NULL when splits == 1. */
int pb2_gemm_bf16(const pb2_gemm_desc* desc, void* workspace, size_t workspace_bytes, cudaStream_t stream);

/* Same GEMM, for callers that cannot supply (or allocate) a workspace: hosts whose memory is owned by
 * someone else's pool, loops that must stay clear of cudaMalloc, plain C callers without an allocator.
 * No counterpart in the reference (it has no native code and no scratch to manage).
 * The split-K scratch is a fixed array inside the library image: PB2_GEMM_STATIC_SCRATCH_BYTES of
 * partial tiles plus the arrival counters, resident on every GPU that loads the library until the
 * process exits.  A problem whose partial tiles need more is refused with an error that names the
 * limit (choose fewer `splits`, or use pb2_gemm_bf16).  Inputs and outputs are still the caller's.
 * The first call on each device also resolves the array's device address (cudaGetSymbolAddress with
 * that device current): issue it at start-up, before any stream capture begins.  Every call uses the
 * same array and inherits its previous contents; the counters are cleared on `stream` before each
 * launch.  Calls that may overlap on the GPU race on it: serialise them.  `stream` must be a stream of the
 * device that is current when the call is made (the scratch address is looked up for the current device).
 * Allocates nothing. */
#define PB2_GEMM_STATIC_SCRATCH_BYTES (16u * 1024u * 1024u)
size_t pb2_gemm_noalloc_max_scratch_bytes(void);
int pb2_gemm_bf16_noalloc(const pb2_gemm_desc* desc, cudaStream_t stream);

/* ---- op chain: dependent skinny GEMMs and the LayerNorms between them as one persistent kernel ----
 * Every GEMM is described exactly as for pb2_gemm_bf16 (same fused epilogue; not available: accumulate, sumsq,
 * bias along n) and must see the same N <= 128 tokens; op i may read what any op j < i wrote.  The weights (`a`) of
 * ALL ops must not be written by the kernel launched before this one: they are fetched ahead of everything else.
 * `workspace` is the GEMM workspace; its last PB2_CHAIN_BARRIER_BYTES (inside the counter block, zero like the rest)
 * hold the grid-barrier counters, which the kernel leaves at zero.  pb2_chain_supported: 1 if the ops can run as a
 * chain, else run them one by one.  No counterpart in the reference (separate cuBLAS / ATen calls, main.py:148). */
#define PB2_CHAIN_MAX_OPS 8
#define PB2_CHAIN_BARRIER_BYTES 256u
typedef struct pb2_chain_ln {
    const void* x;
    const void* h;        /* may be NULL */
    const float* gamma;
    const float* beta;
    void* y;
    float* mean;
    float* rstd;
    int rows, E;
    float eps;
} pb2_chain_ln;
typedef struct pb2_chain_op {
    int kind;             /* 0: GEMM (`gemm`), 1: y = LayerNorm(x + h) (`ln`) */
    pb2_gemm_desc gemm;
    pb2_chain_ln ln;
} pb2_chain_op;
int pb2_chain_supported(const pb2_chain_op* ops, int n_ops);
int pb2_chain_run(const pb2_chain_op* ops, int n_ops, void* workspace, size_t workspace_bytes, cudaStream_t stream);

/* ---- residual add + LayerNorm: y = LN(x + h) (h may be NULL) ------------------------------
 * bf16 rows of width E (multiple of 8, <= 8192), fp32 gamma/beta/statistics.
 * Backward also returns dh = dres * dropout_mask / (1 - p): the gradient of the sub-layer output
 * *before* the dropout that the producing GEMM epilogue applied with the same (seed, rng_stream);
 * dgamma/dbeta are accumulated with fp32 atomics.
 * Reference counterpart: dropout + add + nn.LayerNorm inside nn.TransformerEncoderLayer
 * (main.py:148); SURVEY.md K5. */
int pb2_add_layernorm_fwd(const void* x, const void* h, const float* gamma, const float* beta, void* y,
                          float* mean, float* rstd, int rows, int E, float eps, cudaStream_t stream);
int pb2_add_layernorm_bwd(const void* dy, const void* x, const void* h, const float* gamma, const float* mean,
                          const float* rstd, void* dres, void* dh, float* dgamma, float* dbeta, int rows, int E,
                          float drop_p, unsigned long long seed, unsigned long long rng_stream,
                          cudaStream_t stream);

/* ---- embedding * sqrt(E) + positional table + dropout --------------------------------------
 * tokens int64 [N, S] (batch-first); out bf16 [S*N, E] sequence-first (row = s*N + n); weight and
 * pe fp32.  Backward scatter-adds into the fp32 weight gradient.
 * Reference counterpart: Encoder / PositionalEncoding (main.py:24-40,57-73); SURVEY.md K1. */
int pb2_embed_fwd(const long long* tokens, const float* weight, const float* pe, void* out, int N, int S, int E,
                  float scale, float drop_p, unsigned long long seed, unsigned long long rng_stream,
                  cudaStream_t stream);
int pb2_embed_bwd(const long long* tokens, const void* dout, float* dweight, int N, int S, int E, float scale,
                  float drop_p, unsigned long long seed, unsigned long long rng_stream, cudaStream_t stream);

/* Embedding gradient of a data-parallel hybrid, exchanged as operands instead of all-reduced as a dense V x E matrix.
 * pb2_embed_pack: row r of a micro-batch = dout[r] * scale * dropout mask (the mask of pb2_embed_fwd for the same seed /
 * stream) as fp32 into rows[r], its token id into toks[r], and the same values at the same addresses + peer_delta[q]
 * (this replica's region of the stash in every replica's memory).  pb2_embed_scatter: over R stash slots (toks < 0 =
 * unused), dweight[token] += alpha * (sum of all rows with that token, in slot order), one writer per token — the same
 * bits on every replica; reset_slots = 1 marks every slot unused again afterwards (on `stream`).  Reference counterpart: nn.Embedding's backward (main.py:28) under DDP (pipe.py:290-293). */
int pb2_embed_pack(const long long* tokens, const void* dout, float* rows, long long* toks, int N, int S, int E,
                   float scale, float drop_p, unsigned long long seed, unsigned long long rng_stream, int n_peers,
                   const long long* peer_delta, cudaStream_t stream);
int pb2_embed_scatter(long long* toks, const float* rows, float* dweight, int R, int E, float alpha,
                      int reset_slots, cudaStream_t stream);

/* ---- softmax cross-entropy over bf16 logits [rows, ld] with V valid columns ----------------
 * loss_rows[r] = logsumexp - logit[target] (loss_rows may be NULL); with write_grad the logits are
 * overwritten by (softmax - onehot) * row_scale[r] (row_scale may be NULL = 1; pad columns zeroed),
 * ready to be the dY of the vocabulary GEMMs.
 * Reference counterpart: nn.CrossEntropyLoss (main.py:182,217); SURVEY.md K9. */
int pb2_cross_entropy(void* logits, const long long* targets, float* loss_rows, const float* row_scale, int rows,
                      int V, long long ld, int write_grad, cudaStream_t stream);

/* ---- bf16 transpose dst[c, r] = src[r, c] with optional fused column sums -------------------
 * Produces the K-major operands of the weight-gradient GEMM; colsum[c] += scale * sum_r src[r, c]
 * is the bias gradient.  dst columns [R, ld_dst) are zero-filled. */
int pb2_transpose_colsum(const void* src, long long ld_src, void* dst, long long ld_dst, int R, int C,
                         float* colsum, float scale, cudaStream_t stream);

/* Several transposes in one launch (a layer's backward stacks all its wgrad operands at once).
 * Unlike the single form, dst columns beyond R are left untouched. */
#define PB2_MAX_TRANSPOSES 8
typedef struct pb2_transpose_job {
    const void* src;      /* bf16 [R, C], row stride ld_src */
    void* dst;            /* bf16 [C, >= R], row stride ld_dst */
    long long ld_src, ld_dst;
    int R, C;
    float* colsum;        /* optional fp32 [C]: += scale * column sums */
    float scale;
} pb2_transpose_job;
#define PB2_MAX_PEERS 7
typedef struct pb2_transpose_jobs {
    pb2_transpose_job job[PB2_MAX_TRANSPOSES];
    int count;
    /* Replicated destinations (Pipe-inside-DDP hybrid): every element stored to a job's `dst` is also stored at
     * the same address + peer_delta[q] bytes, q < n_peers — the same slot of the symmetric wgrad stash in each
     * data-parallel replica's memory, mapped into this process.  The gradient "all-reduce" of the weight matrices
     * then happens inside the wgrad GEMM: K runs over all replicas' tokens and the tensor core's fp32 accumulator
     * is the reduction (see pb2_group_barrier for the ordering).  n_peers = 0: local only. */
    int n_peers;
    long long peer_delta[PB2_MAX_PEERS];
} pb2_transpose_jobs;
int pb2_transpose_multi(const pb2_transpose_jobs* jobs, cudaStream_t stream);

/* ---- stage-to-stage mailbox over NVLink peer memory (multi-process pipeline) --------------------
 * A mailbox lives in memory that both ranks of a stage boundary have mapped (symmetric memory or CUDA
 * IPC).  Per message slot the receiver holds the payload area and PB2_MAILBOX_MAX_CTAS flag words, the
 * sender holds PB2_MAILBOX_MAX_CTAS acknowledgement words; all words start at zero and `seq` starts at 1
 * and grows by one per use of the slot.  Both sides must pass the same `bytes` (multiple of 16) and
 * `ctas`.  pb2_mailbox_send: wait until the previous message of the slot was drained, store `src` into
 * the peer's slot, publish flag = seq.  pb2_mailbox_recv: wait for flag == seq, copy the slot into
 * `dst`, acknowledge.  Both are single kernel launches on `stream` (asynchronous, stream-ordered with
 * the kernels that produce `src` / consume `dst`); a peer that never shows up traps the kernel after
 * PB2_MAILBOX_TIMEOUT_S seconds (default 20).  No counterpart in the reference (its stages share a
 * process and use cudaMemcpyPeerAsync); NCCL send/recv is the baseline. */
#define PB2_MAILBOX_MAX_CTAS 32
int pb2_mailbox_send(const void* src, size_t bytes, void* remote_slot, unsigned int* remote_flags,
                     const unsigned int* local_acks, unsigned int seq, int slot, int ctas, cudaStream_t stream);
int pb2_mailbox_recv(void* dst, size_t bytes, const void* local_slot, const unsigned int* local_flags,
                     unsigned int* remote_acks, unsigned int seq, int slot, int ctas, cudaStream_t stream);

/* ---- *out += sum(x^2)  (gradient-norm partial; x fp32, 16-byte aligned) ---------------------- */
int pb2_sumsq(const float* x, size_t n, float* out, cudaStream_t stream);

/* ---- attention core on packed QKV [T, 3E] (row = s*NB + n), out [T, E] ----------------------
 * softmax(QK^T/sqrt(D)) V with probability dropout and optional causal mask.  Single-tile kernels: S <= 128,
 * D <= 256 (the path for BASELINE's 25-token sequences); longer sequences use pb2_flash_attention_*.
 * Backward recomputes the probabilities and the dropout mask.
 * Reference counterpart: nn.MultiheadAttention core (main.py:148); SURVEY.md K3. */
/* 1 when the layer executor routes this shape to the flash kernels below (head_dim 64 / 128 and S >= 64): the
 * caller then provides attn_lse / attn_scratch in pb2_encoder_layer. */
int pb2_attention_uses_flash(int S, int D);
int pb2_attention_fwd(const void* qkv, void* out, int NB, int S, int H, int D, int causal, float drop_p,
                      unsigned long long seed, unsigned long long rng_stream, cudaStream_t stream);
int pb2_attention_bwd(const void* qkv, const void* dout, void* dqkv, int NB, int S, int H, int D, int causal,
                      float drop_p, unsigned long long seed, unsigned long long rng_stream, cudaStream_t stream);

/* ---- flash attention on tcgen05 / TMEM / TMA: any S, head_dim 64 / 128 / 256, same packed layout ------------
 * softmax(QK^T/sqrt(D)) V with probability dropout and optional causal mask, computed in blocks of keys with a
 * running maximum / sum (no S x S matrix anywhere).  `lse` (optional for inference, required by the backward)
 * receives log-sum-exp of the scaled scores per (sequence, head, query): fp32 [NB * H, S].
 * Reference counterpart: nn.MultiheadAttention core (main.py:148) at sequence lengths up to the positional
 * table's 5000 (main.py:59); SURVEY.md K3. */
int pb2_flash_attention_fwd(const void* qkv, void* out, float* lse, int NB, int S, int H, int D, int causal,
                            float drop_p, unsigned long long seed, unsigned long long rng_stream,
                            cudaStream_t stream);

/* Backward of pb2_flash_attention_fwd (head_dim 64 or 128): recomputes the probabilities block by block from
 * `lse`, keeps dK / dV of a key block in TMEM while it walks the query blocks, adds dQ into an fp32 accumulator.
 * `out` is the forward's output, `dout` its gradient ([T, E] bf16), `dqkv` receives the packed gradient [T, 3E].
 * `scratch` (pb2_flash_attention_bwd_scratch_bytes() bytes, 16-byte aligned, contents irrelevant) holds
 * delta = rowsum(dO o O) and the dQ accumulator; three small launches around the main kernel (memset, delta,
 * dQ cast) run on the same stream. */
size_t pb2_flash_attention_bwd_scratch_bytes(int NB, int S, int H, int D);
int pb2_flash_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                            void* scratch, size_t scratch_bytes, int NB, int S, int H, int D, int causal,
                            float drop_p, unsigned long long seed, unsigned long long rng_stream,
                            cudaStream_t stream);

/* ---- multi-tensor fused clip + Adam ---------------------------------------------------------
 * `table_dev` is a device array of n_tensors descriptors sorted by tile_start; a tensor of
 * [rows, cols] owns ceil(rows/32) * ceil(cols/64) consecutive tiles.  w/g/m/v are fp32; w16
 * (optional) is the bf16 shadow, w16t (optional) the transposed bf16 shadow with row stride ld_t.
 * normsq_dev holds the squared global gradient norm (device scalar); the clip coefficient
 * min(1, max_norm / (sqrt(normsq) * grad_scale + 1e-6)) is computed on the device.
 * Reference counterpart: clip_grad_norm_ + optim.Adam (main.py:184,219-220); SURVEY.md K10, K11. */
typedef struct pb2_adam_tensor {
    float* w;
    const float* g;
    float* m;
    float* v;
    void* w16;
    void* w16t;
    int rows, cols, ld_t, tile_start;
} pb2_adam_tensor;

int pb2_adam_step(const pb2_adam_tensor* table_dev, int n_tensors, int total_tiles, const float* normsq_dev,
                  float lr, float beta1, float beta2, float eps, float weight_decay, int step, float max_norm,
                  float grad_scale, cudaStream_t stream);
int pb2_refresh_shadows(const pb2_adam_tensor* table_dev, int n_tensors, int total_tiles, cudaStream_t stream);

/* ---- gradient all-reduce over NVLink / NVSwitch, fused with scale + squared norm + bf16 cast-out ----
 * In place over symmetric memory: peer_bufs[q] is rank q's fp32 buffer of n elements as mapped into
 * THIS process (n a multiple of 4 * world), peer_pads[q] rank q's signal pad of
 * pb2_allreduce_pad_bytes() bytes (zeroed once; `epoch` must be non-zero and increase by one per call
 * on every rank).  Every rank of the group launches the call at the same point of its stream; the
 * kernels synchronise among themselves with flags in the pads (a stuck peer trips a watchdog trap).
 * use_multimem = 1 reduces inside the NVSwitch through `multicast` (the multicast alias of the
 * buffers); otherwise a two-shot P2P pull/push.  On return (stream order) every rank's buffer holds
 * scale * sum, *sumsq (optional) has this rank's share of the squared norm added, and cast_out_bf16
 * (optional, local, n elements) holds a bf16 copy.
 * Reference counterpart: none in code (DDP is only mentioned, pipe.py:290-293); baseline = NCCL
 * all-reduce.  SURVEY.md K18. */
/* Barrier among the ranks of a replica group, as one tiny kernel on `stream`: everything this rank's earlier
 * kernels on the stream stored into peers' memory is visible to a peer's kernels that follow the peer's own
 * barrier call.  `peer_pads[q]` is rank q's signal pad (pb2_allreduce_pad_bytes(), zeroed once) as mapped here;
 * `epoch` is non-zero and grows by one per call on every rank; `phase` (0 or 1) selects one of two independent
 * flag sets so that two barriers can alternate.  A missing peer trips the watchdog trap.
 * No counterpart in the reference. */
int pb2_group_barrier(void* const* peer_pads, int world, int rank, unsigned int epoch, int phase, cudaStream_t stream);

int pb2_allreduce_max_ctas(void);
size_t pb2_allreduce_pad_bytes(void);
int pb2_allreduce_f32(float* const* peer_bufs, void* const* peer_pads, float* multicast, int world, int rank,
                      size_t n, float scale, float* sumsq, void* cast_out_bf16, unsigned int epoch, int ctas,
                      int use_multimem, cudaStream_t stream);

/* ---- whole encoder layer: one host call launches the layer's kernel sequence -----------------
 * The framework's host-side critical path is the per-kernel launch cost:
 * a 75-token micro-batch gives each kernel 5-20 us of GPU work.  These two entry points run the
 * forward (7 launches) and backward (8 launches) of a post-norm Transformer encoder layer from C++,
 * so the Python layer pays one call per layer per pass.  All buffers are the caller's.
 * Forward:  qkv = x Wqkv^T + b;  a = attn(qkv);  o = drop(a Wo^T + b);  x1 = LN1(x + o);
 *           h = drop(relu(x1 W1^T + b));  f = drop(h W2^T + b);  y = LN2(x1 + f)
 * Backward: the reverse, with dgrad GEMMs against the transposed weight shadows, and every wgrad
 *           operand transposed into the caller's stacked K-major stash at column `stash_off`
 *           (the wgrad GEMMs themselves run once per mini-batch, see pb2_gemm_bf16).
 * Reference counterpart: nn.TransformerEncoderLayer forward/backward (main.py:148); SURVEY.md K2-K6. */
typedef struct pb2_encoder_layer {
    int T, NB, S, E, heads, FF;      /* T = S * NB tokens (row = s * NB + n) */
    int causal;
    float drop_p, eps;
    unsigned long long seed;
    /* bf16 weight shadows [out, in] and — optional — transposed shadows [in, ld >= out]: with wt_* = NULL the dgrad GEMMs
     * read w_* as MN-major operands (no transposed copy exists) */
    const void *w_qkv, *w_o, *w_1, *w_2;
    const void *wt_qkv, *wt_o, *wt_1, *wt_2;
    long long ldt_qkv, ldt_o, ldt_1, ldt_2;
    /* fp32 biases and LayerNorm parameters */
    const float *b_qkv, *b_o, *b_1, *b_2, *g1, *be1, *g2, *be2;
    /* forward activations (bf16 unless noted) */
    const void* x;                   /* [T, E] */
    void *qkv, *a, *o, *x1, *hmid, *f, *y;   /* [T,3E] [T,E] [T,E] [T,E] [T,FF] [T,E] [T,E] */
    float *mean1, *rstd1, *mean2, *rstd2;    /* [T] each */
    /* backward */
    const void* dy;                  /* [T, E] */
    void* dx;                        /* [T, E] */
    void *dr2, *df, *dz1, *dx1, *dr1, *do_, *da, *dqkv;   /* scratch: [T,E] [T,E] [T,FF] [T,E] [T,E] [T,E] [T,E] [T,3E] */
    float *dg1, *dbe1, *dg2, *dbe2;  /* LayerNorm gradients (accumulated) */
    float *db_qkv, *db_o, *db_1, *db_2;       /* bias gradients (accumulated) */
    /* stacked wgrad operands: x^T [in, ld] and dy^T [out, ld] per linear, written at column stash_off */
    void *xt_qkv, *dyt_qkv, *xt_o, *dyt_o, *xt_1, *dyt_1, *xt_2, *dyt_2;
    long long ld_stash_qkv, ld_stash_o, ld_stash_1, ld_stash_2;
    int stash_off_qkv, stash_off_o, stash_off_1, stash_off_2;
    /* split-K workspace of pb2_gemm_workspace_bytes() bytes (see pb2_gemm_bf16) */
    void* workspace;
    size_t workspace_bytes;
    /* unused (kept for layout compatibility; a measured-slower L2 prefetch experiment lived here) */
    const void* next_weight;
    size_t next_weight_bytes;
    int prefetch;
    /* optional (backward): a second stream for the launch that stacks x^T / dy^T for the deferred weight
     * gradients.  Nothing in backward reads those stacks, so the launch leaves the critical path: it waits
     * (event) for this call's attention backward and then runs next to the remaining kernels.  The caller
     * keeps x, the activation buffer and the scratch buffer alive until that stream has passed this point,
     * and orders whatever consumes the stacks / bias gradients after it (pb2_stream_wait_stream). */
    void* side_stream;
    /* optional (backward, hybrid): replicate the stacked wgrad operands into the data-parallel peers' stashes
     * (same meaning as in pb2_transpose_jobs) */
    int n_peers;
    long long peer_delta[PB2_MAX_PEERS];
    /* needed when pb2_attention_uses_flash(S, E / heads): log-sum-exp of the attention rows, fp32 [NB * heads * S]
     * (written by the forward, read by the backward), and for the backward a scratch buffer of
     * pb2_flash_attention_bwd_scratch_bytes() bytes */
    float* attn_lse;
    void* attn_scratch;
    size_t attn_scratch_bytes;
} pb2_encoder_layer;

int pb2_encoder_layer_fwd(const pb2_encoder_layer* L, cudaStream_t stream);
int pb2_encoder_layer_bwd(const pb2_encoder_layer* L, cudaStream_t stream);

/* ---- a run of consecutive encoder layers (the layers of one pipeline stage): one host call per pass ----
 * `layers` is a HOST array of n descriptors in which the per-layer constants are filled in: E, heads, FF,
 * causal, eps, weight shadows and strides, biases, LayerNorm parameters, the gradient pointers
 * (dg*, dbe*, db_*) and the wgrad stash bases / strides.  Everything per call comes from this struct: the
 * run's input `x`, one slab `y` of n x [T, E] layer outputs (y[l] feeds layer l + 1; y[n-1] is the result),
 * one slab `act` of pb2_encoder_stack_act_elems() bf16 elements for the saved activations, `stats` of
 * n x 4T floats; for the backward additionally `dy`, `dx`, a scratch slab `tmp` of
 * pb2_encoder_stack_tmp_elems() bf16 elements (kept alive until the side stream has passed it) and the
 * column `stash_off` at which every linear of the run stacks this micro-batch's wgrad operands.
 * All layers share E.  Same kernels, same order, same results as n calls of pb2_encoder_layer_fwd / _bwd.
 * Reference counterpart: one stage's nn.Sequential of nn.TransformerEncoderLayer as run by a worker
 * thread (pipeline.py:225-230, main.py:147-153); SURVEY.md §3.2, §7.4 item 2. */
#define PB2_MAX_STACK_LAYERS 64
typedef struct pb2_encoder_stack {
    int n_layers;
    const pb2_encoder_layer* layers;
    int T, NB, S;
    float drop_p;
    const unsigned long long* seeds;   /* n seeds (host), or NULL when drop_p == 0 */
    const void* x;
    void* y;
    void* act;
    float* stats;
    const void* dy;
    void* dx;
    void* tmp;
    int stash_off;
    void* workspace;
    size_t workspace_bytes;
    void* side_stream;
    /* flash-attention shapes only (pb2_attention_uses_flash): `stats` then holds n x (4T + heads * T) floats — each
     * layer's LayerNorm statistics followed by its attention log-sum-exp — and the backward needs one scratch buffer
     * of pb2_flash_attention_bwd_scratch_bytes() bytes, shared by the layers of the run */
    void* attn_scratch;
    size_t attn_scratch_bytes;
} pb2_encoder_stack;

size_t pb2_encoder_stack_act_elems(const pb2_encoder_layer* layers, int n, int T);
size_t pb2_encoder_stack_tmp_elems(const pb2_encoder_layer* layers, int n, int T);
size_t pb2_encoder_stack_stats_floats(const pb2_encoder_layer* layers, int n, int S, int T);
int pb2_encoder_stack_fwd(const pb2_encoder_stack* s, cudaStream_t stream);
int pb2_encoder_stack_bwd(const pb2_encoder_stack* s, cudaStream_t stream);

/* cudaMemset2DAsync on `stream`: `height` rows of `width` bytes, `pitch` bytes apart, set to `value`. */
int pb2_memset2d(void* ptr, size_t pitch, int value, size_t width, size_t height, cudaStream_t stream);

/* `waiter` waits (on the device) for everything enqueued on `waited` so far; both on the current device. */
int pb2_stream_wait_stream(cudaStream_t waiter, cudaStream_t waited);

#ifdef __cplusplus
}
#endif

#endif /* PIPE_B200_H_ */
