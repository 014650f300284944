// Host-side executor of one Transformer encoder layer: the launch sequence of the forward and
// of the backward, issued from C++ so the Python layer pays one call per layer per pass instead of
// one per kernel.  Pure orchestration: every kernel it launches is a public entry point of this
// library (pb2_gemm_bf16, pb2_attention_*, pb2_add_layernorm_*, pb2_transpose_multi).
//
// Reference counterpart: the Python-level op sequence of nn.TransformerEncoderLayer
// (main.py:148) as driven by the autograd engine; SURVEY.md §3.2 "GPU hot path", §7.4 item 2.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "pipe_b200.h"

namespace {

struct Lin {
    const void* w;        // A operand: [M rows, K], row stride lda — or, with a_mn, its transpose [K, M] (row stride lda)
    long long lda;
    int M, K;
    int a_mn = 0;
};

// dgrad through a Linear with forward weight W [out, in]: dX[T, in] = dY[T, out] W.  With a transposed shadow (wt) that is
// the usual K-major GEMM; without one the forward shadow itself is read as an MN-major operand (M = in, K = out).
Lin dgrad_weight(const void* wt, long long ldt, const void* w, int in_f, int out_f) {
    if (wt) return Lin{wt, ldt, in_f, out_f, 0};
    return Lin{w, in_f, in_f, out_f, 1};
}

pb2_gemm_desc linear_desc(const pb2_encoder_layer* L, const Lin& w, const void* act, int T, void* out, const float* bias,
                          int relu, float drop_p, unsigned long long rng_stream, const void* mask_src, float mask_scale,
                          const void* addend) {
    // out[T, M] = act[T, K] * w[M, K]^T  (weights on the UMMA M axis, transposed store)
    pb2_gemm_desc d;
    memset(&d, 0, sizeof(d));
    d.a = w.w; d.lda = w.lda; d.M = w.M; d.K = w.K;
    d.b = act; d.ldb = w.K; d.N = T;
    d.out = out; d.ld_out = w.M; d.transpose_out = 1;
    d.bias = bias; d.bias_along_m = 1; d.relu = relu;
    d.drop_p = drop_p; d.seed = L->seed; d.rng_stream = rng_stream;
    d.mask_src = mask_src; d.mask_scale = mask_scale; d.addend = addend;
    d.alpha = 1.f;
    d.a_mn_major = w.a_mn;
    d.a_static = 1;       // weight shadows are only written by the optimizer kernel, never by a layer kernel
    return d;
}

int linear(const pb2_encoder_layer* L, const Lin& w, const void* act, int T, void* out, const float* bias, int relu,
           float drop_p, unsigned long long rng_stream, const void* mask_src, float mask_scale, const void* addend,
           cudaStream_t stream) {
    const pb2_gemm_desc d = linear_desc(L, w, act, T, out, bias, relu, drop_p, rng_stream, mask_src, mask_scale, addend);
    return pb2_gemm_bf16(&d, L->workspace, L->workspace_bytes, stream);
}

int attention_fwd(const pb2_encoder_layer* L, cudaStream_t stream) {
    const int E = L->E;
    const float p = L->drop_p;
    if (pb2_attention_uses_flash(L->S, E / L->heads)) {
        if (!L->attn_lse) { pb2::set_error("pb2_encoder_layer_fwd: this shape uses flash attention: attn_lse is required"); return -1; }
        return pb2_flash_attention_fwd(L->qkv, L->a, L->attn_lse, L->NB, L->S, L->heads, E / L->heads, L->causal, p, L->seed, 0, stream);
    }
    return pb2_attention_fwd(L->qkv, L->a, L->NB, L->S, L->heads, E / L->heads, L->causal, p, L->seed, 0, stream);
}

pb2_chain_op gemm_op(const pb2_gemm_desc& d) {
    pb2_chain_op op;
    memset(&op, 0, sizeof(op));
    op.kind = 0;
    op.gemm = d;
    return op;
}
pb2_chain_op ln_op(const void* x, const void* h, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                   int rows, int E, float eps) {
    pb2_chain_op op;
    memset(&op, 0, sizeof(op));
    op.kind = 1;
    op.ln.x = x; op.ln.h = h; op.ln.gamma = gamma; op.ln.beta = beta; op.ln.y = y; op.ln.mean = mean; op.ln.rstd = rstd;
    op.ln.rows = rows; op.ln.E = E; op.ln.eps = eps;
    return op;
}

// Everything of a layer's forward after the attention core — out-projection, add + LayerNorm, the two feed-forward GEMMs,
// add + LayerNorm — and, when `next` is given, the QKV projection of the following layer, as one persistent kernel.
// Returns the number of ops, 0 if these shapes cannot run as a chain.
int post_attention_chain(const pb2_encoder_layer* L, const pb2_encoder_layer* next, pb2_chain_op* ops) {
    const int T = L->T, E = L->E, FF = L->FF;
    const float p = L->drop_p;
    int n = 0;
    ops[n++] = gemm_op(linear_desc(L, Lin{L->w_o, E, E, E}, L->a, T, L->o, L->b_o, 0, p, 1, nullptr, 0.f, nullptr));
    ops[n++] = ln_op(L->x, L->o, L->g1, L->be1, L->x1, L->mean1, L->rstd1, T, E, L->eps);
    ops[n++] = gemm_op(linear_desc(L, Lin{L->w_1, E, FF, E}, L->x1, T, L->hmid, L->b_1, 1, p, 2, nullptr, 0.f, nullptr));
    ops[n++] = gemm_op(linear_desc(L, Lin{L->w_2, FF, E, FF}, L->hmid, T, L->f, L->b_2, 0, p, 3, nullptr, 0.f, nullptr));
    ops[n++] = ln_op(L->x1, L->f, L->g2, L->be2, L->y, L->mean2, L->rstd2, T, E, L->eps);
    if (next)
        ops[n++] = gemm_op(linear_desc(next, Lin{next->w_qkv, next->E, 3 * next->E, next->E}, next->x, T, next->qkv, next->b_qkv, 0,
                                       0.f, 0, nullptr, 0.f, nullptr));
    return pb2_chain_supported(ops, n) ? n : 0;
}

// PB2_CHAIN=1 runs everything after the attention core (and the next layer's QKV projection) as one persistent op-chain
// kernel.  Off by default: measured on config 3 (75 tokens, E = 4096) the chain takes 71 us where the same six ops as separate
// programmatically-launched kernels take about as long, and a stack forward came out at 95 us per layer against 86 us
// (profiles/chain_kernel_r2.txt) — each op is bound by the same serial phases either way (operand load after the producer is
// complete, a DRAM-bound main loop, split-K exchange, epilogue) and a grid barrier costs what a dependent launch costs.
// Read on every call: tests compare the two paths in one process.
bool chain_enabled() {
    const char* e = getenv("PB2_CHAIN");
    return e && e[0] == '1';
}

}  // namespace

extern "C" int pb2_encoder_layer_fwd(const pb2_encoder_layer* L, cudaStream_t stream) {
    if (!L) { pb2::set_error("pb2_encoder_layer_fwd: null descriptor"); return -1; }
    const int T = L->T, E = L->E, FF = L->FF;
    const float p = L->drop_p;
    if (T != L->S * L->NB || E % L->heads) { pb2::set_error("pb2_encoder_layer_fwd: inconsistent shape"); return -1; }

    if (linear(L, Lin{L->w_qkv, E, 3 * E, E}, L->x, T, L->qkv, L->b_qkv, 0, 0.f, 0, nullptr, 0.f, nullptr, stream)) return -1;
    if (attention_fwd(L, stream)) return -1;
    if (linear(L, Lin{L->w_o, E, E, E}, L->a, T, L->o, L->b_o, 0, p, 1, nullptr, 0.f, nullptr, stream)) return -1;
    if (pb2_add_layernorm_fwd(L->x, L->o, L->g1, L->be1, L->x1, L->mean1, L->rstd1, T, E, L->eps, stream)) return -1;
    if (linear(L, Lin{L->w_1, E, FF, E}, L->x1, T, L->hmid, L->b_1, 1, p, 2, nullptr, 0.f, nullptr, stream)) return -1;
    if (linear(L, Lin{L->w_2, FF, E, FF}, L->hmid, T, L->f, L->b_2, 0, p, 3, nullptr, 0.f, nullptr, stream)) return -1;
    if (pb2_add_layernorm_fwd(L->x1, L->f, L->g2, L->be2, L->y, L->mean2, L->rstd2, T, E, L->eps, stream)) return -1;
    return 0;
}

extern "C" int pb2_encoder_layer_bwd(const pb2_encoder_layer* L, cudaStream_t stream) {
    if (!L) { pb2::set_error("pb2_encoder_layer_bwd: null descriptor"); return -1; }
    const int T = L->T, E = L->E, FF = L->FF;
    const float p = L->drop_p;
    const float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;

    // y = LN2(x1 + drop(f))
    if (pb2_add_layernorm_bwd(L->dy, L->x1, L->f, L->g2, L->mean2, L->rstd2, L->dr2, L->df, L->dg2, L->dbe2, T, E, p,
                              L->seed, 3, stream)) return -1;
    // through linear2, then ReLU + dropout (mask recovered from the saved activation)
    if (linear(L, dgrad_weight(L->wt_2, L->ldt_2, L->w_2, FF, E), L->df, T, L->dz1, nullptr, 0, 0.f, 0, L->hmid, inv_keep, nullptr, stream)) return -1;
    // through linear1, plus the residual branch
    if (linear(L, dgrad_weight(L->wt_1, L->ldt_1, L->w_1, E, FF), L->dz1, T, L->dx1, nullptr, 0, 0.f, 0, nullptr, 0.f, L->dr2, stream)) return -1;
    // x1 = LN1(x + drop(o))
    if (pb2_add_layernorm_bwd(L->dx1, L->x, L->o, L->g1, L->mean1, L->rstd1, L->dr1, L->do_, L->dg1, L->dbe1, T, E, p,
                              L->seed, 1, stream)) return -1;
    if (linear(L, dgrad_weight(L->wt_o, L->ldt_o, L->w_o, E, E), L->do_, T, L->da, nullptr, 0, 0.f, 0, nullptr, 0.f, nullptr, stream)) return -1;
    if (pb2_attention_uses_flash(L->S, E / L->heads)) {
        if (!L->attn_lse || !L->attn_scratch) { pb2::set_error("pb2_encoder_layer_bwd: this shape uses flash attention: attn_lse / attn_scratch are required"); return -1; }
        if (pb2_flash_attention_bwd(L->qkv, L->a, L->da, L->attn_lse, L->dqkv, L->attn_scratch, L->attn_scratch_bytes, L->NB, L->S,
                                    L->heads, E / L->heads, L->causal, p, L->seed, 0, stream)) return -1;
    } else if (pb2_attention_bwd(L->qkv, L->da, L->dqkv, L->NB, L->S, L->heads, E / L->heads, L->causal, p, L->seed, 0, stream)) {
        return -1;
    }
    // every source of the wgrad-operand transposes exists once the attention backward is done: from here
    // on they can run on the side stream, next to the QKV dgrad and the next layer's kernels
    cudaStream_t tstream = stream;
    if (L->side_stream) {
        tstream = static_cast<cudaStream_t>(L->side_stream);
        if (pb2::record_and_wait(stream, tstream)) return -1;
    }
    if (linear(L, dgrad_weight(L->wt_qkv, L->ldt_qkv, L->w_qkv, E, 3 * E), L->dqkv, T, L->dx, nullptr, 0, 0.f, 0, nullptr, 0.f, L->dr1, stream)) return -1;

    // stack every wgrad operand (and column-sum the bias gradients) in one launch
    pb2_transpose_jobs jobs;
    memset(&jobs, 0, sizeof(jobs));
    auto add = [&](const void* src, int C, void* dst_base, long long ld_dst, int off, float* colsum) {
        pb2_transpose_job& j = jobs.job[jobs.count++];
        j.src = src; j.ld_src = C; j.R = T; j.C = C;
        j.dst = static_cast<char*>(dst_base) + static_cast<size_t>(off) * 2; j.ld_dst = ld_dst;
        j.colsum = colsum; j.scale = 1.f;
    };
    add(L->x, E, L->xt_qkv, L->ld_stash_qkv, L->stash_off_qkv, nullptr);
    add(L->dqkv, 3 * E, L->dyt_qkv, L->ld_stash_qkv, L->stash_off_qkv, L->db_qkv);
    add(L->a, E, L->xt_o, L->ld_stash_o, L->stash_off_o, nullptr);
    add(L->do_, E, L->dyt_o, L->ld_stash_o, L->stash_off_o, L->db_o);
    add(L->x1, E, L->xt_1, L->ld_stash_1, L->stash_off_1, nullptr);
    add(L->dz1, FF, L->dyt_1, L->ld_stash_1, L->stash_off_1, L->db_1);
    add(L->hmid, FF, L->xt_2, L->ld_stash_2, L->stash_off_2, nullptr);
    add(L->df, E, L->dyt_2, L->ld_stash_2, L->stash_off_2, L->db_2);
    jobs.n_peers = L->n_peers;
    for (int q = 0; q < L->n_peers && q < PB2_MAX_PEERS; ++q) jobs.peer_delta[q] = L->peer_delta[q];
    return pb2_transpose_multi(&jobs, tstream);
}

// ---------------------------------------------------------------------------------------
// A run of consecutive encoder layers of one pipeline stage as ONE host call per pass.
//
// With 75-token micro-batches a layer is ~80 us of GPU work forward, ~140 us backward, and the
// Python side of a per-layer call (buffer allocation, descriptor fill, autograd bookkeeping) costs
// about as much: a rank that owns two to sixteen layers was launch-bound, which stretched every
// pipeline tick.  Here the caller passes the per-layer constants once (an array of
// pb2_encoder_layer with weights, biases, norms, gradient and stash pointers filled in) plus one
// slab per kind of buffer for the whole run; the per-layer pointers are carved out here.
namespace {

inline size_t act_elems(const pb2_encoder_layer& L, int T) { return static_cast<size_t>(T) * (7 * static_cast<size_t>(L.E) + L.FF); }
inline size_t tmp_elems(const pb2_encoder_layer& L, int T) { return static_cast<size_t>(T) * (10 * static_cast<size_t>(L.E) + L.FF); }

// forward activations of one layer inside the slab: qkv | a | o | x1 | hmid | f
void carve_act(pb2_encoder_layer& d, char* base, int T) {
    const size_t E = d.E, FF = d.FF, t = T;
    d.qkv = base;               base += t * 3 * E * 2;
    d.a = base;                 base += t * E * 2;
    d.o = base;                 base += t * E * 2;
    d.x1 = base;                base += t * E * 2;
    d.hmid = base;              base += t * FF * 2;
    d.f = base;
}

// floats of `stats` per layer: LayerNorm statistics, plus the attention log-sum-exp on flash shapes
inline size_t stats_stride(const pb2_encoder_layer& L, int S, int T) {
    return static_cast<size_t>(4) * T + (pb2_attention_uses_flash(S, L.E / L.heads) ? static_cast<size_t>(L.heads) * T : 0);
}

int check_stack(const pb2_encoder_stack* s, const char* who) {
    if (!s || !s->layers || s->n_layers <= 0) { pb2::set_error("%s: empty stack", who); return -1; }
    if (s->T != s->S * s->NB) { pb2::set_error("%s: T != S * NB", who); return -1; }
    for (int l = 1; l < s->n_layers; ++l)
        if (s->layers[l].E != s->layers[0].E) { pb2::set_error("%s: layers of a stack must share the model width", who); return -1; }
    return 0;
}

}  // namespace

extern "C" size_t pb2_encoder_stack_act_elems(const pb2_encoder_layer* layers, int n, int T) {
    size_t total = 0;
    for (int l = 0; l < n; ++l) total += act_elems(layers[l], T);
    return total;
}

extern "C" size_t pb2_encoder_stack_stats_floats(const pb2_encoder_layer* layers, int n, int S, int T) {
    return n > 0 ? static_cast<size_t>(n) * stats_stride(layers[0], S, T) : 0;
}

extern "C" size_t pb2_encoder_stack_tmp_elems(const pb2_encoder_layer* layers, int n, int T) {
    size_t total = 0;
    for (int l = 0; l < n; ++l) total += tmp_elems(layers[l], T);
    return total;
}

extern "C" int pb2_encoder_stack_fwd(const pb2_encoder_stack* s, cudaStream_t stream) {
    if (check_stack(s, "pb2_encoder_stack_fwd")) return -1;
    const int T = s->T, n = s->n_layers;
    const size_t E = s->layers[0].E;
    if (n > PB2_MAX_STACK_LAYERS) { pb2::set_error("pb2_encoder_stack_fwd: more than %d layers", PB2_MAX_STACK_LAYERS); return -1; }
    char* act = static_cast<char*>(s->act);
    char* y = static_cast<char*>(s->y);
    // every layer's descriptor first: a chain reaches into the next layer (its QKV projection)
    static thread_local pb2_encoder_layer d[PB2_MAX_STACK_LAYERS];
    for (int l = 0; l < n; ++l) {
        d[l] = s->layers[l];
        d[l].T = T; d[l].NB = s->NB; d[l].S = s->S;
        d[l].drop_p = s->drop_p;
        d[l].seed = s->seeds ? s->seeds[l] : 0ull;
        d[l].x = l == 0 ? s->x : y + static_cast<size_t>(l - 1) * T * E * 2;
        d[l].y = y + static_cast<size_t>(l) * T * E * 2;
        carve_act(d[l], act, T);
        act += act_elems(d[l], T) * 2;
        float* st = s->stats + static_cast<size_t>(l) * stats_stride(s->layers[0], s->S, T);
        d[l].mean1 = st; d[l].rstd1 = st + T; d[l].mean2 = st + 2 * T; d[l].rstd2 = st + 3 * T;
        d[l].attn_lse = st + 4 * T;
        d[l].workspace = s->workspace; d[l].workspace_bytes = s->workspace_bytes;
    }
    pb2_chain_op ops[PB2_CHAIN_MAX_OPS];
    bool have_qkv = false;          // layer l's QKV projection was part of layer l - 1's chain
    for (int l = 0; l < n; ++l) {
        const pb2_encoder_layer* L = &d[l];
        const int n_ops = chain_enabled() ? post_attention_chain(L, l + 1 < n ? &d[l + 1] : nullptr, ops) : 0;
        if (n_ops == 0) {
            if (have_qkv) { pb2::set_error("pb2_encoder_stack_fwd: chain support changed inside a stack"); return -1; }
            if (pb2_encoder_layer_fwd(L, stream)) return -1;
            continue;
        }
        if (!have_qkv &&
            linear(L, Lin{L->w_qkv, static_cast<long long>(E), 3 * L->E, L->E}, L->x, T, L->qkv, L->b_qkv, 0, 0.f, 0, nullptr, 0.f,
                   nullptr, stream))
            return -1;
        if (attention_fwd(L, stream)) return -1;
        if (pb2_chain_run(ops, n_ops, s->workspace, s->workspace_bytes, stream)) return -1;
        have_qkv = l + 1 < n;
    }
    return 0;
}

extern "C" int pb2_encoder_stack_bwd(const pb2_encoder_stack* s, cudaStream_t stream) {
    if (check_stack(s, "pb2_encoder_stack_bwd")) return -1;
    const int T = s->T, n = s->n_layers;
    const size_t E = s->layers[0].E;
    char* y = static_cast<char*>(s->y);
    // slab offsets of every layer (the walk is last layer first)
    size_t act_off[PB2_MAX_STACK_LAYERS], tmp_off[PB2_MAX_STACK_LAYERS];
    if (n > PB2_MAX_STACK_LAYERS) { pb2::set_error("pb2_encoder_stack_bwd: more than %d layers", PB2_MAX_STACK_LAYERS); return -1; }
    size_t a = 0, t = 0;
    for (int l = 0; l < n; ++l) {
        act_off[l] = a; tmp_off[l] = t;
        a += act_elems(s->layers[l], T) * 2;
        t += tmp_elems(s->layers[l], T) * 2;
    }
    for (int l = n - 1; l >= 0; --l) {
        pb2_encoder_layer d = s->layers[l];
        d.T = T; d.NB = s->NB; d.S = s->S;
        d.drop_p = s->drop_p;
        d.seed = s->seeds ? s->seeds[l] : 0ull;
        d.x = l == 0 ? s->x : y + static_cast<size_t>(l - 1) * T * E * 2;
        d.y = y + static_cast<size_t>(l) * T * E * 2;
        carve_act(d, static_cast<char*>(s->act) + act_off[l], T);
        float* st = s->stats + static_cast<size_t>(l) * stats_stride(s->layers[0], s->S, T);
        d.mean1 = st; d.rstd1 = st + T; d.mean2 = st + 2 * T; d.rstd2 = st + 3 * T;
        d.attn_lse = st + 4 * T;
        d.attn_scratch = s->attn_scratch; d.attn_scratch_bytes = s->attn_scratch_bytes;
        // scratch of this layer: dr2 | df | dz1 | dx1 | dr1 | do | da | dqkv | dx
        char* b = static_cast<char*>(s->tmp) + tmp_off[l];
        const size_t FF = d.FF, tt = T;
        d.dr2 = b;  b += tt * E * 2;
        d.df = b;   b += tt * E * 2;
        d.dz1 = b;  b += tt * FF * 2;
        d.dx1 = b;  b += tt * E * 2;
        d.dr1 = b;  b += tt * E * 2;
        d.do_ = b;  b += tt * E * 2;
        d.da = b;   b += tt * E * 2;
        d.dqkv = b; b += tt * 3 * E * 2;
        void* dx_here = l == 0 ? s->dx : static_cast<void*>(b);
        // the gradient entering layer l is what layer l + 1 wrote into its own dx slot
        if (l == n - 1) {
            d.dy = s->dy;
        } else {
            const pb2_encoder_layer& up = s->layers[l + 1];
            d.dy = static_cast<char*>(s->tmp) + tmp_off[l + 1] + tt * (9 * E + up.FF) * 2;
        }
        d.dx = dx_here;
        d.stash_off_qkv = d.stash_off_o = d.stash_off_1 = d.stash_off_2 = s->stash_off;
        d.workspace = s->workspace; d.workspace_bytes = s->workspace_bytes;
        d.side_stream = s->side_stream;
        if (pb2_encoder_layer_bwd(&d, stream)) return -1;
    }
    return 0;
}
