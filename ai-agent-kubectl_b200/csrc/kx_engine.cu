// Generation engine: orchestrates the kernels of kx_gemm.cu / kx_ops.cu for a
// decoder-only transformer with a paged KV cache.  What it replaces in the reference
// is the remote call behind `llm` (app.py:117, awaited at app.py:183-186); the model
// math is defined by BASELINE.json's north_star, not by reference code.
//
// One engine = one model replica on one GPU.  Prefill runs over whole sequences in
// chunks of at most `prefill_chunk` tokens; every decode step is the same sequence of
// launches, so after the first (eager) step the rest are replayed as one CUDA graph
// with the positions and the step counter advanced on the device.
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>
#include "kx_common.cuh"
#include "kx_gemm.cuh"
#include "kx_ops.cuh"

using namespace kx;

struct LayerPlan {
    kx_layer_weights w;
    // weight maps in both roles: A operand (decode, 128-row box) and B operand (prefill, 128-row box)
    CUtensorMap wqkv, wo, wgu, wd;
    // half-tile boxes for the CTA-pair kernel: [0] = 64 rows (128-wide tiles), [1] = 128 rows (256-wide tiles)
    CUtensorMap wqkv_h[2], wo_h[2], wgu_h[2], wd_h[2];
    void *kv_pages;
};

struct kx_engine {
    kx_model_desc d;
    int qkv_dim, q_dim, max_pages, num_pages, max_tokens;
    int device = 0;                                // the GPU this engine lives on (current device at creation)
    int prefix_len = 0;                            // tokens of shared prefix whose KV is already cached (multiple of 16)
    // decode attention over the shared prefix: per-chunk partial (O, max/sum) of every (sequence, q head)
    float *prefix_o = nullptr, *prefix_ml = nullptr;
    size_t prefix_rows = 0;                        // capacity in (chunk, sequence, head) rows
    bool attn_tc = true;                           // prefill attention on tcgen05 (kx_attn_tc.cu) for head_dim 64
    int64_t attn_tc_launches = 0;
    bool share_prefix = true;                      // decode attention reads a cached prefix once per 64 query rows
    bool fused_norm = true;                        // decode (batch <= 256): RMSNorm deferred into the GEMM epilogues
    float *ssq = nullptr, *col_rinv = nullptr;     // ... per-tile sums of squares [hidden / 128][max_batch]; 1 / rms per column
    int eos_id = -1;                               // kx_engine_set_eos: decode stops once every sequence has emitted it
    int *done = nullptr, *h_done = nullptr;        // device [1 + max_batch]: count, per-sequence flag; pinned host copy of the count
    bool fused_argmax = true;                      // lm_head keeps (max, index) per tile instead of writing logits
    std::vector<LayerPlan> layers;
    const void *embed, *lm_head, *norm;
    const float *rope;
    CUtensorMap lm_head_map, lm_head_h[2];
    // activations
    float *resid = nullptr, *logits = nullptr;
    __nv_bfloat16 *xn = nullptr, *qkv = nullptr, *attn = nullptr, *mlp = nullptr;
    CUtensorMap xn_a, attn_a, mlp_a;               // as A operand (prefill)
    CUtensorMap xn_b[3], attn_b[3], mlp_b[3];      // as B operand (decode), box rows 32 / 64 / 128
    int *tokens = nullptr, *pos = nullptr, *seq = nullptr, *last_rows = nullptr, *next_tok = nullptr;
    int *out_ids = nullptr, *block_table = nullptr, *step = nullptr;
    void *scratch = nullptr;
    size_t scratch_bytes = 0;
    cudaStream_t stream = nullptr;                 // the engine's own stream (the decode loop is captured on it)
    // tensor parallelism: partial sums of the row-parallel layers, argmax pairs, NCCL entry points
    float *tp_buf = nullptr;
    void *tp_pairs = nullptr, *tp_gathered = nullptr;
    // fused P2P all-reduce (decode): receive buffers / flags on this GPU, and the peer's, mapped through CUDA IPC
    float *p2p_recv = nullptr, *p2p_peer_recv = nullptr;
    uint32_t *p2p_flag = nullptr, *p2p_peer_flag = nullptr, *p2p_epoch = nullptr;
    int p2p_rows = 0;
    int64_t p2p_launches = 0;
    void *comm = nullptr;
    int (*allreduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    int (*allgather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    // wide decode batches (256 < M <= 512): the row-parallel projections run as split-K token-major GEMMs that
    // leave per-slice partial sums here; the RMSNorm that follows adds them to the residual (pend_* describe them)
    void *argmax_parts = nullptr;                  // fused lm_head argmax: int2 [parts][max_batch]
    float *splitk_part = nullptr;
    size_t splitk_cap = 0;                         // floats
    const float *pend_part = nullptr;
    int pend_splits = 0;
    size_t pend_stride = 0;
    // instantiated decode-step graphs, reused by later generate calls of the same shape (capture + instantiate
    // cost about a millisecond of idle GPU per call otherwise); dropped whenever buffers they point to may change
    struct StepGraph { int B, S, n_new, P; cudaGraphExec_t exec; int64_t launches; };
    std::vector<StepGraph> step_graphs;
    // continuous batching (kx_engine_session_*): per-slot state, device int [max_batch] each -- tokens produced (< 0: the
    // slot is free), how many it may produce, ended; copies of the live positions / last tokens across an admission's
    // prefill (which overwrites e->pos, e->seq and e->next_tok); the identity sequence map of a decode step
    int *s_cnt = nullptr, *s_limit = nullptr, *s_fin = nullptr, *s_pos = nullptr, *s_tok = nullptr, *s_seq = nullptr;
    bool s_open = false, s_need_input = false;     // ... the deferred-norm path's step input has to be rebuilt
    int s_live = 0;                                // slots admitted and not yet released
    int64_t device_bytes = 0;
    std::vector<void *> owned;
    TraceCtx *trace = nullptr;                     // KX_TRACE=1: per-launch timeline of the last decode step
};

namespace {

constexpr int kP2pSlots = 8192;                    // flag / epoch slots of the fused tensor-parallel all-reduce

void drop_step_graphs(kx_engine *e) {
    for (auto &g : e->step_graphs) cudaGraphExecDestroy(g.exec);
    e->step_graphs.clear();
}

int dev_alloc(kx_engine *e, void **p, size_t bytes) {
    KX_CUDA(cudaMalloc(p, bytes));
    e->owned.push_back(*p);
    e->device_bytes += (int64_t)bytes;
    return KX_OK;
}

#define KX_TRY(x) do { int rc__ = (x); if (rc__ != KX_OK) return rc__; } while (0)

int bn_index(int bn) { return bn == 32 ? 0 : bn == 64 ? 1 : 2; }

// decode split-K factor: enough CTAs to keep every SM streaming weights
int pick_splits(int features, int k, int b_tiles) {
    const int tiles = ((features + 127) / 128) * b_tiles;
    const int kb = (k + 63) / 64;
    // one wave at 2 CTAs per SM would be 296 CTAs, but clusters do not pack that tightly: measured
    // (profiles/r01_sweep_decode_gemm.txt) the time jumps by ~50% once tiles * splits exceeds ~256
    int s = 256 / tiles;
    if (s > 8) {                                    // clusters above the portable size 8 (opt-in, up to 16) pack worse:
        s = 192 / tiles;                            // they pay off only while tiles * splits stays below ~192
        if (s > 16) s = 16;
        if (s < 8) s = 8;
    }
    if (s > kb) s = kb;
    if (s < 1) s = 1;
    return s;
}

__global__ void trace_reset_kernel(unsigned long long *buf, int slots) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < slots) {
        buf[8 * i] = ~0ull; buf[8 * i + 1] = ~0ull;
        for (int k = 2; k < 8; ++k) buf[8 * i + k] = 0;
    }
}

__global__ void advance_step_kernel(int *pos, int *step, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pos[i] += 1;
    if (i == 0) *step += 1;
}

__global__ void store_token_kernel(const int *next_tok, int *out_ids, const int *step, int B, int n_new) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) out_ids[(size_t)b * n_new + *step] = next_tok[b];
}

struct GemmCall {
    const CUtensorMap *a, *b;
    int a_rows, b_rows, k, epi, swap, splits, bn;
    void *out; int ldo; const float *bias;
    const CUtensorMap *b_half = nullptr;           // token-major pair kernel: weight map with a half-tile box
    int pair_bn = 0;                               // ... and its tile width (128 or 256)
    const GemmParams *extra = nullptr;             // EPI_QKV_ROPE: the rope / cache fields
    bool tp_fused = false;                         // EPI_RESID + all-reduce over peer memory in the epilogue
    size_t split_stride = 0;                       // pair kernel with splits > 1: distance between the slices' outputs
    const char *label = "gemm";                    // name in the diagnostic timeline
    bool ssq_in = false;                           // deferred RMSNorm: scale the accumulator columns by 1 / rms (from e->ssq)
    const void *norm_g = nullptr;                  // EPI_RESID_NORM: gain of the next norm
    const float *col_scale = nullptr;
};

int run_gemm(kx_engine *e, const GemmCall &c, cudaStream_t st) {
    GemmParams p{};
    p.a_rows = c.a_rows; p.b_rows = c.b_rows; p.k = c.k; p.out = c.out; p.ldo = c.ldo; p.bias = c.bias;
    p.swap = c.swap; p.epilogue = c.epi; p.splits = c.splits; p.split_stride = c.split_stride;
    p.trace = trace_slot(c.label);
    p.ssq_ld = e->d.max_batch; p.ssq_tiles = e->d.hidden / 128; p.norm_eps = e->d.eps; p.norm_inv_n = 1.0f / (float)e->d.hidden;
    if (c.ssq_in) p.ssq_in = e->ssq;
    if (c.epi == EPI_RESID_NORM) { p.norm_g = c.norm_g; p.norm_out = e->xn; p.ssq_out = e->ssq; }
    p.col_scale = c.col_scale;
    if (c.extra) {
        p.pos = c.extra->pos; p.seq = c.extra->seq; p.rope = c.extra->rope; p.kv_pages = c.extra->kv_pages;
        p.block_table = c.extra->block_table; p.max_pages = c.extra->max_pages;
        p.n_q = c.extra->n_q; p.n_kv = c.extra->n_kv; p.hd = c.extra->hd;
    }
    if (c.tp_fused) {
        p.tp_peer_recv = e->p2p_peer_recv; p.tp_peer_flag = e->p2p_peer_flag; p.tp_recv = e->p2p_recv;
        p.tp_flag = e->p2p_flag; p.tp_epoch = e->p2p_epoch; p.tp_rows = e->p2p_rows;
        p.tp_seq = (int)(e->p2p_launches & 0x7fffffff);
        const int64_t slots = (int64_t)((c.a_rows + 127) / 128) * ((c.b_rows + c.bn - 1) / c.bn) * (c.splits < 1 ? 1 : c.splits);
        KX_REQUIRE(slots <= kP2pSlots, KX_ETOOBIG, "engine: fused all-reduce would use %lld flag slots, %d exist", (long long)slots, kP2pSlots);
    }
    if (!c.swap && c.b_half) return gemm_launch_pair(*c.a, *c.b_half, p, c.pair_bn, st);
    KX_REQUIRE(gemm_scratch_bytes(p, c.bn) <= e->scratch_bytes, KX_ENOMEM, "engine: split-K scratch too small");
    return gemm_launch(*c.a, *c.b, p, c.bn, e->scratch, st);
}

// A linear layer y = x W^T over M token rows with a fused epilogue.  Chooses the kernel:
//   * token-major CTA-pair kernel when the output has enough 256-row x BN tiles to fill the GPU
//     (prefill, and the widest layers of a large decode batch);
//   * otherwise the swap-AB weight-streaming kernel, split-K inside a cluster.
struct Linear {
    const CUtensorMap *w_full, *w_half;            // weight maps: 128-row box; half-pair-tile boxes [64 rows, 128 rows]
    const CUtensorMap *x_a;                        // activation as A operand (128-row box)
    const CUtensorMap *x_b;                        // activation as B operand, indexed by bn_index
    int features, k;
    const char *label = "gemm";
};

// tile width the CTA-pair kernel would use for this layer at M rows (0: another kernel is chosen)
int linear_pair_bn(const Linear &l, int M) {
    if (M < 256) return 0;
    const int m_pairs = (M + 255) / 256;
    const int t256 = m_pairs * ((l.features + 255) / 256), t128 = m_pairs * ((l.features + 127) / 128);
    // 256-wide tiles halve the operand bytes per flop; a partial last tile is worth it up to ~15 % padding
    // (Qwen2.5-0.5B: 896 and 1152 features -> prefill 17.2 -> 16.7 ms)
    constexpr int waste_pct = 15;
    const int padded = ((l.features + 255) / 256) * 256;
    const bool little_waste = (padded - l.features) * 100 <= waste_pct * l.features;   // partial last tile is cheap enough
    if ((l.features % 256 == 0 || l.features >= 2048 || little_waste) && t256 >= 64) return 256;
    // fewer 128-wide tiles than this leave most SMs idle: the weight-streaming kernel with split-K covers them
    constexpr int min128 = 24;     // (QKV of a 256 / 512-wide decode batch through the cluster kernel instead: 2.64 / 3.20 vs 2.47 / 3.07 ms per step)
    return t128 >= min128 ? 128 : 0;
}

int linear(kx_engine *e, const Linear &l, int M, int epi, void *out, int ldo, const float *bias, cudaStream_t st,
           const GemmParams *extra = nullptr) {
    if (const int pbn = linear_pair_bn(l, M)) {
        GemmCall c{l.x_a, l.w_full, M, l.features, l.k, epi, 0, 1, 128, out, ldo, bias, l.w_half + (pbn == 256), pbn, extra};
        c.label = l.label;
        return run_gemm(e, c, st);
    }
    if (M > 512) {   // tall but very narrow: the single-CTA token-major kernel (128 x 128 tiles)
        GemmCall c{l.x_a, l.w_full, M, l.features, l.k, epi, 0, 1, 128, out, ldo, bias};
        c.label = l.label;
        return run_gemm(e, c, st);
    }
    const int bn = gemm_pick_bn(1, M), bi = bn_index(bn);
    const int b_tiles = (M + bn - 1) / bn;
    GemmCall c{l.w_full, l.x_b + bi, l.features, M, l.k, epi, 1, pick_splits(l.features, l.k, b_tiles), bn, out, ldo, bias};
    c.label = l.label;
    return run_gemm(e, c, st);
}

// RMSNorm of the residual rows; folds in (and clears) split-K partials a preceding projection left pending
int norm_rows(kx_engine *e, const void *g, int M, const int *row_map, cudaStream_t st) {
    const int rc = op_rmsnorm(e->resid, g, e->xn, M, e->d.hidden, e->d.eps, row_map, st, e->pend_part, e->pend_splits, e->pend_stride);
    e->pend_part = nullptr; e->pend_splits = 0; e->pend_stride = 0;
    return rc;
}

// Row-parallel linear layer whose result is added to the residual stream.  With tensor
// parallelism every rank holds a partial sum: it goes to tp_buf, is all-reduced, then added.
int linear_resid(kx_engine *e, const Linear &l, int M, cudaStream_t st, bool decode) {
    const kx_model_desc &d = e->d;
    constexpr int splitk_min_m = 256;
    if (d.tp_size <= 1 && decode && M >= splitk_min_m && M <= 512 && d.hidden % 256 == 0 && e->splitk_part) {
        // 2 x (hidden / 256) tiles of 256 x 256 would use a fraction of the SMs: cut k so that ~64 CTA pairs work,
        // slices land in splitk_part and the next RMSNorm adds them to the residual
        const int tiles = ((M + 255) / 256) * (d.hidden / 256);
        int S = 74 / tiles;
        const int kb = (l.k + 63) / 64;
        if (S > 8) S = 8;
        if (S > kb / 4) S = kb / 4;
        if (S >= 2 && (size_t)S * M * d.hidden <= e->splitk_cap) {
            GemmCall c{l.x_a, l.w_full, M, l.features, l.k, EPI_F32, 0, S, 128, e->splitk_part, d.hidden, nullptr, l.w_half + 1, 256};
            c.split_stride = (size_t)M * d.hidden;
            c.label = l.label;
            KX_TRY(run_gemm(e, c, st));
            e->pend_part = e->splitk_part; e->pend_splits = S; e->pend_stride = c.split_stride;
            return KX_OK;
        }
    }
    if (d.tp_size <= 1) return linear(e, l, M, EPI_RESID, e->resid, d.hidden, nullptr, st);
    if (d.tp_size == 2 && e->p2p_peer_recv && M <= e->p2p_rows && M < 256 && linear_pair_bn(l, M) == 0) {
        // decode: the swap-AB cluster GEMM reduces across the two ranks itself, over NVLink peer memory
        const int bn = gemm_pick_bn(1, M), bi = bn_index(bn);
        const int b_tiles = (M + bn - 1) / bn;
        GemmCall c{l.w_full, l.x_b + bi, l.features, M, l.k, EPI_RESID, 1, pick_splits(l.features, l.k, b_tiles), bn,
                   e->resid, d.hidden, nullptr};
        c.tp_fused = true;
        c.label = l.label;
        const int rc = run_gemm(e, c, st);             // its receive half-buffer is picked by the parity of p2p_launches
        ++e->p2p_launches;
        return rc;
    }
    KX_REQUIRE(e->comm && e->allreduce, KX_EINVAL, "engine: tensor parallelism needs kx_engine_set_collectives");
    KX_TRY(linear(e, l, M, EPI_F32, e->tp_buf, d.hidden, nullptr, st));
    const int rc = e->allreduce(e->tp_buf, e->tp_buf, (size_t)M * d.hidden, /*ncclFloat32*/ 7, /*ncclSum*/ 0, e->comm, st);
    KX_REQUIRE(rc == 0, KX_ECUDA, "engine: ncclAllReduce failed with %d", rc);
    return op_add_inplace(e->resid, e->tp_buf, (size_t)M * d.hidden, st);
}

// One transformer layer over M token rows.  decode: rows are one token per sequence.
int layer_forward(kx_engine *e, int li, int M, bool decode, int nb, int S, int b0, int max_T, cudaStream_t st, int pos0 = 0) {
    const kx_model_desc &d = e->d;
    LayerPlan &L = e->layers[li];
    const Linear qkv{&L.wqkv, L.wqkv_h, &e->xn_a, e->xn_b, e->qkv_dim, d.hidden, "qkv"};
    const Linear wo{&L.wo, L.wo_h, &e->attn_a, e->attn_b, d.hidden, e->q_dim, "wo"};
    const Linear wgu{&L.wgu, L.wgu_h, &e->xn_a, e->xn_b, 2 * d.inter, d.hidden, "gate_up"};
    const Linear wd{&L.wd, L.wd_h, &e->mlp_a, e->mlp_b, d.hidden, d.inter, "down"};
    KX_TRY(norm_rows(e, L.w.ln1, M, nullptr, st));
    const int head_w = d.head_dim;
    const int qkv_pbn = linear_pair_bn(qkv, M);
    const bool fused_rope = !decode && qkv_pbn && qkv_pbn % head_w == 0;   // QKV epilogue also rotates and appends
    if (fused_rope) {
        GemmParams x{};
        x.pos = e->pos; x.seq = e->seq; x.rope = e->rope; x.kv_pages = L.kv_pages; x.block_table = e->block_table;
        x.max_pages = e->max_pages; x.n_q = d.n_q; x.n_kv = d.n_kv; x.hd = d.head_dim;
        KX_TRY(linear(e, qkv, M, EPI_QKV_ROPE, e->qkv, e->qkv_dim, L.w.bqkv, st, &x));
    } else {
        KX_TRY(linear(e, qkv, M, EPI_BF16, e->qkv, e->qkv_dim, L.w.bqkv, st));
    }
    if (decode) {
        int key0 = 0, n_part = 0;
        if (e->share_prefix && e->prefix_len >= 64 && M >= 2 && e->prefix_o &&
            (size_t)attn_prefix_parts(M, d.n_q, d.n_kv, e->prefix_len, nullptr) * M * d.n_q <= e->prefix_rows) {
            // the batch shares its first prefix_len keys: read them once per 64 query rows, not once per sequence
            KX_TRY(op_attn_prefix(e->qkv, e->qkv_dim, L.kv_pages, e->block_table + (size_t)d.max_batch * e->max_pages, e->pos,
                                  e->rope, M, d.n_q, d.n_kv, d.head_dim, e->prefix_len, e->prefix_o, e->prefix_ml, &n_part, st));
            key0 = e->prefix_len;
        }
        KX_TRY(op_attn_decode(e->qkv, e->qkv_dim, e->attn, e->q_dim, L.kv_pages, e->block_table, e->max_pages, e->pos,
                              e->rope, M, max_T, d.n_q, d.n_kv, d.head_dim, key0, e->prefix_o, e->prefix_ml, n_part, st));
    } else {
        if (!fused_rope)
            KX_TRY(op_rope_append(e->qkv, e->qkv_dim, e->pos, e->seq, e->rope, L.kv_pages, e->block_table, e->max_pages, M,
                                  d.n_q, d.n_kv, d.head_dim, st));
        if (e->attn_tc && d.head_dim == 64) {
            ++e->attn_tc_launches;
            KX_TRY(op_attn_prefill_tc(e->qkv, e->qkv_dim, e->attn, e->q_dim, L.kv_pages, e->num_pages,
                                      e->block_table + (size_t)b0 * e->max_pages, e->max_pages, nb, S, pos0, d.n_q, d.n_kv, st));
        } else
        KX_TRY(op_attn_prefill(e->qkv, e->qkv_dim, e->attn, e->q_dim, L.kv_pages, e->block_table + (size_t)b0 * e->max_pages,
                               e->max_pages, nb, S, pos0, d.n_q, d.n_kv, d.head_dim, st));
    }
    KX_TRY(linear_resid(e, wo, M, st, decode));
    KX_TRY(norm_rows(e, L.w.ln2, M, nullptr, st));
    KX_TRY(linear(e, wgu, M, EPI_SWIGLU, e->mlp, d.inter, nullptr, st));
    KX_TRY(linear_resid(e, wd, M, st, decode));
    return KX_OK;
}

// Decode on the deferred-RMSNorm path (batch <= 256: every projection is a swap-AB cluster GEMM): single-GPU replicas, and
// tensor-parallel pairs whose row-parallel projections all-reduce inside their epilogue over peer memory.
bool fused_norm_decode(const kx_engine *e, int B) {
    // (Llama-3.2-1B, ms per step on this path / on the token-major pair-kernel path: batch 192: 2.15 / 2.53, 256: 2.35 / 2.60,
    // 320: 2.94 / 2.70, 384: 3.24 / 2.77, 512: 3.28 / 3.02)
    if (!e->fused_norm || B > 256) return false;
    return e->d.tp_size <= 1 || (e->d.tp_size == 2 && e->p2p_peer_recv && B <= e->p2p_rows);
}

// One layer of a decode step with the RMSNorms folded into the projections (kx_gemm.cuh): five launches --
// QKV (scaled by 1 / rms), attention, wo (+ residual, prepares norm 2), gate/up (scaled, SwiGLU), down (+ residual,
// prepares the next layer's norm 1 or the final norm).  e->xn holds bf16(residual * g) of the pending norm, e->ssq
// its per-tile sums of squares.
int layer_forward_fused(kx_engine *e, int li, int B, int max_T, cudaStream_t st) {
    const kx_model_desc &d = e->d;
    LayerPlan &L = e->layers[li];
    // batch tiles of 64 columns (gate/up: see below): a 128-wide tile doubles the split-K exchange and the epilogue of every CTA
    // (Llama-3.2-1B, batch 128: 1.9 ms per step with one 128-wide tile, 1.55 with two 64-wide ones)
    const int bn = B <= 32 ? 32 : 64, bi = bn_index(bn);
    const int b_tiles = (B + bn - 1) / bn;
    const bool last = li + 1 == d.layers;
    auto gemm = [&](const CUtensorMap *w, const CUtensorMap *x, int features, int k, int epi, void *out, int ldo, const float *bias,
                    const char *label, bool ssq_in, const void *norm_g) {
        GemmCall c{w, x + bi, features, B, k, epi, 1, pick_splits(features, k, b_tiles), bn, out, ldo, bias};
        if (bn == 64 && ((features + 127) / 128) * b_tiles > 296) {   // (at batch 128, 256 tiles: 128-wide ones measure 1.65 vs 1.46 ms)
            // more 64-wide tiles than one resident wave holds (gate/up from batch 129): 128-wide tiles, which then need
            // no split-K exchange either (batch 256: 2.64 -> 2.35 ms per step, batch 192: 2.24 -> 2.15)
            const int bt2 = (B + 127) / 128;
            c.b = x + bn_index(128); c.bn = 128; c.splits = pick_splits(features, k, bt2);
        }
        c.label = label; c.ssq_in = ssq_in; c.norm_g = norm_g;
        if (epi == EPI_RESID_NORM && d.tp_size > 1) {      // row-parallel: the two ranks' partial sums meet in the epilogue
            c.tp_fused = true;
            const int rc = run_gemm(e, c, st);             // (receive half-buffer: parity of p2p_launches)
            ++e->p2p_launches;
            return rc;
        }
        return run_gemm(e, c, st);
    };
    KX_TRY(gemm(&L.wqkv, e->xn_b, e->qkv_dim, d.hidden, EPI_BF16, e->qkv, e->qkv_dim, L.w.bqkv, "qkv", true, nullptr));
    int key0 = 0, n_part = 0;
    if (e->share_prefix && e->prefix_len >= 64 && B >= 2 && e->prefix_o &&
        (size_t)attn_prefix_parts(B, d.n_q, d.n_kv, e->prefix_len, nullptr) * B * d.n_q <= e->prefix_rows) {
        KX_TRY(op_attn_prefix(e->qkv, e->qkv_dim, L.kv_pages, e->block_table + (size_t)d.max_batch * e->max_pages, e->pos,
                              e->rope, B, d.n_q, d.n_kv, d.head_dim, e->prefix_len, e->prefix_o, e->prefix_ml, &n_part, st));
        key0 = e->prefix_len;
    }
    KX_TRY(op_attn_decode(e->qkv, e->qkv_dim, e->attn, e->q_dim, L.kv_pages, e->block_table, e->max_pages, e->pos,
                          e->rope, B, max_T, d.n_q, d.n_kv, d.head_dim, key0, e->prefix_o, e->prefix_ml, n_part, st));
    KX_TRY(gemm(&L.wo, e->attn_b, d.hidden, e->q_dim, EPI_RESID_NORM, e->resid, d.hidden, nullptr, "wo", false, L.w.ln2));
    KX_TRY(gemm(&L.wgu, e->xn_b, 2 * d.inter, d.hidden, EPI_SWIGLU, e->mlp, d.inter, nullptr, "gate_up", true, nullptr));
    KX_TRY(gemm(&L.wd, e->mlp_b, d.hidden, d.inter, EPI_RESID_NORM, e->resid, d.hidden, nullptr, "down", false,
                last ? e->norm : e->layers[li + 1].w.ln1));
    return KX_OK;
}

// greedy token per sequence from e->logits (this rank's vocabulary shard under tensor parallelism)
int greedy_pick(kx_engine *e, int B, cudaStream_t st) {
    const kx_model_desc &d = e->d;
    if (d.tp_size <= 1) return op_argmax(e->logits, B, d.vocab, d.vocab, e->next_tok, nullptr, st);
    KX_REQUIRE(e->comm && e->allgather, KX_EINVAL, "engine: tensor parallelism needs kx_engine_set_collectives");
    KX_TRY(op_argmax_pairs(e->logits, B, d.vocab, d.vocab, d.vocab_offset, e->tp_pairs, st));
    const int rc = e->allgather(e->tp_pairs, e->tp_gathered, (size_t)B * 8, /*ncclInt8*/ 0, e->comm, st);
    KX_REQUIRE(rc == 0, KX_ECUDA, "engine: ncclAllGather failed with %d", rc);
    return op_pick(e->tp_gathered, d.tp_size, B, e->next_tok, st);
}

// final norm over `rows` (row_map selects which residual rows), lm_head, argmax, store
// deferred: e->xn already holds bf16(residual * g_final) and e->ssq its sums of squares (decode on the fused-norm path);
// the caller's tail kernel then stores the token instead of store_token_kernel (store = false: so it does on the classic path)
int head_forward(kx_engine *e, int B, const int *row_map, int n_new, float *d_logits_step, cudaStream_t st, bool deferred = false,
                 bool store = true) {
    const kx_model_desc &d = e->d;
    if (!deferred) KX_TRY(norm_rows(e, e->norm, B, row_map, st));
    const Linear head{&e->lm_head_map, e->lm_head_h, &e->xn_a, e->xn_b, d.vocab, d.hidden, "lm_head"};
    // Nobody asked for the logits: the lm_head epilogue keeps only (max, index) per token and tile, a small
    // second kernel finishes the pick; [B, vocab] is never written.  (Small vocabularies take the cluster split-K
    // GEMM, which has no such epilogue, and B > 512 the single-CTA kernel: both use the logits path.)
    const int pbn = linear_pair_bn(head, B);
    const int v_tiles = (d.vocab + 127) / 128;
    const int sbn = gemm_pick_bn(1, B);
    const bool want_fused = e->fused_argmax;
    const bool fused = want_fused && !d_logits_step && e->argmax_parts &&
                       (pbn || (B <= 512 && pick_splits(d.vocab, d.hidden, (B + sbn - 1) / sbn) == 1));
    if (fused) {
        const int n_parts = pbn ? (d.vocab + pbn - 1) / pbn : 4 * v_tiles;
        KX_TRY(linear(e, head, B, EPI_ARGMAX, e->argmax_parts, d.max_batch, nullptr, st));
        if (d.tp_size <= 1) {
            KX_TRY(op_argmax_parts(e->argmax_parts, n_parts, d.max_batch, B, 0, e->next_tok, nullptr, nullptr, st));
        } else {
            KX_REQUIRE(e->comm && e->allgather, KX_EINVAL, "engine: tensor parallelism needs kx_engine_set_collectives");
            KX_TRY(op_argmax_parts(e->argmax_parts, n_parts, d.max_batch, B, d.vocab_offset, nullptr, nullptr, e->tp_pairs, st));
            const int rc = e->allgather(e->tp_pairs, e->tp_gathered, (size_t)B * 8, /*ncclInt8*/ 0, e->comm, st);
            KX_REQUIRE(rc == 0, KX_ECUDA, "engine: ncclAllGather failed with %d", rc);
            KX_TRY(op_pick(e->tp_gathered, d.tp_size, B, e->next_tok, st));
        }
        if (deferred || !store) return KX_OK;          // argmax is unchanged by the positive per-token factor 1 / rms
        store_token_kernel<<<(B + 127) / 128, 128, 0, st>>>(e->next_tok, e->out_ids, e->step, B, n_new);
        count_launch();
        KX_CUDA(cudaGetLastError());
        return KX_OK;
    }
    if (deferred) {                                    // the logits themselves are wanted: scale column b by 1 / rms_b
        KX_TRY(op_col_rinv(e->ssq, d.hidden / 128, d.max_batch, B, d.hidden, d.eps, e->col_rinv, st));
        const int bn = gemm_pick_bn(1, B), bi = bn_index(bn);
        GemmCall c{&e->lm_head_map, e->xn_b + bi, d.vocab, B, d.hidden, EPI_F32, 1, 1, bn, e->logits, d.vocab, nullptr};
        c.label = "lm_head"; c.col_scale = e->col_rinv;
        KX_TRY(run_gemm(e, c, st));
    } else {
        KX_TRY(linear(e, head, B, EPI_F32, e->logits, d.vocab, nullptr, st));
    }
    if (d_logits_step)
        KX_CUDA(cudaMemcpyAsync(d_logits_step, e->logits, sizeof(float) * (size_t)B * d.vocab, cudaMemcpyDeviceToDevice, st));
    KX_TRY(greedy_pick(e, B, st));
    if (deferred || !store) return KX_OK;
    store_token_kernel<<<(B + 127) / 128, 128, 0, st>>>(e->next_tok, e->out_ids, e->step, B, n_new);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

// Prefill of B whole sequences of S tokens each (host int32 [B, S]) into the KV pages of sequences slot0 .. slot0 + B - 1,
// in chunks of at most prefill_chunk tokens; leaves the logits of every sequence's last prompt row in
// e->logits[slot0 + b].  Overwrites the per-token buffers (tokens, pos, seq, activations).
int prefill_sequences(kx_engine *e, const int32_t *h_tokens, int slot0, int B, int S, cudaStream_t st, int64_t *h2d) {
    const kx_model_desc &d = e->d;
    const int P = e->prefix_len;
    int seqs_per_chunk = d.prefill_chunk / S;
    if (seqs_per_chunk < 1) seqs_per_chunk = 1;
    KX_REQUIRE((int64_t)seqs_per_chunk * S <= e->max_tokens || S <= e->max_tokens, KX_ETOOBIG, "kx_engine_generate: prompt too long");
    if ((int64_t)seqs_per_chunk * S > e->max_tokens) seqs_per_chunk = e->max_tokens / S;
    std::vector<int> hpos, hseq;
    for (int b0 = 0; b0 < B; b0 += seqs_per_chunk) {
        const int nb = (B - b0 < seqs_per_chunk) ? B - b0 : seqs_per_chunk;
        const int M = nb * S;
        hpos.resize(M); hseq.resize(M);
        for (int b = 0; b < nb; ++b)
            for (int t = 0; t < S; ++t) { hpos[b * S + t] = P + t; hseq[b * S + t] = slot0 + b0 + b; }
        KX_CUDA(cudaMemcpyAsync(e->tokens, h_tokens + (size_t)b0 * S, (size_t)M * 4, cudaMemcpyHostToDevice, st));
        KX_CUDA(cudaMemcpyAsync(e->pos, hpos.data(), (size_t)M * 4, cudaMemcpyHostToDevice, st));
        KX_CUDA(cudaMemcpyAsync(e->seq, hseq.data(), (size_t)M * 4, cudaMemcpyHostToDevice, st));
        *h2d += (int64_t)3 * M * 4;
        KX_CUDA(cudaStreamSynchronize(st));          // hpos/hseq are reused by the next chunk
        KX_TRY(op_embed(e->tokens, e->embed, e->resid, M, d.hidden, st));
        for (int li = 0; li < d.layers; ++li) KX_TRY(layer_forward(e, li, M, false, nb, S, slot0 + b0, 0, st, P));
        // first token of this chunk's sequences: final norm on each last prompt row
        std::vector<int> rows(nb);
        for (int b = 0; b < nb; ++b) rows[b] = b * S + S - 1;
        KX_CUDA(cudaMemcpyAsync(e->last_rows, rows.data(), (size_t)nb * 4, cudaMemcpyHostToDevice, st));
        *h2d += (int64_t)nb * 4;
        KX_CUDA(cudaStreamSynchronize(st));
        {
            const Linear head{&e->lm_head_map, e->lm_head_h, &e->xn_a, e->xn_b, d.vocab, d.hidden};
            KX_TRY(op_rmsnorm(e->resid, e->norm, e->xn, nb, d.hidden, d.eps, e->last_rows, st));
            KX_TRY(linear(e, head, nb, EPI_F32, e->logits + (size_t)(slot0 + b0) * d.vocab, d.vocab, nullptr, st));
        }
    }
    return KX_OK;
}

}  // namespace

extern "C" int kx_engine_create(const kx_model_desc *desc, const kx_model_weights *w, kx_engine **out) {
    KX_REQUIRE(desc && w && out, KX_EINVAL, "kx_engine_create: NULL argument");
    const kx_model_desc &d = *desc;
    KX_REQUIRE(d.head_dim == 64 || d.head_dim == 128, KX_EINVAL, "kx_engine_create: head_dim must be 64 or 128");
    KX_REQUIRE(d.hidden % 128 == 0 && d.inter % 8 == 0 && d.vocab % 8 == 0, KX_EINVAL,
               "kx_engine_create: hidden must be a multiple of 128, inter and vocab of 8");
    KX_REQUIRE(d.n_q % d.n_kv == 0, KX_EINVAL, "kx_engine_create: n_q must be a multiple of n_kv");
    KX_REQUIRE(d.max_seq <= d.max_pos, KX_EINVAL, "kx_engine_create: max_seq exceeds the rope table");
    kx_engine *e = new kx_engine();
    e->d = d;
    if (cudaGetDevice(&e->device) != cudaSuccess) { delete e; set_error("kx_engine_create: no current CUDA device"); return KX_ECUDA; }
    e->share_prefix = !debug_flag("no-shared-prefix");
    e->attn_tc = !debug_flag("no-attn-tc");
    e->fused_norm = !debug_flag("no-fused-norm");
    e->fused_argmax = !debug_flag("no-fused-argmax");
    if (e->d.prefill_chunk <= 0) e->d.prefill_chunk = 32768;
    e->qkv_dim = (d.n_q + 2 * d.n_kv) * d.head_dim;
    e->q_dim = d.n_q * d.head_dim;
    e->max_pages = (d.max_seq + KX_KV_PAGE - 1) / KX_KV_PAGE;
    e->num_pages = e->max_pages * (d.max_batch + 1);   // + one sequence worth of pages for a shared prefix
    if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess) { delete e; set_error("kx_engine_create: cannot create a stream"); return KX_ECUDA; }
    e->embed = w->embed; e->lm_head = w->lm_head; e->norm = w->norm; e->rope = w->rope;
    // prefill works on whole sequences, so the token buffers hold at least one chunk or one full batch of prompts
    e->max_tokens = e->d.prefill_chunk > d.max_batch ? e->d.prefill_chunk : d.max_batch;
    if (e->max_tokens < d.max_seq) e->max_tokens = d.max_seq;
    const size_t Mx = (size_t)e->max_tokens;
    int rc = KX_OK;
    auto fail = [&](int code) { kx_engine_destroy(e); return code; };
#define KX_E(x) do { rc = (x); if (rc != KX_OK) return fail(rc); } while (0)
    KX_E(dev_alloc(e, (void **)&e->resid, Mx * d.hidden * 4));
    KX_E(dev_alloc(e, (void **)&e->xn, Mx * d.hidden * 2));
    KX_E(dev_alloc(e, (void **)&e->qkv, Mx * e->qkv_dim * 2));
    KX_E(dev_alloc(e, (void **)&e->attn, Mx * e->q_dim * 2));
    KX_E(dev_alloc(e, (void **)&e->mlp, Mx * d.inter * 2));
    KX_E(dev_alloc(e, (void **)&e->logits, (size_t)d.max_batch * d.vocab * 4));
    KX_E(dev_alloc(e, (void **)&e->tokens, Mx * 4));
    KX_E(dev_alloc(e, (void **)&e->pos, Mx * 4));
    KX_E(dev_alloc(e, (void **)&e->seq, Mx * 4));
    KX_E(dev_alloc(e, (void **)&e->last_rows, (size_t)d.max_batch * 4));
    KX_E(dev_alloc(e, (void **)&e->next_tok, (size_t)d.max_batch * 4));
    KX_E(dev_alloc(e, (void **)&e->out_ids, (size_t)d.max_batch * d.max_seq * 4));
    KX_E(dev_alloc(e, (void **)&e->block_table, (size_t)(d.max_batch + 1) * e->max_pages * 4));
    KX_E(dev_alloc(e, (void **)&e->step, 4));
    if (d.tp_size > 1) {
        KX_E(dev_alloc(e, (void **)&e->tp_buf, Mx * d.hidden * 4));
        KX_E(dev_alloc(e, &e->tp_pairs, (size_t)d.max_batch * 8));
        KX_E(dev_alloc(e, &e->tp_gathered, (size_t)d.max_batch * 8 * d.tp_size));
    }
    // split-K scratch for the largest decode GEMM: tiles * 128 * bn floats + tickets
    {
        size_t best = 0;
        const int feats[3] = {e->qkv_dim, d.hidden, 2 * d.inter};   // lm_head never splits
        for (int f : feats) {
            const size_t tiles = (size_t)((f + 127) / 128) * (size_t)((d.max_batch + 127) / 128 + 1);
            const size_t b = tiles * 128 * 128 * 4 + tiles * 4;
            if (b > best) best = b;
        }
        e->scratch_bytes = best;
        KX_E(dev_alloc(e, &e->scratch, best));
        if (cudaMemset(e->scratch, 0, best) != cudaSuccess) return fail(KX_ECUDA);
    }
    // static block table: sequence b owns pages [b * max_pages, (b+1) * max_pages)
    {
        std::vector<int> bt((size_t)(d.max_batch + 1) * e->max_pages);
        for (size_t i = 0; i < bt.size(); ++i) bt[i] = (int)i;
        if (cudaMemcpy(e->block_table, bt.data(), bt.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess) return fail(KX_ECUDA);
    }
    const size_t page_bytes = (size_t)2 * d.n_kv * KX_KV_PAGE * d.head_dim * 2;
    e->layers.resize(d.layers);
    for (int i = 0; i < d.layers; ++i) {
        LayerPlan &L = e->layers[i];
        L.w = w->layers[i];
        KX_E(dev_alloc(e, &L.kv_pages, page_bytes * e->num_pages));
        // Zeroed once: the tcgen05 prefill kernel multiplies whole 128-key V tiles by P, and keys past the end of a
        // sequence get P = 0 exactly -- which only cancels finite values (0 x NaN from never-written memory is NaN).
        if (cudaMemset(L.kv_pages, 0, page_bytes * e->num_pages) != cudaSuccess) return fail(KX_ECUDA);
        // the CTA-pair GEMM multicasts weight tiles: they have to live on this GPU
        KX_E(require_local_device_memory(L.w.wqkv, "kx_engine_create: wqkv"));
        KX_E(require_local_device_memory(L.w.wo, "kx_engine_create: wo"));
        KX_E(require_local_device_memory(L.w.wgu, "kx_engine_create: wgu"));
        KX_E(require_local_device_memory(L.w.wd, "kx_engine_create: wd"));
        KX_E(make_tmap_bf16_2d(&L.wqkv, L.w.wqkv, e->qkv_dim, d.hidden, d.hidden, 128));
        KX_E(make_tmap_bf16_2d(&L.wo, L.w.wo, d.hidden, e->q_dim, e->q_dim, 128));
        KX_E(make_tmap_bf16_2d(&L.wgu, L.w.wgu, 2 * d.inter, d.hidden, d.hidden, 128));
        KX_E(make_tmap_bf16_2d(&L.wd, L.w.wd, d.hidden, d.inter, d.inter, 128));
        for (int j = 0; j < 2; ++j) {              // pair kernel: box = half of its tile width
            const int box = j ? 128 : 64;
            KX_E(make_tmap_bf16_2d(&L.wqkv_h[j], L.w.wqkv, e->qkv_dim, d.hidden, d.hidden, box));
            KX_E(make_tmap_bf16_2d(&L.wo_h[j], L.w.wo, d.hidden, e->q_dim, e->q_dim, box));
            KX_E(make_tmap_bf16_2d(&L.wgu_h[j], L.w.wgu, 2 * d.inter, d.hidden, d.hidden, box));
            KX_E(make_tmap_bf16_2d(&L.wd_h[j], L.w.wd, d.hidden, d.inter, d.inter, box));
        }
    }
    KX_E(require_local_device_memory(e->lm_head, "kx_engine_create: lm_head"));
    KX_E(make_tmap_bf16_2d(&e->lm_head_map, e->lm_head, d.vocab, d.hidden, d.hidden, 128));
    KX_E(make_tmap_bf16_2d(&e->lm_head_h[0], e->lm_head, d.vocab, d.hidden, d.hidden, 64));
    KX_E(make_tmap_bf16_2d(&e->lm_head_h[1], e->lm_head, d.vocab, d.hidden, d.hidden, 128));
    KX_E(make_tmap_bf16_2d(&e->xn_a, e->xn, Mx, d.hidden, d.hidden, 128));
    KX_E(make_tmap_bf16_2d(&e->attn_a, e->attn, Mx, e->q_dim, e->q_dim, 128));
    KX_E(make_tmap_bf16_2d(&e->mlp_a, e->mlp, Mx, d.inter, d.inter, 128));
    const int bns[3] = {32, 64, 128};
    for (int j = 0; j < 3; ++j) {
        KX_E(make_tmap_bf16_2d(&e->xn_b[j], e->xn, Mx, d.hidden, d.hidden, bns[j]));
        KX_E(make_tmap_bf16_2d(&e->attn_b[j], e->attn, Mx, e->q_dim, e->q_dim, bns[j]));
        KX_E(make_tmap_bf16_2d(&e->mlp_b[j], e->mlp, Mx, d.inter, d.inter, bns[j]));
    }
    KX_E(dev_alloc(e, &e->argmax_parts, (size_t)4 * ((d.vocab + 127) / 128) * d.max_batch * 8));
    if (d.tp_size <= 1 && d.max_batch > 128 && d.hidden % 256 == 0) {
        e->splitk_cap = (size_t)8 * (d.max_batch < 512 ? d.max_batch : 512) * d.hidden;
        KX_E(dev_alloc(e, (void **)&e->splitk_part, e->splitk_cap * sizeof(float)));
    }
    if (debug_flag("trace")) {
        e->trace = new TraceCtx();
        e->trace->cap = 2048;
        KX_E(dev_alloc(e, (void **)&e->trace->buf, sizeof(unsigned long long) * 8 * e->trace->cap));
    }
    KX_E(dev_alloc(e, (void **)&e->ssq, sizeof(float) * (size_t)(d.hidden / 128) * d.max_batch));
    KX_E(dev_alloc(e, (void **)&e->col_rinv, sizeof(float) * (size_t)d.max_batch));
    KX_E(dev_alloc(e, (void **)&e->done, sizeof(int) * (size_t)(1 + d.max_batch)));
    if (cudaMallocHost((void **)&e->h_done, sizeof(int)) != cudaSuccess) return fail(KX_ECUDA);
#undef KX_E
    *out = e;
    return KX_OK;
}

// ---- fused P2P all-reduce set-up: export this rank's buffers, import the peer's (CUDA IPC) ----

extern "C" int kx_engine_p2p_export(kx_engine *e, void *handles /* 2 x 64 bytes */) {
    KX_REQUIRE(e && handles, KX_EINVAL, "kx_engine_p2p_export: NULL argument");
    KX_REQUIRE(e->d.tp_size == 2, KX_EINVAL, "kx_engine_p2p_export: the fused all-reduce is built for tp_size 2");
    if (!e->p2p_recv) {
        e->p2p_rows = e->d.max_batch < 256 ? e->d.max_batch : 255;
        const size_t bytes = (size_t)2 * e->p2p_rows * e->d.hidden * sizeof(float);
        KX_TRY(dev_alloc(e, (void **)&e->p2p_recv, bytes));
        KX_TRY(dev_alloc(e, (void **)&e->p2p_flag, kP2pSlots * 4));
        KX_TRY(dev_alloc(e, (void **)&e->p2p_epoch, kP2pSlots * 4));
        KX_CUDA(cudaMemset(e->p2p_flag, 0, kP2pSlots * 4));
        KX_CUDA(cudaMemset(e->p2p_epoch, 0, kP2pSlots * 4));
    }
    cudaIpcMemHandle_t *h = reinterpret_cast<cudaIpcMemHandle_t *>(handles);
    KX_CUDA(cudaIpcGetMemHandle(&h[0], e->p2p_recv));
    KX_CUDA(cudaIpcGetMemHandle(&h[1], e->p2p_flag));
    return KX_OK;
}

extern "C" int64_t kx_engine_p2p_launches(const kx_engine *e) { return e ? e->p2p_launches : 0; }

extern "C" int kx_engine_p2p_import(kx_engine *e, const void *peer_handles /* 2 x 64 bytes */) {
    KX_REQUIRE(e && peer_handles && e->p2p_recv, KX_EINVAL, "kx_engine_p2p_import: export first");
    drop_step_graphs(e);
    const cudaIpcMemHandle_t *h = reinterpret_cast<const cudaIpcMemHandle_t *>(peer_handles);
    KX_CUDA(cudaIpcOpenMemHandle((void **)&e->p2p_peer_recv, h[0], cudaIpcMemLazyEnablePeerAccess));
    KX_CUDA(cudaIpcOpenMemHandle((void **)&e->p2p_peer_flag, h[1], cudaIpcMemLazyEnablePeerAccess));
    return KX_OK;
}

extern "C" int kx_engine_set_collectives(kx_engine *e, void *comm, void *allreduce, void *allgather) {
    KX_REQUIRE(e && comm && allreduce && allgather, KX_EINVAL, "kx_engine_set_collectives: NULL argument");
    drop_step_graphs(e);
    e->comm = comm;
    e->allreduce = reinterpret_cast<decltype(e->allreduce)>(allreduce);
    e->allgather = reinterpret_cast<decltype(e->allgather)>(allgather);
    return KX_OK;
}

extern "C" int kx_engine_set_eos(kx_engine *e, int32_t eos_id) {
    KX_REQUIRE(e, KX_EINVAL, "kx_engine_set_eos: NULL engine");
    KX_REQUIRE(eos_id < e->d.vocab * (e->d.tp_size > 1 ? e->d.tp_size : 1), KX_EINVAL, "kx_engine_set_eos: id %d is outside the vocabulary", eos_id);
    drop_step_graphs(e);                                       // the id is a kernel argument baked into captured steps
    e->eos_id = eos_id < 0 ? -1 : eos_id;
    return KX_OK;
}

extern "C" int kx_engine_set_prefix(kx_engine *e, const int32_t *h_tokens, int32_t n, void *stream) {
    KX_REQUIRE(e, KX_EINVAL, "kx_engine_set_prefix: NULL engine");
    KX_CUDA(cudaSetDevice(e->device));
    const kx_model_desc &d = e->d;
    KX_REQUIRE(n >= 0 && n % KX_KV_PAGE == 0, KX_EINVAL, "kx_engine_set_prefix: length %d is not a multiple of %d", n, KX_KV_PAGE);
    KX_REQUIRE(n < d.max_seq && n <= e->max_tokens, KX_ETOOBIG, "kx_engine_set_prefix: prefix of %d tokens is too long", n);
    KX_REQUIRE(n == 0 || h_tokens, KX_EINVAL, "kx_engine_set_prefix: NULL tokens");
    KX_REQUIRE(e->s_live == 0, KX_EINVAL, "kx_engine_set_prefix: %d session slots are live and attend to the current prefix", e->s_live);
    e->s_open = false;                               // free slots park behind the prefix: sessions start over
    drop_step_graphs(e);
    cudaStream_t st = e->stream;
    cudaEvent_t ev;
    KX_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    KX_CUDA(cudaEventRecord(ev, (cudaStream_t)stream));
    KX_CUDA(cudaStreamWaitEvent(st, ev, 0));
    cudaEventDestroy(ev);
    // block table: every sequence's first n/16 pages alias the prefix sequence's pages (row max_batch)
    const int pp = n / KX_KV_PAGE;
    std::vector<int> bt((size_t)(d.max_batch + 1) * e->max_pages);
    for (int b = 0; b <= d.max_batch; ++b)
        for (int i = 0; i < e->max_pages; ++i)
            bt[(size_t)b * e->max_pages + i] = (b < d.max_batch && i < pp) ? d.max_batch * e->max_pages + i : b * e->max_pages + i;
    KX_CUDA(cudaMemcpyAsync(e->block_table, bt.data(), bt.size() * 4, cudaMemcpyHostToDevice, st));
    KX_CUDA(cudaStreamSynchronize(st));
    e->prefix_len = 0;
    if (n == 0) return KX_OK;
    if (e->share_prefix && n >= 64) {                // partial-result buffers of the shared-prefix decode attention
        size_t need = 0;
        for (int b = 2; b <= d.max_batch; ++b) {
            const size_t r = (size_t)attn_prefix_parts(b, d.n_q, d.n_kv, n, nullptr) * b * d.n_q;
            if (r > need) need = r;
        }
        if (need > e->prefix_rows) {                 // grown, never shrunk; the old buffers stay owned until destroy
            KX_TRY(dev_alloc(e, (void **)&e->prefix_o, need * d.head_dim * sizeof(float)));
            KX_TRY(dev_alloc(e, (void **)&e->prefix_ml, need * 2 * sizeof(float)));
            e->prefix_rows = need;
        }
    }
    // prefill the prefix once, as the extra sequence `max_batch`
    std::vector<int> hpos(n), hseq(n, d.max_batch);
    for (int t = 0; t < n; ++t) hpos[t] = t;
    KX_CUDA(cudaMemcpyAsync(e->tokens, h_tokens, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    KX_CUDA(cudaMemcpyAsync(e->pos, hpos.data(), (size_t)n * 4, cudaMemcpyHostToDevice, st));
    KX_CUDA(cudaMemcpyAsync(e->seq, hseq.data(), (size_t)n * 4, cudaMemcpyHostToDevice, st));
    KX_CUDA(cudaStreamSynchronize(st));
    KX_TRY(op_embed(e->tokens, e->embed, e->resid, n, d.hidden, st));
    for (int li = 0; li < d.layers; ++li) KX_TRY(layer_forward(e, li, n, false, 1, n, d.max_batch, 0, st, 0));
    KX_CUDA(cudaStreamSynchronize(st));
    e->prefix_len = n;
    return KX_OK;
}

extern "C" void kx_engine_destroy(kx_engine *e) {
    if (!e) return;
    for (auto &g : e->step_graphs) cudaGraphExecDestroy(g.exec);
    e->step_graphs.clear();
    if (e->p2p_peer_recv) cudaIpcCloseMemHandle(e->p2p_peer_recv);
    if (e->p2p_peer_flag) cudaIpcCloseMemHandle(e->p2p_peer_flag);
    for (void *p : e->owned) cudaFree(p);
    if (e->stream) cudaStreamDestroy(e->stream);
    if (e->h_done) cudaFreeHost(e->h_done);
    delete e->trace;
    delete e;
}

extern "C" int64_t kx_engine_device_bytes(const kx_engine *e) { return e ? e->device_bytes : 0; }

extern "C" int kx_engine_trace(kx_engine *e, uint64_t *h_slots, int32_t max_slots) {
    KX_REQUIRE(e && h_slots, KX_EINVAL, "kx_engine_trace: NULL argument");
    if (!e->trace) return 0;
    const int n = max_slots < e->trace->n ? max_slots : e->trace->n;
    KX_CUDA(cudaStreamSynchronize(e->stream));
    KX_CUDA(cudaMemcpy(h_slots, e->trace->buf, sizeof(uint64_t) * 8 * n, cudaMemcpyDeviceToHost));
    return n;
}

extern "C" const char *kx_engine_trace_name(const kx_engine *e, int32_t i) {
    return (e && e->trace && i >= 0 && i < e->trace->n) ? e->trace->names[i] : "";
}

extern "C" int64_t kx_engine_attn_tc_launches(const kx_engine *e) { return e ? e->attn_tc_launches : 0; }

extern "C" int kx_engine_generate(kx_engine *e, const int32_t *h_tokens, int32_t B, int32_t S, int32_t n_new,
                                  int32_t *h_out, float *d_logits, kx_gen_stats *stats, void *stream) {
    KX_REQUIRE(e && h_tokens && h_out, KX_EINVAL, "kx_engine_generate: NULL argument");
    KX_CUDA(cudaSetDevice(e->device));               // a host thread serving several engines may have another device current
    const kx_model_desc &d = e->d;
    KX_REQUIRE(B >= 1 && B <= d.max_batch, KX_ETOOBIG, "kx_engine_generate: batch %d exceeds max_batch %d", B, d.max_batch);
    KX_REQUIRE(e->s_live == 0, KX_EINVAL, "kx_engine_generate: %d session slots are live (kx_engine_session_release)", e->s_live);
    e->s_open = false;                               // a static batch overwrites the slots' state: sessions start over
    const int P = e->prefix_len;
    KX_REQUIRE(S >= 1 && n_new >= 1 && P + S + n_new <= d.max_seq, KX_ETOOBIG,
               "kx_engine_generate: %d cached + %d prompt + %d new tokens exceed max_seq %d", P, S, n_new, d.max_seq);
    // Work runs on the engine's own stream (a legacy default stream cannot be captured into a
    // graph); it is ordered after whatever the caller already enqueued on `stream`.
    cudaStream_t st = e->stream;
    const int64_t launches0 = kx_launch_count();
    int64_t h2d = 0, d2h = 0;                        // counted where the copies are issued
    cudaEvent_t ev0, ev1, ev2;
    KX_CUDA(cudaEventCreate(&ev0)); KX_CUDA(cudaEventCreate(&ev1)); KX_CUDA(cudaEventCreate(&ev2));
    KX_CUDA(cudaEventRecord(ev0, (cudaStream_t)stream));
    KX_CUDA(cudaStreamWaitEvent(st, ev0, 0));
    KX_CUDA(cudaMemsetAsync(e->step, 0, 4, st));
    KX_CUDA(cudaEventRecord(ev0, st));

    KX_TRY(prefill_sequences(e, h_tokens, 0, B, S, st, &h2d));
    if (d_logits) KX_CUDA(cudaMemcpyAsync(d_logits, e->logits, sizeof(float) * (size_t)B * d.vocab, cudaMemcpyDeviceToDevice, st));
    KX_TRY(greedy_pick(e, B, st));
    store_token_kernel<<<(B + 127) / 128, 128, 0, st>>>(e->next_tok, e->out_ids, e->step, B, n_new);
    count_launch();
    KX_CUDA(cudaEventRecord(ev1, st));

    // ---- decode: positions S, S+1, ... one token per sequence ----
    {
        std::vector<int> p0(B, P + S - 1), sq(B);      // advance_step_kernel moves to P + S before the first use
        for (int b = 0; b < B; ++b) sq[b] = b;
        KX_CUDA(cudaMemcpyAsync(e->pos, p0.data(), (size_t)B * 4, cudaMemcpyHostToDevice, st));
        KX_CUDA(cudaMemcpyAsync(e->seq, sq.data(), (size_t)B * 4, cudaMemcpyHostToDevice, st));
        h2d += (int64_t)2 * B * 4;
        KX_CUDA(cudaStreamSynchronize(st));
    }
    const int max_T = P + S + n_new;
    const bool fused = fused_norm_decode(e, B);
    auto step_once = [&](float *logits_dst) -> int {
        if (e->trace) {                                // timeline of this step: slots handed out from 0 again
            e->trace->n = 0;
            trace_reset_kernel<<<(e->trace->cap + 255) / 256, 256, 0, st>>>(e->trace->buf, e->trace->cap);
            t_trace = e->trace;
        }
        struct Untrace { ~Untrace() { t_trace = nullptr; } } untrace;
        if (fused) {
            // the previous tail kernel (or `begin` below) left this step's input in resid / xn / ssq and advanced pos
            for (int li = 0; li < d.layers; ++li) KX_TRY(layer_forward_fused(e, li, B, max_T, st));
            KX_TRY(head_forward(e, B, nullptr, n_new, logits_dst, st, true));
            return op_decode_tail(e->next_tok, e->out_ids, n_new, P + S - 1, e->pos, e->embed, e->layers[0].w.ln1, e->resid, e->xn,
                                  e->ssq, d.max_batch, B, d.hidden, 1, e->eos_id, e->done, st);
        }
        // pos[b] currently holds the previous token's position
        advance_step_kernel<<<(B + 127) / 128, 128, 0, st>>>(e->pos, e->step, B);
        count_launch();
        KX_TRY(op_embed(e->next_tok, e->embed, e->resid, B, d.hidden, st));
        for (int li = 0; li < d.layers; ++li) KX_TRY(layer_forward(e, li, B, true, B, 1, 0, max_T, st));
        KX_TRY(head_forward(e, B, nullptr, n_new, logits_dst, st));
        return KX_OK;
    };
    // End of sequence (kx_engine_set_eos, deferred-norm path): the tail kernel counts the sequences that have emitted
    // eos; every 8 replayed steps the host reads the count and stops launching once all have.  (Sequences that
    // finished keep decoding until the last one has.)
    const bool watch_eos = fused && e->eos_id >= 0;
    if (watch_eos) KX_CUDA(cudaMemsetAsync(e->done, 0, sizeof(int) * (size_t)(1 + B), st));
    if (fused && n_new > 1)      // first decode step's input: embedding of the token prefill chose, pos -> P + S
        KX_TRY(op_decode_tail(e->next_tok, e->out_ids, n_new, P + S - 1, e->pos, e->embed, e->layers[0].w.ln1, e->resid, e->xn,
                              e->ssq, d.max_batch, B, d.hidden, 0, e->eos_id, e->done, st));
    auto all_done = [&](int t) -> bool {
        if (!watch_eos || (t & 7) != 0) return false;
        if (cudaMemcpyAsync(e->h_done, e->done, sizeof(int), cudaMemcpyDeviceToHost, st) != cudaSuccess) return false;
        if (cudaStreamSynchronize(st) != cudaSuccess) return false;
        return *e->h_done >= B;
    };
    int steps_done = 0;
    if (n_new > 1 && d_logits) {
        for (int t = 1; t < n_new; ++t) KX_TRY(step_once(d_logits + (size_t)t * B * d.vocab));
        steps_done = n_new - 1;
    } else if (n_new > 1) {
        kx_engine::StepGraph *hit = nullptr;
        for (auto &g : e->step_graphs)
            if (g.B == B && g.S == S && g.n_new == n_new && g.P == P) hit = &g;
        steps_done = n_new - 1;
        if (hit) {                                             // every step replays the cached graph
            int t = 1;
            for (; t < n_new; ++t) {
                KX_CUDA(cudaGraphLaunch(hit->exec, st));
                if (all_done(t)) { steps_done = t; break; }
            }
            count_launch((int)(hit->launches * steps_done));
        } else {
            KX_TRY(step_once(nullptr));                        // eager: also sets kernel attributes outside any capture
            if (n_new > 2) {
                cudaGraph_t graph = nullptr;
                cudaGraphExec_t exec = nullptr;
                const int64_t l0 = kx_launch_count();
                KX_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
                int rc = step_once(nullptr);
                cudaError_t ce = cudaStreamEndCapture(st, &graph);
                if (rc != KX_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
                KX_CUDA(ce);
                const int64_t per_step = kx_launch_count() - l0;
                KX_CUDA(cudaGraphInstantiate(&exec, graph, 0));
                cudaGraphDestroy(graph);
                for (int t = 2; t < n_new; ++t) {
                    KX_CUDA(cudaGraphLaunch(exec, st));
                    if (all_done(t)) { steps_done = t; break; }
                }
                count_launch((int)(per_step * (steps_done - 2)));  // the capture itself was counted once
                if (e->step_graphs.size() >= 8) {              // small cache, oldest entry out
                    cudaGraphExecDestroy(e->step_graphs.front().exec);
                    e->step_graphs.erase(e->step_graphs.begin());
                }
                e->step_graphs.push_back({B, S, n_new, P, exec, per_step});
            }
        }
    }
    KX_CUDA(cudaEventRecord(ev2, st));
    KX_CUDA(cudaMemcpyAsync(h_out, e->out_ids, (size_t)B * n_new * 4, cudaMemcpyDeviceToHost, st));
    d2h += (int64_t)B * n_new * 4;
    KX_CUDA(cudaStreamSynchronize(st));
    if (watch_eos && steps_done < n_new - 1)                   // steps that never ran: every sequence had already ended
        for (int b = 0; b < B; ++b)
            for (int t = steps_done + 1; t < n_new; ++t) h_out[(size_t)b * n_new + t] = e->eos_id;
    if (stats) {
        cudaEventElapsedTime(&stats->prefill_ms, ev0, ev1);
        cudaEventElapsedTime(&stats->decode_ms, ev1, ev2);
        stats->decode_steps = steps_done;
        stats->launches = (int32_t)(kx_launch_count() - launches0);
        stats->h2d_bytes = h2d; stats->d2h_bytes = d2h;
    }
    cudaEventDestroy(ev0); cudaEventDestroy(ev1); cudaEventDestroy(ev2);
    return KX_OK;
}

// --------------------------------------------------------------------------- //
// Continuous batching.  The reference serves requests as they arrive, each one awaiting its own completion
// (app.py:183-186), and cancels the one whose deadline passes (app.py:189-191).  A static batch makes a request that
// arrives mid-batch wait for every remaining step of that batch; here the engine's max_batch sequence SLOTS are
// admitted, stepped and released independently, so the host scheduler (engine.py) can admit at any step boundary
// and free a cancelled or finished request's slot (its KV pages with it) at the next one.
//
// State per slot lives on the device (s_cnt / s_limit / s_fin, e->pos, e->next_tok); a decode step is the same launch
// sequence as kx_engine_generate's with session_tail_kernel keeping the books, replayed as a CUDA graph per batch
// bucket.  Single-GPU replicas only (a tensor-parallel pair would have to take the same decisions on both ranks).
// --------------------------------------------------------------------------- //
namespace {

int session_bucket(const kx_engine *e, int n_slots) {   // decode batch for slots [0, n_slots): whole batch tiles
    // 32 columns, then multiples of 64 (a ragged second tile costs more than a full one: batch 96 1.50 ms, 128 1.46)
    int B = n_slots <= 32 ? 32 : (n_slots + 63) / 64 * 64;
    return B > e->d.max_batch ? e->d.max_batch : B;
}

int session_step_once(kx_engine *e, int B, cudaStream_t st) {
    const kx_model_desc &d = e->d;
    if (fused_norm_decode(e, B)) {
        for (int li = 0; li < d.layers; ++li) KX_TRY(layer_forward_fused(e, li, B, d.max_seq, st));
        KX_TRY(head_forward(e, B, nullptr, 0, nullptr, st, true));
        return op_session_tail(e->next_tok, e->out_ids, d.max_seq, e->pos, e->s_cnt, e->s_limit, e->s_fin, e->eos_id, 1, e->embed,
                               e->layers[0].w.ln1, e->resid, e->xn, e->ssq, d.max_batch, B, d.hidden, st);
    }
    KX_TRY(op_embed(e->next_tok, e->embed, e->resid, B, d.hidden, st));
    for (int li = 0; li < d.layers; ++li) KX_TRY(layer_forward(e, li, B, true, B, 1, 0, d.max_seq, st));
    KX_TRY(head_forward(e, B, nullptr, 0, nullptr, st, false, false));
    return op_session_tail(e->next_tok, e->out_ids, d.max_seq, e->pos, e->s_cnt, e->s_limit, e->s_fin, e->eos_id, 1, nullptr, nullptr,
                           nullptr, nullptr, nullptr, 0, B, d.hidden, st);
}

}  // namespace

extern "C" int kx_engine_session_open(kx_engine *e) {
    KX_REQUIRE(e, KX_EINVAL, "kx_engine_session_open: NULL argument");
    KX_REQUIRE(e->d.tp_size <= 1, KX_EINVAL, "kx_engine_session_open: continuous batching runs on single-GPU replicas only");
    KX_CUDA(cudaSetDevice(e->device));
    const kx_model_desc &d = e->d;
    const size_t n = (size_t)d.max_batch;
    if (!e->s_cnt) {
        int *base = nullptr;
        KX_TRY(dev_alloc(e, (void **)&base, 6 * n * sizeof(int)));
        e->s_cnt = base; e->s_limit = base + n; e->s_fin = base + 2 * n; e->s_pos = base + 3 * n; e->s_tok = base + 4 * n;
        e->s_seq = base + 5 * n;
        std::vector<int> ident(n);
        for (size_t b = 0; b < n; ++b) ident[b] = (int)b;
        KX_CUDA(cudaMemcpy(e->s_seq, ident.data(), n * sizeof(int), cudaMemcpyHostToDevice));
    }
    cudaStream_t st = e->stream;
    KX_CUDA(cudaMemsetAsync(e->next_tok, 0, n * sizeof(int), st));
    KX_CUDA(cudaMemsetAsync(e->s_limit, 0, n * sizeof(int), st));
    KX_CUDA(cudaMemcpyAsync(e->seq, e->s_seq, n * sizeof(int), cudaMemcpyDeviceToDevice, st));
    KX_TRY(op_session_release(e->pos, e->s_cnt, e->s_fin, 0, d.max_batch, e->prefix_len, st));
    KX_CUDA(cudaStreamSynchronize(st));
    e->s_open = true; e->s_need_input = true; e->s_live = 0;
    return KX_OK;
}

extern "C" int kx_engine_session_admit(kx_engine *e, int32_t slot0, int32_t n, const int32_t *h_tokens, int32_t S, int32_t max_new,
                                       kx_gen_stats *stats) {
    KX_REQUIRE(e && h_tokens, KX_EINVAL, "kx_engine_session_admit: NULL argument");
    KX_REQUIRE(e->s_open, KX_EINVAL, "kx_engine_session_admit: no open session (kx_engine_session_open)");
    KX_CUDA(cudaSetDevice(e->device));
    const kx_model_desc &d = e->d;
    const int P = e->prefix_len;
    KX_REQUIRE(slot0 >= 0 && n >= 1 && slot0 + n <= d.max_batch, KX_ETOOBIG, "kx_engine_session_admit: slots %d .. %d outside 0 .. %d",
               slot0, slot0 + n - 1, d.max_batch - 1);
    KX_REQUIRE(S >= 1 && max_new >= 1 && P + S + max_new <= d.max_seq, KX_ETOOBIG,
               "kx_engine_session_admit: %d cached + %d prompt + %d new tokens exceed max_seq %d", P, S, max_new, d.max_seq);
    cudaStream_t st = e->stream;
    const int64_t launches0 = kx_launch_count();
    const size_t nb = sizeof(int) * (size_t)d.max_batch;
    int64_t h2d = 0;
    cudaEvent_t ev0, ev1;
    KX_CUDA(cudaEventCreate(&ev0)); KX_CUDA(cudaEventCreate(&ev1));
    KX_CUDA(cudaEventRecord(ev0, st));
    // the prefill uses pos / seq / next_tok as per-token scratch: the live slots' values wait in s_pos / s_tok
    KX_CUDA(cudaMemcpyAsync(e->s_pos, e->pos, nb, cudaMemcpyDeviceToDevice, st));
    KX_CUDA(cudaMemcpyAsync(e->s_tok, e->next_tok, nb, cudaMemcpyDeviceToDevice, st));
    int rc = prefill_sequences(e, h_tokens, slot0, n, S, st, &h2d);
    if (rc == KX_OK) rc = op_argmax(e->logits + (size_t)slot0 * d.vocab, n, d.vocab, d.vocab, e->s_tok + slot0, nullptr, st);
    if (rc == KX_OK)
        rc = op_session_admit(e->s_tok, e->out_ids, d.max_seq, e->s_pos, e->s_cnt, e->s_limit, e->s_fin, slot0, n, P + S, max_new,
                              e->eos_id, st);
    // (also after a failure: the live slots get their state back)
    cudaMemcpyAsync(e->pos, e->s_pos, nb, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(e->next_tok, e->s_tok, nb, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(e->seq, e->s_seq, nb, cudaMemcpyDeviceToDevice, st);
    cudaEventRecord(ev1, st);
    const cudaError_t ce = cudaStreamSynchronize(st);
    e->s_need_input = true;
    if (rc == KX_OK && ce == cudaSuccess) e->s_live += n;
    if (stats && rc == KX_OK && ce == cudaSuccess) {
        cudaEventElapsedTime(&stats->prefill_ms, ev0, ev1);
        stats->decode_ms = 0.f; stats->decode_steps = 0;
        stats->launches = (int32_t)(kx_launch_count() - launches0);
        stats->h2d_bytes = h2d; stats->d2h_bytes = 0;
    }
    cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    if (rc != KX_OK) return rc;
    KX_CUDA(ce);
    return KX_OK;
}

extern "C" int kx_engine_session_step(kx_engine *e, int32_t n_steps, int32_t n_slots) {
    KX_REQUIRE(e && e->s_open, KX_EINVAL, "kx_engine_session_step: no open session");
    KX_REQUIRE(n_steps >= 1 && n_slots >= 1 && n_slots <= e->d.max_batch, KX_EINVAL, "kx_engine_session_step: bad argument");
    KX_CUDA(cudaSetDevice(e->device));
    const kx_model_desc &d = e->d;
    cudaStream_t st = e->stream;
    const int B = session_bucket(e, n_slots), P = e->prefix_len;
    if (fused_norm_decode(e, B) && e->s_need_input)     // an admission (or a wider bucket) left no step input behind
        KX_TRY(op_session_tail(e->next_tok, e->out_ids, d.max_seq, e->pos, e->s_cnt, e->s_limit, e->s_fin, e->eos_id, 0, e->embed,
                               e->layers[0].w.ln1, e->resid, e->xn, e->ssq, d.max_batch, d.max_batch, d.hidden, st));
    e->s_need_input = !fused_norm_decode(e, B);         // the classic path leaves none either
    kx_engine::StepGraph *hit = nullptr;
    for (auto &g : e->step_graphs)
        if (g.B == B && g.S == -1 && g.n_new == -1 && g.P == P) hit = &g;
    int t = 0;
    if (!hit) {
        KX_TRY(session_step_once(e, B, st));               // eager: also sets kernel attributes outside any capture
        t = 1;
        if (n_steps > 1) {
            cudaGraph_t graph = nullptr;
            cudaGraphExec_t exec = nullptr;
            const int64_t l0 = kx_launch_count();
            KX_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            const int rc = session_step_once(e, B, st);
            const cudaError_t ce = cudaStreamEndCapture(st, &graph);
            if (rc != KX_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
            KX_CUDA(ce);
            const int64_t per_step = kx_launch_count() - l0;
            KX_CUDA(cudaGraphInstantiate(&exec, graph, 0));
            cudaGraphDestroy(graph);
            count_launch(-(int)per_step);                  // the capture launched nothing
            if (e->step_graphs.size() >= 8) {
                cudaGraphExecDestroy(e->step_graphs.front().exec);
                e->step_graphs.erase(e->step_graphs.begin());
            }
            e->step_graphs.push_back({B, -1, -1, P, exec, per_step});
            hit = &e->step_graphs.back();
        }
    }
    for (; t < n_steps; ++t) {
        KX_CUDA(cudaGraphLaunch(hit->exec, st));
        count_launch((int)hit->launches);
    }
    return KX_OK;
}

extern "C" int kx_engine_session_poll(kx_engine *e, int32_t *h_count, int32_t *h_finished) {
    KX_REQUIRE(e && e->s_open && h_count && h_finished, KX_EINVAL, "kx_engine_session_poll: no open session or NULL argument");
    KX_CUDA(cudaSetDevice(e->device));
    const size_t nb = sizeof(int) * (size_t)e->d.max_batch;
    KX_CUDA(cudaMemcpyAsync(h_count, e->s_cnt, nb, cudaMemcpyDeviceToHost, e->stream));
    KX_CUDA(cudaMemcpyAsync(h_finished, e->s_fin, nb, cudaMemcpyDeviceToHost, e->stream));
    KX_CUDA(cudaStreamSynchronize(e->stream));
    return KX_OK;
}

extern "C" int kx_engine_session_read(kx_engine *e, int32_t slot0, int32_t n_slots, int32_t *h_out, int32_t n) {
    KX_REQUIRE(e && e->s_open && h_out, KX_EINVAL, "kx_engine_session_read: no open session or NULL argument");
    KX_REQUIRE(slot0 >= 0 && n_slots >= 1 && slot0 + n_slots <= e->d.max_batch && n >= 1 && n <= e->d.max_seq, KX_EINVAL,
               "kx_engine_session_read: bad slots or count");
    KX_CUDA(cudaSetDevice(e->device));
    KX_CUDA(cudaMemcpy2DAsync(h_out, sizeof(int) * (size_t)n, e->out_ids + (size_t)slot0 * e->d.max_seq, sizeof(int) * (size_t)e->d.max_seq,
                              sizeof(int) * (size_t)n, (size_t)n_slots, cudaMemcpyDeviceToHost, e->stream));
    KX_CUDA(cudaStreamSynchronize(e->stream));
    return KX_OK;
}

extern "C" int kx_engine_session_release(kx_engine *e, int32_t slot0, int32_t n) {
    KX_REQUIRE(e && e->s_open, KX_EINVAL, "kx_engine_session_release: no open session");
    KX_REQUIRE(slot0 >= 0 && n >= 1 && slot0 + n <= e->d.max_batch, KX_EINVAL, "kx_engine_session_release: bad slots");
    KX_CUDA(cudaSetDevice(e->device));
    KX_TRY(op_session_release(e->pos, e->s_cnt, e->s_fin, slot0, n, e->prefix_len, e->stream));
    e->s_live -= n;
    if (e->s_live < 0) e->s_live = 0;
    return KX_OK;
}

extern "C" int32_t kx_engine_session_live(const kx_engine *e) { return (e && e->s_open) ? e->s_live : 0; }
