// Internal interface of the non-GEMM generator kernels (kx_ops.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define KX_KV_PAGE 16   // tokens per KV-cache page

namespace kx {

int op_embed(const int *tokens, const void *table, float *x, int M, int H, cudaStream_t st);
// part / splits / split_stride: split-K partial sums of the preceding projection still to be added to the residual
// rows (x is updated in place before it is normalised); splits = 0: plain RMSNorm
int op_rmsnorm(float *x, const void *g, void *y, int M, int H, float eps, const int *row_map, cudaStream_t st,
               const float *part = nullptr, int splits = 0, size_t split_stride = 0);
int op_rope_append(void *qkv, int qkv_ld, const int *pos, const int *seq, const float *rope, void *pages,
                   const int *block_table, int max_pages, int M, int n_q, int n_kv, int hd, cudaStream_t st);
// decode attention also applies RoPE to the new token's q/k and appends its k, v to the cache.
// key0 > 0: keys [0, key0) are a prefix shared by the batch, already reduced by op_attn_prefix to n_part partials.
int op_attn_decode(const void *qkv, int qkv_ld, void *out, int out_ld, void *pages, const int *block_table,
                   int max_pages, const int *pos, const float *rope, int B, int max_T, int n_q, int n_kv, int hd,
                   int key0, const float *part_o, const float *part_ml, int n_part, cudaStream_t st);
// shared-prefix part of decode attention: the prefix K/V are read once per 64 query rows, not once per sequence
int attn_prefix_parts(int B, int n_q, int n_kv, int P, int *chunk_keys);
int op_attn_prefix(const void *qkv, int qkv_ld, const void *pages, const int *bt_prefix, const int *pos, const float *rope,
                   int B, int n_q, int n_kv, int hd, int P, float *part_o, float *part_ml, int *n_part, cudaStream_t st);
int op_attn_prefill(const void *qkv, int qkv_ld, void *out, int out_ld, const void *pages, const int *block_table,
                    int max_pages, int B, int S, int pos0, int n_q, int n_kv, int hd, cudaStream_t st);
// tcgen05 version (kx_attn_tc.cu): head_dim 64
int op_attn_prefill_tc(const void *qkv, int qkv_ld, void *out, int out_ld, const void *pages, int num_pages,
                       const int *block_table, int max_pages, int B, int S, int pos0, int n_q, int n_kv, cudaStream_t st);
int op_argmax(const float *logits, int B, int V, int ld, int *out, int *out2, cudaStream_t st);
// second stage of the GEMM's EPI_ARGMAX epilogue: parts int2 [n_parts][ld] -> ids (out / out2) or, for a
// vocabulary shard, (value bits, id + id_offset) pairs (out_pairs)
int op_argmax_parts(const void *parts, int n_parts, int ld, int B, int id_offset, int *out, int *out2, void *out_pairs,
                    cudaStream_t st);
int op_advance(int *pos, int n, cudaStream_t st);
// deferred-RMSNorm decode path: end-of-step bookkeeping + embedding of the chosen token as the next step's input
int op_decode_tail(const int *next_tok, int *out_ids, int n_new, int base_pos, int *pos, const void *table, const void *g,
                   float *resid, void *xg, float *ssq, int ssq_ld, int B, int H, int store, int eos_id, int *done, cudaStream_t st);
int op_col_rinv(const float *ssq, int tiles, int ld, int B, int H, float eps, float *out, cudaStream_t st);
// continuous batching: per-slot bookkeeping of a decode step (+ the next step's input on the deferred-norm path when
// `table` is given), admission and release of slots
int op_session_tail(const int *next_tok, int *out, int out_ld, int *pos, int *cnt, const int *limit, int *fin, int eos_id, int book,
                    const void *table, const void *g, float *resid, void *xg, float *ssq, int ssq_ld, int B, int H, cudaStream_t st);
int op_session_admit(const int *tok, int *out, int out_ld, int *pos, int *cnt, int *limit, int *fin, int slot0, int n, int first_pos,
                     int max_new, int eos_id, cudaStream_t st);
int op_session_release(int *pos, int *cnt, int *fin, int slot0, int n, int park, cudaStream_t st);
// tensor-parallel helpers
int op_add_inplace(float *x, const float *d, size_t n, cudaStream_t st);
int op_argmax_pairs(const float *logits, int B, int V, int ld, int id_offset, void *pairs, cudaStream_t st);
int op_pick(const void *gathered, int tp, int B, int *out, cudaStream_t st);

}  // namespace kx
