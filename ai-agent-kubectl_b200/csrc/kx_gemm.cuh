// Internal interface of the tcgen05 GEMM (kx_gemm.cu) used by the engine.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace kx {

// EPI_ARGMAX (lm_head, k8): no [tokens, vocabulary] logits are written; every CTA reduces its tile to one
// (max logit, lowest feature index attaining it) pair per token and part: out = int2 [parts][ldo] with the
// value's bits in .x.  Token-major pair kernel: parts = feature tiles.  Swap-AB kernel: parts = 4 per 128-feature
// tile (one per epilogue warp).  op_argmax_parts finishes the reduction.
enum GemmEpilogue { EPI_BF16 = 0, EPI_RESID = 1, EPI_SWIGLU = 2, EPI_F32 = 3, EPI_QKV_ROPE = 4, EPI_ARGMAX = 5 };

struct GemmParams {
    int a_rows, b_rows, k;   // D[a_rows, b_rows] = A[a_rows, k] * B[b_rows, k]^T
    void *out;
    int ldo;                 // elements between consecutive output rows
    const float *bias;       // per output feature, or NULL
    int swap;                // 0: rows of A are output rows; 1: rows of A are features, out[b * ldo + feature]
    int epilogue;
    int splits;              // split-K factor requested (clamped to the number of 64-wide k blocks)
    int kb_per_split;        // filled by gemm_launch
    size_t split_stride;     // token-major pair kernel with splits > 1 (EPI_F32 only): fp32 elements between the splits' partial outputs
    float *ws;               // split-K scratch, filled by gemm_launch
    uint32_t *tickets;
    // EPI_QKV_ROPE (token-major pair kernel only): bias, bf16 rounding, RoPE on q and k, q -> out,
    // k and v -> the paged KV cache; replaces the separate rope/append kernel in prefill
    const int *pos, *seq;    // per token row: absolute position, sequence (block-table row)
    const void *rope;        // float2 [max_pos][hd/2]
    void *kv_pages;          // this layer's pages
    const int *block_table;
    int max_pages, n_q, n_kv, hd;
    // Tensor-parallel residual epilogue fused with the all-reduce (decode, 2 ranks, swap-AB cluster kernel):
    // each CTA pushes its slice of this rank's partial sum straight into the PEER GPU's receive buffer over
    // NVLink, publishes an epoch flag there, waits for the peer CTA of the same tile to do the same here, and
    // adds both partials to the residual.  Buffers are double-buffered by epoch parity.
    float *tp_peer_recv;     // peer GPU:  float [2][tp_rows * ldo]   (written by this rank)
    uint32_t *tp_peer_flag;  // peer GPU:  uint32 [tiles]             (epoch published by this rank)
    const float *tp_recv;    // this GPU:  the same two arrays, written by the peer
    const uint32_t *tp_flag;
    uint32_t *tp_epoch;      // this GPU:  uint32 [tiles], launches seen so far per tile slot
    int tp_rows;             // rows of one receive buffer (max decode batch)
    unsigned long long *trace;   // diagnostic timeline slot (kx_common.cuh) or NULL
};

int make_tmap_bf16_2d(CUtensorMap *out, const void *ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows);
int make_tmap_f32_box32(CUtensorMap *out, const void *ptr, int64_t rows, int64_t cols, int64_t ld);   // residual rows, TMA reduce-add
int require_local_device_memory(const void *ptr, const char *what);   // operands that are loaded with TMA multicast
int gemm_pick_bn(int swap, int64_t b_rows);
int gemm_pair_bn(int64_t a_rows, int64_t b_rows);   // 0 = use the single-CTA kernel
int gemm_launch_pair(const CUtensorMap &ta, const CUtensorMap &tb_half, const GemmParams &p, int bn, cudaStream_t st);
size_t gemm_scratch_bytes(const GemmParams &p, int bn);
int gemm_launch(const CUtensorMap &ta, const CUtensorMap &tb, GemmParams p, int bn, void *scratch, cudaStream_t st);

}  // namespace kx
