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
//
// Deferred RMSNorm (decode, swap-AB cluster kernel): the row-parallel projections (wo, wd) own whole residual
// elements after their split-K reduction, so their epilogue (EPI_RESID_NORM) also prepares the NEXT norm:
// it writes bf16(resid * g) -- the next GEMM's input, not yet divided by the row's rms -- and, per 128-feature
// tile and batch column, the sum of squares of the updated residual (ssq_out[tile][column]).  The consuming GEMM
// (ssq_in set) adds the tiles' sums in a fixed order and multiplies its accumulator columns by
// 1 / sqrt(sum / hidden + eps) before bias / SwiGLU: W (x g / rms) == (W (x g)) / rms.  No RMSNorm launch.
enum GemmEpilogue { EPI_BF16 = 0, EPI_RESID = 1, EPI_SWIGLU = 2, EPI_F32 = 3, EPI_QKV_ROPE = 4, EPI_ARGMAX = 5,
                    EPI_RESID_NORM = 6 };

struct GemmParams {
    int a_rows, b_rows, k;   // D[a_rows, b_rows] = A[a_rows, k] * B[b_rows, k]^T
    void *out;
    int ldo;                 // elements between consecutive output rows
    const float *bias;       // per output feature, or NULL
    int swap;                // 0: rows of A are output rows; 1: rows of A are features, out[b * ldo + feature]
    int epilogue;
    int splits;              // split-K factor requested (clamped to the number of 64-wide k blocks)
    int kb_per_split;        // filled by gemm_launch
    int tile_band;           // pair kernel: feature tiles per band of its tile walk (filled by its launcher)
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
    // adds both partials to the residual.  The receive buffer has two halves used by alternate launches: a rank can only
    // reach launch n + 1 after the peer published launch n, i.e. after the peer finished reading launch n - 1's half.
    float *tp_peer_recv;     // peer GPU:  float [2][tp_rows * ldo]   (written by this rank)
    uint32_t *tp_peer_flag;  // peer GPU:  uint32 [tiles]             (epoch published by this rank)
    const float *tp_recv;    // this GPU:  the same two arrays, written by the peer
    const uint32_t *tp_flag;
    uint32_t *tp_epoch;      // this GPU:  uint32 [tiles], launches seen so far per tile slot
    int tp_rows;             // rows of one receive buffer (max decode batch)
    int tp_seq;              // sequence number of this fused launch (same on both ranks); its parity picks the receive half-buffer
    unsigned long long *trace;   // diagnostic timeline slot (kx_common.cuh) or NULL
    // deferred RMSNorm (see above)
    const void *norm_g;      // EPI_RESID_NORM: bf16 gain of the next norm [features]
    void *norm_out;          // EPI_RESID_NORM: bf16 [batch][ldo] = bf16(resid * g)
    float *ssq_out;          // EPI_RESID_NORM: fp32 [features / 128][ssq_ld]
    const float *ssq_in;     // consumer: fp32 [ssq_tiles][ssq_ld] (NULL: input is already normalised)
    int ssq_ld, ssq_tiles;
    float norm_eps, norm_inv_n;   // eps and 1 / hidden of the norm being applied
    const float *col_scale;  // single-CTA swap-AB kernel, EPI_F32: logits column b is multiplied by col_scale[b] (or NULL)
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
