// Internal interface of the persistent decode kernel (kx_decode_mega.cu).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace kx {

// one linear layer of the decode step as a set of (weight tile, k slice) work items
struct MegaGemm {
    int feats;      // output features (rows of the weight matrix)
    int tiles;      // 128-row weight tiles
    int splits;     // k slices per tile
    int kb_total;   // 64-wide k blocks
    int kb_per;     // k blocks per slice
    int items;      // tiles * splits
};

struct MegaLayer {
    const void *ln1, *ln2;   // RMSNorm weights (bf16)
    const float *bqkv;       // QKV bias or NULL
    void *kv_pages;
};

struct MegaParams {
    const CUtensorMap *wmaps;        // device array [layers][4]: wqkv, wo, wgu, wd (128-row boxes)
    const MegaLayer *layers_ptr;     // device array [layers]
    MegaGemm g[4];
    float *resid;
    void *xn, *attn, *mlp;           // bf16 activations (also reachable through the three tensor maps)
    float *part_qkv, *part_o, *part_d;   // split-K partial sums [splits][B][feats]
    const int *pos, *block_table;
    const float2 *rope;
    uint32_t *bar;                   // grid barrier: one arrival counter per phase [7 * layers], zeroed before every launch
    int layers, B, hidden, qkv_dim, q_dim, n_q, n_kv, max_pages;
    float eps, scale_log2;
    uint64_t *trace;                 // diagnostic (tag, globaltimer ns) pairs written by CTA 0, or NULL
    int trace_cap;
};

bool mega_supported(int hd, int G, int hidden);
int mega_bn(int B);                                  // batch tile (64 / 128), 0 = batch too large for this kernel
MegaGemm mega_plan_gemm(int feats, int k, int grid, bool may_split);
int mega_launch(const CUtensorMap &xn, const CUtensorMap &at, const CUtensorMap &ml, const MegaParams &p, int bn, int hd, int grid,
                cudaStream_t st);

}  // namespace kx
