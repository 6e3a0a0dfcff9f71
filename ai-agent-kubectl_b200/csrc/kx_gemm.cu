// k1 / k3: bf16 GEMM on the 5th-generation tensor cores.
//
//     D[Ma, Nb] = A[Ma, K] * B[Nb, K]^T      A, B bf16 with K contiguous, fp32 accumulate
//
// One kernel serves prefill and decode:
//   * prefill (swap = 0): A = activations [tokens, K], B = weights [features, K];
//     the accumulator row is a token, the epilogue writes rows of `out`.
//   * decode  (swap = 1): A = weights [features, K], B = activations [batch, K]
//     ("swap-AB": the 128-row MMA operand is the weight tile, the batch is the MMA
//     N), so every weight byte is streamed from HBM exactly once by a full-width
//     tcgen05.mma even when the batch is 64; split-K over gridDim.z gives enough CTAs
//     to pull HBM bandwidth when features/128 is small.
//
// Structure per CTA (192 threads): warp 0 = TMA producer (cp.async.bulk.tensor, 128 B
// swizzle, mbarrier full/empty ring), warp 1 = TMEM allocator + single-thread
// tcgen05.mma issuer (accumulator in TMEM, tcgen05.commit releases smem stages),
// warps 2-5 = epilogue (tcgen05.ld 32x32b -> registers -> fused epilogue -> global).
// Epilogues: bf16 (+bias) | fp32 residual accumulate | SwiGLU over interleaved
// (gate, up) features -> bf16 | fp32 (logits).
// Split-K: partial tiles are reduced with fp32 atomics into library scratch; the CTA
// that takes the last ticket of a tile applies the epilogue and re-zeroes the scratch
// (residual epilogue: atomics go straight into the residual stream, no scratch).
#include <map>
#include <mutex>
#include <tuple>
#include <cuda_bf16.h>
#include "kx_common.cuh"
#include "kx_sm100.cuh"
#include "kx_gemm.cuh"
#include "kx_ops.cuh"

namespace kx {

using namespace sm100;

namespace {

constexpr int BM = 128, BK = 64;
constexpr int kGemmThreads = 192;

template <int BN> struct Cfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    // 96 KB ring: two CTAs per SM.  (Three 72 KB CTAs per SM -- so that a launch and its early-started successor are
    // resident together -- measured slower: Llama-3.2-1B batch 64, 1.20 against 1.05 ms per decode step.)
    static constexpr int RING_KB = 96;
    static constexpr int STAGES = (RING_KB * 1024) / STAGE_BYTES;
    static constexpr int TMEM_COLS = BN < 32 ? 32 : BN;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

// x * sigmoid(x) with the hardware reciprocal (MUFU.RCP, 1 ulp, no slow path): a full fp32 division is ~40 dependent
// instructions plus a call for the denormal case, which the decode epilogues (one warp per scheduler) feel.  The
// result is rounded to bf16 right away, so the last fp32 bit does not matter.
__device__ __forceinline__ float silu(float x) { return __fdividef(x, 1.0f + __expf(-x)); }

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

// Fused epilogue for 32 consecutive accumulator columns [c0, c0+32) of accumulator row `r`.
template <int EPI, bool SWAP>
__device__ __forceinline__ void epilogue_chunk(const GemmParams &p, float (&v)[32], int arow, int bcol0, bool atomic_resid) {
    if (!SWAP) {
        // accumulator row = output row (token), columns = features
        if (arow >= p.a_rows) return;
        if (EPI == EPI_BF16) {
            __nv_bfloat16 *o = reinterpret_cast<__nv_bfloat16 *>(p.out) + (size_t)arow * p.ldo + bcol0;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
                if (bcol0 + j >= p.b_rows) break;
                float b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                if (p.bias) {
#pragma unroll
                    for (int t = 0; t < 8; ++t) b[t] = p.bias[bcol0 + j + t];
                }
                uint4 q;
                q.x = pack_bf16(v[j + 0] + b[0], v[j + 1] + b[1]);
                q.y = pack_bf16(v[j + 2] + b[2], v[j + 3] + b[3]);
                q.z = pack_bf16(v[j + 4] + b[4], v[j + 5] + b[5]);
                q.w = pack_bf16(v[j + 6] + b[6], v[j + 7] + b[7]);
                *reinterpret_cast<uint4 *>(o + j) = q;
            }
        } else if (EPI == EPI_RESID) {
            float *o = reinterpret_cast<float *>(p.out) + (size_t)arow * p.ldo + bcol0;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                if (bcol0 + j >= p.b_rows) break;
                if (atomic_resid) {
#pragma unroll
                    for (int t = 0; t < 4; ++t) atomicAdd(o + j + t, v[j + t]);
                } else {
                    float4 x = *reinterpret_cast<float4 *>(o + j);
                    x.x += v[j]; x.y += v[j + 1]; x.z += v[j + 2]; x.w += v[j + 3];
                    *reinterpret_cast<float4 *>(o + j) = x;
                }
            }
        } else if (EPI == EPI_SWIGLU) {
            __nv_bfloat16 *o = reinterpret_cast<__nv_bfloat16 *>(p.out) + (size_t)arow * p.ldo + (bcol0 >> 1);
#pragma unroll
            for (int j = 0; j < 32; j += 16) {
                if (bcol0 + j >= p.b_rows) break;
                uint4 q;
                q.x = pack_bf16(silu(v[j + 0]) * v[j + 1], silu(v[j + 2]) * v[j + 3]);
                q.y = pack_bf16(silu(v[j + 4]) * v[j + 5], silu(v[j + 6]) * v[j + 7]);
                q.z = pack_bf16(silu(v[j + 8]) * v[j + 9], silu(v[j + 10]) * v[j + 11]);
                q.w = pack_bf16(silu(v[j + 12]) * v[j + 13], silu(v[j + 14]) * v[j + 15]);
                *reinterpret_cast<uint4 *>(o + (j >> 1)) = q;
            }
        } else {
            float *o = reinterpret_cast<float *>(p.out) + (size_t)arow * p.ldo + bcol0;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                if (bcol0 + j >= p.b_rows) break;
                *reinterpret_cast<float4 *>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            }
        }
    } else {
        // accumulator row = feature, columns = batch entries: out[b * ldo + feature]
        const bool ok = arow < p.a_rows;
        const float bias = (ok && p.bias) ? p.bias[arow] : 0.0f;   // plain load: ld.global.nc may be hoisted above the null test
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const int b = bcol0 + j;
            if (b >= p.b_rows) break;                 // warp-uniform
            if (EPI == EPI_SWIGLU) {
                const float other = __shfl_xor_sync(0xffffffffu, v[j], 1);
                if (ok && !(arow & 1))
                    reinterpret_cast<__nv_bfloat16 *>(p.out)[(size_t)b * p.ldo + (arow >> 1)] = __float2bfloat16(silu(v[j]) * other);
            } else if (!ok) {
                continue;
            } else if (EPI == EPI_BF16) {
                reinterpret_cast<__nv_bfloat16 *>(p.out)[(size_t)b * p.ldo + arow] = __float2bfloat16(v[j] + bias);
            } else if (EPI == EPI_RESID) {
                float *o = reinterpret_cast<float *>(p.out) + (size_t)b * p.ldo + arow;
                if (atomic_resid) atomicAdd(o, v[j]); else *o += v[j];
            } else {
                reinterpret_cast<float *>(p.out)[(size_t)b * p.ldo + arow] = p.col_scale ? __fmul_rn(v[j], p.col_scale[b]) : v[j];
            }
        }
    }
}

template <int BN, int EPI, bool SWAP>
__global__ void __launch_bounds__(kGemmThreads, 2)
kx_gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
    using C = Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t *smem = smem_raw + (base - smem_u32(smem_raw));
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + C::STAGES * C::STAGE_BYTES);
    uint64_t *empty = full + C::STAGES;
    uint64_t *tmem_full = empty + C::STAGES;
    uint32_t *tmem_holder = reinterpret_cast<uint32_t *>(tmem_full + 1);
    uint32_t *flag = tmem_holder + 1;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM;
    const int total_kb = (p.k + BK - 1) / BK;
    const int kb0 = blockIdx.z * p.kb_per_split;
    const int nkb = min(total_kb - kb0, p.kb_per_split);

    if (threadIdx.x == 0) {
        trace_min(p.trace, 0);
        trace_min(p.trace, 1);
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < C::STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
        mbar_init(tmem_full, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_holder, C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_holder;

    if (warp == 0) {
        if (lane == 0) {
            // weights are streamed once (evict-first); activations are re-read by other CTAs (evict-last)
            const uint64_t pol_a = SWAP ? policy_evict_first() : policy_evict_last();
            const uint64_t pol_b = SWAP ? policy_evict_last() : policy_evict_first();
            for (int i = 0; i < nkb; ++i) {
                const int s = i % C::STAGES;
                const uint32_t ph = (i / C::STAGES) & 1;
                mbar_wait(empty + s, ph ^ 1);
                mbar_expect_tx(full + s, C::STAGE_BYTES);
                uint8_t *sa = smem + s * C::STAGE_BYTES;
                tma_load_2d_hint(sa, &tmA, full + s, (kb0 + i) * BK, m0, pol_a);
                tma_load_2d_hint(sa + C::A_BYTES, &tmB, full + s, (kb0 + i) * BK, n0, pol_b);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16_f32(BM, BN);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % C::STAGES;
                const uint32_t ph = (i / C::STAGES) & 1;
                mbar_wait(full + s, ph);
                tc_fence_after();
                const uint32_t sa = base + s * C::STAGE_BYTES;
                const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + C::A_BYTES);
#pragma unroll
                for (int k = 0; k < BK / 16; ++k)   // +32 B per 16-element K step inside the swizzle atom
                    umma_bf16(tmem, da + 2 * k, db + 2 * k, idesc, (i | k) != 0);
                umma_commit(empty + s);             // frees the smem stage once these MMAs retire
            }
            umma_commit(tmem_full);
        }
    } else {
        // ---- epilogue warps: TMEM lane quarter = warp % 4 ----
        const int q = warp & 3;
        const int arow = m0 + q * 32 + lane;
        const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16);
        const int etid = threadIdx.x - 64;
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        if (etid == 0) trace_max(p.trace, 2);
        const bool split = gridDim.z > 1;
        if (EPI == EPI_ARGMAX) {
            // swap-AB lm_head: rows = vocabulary entries, columns = tokens.  Per token the warp reduces its 32 rows
            // to (max, lowest index); lane j ends up holding the pair of token column j and writes it.
            // The 32 x 32 block (rows x token columns) goes through this warp's slice of the drained stage ring
            // (padded rows: conflict-free both ways); then lane j scans token column j over the 32 rows in order.
            int2 *parts = reinterpret_cast<int2 *>(p.out) + (size_t)(blockIdx.y * 4 + q) * p.ldo;
            float *tr = reinterpret_cast<float *>(smem) + q * (32 * 33);
            const int row0 = m0 + q * 32;
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                float v[32];
                tmem_ld_32x32(taddr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) tr[lane * 33 + j] = v[j];
                __syncwarp();
                float bv = -INFINITY;
                int bi = 0x7fffffff;
                const int rows = min(32, p.a_rows - row0);
                for (int r = 0; r < rows; ++r) {
                    const float x = tr[r * 33 + lane];
                    if (x > bv) { bv = x; bi = row0 + r; }        // ascending rows: the lowest index wins a tie
                }
                __syncwarp();
                const int tok = n0 + c + lane;
                if (tok < p.b_rows) parts[tok] = make_int2(__float_as_int(bv), bi);
            }
        } else if (!split || EPI == EPI_RESID) {
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                float v[32];
                tmem_ld_32x32(taddr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                tmem_ld_wait();
                epilogue_chunk<EPI == EPI_ARGMAX ? EPI_F32 : EPI, SWAP>(p, v, arow, n0 + c, split);
            }
        } else {
            // split-K: reduce into scratch laid out [tile][col][row] so a warp's atomics are contiguous
            const size_t tile = (size_t)blockIdx.y * gridDim.x + blockIdx.x;
            float *ws = p.ws + tile * (size_t)(BM * BN) + q * 32 + lane;
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                float v[32];
                tmem_ld_32x32(taddr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (n0 + c + j < p.b_rows) atomicAdd(ws + (size_t)(c + j) * BM, v[j]);
            }
            __threadfence();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (etid == 0) *flag = atomicAdd(p.tickets + tile, 1u) == gridDim.z - 1 ? 1u : 0u;
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (*flag) {
                __threadfence();
#pragma unroll 1
                for (int c = 0; c < BN; c += 32) {
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        v[j] = 0.f;
                        if (n0 + c + j < p.b_rows) { v[j] = __ldcg(ws + (size_t)(c + j) * BM); __stcg(ws + (size_t)(c + j) * BM, 0.f); }
                    }
                    epilogue_chunk<EPI == EPI_ARGMAX ? EPI_F32 : EPI, SWAP>(p, v, arow, n0 + c, false);
                }
                if (etid == 0) p.tickets[tile] = 0;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, C::TMEM_COLS);
    if (threadIdx.x == 0) trace_max(p.trace, 3);
}


// ---------------------------------------------------------------------------------
// Prefill GEMM, second generation: persistent, CTA pairs with TMA multicast.
//
// The 128x128 kernel above is L2-bandwidth bound: it pulls 32 KB of operands per
// 128x128x64 MMA block (128 B/cycle/SM at full tensor rate) while the L2 feeds about
// 45 B/cycle/SM.  Here a cluster of two CTAs computes a 256 x BN output tile: each CTA
// owns 128 rows (its own A tile) and the pair shares the B tile -- each CTA fetches half
// of it and TMA multicasts that half into both CTAs' shared memory.  Operand traffic
// per CTA drops to 16 KB (A) + BN/2 x 128 B per k-block.  The kernel is persistent (one
// cluster per SM pair walks the tile list) and the accumulator is double buffered in
// TMEM (2 x BN columns), so the epilogue of tile i overlaps the MMAs of tile i+1.
// ---------------------------------------------------------------------------------
// QKV epilogue with RoPE + KV append for one (row, head, 32-wide half-pair): x1 = columns d0..d0+31 of the
// head, x2 = columns d0+hd/2.. of the same head (rotate-half partners).  Values are rounded to bf16 after the
// bias (what the unfused path stores), rotated in fp32 without FMA contraction, rounded again.
__device__ __forceinline__ void qkv_rope_store(const GemmParams &p, float (&x1)[32], float (&x2)[32], int row, int head, int d0,
                                               int pos, int page) {   // position and KV page of the row, looked up by the caller
    const int half = p.hd >> 1;
    const int fcol = head * p.hd + d0;
    __nv_bfloat16 *dst1, *dst2;
    const bool is_v = head >= p.n_q + p.n_kv;
    if (head < p.n_q) {
        dst1 = reinterpret_cast<__nv_bfloat16 *>(p.out) + (size_t)row * p.ldo + fcol;
    } else {
        const int kvh = head - p.n_q - (is_v ? p.n_kv : 0);
        dst1 = reinterpret_cast<__nv_bfloat16 *>(p.kv_pages) +
               (((size_t)page * 2 + (is_v ? 1 : 0)) * p.n_kv + kvh) * KX_KV_PAGE * p.hd + (size_t)(pos % KX_KV_PAGE) * p.hd + d0;
    }
    dst2 = dst1 + half;
    const float2 *cs = reinterpret_cast<const float2 *>(p.rope) + (size_t)pos * half + d0;
    uint32_t o1[16], o2[16];
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
        float a[2], b[2];
        const float4 c2 = is_v ? make_float4(0.f, 0.f, 0.f, 0.f) : __ldg(reinterpret_cast<const float4 *>(cs + j));   // (cos, sin) of two dims
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            const float bias1 = p.bias ? p.bias[fcol + j + t] : 0.f, bias2 = p.bias ? p.bias[fcol + half + j + t] : 0.f;
            const float v1 = __bfloat162float(__float2bfloat16(x1[j + t] + bias1));
            const float v2 = __bfloat162float(__float2bfloat16(x2[j + t] + bias2));
            if (is_v) { a[t] = v1; b[t] = v2; }
            else {
                const float cx = t ? c2.z : c2.x, cy = t ? c2.w : c2.y;
                a[t] = __fsub_rn(__fmul_rn(v1, cx), __fmul_rn(v2, cy));
                b[t] = __fadd_rn(__fmul_rn(v2, cx), __fmul_rn(v1, cy));
            }
        }
        o1[j >> 1] = pack_bf16(a[0], a[1]);
        o2[j >> 1] = pack_bf16(b[0], b[1]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        reinterpret_cast<uint4 *>(dst1)[j] = make_uint4(o1[4 * j], o1[4 * j + 1], o1[4 * j + 2], o1[4 * j + 3]);
        reinterpret_cast<uint4 *>(dst2)[j] = make_uint4(o2[4 * j], o2[4 * j + 1], o2[4 * j + 2], o2[4 * j + 3]);
    }
}

template <int BN> struct Cfg2 {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = (192 * 1024) / STAGE_BYTES;
    static constexpr int TMEM_COLS = 2 * BN;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
};

// Tile walk of the persistent pair kernel.  Tiles are taken n-fastest inside a band of `band` feature tiles, band after
// band: the clusters that run together share one or two activation row blocks and a band of weight tiles, and the band
// (at most ~32 MB of weights) stays in the L2 while all row blocks pass over it.  With the whole of a 67 MB gate/up matrix
// as one band, each pass over the rows re-read an eighth of it from HBM (1.33 GB read for 201 MB of operands).
__device__ __forceinline__ void pair_tile_coords(int tile, int m_pairs, int n_tiles, int band, int &mp, int &nt) {
    const int per_band = m_pairs * band;
    const int b = tile / per_band, r = tile - b * per_band;
    const int w = min(band, n_tiles - b * band);
    mp = r / w;
    nt = b * band + (r - mp * w);
}

constexpr int kPairOutStage = 8 * 4096;   // residual epilogue staging (8 epilogue warps x 4 KB), EPI_RESID only
constexpr int kPairThreads = 320;   // warp 0 TMA, warp 1 MMA, warps 2-9 epilogue: two warps per TMEM lane quarter, half of the tile's columns each

template <int BN, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPairThreads, 1)
kx_gemm_bf16_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmBhalf,
                         const __grid_constant__ CUtensorMap tmOut, const GemmParams p, int m_pairs, int n_tiles) {
    using C = Cfg2<BN>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t *smem = smem_raw + (base - smem_u32(smem_raw));
    // residual epilogue: one 4 KB staging box (32 rows x 32 fp32, 128 B swizzle) per epilogue warp, after the ring
    constexpr int kOutStage = EPI == EPI_RESID ? kPairOutStage : 0;
    uint8_t *out_stage = smem + C::STAGES * C::STAGE_BYTES;
    uint64_t *full = reinterpret_cast<uint64_t *>(out_stage + kOutStage);
    uint64_t *empty = full + C::STAGES;
    uint64_t *tmem_full = empty + C::STAGES;      // [2]
    uint64_t *tmem_empty = tmem_full + 2;         // [2]
    uint32_t *tmem_holder = reinterpret_cast<uint32_t *>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
    asm volatile("griddepcontrol.launch_dependents;");        // launched with programmatic serialization (see launch_pair)
    if (threadIdx.x == 0) trace_min(p.trace, 0);
    const int total_kb = (p.k + BK - 1) / BK;
    // split-K (EPI_F32 only): a work item is (output tile, k slice); slice ks writes its partial tile to
    // out + ks * split_stride and the consumer (the next RMSNorm) adds the slices up in a fixed order
    const int S = p.splits > 1 ? p.splits : 1;
    const int kbp = S > 1 ? p.kb_per_split : total_kb;
    const int n_work = m_pairs * n_tiles * S;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmBhalf);
        for (int s = 0; s < C::STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 2); }   // empty: both CTAs' MMAs
        for (int b = 0; b < 2; ++b) { mbar_init(tmem_full + b, 1); mbar_init(tmem_empty + b, 8); }  // 8 epilogue warps
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_holder, C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                            // the peer's barriers exist before anything targets them
    tc_fence_after();
    const uint32_t tmem = *tmem_holder;

    if (warp == 0) {
        if (lane == 0) {
            const uint64_t pol_act = policy_evict_last(), pol_w = policy_evict_last();
            int it = 0;
            // The grid may start while its predecessor (which produces the activations) drains: the first ring
            // of weight half-tiles is requested before griddepcontrol.wait, the activation tiles after it.
            int pre = 0;
            if (cluster_id < n_work) {
                const int tile = cluster_id / S, ks = cluster_id - tile * S;
                int mp, nt;
                pair_tile_coords(tile, m_pairs, n_tiles, p.tile_band, mp, nt);
                const int n0 = nt * BN + (int)rank * (BN / 2), kb0 = ks * kbp;
                const int nkb = min(kbp, total_kb - kb0);
                pre = nkb < C::STAGES ? nkb : C::STAGES;
                for (int i = 0; i < pre; ++i) {
                    mbar_expect_tx(full + i, C::STAGE_BYTES);
                    tma_load_2d_mcast(smem + i * C::STAGE_BYTES + C::A_BYTES + rank * (C::B_BYTES / 2), &tmBhalf, full + i, (kb0 + i) * BK, n0,
                                      (uint16_t)3, pol_w);
                }
            }
            asm volatile("griddepcontrol.wait;" ::: "memory");
            trace_min(p.trace, 1);
            for (int wk = cluster_id; wk < n_work; wk += n_clusters) {
                const int tile = wk / S, ks = wk - tile * S;
                int mp, nt;
                pair_tile_coords(tile, m_pairs, n_tiles, p.tile_band, mp, nt);
                const int m0 = (2 * mp + (int)rank) * BM, n0 = nt * BN + (int)rank * (BN / 2), kb0 = ks * kbp;
                const int nkb = min(kbp, total_kb - kb0);
                for (int i = 0; i < nkb; ++i, ++it) {
                    const int s = it % C::STAGES;
                    uint8_t *sa = smem + s * C::STAGE_BYTES;
                    if (it < pre) {                                      // weights of this stage are already on their way
                        tma_load_2d_hint(sa, &tmA, full + s, (kb0 + i) * BK, m0, pol_act);
                        continue;
                    }
                    const uint32_t ph = (it / C::STAGES) & 1;
                    mbar_wait(empty + s, ph ^ 1);                       // both CTAs are done with this stage
                    mbar_expect_tx(full + s, C::STAGE_BYTES);
                    tma_load_2d_hint(sa, &tmA, full + s, (kb0 + i) * BK, m0, pol_act);
                    tma_load_2d_mcast(sa + C::A_BYTES + rank * (C::B_BYTES / 2), &tmBhalf, full + s, (kb0 + i) * BK, n0, (uint16_t)3, pol_w);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16_f32(BM, BN);
            int it = 0, tile_it = 0;
            for (int wk = cluster_id; wk < n_work; wk += n_clusters, ++tile_it) {
                const int ab = tile_it & 1;
                mbar_wait(tmem_empty + ab, ((tile_it >> 1) & 1) ^ 1);   // epilogue drained this accumulator
                tc_fence_after();
                const uint32_t acc = tmem + (uint32_t)(ab * BN);
                const int nkb = min(kbp, total_kb - (wk % S) * kbp);
                for (int i = 0; i < nkb; ++i, ++it) {
                    const int s = it % C::STAGES;
                    const uint32_t ph = (it / C::STAGES) & 1;
                    mbar_wait(full + s, ph);
                    tc_fence_after();
                    const uint32_t sa = base + s * C::STAGE_BYTES;
                    const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + C::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) umma_bf16(acc, da + 2 * k, db + 2 * k, idesc, (i | k) != 0);
                    umma_commit_mcast(empty + s, (uint16_t)3);          // stage free in both CTAs once these retire
                }
                umma_commit(tmem_full + ab);
            }
            trace_max(p.trace, 2);
        }
    } else {
        const int q = warp & 3;
        const int eg = (warp - 2) >> 2;                        // column half this warp drains: [eg BN/2, (eg + 1) BN/2)
        const int c_beg = eg * (BN / 2), c_end = c_beg + BN / 2;
        int tile_it = 0;
        asm volatile("griddepcontrol.wait;" ::: "memory");     // the epilogue reads / overwrites buffers earlier kernels used
        GemmParams ps = p;                                     // per-slice output pointer (split-K)
        auto release_acc = [&](int ab) {                       // this warp has read its last accumulator column of the tile
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tmem_empty + ab);
        };
        for (int wk = cluster_id; wk < n_work; wk += n_clusters, ++tile_it) {
            const int tile = wk / S, ks = wk - tile * S;
            int mp, nt;
            pair_tile_coords(tile, m_pairs, n_tiles, p.tile_band, mp, nt);
            const int arow = (2 * mp + (int)rank) * BM + q * 32 + lane, n0 = nt * BN;
            const int ab = tile_it & 1;
            if (EPI == EPI_F32) ps.out = reinterpret_cast<float *>(p.out) + (size_t)ks * p.split_stride;
            // QKV epilogue: the row's position, KV page and rope row do not depend on the accumulator -- fetch them (and
            // pull the cos/sin row towards L1) while the MMAs of this tile are still running
            int row_pos = 0, row_page = 0;
            if (EPI == EPI_QKV_ROPE && arow < p.a_rows) {
                row_pos = __ldg(p.pos + arow);
                row_page = __ldg(p.block_table + (size_t)__ldg(p.seq + arow) * p.max_pages + row_pos / KX_KV_PAGE);
                const char *cs_row = reinterpret_cast<const char *>(p.rope) + (size_t)row_pos * (p.hd >> 1) * 8;
                for (int o = 0; o < (p.hd >> 1) * 8; o += 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(cs_row + o));
            }
            mbar_wait(tmem_full + ab, (tile_it >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(ab * BN);
            if (EPI == EPI_QKV_ROPE) {
                // head by head: the rotate-half partners of a head live hd/2 columns apart in the same row
                const int hd = p.hd, half = hd >> 1;
                const int heads = BN / hd, h_mid = (heads + 1) >> 1;        // the two warp groups split the tile's heads
                const int h_beg = eg ? h_mid * hd : 0, h_end = eg ? BN : h_mid * hd;
                if (h_beg >= h_end) release_acc(ab);                       // a single head in the tile: nothing for the second group
#pragma unroll 1
                for (int h0 = h_beg; h0 < h_end; h0 += hd) {
#pragma unroll 1
                    for (int d0 = 0; d0 < half; d0 += 32) {
                        float x1[32], x2[32];
                        tmem_ld_32x32(taddr + h0 + d0, reinterpret_cast<uint32_t(&)[32]>(x1));
                        tmem_ld_32x32(taddr + h0 + half + d0, reinterpret_cast<uint32_t(&)[32]>(x2));
                        tmem_ld_wait();
                        if (h0 + hd >= h_end && d0 + 32 >= half) release_acc(ab);   // last read of this accumulator
                        if (arow < p.a_rows && n0 + h0 < p.b_rows) qkv_rope_store(p, x1, x2, arow, (n0 + h0) / hd, d0, row_pos, row_page);
                    }
                }
            } else if (EPI == EPI_RESID) {
                // residual accumulate without reading the residual: each warp stages its 32 rows x 32 columns of fp32 in
                // shared memory (128 B swizzle) and one lane issues a TMA reduce-add of that box; the add happens in the
                // L2, rows / columns past the matrix edge are clipped by the tensor map.  (A read-modify-write from the
                // SM -- one row segment per thread -- made K = 2048 tiles epilogue bound: 19 us against 14 us of MMAs.)
                uint8_t *obuf = out_stage + (warp - 2) * 4096;
                const int row0 = (2 * mp + (int)rank) * BM + q * 32;
#pragma unroll 1
                for (int c = c_beg; c < c_end; c += 32) {
                    float v[32];
                    tmem_ld_32x32(taddr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                    tmem_ld_wait();
                    if (c + 32 >= c_end) release_acc(ab);
                    if (lane == 0) bulk_wait_group_read0();               // the previous box has left the staging buffer
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<float4 *>(obuf + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                            make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0 && row0 < p.a_rows && n0 + c < p.b_rows) {
                        tma_reduce_add_2d(&tmOut, obuf, n0 + c, row0);
                        bulk_commit_group();
                    }
                }
            } else if (EPI == EPI_ARGMAX) {
                // token-major lm_head: this thread's row is a token; scan the tile's vocabulary columns in order
                // (one part per tile: the first warp group scans all columns, the second only hands the accumulator back)
                float bv = -INFINITY;
                int bi = 0x7fffffff;
                if (eg) release_acc(ab);
#pragma unroll 1
                for (int c = 0; c < (eg ? 0 : BN); c += 32) {
                    float v[32];
                    tmem_ld_32x32(taddr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                    tmem_ld_wait();
                    if (c + 32 >= BN) release_acc(ab);
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (n0 + c + j < p.b_rows && v[j] > bv) { bv = v[j]; bi = n0 + c + j; }
                }
                if (!eg && arow < p.a_rows) reinterpret_cast<int2 *>(p.out)[(size_t)nt * p.ldo + arow] = make_int2(__float_as_int(bv), bi);
            } else {
#pragma unroll 1
                for (int c = c_beg; c < c_end; c += 32) {
                    float v[32];
                    tmem_ld_32x32(taddr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                    tmem_ld_wait();
                    if (c + 32 >= c_end) release_acc(ab);                   // last read of this accumulator: hand it back
                    epilogue_chunk<EPI == EPI_ARGMAX ? EPI_F32 : EPI, false>(ps, v, arow, n0 + c, false);
                }
            }
        }
    }

    if (EPI == EPI_RESID && warp >= 2 && lane == 0) bulk_wait_group0();   // every reduce-store of this warp has completed
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                            // nobody leaves while the peer can still reach its smem
    if (warp == 1) tmem_dealloc(tmem, C::TMEM_COLS);
    if (threadIdx.x == 0) trace_max(p.trace, 3);
}

template <int BN, int EPI>
int launch_pair(const CUtensorMap &ta, const CUtensorMap &tbh, const GemmParams &p, cudaStream_t st) {
    auto kern = kx_gemm_bf16_pair_kernel<BN, EPI>;
    constexpr int kSmem = Cfg2<BN>::SMEM_BYTES + (EPI == EPI_RESID ? kPairOutStage : 0);
    static_assert(kSmem <= 232448, "pair kernel shared memory");
    static PerDeviceOnce once;
    static int sms_of[64];
    if (once.first()) {
        KX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        int dev = 0;
        KX_CUDA(cudaGetDevice(&dev));
        KX_CUDA(cudaDeviceGetAttribute(&sms_of[dev & 63], cudaDevAttrMultiProcessorCount, dev));
        once.done();
    }
    const int sms = sms_of[PerDeviceOnce::device()];
    const int m_pairs = (p.a_rows + 2 * BM - 1) / (2 * BM), n_tiles = (p.b_rows + BN - 1) / BN;
    GemmParams pp = p;
    if (pp.splits > 1) {
        KX_REQUIRE(EPI == EPI_F32, KX_EINVAL, "kx_gemm: the token-major kernel splits k only with the fp32 epilogue");
        const int total_kb = (pp.k + BK - 1) / BK;
        if (pp.splits > total_kb) pp.splits = total_kb;
        pp.kb_per_split = (total_kb + pp.splits - 1) / pp.splits;
        pp.splits = (total_kb + pp.kb_per_split - 1) / pp.kb_per_split;
    } else {
        pp.splits = 1;
    }
    {   // bands of at most ~32 MB of weights (BN rows x k bf16 per feature tile), equal widths
        const int64_t tile_bytes = (int64_t)BN * pp.k * 2;
        // (Llama-3.2-1B, 512 x 256 tokens: prefill 236 ms with one band, 234 with 8 MB bands, 228 with 16 - 32 MB, 230 with 48)
        int fit = (int)((32ll << 20) / tile_bytes);
        if (fit < 1) fit = 1;
        const int bands = (n_tiles + fit - 1) / fit;
        pp.tile_band = (n_tiles + bands - 1) / bands;
    }
    int clusters = sms / 2;
    if (clusters > m_pairs * n_tiles * pp.splits) clusters = m_pairs * n_tiles * pp.splits;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(2 * clusters);
    cfg.blockDim = dim3(kPairThreads);
    cfg.dynamicSmemBytes = kSmem;
    cfg.stream = st;
    CUtensorMap tout{};                                 // fp32 view of the residual rows for the TMA reduce-add epilogue
    if (EPI == EPI_RESID) {
        const int rc = make_tmap_f32_box32(&tout, pp.out, pp.a_rows, pp.b_rows, pp.ldo);
        if (rc != KX_OK) return rc;
    }
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // prologue + first weight tiles overlap the predecessor's tail
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    KX_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tbh, tout, pp, m_pairs, n_tiles));
    count_launch();
    return KX_OK;
}

template <int BN>
int launch_pair_epi(const CUtensorMap &ta, const CUtensorMap &tbh, const GemmParams &p, cudaStream_t st) {
    switch (p.epilogue) {
        case EPI_BF16: return launch_pair<BN, EPI_BF16>(ta, tbh, p, st);
        case EPI_RESID: return launch_pair<BN, EPI_RESID>(ta, tbh, p, st);
        case EPI_SWIGLU: return launch_pair<BN, EPI_SWIGLU>(ta, tbh, p, st);
        case EPI_F32: return launch_pair<BN, EPI_F32>(ta, tbh, p, st);
        case EPI_QKV_ROPE: return launch_pair<BN, EPI_QKV_ROPE>(ta, tbh, p, st);
        case EPI_ARGMAX: return launch_pair<BN, EPI_ARGMAX>(ta, tbh, p, st);
    }
    set_error("kx_gemm: unknown epilogue %d", p.epilogue);
    return KX_EINVAL;
}


// ---------------------------------------------------------------------------------
// Decode GEMM with split-K inside a thread-block cluster (swap-AB orientation).
//
// gridDim.z CTAs form one cluster per output tile; CTA z streams its slice of K (its
// share of the weight tile) through the same TMA -> tcgen05 pipeline, parks its fp32
// partial tile in its own shared memory, and after a cluster barrier every CTA reduces
// a slice of the tile's batch columns by reading all partials over distributed shared
// memory, in a fixed order (so the result is deterministic), and applies the epilogue.
// No global scratch, no atomics.
// ---------------------------------------------------------------------------------
template <int EPI>
__device__ __forceinline__ void epilogue_swap_elem(const GemmParams &p, float v, int arow, int b, bool ok, float bias) {
    if (EPI == EPI_SWIGLU) {
        const float other = __shfl_xor_sync(0xffffffffu, v, 1);
        if (ok && !(arow & 1))
            reinterpret_cast<__nv_bfloat16 *>(p.out)[(size_t)b * p.ldo + (arow >> 1)] = __float2bfloat16(silu(v) * other);
    } else if (!ok) {
        return;
    } else if (EPI == EPI_BF16) {
        reinterpret_cast<__nv_bfloat16 *>(p.out)[(size_t)b * p.ldo + arow] = __float2bfloat16(v + bias);
    } else if (EPI == EPI_RESID) {
        reinterpret_cast<float *>(p.out)[(size_t)b * p.ldo + arow] += v;
    } else {
        reinterpret_cast<float *>(p.out)[(size_t)b * p.ldo + arow] = v;
    }
}

constexpr int kMaxSplits = 16;   // cluster size: 8 is the portable limit, 16 needs the non-portable opt-in (see max_cluster_splits)

// bulk copy from this CTA's shared memory into a cluster CTA's shared memory; completes `bytes` on an mbarrier of the
// DESTINATION CTA when the data has landed (both addresses 16-byte aligned, bytes a multiple of 16)
__device__ __forceinline__ void dsmem_bulk_copy(uint32_t dst_cluster_addr, uint32_t src_cta_addr, uint32_t bytes, uint32_t dst_cluster_mbar) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst_cluster_addr), "r"(src_cta_addr), "r"(bytes), "r"(dst_cluster_mbar) : "memory");
}
// execution barrier over the cluster without memory ordering (nothing is handed over through memory)
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait_plain() { asm volatile("barrier.cluster.wait.aligned;" ::: "memory"); }

// Split-K reduction, push model.  Batch column j of the tile is owned by cluster rank j / cols_per.  Once every CTA
// of the cluster has finished its MMAs (one cluster barrier, execution only: nobody's stage ring is in use any more),
// each CTA stages its accumulator (TMEM -> registers -> shared memory, 32 columns at a time, column-major so that a
// rank's columns are contiguous) and sends every owner its columns with ONE bulk shared-to-shared copy per (chunk,
// owner) into the owner's slot [source rank]; the copy engine completes the bytes on the owner's `xchg` mbarrier,
// which the owner armed with the byte count it expects.  An owner whose slices have landed adds them from its OWN
// shared memory in rank order (deterministic) and applies the epilogue -- no second barrier, no fence, CTAs leave
// independently.  (Per-element st.async stores cost 2-3 us per launch: 8,192 four-byte transactions per CTA.)  (The first version pulled: every owner loaded the other CTAs' partials over DSMEM column by
// column, a full round trip each, between two release/acquire cluster barriers -- 5 us per launch, 13 us for the
// 32 columns of a gate/up CTA, more than the weight streaming itself.)
template <int BN, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 2)
kx_gemm_bf16_ksplit_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
    using C = Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t *smem = smem_raw + (base - smem_u32(smem_raw));
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + C::STAGES * C::STAGE_BYTES);
    uint64_t *empty = full + C::STAGES;
    uint64_t *tmem_full = empty + C::STAGES;
    uint64_t *xchg = tmem_full + 1;                            // split-K slices of this CTA's columns have landed
    uint32_t *tmem_holder = reinterpret_cast<uint32_t *>(xchg + 1);
    float *recv = reinterpret_cast<float *>(smem);             // [S][cols_per][128] fp32, reuses the drained stage ring
    constexpr int kRecvBytes = (BN + kMaxSplits) * BM * 4;     // S * cols_per <= BN + S - 1 columns of 512 bytes
    constexpr int kStageBufs = BN <= 64 ? (BN / 32 < 2 ? 1 : 2) : 1;   // send staging: 32 columns x 128 rows fp32 each
    float *stage = reinterpret_cast<float *>(smem + kRecvBytes);
    static_assert(kRecvBytes + kStageBufs * 32 * BM * 4 <= C::STAGES * C::STAGE_BYTES, "exchange buffers must fit in the stage ring");
    __shared__ float rinv_s[BN];                               // deferred RMSNorm: 1 / rms of every batch column of this tile
    __shared__ float sq_s[EPI == EPI_RESID_NORM ? 4 * BN : 1]; // EPI_RESID_NORM: per-warp sums of squares
    __shared__ long long stamp_s[4];                           // diagnostic cycle stamps of CTA (0, 0, 0) (tracing only)
    const bool cta0 = p.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM;
    asm volatile("griddepcontrol.launch_dependents;");        // our own dependents may begin their prologue
    if (threadIdx.x == 0) {
        trace_min(p.trace, 0);
        trace_max(p.trace, 7);
    }
    const int S = gridDim.z, z = blockIdx.z;                  // cluster = the z column; rank == z
    const int total_kb = (p.k + BK - 1) / BK;
    const int kb0 = z * p.kb_per_split;
    const int nkb = min(total_kb - kb0, p.kb_per_split);
    const int cols_per = (BN + S - 1) / S;
    const int j0 = z * cols_per, j1 = min(min(BN, j0 + cols_per), p.b_rows - n0);   // batch columns this CTA owns

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < C::STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
        mbar_init(tmem_full, 1);
        mbar_init(xchg, 1);
        fence_barrier_init();
        if (S > 1) mbar_expect_tx(xchg, (uint32_t)(max(j1 - j0, 0) * S * BM * 4));   // every rank sends 128 floats per owned column
    }
    if (warp == 1) tmem_alloc(tmem_holder, C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_holder;
    const int q = warp & 3, r = q * 32 + lane;                // epilogue warps: TMEM lane quarter, accumulator row
    const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16);
    if (EPI == EPI_RESID_NORM && warp >= 2 && m0 + r < p.a_rows) {
        // The residual rows this thread will update were written two launches ago and ~100 MB of weights have
        // streamed through the L2 since: pull their lines (and the gain) back in now, while the predecessor still
        // runs -- a demand miss in the epilogue queues behind the weight tiles of the next two launches (8 us).
        // Only the L2 is warmed, so data that is still being produced is not a problem.
        for (int j = j0; j < j1; ++j)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const float *>(p.out) + (size_t)(n0 + j) * p.ldo + m0 + r));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const __nv_bfloat16 *>(p.norm_g) + m0 + r));
    }

    if (warp == 0) {
        if (lane == 0) {
            const uint64_t pol_w = policy_evict_first(), pol_x = policy_evict_last();
            // Programmatic dependent launch: this grid may start while its predecessor (which produces
            // the activations) is still running.  Weights do not depend on it, so the first ring of
            // weight tiles is already in flight before griddepcontrol.wait; activations follow after.
            const int pre = nkb < C::STAGES ? nkb : C::STAGES;
            for (int i = 0; i < pre; ++i) {
                mbar_expect_tx(full + i, C::STAGE_BYTES);
                tma_load_2d_hint(smem + i * C::STAGE_BYTES, &tmA, full + i, (kb0 + i) * BK, m0, pol_w);
            }
            asm volatile("griddepcontrol.wait;" ::: "memory");
            trace_min(p.trace, 1);
            trace_max(p.trace, 4);
            if (cta0) stamp_s[0] = clock64();
            for (int i = 0; i < pre; ++i)
                tma_load_2d_hint(smem + i * C::STAGE_BYTES + C::A_BYTES, &tmB, full + i, (kb0 + i) * BK, n0, pol_x);
            for (int i = pre; i < nkb; ++i) {
                const int s = i % C::STAGES;
                mbar_wait(empty + s, ((i / C::STAGES) & 1) ^ 1);
                mbar_expect_tx(full + s, C::STAGE_BYTES);
                uint8_t *sa = smem + s * C::STAGE_BYTES;
                tma_load_2d_hint(sa, &tmA, full + s, (kb0 + i) * BK, m0, pol_w);
                tma_load_2d_hint(sa + C::A_BYTES, &tmB, full + s, (kb0 + i) * BK, n0, pol_x);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16_f32(BM, BN);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % C::STAGES;
                mbar_wait(full + s, (i / C::STAGES) & 1);
                if (cta0 && i == 0) stamp_s[1] = clock64();
                tc_fence_after();
                const uint32_t sa = base + s * C::STAGE_BYTES;
                const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + C::A_BYTES);
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) umma_bf16(tmem, da + 2 * k, db + 2 * k, idesc, (i | k) != 0);
                umma_commit(empty + s);
            }
            umma_commit(tmem_full);
            if (cta0) stamp_s[2] = clock64();
        }
    } else {
        asm volatile("griddepcontrol.wait;" ::: "memory");     // outputs (and the residual read) follow the predecessor
        if (p.ssq_in && r < BN) {                              // deferred RMSNorm: 1 / rms of batch column n0 + r, tiles added in order
            float ss = 0.f;
            if (n0 + r < p.b_rows)
                for (int t = 0; t < p.ssq_tiles; ++t) ss += p.ssq_in[(size_t)t * p.ssq_ld + n0 + r];
            rinv_s[r] = 1.0f / sqrtf(ss * p.norm_inv_n + p.norm_eps);
        }
        mbar_wait(tmem_full, 0);                               // all MMAs retired: the stage ring is free
        tc_fence_after();
        if (threadIdx.x == 64) trace_max(p.trace, 2);
    }

    const bool probe = p.trace && threadIdx.x == 64 && blockIdx.y == 0 && blockIdx.z == 0 && blockIdx.x == 0;
    long long ckb = 0, ckp = 0, ckw = 0, ck0 = 0, ck1 = 0;
    if (probe) ckb = clock64();
    // EPI_RESID_NORM: the residual values of the first group of owned columns, requested now so that their latency
    // hides behind the split-K exchange (the predecessor is complete: the epilogue warps are past griddepcontrol.wait)
    float xres[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (EPI == EPI_RESID_NORM && warp >= 2 && m0 + r < p.a_rows) {
#pragma unroll
        for (int c = 0; c < 8; ++c)
            if (j0 + c < j1) xres[c] = reinterpret_cast<const float *>(p.out)[(size_t)(n0 + j0 + c) * p.ldo + m0 + r];
    }
    if (S > 1) {
        __syncthreads();
        cluster_arrive_relaxed();                              // this CTA no longer reads its stage ring ...
        cluster_wait_plain();                                  // ... and neither does any other: slices may be stored into it
        if (probe) ckp = clock64();
        if (warp >= 2) {
            const uint32_t local = smem_u32(recv), bar_local = smem_u32(xchg);
            const int valid = min(BN, p.b_rows - n0);          // columns past the batch are never sent
            int chunk = 0;
#pragma unroll 1
            for (int c = 0; c < valid; c += 32, ++chunk) {
                float *buf = stage + (chunk % kStageBufs) * (32 * BM);
                if (chunk >= kStageBufs) {                     // the copies that read this buffer have finished reading it
                    if (warp == 2) bulk_wait_group_read0();       // (bulk groups are per thread: every sending lane waits for its own)
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                }
                float v[32];
                tmem_ld_32x32(taddr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) buf[j * BM + r] = v[j];
                fence_proxy_async_smem();                      // generic-proxy writes before the copy engine reads them
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (warp == 2) {                               // lane l sends the l-th owner that has columns in this chunk
                    const int c_end = min(c + 32, valid);
                    const int first = c / cols_per, owner = first + lane;
                    const int a = max(owner * cols_per, c), e = min((owner + 1) * cols_per, c_end);
                    if (owner < S && e > a)
                        dsmem_bulk_copy(dsmem_addr(local, (uint32_t)owner) + (uint32_t)((z * cols_per + (a - owner * cols_per)) * BM * 4),
                                        smem_u32(buf) + (uint32_t)((a - c) * BM * 4), (uint32_t)((e - a) * BM * 4),
                                        dsmem_addr(bar_local, (uint32_t)owner));
                    bulk_commit_group();
                }
            }
            if (probe) ckw = clock64();
            mbar_wait(xchg, 0);                                // all S slices of this CTA's columns have landed
        }
    }
    if (probe) ck0 = clock64();

    if (warp >= 2) {
        const int arow = m0 + r;
        const bool ok = arow < p.a_rows;
        const float bias = (ok && p.bias) ? p.bias[arow] : 0.0f;   // plain load: ld.global.nc may be hoisted above the null test
        // sums of kCG owned columns at a time; slices added in ascending rank order: deterministic.
        // A single-CTA cluster has nothing to add: its columns come straight from TMEM.
        constexpr int kCG = 8;
        auto column_sums = [&](int jg, float (&acc)[kCG]) {
            if (S == 1) {
                uint32_t t[8];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                             : "=r"(t[0]), "=r"(t[1]), "=r"(t[2]), "=r"(t[3]), "=r"(t[4]), "=r"(t[5]), "=r"(t[6]), "=r"(t[7])
                             : "r"(taddr + jg) : "memory");
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < kCG; ++c) acc[c] = __uint_as_float(t[c]);
                return;
            }
#pragma unroll
            for (int c = 0; c < kCG; ++c) acc[c] = 0.f;
            // (columns past j1 read slots nobody wrote; their sums are discarded by the callers' predicates)
            const float *src = recv + (jg - j0) * BM + r;
#pragma unroll 4
            for (int s = 0; s < S; ++s, src += cols_per * BM) {
#pragma unroll
                for (int c = 0; c < kCG; ++c) acc[c] += src[c * BM];
            }
        };
        // ---- fused tensor-parallel all-reduce (2 ranks) over peer memory: this CTA's partial sums go to the peer's receive
        // buffer, a flag per CTA publishes them, and the peer's partial is read back from this rank's own buffer ----
        size_t tp_buf = 0;
        auto tp_exchange = [&]() {
            const uint32_t slot = (blockIdx.y * gridDim.x + blockIdx.x) * gridDim.z + blockIdx.z;
            const uint32_t epoch = p.tp_epoch[slot] + 1;       // same count on both ranks: they run the same launches
            tp_buf = (size_t)(p.tp_seq & 1) * (size_t)p.tp_rows * p.ldo;   // receive half-buffer of this launch
            for (int jg = j0; jg < j1; jg += kCG) {
                float v[kCG];
                column_sums(jg, v);
#pragma unroll
                for (int c = 0; c < kCG; ++c)
                    if (ok && jg + c < j1) p.tp_peer_recv[tp_buf + (size_t)(n0 + jg + c) * p.ldo + arow] = v[c];      // NVLink store
            }
            __threadfence_system();
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (threadIdx.x == 64) {
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.tp_peer_flag + slot), "r"(epoch) : "memory");
                uint32_t seen;
                long long polls = 0;
                do {
                    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.tp_flag + slot) : "memory");
                    // the peer runs the same launch: if its flag does not arrive within seconds the ranks have lost step
                    // (or the peer died); trap instead of hanging both GPUs
                    if (++polls > (1ll << 24)) { printf("kx_gemm: peer rank did not publish its partial sums (slot %u)\n", slot); __trap(); }
                } while ((int32_t)(seen - epoch) < 0);
                p.tp_epoch[slot] = epoch;
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
        };
        if (EPI == EPI_RESID && p.tp_peer_recv) {
            tp_exchange();
            for (int jg = j0; jg < j1; jg += kCG) {
                float v[kCG];
                column_sums(jg, v);                            // this rank's partial again (still in shared memory / TMEM)
#pragma unroll
                for (int c = 0; c < kCG; ++c) {
                    if (ok && jg + c < j1) {
                        const float other = __ldcv(p.tp_recv + tp_buf + (size_t)(n0 + jg + c) * p.ldo + arow);   // the peer's partial
                        reinterpret_cast<float *>(p.out)[(size_t)(n0 + jg + c) * p.ldo + arow] += v[c] + other;
                    }
                }
            }
        } else if (EPI == EPI_RESID_NORM) {
            // residual update + preparation of the next RMSNorm (deferred: see kx_gemm.cuh); under tensor parallelism the
            // peer rank's partial sums are added first (both ranks then hold the same residual, bit for bit)
            const bool tp = p.tp_peer_recv != nullptr;
            if (tp) tp_exchange();
            const float g = ok ? __bfloat162float(reinterpret_cast<const __nv_bfloat16 *>(p.norm_g)[arow]) : 0.f;
            for (int jg = j0; jg < j1; jg += kCG) {
                float x[kCG];
                if (jg == j0) {
#pragma unroll
                    for (int c = 0; c < kCG; ++c) x[c] = xres[c];   // requested before the exchange
                } else {
#pragma unroll
                    for (int c = 0; c < kCG; ++c) {
                        x[c] = 0.f;
                        if (ok && jg + c < j1) x[c] = reinterpret_cast<float *>(p.out)[(size_t)(n0 + jg + c) * p.ldo + arow];
                    }
                }
                float v[kCG];
                column_sums(jg, v);
                if (probe && jg == j0) ck1 = clock64();
                if (tp) {
#pragma unroll
                    for (int c = 0; c < kCG; ++c)
                        if (ok && jg + c < j1) v[c] += __ldcv(p.tp_recv + tp_buf + (size_t)(n0 + jg + c) * p.ldo + arow);
                }
                float sq[kCG];
#pragma unroll
                for (int c = 0; c < kCG; ++c) {
                    x[c] += v[c];
                    sq[c] = ok ? __fmul_rn(x[c], x[c]) : 0.f;
                }
#pragma unroll
                for (int c = 0; c < kCG; ++c) {
                    if (ok && jg + c < j1) {
                        reinterpret_cast<float *>(p.out)[(size_t)(n0 + jg + c) * p.ldo + arow] = x[c];
                        reinterpret_cast<__nv_bfloat16 *>(p.norm_out)[(size_t)(n0 + jg + c) * p.ldo + arow] = __float2bfloat16(__fmul_rn(x[c], g));
                    }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1)               // the kCG columns' butterflies side by side
#pragma unroll
                    for (int c = 0; c < kCG; ++c) sq[c] += __shfl_xor_sync(0xffffffffu, sq[c], o);
                if (lane == 0) {
#pragma unroll
                    for (int c = 0; c < kCG; ++c)
                        if (jg + c < j1) sq_s[q * BN + jg + c] = sq[c];
                }
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            for (int j = j0 + (int)threadIdx.x - 64; j < j1; j += 128)     // the four warps' sums, fixed order
                p.ssq_out[(size_t)blockIdx.y * p.ssq_ld + n0 + j] = (sq_s[j] + sq_s[BN + j]) + (sq_s[2 * BN + j] + sq_s[3 * BN + j]);
        } else if (EPI == EPI_SWIGLU && S > 1) {
            // The slices are in shared memory, so a thread need not keep to "its" TMEM lane: thread t takes the
            // (gate, up) pair t % 64 of the tile (two adjacent floats: one 8-byte load per slice) for every second
            // owned column -- no shuffle, no lane computing a SiLU that is thrown away, 128-byte stores per column.
            const int pr = (int)threadIdx.x & 63, half = ((int)threadIdx.x - 64) >> 6;
            const bool pok = m0 + 2 * pr < p.a_rows;
            __nv_bfloat16 *orow = reinterpret_cast<__nv_bfloat16 *>(p.out) + ((m0 >> 1) + pr);
            for (int jg = j0 + half; jg < j1; jg += 2 * kCG) {
                float2 acc[kCG];
#pragma unroll
                for (int c = 0; c < kCG; ++c) acc[c] = make_float2(0.f, 0.f);
                const float *src = recv + (jg - j0) * BM + 2 * pr;
#pragma unroll 2
                for (int s = 0; s < S; ++s, src += cols_per * BM) {
#pragma unroll
                    for (int c = 0; c < kCG; ++c) {
                        const float2 t = *reinterpret_cast<const float2 *>(src + 2 * c * BM);
                        acc[c].x += t.x; acc[c].y += t.y;
                    }
                }
#pragma unroll
                for (int c = 0; c < kCG; ++c) {
                    const int j = jg + 2 * c;
                    const float sc = p.ssq_in ? rinv_s[min(j, BN - 1)] : 1.0f;
                    const float gte = __fmul_rn(acc[c].x, sc), up = __fmul_rn(acc[c].y, sc);
                    if (pok && j < j1) orow[(size_t)(n0 + j) * p.ldo] = __float2bfloat16(silu(gte) * up);
                }
            }
        } else {
            for (int jg = j0; jg < j1; jg += kCG) {
                float v[kCG];
                column_sums(jg, v);
#pragma unroll
                for (int c = 0; c < kCG; ++c)                  // (no early exit: the columns' dependency chains interleave)
                    epilogue_swap_elem<EPI>(p, p.ssq_in ? __fmul_rn(v[c], rinv_s[min(jg + c, BN - 1)]) : v[c], arow, n0 + jg + c,
                                            ok && jg + c < j1, bias);
            }
        }
    }
    if (probe) {       // cycles of CTA (0,0,0): [5] = predecessor done -> first operands landed | -> last MMA issued,
                       // [6] = -> accumulator complete | split-K exchange, [7] = reduce + epilogue
        (void)ckp; (void)ckw; (void)ck1;
        p.trace[5] = ((unsigned long long)(stamp_s[1] - stamp_s[0]) << 32) | (unsigned long long)(unsigned)(stamp_s[2] - stamp_s[1]);
        p.trace[6] = ((unsigned long long)(ckb - stamp_s[2]) << 32) | (unsigned long long)(unsigned)(ck0 - ckb);
    }
    if (S > 1 && warp == 2) bulk_wait_group_read0();           // the copy engine has read this CTA's staging buffers
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, C::TMEM_COLS);
    if (threadIdx.x == 0) trace_max(p.trace, 3);
}

template <int BN, int EPI>
int launch_ksplit(const CUtensorMap &ta, const CUtensorMap &tb, const GemmParams &p, dim3 grid, cudaStream_t st) {
    auto kern = kx_gemm_bf16_ksplit_kernel<BN, EPI>;
    static PerDeviceOnce once;
    if (once.first()) {
        KX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM_BYTES));
        KX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        once.done();
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = Cfg<BN>::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = grid.z;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // may overlap the tail of its predecessor
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    if (p.tp_peer_recv) {
        // Every CTA waits for a flag the SAME tile's CTA on the peer GPU publishes, so the grid has to be one resident
        // wave on both GPUs: two CTAs of this kernel fit an SM (ncu: shared-memory limit 2), and the measured sweep
        // (profiles/r01_sweep_decode_gemm.txt) shows clusters of up to 8 packing up to ~256 CTAs, larger ones ~192.
        // (cudaOccupancyMaxActiveClusters reports one CTA per SM for this kernel -- 15 clusters of 8 -- which the
        // launch list contradicts, so the bound is the measured one.)  A rank that still never sees its peer's flag
        // traps after a bounded number of polls instead of hanging both GPUs.
        const unsigned ctas = grid.x * grid.y * grid.z, limit = grid.z <= 8 ? 256u : 192u;
        KX_REQUIRE(ctas <= limit, KX_EINVAL, "kx_gemm: fused all-reduce needs its %u CTAs (clusters of %u) in one resident wave (<= %u)",
                   ctas, grid.z, limit);
    }
    KX_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, p));
    count_launch();
    return KX_OK;
}

template <int BN>
int launch_ksplit_epi(const CUtensorMap &ta, const CUtensorMap &tb, const GemmParams &p, dim3 grid, cudaStream_t st) {
    switch (p.epilogue) {
        case EPI_BF16: return launch_ksplit<BN, EPI_BF16>(ta, tb, p, grid, st);
        case EPI_RESID: return launch_ksplit<BN, EPI_RESID>(ta, tb, p, grid, st);
        case EPI_SWIGLU: return launch_ksplit<BN, EPI_SWIGLU>(ta, tb, p, grid, st);
        case EPI_F32: return launch_ksplit<BN, EPI_F32>(ta, tb, p, grid, st);
        case EPI_RESID_NORM: return launch_ksplit<BN, EPI_RESID_NORM>(ta, tb, p, grid, st);
    }
    set_error("kx_gemm: unknown epilogue %d", p.epilogue);
    return KX_EINVAL;
}

// ---- host side ----------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    });
    return fn;
}

template <int BN, int EPI, bool SWAP>
int launch(const CUtensorMap &ta, const CUtensorMap &tb, const GemmParams &p, dim3 grid, cudaStream_t st) {
    auto kern = kx_gemm_bf16_kernel<BN, EPI, SWAP>;
    static PerDeviceOnce once;      // per instantiation
    if (once.first()) {
        KX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM_BYTES));
        once.done();
    }
    kern<<<grid, kGemmThreads, Cfg<BN>::SMEM_BYTES, st>>>(ta, tb, p);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

template <int BN, bool SWAP>
int launch_epi(const CUtensorMap &ta, const CUtensorMap &tb, const GemmParams &p, dim3 grid, cudaStream_t st) {
    switch (p.epilogue) {
        case EPI_BF16: return launch<BN, EPI_BF16, SWAP>(ta, tb, p, grid, st);
        case EPI_RESID: return launch<BN, EPI_RESID, SWAP>(ta, tb, p, grid, st);
        case EPI_SWIGLU: return launch<BN, EPI_SWIGLU, SWAP>(ta, tb, p, grid, st);
        case EPI_F32: return launch<BN, EPI_F32, SWAP>(ta, tb, p, grid, st);
        case EPI_ARGMAX:
            KX_REQUIRE(SWAP && grid.z == 1, KX_EINVAL, "kx_gemm: the argmax epilogue of this kernel needs swap-AB without split-K");
            return launch<BN, EPI_ARGMAX, SWAP>(ta, tb, p, grid, st);
    }
    set_error("kx_gemm: unknown epilogue %d", p.epilogue);
    return KX_EINVAL;
}

}  // namespace

int make_tmap_bf16_2d(CUtensorMap *out, const void *ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    KX_REQUIRE(fn != nullptr, KX_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    KX_REQUIRE(((uintptr_t)ptr & 15) == 0 && (ld * 2) % 16 == 0, KX_EINVAL,
               "GEMM operand must be 16-byte aligned with a row pitch that is a multiple of 16 bytes");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    KX_REQUIRE(r == CUDA_SUCCESS, KX_ECUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return KX_OK;
}

// fp32 matrix [rows, cols] (row pitch ld elements) as 32 x 32 boxes with the 128 B swizzle: the residual epilogue's
// TMA reduce-add destination
int make_tmap_f32_box32(CUtensorMap *out, const void *ptr, int64_t rows, int64_t cols, int64_t ld) {
    EncodeTiledFn fn = encode_fn();
    KX_REQUIRE(fn != nullptr, KX_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
    KX_REQUIRE(((uintptr_t)ptr & 15) == 0 && (ld * 4) % 16 == 0, KX_EINVAL,
               "residual matrix must be 16-byte aligned with a row pitch that is a multiple of 16 bytes");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void *>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    KX_REQUIRE(r == CUDA_SUCCESS, KX_ECUDA, "cuTensorMapEncodeTiled (fp32) failed with CUresult %d", (int)r);
    return KX_OK;
}

// The CTA-pair kernel fetches its B operand (the weights) with TMA multicast.  TMA multicast loads here read only
// this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU: an operand that is
// not plain device memory of the current device (a peer allocation, an IPC mapping, managed memory) is refused.
int require_local_device_memory(const void *ptr, const char *what) {
    cudaPointerAttributes a{};
    int dev = -1;
    KX_CUDA(cudaGetDevice(&dev));
    const cudaError_t e = cudaPointerGetAttributes(&a, ptr);
    if (e != cudaSuccess) (void)cudaGetLastError();
    KX_REQUIRE(e == cudaSuccess && a.type == cudaMemoryTypeDevice && a.device == dev, KX_EINVAL,
               "%s must be device memory allocated on the current GPU (device %d)", what, dev);
    return KX_OK;
}

// Prefill tile width: 256-wide pair tiles when the feature count fills them, else 128-wide.
int gemm_pair_bn(int64_t a_rows, int64_t b_rows) {
    if (a_rows < 256) return 0;                       // too few rows for a CTA pair: single-CTA kernel
    return (b_rows % 256 == 0 || b_rows >= 2048) ? 256 : 128;
}

int gemm_launch_pair(const CUtensorMap &ta, const CUtensorMap &tb_half, const GemmParams &p, int bn, cudaStream_t st) {
    if (bn == 256) return launch_pair_epi<256>(ta, tb_half, p, st);
    if (bn == 128) return launch_pair_epi<128>(ta, tb_half, p, st);
    set_error("kx_gemm: unsupported pair tile width %d", bn);
    return KX_EINVAL;
}

int gemm_pick_bn(int swap, int64_t b_rows) {
    if (!swap) return 128;
    return b_rows <= 32 ? 32 : b_rows <= 64 ? 64 : 128;
}

size_t gemm_scratch_bytes(const GemmParams &p, int bn) {
    if (p.splits <= 1 || p.epilogue == EPI_RESID || p.epilogue == EPI_RESID_NORM || p.swap) return 0;   // swap-AB splits reduce over DSMEM
    const size_t tiles = (size_t)((p.a_rows + BM - 1) / BM) * (size_t)((p.b_rows + bn - 1) / bn);
    return tiles * (size_t)(BM * bn) * sizeof(float) + tiles * sizeof(uint32_t);
}

int gemm_launch(const CUtensorMap &ta, const CUtensorMap &tb, GemmParams p, int bn, void *scratch, cudaStream_t st) {
    const int total_kb = (p.k + BK - 1) / BK;
    int splits = p.splits < 1 ? 1 : p.splits;
    if (splits > total_kb) splits = total_kb;
    p.kb_per_split = (total_kb + splits - 1) / splits;
    splits = (total_kb + p.kb_per_split - 1) / p.kb_per_split;      // no empty slice
    p.splits = splits;
    KX_REQUIRE(p.epilogue != EPI_RESID_NORM || (p.swap && p.norm_g && p.norm_out && p.ssq_out), KX_EINVAL,
               "kx_gemm: EPI_RESID_NORM needs the swap-AB kernel and its norm arguments");
    if (p.swap && (splits > 1 || p.tp_peer_recv || p.ssq_in || p.epilogue == EPI_RESID_NORM)) {
        // decode: split-K inside a cluster (portable cluster size <= 8), reduced over distributed shared memory
        if (splits > kMaxSplits) splits = kMaxSplits;
        p.kb_per_split = (total_kb + splits - 1) / splits;
        splits = (total_kb + p.kb_per_split - 1) / p.kb_per_split;
        p.splits = splits;
        dim3 cgrid((unsigned)((p.b_rows + bn - 1) / bn), (unsigned)((p.a_rows + BM - 1) / BM), (unsigned)splits);
        switch (bn) {
            case 32: return launch_ksplit_epi<32>(ta, tb, p, cgrid, st);
            case 64: return launch_ksplit_epi<64>(ta, tb, p, cgrid, st);
            case 128: return launch_ksplit_epi<128>(ta, tb, p, cgrid, st);
        }
    }
    const size_t tiles = (size_t)((p.a_rows + BM - 1) / BM) * (size_t)((p.b_rows + bn - 1) / bn);
    if (splits > 1 && p.epilogue != EPI_RESID) {
        KX_REQUIRE(scratch != nullptr, KX_EINVAL, "kx_gemm: split-K needs scratch");
        p.ws = reinterpret_cast<float *>(scratch);
        p.tickets = reinterpret_cast<uint32_t *>(p.ws + tiles * (size_t)(BM * bn));
    }
    dim3 grid((unsigned)((p.b_rows + bn - 1) / bn), (unsigned)((p.a_rows + BM - 1) / BM), (unsigned)splits);
    if (p.swap) {
        switch (bn) {
            case 32: return launch_epi<32, true>(ta, tb, p, grid, st);
            case 64: return launch_epi<64, true>(ta, tb, p, grid, st);
            case 128: return launch_epi<128, true>(ta, tb, p, grid, st);
        }
    } else if (bn == 128) {
        return launch_epi<128, false>(ta, tb, p, grid, st);
    }
    set_error("kx_gemm: unsupported tile width %d (swap=%d)", bn, p.swap);
    return KX_EINVAL;
}

}  // namespace kx

// Stand-alone GEMM entry point (tests, and any caller that wants the op by itself).
extern "C" int kx_gemm_bf16(const void *d_a, int64_t a_rows, int64_t lda, const void *d_b, int64_t b_rows, int64_t ldb,
                            int k, int epilogue, int swap, void *d_out, int64_t ldo, const float *d_bias, int splits,
                            void *stream) {
    using namespace kx;
    KX_REQUIRE(d_a && d_b && d_out, KX_EINVAL, "kx_gemm_bf16: NULL operand");
    KX_REQUIRE(a_rows > 0 && b_rows > 0 && k > 0 && k % 8 == 0, KX_EINVAL, "kx_gemm_bf16: bad shape (k must be a multiple of 8)");
    KX_REQUIRE(swap || b_rows % 8 == 0, KX_EINVAL, "kx_gemm_bf16: feature count must be a multiple of 8");
    KX_REQUIRE(epilogue == EPI_ARGMAX || ldo % 8 == 0, KX_EINVAL, "kx_gemm_bf16: ldo must be a multiple of 8 elements");
    KX_REQUIRE(epilogue != EPI_ARGMAX || splits <= 1, KX_EINVAL, "kx_gemm_bf16: the argmax epilogue does not split k");
    KX_REQUIRE(epilogue != EPI_QKV_ROPE, KX_EINVAL, "kx_gemm_bf16: epilogue %d is engine-internal", epilogue);
    KX_REQUIRE(epilogue != EPI_SWIGLU || (swap ? a_rows : b_rows) % 2 == 0, KX_EINVAL, "kx_gemm_bf16: SwiGLU needs (gate, up) pairs");
    cudaStream_t st = (cudaStream_t)stream;
    GemmParams p{};
    p.a_rows = (int)a_rows; p.b_rows = (int)b_rows; p.k = k; p.out = d_out; p.ldo = (int)ldo; p.bias = d_bias;
    p.swap = swap; p.epilogue = epilogue; p.splits = splits;
    CUtensorMap ta, tb;
    int rc = make_tmap_bf16_2d(&ta, d_a, a_rows, k, lda, BM);
    if (rc != KX_OK) return rc;
    const int pbn = swap ? 0 : gemm_pair_bn(a_rows, b_rows);
    if (pbn) {                                        // prefill: persistent CTA-pair kernel
        rc = require_local_device_memory(d_b, "kx_gemm_bf16: the B operand of the token-major kernel");
        if (rc != KX_OK) return rc;
        p.splits = 1;                                 // (its split-K mode writes per-slice partials: engine-internal)
        rc = make_tmap_bf16_2d(&tb, d_b, b_rows, k, ldb, pbn / 2);
        if (rc != KX_OK) return rc;
        return gemm_launch_pair(ta, tb, p, pbn, st);
    }
    const int bn = gemm_pick_bn(swap, b_rows);
    rc = make_tmap_bf16_2d(&tb, d_b, b_rows, k, ldb, bn);
    if (rc != KX_OK) return rc;
    void *scratch = nullptr;
    const size_t sb = gemm_scratch_bytes(p, bn);
    if (sb) {
        rc = stream_scratch(st, SCRATCH_GEMM, sb, &scratch, /*zero_on_alloc=*/true);
        if (rc != KX_OK) return rc;
    }
    return gemm_launch(ta, tb, p, bn, scratch, st);
}
