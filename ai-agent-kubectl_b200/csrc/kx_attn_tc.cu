// Prefill attention on the 5th-generation tensor cores, head_dim 64.
//
// One CTA = one (sequence, query head, block of 128 query rows).  For every 128-key tile up to the diagonal:
//   S = Q K^T      tcgen05.mma  M=128 (query rows) N=128 (keys) K=64, accumulator in TMEM columns [0,128)
//   softmax        one thread per query row: tcgen05.ld of its S row (two passes: max, then exp2 / sum),
//                  P written as bf16 into shared memory in the K-major 128-byte-swizzled operand layout
//   O_j = P V      tcgen05.mma  M=128 N=64 K=128 (keys), V consumed MN-major exactly as TMA lays the paged
//                  K/V rows down ([key][head_dim], 128 B per key), accumulator in TMEM columns [128,192)
//   O += O_j       in registers, after the usual rescale by exp2(m_old - m_new)
// Q, K, V arrive by TMA (Q from the token-major qkv buffer, K/V page by page through the block table).
// Two CTAs fit an SM (80 KB shared memory, 256 TMEM columns each), so one CTA's softmax overlaps the
// other's MMAs; inside a CTA the K tile of step j+1 is fetched while step j is in its softmax.
//
// Roles (192 threads): warp 0 lane 0 = TMA producer, warp 1 lane 0 = MMA issuer, warps 2-5 = softmax
// (TMEM lane quarter = warp % 4).  For head_dim 64, with or without a cached prefix; head_dim 128 takes the
// mma.sync kernel in kx_ops.cu (KX_ATTN_TC=0 forces that kernel everywhere).  Persistent: two CTAs per SM walk
// the (sequence, head, query block) items, barrier parities run on a tile counter that continues across items.
#include <cuda_bf16.h>
#include "kx_common.cuh"
#include "kx_sm100.cuh"
#include "kx_gemm.cuh"
#include "kx_ops.cuh"

namespace kx {

using namespace sm100;

namespace {

constexpr int kHD = 64, kQ = 128, kKeys = 128, kPage = KX_KV_PAGE;
constexpr int kTile = kQ * kHD * 2;                    // 16 KB: Q, K, V tiles and each 64-key half of P
constexpr int kSmem = 1024 + 5 * kTile + 256;          // Q, K, V, P (2 halves), barriers
constexpr int kTmemCols = 256;

enum { B_QFULL = 0, B_KFULL, B_VFULL, B_SDONE, B_ODONE, B_PFULL, B_SFREE, B_OFREE, B_COUNT };

// shared-memory descriptor of an MN-major bf16 operand with the 128-byte swizzle: 64 contiguous MN elements per
// K index (one 128 B row), 8 K indices per 1 KB atom, SBO = 1024 B between 8-row groups along K; LBO unused
// (a single 64-element MN block)
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor, kind::f16, bf16 inputs, fp32 accumulate; bit 16: B operand MN-major
__host__ __device__ constexpr uint32_t idesc_bf16(int M, int N, bool b_mn_major) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((b_mn_major ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ uint32_t pack2bf(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

__global__ void __launch_bounds__(192, 2)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                       __nv_bfloat16 *__restrict__ out, int out_ld, const int *__restrict__ block_table, int max_pages,
                       int S, int pos0, int n_q, int n_kv, float scale_log2, int n_qb, int n_items) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t *smem = smem_raw + (base - smem_u32(smem_raw));
    uint8_t *Qs = smem, *Ks = Qs + kTile, *Vs = Ks + kTile, *Ps = Vs + kTile;      // Ps: two 16 KB halves (keys 0-63, 64-127)
    uint64_t *bar = reinterpret_cast<uint64_t *>(Ps + 2 * kTile);
    uint32_t *tmem_holder = reinterpret_cast<uint32_t *>(bar + B_COUNT);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // persistent: CTA c takes work items c, c + gridDim.x, ...; item -> (sequence, head, query block), the long
    // (late) query blocks first.  Barrier parities run on a tile counter that continues across items, so the
    // next item's Q and first K tile are fetched while the current item is still in its last softmax.
    auto item_of = [&](int w, int &b, int &head, int &qb) {
        qb = n_qb - 1 - w / (n_q * (n_items / (n_q * n_qb)));
        const int rest = w % (n_q * (n_items / (n_q * n_qb)));
        b = rest / n_q;
        head = rest - b * n_q;
    };

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&tmQ);
        tma_prefetch_desc(&tmKV);
        for (int i = 0; i < B_COUNT; ++i) mbar_init(bar + i, (i == B_PFULL || i == B_SFREE || i == B_OFREE) ? 4 : 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_holder, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_holder;
    const uint32_t tmem_s = tmem, tmem_o = tmem + 128;

    if (warp == 0) {
        if (lane == 0) {
            int t = 0;                                           // tiles issued so far by this CTA
            for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
            int b, head, qb;
            item_of(w, b, head, qb);
            const int kvh = head / (n_q / n_kv), n_tiles = (pos0 + (qb + 1) * kQ + kKeys - 1) / kKeys;   // causal: keys 0 .. pos0 + (qb+1) 128 - 1
            const int *bt = block_table + (size_t)b * max_pages;
            for (int j = 0; j < n_tiles; ++j, ++t) {
                // K tile (and a new item's Q): free once the S MMAs of the previous tile have completed; V: once its PV MMAs have
                if (t > 0) mbar_wait(bar + B_SDONE, (t - 1) & 1);
                if (j == 0) {
                    mbar_expect_tx(bar + B_QFULL, kTile);
                    tma_load_2d(Qs, &tmQ, bar + B_QFULL, head * kHD, b * S + qb * kQ);
                }
                mbar_expect_tx(bar + B_KFULL, kTile);
                for (int i = 0; i < kKeys / kPage; ++i) {
                    int pi = (j * kKeys) / kPage + i;
                    if (pi >= max_pages) pi = max_pages - 1;     // keys past the sequence: masked anyway
                    tma_load_2d(Ks + i * kPage * kHD * 2, &tmKV, bar + B_KFULL, 0, ((bt[pi] * 2 + 0) * n_kv + kvh) * kPage);
                }
                if (t > 0) mbar_wait(bar + B_ODONE, (t - 1) & 1);
                mbar_expect_tx(bar + B_VFULL, kTile);
                for (int i = 0; i < kKeys / kPage; ++i) {
                    int pi = (j * kKeys) / kPage + i;
                    if (pi >= max_pages) pi = max_pages - 1;
                    tma_load_2d(Vs + i * kPage * kHD * 2, &tmKV, bar + B_VFULL, 0, ((bt[pi] * 2 + 1) * n_kv + kvh) * kPage);
                }
            }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc_s = idesc_bf16(kQ, kKeys, false), idesc_o = idesc_bf16(kQ, kHD, true);
            const uint32_t q_u = base, k_u = q_u + kTile, v_u = k_u + kTile, p_u = v_u + kTile;
            int t = 0, it = 0;
            for (int w = blockIdx.x; w < n_items; w += gridDim.x, ++it) {
            int b, head, qb;
            item_of(w, b, head, qb);
            const int n_tiles = (pos0 + (qb + 1) * kQ + kKeys - 1) / kKeys;
            mbar_wait(bar + B_QFULL, it & 1);
            for (int j = 0; j < n_tiles; ++j, ++t) {
                mbar_wait(bar + B_KFULL, t & 1);
                if (t > 0) mbar_wait(bar + B_SFREE, (t - 1) & 1);           // the softmax threads have read S of the previous tile
                tc_fence_after();
                const uint64_t dq = umma_desc_sw128(q_u), dk = umma_desc_sw128(k_u);
#pragma unroll
                for (int k = 0; k < kHD / 16; ++k) umma_bf16(tmem_s, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
                umma_commit(bar + B_SDONE);
                mbar_wait(bar + B_PFULL, t & 1);                            // P of this tile is in shared memory
                mbar_wait(bar + B_VFULL, t & 1);
                if (t > 0) mbar_wait(bar + B_OFREE, (t - 1) & 1);           // the previous O tile has been read out of TMEM
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < kKeys / 16; ++k) {
                    // A = P: K-major, 64 keys per 16 KB half, 32 B per 16-key step inside a half
                    const uint64_t dp = umma_desc_sw128(p_u + (k >> 2) * kTile) + 2 * (k & 3);
                    // B = V: MN-major, 16 keys = two 1 KB atoms = 2048 B per step
                    const uint64_t dv = umma_desc_mn_sw128(v_u + k * 16 * kHD * 2);
                    umma_bf16(tmem_o, dp, dv, idesc_o, k != 0);
                }
                umma_commit(bar + B_ODONE);
            }
            }
        }
    } else {
        // ================= softmax: one thread per query row =================
        const int q = warp & 3, r = q * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        int t = 0;
        for (int w = blockIdx.x; w < n_items; w += gridDim.x) {
        int b, head, qb;
        item_of(w, b, head, qb);
        const int n_tiles = (pos0 + (qb + 1) * kQ + kKeys - 1) / kKeys;
        const int qpos = qb * kQ + r;                                   // row inside the sequence's new tokens; absolute position pos0 + qpos
        float o[kHD];
#pragma unroll
        for (int i = 0; i < kHD; ++i) o[i] = 0.f;
        float m = -INFINITY, l = 0.f;
        for (int j = 0; j < n_tiles; ++j, ++t) {
            mbar_wait(bar + B_SDONE, t & 1);
            tc_fence_after();
            const int klim = pos0 + qpos - j * kKeys;                       // keys 0 .. klim of this tile are visible (causal; >= 127: all)
            float mx = m;
#pragma unroll 1
            for (int c = 0; c < kKeys; c += 32) {                           // pass 1: row maximum
                float v[32];
                tmem_ld_32x32(tmem_s + lane_addr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) mx = fmaxf(mx, c + i <= klim ? v[i] : -INFINITY);
            }
            const float corr = ex2((m - mx) * scale_log2), msc = mx * scale_log2;
            l *= corr;
            m = mx;
#pragma unroll 1
            for (int c = 0; c < kKeys; c += 32) {                           // pass 2: P = exp2(S scale - m scale), row sum, bf16 to smem
                float v[32];
                tmem_ld_32x32(tmem_s + lane_addr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                tmem_ld_wait();
                if (c + 32 >= kKeys) {                                      // S has been read: the next tile's MMAs may overwrite it
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar + B_SFREE);
                }
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    const float p0 = c + i <= klim ? ex2(v[i] * scale_log2 - msc) : 0.f;
                    const float p1 = c + i + 1 <= klim ? ex2(v[i + 1] * scale_log2 - msc) : 0.f;
                    l += p0 + p1;
                    pk[i >> 1] = pack2bf(p0, p1);
                }
                uint8_t *prow = Ps + (c >> 6) * kTile + r * 128;
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const int chunk = ((c & 63) >> 3) + t;
                    *reinterpret_cast<uint4 *>(prow + ((chunk ^ (r & 7)) << 4)) = make_uint4(pk[4 * t], pk[4 * t + 1], pk[4 * t + 2], pk[4 * t + 3]);
                }
            }
            fence_proxy_async_smem();                                       // P: generic-proxy writes -> async-proxy (MMA) reads
            __syncwarp();
            if (lane == 0) mbar_arrive(bar + B_PFULL);
            mbar_wait(bar + B_ODONE, t & 1);                                // O_j = P V is in TMEM
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < kHD; c += 32) {
                float v[32];
                tmem_ld_32x32(tmem_o + lane_addr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) o[c + i] = o[c + i] * corr + v[i];
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar + B_OFREE);
        }
        if (qpos < S) {
            const float inv = 1.0f / l;
            __nv_bfloat16 *orow = out + (size_t)(b * S + qpos) * out_ld + (size_t)head * kHD;
#pragma unroll
            for (int i = 0; i < kHD; i += 8)
                *reinterpret_cast<uint4 *>(orow + i) = make_uint4(pack2bf(o[i] * inv, o[i + 1] * inv), pack2bf(o[i + 2] * inv, o[i + 3] * inv),
                                                                 pack2bf(o[i + 4] * inv, o[i + 5] * inv), pack2bf(o[i + 6] * inv, o[i + 7] * inv));
        }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, kTmemCols);
}

}  // namespace

// head_dim 64; the S new tokens of every sequence sit at positions pos0 .. pos0 + S - 1 behind pos0 cached keys.
// `pages` is one layer's page pool of `num_pages` pages.
int op_attn_prefill_tc(const void *qkv, int qkv_ld, void *out, int out_ld, const void *pages, int num_pages,
                       const int *block_table, int max_pages, int B, int S, int pos0, int n_q, int n_kv, cudaStream_t st) {
    CUtensorMap tq, tkv;
    int rc = make_tmap_bf16_2d(&tq, qkv, (int64_t)B * S, qkv_ld, qkv_ld, kQ);
    if (rc != KX_OK) return rc;
    rc = make_tmap_bf16_2d(&tkv, pages, (int64_t)num_pages * 2 * n_kv * kPage, kHD, kHD, kPage);
    if (rc != KX_OK) return rc;
    static PerDeviceOnce once;
    if (once.first()) {
        KX_CUDA(cudaFuncSetAttribute(attn_prefill_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        once.done();
    }
    const float scale_log2 = (1.0f / sqrtf((float)kHD)) * 1.4426950408889634f;
    const int n_qb = (S + kQ - 1) / kQ, n_items = n_qb * n_q * B;
    static int sms_of[64];
    int dev = 0;
    KX_CUDA(cudaGetDevice(&dev));
    if (!sms_of[dev & 63]) KX_CUDA(cudaDeviceGetAttribute(&sms_of[dev & 63], cudaDevAttrMultiProcessorCount, dev));
    const int grid = n_items < 2 * sms_of[dev & 63] ? n_items : 2 * sms_of[dev & 63];     // two CTAs per SM
    attn_prefill_tc_kernel<<<grid, 192, kSmem, st>>>(tq, tkv, (__nv_bfloat16 *)out, out_ld, block_table, max_pages, S, pos0, n_q, n_kv,
                                                     scale_log2, n_qb, n_items);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

}  // namespace kx
