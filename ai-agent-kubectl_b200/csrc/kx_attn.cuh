// Device-side building blocks of the decode attention kernels (kx_ops.cu):
// cp.async / ldmatrix / mma.sync wrappers, the XOR-swizzled K/V tile layout, 16-byte RoPE, and the
// one-warp-per-(sequence, kv head) key loop.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "kx_ops.cuh"

namespace kx {
namespace attn {

constexpr int kPage = KX_KV_PAGE;

__device__ __forceinline__ void cp_async16_zfill(uint32_t smem_dst, const void *src, bool valid) {
    const int n = valid ? 16 : 0;                     // src-size 0: the 16 destination bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_dst), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

// RoPE on 8 consecutive rotate-half pairs (d, d + half), d = 8 c .. 8 c + 7, of one head: 16-byte loads,
// the same separately rounded products as rope_append_kernel (no FMA contraction).
__device__ __forceinline__ void rope8(const __nv_bfloat16 *__restrict__ src, int half, const float2 *__restrict__ cs, int c,
                                      uint4 &lo, uint4 &hi) {
    const uint4 a = *reinterpret_cast<const uint4 *>(src + c * 8), b = *reinterpret_cast<const uint4 *>(src + half + c * 8);
    const __nv_bfloat162 *ah = reinterpret_cast<const __nv_bfloat162 *>(&a), *bh = reinterpret_cast<const __nv_bfloat162 *>(&b);
    uint32_t o1[4], o2[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const float4 cc = *reinterpret_cast<const float4 *>(cs + c * 8 + 2 * t);     // (cos, sin) of two dims
        const float x1a = __low2float(ah[t]), x1b = __high2float(ah[t]), x2a = __low2float(bh[t]), x2b = __high2float(bh[t]);
        o1[t] = pack2(__fsub_rn(__fmul_rn(x1a, cc.x), __fmul_rn(x2a, cc.y)), __fsub_rn(__fmul_rn(x1b, cc.z), __fmul_rn(x2b, cc.w)));
        o2[t] = pack2(__fadd_rn(__fmul_rn(x2a, cc.x), __fmul_rn(x1a, cc.y)), __fadd_rn(__fmul_rn(x2b, cc.z), __fmul_rn(x1b, cc.w)));
    }
    lo = make_uint4(o1[0], o1[1], o1[2], o1[3]);
    hi = make_uint4(o2[0], o2[1], o2[2], o2[3]);
}

template <int HD>
__device__ __forceinline__ uint32_t swz(int row, int chunk) {       // byte offset of a 16 B chunk
    return (uint32_t)(row * HD * 2 + ((chunk ^ (row & 7)) << 4));
}

__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// shared memory of one warp of the warp-per-pair kernel: Q tile, NS stages of (K, V), the new token's k and v
template <int HD, int KB, int NS> struct WarpAttnSmem {
    static constexpr int QT = 16 * HD * 2, ST = KB * HD * 2, NEWB = 2 * HD * 2;
    static constexpr int BYTES = QT + NS * 2 * ST + NEWB;
};

// request keys key0 + blk KB ... of one (sequence, kv head) pair into ring stage blk % NS (keys >= T zero-filled)
template <int HD, int KB, int NS>
__device__ __forceinline__ void warp_attn_fetch(uint32_t ring_u, int blk, int key0, int T, int p, const __nv_bfloat16 *__restrict__ pages,
                                                const int *__restrict__ bt, int n_kv, int kvh, int lane) {
    constexpr int CH = HD / 8, ST = KB * HD * 2;
    const uint32_t kdst = ring_u + (uint32_t)(blk % NS) * 2 * ST, vdst = kdst + ST;
    const int c0 = key0 + blk * KB;
#pragma unroll
    for (int i = lane; i < KB * CH; i += 32) {
        const int r = i / CH, c = i - r * CH;
        const int t = c0 + r;
        const bool ok = t < T;
        const int tt = ok ? t : p;
        const __nv_bfloat16 *kp = pages + (((size_t)bt[tt / kPage] * 2) * n_kv + kvh) * kPage * HD + (size_t)(tt % kPage) * HD + c * 8;
        cp_async16_zfill(kdst + swz<HD>(r, c), kp, ok);
        cp_async16_zfill(vdst + swz<HD>(r, c), kp + (size_t)n_kv * kPage * HD, ok);
    }
}

// The key loop of one (sequence, kv head) pair owned by one warp: Q tile (rows 0..G-1 = the rotated query heads),
// knew / vnew (the new token's k, v) are in the warp's shared memory and the first NS-1 stages have been
// requested (one committed cp.async group each).  Streams the keys key0 .. T-1, merges with n_part
// shared-prefix partials (rows prow0 + h of part_o / part_ml) and stores the G normalised rows at out_pair.
// The loop alone: keys key0 .. T-1 in nblk blocks (the first NS-1 already requested); leaves the unnormalised
// accumulator rows o (rows g and g + 8 of the 16-row tile), their running maxima and (quad-reduced) sums.
template <int HD, int G, int KB, int NS>
__device__ __forceinline__ void warp_attn_loop(uint8_t *Qs, int nblk, int key0, int T, int p, const __nv_bfloat16 *__restrict__ pages,
                                               const int *__restrict__ bt, int n_kv, int kvh, int lane, float scale_log2,
                                               float (&o)[HD / 8][4], float (&mrow)[2], float (&lrow)[2], bool has_new = true) {
    constexpr int CH = HD / 8;
    constexpr int QT = 16 * HD * 2, ST = KB * HD * 2;
    const uint32_t qs_u = (uint32_t)__cvta_generic_to_shared(Qs), ring_u = qs_u + QT;
    const __nv_bfloat16 *knew = reinterpret_cast<const __nv_bfloat16 *>(Qs + QT + NS * 2 * ST), *vnew = knew + HD;
    uint32_t qf[HD / 16][4];
#pragma unroll
    for (int ks = 0; ks < HD / 16; ++ks) ldsm_x4(qf[ks], qs_u + swz<HD>(lane & 15, 2 * ks + (lane >> 4)));

#pragma unroll
    for (int j = 0; j < HD / 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
    mrow[0] = mrow[1] = -INFINITY;
    lrow[0] = lrow[1] = 0.f;
    const int tq = lane & 3;

    for (int blk = 0; blk < nblk; ++blk) {
        if (blk + NS - 1 < nblk) warp_attn_fetch<HD, KB, NS>(ring_u, blk + NS - 1, key0, T, p, pages, bt, n_kv, kvh, lane);
        cp_async_commit();
        cp_async_wait<NS - 1>();                      // block blk has landed (for this lane)
        __syncwarp();                                 // ... and for the whole warp
        const uint32_t ks_u = ring_u + (uint32_t)(blk % NS) * 2 * ST, vs_u = ks_u + ST;
        const int c0 = key0 + blk * KB;
        const int n = min(T - c0, KB);
        if (has_new && p >= c0 && p < c0 + KB) {      // the new token's row comes from this warp, not from the copy
            if (lane < CH) {
                uint8_t *ring = Qs + QT + (blk % NS) * 2 * ST;
                *reinterpret_cast<uint4 *>(ring + swz<HD>(p - c0, lane)) = reinterpret_cast<const uint4 *>(knew)[lane];
                *reinterpret_cast<uint4 *>(ring + ST + swz<HD>(p - c0, lane)) = reinterpret_cast<const uint4 *>(vnew)[lane];
            }
            __syncwarp();
        }
        float s[KB / 8][4];
#pragma unroll
        for (int j = 0; j < KB / 8; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) {
#pragma unroll
            for (int jp = 0; jp < KB / 16; ++jp) {
                uint32_t kf[4];
                ldsm_x4(kf, ks_u + swz<HD>(jp * 16 + (lane & 7) + ((lane >> 4) << 3), 2 * ks + ((lane >> 3) & 1)));
                mma_bf16(s[2 * jp], qf[ks], kf[0], kf[1]);
                mma_bf16(s[2 * jp + 1], qf[ks], kf[2], kf[3]);
            }
        }
        float mx[2] = {mrow[0], mrow[1]};
#pragma unroll
        for (int j = 0; j < KB / 8; ++j) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (j * 8 + 2 * tq + (e & 1) >= n) s[j][e] = -INFINITY;
                mx[e >> 1] = fmaxf(mx[e >> 1], s[j][e]);
            }
        }
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
        }
        float corr[2], msc[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            corr[r] = exp2f((mrow[r] - mx[r]) * scale_log2);
            mrow[r] = mx[r];
            msc[r] = mx[r] * scale_log2;
            lrow[r] *= corr[r];
        }
#pragma unroll
        for (int j = 0; j < HD / 8; ++j) { o[j][0] *= corr[0]; o[j][1] *= corr[0]; o[j][2] *= corr[1]; o[j][3] *= corr[1]; }
        uint32_t pf[KB / 16][4];
#pragma unroll
        for (int j = 0; j < KB / 8; ++j) {
            const float p0 = exp2f(s[j][0] * scale_log2 - msc[0]), p1 = exp2f(s[j][1] * scale_log2 - msc[0]);
            const float p2 = exp2f(s[j][2] * scale_log2 - msc[1]), p3 = exp2f(s[j][3] * scale_log2 - msc[1]);
            lrow[0] += p0 + p1;
            lrow[1] += p2 + p3;
            pf[j >> 1][(j & 1) * 2] = pack2(p0, p1);
            pf[j >> 1][(j & 1) * 2 + 1] = pack2(p2, p3);
        }
#pragma unroll
        for (int kk = 0; kk < KB / 16; ++kk) {
#pragma unroll
            for (int dp = 0; dp < HD / 16; ++dp) {
                uint32_t vf[4];
                ldsm_x4_t(vf, vs_u + swz<HD>(kk * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), 2 * dp + (lane >> 4)));
                mma_bf16(o[2 * dp], pf[kk], vf[0], vf[1]);
                mma_bf16(o[2 * dp + 1], pf[kk], vf[2], vf[3]);
            }
        }
        __syncwarp();                                 // every lane is done with this stage before it is refilled
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 1);
        lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 2);
    }
}

template <int HD, int G, int KB, int NS>
__device__ __forceinline__ void warp_attn_main(uint8_t *Qs, int nblk, int key0, int T, int p, const __nv_bfloat16 *__restrict__ pages,
                                               const int *__restrict__ bt, int n_kv, int kvh, int lane, float scale_log2,
                                               __nv_bfloat16 *__restrict__ out_pair, const float *__restrict__ part_o,
                                               const float *__restrict__ part_ml, int n_part, size_t part_rows, size_t prow0) {
    float o[HD / 8][4], mrow[2], lrow[2];
    warp_attn_loop<HD, G, KB, NS>(Qs, nblk, key0, T, p, pages, bt, n_kv, kvh, lane, scale_log2, o, mrow, lrow);
    const int g = lane >> 2, tq = lane & 3;
    // merge with the shared-prefix partials (if any), normalise, store the G real rows
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int h = g + 8 * r;
        if (h >= G) continue;
        const size_t prow = prow0 + h;
        float M = mrow[r];
        for (int c = 0; c < n_part; ++c) M = fmaxf(M, part_ml[(c * part_rows + prow) * 2]);
        const float fs = exp2f((mrow[r] - M) * scale_log2);
        float den = fs * lrow[r];
        float acc[HD / 8][2];
#pragma unroll
        for (int j = 0; j < HD / 8; ++j) { acc[j][0] = fs * o[j][2 * r]; acc[j][1] = fs * o[j][2 * r + 1]; }
        for (int c = 0; c < n_part; ++c) {
            const float2 ml = *reinterpret_cast<const float2 *>(part_ml + (c * part_rows + prow) * 2);
            const float f = exp2f((ml.x - M) * scale_log2);
            den += f * ml.y;
            const float *po = part_o + (c * part_rows + prow) * HD + 2 * tq;
#pragma unroll
            for (int j = 0; j < HD / 8; ++j) {
                const float2 v = *reinterpret_cast<const float2 *>(po + j * 8);
                acc[j][0] += f * v.x; acc[j][1] += f * v.y;
            }
        }
        const float inv = 1.0f / den;
        __nv_bfloat16 *orow = out_pair + (size_t)h * HD + 2 * tq;
#pragma unroll
        for (int j = 0; j < HD / 8; ++j)
            *reinterpret_cast<uint32_t *>(orow + j * 8) = pack2(acc[j][0] * inv, acc[j][1] * inv);
    }
}

}  // namespace attn
}  // namespace kx
