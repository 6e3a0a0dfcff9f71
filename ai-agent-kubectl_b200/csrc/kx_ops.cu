// Generator kernels other than the GEMM:
// embedding gather, RMSNorm, RoPE + paged-KV append, causal prefill attention
// (flash-style, tensor cores), decode attention over the paged KV cache, greedy argmax.
//
// Numerics contract (mirrored by oracle/cpu_model.py in emulate_bf16 mode): the
// residual stream is fp32; GEMM inputs, q/k/v, the KV cache and the attention output
// are bf16; all reductions accumulate in fp32.
//
// Paged KV cache, per layer: pages[num_pages][2 (k,v)][n_kv][16 tokens][head_dim] bf16,
// block_table[seq][i] = physical page holding tokens 16 i .. 16 i + 15 of that sequence.
#include <cuda_bf16.h>
#include "kx_common.cuh"
#include "kx_ops.cuh"
#include "kx_sm100.cuh"

namespace kx {

namespace {

constexpr int kPage = KX_KV_PAGE;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ---- embedding gather: x[m, :] = float(table[tokens[m], :]) -----------------
__global__ void embed_kernel(const int *__restrict__ tokens, const __nv_bfloat16 *__restrict__ table,
                             float *__restrict__ x, int H) {
    const int m = blockIdx.x;
    const __nv_bfloat16 *row = table + (size_t)tokens[m] * H;
    float *o = x + (size_t)m * H;
    for (int i = threadIdx.x * 8; i < H; i += blockDim.x * 8) {
        const uint4 q = *reinterpret_cast<const uint4 *>(row + i);
        const __nv_bfloat162 *h = reinterpret_cast<const __nv_bfloat162 *>(&q);
        float4 a, b;
        a.x = __low2float(h[0]); a.y = __high2float(h[0]); a.z = __low2float(h[1]); a.w = __high2float(h[1]);
        b.x = __low2float(h[2]); b.y = __high2float(h[2]); b.z = __low2float(h[3]); b.w = __high2float(h[3]);
        *reinterpret_cast<float4 *>(o + i) = a;
        *reinterpret_cast<float4 *>(o + i + 4) = b;
    }
}

// ---- RMSNorm: y = bf16((x * (1/sqrt(mean(x^2)+eps))) * g); one warp per row ----
__global__ void rmsnorm_kernel(const float *__restrict__ x, const __nv_bfloat16 *__restrict__ g,
                               __nv_bfloat16 *__restrict__ y, int M, int H, float eps, const int *__restrict__ row_map) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= M) return;
    const int lane = threadIdx.x & 31;
    const float *xr = x + (size_t)(row_map ? row_map[row] : row) * H;
    float ss = 0.f;
    for (int i = lane * 4; i < H; i += 128) {
        const float4 v = *reinterpret_cast<const float4 *>(xr + i);
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    ss = warp_sum(ss);
    const float r = 1.0f / sqrtf(ss / (float)H + eps);
    __nv_bfloat16 *yr = y + (size_t)row * H;
    for (int i = lane * 4; i < H; i += 128) {
        const float4 v = *reinterpret_cast<const float4 *>(xr + i);
        const uint2 gq = *reinterpret_cast<const uint2 *>(g + i);
        const __nv_bfloat162 *gh = reinterpret_cast<const __nv_bfloat162 *>(&gq);
        __nv_bfloat162 o0 = __floats2bfloat162_rn(__fmul_rn(__fmul_rn(v.x, r), __low2float(gh[0])),
                                                  __fmul_rn(__fmul_rn(v.y, r), __high2float(gh[0])));
        __nv_bfloat162 o1 = __floats2bfloat162_rn(__fmul_rn(__fmul_rn(v.z, r), __low2float(gh[1])),
                                                  __fmul_rn(__fmul_rn(v.w, r), __high2float(gh[1])));
        uint2 out;
        out.x = *reinterpret_cast<uint32_t *>(&o0);
        out.y = *reinterpret_cast<uint32_t *>(&o1);
        *reinterpret_cast<uint2 *>(yr + i) = out;
    }
}

// ---- RoPE (rotate-half pairs i, i + hd/2) on q and k, then append k, v to the cache ----
__global__ void rope_append_kernel(__nv_bfloat16 *__restrict__ qkv, int qkv_ld, const int *__restrict__ pos,
                                   const int *__restrict__ seq, const float2 *__restrict__ rope,
                                   __nv_bfloat16 *__restrict__ pages, const int *__restrict__ block_table,
                                   int max_pages, int n_q, int n_kv, int hd) {
    const int m = blockIdx.x;
    const int p = pos[m], s = seq[m];
    const int half = hd >> 1;
    __nv_bfloat16 *row = qkv + (size_t)m * qkv_ld;
    const int page = block_table[(size_t)s * max_pages + p / kPage];
    __nv_bfloat16 *kdst = pages + ((size_t)page * 2 * n_kv) * kPage * hd + (size_t)(p % kPage) * hd;
    __nv_bfloat16 *vdst = kdst + (size_t)n_kv * kPage * hd;
    const float2 *cs = rope + (size_t)p * half;
    const int n_rot = (n_q + n_kv) * half;
    for (int i = threadIdx.x; i < n_rot; i += blockDim.x) {
        const int head = i / half, d = i - head * half;
        __nv_bfloat16 *h = row + head * hd;
        const float x1 = __bfloat162float(h[d]), x2 = __bfloat162float(h[d + half]);
        const float2 c = cs[d];
        // no FMA contraction: the CPU reference computes x1*cos - x2*sin with separate roundings
        const __nv_bfloat16 o1 = __float2bfloat16(__fsub_rn(__fmul_rn(x1, c.x), __fmul_rn(x2, c.y)));
        const __nv_bfloat16 o2 = __float2bfloat16(__fadd_rn(__fmul_rn(x2, c.x), __fmul_rn(x1, c.y)));
        if (head < n_q) {
            h[d] = o1; h[d + half] = o2;
        } else {
            __nv_bfloat16 *kd = kdst + (size_t)(head - n_q) * kPage * hd;
            kd[d] = o1; kd[d + half] = o2;
        }
    }
    const __nv_bfloat16 *vsrc = row + (n_q + n_kv) * hd;
    for (int i = threadIdx.x; i < n_kv * hd; i += blockDim.x) {
        const int head = i / hd, d = i - head * hd;
        vdst[(size_t)head * kPage * hd + d] = vsrc[i];
    }
}

// ---- decode attention: one query token per sequence over its paged KV ----------
// grid (n_kv, B), 256 threads.  Latency, not arithmetic, decides this kernel, so the CTA
// first puts its whole K and V working set in flight: one warp issues a 1-D bulk copy
// (cp.async.bulk -> mbarrier) per cache page into shared memory, up to 32 KB of K and
// 32 KB of V per chunk, while the other warps apply RoPE to the new token's q heads (of
// this GQA group) and k head and append k, v to the cache (no separate launch).  All
// G = n_q / n_kv query heads are then served from that one copy of K and V: lanes are split
// (key, 8-dim chunk) so every shared-memory access is a conflict-free 16-byte load, dot
// products finish with log2(HD/8) shuffles, softmax is online across chunks.
constexpr int kDecThreads = 256, kDecWarps = 8;

template <int HD, int G>
__global__ void __launch_bounds__(kDecThreads, (G <= 4 ? 3 : 1))
attn_decode_kernel(const __nv_bfloat16 *__restrict__ qkv, int qkv_ld, __nv_bfloat16 *__restrict__ out, int out_ld,
                   __nv_bfloat16 *__restrict__ pages, const int *__restrict__ block_table, int max_pages,
                   const int *__restrict__ pos, const float2 *__restrict__ rope, int n_q, int n_kv, float scale) {
    extern __shared__ __align__(128) uint8_t dsm[];
    constexpr int LPK = HD / 8;                       // lanes per key
    constexpr int KPW = 32 / LPK;                     // keys per warp iteration
    constexpr int HALF = HD / 2;
    constexpr int CH = 32768 / (HD * 2);              // tokens per chunk (32 KB of K, 32 KB of V)
    constexpr int PAGE_BYTES = kPage * HD * 2;
    uint64_t *bar = reinterpret_cast<uint64_t *>(dsm);
    __nv_bfloat16 *Ks = reinterpret_cast<__nv_bfloat16 *>(dsm + 128);
    __nv_bfloat16 *Vs = Ks + CH * HD;
    float *qs = reinterpret_cast<float *>(Vs + CH * HD);   // [G][HD]
    float *sc = qs + G * HD;                               // [G][CH]
    float *ml = sc + G * CH;                               // [G] running max, [G] running sum, [G] rescale
    __nv_bfloat16 *knew = reinterpret_cast<__nv_bfloat16 *>(ml + ((3 * G + 3) & ~3));   // [HD] new k, [HD] new v (16 B aligned)
    __nv_bfloat16 *vnew = knew + HD;
    float *part = reinterpret_cast<float *>(Ks);           // [kDecWarps][G][HD], aliases K/V after the last chunk

    const int kvh = blockIdx.x, b = blockIdx.y;
    const int p = pos[b], T = p + 1;                  // keys 0 .. p (p = the new token)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int *bt = block_table + (size_t)b * max_pages;
    const __nv_bfloat16 *row = qkv + (size_t)b * qkv_ld;

    if (tid == 0) { sm100::mbar_init(bar, 1); sm100::fence_barrier_init(); }
    if (tid < 3 * G) ml[tid] = tid < G ? -INFINITY : 0.f;
    __syncthreads();

    auto issue_chunk = [&](int c0) {                  // warp 0: one bulk copy per page for K and for V
        const int npages = (min(T, c0 + CH) - c0 + kPage - 1) / kPage;
        if (lane == 0) sm100::mbar_expect_tx(bar, (uint32_t)(2 * npages * PAGE_BYTES));
        __syncwarp();
        for (int i = lane; i < 2 * npages; i += 32) {
            const int pg = i >> 1, which = i & 1;
            const __nv_bfloat16 *src = pages + (((size_t)bt[c0 / kPage + pg] * 2 + which) * n_kv + kvh) * kPage * HD;
            sm100::bulk_load((which ? Vs : Ks) + (size_t)pg * kPage * HD, src, PAGE_BYTES, bar);
        }
    };
    if (warp == 0) issue_chunk(0);

    // RoPE on q (-> smem, rounded to bf16 like the stored q of the prefill path) and on k; append k, v
    {
        const float2 *cs = rope + (size_t)p * HALF;
        __nv_bfloat16 *kdst = pages + (((size_t)bt[p / kPage] * 2) * n_kv + kvh) * kPage * HD + (size_t)(p % kPage) * HD;
        __nv_bfloat16 *vdst = kdst + (size_t)n_kv * kPage * HD;
        for (int i = tid; i < (G + 1) * HALF; i += kDecThreads) {
            const int h = i / HALF, d = i - h * HALF;
            const __nv_bfloat16 *src = h < G ? row + (size_t)(kvh * G + h) * HD : row + (size_t)(n_q + kvh) * HD;
            const float x1 = __bfloat162float(src[d]), x2 = __bfloat162float(src[d + HALF]);
            const float2 c = cs[d];
            const __nv_bfloat16 o1 = __float2bfloat16(__fsub_rn(__fmul_rn(x1, c.x), __fmul_rn(x2, c.y)));
            const __nv_bfloat16 o2 = __float2bfloat16(__fadd_rn(__fmul_rn(x2, c.x), __fmul_rn(x1, c.y)));
            if (h < G) { qs[h * HD + d] = __bfloat162float(o1); qs[h * HD + d + HALF] = __bfloat162float(o2); }
            else { kdst[d] = o1; kdst[d + HALF] = o2; knew[d] = o1; knew[d + HALF] = o2; }
        }
        const __nv_bfloat16 *vsrc = row + (size_t)(n_q + n_kv + kvh) * HD;
        for (int i = tid; i < HD; i += kDecThreads) { const __nv_bfloat16 x = vsrc[i]; vdst[i] = x; vnew[i] = x; }
    }
    __syncthreads();

    const int kq = lane / LPK, dc = lane % LPK;
    float qr[G][8], o[G][8];
#pragma unroll
    for (int g = 0; g < G; ++g)
#pragma unroll
        for (int u = 0; u < 8; ++u) { qr[g][u] = qs[g * HD + dc * 8 + u]; o[g][u] = 0.f; }

    uint32_t phase = 0;
    for (int c0 = 0; c0 < T; c0 += CH) {
        const int n = min(T, c0 + CH) - c0;           // valid keys in this chunk
        const int npad = (n + 31) & ~31;
        sm100::mbar_wait(bar, phase);
        phase ^= 1;
        // the new token's row comes from this CTA, not from the copy (which raced with the append)
        if (p >= c0 && p < c0 + CH && tid < HD / 8) {
            reinterpret_cast<uint4 *>(Ks + (size_t)(p - c0) * HD)[tid] = reinterpret_cast<const uint4 *>(knew)[tid];
            reinterpret_cast<uint4 *>(Vs + (size_t)(p - c0) * HD)[tid] = reinterpret_cast<const uint4 *>(vnew)[tid];
        }
        __syncthreads();

        // scores
        for (int t0 = warp * KPW; t0 < npad; t0 += kDecWarps * KPW) {
            const int t = t0 + kq;
            float acc[G];
#pragma unroll
            for (int g = 0; g < G; ++g) acc[g] = 0.f;
            if (t < n) {
                const uint4 kq4 = *reinterpret_cast<const uint4 *>(Ks + (size_t)t * HD + dc * 8);
                const __nv_bfloat162 *kh = reinterpret_cast<const __nv_bfloat162 *>(&kq4);
                float kf[8];
#pragma unroll
                for (int e = 0; e < 4; ++e) { kf[2 * e] = __low2float(kh[e]); kf[2 * e + 1] = __high2float(kh[e]); }
#pragma unroll
                for (int g = 0; g < G; ++g)
#pragma unroll
                    for (int e = 0; e < 8; ++e) acc[g] = fmaf(qr[g][e], kf[e], acc[g]);
            }
#pragma unroll
            for (int g = 0; g < G; ++g) {
#pragma unroll
                for (int ofs = LPK / 2; ofs > 0; ofs >>= 1) acc[g] += __shfl_xor_sync(0xffffffffu, acc[g], ofs);
            }
            if (dc == 0) {
#pragma unroll
                for (int g = 0; g < G; ++g) sc[g * CH + t] = t < n ? acc[g] * scale : -INFINITY;
            }
        }
        __syncthreads();

        // online softmax bookkeeping per query head (warp w handles heads w, w + kDecWarps, ...)
        for (int g = warp; g < G; g += kDecWarps) {
            float mx = -INFINITY;
            for (int t = lane; t < npad; t += 32) mx = fmaxf(mx, sc[g * CH + t]);
            mx = warp_max(mx);
            const float m_old = ml[g], m_new = fmaxf(m_old, mx);
            float sum = 0.f;
            for (int t = lane; t < npad; t += 32) {
                const float e = __expf(sc[g * CH + t] - m_new);
                sc[g * CH + t] = e;
                sum += e;
            }
            sum = warp_sum(sum);
            if (lane == 0) {
                const float f = __expf(m_old - m_new);       // 0 on the first chunk (m_old = -inf)
                ml[g] = m_new;
                ml[G + g] = ml[G + g] * f + sum;
                ml[2 * G + g] = f;
            }
        }
        __syncthreads();

        // O = O * f + P V
#pragma unroll
        for (int g = 0; g < G; ++g) {
            const float f = ml[2 * G + g];
#pragma unroll
            for (int u = 0; u < 8; ++u) o[g][u] *= f;
        }
        for (int t0 = warp * KPW; t0 < n; t0 += kDecWarps * KPW) {
            const int t = t0 + kq;
            if (t < n) {
                const uint4 vq4 = *reinterpret_cast<const uint4 *>(Vs + (size_t)t * HD + dc * 8);
                const __nv_bfloat162 *vh = reinterpret_cast<const __nv_bfloat162 *>(&vq4);
                float vf[8];
#pragma unroll
                for (int e = 0; e < 4; ++e) { vf[2 * e] = __low2float(vh[e]); vf[2 * e + 1] = __high2float(vh[e]); }
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    const float pg = sc[g * CH + t];
#pragma unroll
                    for (int e = 0; e < 8; ++e) o[g][e] = fmaf(pg, vf[e], o[g][e]);
                }
            }
        }
        __syncthreads();                              // K/V/sc of this chunk fully consumed
        if (c0 + CH < T && warp == 0) issue_chunk(c0 + CH);
    }

    // reduce over the KPW key slots of the warp, then over the warps
#pragma unroll
    for (int g = 0; g < G; ++g)
#pragma unroll
        for (int u = 0; u < 8; ++u)
#pragma unroll
            for (int ofs = LPK; ofs < 32; ofs <<= 1) o[g][u] += __shfl_xor_sync(0xffffffffu, o[g][u], ofs);
    if (kq == 0) {
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int u = 0; u < 8; ++u) part[(warp * G + g) * HD + dc * 8 + u] = o[g][u];
    }
    __syncthreads();
    for (int i = tid; i < G * HD; i += kDecThreads) {
        const int g = i / HD, d = i - g * HD;
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < kDecWarps; ++w) s += part[(w * G + g) * HD + d];
        out[(size_t)b * out_ld + (size_t)(kvh * G + g) * HD + d] = __float2bfloat16(s / ml[G + g]);
    }
}

// ---- prefill attention: causal flash attention, GQA, K/V from the paged cache ----
// grid (ceil(S/64), n_q, B), 128 threads; warp w owns query rows 16w .. 16w+15 of the
// 64-row block.  QK^T and PV run on tensor cores (mma.sync m16n8k16 bf16, fp32 acc) with
// the online softmax in fp32 registers.
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

template <int HD>
__device__ __forceinline__ uint32_t swz(int row, int chunk) {       // byte offset of a 16 B chunk
    return (uint32_t)(row * HD * 2 + ((chunk ^ (row & 7)) << 4));
}

template <int HD>
__global__ void __launch_bounds__(128)
attn_prefill_kernel(const __nv_bfloat16 *__restrict__ qkv, int qkv_ld, __nv_bfloat16 *__restrict__ out, int out_ld,
                    const __nv_bfloat16 *__restrict__ pages, const int *__restrict__ block_table, int max_pages,
                    int S, int pos0, int n_q, int n_kv, float scale_log2) {
    extern __shared__ __align__(128) uint8_t smem_attn[];
    constexpr int CH = HD / 8;                       // 16 B chunks per row
    constexpr int TILE = 64 * HD * 2;
    uint8_t *Qs = smem_attn, *Ks = Qs + TILE, *Vs = Ks + TILE;
    const uint32_t qs_u = (uint32_t)__cvta_generic_to_shared(Qs), ks_u = qs_u + TILE, vs_u = ks_u + TILE;
    const int qb = blockIdx.x, head = blockIdx.y, b = blockIdx.z;
    const int kvh = head / (n_q / n_kv);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int *bt = block_table + (size_t)b * max_pages;

    // Q tile -> smem (zero rows past S)
    for (int i = tid; i < 64 * CH; i += 128) {
        const int r = i / CH, c = i - r * CH;
        const int t = qb * 64 + r;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (t < S) v = *reinterpret_cast<const uint4 *>(qkv + (size_t)(b * S + t) * qkv_ld + (size_t)head * HD + c * 8);
        *reinterpret_cast<uint4 *>(Qs + swz<HD>(r, c)) = v;
    }
    __syncthreads();
    uint32_t qf[HD / 16][4];
#pragma unroll
    for (int ks = 0; ks < HD / 16; ++ks)
        ldsm_x4(qf[ks], qs_u + swz<HD>(warp * 16 + (lane & 15), 2 * ks + (lane >> 4)));

    float o[HD / 8][4];
#pragma unroll
    for (int j = 0; j < HD / 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
    float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
    const int g = lane >> 2, tq = lane & 3;
    const int qpos0 = pos0 + qb * 64 + warp * 16 + g;           // absolute position of row g (row g+8: +8)
    const int T = pos0 + S;
    const int last_key = min(T, pos0 + (qb + 1) * 64);          // keys needed by this query block
    const int nkb = (last_key + 63) / 64;

    for (int kb = 0; kb < nkb; ++kb) {
        __syncthreads();                                         // previous tile fully consumed
        for (int i = tid; i < 64 * CH; i += 128) {
            const int r = i / CH, c = i - r * CH;
            const int t = kb * 64 + r;
            uint4 kq = make_uint4(0, 0, 0, 0), vq = kq;
            if (t < T) {
                const __nv_bfloat16 *kp = pages + (((size_t)bt[t / kPage] * 2) * n_kv + kvh) * kPage * HD + (size_t)(t % kPage) * HD + c * 8;
                kq = *reinterpret_cast<const uint4 *>(kp);
                vq = *reinterpret_cast<const uint4 *>(kp + (size_t)n_kv * kPage * HD);
            }
            *reinterpret_cast<uint4 *>(Ks + swz<HD>(r, c)) = kq;
            *reinterpret_cast<uint4 *>(Vs + swz<HD>(r, c)) = vq;
        }
        __syncthreads();

        // S = Q K^T for 64 keys: 8 n-tiles of 8 keys
        float s[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) {
#pragma unroll
            for (int jp = 0; jp < 4; ++jp) {                     // pairs of n-tiles (16 keys)
                uint32_t kf[4];
                ldsm_x4(kf, ks_u + swz<HD>(jp * 16 + (lane & 7) + ((lane >> 4) << 3), 2 * ks + ((lane >> 3) & 1)));
                mma_bf16(s[2 * jp], qf[ks], kf[0], kf[1]);
                mma_bf16(s[2 * jp + 1], qf[ks], kf[2], kf[3]);
            }
        }
        // causal mask + online softmax (rows g and g+8)
        const bool diag = (kb + 1) * 64 > pos0 + qb * 64;       // block may contain masked keys
        float mx[2] = {mrow[0], mrow[1]};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int kpos = kb * 64 + j * 8 + 2 * tq + (e & 1);
                const int qpos = qpos0 + ((e >> 1) << 3);
                if (diag && kpos > qpos) s[j][e] = -INFINITY;
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
        uint32_t pf[4][4];                                       // P as A fragments, 4 k-steps of 16 keys
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float p0 = exp2f(s[j][0] * scale_log2 - msc[0]), p1 = exp2f(s[j][1] * scale_log2 - msc[0]);
            const float p2 = exp2f(s[j][2] * scale_log2 - msc[1]), p3 = exp2f(s[j][3] * scale_log2 - msc[1]);
            lrow[0] += p0 + p1;
            lrow[1] += p2 + p3;
            pf[j >> 1][(j & 1) * 2] = pack2(p0, p1);
            pf[j >> 1][(j & 1) * 2 + 1] = pack2(p2, p3);
        }
        // O += P V
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
            for (int dp = 0; dp < HD / 16; ++dp) {               // pairs of d n-tiles
                uint32_t vf[4];
                ldsm_x4_t(vf, vs_u + swz<HD>(kk * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), 2 * dp + (lane >> 4)));
                mma_bf16(o[2 * dp], pf[kk], vf[0], vf[1]);
                mma_bf16(o[2 * dp + 1], pf[kk], vf[2], vf[3]);
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 1);
        lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 2);
        lrow[r] = 1.0f / lrow[r];
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int t = qb * 64 + warp * 16 + g + r * 8;
        if (t >= S) continue;
        __nv_bfloat16 *orow = out + (size_t)(b * S + t) * out_ld + (size_t)head * HD + 2 * tq;
#pragma unroll
        for (int j = 0; j < HD / 8; ++j)
            *reinterpret_cast<uint32_t *>(orow + j * 8) = pack2(o[j][2 * r] * lrow[r], o[j][2 * r + 1] * lrow[r]);
    }
}

// ---- greedy argmax over fp32 logits; ties -> lowest index (torch.argmax) ------------
__global__ void __launch_bounds__(256)
argmax_kernel(const float *__restrict__ logits, int V, int ld, int *__restrict__ out, int *__restrict__ out2) {
    const int b = blockIdx.x;
    const float *row = logits + (size_t)b * ld;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x * 4; i < V; i += 256 * 4) {
        const float4 v = *reinterpret_cast<const float4 *>(row + i);
        const float a[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (a[u] > best || (a[u] == best && i + u < bi)) { best = a[u]; bi = i + u; }
    }
    __shared__ float sb[8];
    __shared__ int si[8];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
    }
    if ((threadIdx.x & 31) == 0) { sb[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w)
            if (sb[w] > best || (sb[w] == best && si[w] < bi)) { best = sb[w]; bi = si[w]; }
        out[b] = bi;
        if (out2) out2[b] = bi;
    }
}

// positions / bookkeeping for the decode loop, kept on the device so a step can be a CUDA graph
__global__ void advance_kernel(int *pos, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pos[i] += 1;
}

template <int HD>
int launch_decode_g(int G, dim3 grid, size_t smem, cudaStream_t st, const __nv_bfloat16 *qkv, int qkv_ld,
                    __nv_bfloat16 *out, int out_ld, __nv_bfloat16 *pages, const int *bt, int max_pages,
                    const int *pos, const float2 *rope, int n_q, int n_kv, float scale) {
#define KX_DEC(GV)                                                                                          \
    case GV: {                                                                                              \
        auto k = attn_decode_kernel<HD, GV>;                                                                \
        static size_t granted = 48 * 1024;   /* raised outside any stream capture by the first, eager call */  \
        if (smem > granted) {                                                                               \
            KX_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));        \
            granted = smem;                                                                                 \
        }                                                                                                   \
        k<<<grid, kDecThreads, smem, st>>>(qkv, qkv_ld, out, out_ld, pages, bt, max_pages, pos, rope, n_q, n_kv, scale); \
        break;                                                                                              \
    }
    switch (G) {
        KX_DEC(1) KX_DEC(2) KX_DEC(4) KX_DEC(7) KX_DEC(8)
        default:
            set_error("attention: unsupported GQA group size %d", G);
            return KX_EINVAL;
    }
#undef KX_DEC
    return KX_OK;
}

}  // namespace

int op_embed(const int *tokens, const void *table, float *x, int M, int H, cudaStream_t st) {
    embed_kernel<<<M, 128, 0, st>>>(tokens, (const __nv_bfloat16 *)table, x, H);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_rmsnorm(const float *x, const void *g, void *y, int M, int H, float eps, const int *row_map, cudaStream_t st) {
    rmsnorm_kernel<<<(M + 7) / 8, 256, 0, st>>>(x, (const __nv_bfloat16 *)g, (__nv_bfloat16 *)y, M, H, eps, row_map);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_rope_append(void *qkv, int qkv_ld, const int *pos, const int *seq, const float *rope, void *pages,
                   const int *block_table, int max_pages, int M, int n_q, int n_kv, int hd, cudaStream_t st) {
    rope_append_kernel<<<M, 128, 0, st>>>((__nv_bfloat16 *)qkv, qkv_ld, pos, seq, (const float2 *)rope,
                                          (__nv_bfloat16 *)pages, block_table, max_pages, n_q, n_kv, hd);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_attn_decode(const void *qkv, int qkv_ld, void *out, int out_ld, void *pages, const int *block_table,
                   int max_pages, const int *pos, const float *rope, int B, int max_T, int n_q, int n_kv, int hd,
                   cudaStream_t st) {
    const int G = n_q / n_kv;
    (void)max_T;                                      // any context length: K/V are consumed in 32 KB chunks
    const int ch = 32768 / (hd * 2);
    const size_t smem = 128 + (size_t)2 * ch * hd * 2 + sizeof(float) * ((size_t)G * hd + (size_t)G * ch + 3 * G) + 4 * hd + 64;
    const float scale = 1.0f / sqrtf((float)hd);
    dim3 grid(n_kv, B);
    int rc;
    if (hd == 64)
        rc = launch_decode_g<64>(G, grid, smem, st, (const __nv_bfloat16 *)qkv, qkv_ld, (__nv_bfloat16 *)out, out_ld,
                                 (__nv_bfloat16 *)pages, block_table, max_pages, pos, (const float2 *)rope, n_q, n_kv, scale);
    else if (hd == 128)
        rc = launch_decode_g<128>(G, grid, smem, st, (const __nv_bfloat16 *)qkv, qkv_ld, (__nv_bfloat16 *)out, out_ld,
                                  (__nv_bfloat16 *)pages, block_table, max_pages, pos, (const float2 *)rope, n_q, n_kv, scale);
    else {
        set_error("attention: head_dim %d not supported (64 or 128)", hd);
        return KX_EINVAL;
    }
    if (rc != KX_OK) return rc;
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_attn_prefill(const void *qkv, int qkv_ld, void *out, int out_ld, const void *pages, const int *block_table,
                    int max_pages, int B, int S, int pos0, int n_q, int n_kv, int hd, cudaStream_t st) {
    const float scale_log2 = (1.0f / sqrtf((float)hd)) * 1.4426950408889634f;
    dim3 grid((S + 63) / 64, n_q, B);
    if (hd == 64) {
        attn_prefill_kernel<64><<<grid, 128, 3 * 64 * 64 * 2, st>>>((const __nv_bfloat16 *)qkv, qkv_ld, (__nv_bfloat16 *)out, out_ld,
            (const __nv_bfloat16 *)pages, block_table, max_pages, S, pos0, n_q, n_kv, scale_log2);
    } else if (hd == 128) {
        auto k = attn_prefill_kernel<128>;
        static bool granted = false;
        if (!granted) { KX_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 64 * 128 * 2)); granted = true; }
        k<<<grid, 128, 3 * 64 * 128 * 2, st>>>((const __nv_bfloat16 *)qkv, qkv_ld, (__nv_bfloat16 *)out, out_ld,
            (const __nv_bfloat16 *)pages, block_table, max_pages, S, pos0, n_q, n_kv, scale_log2);
    } else {
        set_error("attention: head_dim %d not supported (64 or 128)", hd);
        return KX_EINVAL;
    }
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_argmax(const float *logits, int B, int V, int ld, int *out, int *out2, cudaStream_t st) {
    argmax_kernel<<<B, 256, 0, st>>>(logits, V, ld, out, out2);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_advance(int *pos, int n, cudaStream_t st) {
    advance_kernel<<<(n + 127) / 128, 128, 0, st>>>(pos, n);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

}  // namespace kx
