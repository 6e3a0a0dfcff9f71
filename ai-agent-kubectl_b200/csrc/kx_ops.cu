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
#include "kx_attn.cuh"

namespace kx {

namespace {

using namespace attn;

// Launch with programmatic stream serialization: the grid may start while its predecessor drains;
// the kernel itself calls griddepcontrol.wait before it touches the predecessor's output.
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
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

// Same, one CTA of 128 threads per row: for the few rows of a decode step one warp per row leaves
// the GPU idle and is latency bound; here the row is read once into registers.
template <int PER>   // float4 chunks per thread: H = 512 * PER
__global__ void __launch_bounds__(128)
rmsnorm_row_kernel(float *__restrict__ x, const __nv_bfloat16 *__restrict__ g, __nv_bfloat16 *__restrict__ y,
                   int H, float eps, const int *__restrict__ row_map, const float *__restrict__ part, int splits,
                   size_t split_stride, unsigned long long *trace) {
    asm volatile("griddepcontrol.launch_dependents;");
    if (threadIdx.x == 0) trace_min(trace, 0);
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0) trace_min(trace, 1);
    const int row = blockIdx.x, tid = threadIdx.x;
    const size_t src = (size_t)(row_map ? row_map[row] : row) * H;
    float *xr = x + src;
    float4 v[PER];
    float ss = 0.f;
#pragma unroll
    for (int u = 0; u < PER; ++u) {
        const int i = (u * 128 + tid) * 4;
        v[u] = i < H ? *reinterpret_cast<const float4 *>(xr + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        if (splits > 0 && i < H) {
            // pending split-K partial sums of the preceding row-parallel projection: residual += their sum,
            // slices added in index order (deterministic), and the residual row is written back
            float4 a = *reinterpret_cast<const float4 *>(part + src + i);
            for (int s0 = 1; s0 < splits; s0 += 4) {           // four slices' loads in flight, added in slice order
                float4 c[4];
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (s0 + j < splits) c[j] = *reinterpret_cast<const float4 *>(part + (size_t)(s0 + j) * split_stride + src + i);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (s0 + j < splits) { a.x += c[j].x; a.y += c[j].y; a.z += c[j].z; a.w += c[j].w; }
            }
            v[u].x += a.x; v[u].y += a.y; v[u].z += a.z; v[u].w += a.w;
            *reinterpret_cast<float4 *>(xr + i) = v[u];
        }
        ss += v[u].x * v[u].x + v[u].y * v[u].y + v[u].z * v[u].z + v[u].w * v[u].w;
    }
    ss = warp_sum(ss);
    __shared__ float wsum[4];
    if ((tid & 31) == 0) wsum[tid >> 5] = ss;
    __syncthreads();
    const float r = 1.0f / sqrtf((wsum[0] + wsum[1] + wsum[2] + wsum[3]) / (float)H + eps);
    __nv_bfloat16 *yr = y + (size_t)row * H;
#pragma unroll
    for (int u = 0; u < PER; ++u) {
        const int i = (u * 128 + tid) * 4;
        if (i >= H) continue;
        const uint2 gq = *reinterpret_cast<const uint2 *>(g + i);
        const __nv_bfloat162 *gh = reinterpret_cast<const __nv_bfloat162 *>(&gq);
        __nv_bfloat162 o0 = __floats2bfloat162_rn(__fmul_rn(__fmul_rn(v[u].x, r), __low2float(gh[0])),
                                                  __fmul_rn(__fmul_rn(v[u].y, r), __high2float(gh[0])));
        __nv_bfloat162 o1 = __floats2bfloat162_rn(__fmul_rn(__fmul_rn(v[u].z, r), __low2float(gh[1])),
                                                  __fmul_rn(__fmul_rn(v[u].w, r), __high2float(gh[1])));
        uint2 out;
        out.x = *reinterpret_cast<uint32_t *>(&o0);
        out.y = *reinterpret_cast<uint32_t *>(&o1);
        *reinterpret_cast<uint2 *>(yr + i) = out;
    }
    if (tid == 0) trace_max(trace, 3);
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

// ---- prefill attention: causal flash attention, GQA, K/V from the paged cache ----
// grid (ceil(S/64), n_q, B), 128 threads; warp w owns query rows 16w .. 16w+15 of the
// 64-row block.  QK^T and PV run on tensor cores (mma.sync m16n8k16 bf16, fp32 acc) with
// the online softmax in fp32 registers.
template <int HD>
__global__ void __launch_bounds__(128)
attn_prefill_kernel(const __nv_bfloat16 *__restrict__ qkv, int qkv_ld, __nv_bfloat16 *__restrict__ out, int out_ld,
                    const __nv_bfloat16 *__restrict__ pages, const int *__restrict__ block_table, int max_pages,
                    int S, int pos0, int n_q, int n_kv, float scale_log2) {
    extern __shared__ __align__(128) uint8_t smem_attn[];
    constexpr int CH = HD / 8;                       // 16 B chunks per row
    constexpr int TILE = 64 * HD * 2;
    // Q tile, then two (K, V) buffers: tile kb+1 is fetched with cp.async while tile kb is consumed
    const uint32_t qs_u = (uint32_t)__cvta_generic_to_shared(smem_attn);
    const int qb = blockIdx.x, head = blockIdx.y, b = blockIdx.z;
    const int kvh = head / (n_q / n_kv);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int *bt = block_table + (size_t)b * max_pages;
    const int T = pos0 + S;
    const int last_key = min(T, pos0 + (qb + 1) * 64);          // keys needed by this query block
    const int nkb = (last_key + 63) / 64;

    auto fetch_kv = [&](int kb) {
        const uint32_t kdst = qs_u + TILE + (uint32_t)(kb & 1) * 2 * TILE, vdst = kdst + TILE;
        for (int i = tid; i < 64 * CH; i += 128) {
            const int r = i / CH, c = i - r * CH;
            const int t = kb * 64 + r;
            const bool ok = t < T;
            const int tt = ok ? t : 0;
            const __nv_bfloat16 *kp = pages + (((size_t)bt[tt / kPage] * 2) * n_kv + kvh) * kPage * HD + (size_t)(tt % kPage) * HD + c * 8;
            cp_async16_zfill(kdst + swz<HD>(r, c), kp, ok);
            cp_async16_zfill(vdst + swz<HD>(r, c), kp + (size_t)n_kv * kPage * HD, ok);
        }
    };
    for (int i = tid; i < 64 * CH; i += 128) {                   // Q tile (rows past S are zero-filled)
        const int r = i / CH, c = i - r * CH;
        const int t = qb * 64 + r;
        const bool ok = t < S;
        cp_async16_zfill(qs_u + swz<HD>(r, c), qkv + (size_t)(b * S + (ok ? t : 0)) * qkv_ld + (size_t)head * HD + c * 8, ok);
    }
    fetch_kv(0);
    cp_async_commit();

    uint32_t qf[HD / 16][4];
    float o[HD / 8][4];
#pragma unroll
    for (int j = 0; j < HD / 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
    float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
    const int g = lane >> 2, tq = lane & 3;
    const int qpos0 = pos0 + qb * 64 + warp * 16 + g;           // absolute position of row g (row g+8: +8)

    for (int kb = 0; kb < nkb; ++kb) {
        if (kb + 1 < nkb) { fetch_kv(kb + 1); cp_async_commit(); cp_async_wait<1>(); }
        else cp_async_wait<0>();
        __syncthreads();                                         // tile kb (and Q) visible to every warp
        if (kb == 0) {
#pragma unroll
            for (int ks = 0; ks < HD / 16; ++ks)
                ldsm_x4(qf[ks], qs_u + swz<HD>(warp * 16 + (lane & 15), 2 * ks + (lane >> 4)));
        }
        const uint32_t ks_u = qs_u + TILE + (uint32_t)(kb & 1) * 2 * TILE, vs_u = ks_u + TILE;

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
        __syncthreads();                                         // buffer kb&1 may be refilled by the next iteration's prefetch
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

// ---- decode attention: one query token per sequence over its paged KV ----------
// grid (n_kv, B), 128 threads.  The CTA applies RoPE to the new token's q heads (of this
// GQA group) and k head and appends k, v to the cache (no separate launch); the G query
// heads become the rows of one 16-row MMA tile, so QK^T and PV for all heads of the group
// run on tensor cores (mma.sync m16n8k16 bf16) from ONE copy of K and V.  K/V rows are
// staged with 16-byte cp.async into XOR-swizzled shared memory (conflict-free ldmatrix),
// 256 keys per round; each warp owns a 64-key block of the round with its own online
// softmax state and the four partial results are merged at the end (flash decoding
// inside the CTA).
// NW warps x 64 keys per round: 8 warps hold a whole <= 512-key context in one round (few CTAs, latency
// matters); 4 warps / 256-key rounds keep three CTAs per SM resident (many CTAs, bandwidth matters)

template <int HD, int G, int NW>
__global__ void __launch_bounds__(NW * 32)
attn_decode_kernel(const __nv_bfloat16 *__restrict__ qkv, int qkv_ld, __nv_bfloat16 *__restrict__ out, int out_ld,
                   __nv_bfloat16 *__restrict__ pages, const int *__restrict__ block_table, int max_pages,
                   const int *__restrict__ pos, const float2 *__restrict__ rope, int n_q, int n_kv, float scale_log2,
                   int round_keys, int key0, const float *__restrict__ part_o, const float *__restrict__ part_ml, int n_part,
                   unsigned long long *trace) {
    extern __shared__ __align__(128) uint8_t smem_dec[];
    asm volatile("griddepcontrol.launch_dependents;");   // let a dependent GEMM start prefetching its weights
    if (threadIdx.x == 0) trace_min(trace, 0);
    asm volatile("griddepcontrol.wait;" ::: "memory");   // launched with programmatic serialization: qkv comes from the predecessor
    if (threadIdx.x == 0) trace_min(trace, 1);
    constexpr int CH = HD / 8;                        // 16 B chunks per row
    constexpr int HALF = HD / 2;
    const int ROUND = round_keys;                     // keys staged per round: a multiple of 64, at most 64 per warp
    constexpr int QT = 16 * HD * 2;
    const int KT = ROUND * HD * 2;
    uint8_t *Qs = smem_dec, *Ks = Qs + QT, *Vs = Ks + KT;
    __nv_bfloat16 *knew = reinterpret_cast<__nv_bfloat16 *>(Vs + KT), *vnew = knew + HD;
    const uint32_t qs_u = (uint32_t)__cvta_generic_to_shared(Qs), ks_u = qs_u + QT, vs_u = ks_u + KT;
    float *comb_m = reinterpret_cast<float *>(Ks);    // after the last round: [8][16] max, [8][16] sum, [8][G][HD] O
    float *comb_l = comb_m + NW * 16;
    float *comb_o = comb_l + NW * 16;

    const int kvh = blockIdx.x, b = blockIdx.y;
    const int p = pos[b], T = p + 1;                  // keys 0 .. p (p = the new token)
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int *bt = block_table + (size_t)b * max_pages;
    const __nv_bfloat16 *row = qkv + (size_t)b * qkv_ld;

    // RoPE: q heads -> Q tile rows 0..G-1 (bf16, like the q the prefill path stores), k -> cache; v -> cache
    {
        const float2 *cs = rope + (size_t)p * HALF;
        __nv_bfloat16 *kdst = pages + (((size_t)bt[p / kPage] * 2) * n_kv + kvh) * kPage * HD + (size_t)(p % kPage) * HD;
        __nv_bfloat16 *vdst = kdst + (size_t)n_kv * kPage * HD;
        for (int i = tid; i < (G + 1) * (HALF / 8); i += (NW * 32)) {
            const int h = i / (HALF / 8), c = i - h * (HALF / 8);
            const __nv_bfloat16 *src = h < G ? row + (size_t)(kvh * G + h) * HD : row + (size_t)(n_q + kvh) * HD;
            uint4 lo, hi;
            rope8(src, HALF, cs, c, lo, hi);
            if (h < G) {
                *reinterpret_cast<uint4 *>(Qs + swz<HD>(h, c)) = lo;
                *reinterpret_cast<uint4 *>(Qs + swz<HD>(h, c + HALF / 8)) = hi;
            } else {
                reinterpret_cast<uint4 *>(kdst)[c] = lo; reinterpret_cast<uint4 *>(kdst)[c + HALF / 8] = hi;
                reinterpret_cast<uint4 *>(knew)[c] = lo; reinterpret_cast<uint4 *>(knew)[c + HALF / 8] = hi;
            }
        }
        const __nv_bfloat16 *vsrc = row + (size_t)(n_q + n_kv + kvh) * HD;
        for (int i = tid; i < CH; i += (NW * 32)) {
            const uint4 x = reinterpret_cast<const uint4 *>(vsrc)[i];
            reinterpret_cast<uint4 *>(vdst)[i] = x; reinterpret_cast<uint4 *>(vnew)[i] = x;
        }
        for (int i = tid; i < (16 - G) * CH; i += (NW * 32))          // padding rows of the Q tile
            *reinterpret_cast<uint4 *>(Qs + swz<HD>(G + i / CH, i % CH)) = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    uint32_t qf[HD / 16][4];
#pragma unroll
    for (int ks = 0; ks < HD / 16; ++ks) ldsm_x4(qf[ks], qs_u + swz<HD>(lane & 15, 2 * ks + (lane >> 4)));

    float o[HD / 8][4];
#pragma unroll
    for (int j = 0; j < HD / 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
    float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
    const int g = lane >> 2, tq = lane & 3;

    // keys below key0 belong to a prefix shared by the whole batch: attn_prefix_kernel has already
    // reduced them to n_part partial (max, sum, O) triples per query head, merged in below
    for (int c0 = key0; c0 < T; c0 += ROUND) {
        const int n = min(T, c0 + ROUND) - c0;        // valid keys this round
        const int nfill = (n + 63) & ~63;             // whole 64-key blocks are read by ldmatrix
        for (int i = tid; i < nfill * CH; i += (NW * 32)) {
            const int r = i / CH, c = i - r * CH;
            if (r < n) {
                const int t = c0 + r;
                const __nv_bfloat16 *kp = pages + (((size_t)bt[t / kPage] * 2) * n_kv + kvh) * kPage * HD + (size_t)(t % kPage) * HD + c * 8;
                cp_async16(ks_u + swz<HD>(r, c), kp);
                cp_async16(vs_u + swz<HD>(r, c), kp + (size_t)n_kv * kPage * HD);
            } else {
                *reinterpret_cast<uint4 *>(Ks + swz<HD>(r, c)) = make_uint4(0, 0, 0, 0);
                *reinterpret_cast<uint4 *>(Vs + swz<HD>(r, c)) = make_uint4(0, 0, 0, 0);
            }
        }
        cp_async_wait_all();
        __syncthreads();
        // the new token's row comes from this CTA, not from the copy (which raced with the append)
        if (p >= c0 && p < c0 + ROUND && tid < CH) {
            *reinterpret_cast<uint4 *>(Ks + swz<HD>(p - c0, tid)) = reinterpret_cast<const uint4 *>(knew)[tid];
            *reinterpret_cast<uint4 *>(Vs + swz<HD>(p - c0, tid)) = reinterpret_cast<const uint4 *>(vnew)[tid];
        }
        __syncthreads();

        const int k0 = warp * 64;                     // this warp's 64-key block inside the round
        if (k0 < n) {
            float s[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
#pragma unroll
            for (int ks = 0; ks < HD / 16; ++ks) {
#pragma unroll
                for (int jp = 0; jp < 4; ++jp) {
                    uint32_t kf[4];
                    ldsm_x4(kf, ks_u + swz<HD>(k0 + jp * 16 + (lane & 7) + ((lane >> 4) << 3), 2 * ks + ((lane >> 3) & 1)));
                    mma_bf16(s[2 * jp], qf[ks], kf[0], kf[1]);
                    mma_bf16(s[2 * jp + 1], qf[ks], kf[2], kf[3]);
                }
            }
            float mx[2] = {mrow[0], mrow[1]};
#pragma unroll
            for (int j = 0; j < 8; ++j) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    if (k0 + j * 8 + 2 * tq + (e & 1) >= n) s[j][e] = -INFINITY;
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
            uint32_t pf[4][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float p0 = exp2f(s[j][0] * scale_log2 - msc[0]), p1 = exp2f(s[j][1] * scale_log2 - msc[0]);
                const float p2 = exp2f(s[j][2] * scale_log2 - msc[1]), p3 = exp2f(s[j][3] * scale_log2 - msc[1]);
                lrow[0] += p0 + p1;
                lrow[1] += p2 + p3;
                pf[j >> 1][(j & 1) * 2] = pack2(p0, p1);
                pf[j >> 1][(j & 1) * 2 + 1] = pack2(p2, p3);
            }
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
                for (int dp = 0; dp < HD / 16; ++dp) {
                    uint32_t vf[4];
                    ldsm_x4_t(vf, vs_u + swz<HD>(k0 + kk * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), 2 * dp + (lane >> 4)));
                    mma_bf16(o[2 * dp], pf[kk], vf[0], vf[1]);
                    mma_bf16(o[2 * dp + 1], pf[kk], vf[2], vf[3]);
                }
            }
        }
        __syncthreads();                              // K/V of this round fully consumed
    }

    // merge the warps' (max, sum, O) -- only rows < G are real query heads
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 1);
        lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 2);
    }
    if (tq == 0) {
        comb_m[warp * 16 + g] = mrow[0]; comb_m[warp * 16 + g + 8] = mrow[1];
        comb_l[warp * 16 + g] = lrow[0]; comb_l[warp * 16 + g + 8] = lrow[1];
    }
    if (g < G) {
#pragma unroll
        for (int j = 0; j < HD / 8; ++j) {
            comb_o[(warp * G + g) * HD + j * 8 + 2 * tq] = o[j][0];
            comb_o[(warp * G + g) * HD + j * 8 + 2 * tq + 1] = o[j][1];
        }
    }
    if (G > 8 && g + 8 < G) {
#pragma unroll
        for (int j = 0; j < HD / 8; ++j) {
            comb_o[(warp * G + g + 8) * HD + j * 8 + 2 * tq] = o[j][2];
            comb_o[(warp * G + g + 8) * HD + j * 8 + 2 * tq + 1] = o[j][3];
        }
    }
    __syncthreads();
    const size_t part_rows = (size_t)gridDim.y * n_q;                       // rows of one prefix partial: (sequence, q head)
    for (int i = tid; i < G * HD; i += (NW * 32)) {
        const int h = i / HD, d = i - h * HD;
        const size_t prow = (size_t)b * n_q + kvh * G + h;
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < NW; ++w) M = fmaxf(M, comb_m[w * 16 + h]);
        for (int c = 0; c < n_part; ++c) M = fmaxf(M, part_ml[(c * part_rows + prow) * 2]);
        float num = 0.f, den = 0.f;
#pragma unroll
        for (int w = 0; w < NW; ++w) {
            const float f = exp2f((comb_m[w * 16 + h] - M) * scale_log2);   // 0 for a warp that saw no keys
            num += f * comb_o[(w * G + h) * HD + d];
            den += f * comb_l[w * 16 + h];
        }
        for (int c = 0; c < n_part; ++c) {
            const float2 ml = *reinterpret_cast<const float2 *>(part_ml + (c * part_rows + prow) * 2);
            const float f = exp2f((ml.x - M) * scale_log2);
            num += f * part_o[(c * part_rows + prow) * HD + d];
            den += f * ml.y;
        }
        out[(size_t)b * out_ld + (size_t)(kvh * G + h) * HD + d] = __float2bfloat16(num / den);
    }
    if (tid == 0) trace_max(trace, 3);
}

// ---- decode attention, one WARP per (sequence, kv head) -------------------------------------
// For a wide batch there are more (sequence, kv head) pairs than the GPU has warps to spare, so splitting
// one pair's keys over a CTA (attn_decode_kernel) only adds barriers and a merge.  Here a warp owns a pair:
// it streams the pair's keys through its private NS-stage cp.async ring, KB keys per stage, with the same
// 16-row tensor-core tile for the G query heads, and no block-level synchronisation at all.  The first
// stages are requested before griddepcontrol.wait (cached K/V were written by earlier steps, pos by the
// step's first kernel, both complete before the predecessor GEMM could start).
template <int HD, int G, int KB, int NS>
__global__ void __launch_bounds__(128)
attn_decode_warp_kernel(const __nv_bfloat16 *__restrict__ qkv, int qkv_ld, __nv_bfloat16 *__restrict__ out, int out_ld,
                        __nv_bfloat16 *__restrict__ pages, const int *__restrict__ block_table, int max_pages,
                        const int *__restrict__ pos, const float2 *__restrict__ rope, int n_q, int n_kv, float scale_log2,
                        int key0, const float *__restrict__ part_o, const float *__restrict__ part_ml, int n_part, int n_tasks,
                        unsigned long long *trace) {
    extern __shared__ __align__(128) uint8_t smem_dw[];
    asm volatile("griddepcontrol.launch_dependents;");
    if (threadIdx.x == 0) trace_min(trace, 0);
    using W = WarpAttnSmem<HD, KB, NS>;
    constexpr int CH = HD / 8, HALF = HD / 2;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int task = blockIdx.x * 4 + warp;
    const bool live = task < n_tasks;
    const int b = live ? task / n_kv : 0, kvh = live ? task - b * n_kv : 0;
    uint8_t *Qs = smem_dw + warp * W::BYTES;
    const uint32_t ring_u = (uint32_t)__cvta_generic_to_shared(Qs) + W::QT;
    __nv_bfloat16 *knew = reinterpret_cast<__nv_bfloat16 *>(Qs + W::QT + NS * 2 * W::ST), *vnew = knew + HD;
    const int p = pos[b], T = p + 1;
    const int *bt = block_table + (size_t)b * max_pages;
    const int nblk = live ? (T - key0 + KB - 1) / KB : 0;
#pragma unroll
    for (int i = 0; i < NS - 1; ++i) {                // prologue: NS-1 stages in flight (empty groups keep the count uniform)
        if (i < nblk) warp_attn_fetch<HD, KB, NS>(ring_u, i, key0, T, p, pages, bt, n_kv, kvh, lane);
        cp_async_commit();
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");   // qkv comes from the predecessor
    if (threadIdx.x == 0) trace_min(trace, 1);
    if (!live) return;

    // RoPE: q heads -> Q tile rows 0..G-1, k -> cache and knew, v -> cache and vnew
    {
        const __nv_bfloat16 *row = qkv + (size_t)b * qkv_ld;
        const float2 *cs = rope + (size_t)p * HALF;
        __nv_bfloat16 *kdst = pages + (((size_t)bt[p / kPage] * 2) * n_kv + kvh) * kPage * HD + (size_t)(p % kPage) * HD;
        __nv_bfloat16 *vdst = kdst + (size_t)n_kv * kPage * HD;
        for (int i = lane; i < (G + 1) * (HALF / 8); i += 32) {
            const int h = i / (HALF / 8), c = i - h * (HALF / 8);
            const __nv_bfloat16 *src = h < G ? row + (size_t)(kvh * G + h) * HD : row + (size_t)(n_q + kvh) * HD;
            uint4 lo, hi;
            rope8(src, HALF, cs, c, lo, hi);
            if (h < G) {
                *reinterpret_cast<uint4 *>(Qs + swz<HD>(h, c)) = lo;
                *reinterpret_cast<uint4 *>(Qs + swz<HD>(h, c + HALF / 8)) = hi;
            } else {
                reinterpret_cast<uint4 *>(kdst)[c] = lo; reinterpret_cast<uint4 *>(kdst)[c + HALF / 8] = hi;
                reinterpret_cast<uint4 *>(knew)[c] = lo; reinterpret_cast<uint4 *>(knew)[c + HALF / 8] = hi;
            }
        }
        const __nv_bfloat16 *vsrc = row + (size_t)(n_q + n_kv + kvh) * HD;
        for (int i = lane; i < CH; i += 32) {
            const uint4 x = reinterpret_cast<const uint4 *>(vsrc)[i];
            reinterpret_cast<uint4 *>(vdst)[i] = x; reinterpret_cast<uint4 *>(vnew)[i] = x;
        }
        for (int i = lane; i < (16 - G) * CH; i += 32)
            *reinterpret_cast<uint4 *>(Qs + swz<HD>(G + i / CH, i % CH)) = make_uint4(0, 0, 0, 0);
    }
    __syncwarp();
    warp_attn_main<HD, G, KB, NS>(Qs, nblk, key0, T, p, pages, bt, n_kv, kvh, lane, scale_log2,
                                  out + (size_t)b * out_ld + (size_t)kvh * G * HD, part_o, part_ml, n_part,
                                  (size_t)(n_tasks / n_kv) * n_q, (size_t)b * n_q + kvh * G);
    if (lane == 0) trace_max(trace, 3);
}

// ---- decode attention, WPP (2 or 4) warps per (sequence, kv head) --------------------------------------
// Medium batches (a few hundred pairs): one warp per pair leaves most SMs with four resident warps, each walking its
// pair's key blocks one after the other (9 dependent softmax steps at 272 keys: 12 us per layer at batch 64).  Here
// the WPP warps of a pair take 1 / WPP of its key blocks each -- same private cp.async rings, same prefetch before
// griddepcontrol.wait -- and their (max, sum, O) partials are merged through shared memory.  A CTA of four warps
// holds 4 / WPP pairs; WPP is chosen so that the grid is one resident wave.
template <int HD, int G, int KB, int NS, int WPP>
__global__ void __launch_bounds__(128)
attn_decode_split_kernel(const __nv_bfloat16 *__restrict__ qkv, int qkv_ld, __nv_bfloat16 *__restrict__ out, int out_ld,
                         __nv_bfloat16 *__restrict__ pages, const int *__restrict__ block_table, int max_pages,
                         const int *__restrict__ pos, const float2 *__restrict__ rope, int n_q, int n_kv, float scale_log2,
                         int key0, const float *__restrict__ part_o, const float *__restrict__ part_ml, int n_part, int n_tasks,
                         unsigned long long *trace) {
    extern __shared__ __align__(128) uint8_t smem_dw[];
    asm volatile("griddepcontrol.launch_dependents;");
    if (threadIdx.x == 0) trace_min(trace, 0);
    using W = WarpAttnSmem<HD, KB, NS>;
    constexpr int CH = HD / 8, HALF = HD / 2, PP = 4 / WPP;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
    const int pl = warp / WPP, sub = warp - pl * WPP;                       // pair inside the CTA, slice of its keys
    const int task = blockIdx.x * PP + pl;
    const bool live = task < n_tasks;
    const int b = live ? task / n_kv : 0, kvh = live ? task - b * n_kv : 0;
    uint8_t *Qs = smem_dw + warp * W::BYTES;
    const uint32_t ring_u = (uint32_t)__cvta_generic_to_shared(Qs) + W::QT;
    __nv_bfloat16 *knew = reinterpret_cast<__nv_bfloat16 *>(Qs + W::QT + NS * 2 * W::ST), *vnew = knew + HD;
    float *comb_m = reinterpret_cast<float *>(smem_dw + 4 * W::BYTES);     // [4 warps][16] max, [4][16] sum, [4][G][HD] O
    float *comb_l = comb_m + 4 * 16, *comb_o = comb_l + 4 * 16;
    const int p = pos[b], T = p + 1;
    const int *bt = block_table + (size_t)b * max_pages;
    const int nblk_all = live ? (T - key0 + KB - 1) / KB : 0;
    const int blk0 = sub * nblk_all / WPP, blk1 = (sub + 1) * nblk_all / WPP;   // this warp's blocks
    const int k0 = key0 + blk0 * KB, Tw = min(T, key0 + blk1 * KB), nblk = blk1 - blk0;
    const bool owner = sub == WPP - 1;                                      // its last block holds the new token
#pragma unroll
    for (int i = 0; i < NS - 1; ++i) {                // prologue: NS-1 stages in flight (empty groups keep the count uniform)
        if (i < nblk) warp_attn_fetch<HD, KB, NS>(ring_u, i, k0, Tw, p, pages, bt, n_kv, kvh, lane);
        cp_async_commit();
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");   // qkv comes from the predecessor
    if (threadIdx.x == 0) trace_min(trace, 1);
    float o[HD / 8][4], mrow[2], lrow[2];
    if (live) {
        // RoPE: q heads -> this warp's Q tile; the owner warp also rotates k and appends k, v
        const __nv_bfloat16 *row = qkv + (size_t)b * qkv_ld;
        const float2 *cs = rope + (size_t)p * HALF;
        __nv_bfloat16 *kdst = pages + (((size_t)bt[p / kPage] * 2) * n_kv + kvh) * kPage * HD + (size_t)(p % kPage) * HD;
        __nv_bfloat16 *vdst = kdst + (size_t)n_kv * kPage * HD;
        for (int i = lane; i < (G + (owner ? 1 : 0)) * (HALF / 8); i += 32) {
            const int h = i / (HALF / 8), c = i - h * (HALF / 8);
            const __nv_bfloat16 *src = h < G ? row + (size_t)(kvh * G + h) * HD : row + (size_t)(n_q + kvh) * HD;
            uint4 lo, hi;
            rope8(src, HALF, cs, c, lo, hi);
            if (h < G) {
                *reinterpret_cast<uint4 *>(Qs + swz<HD>(h, c)) = lo;
                *reinterpret_cast<uint4 *>(Qs + swz<HD>(h, c + HALF / 8)) = hi;
            } else {
                reinterpret_cast<uint4 *>(kdst)[c] = lo; reinterpret_cast<uint4 *>(kdst)[c + HALF / 8] = hi;
                reinterpret_cast<uint4 *>(knew)[c] = lo; reinterpret_cast<uint4 *>(knew)[c + HALF / 8] = hi;
            }
        }
        if (owner) {
            const __nv_bfloat16 *vsrc = row + (size_t)(n_q + n_kv + kvh) * HD;
            for (int i = lane; i < CH; i += 32) {
                const uint4 x = reinterpret_cast<const uint4 *>(vsrc)[i];
                reinterpret_cast<uint4 *>(vdst)[i] = x; reinterpret_cast<uint4 *>(vnew)[i] = x;
            }
        }
        for (int i = lane; i < (16 - G) * CH; i += 32)
            *reinterpret_cast<uint4 *>(Qs + swz<HD>(G + i / CH, i % CH)) = make_uint4(0, 0, 0, 0);
        __syncwarp();
        warp_attn_loop<HD, G, KB, NS>(Qs, nblk, k0, Tw, p, pages, bt, n_kv, kvh, lane, scale_log2, o, mrow, lrow, owner);
        const int g = lane >> 2, tq = lane & 3;
        if (tq == 0) {
            comb_m[warp * 16 + g] = mrow[0]; comb_m[warp * 16 + g + 8] = mrow[1];
            comb_l[warp * 16 + g] = lrow[0]; comb_l[warp * 16 + g + 8] = lrow[1];
        }
        if (g < G) {
#pragma unroll
            for (int j = 0; j < HD / 8; ++j) {
                comb_o[(warp * G + g) * HD + j * 8 + 2 * tq] = o[j][0];
                comb_o[(warp * G + g) * HD + j * 8 + 2 * tq + 1] = o[j][1];
            }
        }
        if (G > 8 && g + 8 < G) {
#pragma unroll
            for (int j = 0; j < HD / 8; ++j) {
                comb_o[(warp * G + g + 8) * HD + j * 8 + 2 * tq] = o[j][2];
                comb_o[(warp * G + g + 8) * HD + j * 8 + 2 * tq + 1] = o[j][3];
            }
        }
    }
    __syncthreads();
    if (live) {
        const size_t part_rows = (size_t)(n_tasks / n_kv) * n_q;            // rows of one prefix partial: (sequence, q head)
        const int w0 = pl * WPP;
        for (int i = sub * 32 + lane; i < G * HD; i += WPP * 32) {
            const int h = i / HD, d = i - h * HD;
            const size_t prow = (size_t)b * n_q + kvh * G + h;
            float M = -INFINITY;
#pragma unroll
            for (int w = 0; w < WPP; ++w) M = fmaxf(M, comb_m[(w0 + w) * 16 + h]);
            for (int c = 0; c < n_part; ++c) M = fmaxf(M, part_ml[(c * part_rows + prow) * 2]);
            float num = 0.f, den = 0.f;
#pragma unroll
            for (int w = 0; w < WPP; ++w) {
                const float f = exp2f((comb_m[(w0 + w) * 16 + h] - M) * scale_log2);   // 0 for a warp that saw no keys
                num += f * comb_o[((w0 + w) * G + h) * HD + d];
                den += f * comb_l[(w0 + w) * 16 + h];
            }
            for (int c = 0; c < n_part; ++c) {
                const float2 ml = *reinterpret_cast<const float2 *>(part_ml + (c * part_rows + prow) * 2);
                const float f = exp2f((ml.x - M) * scale_log2);
                num += f * part_o[(c * part_rows + prow) * HD + d];
                den += f * ml.y;
            }
            out[(size_t)b * out_ld + (size_t)(kvh * G + h) * HD + d] = __float2bfloat16(num / den);
        }
    }
    if (tid == 0) trace_max(trace, 3);
}

// ---- decode attention over a prefix shared by the whole batch ("read the prefix once") ----
// Every sequence's first P keys are the same cached pages (kx_engine_set_prefix), so for one kv head the
// B x G query rows of the step attend to them as one tall tile problem: grid (n_kv, ceil(B G / 64), chunks),
// 128 threads, each CTA takes 64 query rows (row = b G + g, rotated here like the decode kernel rotates
// them) against one chunk of the prefix and writes an unnormalised partial (max, sum, O[HD]) per row.
// K/V of the prefix are read once per 64 rows instead of once per sequence.  The prefix pages do not depend
// on the preceding GEMM, so the first K/V tile is requested before griddepcontrol.wait.
template <int HD>
__global__ void __launch_bounds__(128)
attn_prefix_kernel(const __nv_bfloat16 *__restrict__ qkv, int qkv_ld, const __nv_bfloat16 *__restrict__ pages,
                   const int *__restrict__ bt, const int *__restrict__ pos, const float2 *__restrict__ rope,
                   int B, int G, int n_q, int n_kv, int P, int chunk_keys, float scale_log2,
                   float *__restrict__ part_o, float *__restrict__ part_ml, unsigned long long *trace) {
    extern __shared__ __align__(128) uint8_t smem_attn[];
    asm volatile("griddepcontrol.launch_dependents;");
    if (threadIdx.x == 0) trace_min(trace, 0);
    constexpr int CH = HD / 8, HALF = HD / 2;
    constexpr int TILE = 64 * HD * 2;
    const uint32_t qs_u = (uint32_t)__cvta_generic_to_shared(smem_attn);
    const int kvh = blockIdx.x, rt = blockIdx.y, chunk = blockIdx.z;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int kbeg = chunk * chunk_keys, kend = min(P, kbeg + chunk_keys);
    const int nkb = (kend - kbeg + 63) / 64;
    const int rows = B * G;

    auto fetch_kv = [&](int kb) {
        const uint32_t kdst = qs_u + TILE + (uint32_t)(kb & 1) * 2 * TILE, vdst = kdst + TILE;
        for (int i = tid; i < 64 * CH; i += 128) {
            const int r = i / CH, c = i - r * CH;
            const int t = kbeg + kb * 64 + r;
            const bool ok = t < kend;
            const int tt = ok ? t : kbeg;
            const __nv_bfloat16 *kp = pages + (((size_t)bt[tt / kPage] * 2) * n_kv + kvh) * kPage * HD + (size_t)(tt % kPage) * HD + c * 8;
            cp_async16_zfill(kdst + swz<HD>(r, c), kp, ok);
            cp_async16_zfill(vdst + swz<HD>(r, c), kp + (size_t)n_kv * kPage * HD, ok);
        }
    };
    fetch_kv(0);
    cp_async_commit();
    asm volatile("griddepcontrol.wait;" ::: "memory");   // q comes from the QKV GEMM
    if (tid == 0) trace_min(trace, 1);
    for (int i = tid; i < 64 * (HALF / 8); i += 128) {    // rotated q rows -> Q tile (rows past B G are zero)
        const int r = i / (HALF / 8), c = i - r * (HALF / 8);
        const int row = rt * 64 + r;
        uint4 lo = make_uint4(0, 0, 0, 0), hi = lo;
        if (row < rows) {
            const int b = row / G, g = row - b * G;
            rope8(qkv + (size_t)b * qkv_ld + (size_t)(kvh * G + g) * HD, HALF, rope + (size_t)pos[b] * HALF, c, lo, hi);
        }
        *reinterpret_cast<uint4 *>(smem_attn + swz<HD>(r, c)) = lo;
        *reinterpret_cast<uint4 *>(smem_attn + swz<HD>(r, c + HALF / 8)) = hi;
    }

    uint32_t qf[HD / 16][4];
    float o[HD / 8][4];
#pragma unroll
    for (int j = 0; j < HD / 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
    float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
    const int g8 = lane >> 2, tq = lane & 3;

    for (int kb = 0; kb < nkb; ++kb) {
        if (kb + 1 < nkb) { fetch_kv(kb + 1); cp_async_commit(); cp_async_wait<1>(); }
        else cp_async_wait<0>();
        __syncthreads();
        if (kb == 0) {
#pragma unroll
            for (int ks = 0; ks < HD / 16; ++ks)
                ldsm_x4(qf[ks], qs_u + swz<HD>(warp * 16 + (lane & 15), 2 * ks + (lane >> 4)));
        }
        const uint32_t ks_u = qs_u + TILE + (uint32_t)(kb & 1) * 2 * TILE, vs_u = ks_u + TILE;
        float s[8][4];
#pragma unroll
        for (int j = 0; j < 8; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) {
#pragma unroll
            for (int jp = 0; jp < 4; ++jp) {
                uint32_t kf[4];
                ldsm_x4(kf, ks_u + swz<HD>(jp * 16 + (lane & 7) + ((lane >> 4) << 3), 2 * ks + ((lane >> 3) & 1)));
                mma_bf16(s[2 * jp], qf[ks], kf[0], kf[1]);
                mma_bf16(s[2 * jp + 1], qf[ks], kf[2], kf[3]);
            }
        }
        const int nvalid = kend - kbeg - kb * 64;              // keys of this tile inside the chunk
        float mx[2] = {mrow[0], mrow[1]};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (j * 8 + 2 * tq + (e & 1) >= nvalid) s[j][e] = -INFINITY;
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
        uint32_t pf[4][4];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float p0 = exp2f(s[j][0] * scale_log2 - msc[0]), p1 = exp2f(s[j][1] * scale_log2 - msc[0]);
            const float p2 = exp2f(s[j][2] * scale_log2 - msc[1]), p3 = exp2f(s[j][3] * scale_log2 - msc[1]);
            lrow[0] += p0 + p1;
            lrow[1] += p2 + p3;
            pf[j >> 1][(j & 1) * 2] = pack2(p0, p1);
            pf[j >> 1][(j & 1) * 2 + 1] = pack2(p2, p3);
        }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
            for (int dp = 0; dp < HD / 16; ++dp) {
                uint32_t vf[4];
                ldsm_x4_t(vf, vs_u + swz<HD>(kk * 16 + (lane & 7) + (((lane >> 3) & 1) << 3), 2 * dp + (lane >> 4)));
                mma_bf16(o[2 * dp], pf[kk], vf[0], vf[1]);
                mma_bf16(o[2 * dp + 1], pf[kk], vf[2], vf[3]);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 1);
        lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 2);
    }
    const size_t part_rows = (size_t)B * n_q;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const int row = rt * 64 + warp * 16 + g8 + r * 8;
        if (row >= rows) continue;
        const int b = row / G, g = row - b * G;
        const size_t prow = (size_t)chunk * part_rows + (size_t)b * n_q + kvh * G + g;
        if (tq == 0) *reinterpret_cast<float2 *>(part_ml + prow * 2) = make_float2(mrow[r], lrow[r]);
        float *po = part_o + prow * HD + 2 * tq;
#pragma unroll
        for (int j = 0; j < HD / 8; ++j)
            *reinterpret_cast<float2 *>(po + j * 8) = make_float2(o[j][2 * r], o[j][2 * r + 1]);
    }
    if (tid == 0) trace_max(trace, 3);
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

// ---- second stage of the fused lm_head argmax: parts[n_parts][ld] (value bits, index) -> one id per token ----
// block = 8 tokens x 128 part groups (a 64-byte run of one part row per 8 lanes), so batch 64 already spreads over 8 SMs and
// a thread walks n_parts / 128 parts with four loads in flight; ties -> lowest index (what torch.argmax over the full
// row returns).  (32 tokens x 32 groups per block left batch 64 on two SMs with 125 dependent steps per thread: 55 us.)
__device__ __forceinline__ void argmax_take(float &best, int &bi, float f, int i) {
    if (f > best || (f == best && i < bi)) { best = f; bi = i; }
}
__global__ void __launch_bounds__(1024)
argmax_parts_kernel(const int2 *__restrict__ parts, int n_parts, int ld, int B, int id_offset, int *__restrict__ out,
                    int *__restrict__ out2, int2 *__restrict__ out_pairs) {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;      // token within the block, part group 0..127
    const int b = blockIdx.x * 8 + tx;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    if (b < B) {
        int pi = ty;
        for (; pi + 384 < n_parts; pi += 512) {
            const int2 v0 = parts[(size_t)pi * ld + b], v1 = parts[(size_t)(pi + 128) * ld + b];
            const int2 v2 = parts[(size_t)(pi + 256) * ld + b], v3 = parts[(size_t)(pi + 384) * ld + b];
            argmax_take(best, bi, __int_as_float(v0.x), v0.y); argmax_take(best, bi, __int_as_float(v1.x), v1.y);
            argmax_take(best, bi, __int_as_float(v2.x), v2.y); argmax_take(best, bi, __int_as_float(v3.x), v3.y);
        }
        for (; pi < n_parts; pi += 128) {
            const int2 v = parts[(size_t)pi * ld + b];
            argmax_take(best, bi, __int_as_float(v.x), v.y);
        }
    }
    // the four part groups of a warp (lanes tx, tx + 8, tx + 16, tx + 24), then the 32 warps through shared memory
#pragma unroll
    for (int d = 8; d < 32; d <<= 1) {
        const float f = __shfl_xor_sync(0xffffffffu, best, d);
        const int i = __shfl_xor_sync(0xffffffffu, bi, d);
        argmax_take(best, bi, f, i);
    }
    __shared__ float sv[32][8];
    __shared__ int si[32][8];
    const int warp = threadIdx.x >> 5;
    if ((threadIdx.x & 31) < 8) { sv[warp][tx] = best; si[warp][tx] = bi; }
    __syncthreads();
    if (threadIdx.x < 8 && b < B) {
        best = sv[0][tx]; bi = si[0][tx];
        for (int g = 1; g < 32; ++g) argmax_take(best, bi, sv[g][tx], si[g][tx]);
        if (out_pairs) out_pairs[b] = make_int2(__float_as_int(best), bi + id_offset);
        else { out[b] = bi + id_offset; if (out2) out2[b] = bi + id_offset; }
    }
}

// ---- tensor-parallel helpers -----------------------------------------------------
__global__ void add_inplace_kernel(float *__restrict__ x, const float *__restrict__ d, size_t n4) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n4) {
        float4 a = reinterpret_cast<float4 *>(x)[i];
        const float4 b = reinterpret_cast<const float4 *>(d)[i];
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        reinterpret_cast<float4 *>(x)[i] = a;
    }
}
// per-shard argmax -> (value bits, global id) pairs; merged by pick_kernel after an all-gather
__global__ void __launch_bounds__(256)
argmax_pair_kernel(const float *__restrict__ logits, int V, int ld, int id_offset, int2 *__restrict__ pairs) {
    const int b = blockIdx.x;
    const float *row = logits + (size_t)b * ld;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < V; i += 256) {
        const float a = row[i];
        if (a > best || (a == best && i < bi)) { best = a; bi = i; }
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
        pairs[b] = make_int2(__float_as_int(best), bi + id_offset);
    }
}
__global__ void pick_kernel(const int2 *__restrict__ gathered, int tp, int B, int *__restrict__ out) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int r = 0; r < tp; ++r) {
        const int2 p = gathered[(size_t)r * B + b];
        const float v = __int_as_float(p.x);
        if (v > best || (v == best && p.y < bi)) { best = v; bi = p.y; }
    }
    out[b] = bi;
}

// ---- end of a decode step on the deferred-RMSNorm path (kx_gemm.cuh) ---------------------------------
// One CTA per sequence: records the token the step chose (store != 0), moves the sequence to its next position and
// gathers that token's embedding row as the next step's input -- the fp32 residual, bf16(residual * g) for the first
// QKV projection and, per 128-feature tile, the sum of squares the projection turns into 1 / rms.  Replaces the
// store-token, advance, embedding and first RMSNorm launches.
__global__ void __launch_bounds__(128)
decode_tail_kernel(const int *__restrict__ next_tok, int *__restrict__ out_ids, int n_new, int base_pos, int *__restrict__ pos,
                   const __nv_bfloat16 *__restrict__ table, const __nv_bfloat16 *__restrict__ g, float *__restrict__ resid,
                   __nv_bfloat16 *__restrict__ xg, float *__restrict__ ssq, int ssq_ld, int H, int store,
                   int eos_id, int *__restrict__ done) {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int b = blockIdx.x, tid = threadIdx.x;
    const int tok = next_tok[b];
    if (tid == 0) {
        const int p = pos[b];
        if (store) out_ids[(size_t)b * n_new + (p - base_pos)] = tok;
        pos[b] = p + 1;
        // end-of-sequence bookkeeping: done[1 + b] = this sequence has emitted eos, done[0] = how many have
        if (eos_id >= 0 && tok == eos_id && done[1 + b] == 0) { done[1 + b] = 1; atomicAdd(done, 1); }
    }
    const __nv_bfloat16 *row = table + (size_t)tok * H;
    __shared__ float ws[4];
    for (int f0 = 0; f0 < H; f0 += 128) {
        const float v = __bfloat162float(row[f0 + tid]);
        resid[(size_t)b * H + f0 + tid] = v;
        xg[(size_t)b * H + f0 + tid] = __float2bfloat16(__fmul_rn(v, __bfloat162float(g[f0 + tid])));
        const float sq = warp_sum(__fmul_rn(v, v));
        if ((tid & 31) == 0) ws[tid >> 5] = sq;
        __syncthreads();
        if (tid == 0) ssq[(size_t)(f0 >> 7) * ssq_ld + b] = (ws[0] + ws[1]) + (ws[2] + ws[3]);
        __syncthreads();
    }
}

// Continuous batching (kx_engine_session_*): what decode_tail_kernel does, with the bookkeeping per SLOT instead of per
// batch.  cnt[b] = tokens the slot has produced (< 0: the slot is free), limit[b] = how many it may produce, fin[b] = it
// has ended (eos or limit).  A free or finished slot keeps its position: it re-decodes one harmless position of its own
// pages until the host reuses it.  book = 0: only prepare the step's input (after an admission overwrote the buffers);
// table = NULL: only the bookkeeping (the classic decode path embeds at the start of its step).
__global__ void __launch_bounds__(128)
session_tail_kernel(const int *__restrict__ next_tok, int *__restrict__ out, int out_ld, int *__restrict__ pos, int *__restrict__ cnt,
                    const int *__restrict__ limit, int *__restrict__ fin, int eos_id, int book,
                    const __nv_bfloat16 *__restrict__ table, const __nv_bfloat16 *__restrict__ g, float *__restrict__ resid,
                    __nv_bfloat16 *__restrict__ xg, float *__restrict__ ssq, int ssq_ld, int H) {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int b = blockIdx.x, tid = threadIdx.x;
    const int tok = next_tok[b];
    if (tid == 0 && book) {
        const int c = cnt[b];
        if (c >= 0 && fin[b] == 0) {
            out[(size_t)b * out_ld + c] = tok;
            cnt[b] = c + 1;
            if (tok == eos_id || c + 1 >= limit[b]) fin[b] = 1;
            else pos[b] += 1;
        }
    }
    if (table == nullptr) return;
    const __nv_bfloat16 *row = table + (size_t)tok * H;
    __shared__ float ws[4];
    for (int f0 = 0; f0 < H; f0 += 128) {
        const float v = __bfloat162float(row[f0 + tid]);
        resid[(size_t)b * H + f0 + tid] = v;
        xg[(size_t)b * H + f0 + tid] = __float2bfloat16(__fmul_rn(v, __bfloat162float(g[f0 + tid])));
        const float sq = warp_sum(__fmul_rn(v, v));
        if ((tid & 31) == 0) ws[tid >> 5] = sq;
        __syncthreads();
        if (tid == 0) ssq[(size_t)(f0 >> 7) * ssq_ld + b] = (ws[0] + ws[1]) + (ws[2] + ws[3]);
        __syncthreads();
    }
}

// slots [slot0, slot0 + n) start a sequence whose prefill chose tok[b]: token 0 of the slot's output
__global__ void session_admit_kernel(const int *__restrict__ tok, int *__restrict__ out, int out_ld, int *__restrict__ pos,
                                     int *__restrict__ cnt, int *__restrict__ limit, int *__restrict__ fin, int slot0, int n,
                                     int first_pos, int max_new, int eos_id) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int b = slot0 + i, t = tok[b];
    out[(size_t)b * out_ld] = t;
    cnt[b] = 1;
    limit[b] = max_new;
    fin[b] = (t == eos_id || max_new <= 1) ? 1 : 0;
    pos[b] = first_pos;
}

// slots [slot0, slot0 + n) become free; their position parks at `park` (the first position after a cached prefix)
__global__ void session_release_kernel(int *__restrict__ pos, int *__restrict__ cnt, int *__restrict__ fin, int slot0, int n, int park) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    cnt[slot0 + i] = -1;
    fin[slot0 + i] = 0;
    pos[slot0 + i] = park;
}

// 1 / rms per batch column from the per-tile sums of squares (only the logits path needs it as an array: the
// projections compute it themselves)
__global__ void col_rinv_kernel(const float *__restrict__ ssq, int tiles, int ld, int B, float inv_n, float eps, float *__restrict__ out) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    float ss = 0.f;
    for (int t = 0; t < tiles; ++t) ss += ssq[(size_t)t * ld + b];
    out[b] = 1.0f / sqrtf(ss * inv_n + eps);
}

// positions / bookkeeping for the decode loop, kept on the device so a step can be a CUDA graph
__global__ void advance_kernel(int *pos, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pos[i] += 1;
}

template <int HD, int NS, int KBQ = 0>
int launch_decode_warp(int G, int n_tasks, cudaStream_t st, const __nv_bfloat16 *qkv, int qkv_ld, __nv_bfloat16 *out, int out_ld,
                       __nv_bfloat16 *pages, const int *bt, int max_pages, const int *pos, const float2 *rope, int n_q, int n_kv,
                       float scale, int key0, const float *part_o, const float *part_ml, int n_part) {
    constexpr int KB = KBQ ? KBQ : (HD == 64 ? 32 : 16);
    constexpr size_t smem = (size_t)4 * (16 * HD * 2 + NS * 2 * KB * HD * 2 + 2 * HD * 2);
    const dim3 grid((n_tasks + 3) / 4);
#define KX_DECW(GV)                                                                                          \
    case GV: {                                                                                               \
        auto k = attn_decode_warp_kernel<HD, GV, KB, NS>;                                                    \
        static PerDeviceOnce once;     /* first use is the eager decode step, outside any capture */         \
        if (once.first()) { KX_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); once.done(); } \
        KX_CUDA(launch_pdl(k, grid, dim3(128), smem, st, qkv, qkv_ld, out, out_ld, pages, bt, max_pages, pos, rope, n_q, n_kv, scale, \
                           key0, part_o, part_ml, n_part, n_tasks, trace_slot("attn.warp")));                \
        break;                                                                                               \
    }
    switch (G) {
        KX_DECW(1) KX_DECW(2) KX_DECW(4) KX_DECW(7) KX_DECW(8)
        default:
            set_error("attention: unsupported GQA group size %d", G);
            return KX_EINVAL;
    }
#undef KX_DECW
    return KX_OK;
}

template <int HD, int WPP>
int launch_decode_split(int G, int n_tasks, cudaStream_t st, const __nv_bfloat16 *qkv, int qkv_ld, __nv_bfloat16 *out, int out_ld,
                        __nv_bfloat16 *pages, const int *bt, int max_pages, const int *pos, const float2 *rope, int n_q, int n_kv,
                        float scale, int key0, const float *part_o, const float *part_ml, int n_part) {
    constexpr int KB = HD == 64 ? 32 : 16, NS = 3, PP = 4 / WPP;
    const size_t smem = (size_t)4 * (16 * HD * 2 + NS * 2 * KB * HD * 2 + 2 * HD * 2) + sizeof(float) * (2 * 4 * 16 + (size_t)4 * G * HD);
    const dim3 grid((n_tasks + PP - 1) / PP);
#define KX_DECS(GV)                                                                                          \
    case GV: {                                                                                               \
        auto k = attn_decode_split_kernel<HD, GV, KB, NS, WPP>;                                              \
        static size_t granted[64];     /* per device; raised outside any capture by the first, eager call */   \
        size_t &have = granted[PerDeviceOnce::device()];                                                     \
        if (smem > 48 * 1024 && smem > have) { KX_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); have = smem; } \
        KX_CUDA(launch_pdl(k, grid, dim3(128), smem, st, qkv, qkv_ld, out, out_ld, pages, bt, max_pages, pos, rope, n_q, n_kv, scale, \
                           key0, part_o, part_ml, n_part, n_tasks, trace_slot("attn.split")));               \
        break;                                                                                               \
    }
    switch (G) {
        KX_DECS(1) KX_DECS(2) KX_DECS(4) KX_DECS(7) KX_DECS(8)
        default:
            set_error("attention: unsupported GQA group size %d", G);
            return KX_EINVAL;
    }
#undef KX_DECS
    return KX_OK;
}

template <int HD>
int launch_decode_g(int G, dim3 grid, size_t smem, cudaStream_t st, const __nv_bfloat16 *qkv, int qkv_ld,
                    __nv_bfloat16 *out, int out_ld, __nv_bfloat16 *pages, const int *bt, int max_pages,
                    const int *pos, const float2 *rope, int n_q, int n_kv, float scale, int round_keys, int nw,
                    int key0, const float *part_o, const float *part_ml, int n_part) {
#define KX_DEC(GV)                                                                                          \
    case GV: {                                                                                              \
        auto k = nw == 8 ? attn_decode_kernel<HD, GV, 8> : attn_decode_kernel<HD, GV, 4>;                   \
        static size_t granted[64][2];  /* per device; raised outside any capture by the first, eager call */    \
        size_t &have = granted[PerDeviceOnce::device()][nw == 8];                                           \
        if (smem > 48 * 1024 && smem > have) {                                                              \
            KX_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));        \
            have = smem;                                                                                    \
        }                                                                                                   \
        KX_CUDA(launch_pdl(k, grid, dim3(nw * 32), smem, st, qkv, qkv_ld, out, out_ld, pages, bt, max_pages, pos, rope, n_q, n_kv, scale, round_keys, key0, part_o, part_ml, n_part, trace_slot("attn.cta"))); \
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

int op_rmsnorm(float *x, const void *g, void *y, int M, int H, float eps, const int *row_map, cudaStream_t st,
               const float *part, int splits, size_t split_stride) {
    if (M <= 2048 && H <= 8192) {
        const int per = (H + 511) / 512;
        const __nv_bfloat16 *gg = (const __nv_bfloat16 *)g;
        __nv_bfloat16 *yy = (__nv_bfloat16 *)y;
        unsigned long long *tr = trace_slot("rmsnorm");
        if (per <= 2) KX_CUDA(launch_pdl(rmsnorm_row_kernel<2>, M, 128, 0, st, x, gg, yy, H, eps, row_map, part, splits, split_stride, tr));
        else if (per <= 4) KX_CUDA(launch_pdl(rmsnorm_row_kernel<4>, M, 128, 0, st, x, gg, yy, H, eps, row_map, part, splits, split_stride, tr));
        else if (per <= 8) KX_CUDA(launch_pdl(rmsnorm_row_kernel<8>, M, 128, 0, st, x, gg, yy, H, eps, row_map, part, splits, split_stride, tr));
        else KX_CUDA(launch_pdl(rmsnorm_row_kernel<16>, M, 128, 0, st, x, gg, yy, H, eps, row_map, part, splits, split_stride, tr));
        count_launch();
        return KX_OK;
    }
    KX_REQUIRE(splits == 0, KX_EINVAL, "rmsnorm: pending split-K partials are only folded in by the row kernel (M <= 2048)");
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
                   int key0, const float *part_o, const float *part_ml, int n_part, cudaStream_t st) {
    const int G = n_q / n_kv;
    KX_REQUIRE(G <= 16, KX_EINVAL, "attention: GQA group of %d query heads per kv head exceeds 16", G);
    KX_REQUIRE(key0 == 0 || (part_o && part_ml && n_part > 0), KX_EINVAL, "attention: shared-prefix partials missing");
    // one round holds the whole (non-shared) context when it fits (<= 512 keys); shared memory is sized for it
    int round_keys = ((max_T - key0 + 63) / 64) * 64;
    const int nw = ((int64_t)n_kv * B > 600 || round_keys <= 256) ? 4 : 8;
    if (round_keys > 64 * nw) round_keys = 64 * nw;
    const int fit = (int)((160 * 1024) / (4 * hd)) / 64 * 64;          // K and V of one round must fit in shared memory
    if (round_keys > fit) round_keys = fit;
    const size_t comb = sizeof(float) * (2 * 8 * 16 + (size_t)8 * G * hd);       // merge area aliases the K buffer
    size_t kv = (size_t)2 * round_keys * hd * 2;
    if (kv < comb) kv = comb;
    const size_t smem = (size_t)16 * hd * 2 + kv + 4 * hd + 128;
    const float scale = (1.0f / sqrtf((float)hd)) * 1.4426950408889634f;   // softmax scale in log2 units
    dim3 grid(n_kv, B);
    int rc;
    const int n_tasks = n_kv * B;
    if ((hd == 64 || hd == 128) && n_tasks >= 96 && n_tasks <= 1536 && max_T - key0 > 128 && G <= 8) {
        // up to ~1,500 pairs with long key lists: 2 or 4 warps share a pair (measured on Llama-3.2-1B, 272 keys: batch 64
        // 12.1 -> 9.3 us per layer, batch 128 1.67 -> 1.53 ms per step; batch 192 is a tie; from batch 256 the one-warp
        // kernel with 16-key stages, four CTAs per SM, is faster: 2.42 -> 2.13 ms per step.  16-key stages in this kernel
        // changed nothing at batch 128 - 256 and cost 10 % at batch 64.)
        const __nv_bfloat16 *q = (const __nv_bfloat16 *)qkv;
        __nv_bfloat16 *o = (__nv_bfloat16 *)out, *pg = (__nv_bfloat16 *)pages;
        const float2 *rp = (const float2 *)rope;
        if (n_tasks <= 296)
            rc = hd == 64 ? launch_decode_split<64, 4>(G, n_tasks, st, q, qkv_ld, o, out_ld, pg, block_table, max_pages, pos, rp, n_q, n_kv, scale, key0, part_o, part_ml, n_part)
                          : launch_decode_split<128, 4>(G, n_tasks, st, q, qkv_ld, o, out_ld, pg, block_table, max_pages, pos, rp, n_q, n_kv, scale, key0, part_o, part_ml, n_part);
        else
            rc = hd == 64 ? launch_decode_split<64, 2>(G, n_tasks, st, q, qkv_ld, o, out_ld, pg, block_table, max_pages, pos, rp, n_q, n_kv, scale, key0, part_o, part_ml, n_part)
                          : launch_decode_split<128, 2>(G, n_tasks, st, q, qkv_ld, o, out_ld, pg, block_table, max_pages, pos, rp, n_q, n_kv, scale, key0, part_o, part_ml, n_part);
        if (rc != KX_OK) return rc;
        count_launch();
        return KX_OK;
    }
    if ((hd == 64 || hd == 128) && (n_tasks >= 512 || max_T - key0 <= 128)) {
        // many (sequence, kv head) pairs, or few keys each: one warp per pair, no CTA-level split of the keys
        const bool deep = (n_tasks + 3) / 4 <= 148;            // one CTA per SM anyway: spend the shared memory on a deeper ring
        const __nv_bfloat16 *q = (const __nv_bfloat16 *)qkv;
        __nv_bfloat16 *o = (__nv_bfloat16 *)out, *pg = (__nv_bfloat16 *)pages;
        const float2 *rp = (const float2 *)rope;
        if (hd == 64)
            if (!deep)     // more CTAs than SMs: 16-key stages -> 41 KB per CTA -> four CTAs per SM instead of three, i.e. two
                           // rounds of CTAs instead of three for 4,096 pairs (Llama-3.2-1B batch 512: 3.19 -> 3.02 ms per step)
                rc = launch_decode_warp<64, 2, 16>(G, n_tasks, st, q, qkv_ld, o, out_ld, pg, block_table, max_pages, pos, rp, n_q, n_kv, scale, key0, part_o, part_ml, n_part);
            else
            rc = deep ? launch_decode_warp<64, 4>(G, n_tasks, st, q, qkv_ld, o, out_ld, pg, block_table, max_pages, pos, rp, n_q, n_kv, scale, key0, part_o, part_ml, n_part)
                      : launch_decode_warp<64, 2>(G, n_tasks, st, q, qkv_ld, o, out_ld, pg, block_table, max_pages, pos, rp, n_q, n_kv, scale, key0, part_o, part_ml, n_part);
        else
            rc = deep ? launch_decode_warp<128, 4>(G, n_tasks, st, q, qkv_ld, o, out_ld, pg, block_table, max_pages, pos, rp, n_q, n_kv, scale, key0, part_o, part_ml, n_part)
                      : launch_decode_warp<128, 2>(G, n_tasks, st, q, qkv_ld, o, out_ld, pg, block_table, max_pages, pos, rp, n_q, n_kv, scale, key0, part_o, part_ml, n_part);
        if (rc != KX_OK) return rc;
        count_launch();
        return KX_OK;
    }
    if (hd == 64)
        rc = launch_decode_g<64>(G, grid, smem, st, (const __nv_bfloat16 *)qkv, qkv_ld, (__nv_bfloat16 *)out, out_ld,
                                 (__nv_bfloat16 *)pages, block_table, max_pages, pos, (const float2 *)rope, n_q, n_kv, scale, round_keys, nw,
                                 key0, part_o, part_ml, n_part);
    else if (hd == 128)
        rc = launch_decode_g<128>(G, grid, smem, st, (const __nv_bfloat16 *)qkv, qkv_ld, (__nv_bfloat16 *)out, out_ld,
                                  (__nv_bfloat16 *)pages, block_table, max_pages, pos, (const float2 *)rope, n_q, n_kv, scale, round_keys, nw,
                                  key0, part_o, part_ml, n_part);
    else {
        set_error("attention: head_dim %d not supported (64 or 128)", hd);
        return KX_EINVAL;
    }
    if (rc != KX_OK) return rc;
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

// chunks the shared prefix is cut into for a batch of B: enough CTAs for one per SM, 64-key granularity
int attn_prefix_parts(int B, int n_q, int n_kv, int P, int *chunk_keys) {
    const int G = n_q / n_kv;
    const int tiles = n_kv * ((B * G + 63) / 64);
    int nc = (148 + tiles - 1) / tiles;               // about one CTA per SM: the kernel is latency-, not bandwidth-bound
    if (nc > 16) nc = 16;
    const int blocks = (P + 63) / 64;
    if (nc > blocks) nc = blocks;
    if (nc < 1) nc = 1;
    const int ck = ((blocks + nc - 1) / nc) * 64;
    if (chunk_keys) *chunk_keys = ck;
    return (P + ck - 1) / ck;
}

int op_attn_prefix(const void *qkv, int qkv_ld, const void *pages, const int *bt_prefix, const int *pos, const float *rope,
                   int B, int n_q, int n_kv, int hd, int P, float *part_o, float *part_ml, int *n_part, cudaStream_t st) {
    KX_REQUIRE(P > 0 && part_o && part_ml && n_part, KX_EINVAL, "prefix attention: bad argument");
    const int G = n_q / n_kv;
    int ck = 0;
    const int nc = attn_prefix_parts(B, n_q, n_kv, P, &ck);
    *n_part = nc;
    const float scale_log2 = (1.0f / sqrtf((float)hd)) * 1.4426950408889634f;
    dim3 grid(n_kv, (B * G + 63) / 64, nc);
    if (hd == 64) {
        KX_CUDA(launch_pdl(attn_prefix_kernel<64>, grid, dim3(128), (size_t)5 * 64 * 64 * 2, st, (const __nv_bfloat16 *)qkv, qkv_ld,
                           (const __nv_bfloat16 *)pages, bt_prefix, pos, (const float2 *)rope, B, G, n_q, n_kv, P, ck, scale_log2, part_o, part_ml, trace_slot("attn.prefix")));
    } else if (hd == 128) {
        auto k = attn_prefix_kernel<128>;
        static PerDeviceOnce once;                     // first use is the eager decode step, outside any capture
        if (once.first()) { KX_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 5 * 64 * 128 * 2)); once.done(); }
        KX_CUDA(launch_pdl(k, grid, dim3(128), (size_t)5 * 64 * 128 * 2, st, (const __nv_bfloat16 *)qkv, qkv_ld,
                           (const __nv_bfloat16 *)pages, bt_prefix, pos, (const float2 *)rope, B, G, n_q, n_kv, P, ck, scale_log2, part_o, part_ml, trace_slot("attn.prefix")));
    } else {
        set_error("attention: head_dim %d not supported (64 or 128)", hd);
        return KX_EINVAL;
    }
    count_launch();
    return KX_OK;
}

int op_attn_prefill(const void *qkv, int qkv_ld, void *out, int out_ld, const void *pages, const int *block_table,
                    int max_pages, int B, int S, int pos0, int n_q, int n_kv, int hd, cudaStream_t st) {
    const float scale_log2 = (1.0f / sqrtf((float)hd)) * 1.4426950408889634f;
    dim3 grid((S + 63) / 64, n_q, B);
    if (hd == 64) {
        attn_prefill_kernel<64><<<grid, 128, 5 * 64 * 64 * 2, st>>>((const __nv_bfloat16 *)qkv, qkv_ld, (__nv_bfloat16 *)out, out_ld,
            (const __nv_bfloat16 *)pages, block_table, max_pages, S, pos0, n_q, n_kv, scale_log2);
    } else if (hd == 128) {
        auto k = attn_prefill_kernel<128>;
        static PerDeviceOnce once;
        if (once.first()) { KX_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 5 * 64 * 128 * 2)); once.done(); }
        k<<<grid, 128, 5 * 64 * 128 * 2, st>>>((const __nv_bfloat16 *)qkv, qkv_ld, (__nv_bfloat16 *)out, out_ld,
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

int op_argmax_parts(const void *parts, int n_parts, int ld, int B, int id_offset, int *out, int *out2, void *out_pairs,
                    cudaStream_t st) {
    KX_CUDA(launch_pdl(argmax_parts_kernel, dim3((B + 7) / 8), dim3(1024), 0, st, (const int2 *)parts, n_parts, ld, B, id_offset, out,
                       out2, (int2 *)out_pairs));
    count_launch();
    return KX_OK;
}

int op_add_inplace(float *x, const float *d, size_t n, cudaStream_t st) {
    const size_t n4 = n / 4;
    add_inplace_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(x, d, n4);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_argmax_pairs(const float *logits, int B, int V, int ld, int id_offset, void *pairs, cudaStream_t st) {
    argmax_pair_kernel<<<B, 256, 0, st>>>(logits, V, ld, id_offset, (int2 *)pairs);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_pick(const void *gathered, int tp, int B, int *out, cudaStream_t st) {
    pick_kernel<<<(B + 127) / 128, 128, 0, st>>>((const int2 *)gathered, tp, B, out);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_decode_tail(const int *next_tok, int *out_ids, int n_new, int base_pos, int *pos, const void *table, const void *g,
                   float *resid, void *xg, float *ssq, int ssq_ld, int B, int H, int store, int eos_id, int *done, cudaStream_t st) {
    KX_REQUIRE(H % 128 == 0, KX_EINVAL, "decode tail: hidden must be a multiple of 128");
    KX_CUDA(launch_pdl(decode_tail_kernel, dim3(B), dim3(128), 0, st, next_tok, out_ids, n_new, base_pos, pos,
                       (const __nv_bfloat16 *)table, (const __nv_bfloat16 *)g, resid, (__nv_bfloat16 *)xg, ssq, ssq_ld, H, store,
                       eos_id, done));
    count_launch();
    return KX_OK;
}

int op_session_tail(const int *next_tok, int *out, int out_ld, int *pos, int *cnt, const int *limit, int *fin, int eos_id, int book,
                    const void *table, const void *g, float *resid, void *xg, float *ssq, int ssq_ld, int B, int H, cudaStream_t st) {
    KX_REQUIRE(H % 128 == 0, KX_EINVAL, "session tail: hidden must be a multiple of 128");
    KX_CUDA(launch_pdl(session_tail_kernel, dim3(B), dim3(128), 0, st, next_tok, out, out_ld, pos, cnt, limit, fin, eos_id, book,
                       (const __nv_bfloat16 *)table, (const __nv_bfloat16 *)g, resid, (__nv_bfloat16 *)xg, ssq, ssq_ld, H));
    count_launch();
    return KX_OK;
}

int op_session_admit(const int *tok, int *out, int out_ld, int *pos, int *cnt, int *limit, int *fin, int slot0, int n, int first_pos,
                     int max_new, int eos_id, cudaStream_t st) {
    session_admit_kernel<<<(n + 127) / 128, 128, 0, st>>>(tok, out, out_ld, pos, cnt, limit, fin, slot0, n, first_pos, max_new, eos_id);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_session_release(int *pos, int *cnt, int *fin, int slot0, int n, int park, cudaStream_t st) {
    session_release_kernel<<<(n + 127) / 128, 128, 0, st>>>(pos, cnt, fin, slot0, n, park);
    count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

int op_col_rinv(const float *ssq, int tiles, int ld, int B, int H, float eps, float *out, cudaStream_t st) {
    col_rinv_kernel<<<(B + 127) / 128, 128, 0, st>>>(ssq, tiles, ld, B, 1.0f / (float)H, eps, out);
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

extern "C" int kx_attn_decode(const void *d_qkv, int32_t qkv_ld, void *d_out, int32_t out_ld, void *d_pages,
                              const int32_t *d_block_table, int32_t max_pages, const int32_t *d_pos, const float *d_rope,
                              int32_t B, int32_t max_T, int32_t n_q, int32_t n_kv, int32_t hd, int32_t prefix_len,
                              int32_t prefix_row, void *stream) {
    using namespace kx;
    KX_REQUIRE(d_qkv && d_out && d_pages && d_block_table && d_pos && d_rope, KX_EINVAL, "kx_attn_decode: NULL argument");
    KX_REQUIRE(B >= 1 && n_kv >= 1 && n_q % n_kv == 0 && max_T >= 1, KX_EINVAL, "kx_attn_decode: bad shape");
    KX_REQUIRE(prefix_len >= 0 && prefix_len % KX_KV_PAGE == 0 && prefix_len < max_T, KX_EINVAL,
               "kx_attn_decode: prefix_len %d must be a multiple of %d below max_T", prefix_len, KX_KV_PAGE);
    cudaStream_t st = (cudaStream_t)stream;
    int key0 = 0, n_part = 0;
    float *part_o = nullptr, *part_ml = nullptr;
    if (prefix_len > 0) {
        KX_REQUIRE(prefix_len >= 64 && prefix_row >= 0, KX_EINVAL, "kx_attn_decode: a shared prefix needs >= 64 keys and its block-table row");
        const size_t rows = (size_t)attn_prefix_parts(B, n_q, n_kv, prefix_len, nullptr) * B * n_q;
        void *scratch = nullptr;
        int rc = stream_scratch(st, SCRATCH_ATTN, rows * (hd + 2) * sizeof(float), &scratch);
        if (rc != KX_OK) return rc;
        part_o = (float *)scratch;
        part_ml = part_o + rows * hd;
        rc = op_attn_prefix(d_qkv, qkv_ld, d_pages, d_block_table + (size_t)prefix_row * max_pages, d_pos, d_rope, B, n_q, n_kv, hd,
                            prefix_len, part_o, part_ml, &n_part, st);
        if (rc != KX_OK) return rc;
        key0 = prefix_len;
    }
    return op_attn_decode(d_qkv, qkv_ld, d_out, out_ld, d_pages, d_block_table, max_pages, d_pos, d_rope, B, max_T, n_q, n_kv, hd,
                          key0, part_o, part_ml, n_part, st);
}

extern "C" int kx_rmsnorm(float *d_x, const void *d_g, void *d_y, int32_t M, int32_t H, float eps, const int32_t *d_row_map,
                          const float *d_part, int32_t splits, int64_t split_stride, void *stream) {
    using namespace kx;
    KX_REQUIRE(d_x && d_g && d_y && M >= 1 && H >= 8 && H % 8 == 0, KX_EINVAL, "kx_rmsnorm: bad argument");
    KX_REQUIRE(splits == 0 || (d_part && splits >= 1 && split_stride >= (int64_t)M * H), KX_EINVAL, "kx_rmsnorm: bad split-K partials");
    return op_rmsnorm(d_x, d_g, d_y, M, H, eps, d_row_map, (cudaStream_t)stream, d_part, splits, (size_t)split_stride);
}

extern "C" int kx_argmax_parts(const void *d_parts, int32_t n_parts, int32_t ld, int32_t B, int32_t *d_ids, void *stream) {
    using namespace kx;
    KX_REQUIRE(d_parts && d_ids && n_parts >= 1 && B >= 1 && ld >= B, KX_EINVAL, "kx_argmax_parts: bad argument");
    return op_argmax_parts(d_parts, n_parts, ld, B, 0, d_ids, nullptr, nullptr, (cudaStream_t)stream);
}
