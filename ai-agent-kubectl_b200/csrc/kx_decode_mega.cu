// Decode step as ONE persistent kernel (SURVEY 8 rows k3-k7 for a small decode batch).
//
// With a batch of <= 128 sequences every decode kernel is short (weights of one linear layer
// stream in 2-10 us), so a layer run as seven launches spends most of its time in launch, ramp
// and drain latency: ~6 us per kernel against ~19 us of HBM time per Llama-3.2-1B layer.  Here
// all layers of a step run inside one cooperative grid of one CTA per SM:
//
//   phase 0  residual += partial sums of the previous layer's down projection; RMSNorm -> xn
//   phase 1  QKV projection (split-K partial sums, fp32)                       [weights stream]
//   phase 2  attention: sum the partials (+bias), RoPE, append k/v, one warp per (sequence, kv head)
//   phase 3  output projection (split-K partial sums)                          [weights stream]
//   phase 4  residual += partials; RMSNorm -> xn
//   phase 5  gate/up projection, SwiGLU in the epilogue -> mlp                  [weights stream]
//   phase 6  down projection (split-K partial sums)                            [weights stream]
//
// separated by grid barriers (one arrival counter per phase in global memory, so a CTA with nothing to do
// in a phase may run ahead and arrive early without being mistaken for a straggler of the phase before;
// arrive = release add, wait = acquire poll).  The point of the design is the TMA producer warp: it walks the CTA's whole
// schedule of weight tiles for ALL phases and layers and keeps the shared-memory ring full of
// weights no matter which phase the rest of the CTA is in -- weights never depend on the step's
// data -- while the small activation operand of a stage is requested only once the barrier of
// the phase that produces it has been passed.  HBM therefore keeps streaming through the
// barriers, the norms and the attention, and a layer costs about max(HBM time, sum of phase
// latencies) instead of their sum.  Split-K partial sums go to global scratch (L2 resident)
// and are reduced, in a fixed order, by the phase that consumes them -- no clusters, no atomics
// on data, deterministic.
//
// Roles in a CTA (224 threads): warp 0 = TMA producer for weights, warp 6 = TMA producer for activations,
// warp 1 = tcgen05.mma issuer (two accumulators in TMEM), warps 2-5 = workers (GEMM epilogues, norms,
// attention, barrier arrival).
// GEMMs are swap-AB like kx_gemm_bf16_ksplit_kernel: A = 128 weight rows, N = the batch.
#include <cuda_bf16.h>
#include "kx_common.cuh"
#include "kx_sm100.cuh"
#include "kx_attn.cuh"
#include "kx_mega.cuh"

namespace kx {

using namespace sm100;
using namespace attn;

namespace {

constexpr int BM = 128, BK = 64;
constexpr int kThreads = 224, kWorkers = 128;   // warps: 0 weights, 1 MMA, 2-5 workers, 6 activations
constexpr int kPhases = 7;                            // per layer
constexpr int kSmemLimit = 232448;                    // 227 KB per CTA

__host__ __device__ constexpr int gemm_phase(int g) { return g == 0 ? 1 : g == 1 ? 3 : g == 2 ? 5 : 6; }

template <int BN, int HD> struct MegaCfg {
    static constexpr int KB = HD == 64 ? 32 : 16, NS = 2;  // (16-key stages, 3 deep, free one more ring stage but triple the attention time)
    using W = WarpAttnSmem<HD, KB, NS>;
    static constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int ATTN_BYTES = 4 * W::BYTES;
    static constexpr int TAIL_BYTES = 512;            // barriers, TMEM holder, reduction scratch
    static constexpr int STAGES = (kSmemLimit - 1024 - TAIL_BYTES - ATTN_BYTES) / STAGE_BYTES;
    static constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + ATTN_BYTES + TAIL_BYTES;
    static constexpr int TMEM_COLS = 2 * BN;           // two accumulators: the MMAs of the next item overlap the epilogue
    static_assert(STAGES >= 3, "ring too shallow");
};

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + __expf(-x)); }
__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ void workers_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// position in a CTA's schedule of weight tiles: layer, GEMM, work item, k block
struct Cursor {
    int l, g, item, kb, nkb, tile, kb0;
    bool done;
};
__device__ __forceinline__ void cursor_settle(Cursor &c, const MegaParams &p, int cta, int grid) {
    while (!c.done && c.item >= p.g[c.g].items) {     // no (more) items in this GEMM for this CTA
        c.item = cta;
        if (++c.g == 4) { c.g = 0; if (++c.l == p.layers) c.done = true; }
    }
    if (c.done) return;
    const MegaGemm &d = p.g[c.g];
    c.tile = c.item % d.tiles;
    const int split = c.item / d.tiles;
    c.kb0 = split * d.kb_per;
    c.nkb = min(d.kb_per, d.kb_total - c.kb0);
    c.kb = 0;
    (void)grid;
}
__device__ __forceinline__ void cursor_next(Cursor &c, const MegaParams &p, int cta, int grid) {
    if (++c.kb < c.nkb) return;
    c.item += grid;
    cursor_settle(c, p, cta, grid);
}

// sum of the split-K partials of 8 consecutive features of one token, in split order, plus bias, rounded to
// bf16 like the stand-alone QKV epilogue rounds (returned as fp32 values of those bf16 numbers)
__device__ __forceinline__ void gather8(const float *__restrict__ part, size_t split_stride, int splits, const float *__restrict__ bias,
                                        float (&v)[8]) {
    float4 a = *reinterpret_cast<const float4 *>(part), b = *reinterpret_cast<const float4 *>(part + 4);
    for (int s0 = 1; s0 < splits; s0 += 2) {           // two splits' loads in flight together, added in split order
        float4 c[2], d[2];
#pragma unroll
        for (int j = 0; j < 2; ++j)
            if (s0 + j < splits) {
                c[j] = *reinterpret_cast<const float4 *>(part + (s0 + j) * split_stride);
                d[j] = *reinterpret_cast<const float4 *>(part + (s0 + j) * split_stride + 4);
            }
#pragma unroll
        for (int j = 0; j < 2; ++j)
            if (s0 + j < splits) {
                a.x += c[j].x; a.y += c[j].y; a.z += c[j].z; a.w += c[j].w;
                b.x += d[j].x; b.y += d[j].y; b.z += d[j].z; b.w += d[j].w;
            }
    }
    if (bias) {
        const float4 c = *reinterpret_cast<const float4 *>(bias), d = *reinterpret_cast<const float4 *>(bias + 4);
        a.x += c.x; a.y += c.y; a.z += c.z; a.w += c.w;
        b.x += d.x; b.y += d.y; b.z += d.z; b.w += d.w;
    }
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
#pragma unroll
    for (int t = 0; t < 8; ++t) v[t] = __bfloat162float(__float2bfloat16(v[t]));
}

template <int BN, int HD, int G>
__global__ void __launch_bounds__(kThreads, 1)
kx_decode_mega_kernel(const __grid_constant__ CUtensorMap xn_map, const __grid_constant__ CUtensorMap attn_map,
                      const __grid_constant__ CUtensorMap mlp_map, const __grid_constant__ MegaParams p) {
    using C = MegaCfg<BN, HD>;
    using W = typename C::W;
    constexpr int KB = C::KB, NS = C::NS, NST = C::STAGES;
    constexpr int CH = HD / 8, HALF = HD / 2;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t *ring = smem_raw + (base - smem_u32(smem_raw));
    uint8_t *attn_smem = ring + NST * C::STAGE_BYTES;
    uint64_t *full = reinterpret_cast<uint64_t *>(attn_smem + C::ATTN_BYTES);
    uint64_t *empty = full + NST;
    uint64_t *tmem_full = empty + NST;            // [2]
    uint64_t *tmem_empty = tmem_full + 2;         // [2]
    uint32_t *tmem_holder = reinterpret_cast<uint32_t *>(tmem_empty + 2);
    float *red = reinterpret_cast<float *>(tmem_holder + 2);          // [4] sum-of-squares partials of the worker warps

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cta = blockIdx.x, grid = gridDim.x;
    const int B = p.B, H = p.hidden;

    if (threadIdx.x == 0) {
        tma_prefetch_desc(&xn_map); tma_prefetch_desc(&attn_map); tma_prefetch_desc(&mlp_map);
        for (int s = 0; s < NST; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tmem_full + i, 1); mbar_init(tmem_empty + i, 4); }   // empty: one arrival per worker warp
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_holder, C::TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_holder;

    if (warp == 0) {
        // ================= TMA producer, weights: runs ahead of everything, limited only by free ring stages =================
        if (lane == 0) {
            const uint64_t pol_w = policy_evict_first();
            Cursor w{0, 0, cta, 0, 0, 0, 0, false};
            cursor_settle(w, p, cta, grid);
            for (uint32_t it = 0; !w.done; ++it) {
                const int s = it % NST;
                mbar_wait(empty + s, ((it / NST) & 1) ^ 1);                      // stage consumed by the MMA warp
                mbar_expect_tx(full + s, C::STAGE_BYTES);                        // weights + activations of this stage
                tma_load_2d_hint(ring + s * C::STAGE_BYTES, p.wmaps + w.l * 4 + w.g, full + s, (w.kb0 + w.kb) * BK, w.tile * BM, pol_w);
                cursor_next(w, p, cta, grid);
            }
        }
    } else if (warp == 6) {
        // ================= TMA producer, activations: the same stage sequence, gated by the phase barriers =================
        if (lane == 0) {
            const uint64_t pol_x = policy_evict_last();
            Cursor a{0, 0, cta, 0, 0, 0, 0, false};
            cursor_settle(a, p, cta, grid);
            int ready_phase = -1;                                                // highest phase known to be complete
            for (uint32_t it = 0; !a.done; ++it) {
                const int s = it % NST;
                mbar_wait(empty + s, ((it / NST) & 1) ^ 1);
                const int need = a.l * kPhases + gemm_phase(a.g) - 1;            // the phase that writes this GEMM's operand
                if (ready_phase < need) {
                    while (ld_acquire_gpu(p.bar + need) < (uint32_t)grid) __nanosleep(20);
                    ready_phase = need;
                    fence_proxy_async();                                         // generic-proxy writes -> async-proxy (TMA) reads
                }
                const CUtensorMap *am = a.g == 1 ? &attn_map : a.g == 3 ? &mlp_map : &xn_map;
                tma_load_2d_hint(ring + s * C::STAGE_BYTES + C::A_BYTES, am, full + s, (a.kb0 + a.kb) * BK, 0, pol_x);
                cursor_next(a, p, cta, grid);
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16_f32(BM, BN);
            uint32_t it = 0, acc_it = 0;
            for (int l = 0; l < p.layers; ++l) {
                for (int g = 0; g < 4; ++g) {
                    const MegaGemm &d = p.g[g];
                    for (int item = cta; item < d.items; item += grid, ++acc_it) {
                        const int kb0 = (item / d.tiles) * d.kb_per;
                        const int nkb = min(d.kb_per, d.kb_total - kb0);
                        const uint32_t ab = acc_it & 1, acc = tmem + ab * BN;
                        mbar_wait(tmem_empty + ab, ((acc_it >> 1) & 1) ^ 1);        // workers drained this accumulator
                        tc_fence_after();
                        for (int i = 0; i < nkb; ++i, ++it) {
                            const int s = it % NST;
                            mbar_wait(full + s, (it / NST) & 1);
                            tc_fence_after();
                            const uint32_t sa = base + s * C::STAGE_BYTES;
                            const uint64_t da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + C::A_BYTES);
#pragma unroll
                            for (int k = 0; k < BK / 16; ++k) umma_bf16(acc, da + 2 * k, db + 2 * k, idesc, (i | k) != 0);
                            umma_commit(empty + s);
                        }
                        umma_commit(tmem_full + ab);
                    }
                }
            }
        }
    } else if (warp < 6) {
        // ================= workers =================
        const int wq = warp & 3, tw = threadIdx.x - 64;                 // TMEM lane quarter this warp may read (warp id mod 4); worker thread
        uint32_t acc_it = 0;
        int n_trace = 0;
        auto trace = [&](uint32_t tag) {                                // diagnostic: (tag, globaltimer) pairs of CTA 0
            if (p.trace && cta == 0 && n_trace < p.trace_cap) {
                uint64_t t;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                p.trace[2 * n_trace] = tag; p.trace[2 * n_trace + 1] = t;
                ++n_trace;
            }
        };
        auto grid_arrive = [&](uint32_t phase) {                        // this CTA's part of `phase` is written
            workers_sync();
            if (tw == 0) { __threadfence(); red_release_gpu_add(p.bar + phase, 1u); trace(phase * 2); }
        };
        auto grid_wait = [&](uint32_t phase) {                          // every CTA has finished `phase`
            if (tw == 0) { while (ld_acquire_gpu(p.bar + phase) < (uint32_t)grid) __nanosleep(20); trace(phase * 2 + 1); }
            workers_sync();
        };
        if (tw == 0) trace(0xffffffffu);
        // residual += split-K partials (fixed order); optionally RMSNorm -> xn.  One row per CTA at a time.
        auto reduce_norm = [&](const float *part, int splits, const __nv_bfloat16 *gamma) {
            const int n4 = H / 4;
            for (int b = cta; b < B; b += grid) {
                float *xr = p.resid + (size_t)b * H;
                float4 xv[8];
                float ss = 0.f;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int i = u * kWorkers + tw;
                    if (i >= n4) { xv[u] = make_float4(0.f, 0.f, 0.f, 0.f); continue; }
                    float4 x = reinterpret_cast<const float4 *>(xr)[i];
                    if (splits > 0) {
                        float4 v = reinterpret_cast<const float4 *>(part + (size_t)b * H)[i];
                        for (int s0 = 1; s0 < splits; s0 += 4) {        // four splits' loads in flight, added in split order
                            float4 c[4];
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (s0 + j < splits) c[j] = reinterpret_cast<const float4 *>(part + ((size_t)(s0 + j) * B + b) * H)[i];
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (s0 + j < splits) { v.x += c[j].x; v.y += c[j].y; v.z += c[j].z; v.w += c[j].w; }
                        }
                        x.x += v.x; x.y += v.y; x.z += v.z; x.w += v.w;
                        reinterpret_cast<float4 *>(xr)[i] = x;
                    }
                    xv[u] = x;
                    ss += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
                }
                if (!gamma) continue;
                ss = warp_sum_f(ss);
                if (lane == 0) red[wq] = ss;
                workers_sync();
                const float r = 1.0f / sqrtf((red[0] + red[1] + red[2] + red[3]) / (float)H + p.eps);
                __nv_bfloat16 *yr = reinterpret_cast<__nv_bfloat16 *>(p.xn) + (size_t)b * H;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int i = u * kWorkers + tw;
                    if (i >= n4) continue;
                    const uint2 gq = reinterpret_cast<const uint2 *>(gamma)[i];
                    const __nv_bfloat162 *gh = reinterpret_cast<const __nv_bfloat162 *>(&gq);
                    uint2 o;
                    o.x = pack2(__fmul_rn(__fmul_rn(xv[u].x, r), __low2float(gh[0])), __fmul_rn(__fmul_rn(xv[u].y, r), __high2float(gh[0])));
                    o.y = pack2(__fmul_rn(__fmul_rn(xv[u].z, r), __low2float(gh[1])), __fmul_rn(__fmul_rn(xv[u].w, r), __high2float(gh[1])));
                    reinterpret_cast<uint2 *>(yr)[i] = o;
                }
                workers_sync();                                          // red[] is reused by the next row
            }
        };
        // epilogue of this CTA's items of one GEMM: accumulator rows = features, columns = tokens
        auto gemm_epilogue = [&](int g, float *part, __nv_bfloat16 *swiglu_out) {
            const MegaGemm &d = p.g[g];
            for (int item = cta; item < d.items; item += grid, ++acc_it) {
                const uint32_t ab = acc_it & 1;
                const uint32_t taddr = tmem + ((uint32_t)(wq * 32) << 16) + ab * BN;
                const int tile = item % d.tiles, split = item / d.tiles;
                const int f = tile * BM + wq * 32 + lane;
                const bool ok = f < d.feats;
                mbar_wait(tmem_full + ab, (acc_it >> 1) & 1);
                tc_fence_after();
#pragma unroll 1
                for (int c = 0; c < BN; c += 32) {
                    float v[32];
                    tmem_ld_32x32(taddr + c, reinterpret_cast<uint32_t(&)[32]>(v));
                    tmem_ld_wait();
                    if (c + 32 >= BN) {                                  // last read of the accumulator: hand it back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(tmem_empty + ab);
                    }
                    if (swiglu_out) {                                    // features are (gate, up) pairs
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float other = __shfl_xor_sync(0xffffffffu, v[j], 1);
                            if (ok && !(f & 1) && c + j < B)
                                swiglu_out[(size_t)(c + j) * (d.feats >> 1) + (f >> 1)] = __float2bfloat16(silu_f(v[j]) * other);
                        }
                    } else if (ok) {
                        float *dst = part + ((size_t)split * B + c) * d.feats + f;
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (c + j < B) dst[(size_t)j * d.feats] = v[j];
                    }
                }
            }
        };

        for (int l = 0; l < p.layers; ++l) {
            const MegaLayer &L = p.layers_ptr[l];
            const uint32_t ph0 = (uint32_t)l * kPhases;
            // ---- phase 0: (residual += down-projection partials of the previous layer) + RMSNorm ----
            if (l > 0) grid_wait(ph0 - 1);
            reduce_norm(p.part_d, l > 0 ? p.g[3].splits : 0, reinterpret_cast<const __nv_bfloat16 *>(L.ln1));
            grid_arrive(ph0);
            // ---- phase 1: QKV projection ----
            gemm_epilogue(0, p.part_qkv, nullptr);
            grid_arrive(ph0 + 1);
            // ---- phase 2: attention ----
            {
                const int n_tasks = B * p.n_kv, stride = grid * 4;
                uint8_t *Qs = attn_smem + wq * W::BYTES;
                const uint32_t ring_u = smem_u32(Qs) + W::QT;
                __nv_bfloat16 *knew = reinterpret_cast<__nv_bfloat16 *>(Qs + W::QT + NS * 2 * W::ST), *vnew = knew + HD;
                __nv_bfloat16 *pages = reinterpret_cast<__nv_bfloat16 *>(L.kv_pages);
                const size_t sstride = (size_t)B * p.qkv_dim;            // between split-K partials
                for (int task = cta * 4 + wq, round = 0; round == 0 || task < n_tasks; task += stride, ++round) {
                    const bool live = task < n_tasks;
                    const int b = live ? task / p.n_kv : 0, kvh = live ? task - b * p.n_kv : 0;
                    const int pp = p.pos[b], T = pp + 1;
                    const int *bt = p.block_table + (size_t)b * p.max_pages;
                    const int nblk = live ? (T + KB - 1) / KB : 0;
#pragma unroll
                    for (int i = 0; i < NS - 1; ++i) {                   // cached K/V do not depend on this step
                        if (i < nblk) warp_attn_fetch<HD, KB, NS>(ring_u, i, 0, T, pp, pages, bt, p.n_kv, kvh, lane);
                        cp_async_commit();
                    }
                    if (round == 0) grid_wait(ph0 + 1);                  // QKV partials of every CTA are in place
                    if (!live) { cp_async_wait<0>(); continue; }
                    {
                        const float *prow = p.part_qkv + (size_t)b * p.qkv_dim;
                        const float2 *cs = p.rope + (size_t)pp * HALF;
                        __nv_bfloat16 *kdst = pages + (((size_t)bt[pp / kPage] * 2) * p.n_kv + kvh) * kPage * HD + (size_t)(pp % kPage) * HD;
                        __nv_bfloat16 *vdst = kdst + (size_t)p.n_kv * kPage * HD;
                        for (int i = lane; i < (G + 1) * (HALF / 8); i += 32) {
                            const int h = i / (HALF / 8), c = i - h * (HALF / 8);
                            const int f0 = (h < G ? (kvh * G + h) : (p.n_q + kvh)) * HD + c * 8;
                            float x1[8], x2[8];
                            gather8(prow + f0, sstride, p.g[0].splits, L.bqkv ? L.bqkv + f0 : nullptr, x1);
                            gather8(prow + f0 + HALF, sstride, p.g[0].splits, L.bqkv ? L.bqkv + f0 + HALF : nullptr, x2);
                            uint32_t o1[4], o2[4];
#pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                const float4 cc = *reinterpret_cast<const float4 *>(cs + c * 8 + 2 * t);
                                o1[t] = pack2(__fsub_rn(__fmul_rn(x1[2 * t], cc.x), __fmul_rn(x2[2 * t], cc.y)),
                                              __fsub_rn(__fmul_rn(x1[2 * t + 1], cc.z), __fmul_rn(x2[2 * t + 1], cc.w)));
                                o2[t] = pack2(__fadd_rn(__fmul_rn(x2[2 * t], cc.x), __fmul_rn(x1[2 * t], cc.y)),
                                              __fadd_rn(__fmul_rn(x2[2 * t + 1], cc.z), __fmul_rn(x1[2 * t + 1], cc.w)));
                            }
                            const uint4 lo = make_uint4(o1[0], o1[1], o1[2], o1[3]), hi = make_uint4(o2[0], o2[1], o2[2], o2[3]);
                            if (h < G) {
                                *reinterpret_cast<uint4 *>(Qs + swz<HD>(h, c)) = lo;
                                *reinterpret_cast<uint4 *>(Qs + swz<HD>(h, c + HALF / 8)) = hi;
                            } else {
                                reinterpret_cast<uint4 *>(kdst)[c] = lo; reinterpret_cast<uint4 *>(kdst)[c + HALF / 8] = hi;
                                reinterpret_cast<uint4 *>(knew)[c] = lo; reinterpret_cast<uint4 *>(knew)[c + HALF / 8] = hi;
                            }
                        }
                        for (int i = lane; i < CH; i += 32) {
                            const int f0 = (p.n_q + p.n_kv + kvh) * HD + i * 8;
                            float x[8];
                            gather8(prow + f0, sstride, p.g[0].splits, L.bqkv ? L.bqkv + f0 : nullptr, x);
                            const uint4 vv = make_uint4(pack2(x[0], x[1]), pack2(x[2], x[3]), pack2(x[4], x[5]), pack2(x[6], x[7]));
                            reinterpret_cast<uint4 *>(vdst)[i] = vv; reinterpret_cast<uint4 *>(vnew)[i] = vv;
                        }
                        for (int i = lane; i < (16 - G) * CH; i += 32)
                            *reinterpret_cast<uint4 *>(Qs + swz<HD>(G + i / CH, i % CH)) = make_uint4(0, 0, 0, 0);
                    }
                    __syncwarp();
                    warp_attn_main<HD, G, KB, NS>(Qs, nblk, 0, T, pp, pages, bt, p.n_kv, kvh, lane, p.scale_log2,
                                                  reinterpret_cast<__nv_bfloat16 *>(p.attn) + (size_t)b * p.q_dim + (size_t)kvh * G * HD, nullptr, nullptr, 0, 0, 0);
                    __syncwarp();
                }
            }
            grid_arrive(ph0 + 2);
            // ---- phase 3: output projection ----
            gemm_epilogue(1, p.part_o, nullptr);
            grid_arrive(ph0 + 3);
            // ---- phase 4: residual += partials; RMSNorm ----
            grid_wait(ph0 + 3);
            reduce_norm(p.part_o, p.g[1].splits, reinterpret_cast<const __nv_bfloat16 *>(L.ln2));
            grid_arrive(ph0 + 4);
            // ---- phase 5: gate/up projection + SwiGLU ----
            gemm_epilogue(2, nullptr, reinterpret_cast<__nv_bfloat16 *>(p.mlp));
            grid_arrive(ph0 + 5);
            // ---- phase 6: down projection ----
            gemm_epilogue(3, p.part_d, nullptr);
            grid_arrive(ph0 + 6);
        }
        // residual += down-projection partials of the last layer (the head runs as separate kernels)
        grid_wait((uint32_t)p.layers * kPhases - 1);
        reduce_norm(p.part_d, p.g[3].splits, nullptr);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, C::TMEM_COLS);
}

template <int BN, int HD, int G>
int launch_mega(const CUtensorMap &xn, const CUtensorMap &at, const CUtensorMap &ml, const MegaParams &p, int grid, cudaStream_t st) {
    using C = MegaCfg<BN, HD>;
    auto kern = kx_decode_mega_kernel<BN, HD, G>;
    static PerDeviceOnce once;
    if (once.first()) {
        KX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
        once.done();
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = C::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;      // every CTA must be resident: the phases meet at a grid barrier
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    KX_CUDA(cudaLaunchKernelEx(&cfg, kern, xn, at, ml, p));
    count_launch();
    return KX_OK;
}

template <int BN>
int launch_mega_bn(int hd, int G, const CUtensorMap &xn, const CUtensorMap &at, const CUtensorMap &ml, const MegaParams &p, int grid,
                   cudaStream_t st) {
    if (hd == 64 && G == 2) return launch_mega<BN, 64, 2>(xn, at, ml, p, grid, st);
    if (hd == 64 && G == 4) return launch_mega<BN, 64, 4>(xn, at, ml, p, grid, st);
    if (hd == 64 && G == 7) return launch_mega<BN, 64, 7>(xn, at, ml, p, grid, st);
    if (hd == 128 && G == 2) return launch_mega<BN, 128, 2>(xn, at, ml, p, grid, st);
    if (hd == 128 && G == 4) return launch_mega<BN, 128, 4>(xn, at, ml, p, grid, st);
    set_error("decode megakernel: head_dim %d with %d query heads per kv head is not instantiated", hd, G);
    return KX_EINVAL;
}

}  // namespace

bool mega_supported(int hd, int G, int hidden) {
    const bool shape = (hd == 64 && (G == 2 || G == 4 || G == 7)) || (hd == 128 && (G == 2 || G == 4));
    return shape && hidden % 4 == 0 && hidden <= 8 * kWorkers * 4;
}

int mega_bn(int B) { return B <= 64 ? 64 : B <= 128 ? 128 : 0; }

MegaGemm mega_plan_gemm(int feats, int k, int grid, bool may_split) {
    MegaGemm d{};
    d.feats = feats;
    d.tiles = (feats + BM - 1) / BM;
    d.kb_total = (k + BK - 1) / BK;
    int s = may_split ? grid / d.tiles : 1;
    if (s > d.kb_total) s = d.kb_total;
    if (s < 1) s = 1;
    d.kb_per = (d.kb_total + s - 1) / s;
    d.splits = (d.kb_total + d.kb_per - 1) / d.kb_per;
    d.items = d.tiles * d.splits;
    return d;
}

int mega_launch(const CUtensorMap &xn, const CUtensorMap &at, const CUtensorMap &ml, const MegaParams &p, int bn, int hd, int grid,
                cudaStream_t st) {
    const int G = p.n_q / p.n_kv;
    KX_CUDA(cudaMemsetAsync(p.bar, 0, sizeof(uint32_t) * (size_t)p.layers * kPhases, st));
    if (bn == 64) return launch_mega_bn<64>(hd, G, xn, at, ml, p, grid, st);
    if (bn == 128) return launch_mega_bn<128>(hd, G, xn, at, ml, p, grid, st);
    set_error("decode megakernel: batch tile %d not supported", bn);
    return KX_EINVAL;
}

}  // namespace kx
