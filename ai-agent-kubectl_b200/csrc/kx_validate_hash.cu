// k9: fused validator + parser + sanitiser + cache-key hash over a CSR batch of
// strings.  Replaces, per string,
//   is_safe_kubectl_command   /root/reference/app.py:72-88   (shlex.split: CPython Lib/shlex.py:133-277)
//   KubectlOutputParser.parse /root/reference/app.py:95-104
//   sanitize_query            /root/reference/app.py:60-68
//   hash(sanitised)           CPython Python/pyhash.c siphash13, zero key; reached at app.py:312,322
//
// Layout: one CTA of 256 threads owns 256 consecutive strings.  The CTA's byte
// range is contiguous in the CSR buffer, so it is staged into shared memory with
// coalesced 16-byte loads (each HBM byte is read exactly once).  Strings are
// handed to threads sorted by length (a counting sort in shared memory) so the
// lanes of a warp finish together.  Each thread then runs the table-driven ASCII
// fast path of k9_core.cuh on its string.  Strings it cannot finish go to a queue in
// global scratch memory: the ASCII strings that need a byte loop (mixed quoting,
// whitespace to collapse) to the back of it, for pass 1b, where every lane has such a
// loop to run; non-ASCII strings, fenced outputs in PARSE mode and CTAs whose byte range
// exceeds the staging buffer to the front, for pass 2, the general code-point path.
#include <atomic>
#include "kx_common.cuh"
#include "k9_core.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kStageBytes = 32 * 1024;
constexpr int kStageSlack = 16;           // sip_stream may read two words past the last string
constexpr int kGeneralThreads = 128;
constexpr int kLoopThreads = 128;         // pass 1b: strings that need a byte loop
constexpr int kLoopSlotWords = 65;        // per-thread staging slot: 64 words + 1 (slots start in different banks)
constexpr int kLoopMaxLen = 228;          // longest string a slot takes: 15 alignment bytes + 228 + two words of look-ahead <= 256

__device__ const k9::FastTables g_tables = k9::make_fast_tables();

struct __align__(16) Shared {
    k9::FastTables tab;
    int32_t soff[kThreads + 1];           // string offsets relative to the staged base
    int32_t hist[64];
    int32_t perm[kThreads];
    int32_t slowq[kThreads];              // strings for the general pass
    int32_t loopq[kThreads];              // strings for the byte-loop pass
    int32_t nslow, nloop;
    uint32_t qbase, qbase_loop;
};

__device__ __forceinline__ void store_result(const k9::Result &r, int64_t s, uint8_t *verdict, int64_t *hash,
                                             int32_t *span, int32_t *san_len) {
    if (hash) hash[s] = r.hash;
    if (san_len) san_len[s] = r.san_len;
    if (verdict) verdict[s] = (uint8_t)r.verdict;
    if (span) { span[2 * s] = r.span_a; span[2 * s + 1] = r.span_b; }
}

// Pass 1: table-driven ASCII fast path out of shared memory.  Strings it cannot
// finish are appended to `queue` (scratch: queue[0] is the element count, entries
// start at queue[2]; entries are indices relative to this launch's first string).
template <bool PARSE>
__global__ void __launch_bounds__(kThreads)
k9_validate_hash_fast(const uint8_t *__restrict__ data, const int64_t *__restrict__ off, int64_t n,
                      uint8_t *__restrict__ verdict, int64_t *__restrict__ hash, int32_t *__restrict__ span,
                      uint8_t *__restrict__ san, int32_t *__restrict__ san_len, uint32_t *__restrict__ queue) {
    extern __shared__ uint4 stage[];                       // kStageBytes + slack, then Shared
    Shared &sh = *reinterpret_cast<Shared *>(reinterpret_cast<uint8_t *>(stage) + kStageBytes + kStageSlack);
    const int tid = threadIdx.x;
    const int64_t s0 = (int64_t)blockIdx.x * kThreads;
    const int cnt = (int)min((int64_t)kThreads, n - s0);
    const int64_t g_beg = off[s0], g_end = off[s0 + cnt];
    const uintptr_t addr = (uintptr_t)(data + g_beg);
    const uint4 *abase = (const uint4 *)(addr & ~(uintptr_t)15);
    const int lead = (int)(addr & 15);
    const int64_t nbytes = lead + (g_end - g_beg);
    const bool staged = nbytes <= kStageBytes;             // CTA-uniform

    if (tid == 0) { sh.nslow = 0; sh.nloop = 0; }
    if (staged) {
        // ---- stage bytes (coalesced 16 B loads), tables and offsets ----
        const int nvec = (int)((nbytes + 15) >> 4);
        for (int v = tid; v < nvec; v += kThreads) stage[v] = __ldg(abase + v);
        const uint32_t *src = reinterpret_cast<const uint32_t *>(&g_tables);
        uint32_t *dst = reinterpret_cast<uint32_t *>(&sh.tab);
        for (int i = tid; i < (int)(sizeof(k9::FastTables) / 4); i += kThreads) dst[i] = src[i];
        for (int i = tid; i <= cnt; i += kThreads) sh.soff[i] = (int32_t)(off[s0 + i] - g_beg) + lead;
        if (tid < 64) sh.hist[tid] = 0;
        __syncthreads();

        // ---- give each warp strings of similar length (counting sort on len/4: the scan loop runs once per word) ----
        int bucket = 0, rank = 0;
        if (tid < cnt) {
            bucket = min((sh.soff[tid + 1] - sh.soff[tid]) >> 2, 63);
            rank = atomicAdd(&sh.hist[bucket], 1);
        }
        __syncthreads();
        if (tid < 32) {
            const int c0 = sh.hist[2 * tid], c1 = sh.hist[2 * tid + 1];
            int incl = c0 + c1;
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (tid >= d) incl += t;
            }
            sh.hist[2 * tid] = incl - c0 - c1;            // exclusive prefix
            sh.hist[2 * tid + 1] = incl - c1;
        }
        __syncthreads();
        if (tid < cnt) sh.perm[sh.hist[bucket] + rank] = tid;
        __syncthreads();

        if (tid < cnt) {
            const int i = sh.perm[tid];
            k9::Result r;
            const int how = k9::fast_string<PARSE, false>(reinterpret_cast<uint32_t *>(stage), sh.soff[i], sh.soff[i + 1],
                                                          sh.tab.flags, hash != nullptr,
                                                          san ? san + g_beg + (sh.soff[i] - lead) : nullptr, r);
            if (how == 0) store_result(r, s0 + i, verdict, hash, span, san_len);
            else if (how == 2 && sh.soff[i + 1] - sh.soff[i] <= kLoopMaxLen) sh.loopq[atomicAdd(&sh.nloop, 1)] = i;
            else sh.slowq[atomicAdd(&sh.nslow, 1)] = i;
        }
    } else {
        // strings too long to stage: the whole CTA goes to the general pass
        __syncthreads();
        if (tid < cnt) sh.slowq[atomicAdd(&sh.nslow, 1)] = tid;
    }
    __syncthreads();
    const int nslow = sh.nslow, nloop = sh.nloop;
    if (nslow == 0 && nloop == 0) return;
    if (tid == 0) {
        if (nslow) sh.qbase = atomicAdd(queue, (uint32_t)nslow);
        if (nloop) sh.qbase_loop = atomicAdd(queue + 1, (uint32_t)nloop);
    }
    __syncthreads();
    // one array, two queues: the general pass's entries grow from the front, the byte-loop pass's from the back
    // (a string is in at most one of them, so together they never exceed n entries)
    for (int q = tid; q < nslow; q += kThreads) queue[2 + sh.qbase + q] = (uint32_t)(s0 + sh.slowq[q]);
    for (int q = tid; q < nloop; q += kThreads) queue[2 + n - 1 - (sh.qbase_loop + q)] = (uint32_t)(s0 + sh.loopq[q]);
}

// Pass 1b: the ASCII strings that need one of fast_string's byte loops -- the quoting automaton (mixed quotes or a
// backslash) or the whitespace collapse; about one string in fifteen of the synthetic mix.  Inside pass 1 such a string
// made its whole warp walk the loops with one or two lanes active (19 of 32 lanes active on average, a third of the
// warp's instructions); here every lane has one.  Each thread copies its string (at most kLoopMaxLen bytes, a few L2
// hits: pass 1 has just read it) into a private shared-memory slot and runs the same fast_string with the loops enabled.
template <bool PARSE>
__global__ void __launch_bounds__(kLoopThreads)
k9_validate_hash_loops(const uint8_t *__restrict__ data, const int64_t *__restrict__ off, int64_t n,
                       uint8_t *__restrict__ verdict, int64_t *__restrict__ hash, int32_t *__restrict__ span,
                       uint8_t *__restrict__ san, int32_t *__restrict__ san_len, const uint32_t *__restrict__ queue) {
    __shared__ uint32_t slots[kLoopThreads * kLoopSlotWords];
    __shared__ uint8_t lut[256];
    __shared__ int32_t hist[32];
    __shared__ uint32_t order[kLoopThreads];
    const int tid = threadIdx.x;
    for (int i = tid; i < 256; i += kLoopThreads) lut[i] = g_tables.flags[i];
    const uint32_t count = queue[1];
    uint32_t *slot = slots + tid * kLoopSlotWords;
    for (uint32_t base = blockIdx.x * kLoopThreads; base < count; base += gridDim.x * kLoopThreads) {
        // the CTA's 128 queue entries, handed to its threads by length (counting sort on len / 8): both loops are
        // O(length), and the queue arrives in the order pass 1's CTAs finished
        const int cnt = (int)min((uint32_t)kLoopThreads, count - base);
        if (tid < 32) hist[tid] = 0;
        __syncthreads();
        uint32_t sidx = 0;
        int bucket = 0, rank = 0;
        if (tid < cnt) {
            sidx = queue[2 + n - 1 - (base + tid)];
            bucket = (int)min((off[sidx + 1] - off[sidx]) >> 3, (int64_t)31);
            rank = atomicAdd(&hist[bucket], 1);
        }
        __syncthreads();
        if (tid < 32) {
            const int c = hist[tid];
            int incl = c;
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (tid >= d) incl += t;
            }
            hist[tid] = incl - c;
        }
        __syncthreads();
        if (tid < cnt) order[hist[bucket] + rank] = sidx;
        __syncthreads();
        if (tid < cnt) {
            const int64_t s = order[tid];
            const int64_t o0 = off[s], o1 = off[s + 1];
            const uintptr_t sa = (uintptr_t)(data + o0);
            const uint4 *src = (const uint4 *)(sa & ~(uintptr_t)15);
            const int lo = (int)(sa & 15), hi = lo + (int)(o1 - o0);
            const int nvec = (hi + 15) >> 4;
            for (int v = 0; v < nvec; ++v) {
                const uint4 x = __ldg(src + v);
                slot[4 * v] = x.x; slot[4 * v + 1] = x.y; slot[4 * v + 2] = x.z; slot[4 * v + 3] = x.w;
            }
            k9::Result r;
            k9::fast_string<PARSE, true>(slot, lo, hi, lut, hash != nullptr, san ? san + o0 : nullptr, r);
            store_result(r, s, verdict, hash, span, san_len);
        }
        __syncthreads();                                       // order / hist are reused by the next batch
    }
}

// Pass 2: general code-point path (k9::process_string) over the queued strings, read from global memory, one
// string per thread.  A CTA takes 256 consecutive queue entries and hands them to its threads sorted by length
// (counting sort on len / 8, as pass 1 does): the queue arrives in no particular order and the strings are
// 12 - 200 bytes long, so without the sort a warp runs as long as its longest string with a third of its lanes
// active (8.9 of 32 measured); adjacent threads now hold strings of similar length.
constexpr int kGeneralPer = 2;                                  // strings per thread and batch (64-thread CTAs with one string
                                                                // per thread: 0.93 instead of 0.85 ms per 10 M strings)
template <bool PARSE>
__global__ void __launch_bounds__(kGeneralThreads)
k9_validate_hash_general(const uint8_t *__restrict__ data, const int64_t *__restrict__ off,
                         uint8_t *__restrict__ verdict, int64_t *__restrict__ hash, int32_t *__restrict__ span,
                         uint8_t *__restrict__ san, int32_t *__restrict__ san_len,
                         const uint32_t *__restrict__ queue) {
    constexpr int kBatch = kGeneralThreads * kGeneralPer;
    __shared__ int32_t hist[32];
    __shared__ uint32_t order[kBatch];
    __shared__ uint8_t lut[256];
    const int tid = threadIdx.x;
    for (int i = tid; i < 256; i += kGeneralThreads) lut[i] = g_tables.flags[i];
    const uint32_t count = queue[0];
    for (uint32_t base = blockIdx.x * kBatch; base < count; base += gridDim.x * kBatch) {
        const int cnt = (int)min((uint32_t)kBatch, count - base);
        if (tid < 32) hist[tid] = 0;
        __syncthreads();
        uint32_t sidx[kGeneralPer];
        int bucket[kGeneralPer], rank[kGeneralPer];
#pragma unroll
        for (int u = 0; u < kGeneralPer; ++u) {
            const int q = u * kGeneralThreads + tid;
            sidx[u] = 0; bucket[u] = 0; rank[u] = 0;
            if (q < cnt) {
                sidx[u] = queue[2 + base + q];
                bucket[u] = (int)min((off[sidx[u] + 1] - off[sidx[u]]) >> 3, (int64_t)31);
                rank[u] = atomicAdd(&hist[bucket[u]], 1);
            }
        }
        __syncthreads();
        if (tid < 32) {
            const int c = hist[tid];
            int incl = c;
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (tid >= d) incl += t;
            }
            hist[tid] = incl - c;                              // exclusive prefix
        }
        __syncthreads();
#pragma unroll
        for (int u = 0; u < kGeneralPer; ++u)
            if (u * kGeneralThreads + tid < cnt) order[hist[bucket[u]] + rank[u]] = sidx[u];
        __syncthreads();
#pragma unroll 1
        for (int u = 0; u < kGeneralPer; ++u) {
            const int q = u * kGeneralThreads + tid;
            if (q >= cnt) continue;
            const int64_t s = order[q];
            const int64_t o0 = off[s], o1 = off[s + 1];
            const uintptr_t sa = (uintptr_t)(data + o0);
            const uint32_t *w = (const uint32_t *)(sa & ~(uintptr_t)15);
            const int lo = (int)(sa & 15), hi = lo + (int)(o1 - o0);
            const k9::Result r = k9::process_string<PARSE>(w, lo, hi, lut, hash != nullptr, san ? san + o0 : nullptr);
            store_result(r, s, verdict, hash, span, san_len);
        }
        __syncthreads();                                       // order / hist are reused by the next batch
    }
}

constexpr int kSmemBytes = kStageBytes + kStageSlack + (int)sizeof(Shared);
constexpr int64_t kMaxStringsPerLaunch = 1ll << 31;   // queue entries are uint32 string indices

// All passes for n strings whose offsets start at `off`; `queue` holds 2 + n words.
int launch_passes(const uint8_t *d_data, const int64_t *d_off, int64_t n, uint32_t flags, uint8_t *d_verdict,
                  int64_t *d_hash, int32_t *d_span, uint8_t *d_san, int32_t *d_san_len, uint32_t *queue,
                  cudaStream_t st) {
    KX_CUDA(cudaMemsetAsync(queue, 0, 8, st));
    const unsigned blocks = (unsigned)((n + kThreads - 1) / kThreads);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t gbatch = kGeneralThreads * kGeneralPer;
    const unsigned gblocks = (unsigned)min((int64_t)sms * 32, (n + gbatch - 1) / gbatch);
    const unsigned lblocks = (unsigned)min((int64_t)sms * 16, (n + kLoopThreads - 1) / kLoopThreads);
    if (flags & KX_FLAG_PARSE) {
        k9_validate_hash_fast<true><<<blocks, kThreads, kSmemBytes, st>>>(d_data, d_off, n, d_verdict, d_hash, d_span, d_san, d_san_len, queue);
        k9_validate_hash_loops<true><<<lblocks, kLoopThreads, 0, st>>>(d_data, d_off, n, d_verdict, d_hash, d_span, d_san, d_san_len, queue);
        k9_validate_hash_general<true><<<gblocks, kGeneralThreads, 0, st>>>(d_data, d_off, d_verdict, d_hash, d_span, d_san, d_san_len, queue);
    } else {
        k9_validate_hash_fast<false><<<blocks, kThreads, kSmemBytes, st>>>(d_data, d_off, n, d_verdict, d_hash, d_span, d_san, d_san_len, queue);
        k9_validate_hash_loops<false><<<lblocks, kLoopThreads, 0, st>>>(d_data, d_off, n, d_verdict, d_hash, d_span, d_san, d_san_len, queue);
        k9_validate_hash_general<false><<<gblocks, kGeneralThreads, 0, st>>>(d_data, d_off, d_verdict, d_hash, d_span, d_san, d_san_len, queue);
    }
    kx::count_launch(3);
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}

}  // namespace

extern "C" int kx_validate_hash(const uint8_t *d_data, const int64_t *d_off, int64_t n, uint32_t flags,
                                uint8_t *d_verdict, int64_t *d_hash, int32_t *d_span,
                                uint8_t *d_san, int32_t *d_san_len, void *stream) {
    KX_REQUIRE(n >= 0, KX_EINVAL, "kx_validate_hash: n = %lld is negative", (long long)n);
    KX_REQUIRE((flags & ~KX_FLAG_PARSE) == 0, KX_EINVAL, "kx_validate_hash: unknown flags 0x%x", flags);
    if (n == 0) return KX_OK;
    KX_REQUIRE(d_off != nullptr, KX_EINVAL, "kx_validate_hash: d_off is NULL");
    cudaStream_t st = (cudaStream_t)stream;
    // Scratch: the queue of strings handed from pass 1 to pass 2 (4 bytes per string, worst case
    // every string).  Owned by the library, cached per (device, stream), grown on demand.
    for (int64_t r0 = 0; r0 < n; r0 += kMaxStringsPerLaunch) {
        const int64_t rows = min(n - r0, kMaxStringsPerLaunch);
        uint32_t *queue = nullptr;
        int rc = kx::stream_scratch(st, kx::SCRATCH_K9_QUEUE, sizeof(uint32_t) * (size_t)(rows + 2), (void **)&queue);
        if (rc != KX_OK) return rc;
        rc = launch_passes(d_data, d_off + r0, rows, flags, d_verdict ? d_verdict + r0 : nullptr,
                           d_hash ? d_hash + r0 : nullptr, d_span ? d_span + 2 * r0 : nullptr, d_san,
                           d_san_len ? d_san_len + r0 : nullptr, queue, st);
        if (rc != KX_OK) return rc;
    }
    return KX_OK;
}

// ---------------------------------------------------------------------------
// kx_validate_hash_noalloc: the same passes, with the hand-over queue in a
// fixed array that is part of the loaded module instead of cudaMalloc'ed scratch.
//
// Who needs this: a host process that cannot (or must not) allocate device memory
// at call time.  Two cases we know of: a serving process whose framework allocator
// has already reserved all of HBM for weights and KV cache, where a cudaMalloc from
// a plugin simply fails; and a request loop with a latency budget, where
// cudaMalloc's device-wide synchronisation would stall every other stream on the
// GPU.  Such a caller passes inputs and outputs as usual and this entry point
// touches no allocator at all (no cudaMalloc / cudaFree, no library pool).
//
// The price is a fixed capacity (KX_NOALLOC_MAX_STRINGS strings per call, about
// 4 MB of device memory that every GPU loading the library holds until exit) and a
// single shared array: two calls that overlap on one GPU (different streams or
// threads) would race on it, so callers must serialise them.  The array keeps the
// previous call's contents; the only part that has to start from zero, the queue's
// element counter, is cleared on the stream at the start of every call.
//
// In host code `g_noalloc_queue` names a host-side shadow of the symbol, not device
// memory, so the device address is looked up once per device with
// cudaGetSymbolAddress (with that device current) and that pointer is what the
// kernels receive.  The lookup happens in the first call on each device; make that
// first call before any stream capture begins.
// ---------------------------------------------------------------------------
static __device__ __align__(256) uint32_t g_noalloc_queue[2 + KX_NOALLOC_MAX_STRINGS];

namespace {
constexpr int kMaxDevices = 64;
std::atomic<uint32_t *> g_noalloc_ptr[kMaxDevices];
}  // namespace

extern "C" int kx_validate_hash_noalloc(const uint8_t *d_data, const int64_t *d_off, int64_t n, uint32_t flags,
                                        uint8_t *d_verdict, int64_t *d_hash, int32_t *d_span,
                                        uint8_t *d_san, int32_t *d_san_len, void *stream) {
    KX_REQUIRE(n >= 0, KX_EINVAL, "kx_validate_hash_noalloc: n = %lld is negative", (long long)n);
    KX_REQUIRE((flags & ~KX_FLAG_PARSE) == 0, KX_EINVAL, "kx_validate_hash_noalloc: unknown flags 0x%x", flags);
    KX_REQUIRE(n <= KX_NOALLOC_MAX_STRINGS, KX_ETOOBIG,
               "kx_validate_hash_noalloc: n = %lld exceeds the %d strings its fixed scratch holds; "
               "split the batch or call kx_validate_hash", (long long)n, KX_NOALLOC_MAX_STRINGS);
    if (n == 0) return KX_OK;
    KX_REQUIRE(d_off != nullptr, KX_EINVAL, "kx_validate_hash_noalloc: d_off is NULL");
    int dev = 0;
    KX_CUDA(cudaGetDevice(&dev));
    KX_REQUIRE(dev >= 0 && dev < kMaxDevices, KX_EINVAL, "kx_validate_hash_noalloc: device ordinal %d not supported", dev);
    uint32_t *queue = g_noalloc_ptr[dev].load(std::memory_order_acquire);
    if (queue == nullptr) {
        // once per device, with that device current: the symbol's address in device memory
        KX_CUDA(cudaGetSymbolAddress((void **)&queue, g_noalloc_queue));
        g_noalloc_ptr[dev].store(queue, std::memory_order_release);
    }
    return launch_passes(d_data, d_off, n, flags, d_verdict, d_hash, d_span, d_san, d_san_len, queue,
                         (cudaStream_t)stream);
}
