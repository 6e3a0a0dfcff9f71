// k9: fused validator + parser + sanitiser + cache-key hash over a CSR batch of
// strings.  Replaces, per string,
//   is_safe_kubectl_command   /root/reference/app.py:72-88   (shlex.split: CPython Lib/shlex.py:133-277)
//   KubectlOutputParser.parse /root/reference/app.py:95-104
//   sanitize_query            /root/reference/app.py:60-68
//   hash(sanitised)           CPython Python/pyhash.c siphash13, zero key; reached at app.py:312,322
//
// Layout: one CTA of 256 threads owns 256 consecutive strings.  The CTA's byte
// range is contiguous in the CSR buffer, so it is staged into shared memory with
// coalesced 16-byte loads (each HBM byte is read exactly once); then every thread
// walks its own string out of shared memory, one code point at a time, carrying
//   - the validator state (prefix match, metacharacter flag, 5-state quoting DFA),
//     snapshotted at the last non-whitespace character so that strip() needs no
//     backward scan,
//   - the sanitiser (token/space state) feeding SipHash-1-3 eight bytes at a time
//     in the buffer width (1/2/4 bytes per code point) CPython would have chosen.
// A CTA whose byte range exceeds the staging buffer reads global memory directly.
#include "kx_common.cuh"
#include "k9_core.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kStageBytes = 32 * 1024;

template <bool PARSE>
__global__ void __launch_bounds__(kThreads)
k9_validate_hash(const uint8_t *__restrict__ data, const int64_t *__restrict__ off, int64_t n,
                 uint8_t *__restrict__ verdict, int64_t *__restrict__ hash, int32_t *__restrict__ span,
                 uint8_t *__restrict__ san, int32_t *__restrict__ san_len) {
    extern __shared__ uint4 stage[];
    const int64_t s0 = (int64_t)blockIdx.x * kThreads;
    const int64_t s1 = min(n, s0 + kThreads);
    const int64_t g_beg = off[s0], g_end = off[s1];
    const uintptr_t addr = (uintptr_t)(data + g_beg);
    const uint4 *abase = (const uint4 *)(addr & ~(uintptr_t)15);
    const int lead = (int)(addr & 15);
    const int64_t nbytes = lead + (g_end - g_beg);
    const bool staged = nbytes <= kStageBytes;   // CTA-uniform
    if (staged) {
        const int nvec = (int)((nbytes + 15) >> 4);
        for (int v = threadIdx.x; v < nvec; v += kThreads) stage[v] = __ldg(abase + v);
        __syncthreads();
    }
    const int64_t s = s0 + threadIdx.x;
    if (s >= s1) return;
    const int64_t o0 = off[s], o1 = off[s + 1];
    const uint32_t *w;
    int lo, hi;
    if (staged) {
        w = (const uint32_t *)stage;
        lo = (int)(o0 - g_beg) + lead;
        hi = (int)(o1 - g_beg) + lead;
    } else {
        const uintptr_t sa = (uintptr_t)(data + o0);
        w = (const uint32_t *)(sa & ~(uintptr_t)15);
        lo = (int)(sa & 15);
        hi = lo + (int)(o1 - o0);
    }

    const k9::Result r = k9::process_string<PARSE>(w, lo, hi, hash != nullptr, san ? san + o0 : nullptr);
    if (hash) hash[s] = r.hash;
    if (san_len) san_len[s] = r.san_len;
    if (verdict) verdict[s] = (uint8_t)r.verdict;
    if (span) { span[2 * s] = r.span_a; span[2 * s + 1] = r.span_b; }
}

}  // namespace

extern "C" int kx_validate_hash(const uint8_t *d_data, const int64_t *d_off, int64_t n, uint32_t flags,
                                uint8_t *d_verdict, int64_t *d_hash, int32_t *d_span,
                                uint8_t *d_san, int32_t *d_san_len, void *stream) {
    KX_REQUIRE(n >= 0, KX_EINVAL, "kx_validate_hash: n = %lld is negative", (long long)n);
    KX_REQUIRE((flags & ~KX_FLAG_PARSE) == 0, KX_EINVAL, "kx_validate_hash: unknown flags 0x%x", flags);
    if (n == 0) return KX_OK;
    KX_REQUIRE(d_off != nullptr, KX_EINVAL, "kx_validate_hash: d_off is NULL");
    const int64_t blocks = (n + kThreads - 1) / kThreads;
    KX_REQUIRE(blocks <= 0x7fffffff, KX_ETOOBIG, "kx_validate_hash: n = %lld exceeds 2^31 * 256 strings", (long long)n);
    cudaStream_t st = (cudaStream_t)stream;
    if (flags & KX_FLAG_PARSE)
        k9_validate_hash<true><<<(unsigned)blocks, kThreads, kStageBytes, st>>>(d_data, d_off, n, d_verdict, d_hash, d_span, d_san, d_san_len);
    else
        k9_validate_hash<false><<<(unsigned)blocks, kThreads, kStageBytes, st>>>(d_data, d_off, n, d_verdict, d_hash, d_span, d_san, d_san_len);
    kx::count_launch();
    KX_CUDA(cudaGetLastError());
    return KX_OK;
}
