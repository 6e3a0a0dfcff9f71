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

namespace {

constexpr int kThreads = 256;
constexpr int kStageBytes = 32 * 1024;

__device__ __forceinline__ bool ws_ascii(uint32_t c) { return (c - 0x09u) <= 4u || (c - 0x1Cu) <= 4u; }

__device__ __forceinline__ bool ws_wide(uint32_t c) {  // c >= 0x80, str.isspace()
    return c == 0x85u || c == 0xA0u || c == 0x1680u || (c - 0x2000u) <= 0xAu || c == 0x2028u ||
           c == 0x2029u || c == 0x202Fu || c == 0x205Fu || c == 0x3000u;
}

__device__ __forceinline__ uint64_t rotl64(uint64_t x, int b) { return (x << b) | (x >> (64 - b)); }

struct Sip {  // CPython pyhash.c siphash13, k0 = k1 = 0
    uint64_t v0, v1, v2, v3, m;
    uint32_t sh, nbytes;
    __device__ __forceinline__ void init() {
        v0 = 0x736f6d6570736575ULL; v1 = 0x646f72616e646f6dULL;
        v2 = 0x6c7967656e657261ULL; v3 = 0x7465646279746573ULL;
        m = 0; sh = 0; nbytes = 0;
    }
    __device__ __forceinline__ void round() {
        v0 += v1; v2 += v3; v1 = rotl64(v1, 13) ^ v0; v3 = rotl64(v3, 16) ^ v2; v0 = rotl64(v0, 32);
        v2 += v1; v0 += v3; v1 = rotl64(v1, 17) ^ v2; v3 = rotl64(v3, 21) ^ v0; v2 = rotl64(v2, 32);
    }
    __device__ __forceinline__ void emit(uint32_t c, uint32_t wbits) {
        m |= (uint64_t)c << sh;
        sh += wbits;
        nbytes += wbits >> 3;
        if (sh == 64) { v3 ^= m; round(); v0 ^= m; m = 0; sh = 0; }
    }
    __device__ __forceinline__ int64_t finish() {
        uint64_t b = ((uint64_t)nbytes << 56) | m;
        v3 ^= b; round(); v0 ^= b;
        v2 ^= 0xff; round(); round(); round();
        int64_t h = (int64_t)((v0 ^ v1) ^ (v2 ^ v3));
        return h == -1 ? -2 : h;
    }
};

// Byte cursor over a word-aligned buffer (shared or global) with a one-word cache.
struct Cursor {
    const uint32_t *w;
    int p, hi;
    uint32_t cur;
    __device__ __forceinline__ void init(const uint32_t *w_, int lo, int hi_) {
        w = w_; p = lo; hi = hi_;
        cur = lo < hi_ ? w[lo >> 2] : 0u;
    }
    __device__ __forceinline__ uint32_t next() {
        uint32_t b = __byte_perm(cur, 0u, 0x4440u + (p & 3));
        ++p;
        if ((p & 3) == 0 && p < hi) cur = w[p >> 2];
        return b;
    }
};

struct Walk {
    int a, b;               // byte range of the stripped string
    uint32_t verdict;
    int lead_ticks, trail_ticks;
    int san_len;
    bool any;               // at least one non-whitespace character
};

// quoting DFA (shlex.py:149-267 reduced): states A=0 SQ=1 DQ=2 EA=3 ED=4,
// classes other=0 backslash=1 squote=2 dquote=3; 3 bits per entry.
__device__ __forceinline__ constexpr uint64_t dfa_table() {
    const int t[5][4] = {{0, 3, 1, 2}, {1, 1, 0, 1}, {2, 4, 2, 0}, {0, 0, 0, 0}, {2, 2, 2, 2}};
    uint64_t r = 0;
    for (int s = 0; s < 5; ++s)
        for (int c = 0; c < 4; ++c) r |= (uint64_t)t[s][c] << (3 * (s * 4 + c));
    return r;
}

// One forward pass over bytes [lo, hi) of w.  HASH: feed the sanitised code points
// to `sip` (and optionally write the sanitised bytes); TICKS: count ``` fences.
template <bool HASH, bool TICKS>
__device__ __forceinline__ void walk(const uint32_t *w, int lo, int hi, uint32_t wbits, Sip &sip,
                                     uint8_t *san, Walk &o) {
    constexpr uint64_t T = dfa_table();
    constexpr uint64_t kPrefix = 0x206c746365627562ULL;  // "kubectl " little-endian
    Cursor cu;
    cu.init(w, lo, hi);
    int a = -1, b = hi;
    uint32_t st = 0, snap_st = 0, prev = 0;
    bool meta = false, snap_meta = false, pfx_ok = true, in_tok = false, emitted = false, all_ticks = true;
    int lead = 0, trail = 0, snap_trail = 0, sn = 0;

    while (cu.p < hi) {
        const int cs = cu.p;
        const uint32_t b0 = cu.next();
        uint32_t c = b0, raw = b0;
        int nact = 1;
        bool ws;
        if (b0 < 0x80u) {
            ws = ws_ascii(b0);
        } else {
            if (b0 >= 0xC0u && b0 < 0xF8u) {
                const int n = b0 < 0xE0u ? 2 : b0 < 0xF0u ? 3 : 4;
                c = b0 & (n == 2 ? 0x1Fu : n == 3 ? 0x0Fu : 0x07u);
                for (int j = 1; j < n; ++j) {
                    uint32_t bj = 0;
                    if (cu.p < hi) { bj = cu.next(); raw |= bj << (8 * j); ++nact; }
                    c = (c << 6) | (bj & 0x3Fu);
                }
            }
            ws = c >= 0x80u ? ws_wide(c) : ws_ascii(c);
        }
        if (!ws && a < 0) a = cs;
        if (a >= 0) {
            // validator, byte level (app.py:75-87)
            for (int j = 0; j < nact; ++j) {
                const uint32_t x = (raw >> (8 * j)) & 0xFFu;
                const int rel = cs + j - a;
                if (rel < 8) pfx_ok = pfx_ok && (x == (uint32_t)((kPrefix >> (8 * rel)) & 0xFFu));
                const bool single = x == ';' || x == '`' || x == '$' || x == '(' || x == ')' || x == '<' || x == '>';
                const bool dbl = (x == '&' || x == '|') && prev == x;
                meta = meta || single || dbl;
                prev = x;
                const uint32_t cls = x == '\\' ? 1u : x == '\'' ? 2u : x == '"' ? 3u : 0u;
                st = (uint32_t)(T >> (3u * (st * 4u + cls))) & 7u;
            }
            if (TICKS) {
                const bool tick = nact == 1 && b0 == 0x60u;
                if (tick) { ++trail; if (all_ticks) ++lead; } else { trail = 0; all_ticks = false; }
            }
        }
        if (!ws) {
            b = cu.p; snap_st = st; snap_meta = meta; snap_trail = trail;
            if (HASH) {
                if (!in_tok && emitted) { sip.emit(0x20u, wbits); if (san) san[sn] = 0x20; ++sn; }
                in_tok = true; emitted = true;
                sip.emit(c, wbits);
                if (san) for (int j = 0; j < nact; ++j) san[sn + j] = (uint8_t)(raw >> (8 * j));
                sn += nact;
            }
        } else if (HASH) {
            in_tok = false;
        }
    }
    if (a < 0) { a = hi; b = hi; }
    o.a = a; o.b = b;
    o.verdict = !(pfx_ok && b - a >= 8) ? KX_BAD_PREFIX : snap_meta ? KX_METACHAR : snap_st != 0 ? KX_BAD_QUOTING : KX_SAFE;
    o.lead_ticks = lead; o.trail_ticks = snap_trail;
    o.san_len = sn; o.any = emitted;
}

// Width in bits of one code point in the compact buffer CPython would allocate for
// the sanitised string: decided by the widest NON-whitespace code point.
__device__ __forceinline__ uint32_t hash_width_bits(const uint32_t *w, int lo, int hi) {
    if (lo >= hi) return 8;
    const int w0 = lo >> 2, w1 = (hi + 3) >> 2;
    uint32_t acc = 0;
    for (int i = w0; i < w1; ++i) {
        uint32_t x = w[i];
        if (i == w0) x &= 0xFFFFFFFFu << (8 * (lo & 3));
        if (i == w1 - 1 && (hi & 3)) x &= 0xFFFFFFFFu >> (8 * (4 - (hi & 3)));
        acc |= x;
    }
    if ((acc & 0x80808080u) == 0) return 8;   // pure ASCII
    Cursor cu;
    cu.init(w, lo, hi);
    uint32_t maxc = 0;
    while (cu.p < hi) {
        const uint32_t b0 = cu.next();
        uint32_t c = b0;
        if (b0 >= 0xC0u && b0 < 0xF8u) {
            const int n = b0 < 0xE0u ? 2 : b0 < 0xF0u ? 3 : 4;
            c = b0 & (n == 2 ? 0x1Fu : n == 3 ? 0x0Fu : 0x07u);
            for (int j = 1; j < n; ++j) c = (c << 6) | ((cu.p < hi ? cu.next() : 0u) & 0x3Fu);
        }
        if (c >= 0x100u && !ws_wide(c)) maxc = max(maxc, c);
    }
    return maxc < 0x100u ? 8u : maxc < 0x10000u ? 16u : 32u;
}

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

    const bool want_hash = hash != nullptr;
    const uint32_t wbits = want_hash ? hash_width_bits(w, lo, hi) : 8u;
    Sip sip;
    sip.init();
    Walk o;
    walk<true, PARSE>(w, lo, hi, wbits, sip, san ? san + o0 : nullptr, o);
    if (want_hash) hash[s] = o.any ? sip.finish() : 0;
    if (san_len) san_len[s] = o.san_len;
    if (PARSE && o.lead_ticks >= 3 && o.trail_ticks >= 3) {   // app.py:99-100
        const int a2 = o.a + 3, b2 = max(o.a + 3, o.b - 3);
        walk<false, false>(w, a2, b2, 8u, sip, nullptr, o);
    }
    if (verdict) verdict[s] = (uint8_t)o.verdict;
    if (span) { span[2 * s] = o.a - lo; span[2 * s + 1] = o.b - lo; }
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
