// k9 core: per-string logic of the fused validator + parser + sanitiser + hash.
// Written once for device and host so that tests/emul can run the exact kernel
// logic on a CPU (test infrastructure; the product only ever calls it from the
// CUDA kernel in kx_validate_hash.cu).  See that file for the references.
#pragma once
#include <stdint.h>
#include "../../include/kx.h"

#ifdef __CUDACC__
#define KX_HD __host__ __device__ __forceinline__
#else
#define KX_HD inline
#endif

namespace k9 {

KX_HD uint32_t kx_byte(uint32_t word, int k) {
#ifdef __CUDA_ARCH__
    return __byte_perm(word, 0u, 0x4440u + k);
#else
    return (word >> (8 * k)) & 0xFFu;
#endif
}

KX_HD bool ws_ascii(uint32_t c) { return (c - 0x09u) <= 4u || (c - 0x1Cu) <= 4u; }

KX_HD bool ws_wide(uint32_t c) {  // c >= 0x80, str.isspace()
    return c == 0x85u || c == 0xA0u || c == 0x1680u || (c - 0x2000u) <= 0xAu || c == 0x2028u ||
           c == 0x2029u || c == 0x202Fu || c == 0x205Fu || c == 0x3000u;
}

KX_HD uint64_t rotl64(uint64_t x, int b) { return (x << b) | (x >> (64 - b)); }

struct Sip {  // CPython pyhash.c siphash13, k0 = k1 = 0
    uint64_t v0, v1, v2, v3, m;
    uint32_t sh, nbytes;
    KX_HD void init() {
        v0 = 0x736f6d6570736575ULL; v1 = 0x646f72616e646f6dULL;
        v2 = 0x6c7967656e657261ULL; v3 = 0x7465646279746573ULL;
        m = 0; sh = 0; nbytes = 0;
    }
    KX_HD void round() {
        v0 += v1; v2 += v3; v1 = rotl64(v1, 13) ^ v0; v3 = rotl64(v3, 16) ^ v2; v0 = rotl64(v0, 32);
        v2 += v1; v0 += v3; v1 = rotl64(v1, 17) ^ v2; v3 = rotl64(v3, 21) ^ v0; v2 = rotl64(v2, 32);
    }
    KX_HD void emit(uint32_t c, uint32_t wbits) {
        m |= (uint64_t)c << sh;
        sh += wbits;
        nbytes += wbits >> 3;
        if (sh == 64) { v3 ^= m; round(); v0 ^= m; m = 0; sh = 0; }
    }
    KX_HD int64_t finish() {
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
    KX_HD void init(const uint32_t *w_, int lo, int hi_) {
        w = w_; p = lo; hi = hi_;
        cur = lo < hi_ ? w[lo >> 2] : 0u;
    }
    KX_HD uint32_t next() {
        uint32_t b = kx_byte(cur, p & 3);
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
KX_HD constexpr uint64_t dfa_table() {
    const int t[5][4] = {{0, 3, 1, 2}, {1, 1, 0, 1}, {2, 4, 2, 0}, {0, 0, 0, 0}, {2, 2, 2, 2}};
    uint64_t r = 0;
    for (int s = 0; s < 5; ++s)
        for (int c = 0; c < 4; ++c) r |= (uint64_t)t[s][c] << (3 * (s * 4 + c));
    return r;
}

// One forward pass over bytes [lo, hi) of w.  HASH: feed the sanitised code points
// to `sip` (and optionally write the sanitised bytes); TICKS: count ``` fences.
template <bool HASH, bool TICKS>
KX_HD void walk(const uint32_t *w, int lo, int hi, uint32_t wbits, Sip &sip,
                                     uint8_t *san, Walk &o) {
    constexpr uint64_t T = dfa_table();
    constexpr uint64_t kPrefix = 0x206c74636562756bULL;  // "kubectl " little-endian
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
KX_HD uint32_t hash_width_bits(const uint32_t *w, int lo, int hi) {
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
        if (c >= 0x100u && !ws_wide(c)) maxc = c > maxc ? c : maxc;
    }
    return maxc < 0x100u ? 8u : maxc < 0x10000u ? 16u : 32u;
}


struct Result {
    int64_t hash;
    uint32_t verdict;
    int span_a, span_b, san_len;
};

// Everything k9 computes for the string occupying bytes [lo, hi) of the word
// buffer w.  `san` (may be NULL) receives the sanitised bytes.
template <bool PARSE>
KX_HD Result process_string(const uint32_t *w, int lo, int hi, bool want_hash, uint8_t *san) {
    const uint32_t wbits = want_hash ? hash_width_bits(w, lo, hi) : 8u;
    Sip sip;
    sip.init();
    Walk o;
    walk<true, PARSE>(w, lo, hi, wbits, sip, san, o);
    Result r;
    r.hash = (want_hash && o.any) ? sip.finish() : 0;
    r.san_len = o.san_len;
    if (PARSE && o.lead_ticks >= 3 && o.trail_ticks >= 3) {   // app.py:99-100
        const int a2 = o.a + 3, b2 = o.b - 3 > a2 ? o.b - 3 : a2;
        walk<false, false>(w, a2, b2, 8u, sip, nullptr, o);
    }
    r.verdict = o.verdict;
    r.span_a = o.a - lo;
    r.span_b = o.b - lo;
    return r;
}


// --------------------------------------------------------------------------- //
// Fast path: ASCII strings staged in (mutable) shared memory.
//
// The string is stripped by scanning ASCII whitespace inwards from both ends, the
// "kubectl " prefix is compared on the raw bytes, and then ONE table-driven pass
// over the stripped bytes advances a combined automaton
//     state = quoting state (A,SQ,DQ,EA,ED) x previous byte class (other,&,|)
// whose table entries also carry the sticky METACHAR / NON-ASCII flags, while the
// sanitised form (whitespace runs -> one 0x20) is compacted IN PLACE at the front
// of the string's own bytes (stores only happen once the output position differs
// from the input position, i.e. never for an already-clean string).  SipHash-1-3
// then runs over that compacted byte stream eight bytes per step, which keeps the
// lanes of a warp in lockstep.  A string with any byte >= 0x80 (or, in PARSE mode,
// starting with a back-tick) is reported back for the general code-point path.
// --------------------------------------------------------------------------- //
enum : uint32_t { F_STATE = 15u, F_META = 16u, F_NONASCII = 32u };

struct FastTables {
    uint8_t cls[256];     // byte -> class: 0 other 1 space 2 other-ws 3 meta 4 & 5 | 6 backslash 7 ' 8 " 9 >=0x80
    uint8_t trans[256];   // (state << 4 | class) -> next state | flags
};

KX_HD constexpr FastTables make_fast_tables() {
    FastTables t{};
    for (int b = 0; b < 256; ++b) {
        uint8_t c = 0;
        if (b >= 0x80) c = 9;
        else if (b == 0x20) c = 1;
        else if ((b >= 0x09 && b <= 0x0D) || (b >= 0x1C && b <= 0x1F)) c = 2;
        else if (b == ';' || b == '`' || b == '$' || b == '(' || b == ')' || b == '<' || b == '>') c = 3;
        else if (b == '&') c = 4;
        else if (b == '|') c = 5;
        else if (b == '\\') c = 6;
        else if (b == '\'') c = 7;
        else if (b == '"') c = 8;
        t.cls[b] = c;
    }
    const int dfa[5][4] = {{0, 3, 1, 2}, {1, 1, 0, 1}, {2, 4, 2, 0}, {0, 0, 0, 0}, {2, 2, 2, 2}};
    for (int st = 0; st < 16; ++st) {
        for (int c = 0; c < 16; ++c) {
            uint8_t e = 0;
            if (st < 15 && c < 10) {
                const int q = st / 3, prev = st % 3;
                const int qc = c == 6 ? 1 : c == 7 ? 2 : c == 8 ? 3 : 0;
                const int nq = dfa[q][qc];
                const int np = c == 4 ? 1 : c == 5 ? 2 : 0;
                e = (uint8_t)(nq * 3 + np);
                if (c == 3 || (c == 4 && prev == 1) || (c == 5 && prev == 2)) e |= F_META;
                if (c == 9) e |= F_NONASCII;
            }
            t.trans[st * 16 + c] = e;
        }
    }
    return t;
}

KX_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t shift_bits) {   // shift_bits in {0,8,16,24}
#ifdef __CUDA_ARCH__
    return __funnelshift_r(lo, hi, shift_bits);
#else
    return (uint32_t)((((uint64_t)hi << 32) | lo) >> shift_bits);
#endif
}

// SipHash-1-3 (zero key) of the n bytes starting at byte offset `lo` of word buffer W.
KX_HD int64_t sip_stream(const uint32_t *W, int lo, int n) {
    if (n == 0) return 0;
    Sip s;
    s.init();
    const uint32_t sh = (uint32_t)(lo & 3) * 8u;
    int wi = lo >> 2;
    uint32_t w0 = W[wi];
    int left = n;
    for (; left >= 8; left -= 8, wi += 2) {
        const uint32_t w1 = W[wi + 1], w2 = W[wi + 2];
        const uint64_t m = ((uint64_t)funnel_r(w1, w2, sh) << 32) | funnel_r(w0, w1, sh);
        s.v3 ^= m; s.round(); s.v0 ^= m;
        w0 = w2;
    }
    uint64_t m = 0;
    if (left > 0) {
        const uint32_t w1 = W[wi + 1];
        const uint32_t w2 = left + (int)(sh >> 3) > 8 ? W[wi + 2] : 0u;
        m = ((uint64_t)funnel_r(w1, w2, sh) << 32) | funnel_r(w0, w1, sh);
        m &= ~0ULL >> (64 - 8 * left);
    }
    s.m = m;
    s.nbytes = (uint32_t)n;
    return s.finish();
}

// Length (2 or 3) of the UTF-8 encoding of a non-ASCII whitespace code point that starts at B[p], else 0.
// Exactly the 19 code points of ws_wide: U+0085 U+00A0 | U+1680 | U+2000..200A U+2028 U+2029 U+202F | U+205F | U+3000.
KX_HD int uws_at(const uint8_t *B, int p, int hi) {
    const uint32_t b0 = B[p];
    if (b0 == 0xC2u) return (p + 1 < hi && (B[p + 1] == 0x85u || B[p + 1] == 0xA0u)) ? 2 : 0;
    if (b0 < 0xE1u || b0 > 0xE3u || p + 2 >= hi) return 0;
    const uint32_t b1 = B[p + 1], b2 = B[p + 2];
    if (b0 == 0xE1u) return (b1 == 0x9Au && b2 == 0x80u) ? 3 : 0;
    if (b0 == 0xE3u) return (b1 == 0x80u && b2 == 0x80u) ? 3 : 0;
    if (b1 == 0x80u) return ((b2 >= 0x80u && b2 <= 0x8Au) || b2 == 0xA8u || b2 == 0xA9u || b2 == 0xAFu) ? 3 : 0;
    return (b1 == 0x81u && b2 == 0x9Fu) ? 3 : 0;
}

// Returns true if the string must be redone by the general path; otherwise fills r.
//
// Non-ASCII whitespace at the two ENDS of the string (str.strip() removes it, app.py:74,97) is stripped here
// as well, so a padded but otherwise ASCII string stays on this path.  Front: exact encodings, decoded
// forwards from the string start.  Back: the same encodings matched backwards; that agrees with the general
// path's forward decoding because what is left in between is pure ASCII (checked by the main loop: any byte
// >= 0x80 there sends the string to the general path), so the trailing run starts on a code point boundary
// and each matched unit is one complete encoding.
template <bool PARSE>
KX_HD bool fast_string(uint32_t *W, int lo, int hi, const uint8_t *cls_lut, const uint8_t *trans,
                       bool want_hash, uint8_t *san, Result &r) {
    uint8_t *B = (uint8_t *)W;
    int a = lo, b = hi;
    for (;;) {
        while (a < hi && ws_ascii(B[a])) ++a;
        if (a >= hi || B[a] < 0x80u) break;
        const int n = uws_at(B, a, hi);
        if (n == 0) break;
        a += n;
    }
    for (;;) {
        while (b > a && ws_ascii(B[b - 1])) --b;
        if (b <= a || B[b - 1] < 0x80u) break;
        int n = 0;
        if (b - 2 >= a && uws_at(B, b - 2, b) == 2) n = 2;
        else if (b - 3 >= a && uws_at(B, b - 3, b) == 3) n = 3;
        if (n == 0) break;
        b -= n;
    }
    if (PARSE && a < b && B[a] == 0x60) return true;

    bool pfx = false;
    if (b - a >= 8) {
        const uint32_t sh = (uint32_t)(a & 3) * 8u;
        const int wi = a >> 2;
        const uint32_t w0 = W[wi], w1 = W[wi + 1], w2 = W[wi + 2];
        pfx = funnel_r(w0, w1, sh) == 0x6562756bu && funnel_r(w1, w2, sh) == 0x206c7463u;   // "kube" "ctl "
    }

    int o = lo;
    uint32_t sf = 0;
    bool prev_nonws = false;
    uint32_t cur = a < b ? W[a >> 2] : 0u;
    for (int p = a; p < b; ++p) {
        const uint32_t x = kx_byte(cur, p & 3);
        if ((p & 3) == 3 && p + 1 < b) cur = W[(p + 1) >> 2];
        const uint32_t c = cls_lut[x];
        sf = trans[((sf & F_STATE) << 4) | c] | (sf & (F_META | F_NONASCII));
        const bool ws = (c - 1u) <= 1u;
        const bool emit = !ws || prev_nonws;
        const uint32_t val = ws ? 0x20u : x;
        if (emit && (o != p || val != x)) B[o] = (uint8_t)val;
        o += emit ? 1 : 0;
        prev_nonws = !ws;
    }
    if (sf & F_NONASCII) return true;

    r.verdict = !pfx ? KX_BAD_PREFIX : (sf & F_META) ? KX_METACHAR : (sf & F_STATE) >= 3u ? KX_BAD_QUOTING : KX_SAFE;
    r.span_a = a - lo;
    r.span_b = b - lo;
    r.san_len = o - lo;
    r.hash = want_hash ? sip_stream(W, lo, o - lo) : 0;
    if (san) for (int i = 0; i < o - lo; ++i) san[i] = B[lo + i];
    return false;
}

}  // namespace k9
