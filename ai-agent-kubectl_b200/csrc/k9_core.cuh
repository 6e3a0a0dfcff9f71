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

KX_HD uint32_t pack4(uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3) {   // four values < 256 into one word
#ifdef __CUDA_ARCH__
    return __byte_perm(__byte_perm(b0, b1, 0x0040), __byte_perm(b2, b3, 0x0040), 0x5410);
#else
    return b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
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

enum : uint32_t { B_META = 1u, B_AMP = 2u, B_PIPE = 4u, B_SQ = 8u, B_DQ = 16u, B_BS = 32u, B_WS = 64u, B_CTRL = 128u };

struct FastTables {
    uint8_t flags[256];   // byte -> B_* flags (0 for every byte >= 0x80 and for ordinary characters)
};

KX_HD constexpr FastTables make_fast_tables() {
    FastTables t{};
    for (int b = 0; b < 128; ++b) {
        uint8_t f = 0;
        if (b == 0x20) f = B_WS;
        else if ((b >= 0x09 && b <= 0x0D) || (b >= 0x1C && b <= 0x1F)) f = B_WS | B_CTRL;
        else if (b == ';' || b == '`' || b == '$' || b == '(' || b == ')' || b == '<' || b == '>') f = B_META;
        else if (b == '&') f = B_AMP;
        else if (b == '|') f = B_PIPE;
        else if (b == '\\') f = B_BS;
        else if (b == '\'') f = B_SQ;
        else if (b == '"') f = B_DQ;
        t.flags[b] = f;
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
// The validator's rules (app.py:75-87) are byte rules; per byte they cost one table look-up (lut = FastTables::flags)
// and one step of the quoting automaton.  Bytes >= 0x80 are ordinary characters to it: one "other" step per code
// point is enough (after one such step the automaton is in A, SQ or DQ, which further ordinary bytes leave alone).
template <bool HASH, bool TICKS>
KX_HD void walk(const uint32_t *w, int lo, int hi, uint32_t wbits, const uint8_t *lut, Sip &sip,
                                     uint8_t *san, Walk &o) {
    constexpr uint64_t T = dfa_table();
    Cursor cu;
    cu.init(w, lo, hi);
    int a = -1, b = hi;
    uint32_t st = 0, snap_st = 0, prevf = 0;
    bool meta = false, snap_meta = false, in_tok = false, emitted = false, all_ticks = true;
    int lead = 0, trail = 0, snap_trail = 0, sn = 0;

    while (cu.p < hi) {
        const int cs = cu.p;
        const uint32_t b0 = cu.next();
        uint32_t c = b0, raw = b0, f = 0, fx = 0;              // fx: flags of ASCII bytes a (broken) sequence swallowed
        int nact = 1;
        bool ws;
        if (b0 < 0x80u) {
            f = lut[b0];
            ws = (f & B_WS) != 0;
        } else {
            if (b0 >= 0xC0u && b0 < 0xF8u) {
                const int n = b0 < 0xE0u ? 2 : b0 < 0xF0u ? 3 : 4;
                c = b0 & (n == 2 ? 0x1Fu : n == 3 ? 0x0Fu : 0x07u);
                for (int j = 1; j < n; ++j) {
                    uint32_t bj = 0;
                    if (cu.p < hi) { bj = cu.next(); raw |= bj << (8 * j); ++nact; fx |= bj < 0x80u ? 1u : 0u; }
                    c = (c << 6) | (bj & 0x3Fu);
                }
            }
            ws = c >= 0x80u ? ws_wide(c) : ws_ascii(c);
        }
        if (!ws && a < 0) a = cs;
        if (a >= 0) {
            auto vstep = [&](uint32_t fl) {
                meta = meta || (fl & B_META) || (fl & prevf & (B_AMP | B_PIPE));
                prevf = fl;
                const uint32_t cls = (fl & B_BS) ? 1u : (fl & B_SQ) ? 2u : (fl & B_DQ) ? 3u : 0u;
                st = (uint32_t)(T >> (3u * (st * 4u + cls))) & 7u;
            };
            vstep(f);
            if (fx)       // invalid UTF-8 only: the validator's rules are byte rules, the swallowed ASCII bytes still count
                for (int j = 1; j < nact; ++j) { const uint32_t bj = (raw >> (8 * j)) & 0xFFu; vstep(bj < 0x80u ? (uint32_t)lut[bj] : 0u); }
            if (TICKS) {
                const bool tick = b0 == 0x60u;
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
    bool pfx = false;                                          // "kubectl " on the raw bytes at a (app.py:75)
    if (b - a >= 8) {
        const uint32_t sh = (uint32_t)(a & 3) * 8u;
        const int wi = a >> 2;
        const uint32_t w0 = w[wi], w1 = w[wi + 1], w2 = sh ? w[wi + 2] : 0u;   // (the third word only when it holds bytes of the string)
        pfx = funnel_r(w0, w1, sh) == 0x6562756bu && funnel_r(w1, w2, sh) == 0x206c7463u;
    }
    o.a = a; o.b = b;
    o.verdict = !pfx ? KX_BAD_PREFIX : snap_meta ? KX_METACHAR : snap_st != 0 ? KX_BAD_QUOTING : KX_SAFE;
    o.lead_ticks = lead; o.trail_ticks = snap_trail;
    o.san_len = sn; o.any = emitted;
}

// Width in bits of one code point in the compact buffer CPython would allocate for
// the sanitised string: decided by the widest NON-whitespace code point.
KX_HD uint32_t hash_width_bits(const uint32_t *w, int lo, int hi) {
    const uint8_t *B = (const uint8_t *)w;
    uint32_t maxc = 0;
    int p = lo;
    while (p < hi) {
        if ((p & 3) == 0 && p + 4 <= hi && (w[p >> 2] & 0x80808080u) == 0) { p += 4; continue; }   // four ASCII bytes
        const uint32_t b0 = B[p++];
        if (b0 < 0xC0u || b0 >= 0xF8u) continue;              // ASCII, or a byte that is its own code point (< 0x100)
        const int n = b0 < 0xE0u ? 2 : b0 < 0xF0u ? 3 : 4;
        uint32_t c = b0 & (n == 2 ? 0x1Fu : n == 3 ? 0x0Fu : 0x07u);
        for (int j = 1; j < n; ++j) c = (c << 6) | ((p < hi ? (uint32_t)B[p++] : 0u) & 0x3Fu);
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
KX_HD Result process_string(const uint32_t *w, int lo, int hi, const uint8_t *lut, bool want_hash, uint8_t *san) {
    const uint32_t wbits = want_hash ? hash_width_bits(w, lo, hi) : 8u;
    Sip sip;
    sip.init();
    Walk o;
    walk<true, PARSE>(w, lo, hi, wbits, lut, sip, san, o);
    Result r;
    r.hash = (want_hash && o.any) ? sip.finish() : 0;
    r.san_len = o.san_len;
    if (PARSE && o.lead_ticks >= 3 && o.trail_ticks >= 3) {   // app.py:99-100
        const int a2 = o.a + 3, b2 = o.b - 3 > a2 ? o.b - 3 : a2;
        walk<false, false>(w, a2, b2, 8u, lut, sip, nullptr, o);
    }
    r.verdict = o.verdict;
    r.span_a = o.a - lo;
    r.span_b = o.b - lo;
    return r;
}


// --------------------------------------------------------------------------- //
// Fast path: ASCII strings staged in (mutable) shared memory.
//
// The string is stripped by scanning whitespace inwards from both ends and the
// "kubectl " prefix is compared on the raw bytes.  The body is then scanned a WORD at a
// time: the four bytes' flag bytes (one 256-entry table) are packed into one register F
// and three running words summarise everything the validator and the sanitiser need:
//     any |= F                        which kinds of byte occur at all
//     par ^= F                        parity of every kind (quote counts)
//     dbl |= F & (F << 8 | prev >> 24)   a kind directly followed by the same kind ("&&", "||", whitespace runs)
// plus the OR of the data words (any byte >= 0x80 sends the string to the general path).
// From these: metacharacter verdict = META seen or AMP/PIPE doubled; quoting: no quote or
// backslash -> closed; only ' (or only ") -> closed iff its count is even; anything mixed
// or a backslash -> the byte-level automaton (rare).  Sanitising: if no whitespace run and
// no control whitespace occurs, the stripped bytes ARE the sanitised string and SipHash
// reads them where they lie; otherwise (rare) they are compacted in place first.
// SipHash-1-3 runs eight bytes per step, which keeps the lanes of a warp in lockstep.
// --------------------------------------------------------------------------- //
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
// path's forward decoding because what is left in between is pure ASCII (checked by the scan: any byte
// >= 0x80 there sends the string to the general path), so the trailing run starts on a code point boundary
// and each matched unit is one complete encoding.
// Strips both ends of the string in B[lo, hi): ASCII whitespace and the exact UTF-8 encodings of the non-ASCII
// whitespace code points (see above).
KX_HD void strip_ends(const uint8_t *B, int lo, int hi, int &a, int &b) {
    a = lo; b = hi;
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
}

KX_HD bool kubectl_prefix(const uint32_t *W, int a, int b) {
    if (b - a < 8) return false;
    const uint32_t sh = (uint32_t)(a & 3) * 8u;
    const int wi = a >> 2;
    const uint32_t w0 = W[wi], w1 = W[wi + 1], w2 = sh ? W[wi + 2] : 0u;
    return funnel_r(w0, w1, sh) == 0x6562756bu && funnel_r(w1, w2, sh) == 0x206c7463u;   // "kube" "ctl "
}

// Returns 1 if the string must be redone by the general path (a byte >= 0x80 inside, or a fenced output in PARSE
// mode); otherwise fills r.  The byte loops (the quoting automaton for mixed quotes / backslashes, the whitespace
// collapse) run only for the strings that need them, about one in fifteen of the synthetic mix.  (Moving those
// strings to a second phase of the same CTA, or splitting their loops, executed fewer instructions but measured slower:
// 1.15 - 1.32 against 1.04 ms per 10 M strings -- seven of eight warps then wait for one.)
// LOOPS = false: a string that needs one of the byte loops is not touched and 2 is returned (the kernel hands those to a
// pass of their own, where every lane runs the loops, instead of letting one lane in fifteen hold its warp back).
// Returns 0: r is filled; 1: general path; 2: byte loops needed.
template <bool PARSE, bool LOOPS = true>
KX_HD int fast_string(uint32_t *W, int lo, int hi, const uint8_t *lut, bool want_hash, uint8_t *san, Result &r) {
    uint8_t *B = (uint8_t *)W;
    int a, b;
    strip_ends(B, lo, hi, a, b);
    if (PARSE && a < b && B[a] == 0x60) return 1;
    const bool pfx = kubectl_prefix(W, a, b);

    // ---- word-at-a-time scan of [a, b) ----
    uint32_t any = 0, par = 0, dbl = 0, prevF = 0, hibits = 0;
    auto step = [&](uint32_t w) {
        hibits |= w;
        const uint32_t F = pack4(lut[kx_byte(w, 0)], lut[kx_byte(w, 1)], lut[kx_byte(w, 2)], lut[kx_byte(w, 3)]);
        any |= F;
        par ^= F;
        dbl |= F & funnel_r(prevF, F, 24);                     // (F << 8) | (prevF >> 24): every byte next to its predecessor
        prevF = F;
    };
    if (a < b) {
        const int w_first = a >> 2, w_last = (b - 1) >> 2;
        const uint32_t m0 = 0xFFFFFFFFu << (8 * (a & 3));       // bytes outside [a, b) read as 0x00: no flags
        const uint32_t m1 = 0xFFFFFFFFu >> (8 * (3 - ((b - 1) & 3)));
        if (w_first == w_last) {
            step(W[w_first] & m0 & m1);
        } else {
            step(W[w_first] & m0);
            for (int wi = w_first + 1; wi < w_last; ++wi) step(W[wi]);
            step(W[w_last] & m1);
        }
    }
    if (hibits & 0x80808080u) return 1;
    any |= any >> 16; any |= any >> 8;
    par ^= par >> 16; par ^= par >> 8;
    dbl |= dbl >> 16; dbl |= dbl >> 8;
    const bool meta = (any & B_META) || (dbl & (B_AMP | B_PIPE));
    const uint32_t qk = any & (B_SQ | B_DQ | B_BS);
    if (!LOOPS && ((qk != 0 && qk != B_SQ && qk != B_DQ) || (dbl & B_WS) || (any & B_CTRL))) return 2;
    bool open_quote = false;
    if (qk == B_SQ || qk == B_DQ) {
        open_quote = (par & qk) != 0;                          // one kind of quote, no escapes: closed iff the count is even
    } else if (qk != 0) {                                      // mixed quotes or a backslash: the automaton itself (shlex.py:149-267)
        constexpr uint64_t T = dfa_table();
        uint32_t st = 0;
        for (int p = a; p < b; ++p) {
            const uint32_t f = lut[B[p]];
            const uint32_t cls = (f & B_BS) ? 1u : (f & B_SQ) ? 2u : (f & B_DQ) ? 3u : 0u;
            st = (uint32_t)(T >> (3u * (st * 4u + cls))) & 7u;
        }
        open_quote = st != 0;
    }
    r.verdict = !pfx ? KX_BAD_PREFIX : meta ? KX_METACHAR : open_quote ? KX_BAD_QUOTING : KX_SAFE;
    r.span_a = a - lo;
    r.span_b = b - lo;

    int s0 = a, sn = b - a;                                    // where the sanitised bytes lie, and how many
    if ((dbl & B_WS) || (any & B_CTRL)) {
        // whitespace runs or control whitespace inside: collapse to single 0x20, in place at the front of the string
        int o = lo;
        bool prev_nonws = false;
        for (int p = a; p < b; ++p) {
            const uint32_t x = B[p];
            const bool ws = (lut[x] & B_WS) != 0;
            if (!ws || prev_nonws) B[o++] = (uint8_t)(ws ? 0x20u : x);
            prev_nonws = !ws;
        }
        s0 = lo; sn = o - lo;
    }
    r.san_len = sn;
    r.hash = want_hash ? sip_stream(W, s0, sn) : 0;
    if (san) for (int i = 0; i < sn; ++i) san[i] = B[s0 + i];
    return 0;
}

}  // namespace k9
