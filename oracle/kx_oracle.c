/*
 * TEST INFRASTRUCTURE ONLY -- plain-C CPU restatement of the reference's string
 * path over a CSR batch.  Nothing on the product path links or loads this file;
 * only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline leg do, and
 * only as the checker.
 *
 * Restates, at byte / code-point level:
 *   - sanitize_query            /root/reference/app.py:60-68
 *   - is_safe_kubectl_command   /root/reference/app.py:72-88
 *       (its shlex.split call is CPython 3.12 Lib/shlex.py:133-277 reduced to
 *        the states whose end-of-input raises ValueError)
 *   - KubectlOutputParser.parse /root/reference/app.py:95-104 (fence removal)
 *   - str.__hash__ reached from cache.get / cache[...] at app.py:312,322:
 *     CPython 3.12 Python/pyhash.c siphash13 with a zero key (PYTHONHASHSEED=0)
 *     over the compact Latin-1 / UCS-2 / UCS-4 buffer of the sanitised string.
 *
 * Pinned by tests/test_oracle.py against tests/golden/strings.json, which was
 * produced by executing the reference's own app.py (tests/golden/make_golden.py).
 *
 * Structure (deliberately unlike the CUDA kernel so the two cross-check): each
 * string is first decoded into arrays of code points and byte offsets, then every
 * reference function is evaluated on those arrays.
 *
 * UTF-8 decode rule shared with the kernel ("lenient"): identical to a strict
 * decoder on valid UTF-8; otherwise the lead byte fixes the length, continuation
 * bytes give their low six bits unchecked, a sequence cut by the end of the
 * string is zero-padded, and a byte that cannot lead is its own code point.
 * Whitespace is decided on code points; the validator's prefix, metacharacter
 * and quoting rules are decided on raw bytes (the same thing on valid UTF-8).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define KX_FLAG_PARSE 1u

static int is_ws(uint32_t c) {
    if (c < 0x80) return (c >= 0x09 && c <= 0x0D) || (c >= 0x1C && c <= 0x20);
    if (c == 0x85 || c == 0xA0 || c == 0x1680) return 1;
    if (c >= 0x2000 && c <= 0x200A) return 1;
    return c == 0x2028 || c == 0x2029 || c == 0x202F || c == 0x205F || c == 0x3000;
}

static uint64_t rotl(uint64_t x, int b) { return (x << b) | (x >> (64 - b)); }

#define SIPROUND do { \
    v0 += v1; v2 += v3; v1 = rotl(v1, 13) ^ v0; v3 = rotl(v3, 16) ^ v2; v0 = rotl(v0, 32); \
    v2 += v1; v0 += v3; v1 = rotl(v1, 17) ^ v2; v3 = rotl(v3, 21) ^ v0; v2 = rotl(v2, 32); \
} while (0)

static uint64_t siphash13_zero_key(const uint8_t *p, uint64_t n) {
    uint64_t v0 = 0x736f6d6570736575ULL, v1 = 0x646f72616e646f6dULL;
    uint64_t v2 = 0x6c7967656e657261ULL, v3 = 0x7465646279746573ULL;
    uint64_t b = n << 56, i = 0;
    for (; i + 8 <= n; i += 8) {
        uint64_t m = 0;
        for (int j = 0; j < 8; ++j) m |= (uint64_t)p[i + j] << (8 * j);
        v3 ^= m; SIPROUND; v0 ^= m;
    }
    for (int j = 0; i + j < n; ++j) b |= (uint64_t)p[i + j] << (8 * j);
    v3 ^= b; SIPROUND; v0 ^= b;
    v2 ^= 0xff; SIPROUND; SIPROUND; SIPROUND;
    return (v0 ^ v1) ^ (v2 ^ v3);
}

/* lenient decode; cp[k] with byte range [bo[k], bo[k+1]) */
static int64_t decode(const uint8_t *s, int64_t len, uint32_t *cp, int64_t *bo) {
    int64_t i = 0, k = 0;
    while (i < len) {
        uint32_t b0 = s[i];
        int n = 1;
        uint32_t c = b0;
        if (b0 >= 0xC0 && b0 < 0xF8) {
            n = b0 < 0xE0 ? 2 : b0 < 0xF0 ? 3 : 4;
            c = b0 & (n == 2 ? 0x1F : n == 3 ? 0x0F : 0x07);
            for (int j = 1; j < n; ++j)
                c = (c << 6) | (i + j < len ? (uint32_t)(s[i + j] & 0x3F) : 0u);
            if (i + n > len) n = (int)(len - i);
        }
        cp[k] = c; bo[k] = i; ++k; i += n;
    }
    bo[k] = len;
    return k;
}

static void strip(const uint32_t *cp, int64_t *a, int64_t *b) {
    while (*a < *b && is_ws(cp[*a])) ++*a;
    while (*b > *a && is_ws(cp[*b - 1])) --*b;
}

/* app.py:72-88 on the byte range [lo, hi) (already stripped) */
static uint8_t verdict_bytes(const uint8_t *s, int64_t lo, int64_t hi) {
    if (hi - lo < 8 || memcmp(s + lo, "kubectl ", 8) != 0) return 1;
    for (int64_t i = lo; i < hi; ++i) {
        uint8_t c = s[i];
        if (c == ';' || c == '`' || c == '$' || c == '(' || c == ')' || c == '<' || c == '>') return 2;
        if ((c == '&' || c == '|') && i + 1 < hi && s[i + 1] == c) return 2;
    }
    enum { A, SQ, DQ, EA, ED } st = A;
    for (int64_t i = lo; i < hi; ++i) {
        uint8_t c = s[i];
        switch (st) {
        case A:  st = c == '\\' ? EA : c == '\'' ? SQ : c == '"' ? DQ : A; break;
        case SQ: st = c == '\'' ? A : SQ; break;
        case DQ: st = c == '"' ? A : c == '\\' ? ED : DQ; break;
        case EA: st = A; break;
        case ED: st = DQ; break;
        }
    }
    return st == A ? 0 : 3;
}

static int is_tick(const uint32_t *cp, const int64_t *bo, int64_t k) {
    return cp[k] == 0x60 && bo[k + 1] - bo[k] == 1;
}

static void one_string(const uint8_t *s, int64_t len, uint32_t flags, uint32_t *cp, int64_t *bo,
                       uint8_t *hbuf, uint8_t *verdict, int64_t *hash, int32_t *span,
                       uint8_t *san, int32_t *san_len) {
    int64_t nc = decode(s, len, cp, bo);

    /* --- validator (optionally behind the parser's fence removal) --- */
    int64_t a = 0, b = nc;
    strip(cp, &a, &b);
    if ((flags & KX_FLAG_PARSE) && b - a >= 3 &&
        is_tick(cp, bo, a) && is_tick(cp, bo, a + 1) && is_tick(cp, bo, a + 2) &&
        is_tick(cp, bo, b - 1) && is_tick(cp, bo, b - 2) && is_tick(cp, bo, b - 3)) {
        a += 3;
        b = b - 3 > a ? b - 3 : a;
        strip(cp, &a, &b);
    }
    if (verdict) *verdict = verdict_bytes(s, bo[a], bo[b]);
    if (span) { span[0] = (int32_t)bo[a]; span[1] = (int32_t)bo[b]; }

    /* --- sanitise + hash --- */
    uint32_t maxc = 0;
    int64_t nout = 0, sb = 0;
    int in_tok = 0;
    /* first pass: widest code point, sanitised length, sanitised bytes */
    for (int64_t k = 0; k < nc; ++k) {
        uint32_t c = cp[k];
        if (is_ws(c)) { in_tok = 0; continue; }
        if (!in_tok && nout) {
            if (san) san[sb] = 0x20;
            ++sb;
            ++nout;
        }
        in_tok = 1;
        if (c > maxc) maxc = c;
        if (san) memcpy(san + sb, s + bo[k], (size_t)(bo[k + 1] - bo[k]));
        sb += bo[k + 1] - bo[k];
        ++nout;
    }
    if (san_len) *san_len = (int32_t)sb;
    if (!hash) return;
    if (nout == 0) { *hash = 0; return; }
    int w = maxc < 0x100 ? 1 : maxc < 0x10000 ? 2 : 4;
    int64_t o = 0;
    in_tok = 0;
    for (int64_t k = 0; k < nc; ++k) {
        uint32_t c = cp[k];
        if (is_ws(c)) { in_tok = 0; continue; }
        if (!in_tok && o) { uint32_t sp = 0x20; for (int j = 0; j < w; ++j) hbuf[o++] = (uint8_t)(sp >> (8 * j)); }
        in_tok = 1;
        for (int j = 0; j < w; ++j) hbuf[o++] = (uint8_t)(c >> (8 * j));
    }
    uint64_t h = siphash13_zero_key(hbuf, (uint64_t)o);
    int64_t sh = (int64_t)h;
    *hash = sh == -1 ? -2 : sh;
}

/*
 * Batch entry: rows [row0, row1) of a CSR batch.  Any output pointer may be NULL.
 * span is [2n] (begin,end byte offsets of the validated command inside its string);
 * san receives sanitised bytes at the row's own offset, san_len its length.
 * Returns 0, or -1 on allocation failure.
 */
int kx_oracle_validate_hash(const uint8_t *data, const int64_t *off, int64_t row0, int64_t row1,
                            uint32_t flags, uint8_t *verdict, int64_t *hash, int32_t *span,
                            uint8_t *san, int32_t *san_len) {
    int64_t maxlen = 0;
    for (int64_t i = row0; i < row1; ++i)
        if (off[i + 1] - off[i] > maxlen) maxlen = off[i + 1] - off[i];
    uint32_t *cp = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(maxlen + 1));
    int64_t *bo = (int64_t *)malloc(sizeof(int64_t) * (size_t)(maxlen + 2));
    uint8_t *hb = (uint8_t *)malloc((size_t)(4 * maxlen + 8));
    if (!cp || !bo || !hb) { free(cp); free(bo); free(hb); return -1; }
    for (int64_t i = row0; i < row1; ++i) {
        one_string(data + off[i], off[i + 1] - off[i], flags, cp, bo, hb,
                   verdict ? verdict + i : 0, hash ? hash + i : 0, span ? span + 2 * i : 0,
                   san ? san + off[i] : 0, san_len ? san_len + i : 0);
    }
    free(cp); free(bo); free(hb);
    return 0;
}
