// TEST INFRASTRUCTURE: runs the kernel's own per-string logic (csrc/k9_core.cuh,
// written __host__ __device__) on the CPU so that logic errors are caught in the
// CPU test suite before a GPU is involved.  Addressing mimics the kernel's
// unstaged path (16-byte aligned word base + byte offset).
#include <stdint.h>
#include <string.h>
#include <vector>
#include "../../ai-agent-kubectl_b200/csrc/k9_core.cuh"

// mode 0: fast path with general fallback (what the kernel does); mode 1: general path only.
// Returns the number of strings that took the general path.
extern "C" int64_t k9_emul(const uint8_t *data, const int64_t *off, int64_t n, uint32_t flags, uint8_t *verdict,
                           int64_t *hash, int32_t *span, uint8_t *san, int32_t *san_len, int mode) {
    static const k9::FastTables tab = k9::make_fast_tables();
    int64_t nslow = 0;
    std::vector<uint32_t> buf;
    for (int64_t s = 0; s < n; ++s) {
        const int64_t len = off[s + 1] - off[s];
        const int lead = (int)(s % 16);                       // exercise every misalignment
        buf.assign((size_t)((lead + len + 15) / 16 * 4 + 4), 0xA5A5A5A5u);   // poison around the string
        memcpy((uint8_t *)buf.data() + lead, data + off[s], (size_t)len);
        const bool parse = flags & KX_FLAG_PARSE;
        uint8_t *sp = san ? san + off[s] : nullptr;
        k9::Result r;
        bool slow = true;
        if (mode != 1)     // the kernel's staged fast path (mutates its buffer in place)
            slow = parse ? k9::fast_string<true>(buf.data(), lead, lead + (int)len, tab.flags, hash != nullptr, sp, r)
                         : k9::fast_string<false>(buf.data(), lead, lead + (int)len, tab.flags, hash != nullptr, sp, r);
        if (slow) {        // general path re-reads the pristine bytes, as the kernel does from global memory
            buf.assign(buf.size(), 0x5A5A5A5Au);
            memcpy((uint8_t *)buf.data() + lead, data + off[s], (size_t)len);
            r = parse ? k9::process_string<true>(buf.data(), lead, lead + (int)len, tab.flags, hash != nullptr, sp)
                      : k9::process_string<false>(buf.data(), lead, lead + (int)len, tab.flags, hash != nullptr, sp);
            ++nslow;
        }
        if (hash) hash[s] = r.hash;
        if (san_len) san_len[s] = r.san_len;
        if (verdict) verdict[s] = (uint8_t)r.verdict;
        if (span) { span[2 * s] = r.span_a; span[2 * s + 1] = r.span_b; }
    }
    return nslow;
}
