// Developer logic check (test infrastructure): key_pack8 (pfbwt-f_b200/csrc/dictkeys.cuh), the word-wide packing the key
// kernel uses, against the byte-by-byte definition (sym3_fast / sym_field / sym_stops) for EVERY 8-byte word over the
// dictionary alphabet {0x00, 0x01, 0x02, '-', 'A', 'C', 'G', 'N', 'T'}.
#define PFB_EMU 1
#include "engine.hpp"
namespace pfb {
size_t &dev_bytes_live() { static size_t v = 0; return v; }
size_t &dev_bytes_peak() { static size_t v = 0; return v; }
}
int main() {
    const unsigned char al[9] = {0, 1, 2, '-', 'A', 'C', 'G', 'N', 'T'};
    unsigned long long n = 0;
    int idx[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (;;) {
        pfb::u64 x = 0;
        pfb::u32 f = 0, st = 0, en = 0;
        for (int j = 0; j < 8; ++j) {
            const pfb::u32 c = al[idx[j]];
            x |= (pfb::u64)c << (8 * j);
            const pfb::u32 s = pfb::sym3_fast(c);
            f |= pfb::sym_field(s) << (14 - 2 * j);
            st |= pfb::sym_stops(s) << (7 - j);
            en |= (s == 0 ? 1u : 0u) << (7 - j);
        }
        const pfb::u32 want = f | (st << 16) | (en << 24), got = pfb::key_pack8(x);
        if (want != got) { printf("MISMATCH x=%016llx want=%08x got=%08x\n", (unsigned long long)x, want, got); return 1; }
        ++n;
        int j = 0;
        while (j < 8 && ++idx[j] == 9) idx[j++] = 0;
        if (j == 8) break;
    }
    printf("key_pack8 ok on %llu words\n", n);
    return 0;
}
