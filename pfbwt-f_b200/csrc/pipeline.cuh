// pipeline.cuh -- the PFP-BWT hot path as device stages (B200-native re-design of
// the path SURVEY.md §8(a) lists; reference file:line cited per stage).
//
//   parse_text()        K0-K5a : normalise, Wang-hash trigger scan, compaction,
//                                fingerprint dedup (GPU hash table, byte-verified),
//                                phrase sort, ranks, .dict/.occ/.parse images
//                                (include/pfparser.hpp:299-369,484-517,595-601)
//   parse_bwt()         K5b-c  : suffix array of the parse by prefix doubling,
//                                bwlast / bwsai / ilist (include/pfparser.hpp:379-467)
//   build_bwt()         K6-K9  : dictionary suffix sort by prefix doubling, group
//                                detection, segmented emission of BWT/SA, run
//                                boundaries (include/pfbwt.hpp:96-223,
//                                src/pfbwt-f.cpp:298-328)
//
// All stages are written against prim.cuh.  Data layout: byte text T' in HBM
// (Dollar . S . Dollar^w, S 16-byte aligned), 32-bit dictionary / parse indices,
// 64-bit text positions.
#pragma once
#include "prim.cuh"
#include "comm.hpp"
#include "../../include/pfbwt_b200.h"

namespace pfb {


// error codes: PFB_OK / PFB_ERR_* from include/pfbwt_b200.h

static const u8 DOLLAR = 2, ENDOFWORD = 1, ENDOFDICT = 0;   // include/utils.h:8-10

struct Params {
    int w = 10;
    u64 p = 100;
    int bits = 32;              // width of uint_t in the reference binary being replaced
    bool non_acgt_to_a = false;
    bool get_sai = false;
    bool verbose = false;
};

struct StageTimes {   // milliseconds, CUDA events on the pipeline stream
    float normalise = 0, trigger = 0, compact = 0, fingerprint = 0, dedup = 0, phrase_sort = 0, dict_build = 0,
          parse_sa = 0, parse_bwt = 0, dict_sort = 0, groups = 0, emit = 0, runs = 0;
    u32 dedup_attempts = 0, phrase_sort_rounds = 0, parse_sa_rounds = 0, dict_sort_rounds = 0;
};

// the device-resident parse: what PfParser holds after finalize()+bwt_of_parse()
struct DeviceParse {
    u64 n = 0, P = 0, Dw = 0, dsize = 0;
    DevBuf<u8> dict;        // dsize           .dict image
    DevBuf<u64> wstart;     // Dw+1            word start offsets in dict (64-bit: a dictionary may pass 2^32 bytes)
    DevBuf<u32> occ;        // Dw              .occ
    DevBuf<u32> parse;      // P+1             .parse (+ EOS slot)
    DevBuf<u8> last;        // P               last_
    DevBuf<u64> sai;        // P               sai_
    DevBuf<u8> bwlast;      // P+1
    DevBuf<u32> ilist;      // P+1
    DevBuf<u64> bwsai;      // P+1
    bool has_parse_bwt = false;
};

// With a communicator (multi-GPU job) a rank holds one contiguous slice of every output: BWT
// positions [off, off+len), .ssa pairs [ssa_off, ssa_off+ssa_cnt), .esa pairs [esa_off, ..); r, n
// are the totals of the whole job.  Without one the slice is everything.
struct DeviceBwt {
    u64 n = 0, r = 0, easy = 0, hard = 0;
    u64 off = 0, len = 0, ssa_off = 0, ssa_cnt = 0, esa_off = 0, esa_cnt = 0;
    DevBuf<u8> bwt;         // len
    DevBuf<u8> sa;          // len*U bytes, U=bits/8 (empty unless any_sa)
    DevBuf<u8> ssa, esa;    // ssa_cnt / esa_cnt pairs of U-byte ints each
};

// PFB_TRACE=1: every lap() is timed (CUDA events) and printed to stderr at the end of a run;
// without it lap() costs nothing (no event, no synchronisation).
struct TraceLog {
    bool on = getenv("PFB_TRACE") != nullptr;
    std::vector<std::pair<std::string, float>> v;
    void dump(const char *title) {
        if (!on) return;
        float tot = 0;
        for (auto &e : v) tot += e.second;
        fprintf(stderr, "[pfb trace] %s: %.3f ms in %zu laps\n", title, tot, v.size());
        for (auto &e : v) fprintf(stderr, "[pfb trace]   %-28s %9.3f ms\n", e.first.c_str(), e.second);
        v.clear();
    }
};
inline TraceLog &trace_log() { static TraceLog t; return t; }

struct Timer {
#ifndef PFB_EMU
    cudaEvent_t a, b, c;
    pfbStream_t s;
    Timer(pfbStream_t st) : s(st) {
        cudaEventCreate(&a); cudaEventCreate(&b); cudaEventCreate(&c);
        cudaEventRecord(a, s);
        if (trace_log().on) cudaEventRecord(c, s);
    }
    float stop() { cudaEventRecord(b, s); cudaEventSynchronize(b); float ms = 0; cudaEventElapsedTime(&ms, a, b); cudaEventRecord(a, s); return ms; }
    void lap(const char *name) {
        if (!trace_log().on) return;
        cudaEventRecord(b, s); cudaEventSynchronize(b);
        float ms = 0; cudaEventElapsedTime(&ms, c, b);
        trace_log().v.emplace_back(name, ms);
        cudaEventRecord(c, s);
    }
    ~Timer() { cudaEventDestroy(a); cudaEventDestroy(b); cudaEventDestroy(c); }
#else
    Timer(pfbStream_t) {}
    float stop() { return 0.f; }
    void lap(const char *name) { if (trace_log().on) trace_log().v.emplace_back(name, 0.f); }      // the steps taken, untimed
#endif
};

// ---------------------------------------------------------------- symbols --
// include/hash.hpp:12-21
PFB_HD inline u64 wang_hash(u64 key) {
    key = (~key) + (key << 21);
    key = key ^ (key >> 24);
    key = (key + (key << 3)) + (key << 8);
    key = key ^ (key >> 14);
    key = (key + (key << 2)) + (key << 4);
    key = key ^ (key >> 28);
    key = key + (key << 31);
    return key;
}
// seq_nt4_ntoa_table restricted to the normalised alphabet (src/utils.c:139-161):
// A,N -> 0  C -> 1  G -> 2  T,'-' -> 3
PFB_HD inline u32 nt_code(u8 c) {
    u32 x = (c >> 1) & 3;       // A->0 C->1 G->3 T->2 '-'->2
    x ^= x >> 1;                // A->0 C->1 G->2 T->3 '-'->3
    return c == 'N' ? 0u : x;
}
// order-preserving 3-bit code of the 8 bytes that can occur in T'/dict:
// 0x01<0x02<'-'<'A'<'C'<'G'<'N'<'T'  ->  0..7   (0 doubles as "past the end")
PFB_HD inline u32 sym3(u8 c) {
    switch (c) {
    case 1: return 0;
    case 2: return 1;
    case '-': return 2;
    case 'A': return 3;
    case 'C': return 4;
    case 'G': return 5;
    case 'N': return 6;
    case 'T': return 7;
    default: return 0;   // EndOfDict (0x00)
    }
}

// divisibility of a 64-bit value by a runtime constant without a division:
// d = 2^k * d', d' odd:  d | x  <=>  ror(x * inv(d'), k) <= floor((2^64-1)/d)
struct DivTest {
    u64 inv = 1, lim = ~0ull;
    u32 k = 0;
    static DivTest make(u64 d) {
        DivTest t;
        u32 k = 0;
        u64 dd = d;
        while ((dd & 1) == 0) { dd >>= 1; ++k; }
        u64 x = dd;   // Newton: x*dd = 1 mod 2^3 initially
        for (int i = 0; i < 6; ++i) x *= 2 - dd * x;
        t.inv = x; t.k = k; t.lim = ~0ull / d;
        return t;
    }
    PFB_HD inline bool divides(u64 x) const {
        u64 m = x * inv;
        m = k ? ((m >> k) | (m << (64 - k))) : m;
        return m <= lim;
    }
};

// ---------------------------------------------------------------------------
// K0  normalise the record bytes in place: toupper, optional non-ACGT -> 'A',
//     validate against the hasher's alphabet.   (include/pfparser.hpp:337-346,
//     include/hash.hpp:30-31).  8 bytes per thread, S is 16-byte aligned.
// ---------------------------------------------------------------------------
PFB_DEV inline u64 normalise_word(u64 x, u64 lim, bool non_acgt_to_a, u64 base, u64 *d_errpos) {
    u64 y = 0;
    for (u64 j = 0; j < lim; ++j) {
        u32 c = (u32)(x >> (8 * j)) & 0xff;
        if (c >= 'a' && c <= 'z') c -= 32;
        bool acgt = (c == 'A') | (c == 'C') | (c == 'G') | (c == 'T');
        if (non_acgt_to_a && !acgt) { c = 'A'; acgt = true; }
        if (!(acgt | (c == 'N') | (c == '-'))) atomic_min(d_errpos, base + j);
        y |= (u64)c << (8 * j);
    }
    return y;
}
inline void k_normalise(u8 *S, u64 n, bool non_acgt_to_a, u64 *d_errpos, pfbStream_t st) {
    u64 nchunks = (n + 15) / 16;
    launch_for(nchunks, st, PFB_LAMBDA(u64 i) {
        u64 base = i * 16;
        if (base + 16 <= n) {
#ifdef __CUDA_ARCH__
            uint4 v = *reinterpret_cast<const uint4 *>(S + base);
            u64 x0 = (u64)v.x | ((u64)v.y << 32), x1 = (u64)v.z | ((u64)v.w << 32);
#else
            u64 x0 = *reinterpret_cast<const u64 *>(S + base), x1 = *reinterpret_cast<const u64 *>(S + base + 8);
#endif
            // fast path: a word of upper-case ACGT needs no work (the common case); everything else goes byte by byte
            u64 y0 = normalise_word(x0, 8, non_acgt_to_a, base, d_errpos);
            u64 y1 = normalise_word(x1, 8, non_acgt_to_a, base + 8, d_errpos);
            if (y0 != x0 || y1 != x1) {
#ifdef __CUDA_ARCH__
                *reinterpret_cast<uint4 *>(S + base) = make_uint4((u32)y0, (u32)(y0 >> 32), (u32)y1, (u32)(y1 >> 32));
#else
                *reinterpret_cast<u64 *>(S + base) = y0; *reinterpret_cast<u64 *>(S + base + 8) = y1;
#endif
            }
            return;
        }
        for (u64 b = base; b < n; b += 8) {
            u64 lim = n - b < 8 ? n - b : 8;
            u64 x = 0;
            for (u64 j = 0; j < lim; ++j) x |= (u64)S[b + j] << (8 * j);
            u64 y = normalise_word(x, lim, non_acgt_to_a, b, d_errpos);
            for (u64 j = 0; j < lim; ++j) S[b + j] = (u8)(y >> (8 * j));
        }
    });
}

// ---------------------------------------------------------------------------
// K1  trigger bitmap: bit q of word q/32 set iff q > w and
//     wang_hash(kmer ending at T'[q]) % p == 0          (include/pfparser.hpp:347)
//     Tp points at T'[0]; valid q range is [w+1, n].
// ---------------------------------------------------------------------------
//     A piece of T' handled by one rank of a multi-GPU job starts at a trigger window, so its first
//     admissible position is qmin = w instead of w+1 (the global rule q > w holds for all of it).
inline void k_trigger_words(const u8 *Tp, u64 n, int w, DivTest dt, u32 *trig, u32 *cnt, u64 nwords, pfbStream_t st, u64 qmin = 0) {
    const u64 mask = (w >= 32) ? ~0ull : ((1ull << (2 * w)) - 1);
    if (qmin == 0) qmin = (u64)w + 1;
    launch_for(nwords, st, PFB_LAMBDA(u64 t) {
        u64 q0 = t * 32, q1 = q0 + 31;
        if (q1 > n) q1 = n;
        u64 qs = q0 < qmin ? qmin : q0;
        u32 bits = 0;
        if (qs <= q1) {
            u64 kmer = 0;
            for (u64 j = qs - (u64)w + 1; j < qs; ++j) kmer = (kmer << 2) | nt_code(Tp[j]);
            for (u64 q = qs; q <= q1; ++q) {
                kmer = ((kmer << 2) | nt_code(Tp[q])) & mask;
                if (dt.divides(wang_hash(kmer))) bits |= 1u << (q - q0);
            }
        }
        trig[t] = bits;
        cnt[t] = (u32)popc32(bits);
    });
}

// K2  scatter trigger positions -> ends[] (sorted), ends[P-1] = n+w
inline void k_scatter_ends(const u32 *trig, const u32 *wordoff, u64 nwords, u64 *ends, u64 P, u64 n, int w, pfbStream_t st,
                           bool dollar_end = true) {
    launch_for(nwords, st, PFB_LAMBDA(u64 t) {
        u32 bits = trig[t];
        u64 o = wordoff[t];
        while (bits) {
            int b = ffs32(bits) - 1;
            ends[o++] = t * 32 + (u64)b;
            bits &= bits - 1;
        }
        if (t == 0 && dollar_end) ends[P - 1] = n + (u64)w;
    });
}

PFB_HD inline u64 phrase_start(const u64 *ends, u64 k, int w) { return k ? ends[k - 1] - (u64)w + 1 : 0; }

// ---------------------------------------------------------------------------
// K3  phrase fingerprints: fp_b(k) = sum_j T'[s_k+j] * B^(len-1-j)  (mod 2^64) for two
//     odd bases.  One logical thread per 64-byte slice of T'; a slice adds its
//     share of every phrase it overlaps with an atomicAdd, so Mbp-long phrases
//     are spread over many threads.
// ---------------------------------------------------------------------------
static const u64 FP_B1 = 0x9E3779B97F4A7C15ull, FP_B2 = 0xC2B2AE3D27D4EB4Full;
// 8 bytes at any address (little endian) from the two aligned words around it (the text buffers carry >= 16 spare bytes)
PFB_HD inline u64 load8_any(const u8 *p) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    const u64 *w = reinterpret_cast<const u64 *>(a & ~(uintptr_t)7);
    const u32 sh = (u32)(a & 7) * 8;
    const u64 lo = w[0];
    return sh ? (lo >> sh) | (w[1] << (64 - sh)) : lo;
}
// bytes from .. to (inclusive, 0..7) of a little-endian word
PFB_HD inline u64 bytes_mask(int from, int to) { return (~0ull << (8 * from)) & (~0ull >> (8 * (7 - to))); }
PFB_HD inline u64 pow_u64(u64 b, u64 e) {
    u64 r = 1;
    while (e) { if (e & 1) r *= b; b *= b; e >>= 1; }
    return r;
}
inline void k_fingerprint(const u8 *Tp, u64 tlen, const u64 *ends, u64 P, int w, const u32 *wordoff, u64 nwords, u64 *fp1,
                          u64 *fp2, pfbStream_t st) {
    u64 nslices = (tlen + 63) / 64;
    launch_for(nslices, st, PFB_LAMBDA(u64 c) {
        u64 lo = c * 64, hi = lo + 63;
        if (hi > tlen - 1) hi = tlen - 1;
        u64 k = (lo / 32 < nwords) ? (u64)wordoff[lo / 32] : P - 1;
        for (; k < P; ++k) {
            u64 s = phrase_start(ends, k, w), e = ends[k];
            if (s > hi) break;
            u64 a = s > lo ? s : lo, b = e < hi ? e : hi;
            // the phrase as 8-byte words counted from ITS first byte (the last one zero padded past e): a word belongs to the
            // slice that holds its first byte.  Every word is mixed before it enters the polynomial: two phrases that differ
            // in the top byte of a word (one substituted base) would otherwise differ in the top bits of both sums only
            const u64 q0 = (a - s + 7) >> 3, q1 = (b - s) >> 3, nw = ((e - s) >> 3) + 1;
            if (q0 > q1) continue;
            u64 h1 = 0, h2 = 0;
            for (u64 q = q0; q <= q1; ++q) {
                const u64 pq = s + 8 * q;
                const u64 x = load8_any(Tp + pq) & bytes_mask(0, e < pq + 7 ? (int)(e - pq) : 7);
                u64 f = (x ^ (x >> 32)) * 0x9FB21C651E98DF25ull;
                u64 g = (x ^ (x >> 27)) * 0xD6E8FEB86659FD93ull;
                h1 = h1 * FP_B1 + (f ^ (f >> 29)); h2 = h2 * FP_B2 + (g ^ (g >> 31));
            }
            u64 rem = nw - 1 - q1;
            if (rem) { h1 *= pow_u64(FP_B1, rem); h2 *= pow_u64(FP_B2, rem); }
            atomic_add(&fp1[k], h1);
            atomic_add(&fp2[k], h2);
        }
    });
}

PFB_HD inline u64 mix64(u64 x) {   // splitmix64 finaliser
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 27; x *= 0x94D049BB133111EBull;
    x ^= x >> 31;
    return x;
}

// ---------------------------------------------------------------------------
// K4  dedup: open-addressing table keyed by a 64-bit mix of (fp1, fp2, len).
//     The representative of a phrase is the smallest text-order index with the
//     same key; every other occurrence is byte-verified against it (slice-wise),
//     so a fingerprint collision can never change the output: it only sets a flag
//     and the table is rebuilt with another seed.   Replaces the std::map insert of
//     process_phrase (include/pfparser.hpp:595-598).
// ---------------------------------------------------------------------------
inline void k_ht_insert(const u64 *fp1, const u64 *fp2, const u64 *ends, u64 P, int w, u64 seed, u64 *tkey, u32 *trep,
                        u64 tmask, u32 *slot, pfbStream_t st) {
    launch_for(P, st, PFB_LAMBDA(u64 k) {
        u64 len = ends[k] - phrase_start(ends, k, w) + 1;
        u64 key = mix64(fp1[k] + seed * 0x632BE59BD9B4E019ull) ^ mix64(fp2[k] ^ (len * 0xD6E8FEB86659FD93ull) ^ seed);
        if (key == 0) key = 1;
        u64 h = (key >> 17) & tmask;
        for (;;) {
            u64 old = atomic_cas(&tkey[h], (u64)0, key);
            if (old == 0 || old == key) {
                atomic_min(&trep[h], (u32)k);
                slot[k] = (u32)h;
                break;
            }
            h = (h + 1) & tmask;
        }
    });
}
inline void k_verify(const u8 *Tp, u64 tlen, const u64 *ends, u64 P, int w, const u32 *wordoff, u64 nwords, const u32 *rep,
                     u32 *d_flag, pfbStream_t st) {
    u64 nslices = (tlen + 63) / 64;
    launch_for(nslices, st, PFB_LAMBDA(u64 c) {
        u64 lo = c * 64, hi = lo + 63;
        if (hi > tlen - 1) hi = tlen - 1;
        u64 k = (lo / 32 < nwords) ? (u64)wordoff[lo / 32] : P - 1;
        for (; k < P; ++k) {
            u64 s = phrase_start(ends, k, w), e = ends[k];
            if (s > hi) break;
            u64 r = rep[k];
            if (r == k) continue;
            u64 rs = phrase_start(ends, r, w), re = ends[r];
            bool bad = (re - rs) != (e - s);
            if (!bad) {
                u64 a = s > lo ? s : lo, b = e < hi ? e : hi;
                for (u64 pq = a; pq <= b; pq += 8) {                 // 8 bytes per step, the tail masked
                    const u64 mk = bytes_mask(0, b < pq + 7 ? (int)(b - pq) : 7);
                    bad |= ((load8_any(Tp + pq) ^ load8_any(Tp + rs + (pq - s))) & mk) != 0;
                }
            }
            if (bad) atomic_or(d_flag, 1u);
        }
    });
}

// pack `cnt` (<=21) 3-bit symbols of a string starting at p[off], total length len
// (symbols past the end are 0), most significant symbol first.
PFB_HD inline u64 pack_syms(const u8 *p, u64 len, u64 off, int cnt) {
    u64 key = 0;
    for (int j = 0; j < cnt; ++j) {
        u64 o = off + (u64)j;
        key = (key << 3) | (o < len ? sym3(p[o]) : 0u);
    }
    return key;
}

// ---- pivot keys (shared by the phrase sort below and the dictionary suffix sort of build.cuh) ----------
// A member of a tie group is compared with the group's first member (the pivot) from the group's depth on;
// l = common length, x / y = the member's / pivot's byte at the first difference (0 = string ended).
static const int PIVOT_LBITS = 12;
static const u32 PIVOT_CAP = (1u << PIVOT_LBITS) - 1;
// lb = bits of the common-length field (PIVOT_LBITS unless the dictionary holds long single-symbol runs, build.cuh)
PFB_HD inline u64 pivot_key_result(u32 l, u32 x, u32 y, int lb = PIVOT_LBITS) {
    if (x == y) return (1ull << (lb + 4)) | 1ull;               // identical up to and including EndOfWord: closed
    const u64 cap = (1ull << lb) - 1;
    const u64 side = x < y ? 0ull : 2ull;
    const u64 lf = x < y ? (u64)l : cap - (u64)l;
    return (side << (lb + 4)) | (lf << 4) | ((u64)sym3((u8)x) << 1);
}
PFB_HD inline u64 pivot_key_clamped(int lb = PIVOT_LBITS) { return (1ull << (lb + 4)) | (1ull << 1); }   // open, tied with the pivot

// symbols consumed by the occupant of a slot after a pivot round
PFB_HD inline u32 pivot_consumed(u64 key, int lb = PIVOT_LBITS) {
    const u32 cap = (u32)((1ull << lb) - 1);
    const u32 side = (u32)(key >> (lb + 4)) & 3u, lf = (u32)(key >> 4) & cap;
    if (side == 0) return lf + 1;
    if (side == 2) return (cap - lf) + 1;
    return (key & 1) ? 0u : cap;
}

// strings buf[sa, sa+la) and buf[sp, sp+lp), compared from offset `depth` on
PFB_HD inline u64 pivot_key_len(const u8 *buf, u64 sa, u64 la, u64 sp, u64 lp, u64 depth) {
    const u64 ra = la > depth ? la - depth : 0, rp = lp > depth ? lp - depth : 0;
    const u64 mn = ra < rp ? ra : rp;
    const u8 *x = buf + sa + depth, *y = buf + sp + depth;
    u32 l = 0;
    for (; l < PIVOT_CAP && l < mn; ++l)
        if (x[l] != y[l]) return pivot_key_result(l, x[l], y[l]);
    if (l >= PIVOT_CAP) return (1ull << (PIVOT_LBITS + 4)) | (1ull << 1);       // clamped
    return pivot_key_result(l, ra > l ? x[l] : 0u, rp > l ? y[l] : 0u);         // both ended: identical (closed)
}

}  // namespace pfb
#include "parsekernels.cuh"
namespace pfb {

struct Workspace {
    Scratch sc;
    pfbStream_t st = 0;
    TriggerLut lut;          // trigger table of the last (w, p): parsekernels.cuh
};

// compaction helper: out positions for flagged items; returns number selected
inline u64 scan_flags(const u8 *flags, u32 *idx, u64 n, Workspace &ws) {
    if (n == 0) return 0;
    exclusive_sum<u8, u32>(flags, idx, n, ws.sc, ws.st);
    u32 last_idx = read_scalar(idx + (n - 1), ws.st);
    u8 last_flag = read_scalar(flags + (n - 1), ws.st);
    return (u64)last_idx + last_flag;
}

// Orders the members of every group of an active list by key.  Groups are contiguous in the list; the group id leads the
// key (bits >= gshift) and hs[g] = list slot of group g's first member, hs[G] = m.  Members of a group of at most `cap`
// members rank themselves by counting the smaller keys of their group - one pass, the group's keys come out of L1/L2 -;
// larger groups are compacted and take the radix sort (whole groups, group id in the key: the i-th sorted pair belongs
// into the i-th of their slots).  Tie groups are small where refinement rounds cost the most (repetitive inputs: the
// variants of one phrase, the copies of one phrase sequence), so rounds become linear passes instead of 6-7 radix passes.
// Result in k / v (buffers swapped); hl[a]: 1 = first, 2 = last member with its key (bits >= kshift) inside its group.
static const u32 SMALL_GROUP_DEFAULT = 48;
inline u32 small_group_cap() {
    static const u32 c = getenv("PFB_SMALL_GROUP") ? (u32)atoi(getenv("PFB_SMALL_GROUP")) : SMALL_GROUP_DEFAULT;   // test knob: 0 = radix sort only
    return c;
}
inline void order_groups(DevBuf<u64> &k, DevBuf<u64> &kb, DevBuf<u32> &v, DevBuf<u32> &vb, DevBuf<u8> &hl, u64 m, const u32 *hs,
                         int gshift, int kshift, int sortbits, Workspace &ws) {
    pfbStream_t st = ws.st;
    DevBuf<u8> bigf(m);
    {
        const u64 *kp = k.get(); const u32 *vp = v.get(); u64 *ko = kb.get(); u32 *vo = vb.get(); u8 *hp = hl.get(); u8 *bf = bigf.get();
        const u32 cap = small_group_cap();
        launch_for(m, st, PFB_LAMBDA(u64 a) {
            const u64 gid = kp[a] >> gshift;
            const u32 s = hs[gid], e = hs[gid + 1];
            if (e - s > cap) { bf[a] = 1; return; }
            bf[a] = 0;
            const u64 mine = kp[a] >> kshift;
            u32 less = 0, eqb = 0, eqa = 0;
            for (u32 j = s; j < e; ++j) {
                const u64 o = kp[j] >> kshift;
                less += o < mine ? 1u : 0u;
                if (o == mine) { if (j < (u32)a) ++eqb; else if (j > (u32)a) ++eqa; }
            }
            const u32 dest = s + less + eqb;               // stable
            ko[dest] = kp[a]; vo[dest] = vp[a];
            hp[dest] = (u8)((eqb == 0 ? 1 : 0) | (eqa == 0 ? 2 : 0));
        });
    }
    DevBuf<u32> bidx(m);
    const u64 mb = scan_flags(bigf.get(), bidx.get(), m, ws);
    if (mb) {
        DevBuf<u64> bk(mb), bkb(mb);
        DevBuf<u32> bv(mb), bvb(mb), bs(mb);
        {
            const u8 *bf = bigf.get(); const u32 *bi = bidx.get(); const u64 *kp = k.get(); const u32 *vp = v.get();
            u64 *bkp = bk.get(); u32 *bvp = bv.get(); u32 *bsp = bs.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) { if (bf[a]) { const u32 i = bi[a]; bkp[i] = kp[a]; bvp[i] = vp[a]; bsp[i] = (u32)a; } });
        }
        sort_pairs(bk, bkb, bv, bvb, mb, kshift, sortbits, ws.sc, st);
        {
            const u32 *bsp = bs.get(); const u64 *bkp = bk.get(); const u32 *bvp = bv.get(); u64 *ko = kb.get(); u32 *vo = vb.get(); u8 *hp = hl.get();
            launch_for(mb, st, PFB_LAMBDA(u64 i) {
                const u64 x = bkp[i] >> kshift;
                const bool head = i == 0 || x != (bkp[i - 1] >> kshift);
                const bool last = i + 1 == mb || x != (bkp[i + 1] >> kshift);
                const u32 dest = bsp[i];
                ko[dest] = bkp[i]; vo[dest] = bvp[i];
                hp[dest] = (u8)((head ? 1 : 0) | (last ? 2 : 0));
            });
        }
    }
    k.swap(kb); v.swap(vb);
}

// ---------------------------------------------------------------------------
// K5a  lexicographic sort of the distinct phrases (strcmp order,
//      include/pfparser.hpp:494-502) by MSD rounds of packed 3-bit symbols:
//      round 0 sorts on 21 symbols; later rounds re-sort only the members of tie
//      groups on (dense group id, next symbols).
//      order[r] = text-order index k of the phrase of rank r.
// ---------------------------------------------------------------------------
//      The strings are given by an accessor str(k, &offset, &len) into the byte buffer Tp.  With
//      flag_out != nullptr equal strings are allowed (union of the ranks' dictionaries in a multi-GPU
//      job): a tie group whose members are exhausted is a set of identical strings and is left alone;
//      flag_out[j] = 1 at the first member of every group of identical strings in the sorted order.
struct PhraseStr {      // phrase k of the parse of Tp
    const u64 *ends; int w;
    PFB_HD inline void operator()(u64 k, u64 &s, u64 &len) const { s = phrase_start(ends, k, w); len = ends[k] - s + 1; }
};
struct WordStr {        // word k of a dictionary image (EndOfWord excluded)
    const u64 *wstart;  // Dw+1 offsets
    PFB_HD inline void operator()(u64 k, u64 &s, u64 &len) const { s = wstart[k]; len = wstart[k + 1] - 1 - s; }
};

template <class Str>
inline int sort_strings(const u8 *Tp, Str str, DevBuf<u32> &order, u64 Dw, Workspace &ws, u32 *rounds_out, std::string &err,
                        DevBuf<u8> *flag_out = nullptr) {
    pfbStream_t st = ws.st;
    const bool dups = flag_out != nullptr;
    DevBuf<u64> key_a(Dw), key_b(Dw);
    DevBuf<u32> val_b(Dw);
    DevBuf<u8> flag(Dw);
    {
        u64 *ka = key_a.get(); const u32 *ord = order.get();
        launch_for(Dw, st, PFB_LAMBDA(u64 j) {
            u64 s, len;
            str(ord[j], s, len);
            ka[j] = pack_syms(Tp + s, len, 0, 21);
        });
    }
    sort_pairs(key_a, key_b, order, val_b, Dw, 0, 63, ws.sc, st);
    {
        const u64 *ka = key_a.get(); u8 *fl = flag.get();
        launch_for(Dw, st, PFB_LAMBDA(u64 j) { fl[j] = (j == 0 || ka[j] != ka[j - 1]) ? 1 : 0; });
    }
    // dep[j]: symbols on which the string at sorted position j is known to agree with its tie group
    // (uniform over a group; ~0 = member of a closed group of identical strings)
    const u32 CLOSED = 0xFFFFFFFFu;
    DevBuf<u32> dep(Dw);
    { u32 *dp = dep.get(); launch_for(Dw, st, PFB_LAMBDA(u64 j) { dp[j] = 21u; }); }
    u32 rounds = 1, pivot_rounds = 0;
    int pivot_knob = -1;
    if (const char *e = getenv("PFB_PIVOT")) pivot_knob = atoi(e);          // test knob (see build.cuh)
    DevBuf<u8> act(Dw), acthead(Dw);
    DevBuf<u32> aidx(Dw), gidx(Dw);
    for (;;) {
        {
            const u8 *fl = flag.get(); u8 *ac = act.get(); u8 *ah = acthead.get(); const u32 *ord = order.get(); const u32 *dp = dep.get();
            launch_for(Dw, st, PFB_LAMBDA(u64 j) {
                bool single = fl[j] && (j + 1 == Dw || fl[j + 1]);
                if (!single && dp[j] == CLOSED) single = true;
                if (dups && !single) {      // tied on dep symbols and exhausted: identical strings (uniform over the group)
                    u64 s, len;
                    str(ord[j], s, len);
                    single = len <= (u64)dp[j];
                }
                ac[j] = single ? 0 : 1;
                ah[j] = (!single && fl[j]) ? 1 : 0;
            });
        }
        u64 m = scan_flags(act.get(), aidx.get(), Dw, ws);
        if (m == 0) break;
        u64 G = scan_flags(acthead.get(), gidx.get(), Dw, ws);
        int gbits = bits_for(G ? G - 1 : 0);
        // long ties (variant phrases of a pangenome differ hundreds of symbols in): pivot rounds, as in the
        // dictionary suffix sort (build.cuh) - the whole group is ordered by one comparison with its first member
        const bool pivot = pivot_knob != 0 && pivot_rounds < 4096 && (pivot_knob == 1 || rounds >= 2 || m > Dw / 8);
        int nsym = (63 - gbits) / 3;
        if (nsym > 21) nsym = 21;
        if (nsym < 1 || gbits + PIVOT_LBITS + 6 > 64) { err = "phrase sort: too many tie groups"; return PFB_ERR_INTERNAL; }
        if (rounds > 200000) { err = "phrase sort: distinct phrases share a prefix longer than supported"; return PFB_ERR_LIMIT; }
        DevBuf<u32> apos(m), aval(m), aval_b(m);
        DevBuf<u64> akey(m), akey_b(m);
        {
            const u8 *ac = act.get(); const u8 *fl = flag.get();
            const u32 *ai = aidx.get(); const u32 *gi = gidx.get(); const u32 *ord = order.get(); const u32 *dp = dep.get();
            u32 *ap = apos.get(); u32 *av = aval.get(); u64 *ak = akey.get();
            DevBuf<u32> hpos(pivot ? G + 1 : 1);                 // sorted position of every group's first member
            u32 *hp = hpos.get();
            if (pivot) launch_for(Dw, st, PFB_LAMBDA(u64 j) { if (ac[j] && fl[j]) hp[gi[j]] = (u32)j; });
            launch_for(Dw, st, PFB_LAMBDA(u64 j) {
                if (!ac[j]) return;
                u32 a = ai[j];
                u64 gid = (u64)gi[j] + (fl[j] ? 1 : 0) - 1;
                u64 k = ord[j];
                u64 s, len;
                str(k, s, len);
                ap[a] = (u32)j; av[a] = (u32)k;
                if (pivot) {
                    u64 ps, pl;
                    str(ord[hp[gid]], ps, pl);
                    ak[a] = (gid << (PIVOT_LBITS + 6)) | pivot_key_len(Tp, s, len, ps, pl, dp[j]);
                } else ak[a] = (gid << (3 * nsym)) | pack_syms(Tp + s, len, dp[j], nsym);
            });
        }
        DevBuf<u8> hl(m);
        {
            DevBuf<u32> hslot(G + 1);
            const u8 *ah = acthead.get(); const u32 *ai = aidx.get(); const u32 *gi = gidx.get(); u32 *hs = hslot.get();
            const u32 m32 = (u32)m; const u64 G_ = G, Dw_ = Dw;
            launch_for(Dw + 1, st, PFB_LAMBDA(u64 j) { if (j == Dw_) hs[G_] = m32; else if (ah[j]) hs[gi[j]] = ai[j]; });
            order_groups(akey, akey_b, aval, aval_b, hl, m, hslot.get(), pivot ? PIVOT_LBITS + 6 : 3 * nsym, pivot ? 1 : 0,
                         pivot ? gbits + PIVOT_LBITS + 6 : gbits + 3 * nsym, ws);
        }
        {
            const u32 *ap = apos.get(); const u32 *av = aval.get(); const u64 *ak = akey.get(); const u8 *hlp = hl.get();
            u32 *ord = order.get(); u8 *fl = flag.get(); u32 *dp = dep.get();
            launch_for(m, st, PFB_LAMBDA(u64 a) {
                u32 j = ap[a];
                ord[j] = av[a];
                fl[j] = (hlp[a] & 1) ? 1 : 0;
                // position j stays inside its old group, so dp[j] still holds that group's depth
                if (!pivot) dp[j] += (u32)nsym;
                else dp[j] = (ak[a] & 1) ? CLOSED : dp[j] + pivot_consumed(ak[a]);
            });
        }
        if (pivot) ++pivot_rounds;
        ++rounds;
    }
    if (rounds_out) *rounds_out = rounds;
    if (flag_out) *flag_out = std::move(flag);
    return PFB_OK;
}
inline int sort_phrases(const u8 *Tp, const u64 *ends, int w, DevBuf<u32> &order, u64 Dw, Workspace &ws, u32 *rounds_out,
                        std::string &err) {
    PhraseStr ps;
    ps.ends = ends; ps.w = w;
    return sort_strings(Tp, ps, order, Dw, ws, rounds_out, err);
}

// ---------------------------------------------------------------------------
// Prefix doubling (Larsson-Sadakane numbering: rank = position of the group head
// in sa).  On entry sa is sorted by the first h symbols, flag[j]=1 at group heads,
// rank[i] = head position of suffix i's group.  `done(idx,h)` says whether a tied
// suffix is already complete within h symbols (dictionary suffixes end at their
// EndOfWord; equal complete suffixes stay one group).  Replaces sacak_int /
// gsacak (gsa/gsacak.c:2499-2524).
// ---------------------------------------------------------------------------
template <class DoneFn>
inline int prefix_doubling(DevBuf<u32> &sa, DevBuf<u32> &rank, DevBuf<u8> &flag, u64 m, u64 h, DoneFn done, Workspace &ws,
                           u32 *rounds_out, std::string &err) {
    pfbStream_t st = ws.st;
    const int rbits = bits_for(m ? m - 1 : 0);
    u32 rounds = 0;
    // the open list: sorted positions j of the members of tie groups that still have to be refined, ascending.  Built
    // once from the flags; after that a round only touches the list (it shrinks geometrically on typical inputs), never
    // all m positions again.
    DevBuf<u32> apos;
    u64 ma;
    {
        DevBuf<u8> act(m);
        const u8 *fl = flag.get(); const u32 *sp = sa.get(); u8 *ac = act.get();
        launch_for(m, st, PFB_LAMBDA(u64 j) {
            bool single = fl[j] && (j + 1 == m || fl[j + 1]);
            ac[j] = (single || done(sp[j], h)) ? 0 : 1;
        });
        DevBuf<u32> scratch(m);
        ma = scan_flags(act.get(), scratch.get(), m, ws);
        apos.alloc(ma);
        if (ma) {
            const u32 *ix = scratch.get(); u32 *ap = apos.get();
            launch_for(m, st, PFB_LAMBDA(u64 j) { if (ac[j]) ap[ix[j]] = (u32)j; });
        }
    }
    while (ma) {
        if (rounds > 64) { err = "prefix doubling did not converge"; return PFB_ERR_INTERNAL; }
        // a group is active as a whole (done() is uniform over a tie group): the first listed member of a group is its head
        DevBuf<u8> ah(ma);
        DevBuf<u32> gidx(ma);
        {
            const u32 *ap = apos.get(); const u8 *fl = flag.get(); u8 *a_ = ah.get();
            launch_for(ma, st, PFB_LAMBDA(u64 a) { a_[a] = fl[ap[a]] ? 1 : 0; });
        }
        const u64 G = scan_flags(ah.get(), gidx.get(), ma, ws);
        const int gbits = bits_for(G ? G - 1 : 0);
        if (gbits + rbits > 64) { err = "prefix doubling: key overflow"; return PFB_ERR_INTERNAL; }
        DevBuf<u32> aval(ma), aval_b(ma), hslot(G + 1);
        DevBuf<u64> akey(ma), akey_b(ma);
        {
            const u32 *ap = apos.get(); const u8 *a_ = ah.get(); const u32 *gi = gidx.get(); const u32 *sp = sa.get(); const u32 *rk = rank.get();
            u32 *av = aval.get(); u64 *ak = akey.get(); u32 *hs = hslot.get();
            const u32 ma32 = (u32)ma; const u64 G_ = G;
            launch_for(ma + 1, st, PFB_LAMBDA(u64 a) {
                if (a == ma32) { hs[G_] = ma32; return; }
                const u64 gid = (u64)gi[a] + a_[a] - 1;
                if (a_[a]) hs[gid] = (u32)a;
                const u64 i = sp[ap[a]];
                const u64 nx = i + h;
                const u64 r2 = nx < m ? (u64)rk[nx] : 0;
                av[a] = (u32)i;
                ak[a] = (gid << rbits) | r2;
            });
        }
        DevBuf<u8> hl(ma);
        order_groups(akey, akey_b, aval, aval_b, hl, ma, hslot.get(), rbits, 0, gbits + rbits, ws);
        DevBuf<u32> hd(ma), hd2(ma);
        {
            const u32 *ap = apos.get(); const u32 *av = aval.get(); const u8 *hlp = hl.get();
            u32 *sp = sa.get(); u8 *fl = flag.get(); u32 *hp = hd.get();
            launch_for(ma, st, PFB_LAMBDA(u64 a) {
                u32 j = ap[a];
                sp[j] = av[a];
                bool head = (hlp[a] & 1) != 0;
                fl[j] = head ? 1 : 0;
                hp[a] = head ? j : 0;
            });
        }
        inclusive_max(hd.get(), hd2.get(), ma, ws.sc, st);
        h *= 2;
        ++rounds;
        // new ranks; what stays open: members of groups that are still tied and not complete within the doubled depth
        DevBuf<u8> keep(ma);
        {
            const u32 *av = aval.get(); const u32 *hp = hd2.get(); const u8 *hlp = hl.get(); u32 *rk = rank.get(); u8 *kp = keep.get();
            const u64 hn = h;
            launch_for(ma, st, PFB_LAMBDA(u64 a) {
                rk[av[a]] = hp[a];
                kp[a] = ((hlp[a] & 3) == 3 || done(av[a], hn)) ? 0 : 1;
            });
        }
        DevBuf<u32> kidx(ma);
        const u64 m2 = scan_flags(keep.get(), kidx.get(), ma, ws);
        DevBuf<u32> apos2(m2);
        if (m2) {
            const u32 *ap = apos.get(); const u8 *kp = keep.get(); const u32 *ki = kidx.get(); u32 *a2 = apos2.get();
            launch_for(ma, st, PFB_LAMBDA(u64 a) { if (kp[a]) a2[ki[a]] = ap[a]; });
        }
        apos = std::move(apos2);
        ma = m2;
    }
    if (rounds_out) *rounds_out = rounds;
    return PFB_OK;
}


// heads/ranks from sorted keys: flag[j] = key differs from predecessor,
// rank[sa[j]] = position of the group head
inline void ranks_from_flags(const DevBuf<u32> &sa, DevBuf<u8> &flag, DevBuf<u32> &rank, u64 m, Workspace &ws) {
    pfbStream_t st = ws.st;
    DevBuf<u32> hd(m), hd2(m);
    {
        const u8 *fl = flag.get(); u32 *hp = hd.get();
        launch_for(m, st, PFB_LAMBDA(u64 j) { hp[j] = fl[j] ? (u32)j : 0; });
    }
    inclusive_max(hd.get(), hd2.get(), m, ws.sc, st);
    {
        const u32 *sp = sa.get(); const u32 *hp = hd2.get(); u32 *rk = rank.get();
        launch_for(m, st, PFB_LAMBDA(u64 j) { rk[sp[j]] = hp[j]; });
    }
}

// largest d with wstart[d] <= pos   (wstart has Dw+1 entries, wstart[Dw] = dsize-1)
PFB_HD inline u32 word_of(const u64 *wstart, u32 Dw, u64 pos) {
    u32 lo = 0, hi = Dw;   // answer in [lo, hi)
    while (hi - lo > 1) {
        u32 mid = lo + (hi - lo) / 2;
        if (wstart[mid] <= pos) lo = mid; else hi = mid;
    }
    return lo;
}

// ===========================================================================
// parse_text: T' bytes already on the device and normalised
// ===========================================================================
struct TextOnDevice {
    DevBuf<u8> buf;     // 64 front room | S (n) | Dollar^w | tail room   (a rank of a multi-GPU job puts the w+1 bytes
    u64 n = 0;          //  before its records into the front room and the bytes up to the next trigger into the tail room)
    u64 tail_room = 0;
    static const u64 FRONT = 64;
    u8 *Tp() const { return buf.get() + FRONT - 1; }   // T'[0] (the leading Dollar)
    u8 *S() const { return buf.get() + FRONT; }
};

// A piece of T' (multi-GPU parse): bytes Tp[0, tlen); positions qmin.. may trigger; the piece either
// ends with Dollar^w (end of T') or exactly at a trigger position (the first trigger owned by a later rank).
static const u64 DICT_MAX_BYTES = 1ull << 40;      // dictionary positions are u32 below 2^32 bytes and u64 above (build.cuh)
struct PieceSpec {
    u64 tlen = 0, qmin = 0; bool dollar_end = true;
    const u8 *mem_lo = nullptr, *mem_hi = nullptr;      // the allocation Tp lies in (the tiled trigger scan loads whole aligned tiles)
};

inline int parse_piece(const u8 *Tp, const PieceSpec &ps, const Params &prm, DeviceParse &out, Workspace &ws, StageTimes &tm, std::string &err) {
    pfbStream_t st = ws.st;
    const int w = prm.w;
    const u64 tlen = ps.tlen;
    const bool dollar_end = ps.dollar_end;
    const u64 n = dollar_end ? tlen - (u64)w - 1 : tlen - 1;      // last position that may trigger
    if (w < 1 || w > 31) { err = "window size w must be < 32!"; return PFB_ERR_ARG; }
    if (prm.p < 1) { err = "modulus p must be >= 1"; return PFB_ERR_ARG; }
    Timer timer(st);

    // K1 trigger bitmap ------------------------------------------------------
    const u64 nwords = n / 32 + 1;   // bits 0..n
    DevBuf<u32> trig(nwords), cnt(nwords), wordoff(nwords);
    bool tiled = false;
#ifndef PFB_EMU
    if (ps.mem_lo) tiled = k_trigger_tiled(Tp, n, w, prm.p, trig.get(), cnt.get(), nwords, ps.qmin, ps.mem_lo, ps.mem_hi, ws.lut, st);
#endif
    if (!tiled) k_trigger_words(Tp, n, w, DivTest::make(prm.p), trig.get(), cnt.get(), nwords, st, ps.qmin);
    tm.trigger = timer.stop();

    // K2 compaction ------------------------------------------------------------
    exclusive_sum<u32, u32>(cnt.get(), wordoff.get(), nwords, ws.sc, st);
    u64 ntrig = (u64)read_scalar(wordoff.get() + (nwords - 1), st) + read_scalar(cnt.get() + (nwords - 1), st);
    const u64 P = dollar_end ? ntrig + 1 : ntrig;
    if (P == 0) { err = "internal: piece without a phrase"; return PFB_ERR_INTERNAL; }
    if (prm.bits == 32 && P > 0x7FFFFFFEull) { err = "Input containing more than 2^31-2 phrases! Please use 64 bit version"; return PFB_ERR_LIMIT; }
    if (P > 0xFFFFFFFEull) { err = "Input containing more than 2^32-2 phrases! This is currently a hard limit"; return PFB_ERR_LIMIT; }
    cnt.release();
    out.sai.alloc(P);   // sai_[k] == end position of phrase k (include/pfparser.hpp:600)
    u64 *ends = out.sai.get();
    k_scatter_ends(trig.get(), wordoff.get(), nwords, ends, P, n, w, st, dollar_end);
    trig.release();
    out.last.alloc(P);
    {
        u8 *last = out.last.get();
        launch_for(P, st, PFB_LAMBDA(u64 k) { last[k] = Tp[ends[k] - (u64)w]; });   // include/pfparser.hpp:599
    }
    tm.compact = timer.stop();

    // K3 fingerprints ------------------------------------------------------------
    DevBuf<u64> fp1(P), fp2(P);
    dev_memset(fp1.get(), 0, P * 8, st);
    dev_memset(fp2.get(), 0, P * 8, st);
    k_fingerprint(Tp, tlen, ends, P, w, wordoff.get(), nwords, fp1.get(), fp2.get(), st);
    tm.fingerprint = timer.stop();

    // K4 dedup -------------------------------------------------------------------
    u64 tsize = 1024;
    while (tsize < 2 * P) tsize <<= 1;
    DevBuf<u32> rep(P);
    {
        DevBuf<u64> tkey(tsize);
        DevBuf<u32> trep(tsize), slot(P), flagbuf(1);
        u32 attempt = 0;
        for (;; ++attempt) {
            if (attempt >= 8) { err = "phrase fingerprint collisions could not be resolved"; return PFB_ERR_INTERNAL; }
            dev_memset(tkey.get(), 0, tsize * 8, st);
            dev_memset(trep.get(), 0xff, tsize * 4, st);
            dev_memset(flagbuf.get(), 0, 4, st);
            k_ht_insert(fp1.get(), fp2.get(), ends, P, w, attempt, tkey.get(), trep.get(), tsize - 1, slot.get(), st);
            {
                const u32 *sl = slot.get(); const u32 *tr = trep.get(); u32 *rp = rep.get();
                launch_for(P, st, PFB_LAMBDA(u64 k) { rp[k] = tr[sl[k]]; });
            }
            k_verify(Tp, tlen, ends, P, w, wordoff.get(), nwords, rep.get(), flagbuf.get(), st);
            if (read_scalar(flagbuf.get(), st) == 0) break;
        }
        tm.dedup_attempts = attempt + 1;
    }
    fp1.release(); fp2.release(); wordoff.release();
    // distinct phrases (representatives) and their occurrence counts
    DevBuf<u32> cntk(P);
    dev_memset(cntk.get(), 0, P * 4, st);
    DevBuf<u32> order;
    u64 Dw;
    {
        DevBuf<u8> isrep(P);
        DevBuf<u32> didx(P);
        const u32 *rp = rep.get(); u8 *ir = isrep.get(); u32 *ck = cntk.get();
        launch_for(P, st, PFB_LAMBDA(u64 k) {
            ir[k] = rp[k] == (u32)k ? 1 : 0;
            atomic_add(&ck[rp[k]], 1u);
        });
        Dw = scan_flags(isrep.get(), didx.get(), P, ws);
        order.alloc(Dw);
        u32 *ord = order.get(); const u32 *di = didx.get();
        launch_for(P, st, PFB_LAMBDA(u64 k) { if (ir[k]) ord[di[k]] = (u32)k; });
    }
    tm.dedup = timer.stop();

    // K5a phrase sort --------------------------------------------------------------
    int rc = sort_phrases(Tp, ends, w, order, Dw, ws, &tm.phrase_sort_rounds, err);
    if (rc) return rc;
    tm.phrase_sort = timer.stop();

    // ranks, occ, dict image ---------------------------------------------------------
    out.n = n; out.P = P; out.Dw = Dw;
    out.parse.alloc(P + 1);
    out.occ.alloc(Dw);
    {
        DevBuf<u32> rankc(P);
        u32 *rc_ = rankc.get(); const u32 *ord = order.get(); const u32 *ck = cntk.get(); u32 *oc = out.occ.get();
        launch_for(Dw, st, PFB_LAMBDA(u64 r) { rc_[ord[r]] = (u32)r + 1; oc[r] = ck[ord[r]]; });
        u32 *pr = out.parse.get(); const u32 *rp = rep.get();
        launch_for(P + 1, st, PFB_LAMBDA(u64 k) { pr[k] = k < P ? rc_[rp[k]] : 0u; });   // include/pfparser.hpp:504-517
    }
    rep.release(); cntk.release();
    DevBuf<u64> wl(Dw + 1), ws64(Dw + 1);
    {
        u64 *wlp = wl.get(); const u32 *ord = order.get();
        launch_for(Dw + 1, st, PFB_LAMBDA(u64 r) {
            if (r == Dw) { wlp[r] = 0; return; }
            u64 k = ord[r];
            wlp[r] = ends[k] - phrase_start(ends, k, w) + 2;   // phrase + EndOfWord
        });
    }
    exclusive_sum<u64, u64>(wl.get(), ws64.get(), Dw + 1, ws.sc, st);
    u64 dsize = read_scalar(ws64.get() + Dw, st) + 1;
    if (dsize >= DICT_MAX_BYTES) { err = "dictionary larger than 2^40 bytes is not supported"; return PFB_ERR_LIMIT; }
    out.dsize = dsize;
    out.wstart = std::move(ws64);
    out.dict.alloc(dsize);
    {
        const u64 *w64 = out.wstart.get();
        u8 *dict = out.dict.get(); const u32 *ord = order.get();
        // the words are copied into the image (include/pfbwt_io.hpp:71-82); the few words longer than
        // LONGW (phrases spanning an N run: up to tens of Mbp) are listed and copied by the copy engine instead
        const u64 LONGW = 1u << 13;
        const u64 lcap = tlen / LONGW + 2;
        DevBuf<u64> longw(3 * lcap);
        DevBuf<u32> nlong(1);
        dev_memset(nlong.get(), 0, 4, st);
        u64 *lw = longw.get(); u32 *nl = nlong.get();
        // 8 lanes per word, 8 bytes per lane and step: aligned 8-byte stores into the image, the source read with load8_any
        // (phrases start anywhere); the ragged head / tail of a word byte by byte
        launch_for(Dw * 8, st, PFB_LAMBDA(u64 i) {
            u64 r = i / 8, lane = i % 8;
            u64 k = ord[r];
            u64 s = phrase_start(ends, k, w), len = ends[k] - s + 1;
            u64 d0 = w64[r];
            if (len <= LONGW) {
                u64 head = (8 - (d0 & 7)) & 7;
                if (head > len) head = len;
                const u64 nfull = (len - head) >> 3;
                if (lane == 0) for (u64 j = 0; j < head; ++j) dict[d0 + j] = Tp[s + j];
                for (u64 j = lane; j < nfull; j += 8) *reinterpret_cast<u64 *>(dict + d0 + head + 8 * j) = load8_any(Tp + s + head + 8 * j);
                if (lane == 7) for (u64 j = head + 8 * nfull; j < len; ++j) dict[d0 + j] = Tp[s + j];
            } else if (lane == 0) { u64 q = atomic_add(nl, 1u); lw[3 * q] = s; lw[3 * q + 1] = d0; lw[3 * q + 2] = len; }
            if (lane == 0) dict[d0 + len] = ENDOFWORD;
            if (i == 0) dict[dsize - 1] = ENDOFDICT;
        });
        const u32 nlw = read_scalar(nl, st);
        if (nlw) {
            std::vector<u64> hl(3 * (size_t)nlw);
            d2h(hl.data(), lw, hl.size() * 8, st);
            stream_sync(st);
            for (u32 q = 0; q < nlw; ++q) d2d(dict + hl[3 * q + 1], Tp + hl[3 * q], hl[3 * q + 2], st);
        }
    }
    tm.dict_build = timer.stop();
    return PFB_OK;
}

inline int parse_text(const TextOnDevice &T, const Params &prm, DeviceParse &out, Workspace &ws, StageTimes &tm, std::string &err) {
    if (T.n == 0) { err = "empty input"; return PFB_ERR_ARG; }
    PieceSpec ps;
    ps.tlen = T.n + (u64)prm.w + 1; ps.qmin = (u64)prm.w + 1; ps.dollar_end = true;
    ps.mem_lo = T.buf.get(); ps.mem_hi = T.buf.get() + T.buf.size();
    return parse_piece(T.Tp(), ps, prm, out, ws, tm, err);
}

// ===========================================================================
// parse_bwt: SA of parse.0, then bwlast / bwsai / ilist (include/pfparser.hpp:379-467)
// ===========================================================================
inline int parse_bwt(DeviceParse &pp, const Params &prm, Workspace &ws, StageTimes &tm, std::string &err) {
    pfbStream_t st = ws.st;
    const u64 P = pp.P, Dw = pp.Dw;
    if (P == 1) { err = "error: only one dict word total. Re-run with a smaller p modulus"; return PFB_ERR_ONE_WORD; }
    const u64 m = P + 1;
    Timer timer(st);
    DevBuf<u32> sa(m), rank(m);
    DevBuf<u8> flag(m);
    {
        // initial order on two symbols: key = R[i] << sbits | R[i+1]
        const int sbits = bits_for(Dw);
        DevBuf<u64> key_a(m), key_b(m);
        DevBuf<u32> val_b(m);
        const u32 *R = pp.parse.get(); u64 *ka = key_a.get(); u32 *sp = sa.get();
        launch_for(m, st, PFB_LAMBDA(u64 i) {
            u64 a = R[i], b = i + 1 < m ? R[i + 1] : 0;
            ka[i] = (a << sbits) | b;
            sp[i] = (u32)i;
        });
        sort_pairs(key_a, key_b, sa, val_b, m, 0, 2 * sbits, ws.sc, st);
        const u64 *ks = key_a.get(); u8 *fl = flag.get();
        launch_for(m, st, PFB_LAMBDA(u64 j) { fl[j] = (j == 0 || ks[j] != ks[j - 1]) ? 1 : 0; });
    }
    ranks_from_flags(sa, flag, rank, m, ws);
    int rc = prefix_doubling(sa, rank, flag, m, 2, PFB_LAMBDA(u32, u64) { return false; }, ws, &tm.parse_sa_rounds, err);
    if (rc) return rc;
    rank.release(); flag.release();
    tm.parse_sa = timer.stop();

    pp.bwlast.alloc(m);
    pp.bwsai.alloc(m);
    pp.ilist.alloc(m);
    {
        DevBuf<u64> bwtp(m), bwtp_b(m);          // 64-bit keys: the onesweep sort of radix.cuh takes (u64, u32) pairs
        DevBuf<u32> il_b(m);
        const u32 *sp = sa.get(); const u32 *R = pp.parse.get(); const u8 *last = pp.last.get(); const u64 *sai = pp.sai.get();
        u8 *bl = pp.bwlast.get(); u64 *bs = pp.bwsai.get(); u64 *bp = bwtp.get(); u32 *il = pp.ilist.get();
        launch_for(m, st, PFB_LAMBDA(u64 i) {     // include/pfparser.hpp:429-451
            u64 s = sp[i];
            il[i] = (u32)i;
            if (s == 0) { bp[i] = 0; bl[i] = 0; bs[i] = 0; }
            else {
                bl[i] = s == 1 ? last[P - 1] : last[s - 2];
                bs[i] = sai[s - 1];
                bp[i] = R[s - 1];
            }
        });
        // ilist = positions grouped by BWT(P) symbol, ascending inside a symbol
        // (the F[]-driven counting sort of include/pfparser.hpp:452-462) = stable sort by symbol
        sort_pairs(bwtp, bwtp_b, pp.ilist, il_b, m, 0, bits_for(Dw), ws.sc, st);
    }
    pp.has_parse_bwt = true;
    tm.parse_bwt = timer.stop();
    return PFB_OK;
}

}  // namespace pfb

#include "build.cuh"
