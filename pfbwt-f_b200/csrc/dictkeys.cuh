// dictkeys.cuh -- sort keys of the dictionary suffixes.
//
// A key holds an ORDER-PRESERVING ENUMERATIVE CODE of the first L symbols of a suffix.
// The dictionary alphabet has 8 byte values (0x01<0x02<'-'<'A'<'C'<'G'<'N'<'T') but almost
// every symbol is one of A,C,G,T, so instead of 3 bits per symbol the window is numbered
// among the strings that can actually be told apart by one key:
//   * EndOfWord / EndOfDict ends the string                      (closed: nothing follows)
//   * 0x02, '-' and 'N' are "stoppers": the key stops after them (open: more symbols follow)
//   * A,C,G,T continue
// With N_0 = 1, N_L = 4 + 4 N_{L-1} = (7*4^L - 4)/3 such windows of length L, the code is
//   V_L(x) = 0 | 1 | 2                       x0 = EndOfWord | 0x02 | '-'
//          = 3 + a*N_{L-1} + [a==3] + V_{L-1}(x1..)   x0 = "ACGT"[a]
//          = 3 + 3 N_{L-1}                   x0 = 'N'
// which for a window of j leading ACGT symbols (2-bit digits a_i, B = sum a_i 4^(L-1-i),
// S = sum a_i, nT = #{a_i == 3}) followed by a stopper has the closed form
//   V = 3j + nT + (7B - 4S)/3 + stop,   stop = 0,1,2 or 7*4^(L-1-j) - 1 for 'N'.
// 2.06 bits per symbol instead of 3: 19 symbols fit 40 bits (5 radix passes instead of 8), and
// a 54-bit field holds 26 symbols.  Equal codes = identical windows (same stopper, same
// position); two suffixes tied on a closed window are identical strings (a group of
// include/pfbwt.hpp:136-145), tied on an open window they are refined further.
//
// key layout (u64):  [0,4) code of the preceding byte   [4] closed   [5,10) symbols consumed
//                    [10, 10+bits) enumerative code
#pragma once

namespace pfb {

// 3-bit order-preserving code of a dictionary byte without a branch:
// perfect hash ((c*7)>>3)&15 into a nibble table (same mapping as sym3()).
PFB_HD inline u32 sym3_fast(u32 c) { return (u32)(0x0500047320060010ull >> (4 * (((c * 7u) >> 3) & 15u))) & 7u; }
// code -> byte (code 0 -> 0x00: the byte "before" S[0], include/pfbwt.hpp:132,140)
PFB_HD inline u8 sym3_byte(u32 code) { return (u8)(0x544E4743412D0200ull >> (8 * code)); }
// sym3 code -> 2-bit field (ACGT digit, or stopper kind 0 EndOfWord, 1 0x02, 2 '-', 3 'N') and "is a stopper"
PFB_HD inline u32 sym_field(u32 s) { return (0x33210210u >> (4 * s)) & 3u; }
PFB_HD inline u32 sym_stops(u32 s) { return (0x47u >> s) & 1u; }

static const int KEY_CODE_SHIFT = 10;
static const int KEY_MAX_SYMS = 26;     // N_26 < 2^54

PFB_HD inline u64 enum_count(int L) { return (7ull * (1ull << (2 * L)) - 4ull) / 3ull; }   // N_L
inline int enum_bits(int L) { return bits_for(enum_count(L) - 1); }

// position -> (word, suffix length) through the sample table.  DP = type of a dictionary position: u32 while the
// dictionary has fewer than 2^32 bytes, u64 beyond (pfbwt-f64 indexes the dictionary with 64-bit uint_t, gsa/gsacak.h:44-58)
template <class DP> struct WordLookupT;
template <> struct WordLookupT<u32> {
    const u64 *wc;      // Dw+2 entries, low 32 bits = word start; wc[Dw] = D-1, wc[Dw+1] = 2^32-1
    const u32 *wsamp;
    u32 Dw;
    PFB_HD inline u32 word(u32 idx) const {
        u32 d = wsamp[idx >> 7];
        while ((u32)wc[d + 1] <= idx) ++d;
        return d;
    }
    PFB_HD inline u32 start(u32 d) const { return (u32)wc[d]; }
    PFB_HD inline u32 cum(u32 d) const { return (u32)(wc[d] >> 32); }
    // length of the suffix starting at idx inside word d, EndOfWord excluded (0 for the EndOfDict slot)
    PFB_HD inline u64 suflen(u32 d, u32 idx) const { return d < Dw ? (u64)((u32)wc[d + 1] - 1 - idx) : 0; }
};
template <> struct WordLookupT<u64> {
    const u64 *ws;      // Dw+2 word starts; ws[Dw] = D-1, ws[Dw+1] = 2^64-1
    const u32 *cm;      // Dw+2 first ilist slots - 1
    const u32 *wsamp;
    u32 Dw;
    PFB_HD inline u32 word(u64 idx) const {
        u32 d = wsamp[idx >> 7];
        while (ws[d + 1] <= idx) ++d;
        return d;
    }
    PFB_HD inline u64 start(u32 d) const { return ws[d]; }
    PFB_HD inline u32 cum(u32 d) const { return cm[d]; }
    PFB_HD inline u64 suflen(u32 d, u64 idx) const { return d < Dw ? ws[d + 1] - 1 - idx : 0; }
};
typedef WordLookupT<u32> WordLookup;

// The bits of a key above the code carry what the assembly needs to know about the suffix, so that
// the sorted order can be classified without a random position -> word lookup per suffix:
//   bit 63            full: the suffix is a whole dictionary word
//   [cshift, 63)      occurrences it emits (occ of its word; 0 = suffix length <= w, all ones = look it up)
struct KeyExtra {
    int cshift = 63;      // first bit of the count field (10 + code bits)
    u64 cmax = 0;         // all-ones value of the field (0: field absent)
    PFB_HD inline u64 pack(u64 cnt, bool full) const {
        if (!cmax) return 0;          // no room above the code: classification falls back to lookups
        u64 c = cnt < cmax ? cnt : cmax;
        return (c << cshift) | (full ? 1ull << 63 : 0ull);
    }
    PFB_HD inline u64 count(u64 key) const { return (key >> cshift) & cmax; }
    u64 cmask = ~0ull;    // mask of the code after shifting the key right by KEY_CODE_SHIFT
    PFB_HD inline u64 code(u64 key) const { return (key >> 10) & cmask; }
    PFB_HD inline bool full(u64 key) const { return (key >> 63) != 0; }
};
inline KeyExtra key_extra_for(int L) {
    KeyExtra x;
    int top = 10 + enum_bits(L);
    x.cmask = (1ull << enum_bits(L)) - 1;
    int room = 63 - top;
    if (const char *e = getenv("PFB_KEY_COUNT_BITS")) { int r = atoi(e); if (r < room) room = r; }   // test knob: force count overflow / absence
    if (room >= 1) { if (room > 24) room = 24; x.cshift = top; x.cmax = (1ull << room) - 1; }
    return x;
}


// B: the j leading ACGT digits left-aligned in a field of L digits (zero below); kind: stopper when j < L
PFB_HD inline u64 enum_code(u64 B, int j, int L, u32 kind) {
    const u64 lo = 0x5555555555555555ull;
    u64 S = (u64)popc64(B & lo) + 2ull * (u64)popc64(B & (lo << 1));
    u64 nT = (u64)popc64(B & (B >> 1) & lo);
    u64 v = 3ull * (u64)j + nT + (7ull * B - 4ull * S) * 0xAAAAAAAAAAAAAAABull;   // exact division by 3
    if (j < L) v += kind < 3 ? (u64)kind : (7ull << (2 * (L - 1 - j))) - 1ull;
    return v;
}

PFB_HD inline u64 pack_key(u64 code, u32 consumed, bool closed, u32 pc) {
    return (code << KEY_CODE_SHIFT) | ((u64)consumed << 5) | (closed ? 16ull : 0ull) | (u64)pc;
}
PFB_HD inline bool key_closed(u64 k) { return (k >> 4) & 1; }
PFB_HD inline u32 key_consumed(u64 k) { return (u32)(k >> 5) & 31u; }

// byte-by-byte form (refinement rounds, emulator): window of L symbols starting at dict[pos]
struct EnumKey { u64 code; u32 consumed; bool closed; };
PFB_HD inline EnumKey enum_scan(const u8 *dict, u64 D, u64 pos, int L) {
    u64 B = 0;
    int j = 0;
    u32 kind = 0;
    for (; j < L; ++j) {
        u64 o = pos + (u64)j;
        u32 s = o < D ? sym3_fast(dict[o]) : 0u;
        if (sym_stops(s)) { kind = sym_field(s); break; }
        B = (B << 2) | sym_field(s);
    }
    EnumKey r;
    if (j < L) { B <<= 2 * (L - j); r.consumed = (u32)j + 1; r.closed = kind == 0; }
    else { r.consumed = (u32)L; r.closed = false; }
    r.code = enum_code(B, j, L, kind);
    return r;
}

// code of the byte preceding dictionary position idx (0 for idx <= 1, include/pfbwt.hpp:132,140;
// 8 = EndOfWord precedes, i.e. idx is a word start)
PFB_HD inline u64 prev_code(const u8 *dict, u64 idx) {
    if (idx <= 1) return 0;
    u64 c = sym3_fast(dict[idx - 1]);
    return c ? c : 8;
}
// the byte a 4-bit preceding-byte code stands for.  A word start inside a group of equal suffixes
// is emitted as EndOfWord (0x01), exactly like dict[gsa[j]-1] in include/pfbwt.hpp:140
PFB_HD inline u8 prev_byte(u64 key) { return (key & 8) ? ENDOFWORD : sym3_byte((u32)key & 7); }

static const int KEYS_TILE = 2048;
// K6a: one CTA = 2048 consecutive dictionary positions, 8 per thread.  The tile (+40 bytes of
// look-ahead) is staged in shared memory with 16-byte loads; each thread packs its 40 symbols
// once into 2-bit fields + a stopper mask and derives its 8 windows with shifts; keys leave
// through shared memory so that global stores are unit-stride.
// MODE 0: keys of all positions -> key[].  The other modes serve the range-restricted sort of the
// multi-GPU path (a rank sorts only the valid suffixes whose code lies in [lo, hi)):
// MODE 1: weighted histogram of the top `hshift`-shifted code bits over sampled tiles (splitter choice)
// MODE 2: number of selected suffixes per tile -> tilecnt[]
// MODE 3: selected (key, position) pairs, in dictionary order, at tilebase[tile] + rank inside the tile
// MODE 4 / 5: the position-sliced form of the multi-GPU path - a rank scans only ITS slice of the dictionary positions
//   (tiles [tile0, tile0 + ntl)) and routes every valid suffix to the rank that owns its code range (bound[d] <= code <
//   bound[d+1]); 4 counts per (destination, tile), 5 writes the pairs grouped by destination, dictionary order inside
static const int KEY_MAX_DEST = 64;
template <class DP>
struct KeySelectT {
    u64 lo = 0, hi = ~0ull;       // code range
    u64 *hist = nullptr; int hshift = 0; u32 tile_stride = 1;
    u32 *tilecnt = nullptr; const u64 *tilebase = nullptr;
    u64 *key_out = nullptr; DP *idx_out = nullptr;
    u64 tile0 = 0, ntl = 0;       // modes 4 / 5: first tile and number of tiles of this rank's position slice
    int nd = 0;                   // destinations
    u64 bound[KEY_MAX_DEST + 1];  // bound[0] = 0 <= ... <= bound[nd] = ~0
    // mode 5: where destination d's pairs go - a staging buffer of this rank, or the receive buffer of rank d itself mapped
    // over NVLink (the part this rank contributes) - and tilebase[d * ntl] (pairs for destinations before d)
    u64 *key_outs[KEY_MAX_DEST]; DP *idx_outs[KEY_MAX_DEST]; u64 destbase[KEY_MAX_DEST];
};
typedef KeySelectT<u32> KeySelect;
PFB_HD inline int key_dest(const u64 *bound, int nd, u64 code) {      // largest d with bound[d] <= code
    int lo = 0, hi = nd;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (bound[mid] <= code) lo = mid; else hi = mid; }
    return lo;
}
// 8 dictionary bytes (little endian: symbol 0 in the low byte) -> [0,16) their 2-bit fields, [16,24) "is a stopper",
// [24,32) "is EndOfWord / EndOfDict", symbol 0 in the most significant position of each part.  Word-wide arithmetic on the
// 8 byte values of the alphabet: field = ((c >> 1) & 3) ^ bit1-of-that ^ bit 3 of c (ACGT -> 0..3; 0x00/0x01 -> 0, 0x02 -> 1,
// '-' -> 2, 'N' -> 3 = sym_field(sym3(c))), stopper = bit 6 clear or bit 3 set, end = c < 2; bits gathered by multiplication.
PFB_HD inline u32 key_pack8(u64 x) {
    // byte swap: symbol 0 to the most significant byte, so that the gathers below keep the order
    u64 z = ((x & 0x00ff00ff00ff00ffull) << 8) | ((x >> 8) & 0x00ff00ff00ff00ffull);
    z = ((z & 0x0000ffff0000ffffull) << 16) | ((z >> 16) & 0x0000ffff0000ffffull);
    z = (z << 32) | (z >> 32);
    const u64 ONE = 0x0101010101010101ull;
    u64 y = (z >> 1) & (3 * ONE);
    y ^= (y >> 1) & ONE;
    y ^= (z >> 3) & ONE;
    u64 a = (y | (y >> 6)) & 0x000F000F000F000Full;          // stride 8 -> 4 -> 2 bits
    a = (a | (a >> 12)) & 0x000000FF000000FFull;
    a = (a | (a >> 24)) & 0xFFFFull;
    const u64 st = ((~z >> 6) | (z >> 3)) & ONE;
    const u32 stops = (u32)((st * 0x0102040810204080ull) >> 56);
    const u64 v = z & (0xFEull * ONE);
    const u64 e = ~((((v & (0x7Full * ONE)) + 0x7Full * ONE) | v) | (0x7Full * ONE));     // 0x80 in every byte that is < 2
    const u32 ends = (u32)(((e >> 7) * 0x0102040810204080ull) >> 56);
    return (u32)a | (stops << 16) | (ends << 24);
}
#ifndef PFB_EMU
template <int MODE, class DP>
__global__ void __launch_bounds__(256) k_suffix_keys_kernel(const u8 *__restrict__ dict, u64 D, int L, u64 *__restrict__ key,
                                                            WordLookupT<DP> wl, const u32 *__restrict__ occ, int w, KeyExtra kx, KeySelectT<DP> ks) {
    __shared__ __align__(16) u8 sb[KEYS_TILE + 48];
    __shared__ u64 sk[KEYS_TILE + KEYS_TILE / 8];      // one pad slot per 8 keys: a thread's 8 consecutive keys spread over banks
    const u64 tile = MODE == 1 ? (u64)blockIdx.x * ks.tile_stride : MODE >= 4 ? ks.tile0 + (u64)blockIdx.x : (u64)blockIdx.x;
    const u64 base = tile * KEYS_TILE;
    for (int o = threadIdx.x * 16; o < KEYS_TILE + 48; o += 256 * 16) {
        u64 g = base + (u64)o;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (g + 16 <= D) v = *reinterpret_cast<const uint4 *>(dict + g);
        else if (g < D) {
            u32 t[4] = {0, 0, 0, 0};
            for (int b = 0; b < 16 && g + b < D; ++b) t[b >> 2] |= (u32)dict[g + b] << (8 * (b & 3));
            v = make_uint4(t[0], t[1], t[2], t[3]);
        }
        *reinterpret_cast<uint4 *>(sb + o) = v;
    }
    __shared__ u32 sprev;
    __shared__ u32 spk[KEYS_TILE / 8 + 8];
    if (threadIdx.x == 0) sprev = base ? (u32)dict[base - 1] : 0u;
    __syncthreads();
    // every 8 bytes of the tile are packed ONCE (word-wide arithmetic, key_pack8) and shared; a thread then assembles
    // the 40-symbol window of its 8 positions from five packed words instead of packing 40 symbols itself
    for (int q = threadIdx.x; q < KEYS_TILE / 8 + 6; q += 256) spk[q] = key_pack8(*reinterpret_cast<const u64 *>(sb + 8 * q));
    __syncthreads();
    const int i0 = threadIdx.x * 8;
    u64 xh, xl, rm, tmk;                    // 2-bit fields of symbols 0..31 / 32..39, stopper mask, EndOfWord / EndOfDict mask (MSB first)
    {
        const u32 w0 = spk[threadIdx.x], w1 = spk[threadIdx.x + 1], w2 = spk[threadIdx.x + 2], w3 = spk[threadIdx.x + 3], w4 = spk[threadIdx.x + 4];
        xh = ((u64)(w0 & 0xffffu) << 48) | ((u64)(w1 & 0xffffu) << 32) | ((u64)(w2 & 0xffffu) << 16) | (u64)(w3 & 0xffffu);
        xl = (u64)(w4 & 0xffffu) << 48;
        rm = ((u64)((w0 >> 16) & 0xffu) << 56) | ((u64)((w1 >> 16) & 0xffu) << 48) | ((u64)((w2 >> 16) & 0xffu) << 40) |
             ((u64)((w3 >> 16) & 0xffu) << 32) | ((u64)((w4 >> 16) & 0xffu) << 24);
        tmk = ((u64)(w0 >> 24) << 56) | ((u64)(w1 >> 24) << 48) | ((u64)(w2 >> 24) << 40) | ((u64)(w3 >> 24) << 32) | ((u64)(w4 >> 24) << 24);
    }
    u32 pcs[8];
    pcs[0] = sym3_fast(i0 ? (u32)sb[i0 - 1] : sprev);
    {
        const u64 wd = *reinterpret_cast<const u64 *>(sb + i0);
#pragma unroll
        for (int k = 0; k < 7; ++k) pcs[k + 1] = sym3_fast((u32)(wd >> (8 * k)) & 0xffu);
    }
    // word of the thread's first position (one sampled lookup per 8 positions); later positions add
    // the EndOfWord bytes passed
    u32 d = (base + (u64)i0 < D) ? wl.word((DP)(base + (u64)i0)) : wl.Dw;
    u32 selmask = 0;
    u64 dstv = ~0ull;                       // modes 4 / 5: destination of each of the 8 positions, one byte each (0xff = none)
    u64 mykeys[(MODE == 3 || MODE == 5) ? 8 : 1];
#pragma unroll
    for (int p = 0; p < 8; ++p) {
        u64 W = p ? ((xh << (2 * p)) | (xl >> (64 - 2 * p))) : xh;
        u64 Xw = W >> (64 - 2 * L);
        u64 Rw = (rm << p) >> (64 - L);
        int j = Rw ? L - 1 - (63 - __clzll((long long)Rw)) : L;
        u32 kind = 0;
        u64 B = Xw;
        if (j < L) { kind = (u32)(Xw >> (2 * (L - 1 - j))) & 3u; B = (Xw >> (2 * (L - j))) << (2 * (L - j)); }
        u64 code = enum_code(B, j, L, kind);
        u64 i = base + (u64)(i0 + p);
        u32 pc = i <= 1 ? 0u : (pcs[p] ? pcs[p] : 8u);
        const bool valid = ((tmk << p) >> (63 - w)) == 0;          // no EndOfWord within the next w+1 bytes: suffix length > w
        const bool full = i == 0 || (i > 1 && pcs[p] == 0) ;       // preceded by EndOfWord (or the very first byte)
        u64 c = (valid && d < wl.Dw) ? (u64)occ[d] : 0ull;
        const u64 kk = pack_key(code, j < L ? (u32)j + 1 : (u32)L, j < L && kind == 0, pc) | kx.pack(c, valid && full);
        if (MODE == 0) sk[i0 + p + (i0 >> 3)] = kk;
        else {
            const bool sel = valid && d < wl.Dw && i < D && code >= ks.lo && code < ks.hi;
            if (MODE == 1) { if (valid && d < wl.Dw && i < D) atomic_add(&ks.hist[code >> ks.hshift], 4ull + c); }
            else if (MODE >= 4) {
                if (valid && d < wl.Dw && i < D) dstv = (dstv & ~(0xffull << (8 * p))) | ((u64)key_dest(ks.bound, ks.nd, code) << (8 * p));
                if (MODE == 5) mykeys[MODE == 5 ? p : 0] = kk;
            }
            else { selmask |= sel ? 1u << p : 0u; if (MODE == 3) mykeys[MODE == 3 ? p : 0] = kk; }
        }
        d += (u32)((tmk >> (63 - p)) & 1);
    }
    if (MODE == 0) {
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            int o = k * 256 + threadIdx.x;
            u64 g = base + (u64)o;
            if (g < D) key[g] = sk[o + (o >> 3)];
        }
    } else if (MODE >= 4) {
        // per destination: exclusive scan of the per-thread counts over the CTA -> rank of every routed position inside
        // the (tile, destination) run, in dictionary order
        __shared__ u32 rsum[8][KEY_MAX_DEST];
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
        u32 offp[8];
#pragma unroll
        for (int p = 0; p < 8; ++p) offp[p] = 0;
        for (int dd = 0; dd < ks.nd; ++dd) {
            const u64 x = dstv ^ ((u64)dd * 0x0101010101010101ull);                          // zero byte <=> destination dd
            const u64 z = ~((((x & 0x7f7f7f7f7f7f7f7full) + 0x7f7f7f7f7f7f7f7full) | x) | 0x7f7f7f7f7f7f7f7full);
            const u32 c = (u32)__popcll((unsigned long long)z);
            u32 inc = c;
#pragma unroll
            for (int q = 1; q < 32; q <<= 1) { u32 o = __shfl_up_sync(0xffffffffu, inc, q); if (lane >= q) inc += o; }
            if (lane == 31) rsum[wid][dd] = inc;
            if (MODE == 5 && c) {
                u32 e = inc - c;
#pragma unroll
                for (int p = 0; p < 8; ++p) if ((z >> (8 * p + 7)) & 1) offp[p] = e++;
            }
        }
        __syncthreads();
        if (MODE == 4) {
            if ((int)threadIdx.x < ks.nd) {
                u32 t = 0;
#pragma unroll
                for (int k = 0; k < 8; ++k) t += rsum[k][threadIdx.x];
                ks.tilecnt[(u64)threadIdx.x * ks.ntl + blockIdx.x] = t;
            }
        } else {
            // the tile's pairs are grouped by destination in shared memory, then every destination's run leaves as
            // unit-stride stores: towards a peer GPU these are full 128-byte NVLink writes instead of scattered 8-byte ones
            __shared__ u32 dstart[KEY_MAX_DEST + 1];
            __shared__ u64 dglob[KEY_MAX_DEST];
            __shared__ DP sidx[MODE == 5 ? KEYS_TILE : 1];
            __shared__ u8 sdst[MODE == 5 ? KEYS_TILE : 1];
            if (threadIdx.x == 0) {
                u32 run = 0;
                for (int dd = 0; dd < ks.nd; ++dd) {
                    u32 t = 0;
#pragma unroll
                    for (int k = 0; k < 8; ++k) t += rsum[k][dd];
                    dstart[dd] = run; run += t;
                    dglob[dd] = ks.tilebase[(u64)dd * ks.ntl + blockIdx.x] - ks.destbase[dd];
                }
                dstart[ks.nd] = run;
            }
            __syncthreads();
#pragma unroll
            for (int p = 0; p < 8; ++p) {
                const u32 dd = (u32)(dstv >> (8 * p)) & 0xffu;
                if (dd == 0xffu) continue;
                u32 wb = 0;
                for (int k = 0; k < wid; ++k) wb += rsum[k][dd];
                const u32 q = dstart[dd] + wb + offp[p];
                sk[q] = mykeys[MODE == 5 ? p : 0];
                sidx[MODE == 5 ? q : 0] = (DP)(base + (u64)(i0 + p));
                sdst[MODE == 5 ? q : 0] = (u8)dd;
            }
            __syncthreads();
            const u32 T = dstart[ks.nd];
            for (u32 i = threadIdx.x; i < T; i += 256) {
                const u32 dd = sdst[MODE == 5 ? i : 0];
                const u64 o = dglob[dd] + (u64)(i - dstart[dd]);
                ks.key_outs[dd][o] = sk[i];
                ks.idx_outs[dd][o] = sidx[MODE == 5 ? i : 0];
            }
        }
    } else if (MODE >= 2) {
        // ordered compaction inside the tile: exclusive scan of the per-thread selected counts
        __shared__ u32 wsum[8];
        const u32 mine = (u32)__popc(selmask);
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
        u32 inc = mine;
#pragma unroll
        for (int dd = 1; dd < 32; dd <<= 1) { u32 o = __shfl_up_sync(0xffffffffu, inc, dd); if (lane >= dd) inc += o; }
        if (lane == 31) wsum[wid] = inc;
        __syncthreads();
        u32 before = 0, tot = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) { u32 s_ = wsum[k]; if (k < wid) before += s_; tot += s_; }
        if (MODE == 2) { if (threadIdx.x == 0) ks.tilecnt[blockIdx.x] = tot; }
        else {
            u64 o = ks.tilebase[blockIdx.x] + before + inc - mine;
#pragma unroll
            for (int p = 0; p < 8; ++p)
                if (selmask & (1u << p)) { ks.key_out[o] = mykeys[MODE == 3 ? p : 0]; ks.idx_out[o] = (DP)(base + (u64)(i0 + p)); ++o; }
        }
    }
}
#endif

// keys of all D dictionary positions, L <= KEY_MAX_SYMS symbols per key (the identity permutation is
// implied: the radix sort's first pass takes value = position)
template <class DP>
inline void k_suffix_keys(const u8 *dict, u64 D, int L, u64 *key, const WordLookupT<DP> &wl, const u32 *occ, int w, const KeyExtra &kx,
                          pfbStream_t st) {
#ifdef PFB_EMU
    (void)st;
    for (u64 i = 0; i < D; ++i) {
        EnumKey e = enum_scan(dict, D, i, L);
        u32 d = wl.word((DP)i);
        bool valid = wl.suflen(d, (DP)i) > (u64)w;
        bool full = valid && (DP)i == wl.start(d);
        key[i] = pack_key(e.code, e.consumed, e.closed, (u32)prev_code(dict, i)) | kx.pack(valid ? occ[d] : 0, full);
    }
#else
    if (D == 0) return;
    k_suffix_keys_kernel<0, DP><<<(unsigned)((D + KEYS_TILE - 1) / KEYS_TILE), 256, 0, st>>>(dict, D, L, key, wl, occ, w, kx, KeySelectT<DP>());
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
#endif
}

// ---- range-restricted keys (multi-GPU path): only VALID suffixes (length > w; the others emit
// nothing and cannot tie with a valid one) whose code lies in [lo, hi) are handed to the sort ----
template <class DP>
PFB_HD inline bool key_is_valid(const WordLookupT<DP> &wl, DP i, int w, u32 &d) {
    d = wl.word(i);
    return d < wl.Dw && wl.suflen(d, i) > (u64)w;
}
// weighted histogram (4 + occurrences per valid suffix: sort work + emission work) of code >> hshift.
// The CUDA build samples every `stride`-th tile: splitters only have to balance, not to be exact.
template <class DP>
inline void k_suffix_key_hist(const u8 *dict, u64 D, int L, const WordLookupT<DP> &wl, const u32 *occ, int w, const KeyExtra &kx,
                              u64 *hist, int hshift, pfbStream_t st) {
#ifdef PFB_EMU
    (void)st; (void)kx;
    for (u64 i = 0; i < D; ++i) {
        u32 d;
        if (!key_is_valid<DP>(wl, (DP)i, w, d)) continue;
        EnumKey e = enum_scan(dict, D, i, L);
        hist[e.code >> hshift] += 4ull + occ[d];
    }
#else
    if (D == 0) return;
    const u64 nt = (D + KEYS_TILE - 1) / KEYS_TILE;
    KeySelectT<DP> ks;
    ks.hist = hist; ks.hshift = hshift;
    ks.tile_stride = nt > 16384 ? (u32)(nt / 16384) : 1u;
    k_suffix_keys_kernel<1, DP><<<(unsigned)((nt + ks.tile_stride - 1) / ks.tile_stride), 256, 0, st>>>(dict, D, L, nullptr, wl, occ, w, kx, ks);
    PFB_CUDA_CHECK(cudaGetLastError());
    launch_stats().own++;
#endif
}
// selected (key, position) pairs in dictionary order; returns how many.  key_out / idx_out are (re)allocated.
template <class DP, class WS>
inline u64 k_suffix_keys_select(const u8 *dict, u64 D, int L, const WordLookupT<DP> &wl, const u32 *occ, int w, const KeyExtra &kx,
                                u64 lo, u64 hi, DevBuf<u64> &key_out, DevBuf<DP> &idx_out, u64 (*pad)(u64), WS &ws) {
    pfbStream_t st = ws.st;
#ifdef PFB_EMU
    std::vector<u64> ks_; std::vector<DP> is_;
    for (u64 i = 0; i < D; ++i) {
        u32 d;
        if (!key_is_valid<DP>(wl, (DP)i, w, d)) continue;
        EnumKey e = enum_scan(dict, D, i, L);
        if (e.code < lo || e.code >= hi) continue;
        bool full = (DP)i == wl.start(d);
        ks_.push_back(pack_key(e.code, e.consumed, e.closed, (u32)prev_code(dict, i)) | kx.pack(occ[d], full));
        is_.push_back((DP)i);
    }
    const u64 m = ks_.size();
    key_out.alloc(m); idx_out.alloc(pad(m));
    if (m) { memcpy(key_out.get(), ks_.data(), m * 8); memcpy(idx_out.get(), is_.data(), m * sizeof(DP)); }
    (void)st;
    return m;
#else
    const u64 nt = (D + KEYS_TILE - 1) / KEYS_TILE;
    if (nt == 0) { key_out.alloc(0); idx_out.alloc(pad(0)); return 0; }
    DevBuf<u32> tcnt(nt + 1);
    DevBuf<u64> tbase(nt + 1);
    KeySelectT<DP> ks;
    ks.lo = lo; ks.hi = hi; ks.tilecnt = tcnt.get();
    k_suffix_keys_kernel<2, DP><<<(unsigned)nt, 256, 0, st>>>(dict, D, L, nullptr, wl, occ, w, kx, ks);
    PFB_CUDA_CHECK(cudaGetLastError());
    dev_memset(tcnt.get() + nt, 0, 4, st);
    exclusive_sum<u32, u64>(tcnt.get(), tbase.get(), nt + 1, ws.sc, st);
    const u64 m = read_scalar(tbase.get() + nt, st);
    key_out.alloc(m); idx_out.alloc(pad(m));
    if (m) {
        ks.tilebase = tbase.get(); ks.key_out = key_out.get(); ks.idx_out = idx_out.get();
        k_suffix_keys_kernel<3, DP><<<(unsigned)nt, 256, 0, st>>>(dict, D, L, nullptr, wl, occ, w, kx, ks);
        PFB_CUDA_CHECK(cudaGetLastError());
    }
    launch_stats().own += 2;
    return m;
#endif
}

// ---- position-sliced routing (multi-GPU path): this rank scans the tiles [tile0, tile0 + ntl) of the dictionary and
// sends the (key, position) pair of every valid suffix to the rank that owns its code range.  Two steps, because the
// receivers' buffers can only be sized (and their addresses exchanged) once every rank knows what it will get:
//   count()    keys of the slice, pairs per destination
//   scatter()  keys again, every pair written to its place: key_outs[d] / idx_outs[d] point at the start of THIS rank's
//              part of destination d's pairs - inside a local staging buffer, or inside rank d's receive buffer mapped over
//              NVLink (the stores of the key kernel are then the exchange).  A destination's part is in dictionary order.
template <class DP>
struct KeyRouterT {
    const u8 *dict; u64 D; int L; WordLookupT<DP> wl; const u32 *occ; int w; KeyExtra kx;
    std::vector<u64> bound; u64 tile0 = 0, ntl = 0;
    std::vector<u64> counts, destbase;        // pairs for destination d / pairs for the destinations before d
#ifdef PFB_EMU
    std::vector<std::vector<u64>> ks_;
    std::vector<std::vector<DP>> is_;
#else
    DevBuf<u32> tcnt;
    DevBuf<u64> tbase;
    KeySelectT<DP> ks;
#endif
    template <class WS>
    void count(WS &ws) {
        pfbStream_t st = ws.st;
        const int nd = (int)bound.size() - 1;
        counts.assign((size_t)nd, 0); destbase.assign((size_t)nd + 1, 0);
#ifdef PFB_EMU
        (void)st;
        ks_.assign((size_t)nd, std::vector<u64>()); is_.assign((size_t)nd, std::vector<DP>());
        const u64 p0 = tile0 * KEYS_TILE, p1 = (tile0 + ntl) * KEYS_TILE < D ? (tile0 + ntl) * KEYS_TILE : D;
        for (u64 i = p0; i < p1; ++i) {
            u32 d;
            if (!key_is_valid<DP>(wl, (DP)i, w, d)) continue;
            EnumKey e = enum_scan(dict, D, i, L);
            const int dst = key_dest(bound.data(), nd, e.code);
            bool full = (DP)i == wl.start(d);
            ks_[(size_t)dst].push_back(pack_key(e.code, e.consumed, e.closed, (u32)prev_code(dict, i)) | kx.pack(occ[d], full));
            is_[(size_t)dst].push_back((DP)i);
        }
        for (int d = 0; d < nd; ++d) { counts[(size_t)d] = ks_[(size_t)d].size(); destbase[(size_t)d + 1] = destbase[(size_t)d] + counts[(size_t)d]; }
#else
        if (ntl == 0) return;
        const u64 cells = (u64)nd * ntl;
        tcnt.alloc(cells + 1); tbase.alloc(cells + 1);
        DevBuf<u64> ends((size_t)nd + 1);
        ks = KeySelectT<DP>();
        ks.tile0 = tile0; ks.ntl = ntl; ks.nd = nd;
        for (int d = 0; d <= nd; ++d) ks.bound[d] = bound[(size_t)d];
        ks.tilecnt = tcnt.get();
        k_suffix_keys_kernel<4, DP><<<(unsigned)ntl, 256, 0, st>>>(dict, D, L, nullptr, wl, occ, w, kx, ks);
        PFB_CUDA_CHECK(cudaGetLastError());
        launch_stats().own++;
        dev_memset(tcnt.get() + cells, 0, 4, st);
        exclusive_sum<u32, u64>(tcnt.get(), tbase.get(), cells + 1, ws.sc, st);
        {
            const u64 *tb = tbase.get(); u64 *ep = ends.get(); const u64 ntl_ = ntl;
            launch_for((u64)nd + 1, st, PFB_LAMBDA(u64 d) { ep[d] = tb[d * ntl_]; });
        }
        d2h(destbase.data(), ends.get(), ((size_t)nd + 1) * 8, st);
        stream_sync(st);
        for (int d = 0; d < nd; ++d) counts[(size_t)d] = destbase[(size_t)d + 1] - destbase[(size_t)d];
#endif
    }
    template <class WS>
    void scatter(u64 *const *key_outs, DP *const *idx_outs, WS &ws) {
        pfbStream_t st = ws.st;
        const int nd = (int)bound.size() - 1;
#ifdef PFB_EMU
        (void)st;
        for (int d = 0; d < nd; ++d)
            if (counts[(size_t)d]) { memcpy(key_outs[d], ks_[(size_t)d].data(), counts[(size_t)d] * 8); memcpy(idx_outs[d], is_[(size_t)d].data(), counts[(size_t)d] * sizeof(DP)); }
        ks_.clear(); is_.clear();
#else
        if (ntl == 0 || destbase[(size_t)nd] == 0) return;
        ks.tilebase = tbase.get();
        for (int d = 0; d < nd; ++d) { ks.key_outs[d] = key_outs[d]; ks.idx_outs[d] = idx_outs[d]; ks.destbase[d] = destbase[(size_t)d]; }
        k_suffix_keys_kernel<5, DP><<<(unsigned)ntl, 256, 0, st>>>(dict, D, L, nullptr, wl, occ, w, kx, ks);
        PFB_CUDA_CHECK(cudaGetLastError());
        launch_stats().own++;
#endif
    }
};

}  // namespace pfb
