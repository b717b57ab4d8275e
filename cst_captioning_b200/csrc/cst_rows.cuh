// Rows of token ids -> canonical token codes -> n-gram instances, and the shared-memory table of
// packed n-gram keys: the device helpers the scoring kernels share (k_score2 in cst_score2.cuh; the
// BLEU / ROUGE-L scorers in cst_metrics.cu).  Needs cst_common.cuh.
#pragma once

#define CST2_ROWS 32                 /* rows per batch: presence bits of a chunk, lanes of the prefix scan */
#define CST2_TOK_STRIDE 32
typedef unsigned short u16;

// One warp, one row, lane = raw position (utils.py:135-152).  `u` = token id + 1 as an unsigned
// (0: padding id -1 or a lane past the row's width, 1: <end>, 2: <start>, > 65535: id out of range).
// Writes the canonical token codes to out[0 .. L) and returns L | bad << 8, the same in every lane.
// Branch-free on purpose: predicates combined with & and |, one pass of three ballots.
__device__ __forceinline__ int v2_canon(u32 u, int use_eos, int lane, u16* out) {
    const u32 isbad = u > 65535u;
    const u32 term = (u - 2u) > 65533u;                        // padding, <end> or a bad id end the row
    const u32 tmask = __ballot_sync(CST_FULL, term);
    const u32 bmask = __ballot_sync(CST_FULL, isbad);
    const int end = __clz(__brev(tmask));                      // first terminator; 32 if there is none
    const u32 keep = ((u32)(lane < end) & (u32)(u != 2u)) | ((u32)(use_eos != 0) & (u32)(lane == end) & (u32)(u == 1u));
    const u32 kmask = __ballot_sync(CST_FULL, keep);
    if (keep) out[__popc(kmask & lanemask_lt(lane))] = (u16)u;
    const u32 bad = (u32)(((u64)bmask >> end) & 1ull);
    return __popc(kmask) | (int)(bad << 8);
}

// List layout: ORDER-MAJOR.  All unigram instances of the batch come first, then the bigrams, the
// trigrams, the 4-grams; inside an order, row after row, position after position.  A warp of the
// processing pass therefore works on ONE order: unigram warps take their idf by a direct load and
// walk long presence masks together, 4-gram warps probe the df table and hardly walk at all --
// with a row-major list every warp paid for the slowest of both.  (row, order k) still owns a
// contiguous segment: base[k] + excl_k(row), max(L - k, 0) slots.
//
// Every warp scans the row lengths itself (no barrier, no shared offsets): the four per-order
// counts of a row are <= 32, so two packed 16-bit pairs ride two shuffle scans.
struct V2Scan {
    u32 exclA, exclB;     // lane r: slots of rows < r, orders {0, 1} and {2, 3}, packed 16 | 16
    int L;                // lane r: canonical length of row r
    int base1, base2, base3, n_ent;   // uniform: first slot of each order (order 0 starts at 0), total
    int nfit;             // uniform: rows of the batch that fit the list (>= 1)
};
__device__ __forceinline__ V2Scan v2_scan(const int* sL, int nrows, int budget, int lane) {
    V2Scan s;
    s.L = (lane < nrows) ? (sL[lane] & 0xff) : 0;
    const int L = s.L;
    const u32 c0 = (u32)L, c1 = (u32)(L > 1 ? L - 1 : 0), c2 = (u32)(L > 2 ? L - 2 : 0), c3 = (u32)(L > 3 ? L - 3 : 0);
    const u32 a = c0 | (c1 << 16), b = c2 | (c3 << 16);
    u32 ia = a, ib = b;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const u32 va = __shfl_up_sync(CST_FULL, ia, o);
        const u32 vb = __shfl_up_sync(CST_FULL, ib, o);
        if (lane >= o) { ia += va; ib += vb; }
    }
    const u32 tot = (ia & 0xffffu) + (ia >> 16) + (ib & 0xffffu) + (ib >> 16);
    const u32 fit = __ballot_sync(CST_FULL, lane < nrows && tot <= (u32)budget);
    s.nfit = __popc(fit);                     // prefix sums rise: the rows that fit are the first ones
    if (s.nfit < 1) s.nfit = 1;               // (a single row always fits: budget >= 4 * 32 - 6)
    const u32 ta = __shfl_sync(CST_FULL, ia, s.nfit - 1), tb = __shfl_sync(CST_FULL, ib, s.nfit - 1);
    s.base1 = (int)(ta & 0xffffu);
    s.base2 = s.base1 + (int)(ta >> 16);
    s.base3 = s.base2 + (int)(tb & 0xffffu);
    s.n_ent = s.base3 + (int)(tb >> 16);
    s.exclA = ia - a;
    s.exclB = ib - b;
    return s;
}
// first slot of (row, order k); `row` may differ per lane
__device__ __forceinline__ int v2_seg(const V2Scan& s, int row, int k) {
    const u32 ea = __shfl_sync(CST_FULL, s.exclA, row), eb = __shfl_sync(CST_FULL, s.exclB, row);
    const u32 e = (k & 2) ? eb : ea;
    const int within = (int)((k & 1) ? (e >> 16) : (e & 0xffffu));
    const int base = (k == 0) ? 0 : (k == 1) ? s.base1 : (k == 2) ? s.base2 : s.base3;
    return base + within;
}

// n-gram instances of one row (G = 16: two rows per warp, one per half; G = 32: rows longer than 16).
// Restates precook (compute_ciderdf.py:28-33) on packed keys: the slot of (order k, position i) holds
// {key, row | k << 5 | tf << 8} for the first occurrence of an n-gram in the row, meta 0 for repeats.
template <int G>
__device__ __forceinline__ void v2_emit(const u16* ctok, int row, int L, const V2Scan& s, uint4* list, int lane) {
    const int sl = lane & (G - 1);
    const int sh = lane & (32 - G);
    const u32 window = (G == 32) ? CST_FULL : 0xffffu;
    const u32 ea = __shfl_sync(CST_FULL, s.exclA, row), eb = __shfl_sync(CST_FULL, s.exclB, row);
    const u32 e0 = (sl < L) ? (u32)ctok[row * CST2_TOK_STRIDE + sl] : 0u;
    u32 e1 = __shfl_down_sync(CST_FULL, e0, 1, G);
    u32 e2 = __shfl_down_sync(CST_FULL, e0, 2, G);
    u32 e3 = __shfl_down_sync(CST_FULL, e0, 3, G);
    if (sl + 1 >= G) e1 = 0;                  // (positions past L already hold 0)
    if (sl + 2 >= G) e2 = 0;
    if (sl + 3 >= G) e3 = 0;
    // E1(i) = positions of the row holding the token of position i; E_k(i) = E_{k-1}(i) & (E1(i+k) >> k)
    const u32 m1 = (__match_any_sync(CST_FULL, e0) >> sh) & window;
    u32 n1 = __shfl_down_sync(CST_FULL, m1, 1, G);
    u32 n2 = __shfl_down_sync(CST_FULL, m1, 2, G);
    u32 n3 = __shfl_down_sync(CST_FULL, m1, 3, G);
    if (sl + 1 >= G) n1 = 0;
    if (sl + 2 >= G) n2 = 0;
    if (sl + 3 >= G) n3 = 0;
    const u32 ek0 = m1;
    const u32 ek1 = ek0 & (n1 >> 1);
    const u32 ek2 = ek1 & (n2 >> 2);
    const u32 ek3 = ek2 & (n3 >> 3);
    const u32 lo1 = e0 | (e1 << 16);
    const u32 hi3 = e2 | (e3 << 16);
    const u32 rowbits = (u32)row;
    uint4* dst = list + sl;
    // meta = 0 marks a repeat (tf >= 1 otherwise); the key words are stored either way
    const u32 sbit = 1u << sl;
    const u32 m0 = ((ek0 & (0u - ek0)) == sbit) ? (rowbits | ((u32)__popc(ek0) << 8)) : 0u;
    const u32 m1k = ((ek1 & (0u - ek1)) == sbit) ? (rowbits | (1u << 5) | ((u32)__popc(ek1) << 8)) : 0u;
    const u32 m2k = ((ek2 & (0u - ek2)) == sbit) ? (rowbits | (2u << 5) | ((u32)__popc(ek2) << 8)) : 0u;
    const u32 m3k = ((ek3 & (0u - ek3)) == sbit) ? (rowbits | (3u << 5) | ((u32)__popc(ek3) << 8)) : 0u;
    if (sl < L) dst[ea & 0xffffu] = make_uint4(e0, 0u, m0, 0u);
    if (sl + 1 < L) dst[s.base1 + (ea >> 16)] = make_uint4(lo1, 0u, m1k, 0u);
    if (sl + 2 < L) dst[s.base2 + (eb & 0xffffu)] = make_uint4(lo1, e2, m2k, 0u);
    if (sl + 3 < L) dst[s.base3 + (eb >> 16)] = make_uint4(lo1, hi3, m3k, 0u);
}

// The per-video table, capacity T (a compile-time constant).  Double hashing: the step of a probe
// sequence comes from the key's hash as well, so keys that collide in one slot part ways at once and
// no clusters build up -- what a warp waits for is the LONGEST probe sequence among its 32 lanes, and
// linear probing's long tail showed there.  T is prime, so every step in [1, T) visits all slots.
#ifndef CST2_DOUBLE_HASH
#define CST2_DOUBLE_HASH 1
#endif
template <u32 T>
__device__ __forceinline__ u32 v2_step(u32 h) {
    return CST2_DOUBLE_HASH ? 1u + __umulhi(h * 0x9E3779B1u, T - 1u) : 1u;
}
template <u32 T>
__device__ __forceinline__ u32 v2_next(u32 idx, u32 step) {
    idx += step;
    return (idx >= T) ? idx - T : idx;
}
template <u32 T>
__device__ __forceinline__ u32 v2_insert(VidSlot* tab, u32 h, u64 key) {
    u32 idx = vid_home(h, T);
    const u32 step = v2_step<T>(h);
#pragma unroll 1
    for (;;) {
        const u64 cur = *reinterpret_cast<volatile u64*>(&tab[idx].key);
        if (cur == key) return idx;
        if (cur == 0ull) {
            const u64 old = atomicCAS(reinterpret_cast<u64*>(&tab[idx].key), 0ull, key);
            if (old == 0ull || old == key) return idx;
        }
        idx = v2_next<T>(idx, step);
    }
}
// BOUNDED: the table came from a caller's buffer (a staged image): a damaged one may hold no free
// slot on a probe path, so the walk gives up after T steps instead of spinning.
template <u32 T, bool BOUNDED>
__device__ __forceinline__ int v2_find(const VidSlot* tab, u32 h, u64 key) {
    u32 idx = vid_home(h, T);
    const u32 step = v2_step<T>(h);
#pragma unroll 1
    for (u32 n = 0; !BOUNDED || n < T; ++n) {
        const u64 cur = tab[idx].key;
        if (cur == key) return (int)idx;
        if (cur == 0ull) return -1;
        idx = v2_next<T>(idx, step);
    }
    return -1;
}
