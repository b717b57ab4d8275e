// k_score2 -- the raw-row scoring kernel of round 2 (included by cst_reward.cu after ScoreParams).
//
// Same arithmetic as k_score (CIDEr-D of every candidate against its video's references, then the
// baseline of utils.py:292-322), reorganised around ONE idea: the expensive per-n-gram work -- the df
// probe into the L2-resident table, the insert / find in the per-video shared-memory table, the walk
// over the presence bits -- runs over a dense LIST of n-gram instances with every lane busy, instead
// of "lane = token position, four orders per lane" where a third of the lanes idle and the four
// order bodies diverge.  One CTA per video:
//
//   canon    warp per row, lane = raw position: ballot/popc compaction of utils.py:135-152
//            (array_to_str) -> canonical token codes (u16) + length per row in shared memory
//   emit     half-warp per row (full warp for rows longer than 16 tokens), lane = start position:
//            packed 1..4-gram keys, term frequency and "first occurrence" from one MATCH.ANY
//            (compute_ciderdf.py:19-34, precook); instance (row, order k, position) owns list slot
//            off[row] + koff(k) + position -- a layout known from the row lengths alone, so no
//            compaction pass and no atomics; repeats inside a row leave their slot empty
//   process  thread per list slot: idf (direct load for unigrams, one 32-byte bucket otherwise);
//            references: insert into the video table, OR the row's presence bit; candidates: find,
//            walk the presence bits.  The slot is overwritten with {(tf idf)^2, idf^2 * inner}
//   sums     thread per (row, order): its list segment is contiguous, summed front to back ->
//            norms, numerators.  No floating-point atomics, no shuffle trees: deterministic.
//
// Shared memory is a fixed 44 KB whatever the row widths and counts: a video whose references do
// not fit the list (CST2_LIST_CAP instances) or need more than 32 presence bits is scored in several
// reference chunks (CIDEr-D adds over references), candidates go through in batches likewise.
#pragma once

#define CST2_ROWS 32                 /* rows per batch: presence bits of a chunk, lanes of the prefix scan */
#define CST2_TOK_STRIDE 32
#define CST2_LIST_CAP 1024           /* n-gram instances per batch (16 B each) */
#ifndef CST2_TAB_SLOTS
#define CST2_TAB_SLOTS 1472          /* per-video table; a chunk inserts at most CST2_LIST_CAP keys */
#endif
#ifndef CST2_WARPS
#define CST2_WARPS 8
#endif
#ifndef CST2_CTAS
#define CST2_CTAS 5
#endif

typedef unsigned short u16;

struct __align__(16) RefA {          // what a presence bit needs of its reference, one 16-byte load
    double inv;                      // 1 / norm of the ref's order-k vector (0 if the norm is 0)
    int len;                         // the scorer's "length": number of bigrams
    int pad;
};

// One warp, one row, lane = raw position (utils.py:135-152).  Writes the canonical token codes
// (id + 1) to out[0 .. L) and returns L | bad << 8, the same value in every lane.
__device__ __forceinline__ int v2_canon(int t, bool in, int width, int use_eos, int lane, u16* out) {
    const bool isbad = in && (t < -1 || t > 65534);
    const bool term = in && (t == 0 || t == -1 || isbad);      // -1 = padding: ends the row, never emitted
    const u32 tmask = __ballot_sync(CST_FULL, term);
    const u32 bmask = __ballot_sync(CST_FULL, isbad);
    const int end = tmask ? (__ffs(tmask) - 1) : width;
    const bool keep = (lane < end && t != 1) || (use_eos && lane == end && in && t == 0);
    const u32 kmask = __ballot_sync(CST_FULL, keep);
    if (keep) out[__popc(kmask & lanemask_lt(lane))] = (u16)((u32)t + 1u);
    const int bad = (end < 32) ? (int)((bmask >> end) & 1u) : 0;
    return __popc(kmask) | (bad << 8);
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
// {key, row | k << 5 | tf << 8} for the first occurrence of an n-gram in the row, zeros for repeats.
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
    const uint4 zero = make_uint4(0, 0, 0, 0);
    if (sl < L) {
        const bool lead = (__ffs(ek0) - 1) == sl;
        dst[ea & 0xffffu] = lead ? make_uint4(e0, 0u, rowbits | ((u32)__popc(ek0) << 8), 0u) : zero;
    }
    if (sl + 1 < L) {
        const bool lead = (__ffs(ek1) - 1) == sl;
        dst[s.base1 + (ea >> 16)] =
            lead ? make_uint4(lo1, 0u, rowbits | (1u << 5) | ((u32)__popc(ek1) << 8), 0u) : zero;
    }
    if (sl + 2 < L) {
        const bool lead = (__ffs(ek2) - 1) == sl;
        dst[s.base2 + (eb & 0xffffu)] =
            lead ? make_uint4(lo1, e2, rowbits | (2u << 5) | ((u32)__popc(ek2) << 8), 0u) : zero;
    }
    if (sl + 3 < L) {
        const bool lead = (__ffs(ek3) - 1) == sl;
        dst[s.base3 + (eb >> 16)] =
            lead ? make_uint4(lo1, hi3, rowbits | (3u << 5) | ((u32)__popc(ek3) << 8), 0u) : zero;
    }
}

// which of the bucket's four keys equals `key` (-1: none), without branches
__device__ __forceinline__ int v2_match4(const u64 kk[4], u64 key) {
    int j = -1;
    if (kk[0] == key) j = 0;
    if (kk[1] == key) j = 1;
    if (kk[2] == key) j = 2;
    if (kk[3] == key) j = 3;
    return j;
}

// idf of a packed n-gram of order k >= 1: one 32-byte bucket of the df table.  A probe ends at a
// bucket with a free slot, or at a full one nothing overflowed from (its first two keys ascend).
__device__ __forceinline__ double v2_idf_hashed(const DfView& t, u64 key, u32 h) {
    u64 kk[4];
    u32 b = cst_slot(h, t.n_buckets);
    ld_bucket(t.keys + (size_t)b * CST_BUCKET, kk);
    int j = v2_match4(kk, key);
    bool more = j < 0 && kk[3] != 0ull && !(kk[0] < kk[1]);
#pragma unroll 1
    while (more) {                                             // a few percent of the probes
        b = (b + 1 == t.n_buckets) ? 0u : b + 1;
        ld_bucket(t.keys + (size_t)b * CST_BUCKET, kk);
        j = v2_match4(kk, key);
        more = j < 0 && kk[3] != 0ull && !(kk[0] < kk[1]);
    }
    return (j >= 0) ? __ldg(t.idfv + (size_t)b * CST_BUCKET + j) : t.unseen;
}

// the per-video table with a compile-time capacity
template <u32 T>
__device__ __forceinline__ u32 v2_next(u32 idx) {
    if ((T & (T - 1)) == 0) return (idx + 1) & (T - 1);
    return (idx + 1 == T) ? 0u : idx + 1;
}
template <u32 T>
__device__ __forceinline__ u32 v2_insert(VidSlot* tab, u32 idx, u64 key) {
#pragma unroll 1
    for (;;) {
        const u64 cur = *reinterpret_cast<volatile u64*>(&tab[idx].key);
        if (cur == key) return idx;
        if (cur == 0ull) {
            const u64 old = atomicCAS(reinterpret_cast<u64*>(&tab[idx].key), 0ull, key);
            if (old == 0ull || old == key) return idx;
        }
        idx = v2_next<T>(idx);
    }
}
template <u32 T>
__device__ __forceinline__ int v2_find(const VidSlot* tab, u32 idx, u64 key) {
#pragma unroll 1
    for (;;) {
        const u64 cur = tab[idx].key;
        if (cur == key) return (int)idx;
        if (cur == 0ull) return -1;
        idx = v2_next<T>(idx);
    }
}

// One pass over the list: thread per slot.
//   is_r:  publish the instance into the video table (presence bit of its row; exact count under a
//          synthetic key when the row holds the n-gram more than once), slot <- {(tf idf)^2, 0}
//   else:  find it, walk the presence bits, slot <- {(tf idf)^2, idf^2 * sum_r min(tf_h, tf_r) tf_r inv_r g_r}
template <int NT, u32 T>
__device__ __forceinline__ void v2_process(const ScoreParams& p, bool is_r, uint4* list, int n_ent, int base1,
                                           VidSlot* tab, const RefA* refA, const int* sL, const double* s_gauss) {
#pragma unroll 1
    for (int i0 = 0; i0 < n_ent; i0 += NT) {
        const int i = i0 + (int)threadIdx.x;
        if (i >= n_ent) break;
        const uint4 e = list[i];
        const u64 key = (u64)e.x | ((u64)e.y << 32);
        double w2 = 0.0, num = 0.0;
        if (key != 0ull) {
            const int row = (int)(e.z & 31u);
            const int k = (int)((e.z >> 5) & 3u);
            const int tf = (int)(e.z >> 8);
            const u32 h = cst_hash(key);
            double idf;
            if (i < base1) idf = __ldg(p.df.uni + (e.x - 1u));      // unigram: token codes are 1 .. 65535
            else idf = v2_idf_hashed(p.df, key, h);
            const double w = (double)tf * idf;                 // counts2vec: vec = tf * idf
            w2 = w * w;                                        // norm += vec ** 2
            // an n-gram every document holds has idf = 0: exact zeros on both sides, no table entry
            if (idf != 0.0) {
                if (is_r) {
                    const u32 idx = v2_insert<T>(tab, vid_home(h, T), key);
                    const u32 bit = 1u << row;
                    atomicOr(&tab[idx].P, bit);
                    if (tf > 1) {
                        atomicOr(&tab[idx].M, bit);
                        const u64 ck = count_key(idx, (u32)row);
                        const u32 j = v2_insert<T>(tab, vid_home(cst_hash(ck), T), ck);
                        tab[j].P = (u32)tf;
                    }
                } else {
                    const int idx = v2_find<T>(tab, vid_home(h, T), key);
                    if (idx >= 0) {
                        const u32 M = tab[idx].M;
                        u32 P = tab[idx].P & ~M;
                        const RefA* a_k = refA + k * CST2_ROWS;
                        const int lenh = (sL[row] & 0xff) - 1 + CST_GAUSS_HALF;   // (L >= 1 here)
                        double inner = 0.0;
#pragma unroll 1
                        while (P) {                // refs holding the n-gram once: min(tf_h, 1) * 1
                            const int r = __ffs(P) - 1;
                            P &= P - 1u;
                            const RefA a = a_k[r];
                            inner += a.inv * s_gauss[lenh - a.len];
                        }
                        u32 Q = M;
#pragma unroll 1
                        while (Q) {                // refs holding it more than once: exact count
                            const int r = __ffs(Q) - 1;
                            Q &= Q - 1u;
                            const u64 ck = count_key((u32)idx, (u32)r);
                            const int j = v2_find<T>(tab, vid_home(cst_hash(ck), T), ck);
                            const int c = (int)tab[j].P;
                            const int m = (tf < c ? tf : c) * c;             // min(tf_h, tf_r) * tf_r
                            const RefA a = a_k[r];
                            inner += (double)m * (a.inv * s_gauss[lenh - a.len]);
                        }
                        num = idf * idf * inner;
                    }
                }
            }
        }
        double2* out = reinterpret_cast<double2*>(list + i);
        *out = make_double2(w2, num);
    }
}

// shared-memory layout: everything the hot loops touch sits at a compile-time offset
#define CST2_OFF_LIST ((size_t)CST2_TAB_SLOTS * sizeof(VidSlot))
#define CST2_OFF_REFA (CST2_OFF_LIST + (size_t)CST2_LIST_CAP * sizeof(uint4))
#define CST2_OFF_CTOK (CST2_OFF_REFA + 4 * CST2_ROWS * sizeof(RefA))
#define CST2_OFF_SLR (CST2_OFF_CTOK + (CST2_ROWS * CST2_TOK_STRIDE + 8) * sizeof(u16))
#define CST2_OFF_SLC (CST2_OFF_SLR + CST2_ROWS * sizeof(int))
#define CST2_OFF_SC (CST2_OFF_SLC + CST2_ROWS * sizeof(int))

template <int NW, int CTAS>
__global__ void __launch_bounds__(NW * 32, CTAS) k_score2(const ScoreParams p) {
    constexpr int NT = NW * 32;
    constexpr u32 T = CST2_TAB_SLOTS;
    constexpr int PER = (CST2_ROWS + NW - 1) / NW;      // rows a warp canonicalises per batch
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ double s_gauss[CST_GAUSS_N];
    const int lane = threadIdx.x & 31;
    // (through a shuffle so that the compiler knows the value is the same in all lanes of the warp)
    const int warp = __shfl_sync(CST_FULL, (int)(threadIdx.x >> 5), 0);
    const int S = p.cands_per_video;
    const int n_local = S * (p.greedy ? 2 : 1);

    VidSlot* tab = reinterpret_cast<VidSlot*>(smem);                          // [T]
    uint4* list = reinterpret_cast<uint4*>(smem + CST2_OFF_LIST);             // [CST2_LIST_CAP]
    RefA* refA = reinterpret_cast<RefA*>(smem + CST2_OFF_REFA);               // [4][32]
    u16* ctok = reinterpret_cast<u16*>(smem + CST2_OFF_CTOK);                 // [32][32] (+ tail pad)
    int* sLr = reinterpret_cast<int*>(smem + CST2_OFF_SLR);                   // [32] refs: L | bad << 8
    int* sLc = reinterpret_cast<int*>(smem + CST2_OFF_SLC);                   // [32] candidates
    double* sc = reinterpret_cast<double*>(smem + CST2_OFF_SC);               // [n_local]
    double* low = sc + n_local;                                               // [S]
    double* red = low + S;                                                    // [2 * NW]

#pragma unroll 1
    for (int i = threadIdx.x; i < CST_GAUSS_N; i += NT) s_gauss[i] = p.gauss[i];
    constexpr int budget = CST2_LIST_CAP;

#pragma unroll 1
    for (int64_t video = blockIdx.x; video < p.n_videos; video += gridDim.x) {
        int64_t r0, r1;
        if (p.ref_offsets) {
            r0 = p.ref_offsets[video];
            r1 = p.ref_offsets[video + 1];
        } else {
            r0 = video * p.refs_per_video;
            r1 = r0 + p.refs_per_video;
        }
        const int R = (int)(r1 - r0);
#pragma unroll 1
        for (int c = threadIdx.x; c < n_local; c += NT) sc[c] = 0.0;
        u32 bad_rows = 0;

#pragma unroll 1
        for (int64_t rbase = r0; rbase < r1;) {
            // a reference chunk (up to 32 rows / CST2_LIST_CAP instances), then every candidate batch
            // against it: both kinds of batch run through the same code below (cb < 0: the chunk)
            {
                uint4* z = reinterpret_cast<uint4*>(tab);
#pragma unroll 2
                for (u32 i = threadIdx.x; i < T; i += NT) z[i] = make_uint4(0, 0, 0, 0);
            }
            V2Scan sr;
            sr.nfit = 0;
#pragma unroll 1
            for (int cb = -1; cb < n_local;) {
                const bool is_r = cb < 0;
                if (cb == 0 && warp < 4) {
                    // norms of the chunk's references (their list segments are still in place):
                    // thread per (row, order), front to back
                    const int row = threadIdx.x >> 2, k = threadIdx.x & 3;
                    const int L = __shfl_sync(CST_FULL, sr.L, row);
                    const int seg0 = v2_seg(sr, row, k);
                    if (row < sr.nfit) {
                        const int n = (L > k) ? L - k : 0;
                        const double2* seg = reinterpret_cast<const double2*>(list + seg0);
                        double a = 0.0;
#pragma unroll 2
                        for (int j = 0; j < n; ++j) a += seg[j].x;
                        RefA ra;
                        ra.inv = inv_norm_of(a);
                        ra.len = (L > 1) ? L - 1 : 0;                        // "length" = bigram count
                        ra.pad = 0;
                        refA[k * CST2_ROWS + row] = ra;
                    }
                }
                int n_try, esz, width;
                int* sL;
                const char* rows;       // first row of the batch (candidates: the video's first sampled row)
                size_t pitch;
                if (is_r) {
                    n_try = (int)((r1 - rbase) < CST2_ROWS ? (r1 - rbase) : CST2_ROWS);
                    esz = p.ref_esz;
                    width = p.ref_width;
                    sL = sLr;
                    pitch = (size_t)p.ref_stride * (size_t)esz;
                    rows = reinterpret_cast<const char*>(p.refs) + (size_t)rbase * pitch;
                } else {
                    n_try = (n_local - cb) < CST2_ROWS ? (n_local - cb) : CST2_ROWS;
                    esz = p.cand_esz;
                    width = p.cand_width;
                    sL = sLc;
                    pitch = (size_t)p.cand_stride * (size_t)esz;
                    rows = reinterpret_cast<const char*>(p.cand) + (size_t)(video * S) * pitch;
                }
                {
                    // canonicalise: warp w takes rows w, w + NW, ...; their loads are issued together
                    const bool in = lane < width;
                    const size_t lane_off = (size_t)(lane * esz);
                    int t[PER];
#pragma unroll
                    for (int j = 0; j < PER; ++j) {
                        const int r = warp + j * NW;
                        t[j] = 0;
                        if (r < n_try) {
                            const char* row;
                            if (is_r) {
                                row = rows + (size_t)r * pitch;
                            } else {
                                const int c = cb + r;                        // sampled rows first, then greedy
                                row = (c >= S) ? reinterpret_cast<const char*>(p.greedy) +
                                                     (size_t)(video * S + (c - S)) * pitch
                                               : rows + (size_t)c * pitch;
                            }
                            if (in) {
                                if (esz == 4) {
                                    t[j] = __ldcs(reinterpret_cast<const int*>(row + lane_off));
                                } else if (esz == 8) {
                                    const long long v = __ldcs(reinterpret_cast<const long long*>(row + lane_off));
                                    t[j] = (v < -1 || v > 65534) ? 0x7fffffff : (int)v;   // out of range: bad
                                } else {
                                    t[j] = (int)__ldcs(reinterpret_cast<const short*>(row + lane_off));
                                }
                            }
                        }
                    }
#pragma unroll
                    for (int j = 0; j < PER; ++j) {
                        const int r = warp + j * NW;
                        if (r < n_try) {
                            const int v = v2_canon(t[j], in, width, p.use_eos, lane, ctok + r * CST2_TOK_STRIDE);
                            if (lane == 0) sL[r] = v;
                        }
                    }
                }
                __syncthreads();
                const V2Scan sb = v2_scan(sL, n_try, budget, lane);
                const int nfit = sb.nfit;
                // rows with ids outside [-1, 65534]: every reference row once, candidates in the first chunk
                if (warp == 0 && lane < nfit && (is_r || rbase == r0)) bad_rows += (u32)(sL[lane] >> 8);
#pragma unroll 1
                for (int pair = warp; 2 * pair < nfit; pair += NW) {
                    const int ra = 2 * pair, rb = ra + 1;
                    const int La = __shfl_sync(CST_FULL, sb.L, ra);
                    const int Lb = (rb < nfit) ? __shfl_sync(CST_FULL, sb.L, rb & 31) : 0;
                    if (La <= 16 && Lb <= 16) {
                        const bool second = (lane >> 4) != 0;
                        v2_emit<16>(ctok, second ? (rb & 31) : ra, second ? Lb : La, sb, list, lane);
                    } else {
#pragma unroll 1
                        for (int q = 0; q < 2; ++q) {
                            const int Lq = q ? Lb : La;
                            if (Lq > 0) v2_emit<32>(ctok, ra + q, Lq, sb, list, lane);
                        }
                    }
                }
                __syncthreads();
                v2_process<NT, T>(p, is_r, list, sb.n_ent, sb.base1, tab, refA, sL, s_gauss);
                __syncthreads();
                if (is_r) {
                    sr = sb;
                    cb = 0;
                    continue;
                }
                if (warp * 8 < nfit) {
                    // thread per (candidate, order): norm and numerator of its segment, front to back
                    const int row = (threadIdx.x >> 2) & 31, k = threadIdx.x & 3;
                    const int L = __shfl_sync(CST_FULL, sb.L, row);
                    const int seg0 = v2_seg(sb, row, k);
                    double val = 0.0;
                    if (row < nfit) {
                        const int n = (L > k) ? L - k : 0;
                        const double2* seg = reinterpret_cast<const double2*>(list + seg0);
                        double a = 0.0, b = 0.0;
#pragma unroll 2
                        for (int j = 0; j < n; ++j) {
                            const double2 v = seg[j];
                            a += v.x;
                            b += v.y;
                        }
                        val = b * inv_norm_of(a);
                    }
                    val += __shfl_xor_sync(CST_FULL, val, 1);
                    val += __shfl_xor_sync(CST_FULL, val, 2);
                    if (row < nfit && k == 0) sc[cb + row] += val;
                }
                __syncthreads();
                cb += nfit;
            }
            rbase += sr.nfit;
        }
        if (R == 0 || n_local == 0) __syncthreads();
        if (bad_rows && p.bad_out) atomicAdd(p.bad_out, bad_rows);

        // ---- phase B: scores, baseline (utils.py:292-322), rewards, per-video statistics ----
        // np.mean over the 4 orders, / len(refs), * 10 (compute_cider)
        const double inv_R = (R > 0) ? __drcp_rn((double)R) : 0.0;
#pragma unroll 1
        for (int c = threadIdx.x; c < n_local; c += NT) {
            const double s = (sc[c] * 0.25) * inv_R * 10.0;
            sc[c] = s;
            if (p.scores) {
                const bool is_greedy = c >= S;
                p.scores[(is_greedy ? p.n_cand : 0) + video * S + (is_greedy ? c - S : c)] = s;
            }
        }
        __syncthreads();
        const int k_mode = p.mode;
        if (k_mode >= 0) {
            const bool scb = (k_mode == CST_BASE_SCB_GT || k_mode == CST_BASE_SCB_SAMPLE);
            const double* x = (k_mode == CST_BASE_SCB_GT) ? (p.bcmr + video * S) : sc;
            if (scb) {
                // k lowest of the row (utils.py:294-312): element i is ranked by one warp with ballots
                // (ties by index = stable sort); low[i] = x_i if its rank < k.  k >= S keeps them all.
                if (p.scb_k >= S) {
#pragma unroll 1
                    for (int i = threadIdx.x; i < S; i += NT) low[i] = x[i];
                } else {
#pragma unroll 1
                    for (int i = warp; i < S; i += NW) {
                        const double xi = x[i];
                        int rank = 0;
#pragma unroll 1
                        for (int j0 = 0; j0 < S; j0 += 32) {
                            const int j = j0 + lane;
                            const double xj = (j < S) ? x[j] : 0.0;
                            rank += __popc(__ballot_sync(CST_FULL, j < S && (xj < xi || (xj == xi && j < i))));
                        }
                        if (lane == 0) low[i] = (rank < p.scb_k) ? xi : 0.0;
                    }
                }
                __syncthreads();
            }
            double lowsum = 0.0;
            if (scb) {
#pragma unroll 1
                for (int i = lane; i < S; i += 32) lowsum += low[i];
                lowsum = warp_sum(lowsum);                       // same fixed tree in every warp
            }
            const double base = scb ? lowsum * __drcp_rn((double)p.scb_k) : 0.0;
#pragma unroll 1
            for (int i = threadIdx.x; i < S; i += NT) {
                const double r = (k_mode == CST_BASE_GREEDY) ? sc[i] - sc[S + i] : sc[i] - base;
                p.reward[video * S + i] = (float)r;
            }
            if (warp == 0) {                                      // the video's share of the statistics
                double am = 0.0, ab = 0.0;
#pragma unroll 1
                for (int i = lane; i < S; i += 32) {
                    am += sc[i];
                    if (k_mode == CST_BASE_GREEDY) ab += sc[S + i];
                    if (k_mode == CST_BASE_SCB_GT) ab += x[i];
                }
                am = warp_sum(am);
                ab = warp_sum(ab);
                if (k_mode == CST_BASE_SCB_SAMPLE) ab = lowsum;
                if (lane == 0) {
                    p.partials[2 * video] = am;
                    p.partials[2 * video + 1] = ab;
                }
            }
        }
        __syncthreads();
    }
    if (p.mode >= 0 && !p.fold_apart) fold_stats_if_last(p.ctl, p.partials, p.n_videos, p.n_m, p.n_b, p.stats, red);
}

static size_t score2_smem(int n_local, int cands_per_video, int nwarps) {
    return (size_t)CST2_TAB_SLOTS * sizeof(VidSlot) + (size_t)CST2_LIST_CAP * sizeof(uint4) +
           4 * CST2_ROWS * sizeof(RefA) + (CST2_ROWS * CST2_TOK_STRIDE + 8) * sizeof(u16) +
           2 * CST2_ROWS * sizeof(int) + (size_t)n_local * sizeof(double) + (size_t)cands_per_video * sizeof(double) +
           2 * (size_t)nwarps * sizeof(double);
}
