// Shared device helpers of the CIDEr-D reward kernels (sm_100a).
//
// Data layout in HBM / shared memory is described in DESIGN.md; the packed key
// format and the C ABI in include/cst_reward.h.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

typedef unsigned long long u64;
typedef unsigned int u32;

#define CST_FULL 0xffffffffu
#define CST_SIGMA 6.0          /* CiderD default sigma (train.py:121 builds CiderD with defaults) */
#define CST_GAUSS_HALF 32      /* LUT covers length deltas in [-32, 32] */
#define CST_GAUSS_N (2 * CST_GAUSS_HALF + 1)

// ---------------------------------------------------------------------------
// global df table (HBM, meant to stay L2-resident)
//
//   keys     [n_buckets][4] u64   open addressing over 32-byte buckets: one
//                                 256-bit load (one L2 sector) fetches a whole
//                                 bucket; a bucket with an empty slot ends a probe,
//                                 and so does a FULL bucket no insertion ever had to
//                                 step past -- marked without a bit of storage by the
//                                 order of its first two keys (key0 < key1; an
//                                 overflowed bucket holds them the other way round).
//                                 Most probes are for n-grams the corpus never saw;
//                                 without the mark 5-16 % of them walk on, with it 2-7 %,
//                                 and a warp waits for its longest walk
//   idfv     [n_buckets][4] f64   idf of the key in the same slot = lut[df]: ONE dependent
//                                 load after the bucket (df -> lut would be two L2 trips)
//   dfc      [n_buckets][4] u32   its document frequency (cold: parity probe, export)
//   lut      [n_documents + 1]    idf by document frequency, float64 computed by the
//                                 caller (bit-identical to the scorer's numpy value)
//   uni      [CST_UNI_N] f64      idf of unigram `token` directly (no hashing)
// ---------------------------------------------------------------------------
#define CST_BUCKET 4
#define CST_UNI_N 65535

struct DfView {
    const u64* keys;
    const double* idfv;
    const u32* dfc;
    const double* lut;
    const double* uni;
    u32 n_buckets;
    double unseen;   // idf of an n-gram absent from the table = lut[0]
};

// Fibonacci hash: bits 32..63 of key * phi.  Both tables (global df table,
// per-video shared table) are indexed from this one value.
__host__ __device__ __forceinline__ u32 cst_hash(u64 key) {
    return (u32)((key * 0x9E3779B97F4A7C15ull) >> 32);
}
// Lemire range reduction, valid for any capacity < 2^32.
__host__ __device__ __forceinline__ u32 cst_slot(u32 h, u32 cap) {
    return (u32)(((u64)h * (u64)cap) >> 32);
}

#ifdef __CUDACC__

// One bucket = 4 keys = 32 bytes = one sector, fetched by a single LDG.256.  Not allocated in L1:
// probes land all over a 36 MB array, and the little L1 an SM has left next to the per-video
// tables is better spent on the token rows and the unigram idf (round 1: 3.22 -> 3.11 ms).
__device__ __forceinline__ void ld_bucket(const u64* p, u64 k[4]) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
                 : "=l"(k[0]), "=l"(k[1]), "=l"(k[2]), "=l"(k[3])
                 : "l"(p));
}

// slot index of `key` (>= 0), or -1 if absent.  `b` is the home bucket, `k` its
// already loaded keys.  Overflow into the next bucket is rare (load factor 0.5).
// MARK: also stop at a full bucket whose first two keys are in ascending order (see the layout
// note above).  (The scoring kernel has its own compact form of this walk, v2_resolve.)
template <bool MARK = true>
__device__ __forceinline__ int bucket_resolve(const u64* __restrict__ keys, u32 nb, u32 b, u64 k[4], u64 key) {
    for (;;) {
#pragma unroll
        for (int j = 0; j < CST_BUCKET; ++j)
            if (k[j] == key) return (int)(b * CST_BUCKET + j);
        if (k[3] == 0ull) return -1;          // slots fill front to back: an empty tail ends the probe
        if (MARK && k[0] < k[1]) return -1;   // full, but nothing ever overflowed from here (build-time order)
        b = (b + 1 == nb) ? 0u : b + 1;
        ld_bucket(keys + (size_t)b * CST_BUCKET, k);
    }
}

__device__ __forceinline__ u32 lanemask_lt(int lane) { return (1u << lane) - 1u; }

// ---------------------------------------------------------------------------
// Row canonicalisation + n-gram keys for the df build and the n-gram parity probe (the scoring
// kernel has its own, leaner forms: v2_canon / v2_emit).  A row is handled by a group of G lanes:
// G = 32 (one row per warp) or G = 16 (two rows per warp, one per half); G and the token width are
// run-time values.
// Restates utils.py:135-152 (array_to_str) and compute_ciderdf.py:28-33
// (precook's k/i loops) on packed keys.
//   key[k]   packed (k+1)-gram starting at the lane's position, 0 if none
//   L        canonical token count of the group's row
//   returns  true if the row held an id outside [-1, 65534] before its end
// tokbuf: 32 u32 of warp-private shared memory (G = 16: 16 per half).
// ---------------------------------------------------------------------------
__device__ __forceinline__ int grp_sl(int lane, int G) { return lane & (G - 1); }      // position in the group
__device__ __forceinline__ int grp_shift(int lane, int G) { return lane & (32 - G); }  // first lane of the group
__device__ __forceinline__ u32 grp_mask(int lane, int G) {
    return (G == 32) ? CST_FULL : (0xffffu << grp_shift(lane, G));
}
// 32-bit mask over the row's (up to 32) raw positions: pa speaks for position sl, pb
// for position sl + 16 (G = 16 only; each lane holds two raw tokens there).
__device__ __forceinline__ u32 grp_vote(bool pa, bool pb, int lane, int G) {
    const u32 a = __ballot_sync(CST_FULL, pa);
    const u32 b = __ballot_sync(CST_FULL, pb);
    const int sh = grp_shift(lane, G);
    return (G == 32) ? a : (((a >> sh) & 0xffffu) | (((b >> sh) & 0xffffu) << 16));
}

// token `i` of a row of int16, int32 or int64 ids (`esz` = bytes per id); rows stream
// through once: evict-first
__device__ __forceinline__ long long ld_tok(const void* row, int esz, int i) {
    if (esz == 4) return (long long)__ldcs(reinterpret_cast<const int*>(row) + i);
    if (esz == 8) return (long long)__ldcs(reinterpret_cast<const long long*>(row) + i);
    return (long long)__ldcs(reinterpret_cast<const short*>(row) + i);
}

__device__ __forceinline__ bool canon_row(const void* row, int esz, int width, int use_eos, int lane, int G,
                                          u32* tokbuf, u64 key[4], int& L) {
    const int sl = grp_sl(lane, G);
    const bool half = (G == 16);
    const int pa = sl, pb = sl + 16;                       // raw positions this lane owns
    const bool ina = pa < width;
    const bool inb = half && pb < width;
    const long long ta = ina ? ld_tok(row, esz, pa) : 0ll;
    const long long tb = inb ? ld_tok(row, esz, pb) : 0ll;
    const bool bada = ina && (ta < -1 || ta > 65534);
    const bool badb = inb && (tb < -1 || tb > 65534);
    // -1 = padding: ends the row, never emitted
    const u32 tmask = grp_vote(ina && (ta == 0 || ta == -1 || bada), inb && (tb == 0 || tb == -1 || badb), lane, G);
    const u32 bmask = grp_vote(bada, badb, lane, G);
    const int end = tmask ? (__ffs(tmask) - 1) : width;    // first <end> (or bad id), else row width
    const u32 upto_end = (end >= 31) ? CST_FULL : ((2u << end) - 1u);
    const bool keepa = (pa < end && ta != 1) || (use_eos && pa == end && ina && ta == 0);
    const bool keepb = half && ((pb < end && tb != 1) || (use_eos && pb == end && inb && tb == 0));
    const u32 kmask = grp_vote(keepa, keepb, lane, G);
    L = __popc(kmask);
    const bool fits = L <= G;                              // G = 16: longer rows are redone by the caller
    u32* buf = tokbuf + (grp_shift(lane, G) & 16);
    if (fits) {
        if (keepa) buf[__popc(kmask & lanemask_lt(pa))] = (u32)ta + 1u;
        if (keepb) buf[__popc(kmask & lanemask_lt(pb))] = (u32)tb + 1u;
    }
    __syncwarp();
    const u32 e0 = (sl < L && fits) ? buf[sl] : 0u;
    __syncwarp();
    u32 e1 = __shfl_down_sync(CST_FULL, e0, 1, G);
    u32 e2 = __shfl_down_sync(CST_FULL, e0, 2, G);
    u32 e3 = __shfl_down_sync(CST_FULL, e0, 3, G);
    if (sl + 1 >= L || sl + 1 >= G) e1 = 0;
    if (sl + 2 >= L || sl + 2 >= G) e2 = 0;
    if (sl + 3 >= L || sl + 3 >= G) e3 = 0;
    const u64 k1 = (u64)e0;
    const u64 k2 = k1 | ((u64)e1 << 16);
    const u64 k3 = k2 | ((u64)e2 << 32);
    const u64 k4 = k3 | ((u64)e3 << 48);
    key[0] = k1;
    key[1] = e1 ? k2 : 0ull;
    key[2] = e2 ? k3 : 0ull;
    key[3] = e3 ? k4 : 0ull;
    return (bmask & upto_end) != 0u;
}

// Term frequency of each lane's n-gram inside its own row; the lowest lane of every set
// of equal n-grams is the leader (the dict entry precook would create at that position).
// One MATCH.ANY on the unigrams gives E1(i) = positions holding the same token as i; the
// higher orders follow from it without further matching:
//     E_k(i) = E_{k-1}(i) & (E_1(i + k - 1) >> (k - 1))
// (positions j whose tokens j .. j+k-1 equal tokens i .. i+k-1).  Positions past the
// end hold token code 0, which never equals a real token code, so partial n-grams drop
// out by themselves.  `only` >= 0: this item covers a single order (warp-uniform).
__device__ __forceinline__ void row_tf(const u64 key[4], int lane, int G, int only, int tf[4], bool lead[4]) {
    const int sh = grp_shift(lane, G);
    const int sl = grp_sl(lane, G);
    const u32 window = (G == 32) ? CST_FULL : 0xffffu;
    const u32 e1 = (__match_any_sync(CST_FULL, (u32)key[0]) >> sh) & window;   // group-relative positions
    u32 ek = e1;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (k > 0) {
            u32 nb = __shfl_down_sync(CST_FULL, e1, k, G);        // E1 of position sl + k
            if (sl + k >= G) nb = 0u;
            ek &= nb >> k;
        }
        const bool mine = (only < 0 || k == only);
        tf[k] = mine ? __popc(ek) : 0;
        lead[k] = mine && (key[k] != 0ull) && ((__ffs(ek) - 1) == sl);
    }
}

__device__ __forceinline__ double inv_norm_of(double sumsq) {
    return sumsq > 0.0 ? rsqrt(sumsq) : 0.0;
}

// ---------------------------------------------------------------------------
// per-video shared-memory table: union of the n-grams of (up to 32) refs.
//   key   packed n-gram, 0 = empty
//   P     bit r set  <=>  ref r of the chunk contains the n-gram
//   M     bit r set  <=>  ref r contains it more than once; the exact count then
//         lives under the synthetic key count_key(slot, r) (field P)
// ---------------------------------------------------------------------------
struct __align__(16) VidSlot {
    u64 key;
    u32 P;
    u32 M;
};

// Real keys always have a non-zero low 16-bit lane, synthetic ones a zero one.
__device__ __forceinline__ u64 count_key(u32 slot, u32 r) {
    return (1ull << 63) | ((u64)slot << 24) | ((u64)r << 16);
}

// The table capacity is any number (sized to the worst case of the staged refs, not
// rounded to a power of two: shared memory decides how many CTAs fit an SM).
__device__ __forceinline__ u32 vid_home(u32 h, u32 slots) { return __umulhi(h, slots); }

#endif  // __CUDACC__
