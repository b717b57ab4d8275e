// k_score2 -- the scoring kernel (included by cst_reward.cu after ScoreParams).
//
// CIDEr-D of every candidate against its video's references, then the baseline of utils.py:292-322.
// The kernel is organised around ONE idea: the expensive per-n-gram work -- the df probe into the
// L2-resident table, the insert / find in the per-video shared-memory table, the walk over the
// presence bits -- runs over a dense LIST of n-gram instances with every lane busy (round 1's kernel
// put a token position on a lane and unrolled the four orders per lane: a third of the lanes idled
// and the order bodies diverged).  One CTA of 8 warps per video, 6 CTAs per SM:
//
//   canon    warp per row, lane = raw position: ballot/popc compaction of utils.py:135-152
//            (array_to_str) -> canonical token codes (u16) + length per row in shared memory
//   emit     half-warp per row (full warp for rows longer than 16 tokens), lane = start position:
//            packed 1..4-gram keys, term frequency and "first occurrence" from one MATCH.ANY
//            (compute_ciderdf.py:19-34, precook); instance (order k, row, position) owns a list slot
//            known from the row lengths alone (order-major, see V2Scan), so no compaction pass and
//            no atomics; repeats inside a row leave their slot marked empty
//   process  thread per list slot: idf (direct load for unigrams, one 32-byte bucket otherwise);
//            references: insert into the video table, OR the row's presence bit; candidates: find,
//            walk the presence bits.  The slot is overwritten with {(tf idf)^2, idf^2 * inner}
//   sums     thread per (row, order): its list segment is contiguous, summed front to back ->
//            norms, numerators.  No floating-point atomics, no shuffle trees: deterministic.
//
// Shared memory is a fixed 37 KB whatever the row widths and counts.  References enter the video
// table in stages of at most CST2_LIST_CAP instances; a video with more than 32 references or more
// than CST2_TAB_BUDGET instances is scored in several chunks, each with a table of its own
// (CIDEr-D adds over references); candidates go through in batches likewise.
//
// Instantiations of the one source (MODE): raw rows; COOK -- the reference stages only, their
// result (header, norms and lengths, table) written out as the video's image; STAGED -- those
// stages replaced by one bulk asynchronous copy of the image (cst_refs_cook / cst_reward_cooked);
// EXCL -- the reference stages counting leave-one-out document frequencies (cst_df_sole_holders).
// The sizes below were settled by measurement (profiles/README.md, round 2).
#pragma once

#include "cst_rows.cuh"

#ifndef CST2_LIST_CAP
#define CST2_LIST_CAP 832            /* n-gram instances per batch (16 B each) */
#endif
#ifndef CST2_TAB_SLOTS
#define CST2_TAB_SLOTS 1193          /* per-video table (prime: double hashing); a chunk inserts <= CST2_TAB_BUDGET keys */
#endif
#ifndef CST2_TAB_BUDGET
#define CST2_TAB_BUDGET 1100         /* n-gram instances a table takes before a new chunk starts */
#endif
#define CST2_MODE_RAW 0
#define CST2_MODE_STAGED 1
#define CST2_MODE_COOK 2
#define CST2_MODE_EXCL 3    /* leave-one-out document frequencies (compute_scores.py:81-92), see k_score2 */
#define CST2_IMG_MAGIC 0x43535432u   /* "CST2": layout version of a staged-reference image */
#ifndef CST2_WARPS
#define CST2_WARPS 8
#endif
#ifndef CST2_CTAS
#define CST2_CTAS 6
#endif

#ifndef CST2_PREFETCH
#define CST2_PREFETCH 1     /* L2 prefetch of the rows a later CTA will start on (measured: 2 % of the step) */
#endif
struct __align__(16) RefA {          // what a presence bit needs of its reference, one 16-byte load
    double inv;                      // 1 / norm of the ref's order-k vector (0 if the norm is 0)
    int len;                         // the scorer's "length": number of bigrams
    int pad;
};


// which of the bucket's four keys equals `key` (-1: none), without branches
__device__ __forceinline__ int v2_match4(const u64 kk[4], u64 key) {
    int j = -1;
    if (kk[0] == key) j = 0;
    if (kk[1] == key) j = 1;
    if (kk[2] == key) j = 2;
    if (kk[3] == key) j = 3;
    return j;
}

// Slot of `key` in the df table (-1: absent) given its home bucket `b`, already loaded into kk.  A
// probe ends at a bucket with a free slot, or at a full one nothing overflowed from (its first two
// keys ascend).
__device__ __forceinline__ int v2_resolve(const DfView& t, u64 key, u32 b, u64 kk[4]) {
    int j = v2_match4(kk, key);
    bool more = j < 0 && kk[3] != 0ull && !(kk[0] < kk[1]);
#pragma unroll 1
    while (more) {                                             // a few percent of the probes
        b = (b + 1 == t.n_buckets) ? 0u : b + 1;
        ld_bucket(t.keys + (size_t)b * CST_BUCKET, kk);
        j = v2_match4(kk, key);
        more = j < 0 && kk[3] != 0ull && !(kk[0] < kk[1]);
    }
    return (j >= 0) ? (int)(b * CST_BUCKET) + j : -1;
}


// One pass over the list: thread per slot.
//   is_r:  publish the instance into the video table (presence bit of its row; exact count under a
//          synthetic key when the row holds the n-gram more than once), slot <- {(tf idf)^2, 0}
//   else:  find it, walk the presence bits, slot <- {(tf idf)^2, idf^2 * sum_r min(tf_h, tf_r) tf_r inv_r g_r}
// (Two slots per thread and round, so that their df-table loads travel together, was measured: slower.)
// row_off: references already in the table before this sub-stage (the presence bit of row r is
// row_off + r).  EXCL: every n-gram is published, with its slot in the df table in the M field.
template <int NT, u32 T, bool BOUNDED, bool EXCL>
__device__ __forceinline__ void v2_process(const ScoreParams& p, bool is_r, uint4* list, int n_ent, int base1,
                                           int row_off, VidSlot* tab, const RefA* refA, const int* sL,
                                           const double* s_gauss) {
#pragma unroll 1
    for (int i = (int)threadIdx.x; i < n_ent; i += NT) {
        const uint4 e = list[i];
        const u64 key = (u64)e.x | ((u64)e.y << 32);
        double w2 = 0.0, num = 0.0;
        if (e.z != 0u) {                                       // (0: a repeat of an n-gram inside its row)
            const int row = (int)(e.z & 31u);
            const int k = (int)((e.z >> 5) & 3u);
            const int tf = (int)(e.z >> 8);
            const u32 h = cst_hash(key);
            double idf;
            int slot = -1;
            if (!EXCL && i < base1) {                          // unigram: token codes are 1 .. 65535
                idf = __ldg(p.df.uni + (e.x - 1u));
            } else {
                u64 kk[4];
                const u32 b = cst_slot(h, p.df.n_buckets);
                ld_bucket(p.df.keys + (size_t)b * CST_BUCKET, kk);
                slot = v2_resolve(p.df, key, b, kk);
                idf = (slot >= 0) ? __ldg(p.df.idfv + slot) : p.df.unseen;
            }
            const double w = (double)tf * idf;                 // counts2vec: vec = tf * idf
            w2 = w * w;                                        // norm += vec ** 2
            // an n-gram every document holds has idf = 0: exact zeros on both sides, no table entry
            if (EXCL) {
                const u32 idx = v2_insert<T>(tab, h, key);
                atomicOr(&tab[idx].P, 1u << (row_off + row));
                tab[idx].M = (u32)(slot + 1);                  // (every holder stores the same value)
            } else if (idf != 0.0) {
                if (is_r) {
                    const u32 idx = v2_insert<T>(tab, h, key);
                    const u32 bit = 1u << (row_off + row);
                    atomicOr(&tab[idx].P, bit);
                    if (tf > 1) {
                        atomicOr(&tab[idx].M, bit);
                        const u64 ck = count_key(idx, (u32)(row_off + row));
                        const u32 j = v2_insert<T>(tab, cst_hash(ck), ck);
                        tab[j].P = (u32)tf;
                    }
                } else {
                    const int idx = v2_find<T, BOUNDED>(tab, h, key);
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
                            const int j = v2_find<T, BOUNDED>(tab, cst_hash(ck), ck);
                            const int c = (j >= 0) ? (int)tab[j].P : 1;
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

__device__ __forceinline__ void v2_prefetch_l2(const void* p, u32 bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

// shared-memory layout: everything the hot loops touch sits at a compile-time offset.  The first
// three pieces -- header, reference norms / lengths, table -- are what the reference stages of a
// video leave behind: contiguous, so that a staged-reference image is one copy in either direction.
#define CST2_CTOK_BYTES ((CST2_ROWS * CST2_TOK_STRIDE + 8) * sizeof(u16))
#define CST2_OFF_REFA ((size_t)16)                                   /* after hdr: magic, T, refs, use_eos */
#define CST2_OFF_TAB (CST2_OFF_REFA + 4 * CST2_ROWS * sizeof(RefA))
#define CST2_IMG_BYTES (CST2_OFF_TAB + (size_t)CST2_TAB_SLOTS * sizeof(VidSlot))
#define CST2_OFF_LIST CST2_IMG_BYTES
#define CST2_OFF_CTOK (CST2_OFF_LIST + (size_t)CST2_LIST_CAP * sizeof(uint4))
#define CST2_OFF_SL (CST2_OFF_CTOK + CST2_CTOK_BYTES)
#define CST2_OFF_SC (CST2_OFF_SL + CST2_ROWS * sizeof(int))
static_assert(CST2_IMG_BYTES % 16 == 0 && CST2_OFF_TAB % 16 == 0, "the image must stay 16-byte aligned");

// -DCST2_TIMING: cycles per warp and stage, added up in the scratch buffer (tools/phase_timing2.py;
// the counters themselves cost time, read the numbers as proportions)
#ifdef CST2_TIMING
#define CST2_TICK(id)                        \
    do {                                     \
        const long long now_ = clock64();    \
        acc_[(id)] += now_ - tick_;          \
        tick_ = now_;                        \
    } while (0)
#else
#define CST2_TICK(id)
#endif

// Stages of a video.  A reference stage takes up to 32 rows / CST2_LIST_CAP instances into the video
// table; further reference stages follow while the chunk has room (32 presence bits,
// CST2_TAB_BUDGET instances).  Then every candidate batch meets the chunk, then the next chunk
// starts on a cleared table, if the video has more references.  Both kinds of stage run through ONE
// copy of the code:   load + canonicalise | scan + emit | process   with a CTA barrier after each;
// a reference stage ends with the norms of its rows, a candidate stage with the sums of its
// segments.  A video of the training shapes is one stage of each kind.
// ESZ: bytes per token id when the launch fixes them for both row sets (4: int32 rows), 0: run-time
template <int MODE, int NW, int CTAS, int ESZ>
__global__ void __launch_bounds__(NW * 32, CTAS) k_score2(const ScoreParams p) {
    constexpr int NT = NW * 32;
    constexpr u32 T = CST2_TAB_SLOTS;
    constexpr int PER = (CST2_ROWS + NW - 1) / NW;      // rows a warp canonicalises per batch
    constexpr bool STAGED = MODE == CST2_MODE_STAGED, EXCL = MODE == CST2_MODE_EXCL;
    constexpr bool COOK = MODE == CST2_MODE_COOK || EXCL;      // reference stages only
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ double s_gauss[CST_GAUSS_N];
    __shared__ unsigned long long s_img_bar;            // staged references: completion of the image fetch
    const int lane = threadIdx.x & 31;
    // (through a shuffle so that the compiler knows the value is the same in all lanes of the warp)
    const int warp = __shfl_sync(CST_FULL, (int)(threadIdx.x >> 5), 0);
    const int S = p.cands_per_video;
    // (both computed by the host: the compiler would rather recompute than keep them)
    const int n_local = p.n_local;                      // candidate rows per video: S, or 2 S with greedy rows
    const int cand_rows = p.cand_rows;                  // per batch, in even shares (40 rows go as 20 + 20, not 32 + 8)

    u32* hdr = reinterpret_cast<u32*>(smem);                                  // [4]
    RefA* refA = reinterpret_cast<RefA*>(smem + CST2_OFF_REFA);               // [4][32]
    VidSlot* tab = reinterpret_cast<VidSlot*>(smem + CST2_OFF_TAB);           // [T]
    uint4* list = reinterpret_cast<uint4*>(smem + CST2_OFF_LIST);             // [CST2_LIST_CAP]
    u16* ctok = reinterpret_cast<u16*>(smem + CST2_OFF_CTOK);                 // [32][32] (+ tail pad)
    int* sL = reinterpret_cast<int*>(smem + CST2_OFF_SL);                     // [32] L | bad << 8 of the batch's rows
    double* sc = reinterpret_cast<double*>(smem + CST2_OFF_SC);               // [n_local]
    double* low = sc + n_local;                                               // [S]
    double* red = low + S;                                                    // [2 * NW]

#pragma unroll 1
    for (int i = threadIdx.x; i < CST_GAUSS_N; i += NT) s_gauss[i] = p.gauss[i];
    const u32 pitch_r = (u32)p.ref_stride * (u32)p.ref_esz;     // (the host checked that both fit 32 bits)
    const u32 pitch_c = (u32)p.cand_stride * (u32)p.cand_esz;
    u32 img_parity = 0;
    if (STAGED) {
        if (threadIdx.x == 0) mbar_init(&s_img_bar, 1);
        __syncthreads();
    }

#pragma unroll 1
    for (int64_t video = blockIdx.x; video < p.n_videos; video += gridDim.x) {
        int64_t r0 = 0, r1 = 0;
        if (!STAGED) {
            if (p.ref_offsets) {
                r0 = p.ref_offsets[video];
                r1 = p.ref_offsets[video + 1];
            } else {
                r0 = video * p.refs_per_video;
                r1 = r0 + p.refs_per_video;
            }
        }
        int R = (int)(r1 - r0);
        bool bad_image = false;
        if (STAGED && threadIdx.x == 0) {
            // one bulk copy by the copy engine: header, norms / lengths and table land while the warps
            // load and canonicalise the candidates; the wait is just before the table's first use
            const int64_t img = p.cooked_index ? p.cooked_index[video] : video;
            bulk_fetch(hdr, p.cooked + (size_t)img * (size_t)p.cooked_stride, (u32)CST2_IMG_BYTES, &s_img_bar);
        }
#if CST2_PREFETCH
        // CTAs start in blockIdx order, so the video `prefetch_ahead` (= resident CTAs) further on is
        // the one a CTA will start on about when this one ends: pull its rows into L2 now, and that
        // CTA's first loads -- a DRAM round trip nothing else in a video can hide -- become L2 hits.
        if (threadIdx.x == 0 && !p.ref_offsets) {
            const int64_t v2 = video + p.prefetch_ahead;
            if (v2 < p.n_videos) {
                if (!STAGED) {
                    const char* a = reinterpret_cast<const char*>(p.refs) + (size_t)(v2 * p.refs_per_video) * pitch_r;
                    const u32 na = ((u32)p.refs_per_video * pitch_r) & ~15u;
                    if (na && !(reinterpret_cast<uintptr_t>(a) & 15u)) v2_prefetch_l2(a, na);
                }
                if (!COOK) {
                    const char* b = reinterpret_cast<const char*>(p.cand) + (size_t)(v2 * S) * pitch_c;
                    const u32 nb = ((u32)S * pitch_c) & ~15u;
                    if (nb && !(reinterpret_cast<uintptr_t>(b) & 15u)) v2_prefetch_l2(b, nb);
                }
            }
        }
#endif
#ifdef CST2_TIMING
        long long acc_[16];
        for (int i = 0; i < 16; ++i) acc_[i] = 0;
        long long tick_ = clock64();
#endif
#pragma unroll 1
        for (int c = threadIdx.x; c < n_local; c += NT) sc[c] = 0.0;   // (stage barriers lie before its first use)
        u32 bad_rows = 0;

        int64_t rbase = r0;
        int cb = STAGED ? 0 : -1;       // < 0: the stage takes references from rbase on; else candidates from row cb
        int chunk_rows = 0, chunk_inst = 0;              // references / instances in the table so far
        bool img_ready = !STAGED;
        bool first_chunk = true;
        bool more = STAGED ? n_local > 0 : rbase < r1;
#pragma unroll 1
        while (more) {
            const bool is_r = COOK || (!STAGED && cb < 0);
            // ---- load + canonicalise the stage's rows: warp w takes rows w, w + NW, ...; all loads are
            // issued before any is used ----
            const int room = CST2_ROWS - chunk_rows;
            const int n_try = is_r ? (int)((r1 - rbase) < room ? (r1 - rbase) : room)
                                   : ((n_local - cb) < cand_rows ? (n_local - cb) : cand_rows);
            {
                const int esz = ESZ ? ESZ : (is_r ? p.ref_esz : p.cand_esz);
                const u32 pitch = is_r ? pitch_r : pitch_c;
                const char* base = is_r ? reinterpret_cast<const char*>(p.refs) + (size_t)rbase * pitch_r
                                        : reinterpret_cast<const char*>(p.cand) + (size_t)(video * S) * pitch_c;
                const char* base2 = reinterpret_cast<const char*>(p.greedy) + (size_t)(video * S) * pitch_c;
                const int first = is_r ? 0 : cb;
                const int split = is_r ? 0x7fffffff : S;                     // sampled rows first, then greedy
                const bool in = lane < (is_r ? p.ref_width : p.cand_width);
                const u32 lane_off = (u32)(lane * esz);
                u32 u[PER];
#pragma unroll
                for (int j = 0; j < PER; ++j) {
                    const int r = warp + j * NW;
                    int c = first + r;
                    const char* src = base;
                    if (c >= split) { c -= split; src = base2; }
                    const char* tok = src + ((unsigned long long)(u32)c * (unsigned long long)pitch + lane_off);
                    u32 v = 0u;
                    if (r < n_try && in) {
                        if (esz == 4) {
                            v = (u32)__ldcs(reinterpret_cast<const int*>(tok)) + 1u;
                        } else if (esz == 8) {
                            const long long w = __ldcs(reinterpret_cast<const long long*>(tok));
                            v = (w < -1 || w > 65534) ? 0xffffffffu : (u32)w + 1u;           // out of range: bad
                        } else {
                            v = (u32)(int)__ldcs(reinterpret_cast<const short*>(tok)) + 1u;
                        }
                    }
                    u[j] = v;
                }
                if (!STAGED && is_r && chunk_rows == 0) {                    // a new chunk: empty table
                    uint4* z = reinterpret_cast<uint4*>(tab);
#pragma unroll 2
                    for (u32 i = threadIdx.x; i < T; i += NT) z[i] = make_uint4(0, 0, 0, 0);
                }
#pragma unroll
                for (int j = 0; j < PER; ++j) {
                    const int r = warp + j * NW;
                    if (r < n_try) {
                        const int v = v2_canon(u[j], p.use_eos, lane, ctok + r * CST2_TOK_STRIDE);
                        if (lane == 0) sL[r] = v;
                    }
                }
            }
            CST2_TICK(is_r ? 0 : 6);
            __syncthreads();
            CST2_TICK(is_r ? 1 : 7);
            // ---- scan + emit (the budgets leave room for one row of 32 tokens: at least one row fits) ----
            const int left = CST2_TAB_BUDGET - chunk_inst;
            const V2Scan sb = v2_scan(sL, n_try, (is_r && left < CST2_LIST_CAP) ? left : CST2_LIST_CAP, lane);
            const int nfit = sb.nfit;
            // rows with ids outside [-1, 65534]: every reference row once, candidates in the first chunk
            if (warp == 0 && lane < nfit && (is_r || first_chunk)) bad_rows += (u32)(sL[lane] >> 8);
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
            CST2_TICK(is_r ? 2 : 8);
            __syncthreads();
            CST2_TICK(is_r ? 3 : 9);
            if (STAGED && !img_ready) {
                // the reference half was run once, by cst_refs_cook: its image must have landed by now
                while (!mbar_try_wait(&s_img_bar, img_parity)) {}
                img_parity ^= 1u;
                img_ready = true;
                bad_image = hdr[0] != CST2_IMG_MAGIC || hdr[1] != T || hdr[3] != (u32)p.use_eos || hdr[2] > 32u;
                R = bad_image ? 0 : (int)hdr[2];
                chunk_rows = R;
            }
            // ---- process ----
            if (!(STAGED && bad_image))
                v2_process<NT, T, STAGED, EXCL>(p, is_r, list, sb.n_ent, sb.base1, chunk_rows, tab, refA, sL, s_gauss);
            CST2_TICK(is_r ? 4 : 10);
            __syncthreads();
            CST2_TICK(is_r ? 5 : 11);
            // ---- thread per (row, order): its list segment, front to back -> reference norms, or the
            // candidate's norm and numerator.  No barrier follows: the next stage starts by overwriting
            // ctok / sL, which nothing here reads (the lengths are in sb), and reaches the list only
            // after its own first barrier.
            if (warp * 8 < nfit) {
                const int row = (threadIdx.x >> 2) & 31, k = threadIdx.x & 3;
                const int L = __shfl_sync(CST_FULL, sb.L, row);
                const int seg0 = v2_seg(sb, row, k);
                double a = 0.0, b = 0.0;
                if (row < nfit) {
                    const int n = (L > k) ? L - k : 0;
                    const double2* seg = reinterpret_cast<const double2*>(list + seg0);
#pragma unroll 2
                    for (int j = 0; j < n; ++j) {
                        const double2 v = seg[j];
                        a += v.x;
                        b += v.y;
                    }
                }
                const double inv = inv_norm_of(a);
                if (is_r) {
                    if (row < nfit) {
                        RefA ra;
                        ra.inv = inv;
                        ra.len = (L > 1) ? L - 1 : 0;                        // "length" = bigram count
                        ra.pad = 0;
                        refA[k * CST2_ROWS + chunk_rows + row] = ra;
                    }
                } else {
                    double val = b * inv;
                    val += __shfl_xor_sync(CST_FULL, val, 1);
                    val += __shfl_xor_sync(CST_FULL, val, 2);
                    if (row < nfit && k == 0) sc[cb + row] += val;
                }
            }
            CST2_TICK(12);
            if (is_r) {
                chunk_rows += nfit;
                chunk_inst += sb.n_ent;
                rbase += nfit;
                // another reference stage into the same table while a row of any length still fits
                if (!(rbase < r1 && chunk_rows < CST2_ROWS && CST2_TAB_BUDGET - chunk_inst >= 4 * 32 - 6)) {
                    if (COOK) more = false;         // the image holds ONE chunk (the host checks rbase == r1)
                    cb = 0;
                }
            } else {
                cb += nfit;
                if (cb >= n_local) {                // the chunk has met every candidate: next chunk, if any
                    more = !STAGED && rbase < r1;
                    cb = -1;
                    chunk_rows = 0;
                    chunk_inst = 0;
                    first_chunk = false;
                }
            }
        }
        if (STAGED && !img_ready) {                 // (no candidate stage ran: still take the fetch off the barrier)
            while (!mbar_try_wait(&s_img_bar, img_parity)) {}
            img_parity ^= 1u;
        }
        if (EXCL) {
            // Leave-one-out document frequencies.  compute_scores.py:81-92 scores caption i of every video
            // against the others with a df recomputed WITHOUT caption i; that df differs from the df of
            // all captions only where caption i is the video's sole holder of an n-gram:
            //     df_i(g) = df_all(g) - #{videos whose caption i alone holds g}.
            // The table knows the holders of each n-gram (P): count the sole ones, per (caption, df slot).
            __syncthreads();
            if (rbase < r1) {                       // more captions than one table takes: the caller falls back
                if (threadIdx.x == 0 && p.bad_out) atomicAdd(p.bad_out, 1u);
            } else {
#pragma unroll 1
                for (u32 i = threadIdx.x; i < T; i += NT) {
                    const u32 P = tab[i].P, g = tab[i].M;
                    if (tab[i].key != 0ull && g != 0u && (P & (P - 1u)) == 0u)
                        atomicAdd(p.excl + (size_t)(__ffs(P) - 1) * p.excl_stride + (g - 1u), 1u);
                }
            }
            __syncthreads();
            continue;
        }
        if (COOK) {
            // what the reference stages left in shared memory is the video's image.  A video whose
            // references did not fit one chunk cannot be staged: invalid header (scores from it are NaN).
            __syncthreads();
            const bool fits = rbase >= r1;
            if (R == 0) {                           // no references: an empty table
                uint4* z = reinterpret_cast<uint4*>(tab);
#pragma unroll 1
                for (u32 i = threadIdx.x; i < T; i += NT) z[i] = make_uint4(0, 0, 0, 0);
            }
            if (threadIdx.x == 0) {
                if (!fits) ++bad_rows;
                hdr[0] = fits ? CST2_IMG_MAGIC : 0u;
                hdr[1] = T;
                hdr[2] = (u32)R;
                hdr[3] = (u32)p.use_eos;
            }
            if (bad_rows && p.bad_out) atomicAdd(p.bad_out, bad_rows);
            __syncthreads();
            const uint4* img = reinterpret_cast<const uint4*>(smem);
            uint4* out = reinterpret_cast<uint4*>(p.cook_out + (size_t)video * (size_t)p.cooked_stride);
#pragma unroll 1
            for (u32 i = threadIdx.x; i < (u32)(CST2_IMG_BYTES / 16); i += NT) out[i] = img[i];
            __syncthreads();
            continue;
        }
        if (STAGED && bad_image && threadIdx.x == 0) ++bad_rows;
        if (bad_rows && p.bad_out) atomicAdd(p.bad_out, bad_rows);
        // (the GT-consensus score of this lane's caption, fetched ahead of the barrier it would otherwise wait behind)
        double x_gt = 0.0;
        if (p.mode == CST_BASE_SCB_GT && warp == 0 && lane < S && S <= 32) x_gt = __ldg(p.bcmr + video * S + lane);
        __syncthreads();

        // ---- scores, baseline (utils.py:292-322), rewards, per-video statistics ----
        // np.mean over the 4 orders, / len(refs), * 10 (compute_cider)
        double inv_R = (R > 0) ? __drcp_rn((double)R) : 0.0;
        if (STAGED && bad_image) inv_R = __longlong_as_double(0x7ff8000000000000ll);   // not this build's image: NaN, loudly
        const int k_mode = p.mode;
        if (S <= 32) {
            // up to 32 captions per video (every training shape): warp 0 finishes the video by itself,
            // lane = caption -- no further barrier, no shared-memory round trips; the sums below are the
            // CTA-wide path's, term for term
            if (warp == 0) {
                const bool on = lane < S;
                double sv = 0.0, gv = 0.0;
                if (on) {
                    sv = (sc[lane] * 0.25) * inv_R * 10.0;
                    if (p.scores) p.scores[video * S + lane] = sv;
                    if (n_local > S) {
                        gv = (sc[S + lane] * 0.25) * inv_R * 10.0;
                        if (p.scores) p.scores[p.n_cand + video * S + lane] = gv;
                    }
                }
                if (k_mode >= 0) {
                    const bool scb = (k_mode == CST_BASE_SCB_GT || k_mode == CST_BASE_SCB_SAMPLE);
                    double x = sv;
                    if (k_mode == CST_BASE_SCB_GT) x = x_gt;
                    double lowsum = 0.0;
                    if (scb) {
                        // k lowest of the row (utils.py:294-312): rank by counting, ties by index (= stable
                        // sort); k >= S keeps them all
                        int rank = 0;
                        if (p.scb_k < S) {
#pragma unroll 1
                            for (int j = 0; j < S; ++j) {
                                const double xj = __shfl_sync(CST_FULL, x, j);
                                rank += (xj < x || (xj == x && j < lane)) ? 1 : 0;
                            }
                        }
                        lowsum = warp_sum((on && rank < p.scb_k) ? x : 0.0);
                    }
                    const double base = scb ? lowsum * __drcp_rn((double)p.scb_k) : 0.0;
                    if (on) p.reward[video * S + lane] = (float)((k_mode == CST_BASE_GREEDY) ? sv - gv : sv - base);
                    const double am = warp_sum(on ? sv : 0.0);
                    double ab = 0.0;
                    if (k_mode == CST_BASE_GREEDY) ab = warp_sum(on ? gv : 0.0);
                    if (k_mode == CST_BASE_SCB_GT) ab = warp_sum(on ? x : 0.0);
                    if (k_mode == CST_BASE_SCB_SAMPLE) ab = lowsum;
                    if (lane == 0) {
                        p.partials[2 * video] = am;
                        p.partials[2 * video + 1] = ab;
                    }
                }
            }
        } else {
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
        }
#ifdef CST2_TIMING
        CST2_TICK(13);
        if (lane == 0 && p.ctl)
            for (int i = 0; i < 16; ++i)
                if (acc_[i]) atomicAdd(reinterpret_cast<unsigned long long*>(p.ctl) + 8 + i, (unsigned long long)acc_[i]);
#endif
        __syncthreads();
    }
    if (p.mode >= 0 && !p.fold_apart) fold_stats_if_last(p.ctl, p.partials, p.n_videos, p.n_m, p.n_b, p.stats, red);
}

static size_t score2_smem(int n_local, int cands_per_video, int nwarps) {
    return CST2_OFF_SC + (size_t)n_local * sizeof(double) + (size_t)cands_per_video * sizeof(double) +
           2 * (size_t)nwarps * sizeof(double);
}
