// The other two reward scorers the reference can be run with (train.py:119-124, --eval_metric):
// sentence BLEU (Bleu_4 is what the reward reads, utils.py:210-218, 276-282) and ROUGE-L, on the same
// token rows, canonicalisation and candidate -> video map as the CIDEr-D path.
// Both restate pycocoevalcap (bleu/bleu_scorer.py, rouge/rouge.py), which the reference imports from an
// un-vendored checkout; METEOR there is a Java program and has no counterpart here.
//
// One CTA of 8 warps per video, like k_score2, built from its helpers (cst_rows.cuh):
//
//   k_bleu   "the candidates own the table, the references stream through it".  BLEU needs, per candidate
//            n-gram, the LARGEST count of that n-gram in any one reference -- a maximum, not a sum, so the
//            reference chunks of k_score2 would not add up.  Turned round it needs no chunks at all: the
//            n-grams of a batch of candidates (<= 32 rows, <= 832 instances) go into the shared-memory
//            table first, then reference rows pass through in stages of any number, each instance
//            raising the maximum of the slot it finds (most find nothing: one probe), and the closest
//            reference length per candidate is kept on the way.  Integer work end to end -- counts,
//            maxima and clipped sums are exact -- followed by four pow() and one exp() per candidate.
//   k_rouge  LCS of every (candidate, reference) pair with the bit-vector recurrence (a reference of up to
//            32 tokens is one machine word): V' = (V + (V & M)) | (V & ~M), LCS = zeros of V, where M marks
//            the reference positions holding the candidate's token.  The masks come from a small
//            shared-memory table keyed by (token, reference), filled once per batch of references; a
//            thread owns a pair.  Precision / recall maxima are integer / IEEE-ordered atomics, the
//            F-measure is evaluated in the scorer's own operation order without contraction: bit-exact.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "cst_common.cuh"
#include "cst_internal.h"
#include "cst_rows.cuh"

#define MT_WARPS 8
#define MT_THREADS (MT_WARPS * 32)
#define MT_LIST_CAP 832
#define MT_CTOK_U16 (CST2_ROWS * CST2_TOK_STRIDE + 8)

struct RowsView {
    const void* cand;
    const void* greedy;      // second row set of the same geometry (scored after the first), or NULL
    const void* refs;
    const int64_t* ref_offsets;
    int64_t n_cand, n_videos;
    int cand_esz, ref_esz, cand_width, ref_width;
    u32 pitch_c, pitch_r;
    int refs_per_video, cands_per_video, use_eos;
    u32* bad_out;
};

// rows [first, first + n) of one set -> canonical token codes + lengths (warp per row, all of a warp's
// loads issued before the first is used).  Candidate rows of a video: local index c < split is a
// sampled row, c >= split the greedy row c - split.
#define MT_PER ((CST2_ROWS + MT_WARPS - 1) / MT_WARPS)
#ifndef BL_LOADS_FIRST
#define BL_LOADS_FIRST 0   /* measured: k_bleu is faster with the plain loop (1.49 vs 1.53 ms) */
#endif
#ifndef BL_BLOOM
#define BL_BLOOM 1
#endif
#ifndef BL_PREFETCH
#define BL_PREFETCH 0      /* measured: no gain for k_bleu (1.49 vs 1.51 ms); k_rouge: 0.95 -> 0.92 ms */
#endif
template <bool FIRST = true>
__device__ __forceinline__ void mt_load_rows(const char* base, const char* base2, int split, int first, int n,
                                             u32 pitch, int esz, int width, int use_eos, u16* ctok, int* sL,
                                             int warp, int lane) {
    if (!FIRST) {
#pragma unroll 2
        for (int r = warp; r < n; r += MT_WARPS) {
            int c = first + r;
            const char* src = base;
            if (c >= split) {
                c -= split;
                src = base2;
            }
            const char* tok = src + ((unsigned long long)(u32)c * (unsigned long long)pitch + (u32)(lane * esz));
            u32 v = 0u;
            if (lane < width) {
                if (esz == 4) {
                    v = (u32)__ldcs(reinterpret_cast<const int*>(tok)) + 1u;
                } else if (esz == 8) {
                    const long long w = __ldcs(reinterpret_cast<const long long*>(tok));
                    v = (w < -1 || w > 65534) ? 0xffffffffu : (u32)w + 1u;
                } else {
                    v = (u32)(int)__ldcs(reinterpret_cast<const short*>(tok)) + 1u;
                }
            }
            const int res = v2_canon(v, use_eos, lane, ctok + r * CST2_TOK_STRIDE);
            if (lane == 0) sL[r] = res;
        }
        return;
    }
    u32 u[MT_PER];
    const bool in = lane < width;
    const u32 lane_off = (u32)(lane * esz);
#pragma unroll
    for (int j = 0; j < MT_PER; ++j) {
        const int r = warp + j * MT_WARPS;
        int c = first + r;
        const char* src = base;
        if (c >= split) {
            c -= split;
            src = base2;
        }
        const char* tok = src + ((unsigned long long)(u32)c * (unsigned long long)pitch + lane_off);
        u32 v = 0u;
        if (r < n && in) {
            if (esz == 4) {
                v = (u32)__ldcs(reinterpret_cast<const int*>(tok)) + 1u;
            } else if (esz == 8) {
                const long long w = __ldcs(reinterpret_cast<const long long*>(tok));
                v = (w < -1 || w > 65534) ? 0xffffffffu : (u32)w + 1u;
            } else {
                v = (u32)(int)__ldcs(reinterpret_cast<const short*>(tok)) + 1u;
            }
        }
        u[j] = v;
    }
#pragma unroll
    for (int j = 0; j < MT_PER; ++j) {
        const int r = warp + j * MT_WARPS;
        if (r < n) {
            const int res = v2_canon(u[j], use_eos, lane, ctok + r * CST2_TOK_STRIDE);
            if (lane == 0) sL[r] = res;
        }
    }
}

// The rows of the video this CTA takes next (grid-stride) into L2 while it works on the current one:
// the first loads of a video are a DRAM round trip nothing else in it can hide.
__device__ __forceinline__ void mt_prefetch_video(const RowsView& p, int64_t v) {
    if (v >= p.n_videos || p.ref_offsets) return;
    const char* a = reinterpret_cast<const char*>(p.refs) + (size_t)(v * p.refs_per_video) * p.pitch_r;
    const u32 na = ((u32)p.refs_per_video * p.pitch_r) & ~15u;
    if (na && !(reinterpret_cast<uintptr_t>(a) & 15u))
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a), "r"(na) : "memory");
    const u32 nb = ((u32)p.cands_per_video * p.pitch_c) & ~15u;
    const char* b = reinterpret_cast<const char*>(p.cand) + (size_t)(v * p.cands_per_video) * p.pitch_c;
    if (nb && !(reinterpret_cast<uintptr_t>(b) & 15u))
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(b), "r"(nb) : "memory");
    if (p.greedy) {
        const char* g = reinterpret_cast<const char*>(p.greedy) + (size_t)(v * p.cands_per_video) * p.pitch_c;
        if (nb && !(reinterpret_cast<uintptr_t>(g) & 15u))
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(g), "r"(nb) : "memory");
    }
}

// n-gram instances of the batch's rows into `list` (see v2_emit): two short rows per warp
__device__ __forceinline__ void mt_emit_rows(const u16* ctok, const V2Scan& sb, uint4* list, int warp, int lane) {
    const int nfit = sb.nfit;
#pragma unroll 1
    for (int pair = warp; 2 * pair < nfit; pair += MT_WARPS) {
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
}

__device__ __forceinline__ void mt_video_refs(const RowsView& p, int64_t video, int64_t& r0, int64_t& r1) {
    if (p.ref_offsets) {
        r0 = p.ref_offsets[video];
        r1 = p.ref_offsets[video + 1];
    } else {
        r0 = video * p.refs_per_video;
        r1 = r0 + p.refs_per_video;
    }
}

// ---------------------------------------------------------------------------
// BLEU
// ---------------------------------------------------------------------------
// Shared memory, 37 KB so that six CTAs share an SM (at four the kernel waited on latency with half the
// issue slots idle): the table, ONE list that candidates and references take turns in, the candidates'
// instances boiled down to 4 bytes each (table slot, row, order, count) for the pass that follows the
// references, one token buffer.  (Measured and dropped: no list at all -- the lanes that form a row's
// n-grams insert / look them up themselves, four barriers per video instead of six and a table twice the
// size: 1.59 ms against 1.49; half of those lanes idle and each takes its four orders in turn.)
#define BL_TAB_SLOTS 1103u            /* prime; holds the <= MT_LIST_CAP instances of one candidate batch */
#define BL_OFF_LIST ((size_t)BL_TAB_SLOTS * sizeof(VidSlot))
#define BL_OFF_CLIST (BL_OFF_LIST + (size_t)MT_LIST_CAP * sizeof(uint4))
#define BL_OFF_CTOK (BL_OFF_CLIST + (size_t)MT_LIST_CAP * sizeof(u32))
#define BL_SMEM (BL_OFF_CTOK + MT_CTOK_U16 * sizeof(u16))
#define BL_NONE 0xffffffffu

// bleu_out [4][n_total] (order-major: row 3 is the BLEU-4 vector the reward reads), comps_out [n_total][6]
// = testlen, reflen, correct[4] (what the corpus-level score adds up; guess[k] = max(testlen - k, 0)).
template <int ESZ>
__global__ void __launch_bounds__(MT_THREADS, 6) k_bleu(const RowsView p, double* __restrict__ bleu_out,
                                                        int* __restrict__ comps_out) {
    constexpr u32 T = BL_TAB_SLOTS;
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ int sL[CST2_ROWS];
    __shared__ int s_correct[CST2_ROWS][4];          // zero at kernel start, re-zeroed by its reader
#if BL_BLOOM
    __shared__ u32 s_bloom[128];                     // 4,096 bits, one per candidate key (12 low bits of its hash)
#endif
    VidSlot* tab = reinterpret_cast<VidSlot*>(smem);
    uint4* list = reinterpret_cast<uint4*>(smem + BL_OFF_LIST);
    u32* clist = reinterpret_cast<u32*>(smem + BL_OFF_CLIST);
    u16* ctok = reinterpret_cast<u16*>(smem + BL_OFF_CTOK);
    const int lane = threadIdx.x & 31;
    const int warp = __shfl_sync(CST_FULL, (int)(threadIdx.x >> 5), 0);
    const int S = p.cands_per_video;
    const int n_local = p.greedy ? 2 * S : S;
    const int64_t n_total = p.greedy ? 2 * p.n_cand : p.n_cand;
    if (threadIdx.x < CST2_ROWS * 4) (&s_correct[0][0])[threadIdx.x] = 0;

#pragma unroll 1
    for (int64_t video = blockIdx.x; video < p.n_videos; video += gridDim.x) {
        int64_t r0, r1;
        mt_video_refs(p, video, r0, r1);
        if (BL_PREFETCH && threadIdx.x == 0) mt_prefetch_video(p, video + gridDim.x);
        const char* cbase = reinterpret_cast<const char*>(p.cand) + (size_t)(video * S) * p.pitch_c;
        const char* gbase = reinterpret_cast<const char*>(p.greedy) + (size_t)(video * S) * p.pitch_c;
        u32 bad_rows = 0;
        bool first_batch = true;
#pragma unroll 1
        for (int cb = 0; cb < n_local;) {
            // ---- the candidate batch: rows -> instances -> table (maxima start at 0) ----
            {
                uint4* z = reinterpret_cast<uint4*>(tab);
#pragma unroll 2
                for (u32 i = threadIdx.x; i < T; i += MT_THREADS) z[i] = make_uint4(0, 0, 0, 0);
#if BL_BLOOM
                if (threadIdx.x < 128) s_bloom[threadIdx.x] = 0u;
#endif
            }
            const int n_try = (n_local - cb) < CST2_ROWS ? (n_local - cb) : CST2_ROWS;
            mt_load_rows<BL_LOADS_FIRST != 0>(cbase, gbase, S, cb, n_try, p.pitch_c, ESZ ? ESZ : p.cand_esz, p.cand_width, p.use_eos, ctok, sL, warp, lane);
            __syncthreads();
            int nC, n_entC, Lh;
            {
                const V2Scan sc = v2_scan(sL, n_try, MT_LIST_CAP, lane);
                nC = sc.nfit;
                n_entC = sc.n_ent;
                Lh = sc.L;                           // lane h: testlen of candidate row h
                if (warp == 0 && lane < nC) bad_rows += (u32)(sL[lane] >> 8);
                mt_emit_rows(ctok, sc, list, warp, lane);
            }
            __syncthreads();
#pragma unroll 1
            for (int i = (int)threadIdx.x; i < n_entC; i += MT_THREADS) {
                const uint4 e = list[i];
                u32 c = BL_NONE;
                if (e.z != 0u) {                     // first occurrence in its row: slot | row | order | count
                    const u64 key = (u64)e.x | ((u64)e.y << 32);
                    const u32 h = cst_hash(key);
#if BL_BLOOM
                    atomicOr(&s_bloom[(h >> 5) & 127u], 1u << (h & 31u));
#endif
                    c = v2_insert<T>(tab, h, key) | ((e.z & 0x7fu) << 11) | ((e.z >> 8) << 18);
                }
                clist[i] = c;
            }
            // closest reference length of the lane's candidate (warp 0): (|l - testlen|, l) smallest
            u32 best = 0xffffffffu;
            // ---- the references stream through (their loads start while other warps still insert: the
            // token buffer is free once the candidates' instances are in the list) ----
#pragma unroll 1
            for (int64_t rb = r0; rb < r1;) {
                const int left = (int)((r1 - rb) < CST2_ROWS ? (r1 - rb) : CST2_ROWS);
                mt_load_rows<BL_LOADS_FIRST != 0>(reinterpret_cast<const char*>(p.refs) + (size_t)rb * p.pitch_r, nullptr, 0x7fffffff, 0, left,
                             p.pitch_r, ESZ ? ESZ : p.ref_esz, p.ref_width, p.use_eos, ctok, sL, warp, lane);
                __syncthreads();                     // (also: every candidate key is in the table, the list is free,
                                                     //  and the previous stage's finds are done with it)
                int n_entR, nR;
                {
                    const V2Scan sr = v2_scan(sL, left, MT_LIST_CAP, lane);
                    nR = sr.nfit;
                    n_entR = sr.n_ent;
                    if (warp == 0 && lane < nR && first_batch) bad_rows += (u32)(sL[lane] >> 8);
                    mt_emit_rows(ctok, sr, list, warp, lane);
                    if (warp == 0) {
#pragma unroll 1
                        for (int j = 0; j < nR; ++j) {
                            const int l = __shfl_sync(CST_FULL, sr.L, j);
                            const int d = (l > Lh) ? l - Lh : Lh - l;
                            const u32 cand = ((u32)d << 8) | (u32)l;
                            best = cand < best ? cand : best;
                        }
                    }
                }
                __syncthreads();
#pragma unroll 1
                for (int i = (int)threadIdx.x; i < n_entR; i += MT_THREADS) {
                    const uint4 e = list[i];
                    if (e.z != 0u) {
                        const u64 key = (u64)e.x | ((u64)e.y << 32);
                        const u32 h = cst_hash(key);
#if BL_BLOOM
                        // most reference n-grams are in no candidate: four in five of those end here, without a probe
                        if (!((s_bloom[(h >> 5) & 127u] >> (h & 31u)) & 1u)) continue;
#endif
                        const int idx = v2_find<T, false>(tab, h, key);
                        if (idx >= 0) atomicMax(&tab[idx].P, e.z >> 8);       // cook_refs: maxcounts
                    }
                }
                rb += nR;
            }
            __syncthreads();                         // the maxima are final (no references: the inserts have landed)
            // ---- clipped counts: min(count in the candidate, largest count in a reference) ----
#pragma unroll 1
            for (int i = (int)threadIdx.x; i < n_entC; i += MT_THREADS) {
                const u32 c = clist[i];
                if (c != BL_NONE) {
                    const u32 tf = c >> 18, mx = tab[c & 0x7ffu].P;
                    const u32 m = tf < mx ? tf : mx;
                    if (m) atomicAdd(&s_correct[(c >> 11) & 31u][(c >> 16) & 3u], (int)m);
                }
            }
            __syncthreads();
            // ---- BleuScorer.compute_score, the per-sentence part (warp 0; the others go on to the next
            // batch: nothing they touch before its first barrier is read here) ----
            if (warp == 0) {
                int cor[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    cor[k] = s_correct[lane][k];
                    s_correct[lane][k] = 0;
                }
                if (lane < nC) {
                    const int c = cb + lane;
                    const int64_t out = (c >= S) ? p.n_cand + video * S + (c - S) : video * S + c;
                    const int testlen = Lh;
                    const int reflen = (best == 0xffffffffu) ? 0 : (int)(best & 0xffu);
                    const double small = 1e-9, tiny = 1e-15;
                    const double ratio = ((double)testlen + tiny) / ((double)reflen + small);
                    const double bp = (ratio < 1.0) ? exp(1.0 - 1.0 / ratio) : 1.0;
                    double bleu = 1.0;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int guess = testlen > k ? testlen - k : 0;
                        bleu *= ((double)cor[k] + tiny) / ((double)guess + small);
                        double b = (k == 0) ? bleu : (k == 1) ? sqrt(bleu) : pow(bleu, 1.0 / (double)(k + 1));
                        if (ratio < 1.0) b *= bp;
                        bleu_out[(size_t)k * (size_t)n_total + (size_t)out] = b;
                    }
                    if (comps_out) {
                        int* q = comps_out + (size_t)out * 6;
                        q[0] = testlen;
                        q[1] = reflen;
                        q[2] = cor[0];
                        q[3] = cor[1];
                        q[4] = cor[2];
                        q[5] = cor[3];
                    }
                }
            }
            cb += nC;
            first_batch = false;
        }
        if (bad_rows && p.bad_out) atomicAdd(p.bad_out, bad_rows);
    }
}

// ---------------------------------------------------------------------------
// ROUGE-L
// ---------------------------------------------------------------------------
// Per batch of references: a dictionary of their distinct tokens (open addressing, the slot is the
// token's dense id) and a dense matrix mask[slot][reference] of position bits.  A candidate token is
// looked up ONCE (-> its slot, or "in no reference"), not once per reference; the pair threads then
// read their masks by index.  Capacity C = RG_MASK_WORDS / n slots for a batch of n references, and a
// batch takes as many references as keep the dictionary at most 80 % full even if all tokens differ.
#define RG_MASK_WORDS 6144            /* 24 KB */
#define RG_DICT_MAX 1024u
#define RG_NONE 0xffffffffu

__device__ __forceinline__ u32 rg_home(u32 code, u32 C) { return __umulhi(code * 0x9E3779B1u, C); }

template <int ESZ>
__global__ void __launch_bounds__(MT_THREADS, 6) k_rouge(const RowsView p, double* __restrict__ scores) {
    __shared__ __align__(16) u32 mask[RG_MASK_WORDS];
    __shared__ __align__(16) u32 dict[RG_DICT_MAX];
    __shared__ __align__(16) u16 ctokC[MT_CTOK_U16];
    __shared__ __align__(16) u16 ctokR[MT_CTOK_U16];
    __shared__ int sLc[CST2_ROWS], sLr[CST2_ROWS];
    __shared__ int s_lcs[CST2_ROWS];
    __shared__ u32 s_key[CST2_ROWS];                 // best recall as floor(4096 lcs / len) | lcs | len
    const int lane = threadIdx.x & 31;
    const int warp = __shfl_sync(CST_FULL, (int)(threadIdx.x >> 5), 0);
    const int S = p.cands_per_video;
    const int n_local = p.greedy ? 2 * S : S;
    if (threadIdx.x < CST2_ROWS) {                   // (re-zeroed by their reader, warp 0)
        s_lcs[threadIdx.x] = 0;
        s_key[threadIdx.x] = 0u;
    }

#pragma unroll 1
    for (int64_t video = blockIdx.x; video < p.n_videos; video += gridDim.x) {
        int64_t r0, r1;
        mt_video_refs(p, video, r0, r1);
        if (threadIdx.x == 0) mt_prefetch_video(p, video + gridDim.x);
        const char* cbase = reinterpret_cast<const char*>(p.cand) + (size_t)(video * S) * p.pitch_c;
        const char* gbase = reinterpret_cast<const char*>(p.greedy) + (size_t)(video * S) * p.pitch_c;
        u32 bad_rows = 0;
#pragma unroll 1
        for (int cb = 0; cb < n_local; cb += CST2_ROWS) {
            const int nC = (n_local - cb) < CST2_ROWS ? (n_local - cb) : CST2_ROWS;
            mt_load_rows(cbase, gbase, S, cb, nC, p.pitch_c, ESZ ? ESZ : p.cand_esz, p.cand_width, p.use_eos, ctokC, sLc, warp, lane);
            bool cand_counted = false;
            int Lc_mine = 0;                         // warp 0, lane h: L | bad << 8 of candidate row h (kept for the end)
#pragma unroll 1
            for (int64_t rb = r0; rb < r1;) {
                const int left = (int)((r1 - rb) < CST2_ROWS ? (r1 - rb) : CST2_ROWS);
                {
                    uint4* z = reinterpret_cast<uint4*>(mask);
#pragma unroll
                    for (int i = 0; i < RG_MASK_WORDS / 4 / MT_THREADS; ++i) z[threadIdx.x + i * MT_THREADS] = make_uint4(0, 0, 0, 0);
                    uint4* d = reinterpret_cast<uint4*>(dict);
                    d[threadIdx.x] = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
                }
                mt_load_rows(reinterpret_cast<const char*>(p.refs) + (size_t)rb * p.pitch_r, nullptr, 0x7fffffff, 0, left,
                             p.pitch_r, ESZ ? ESZ : p.ref_esz, p.ref_width, p.use_eos, ctokR, sLr, warp, lane);
                __syncthreads();
                // how many of the loaded references this batch takes (every warp works it out for itself)
                int nR;
                bool any_long;
                {
                    const int L = (lane < left) ? (sLr[lane] & 0xff) : 0;
                    int tok = (lane < left) ? (L > 0 ? L : 1) : 0;       // an empty sentence is ONE empty token
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const int v = __shfl_up_sync(CST_FULL, tok, o);
                        if (lane >= o) tok += v;
                    }
                    // tokens(n) <= 0.8 * min(RG_DICT_MAX, RG_MASK_WORDS / n)
                    nR = __popc(__ballot_sync(CST_FULL, lane < left && (u32)tok * 5u <= 4u * RG_DICT_MAX &&
                                                            (u32)tok * 5u * (u32)(lane + 1) <= 4u * RG_MASK_WORDS));
                    if (nR < 1) nR = 1;                                  // (one row of 32 tokens always fits)
                    any_long = __ballot_sync(CST_FULL, lane < nR && L > 16) != 0u;
                }
                const u32 C = min((u32)RG_DICT_MAX, (u32)RG_MASK_WORDS / (u32)nR);
                if (warp == 0 && cb == 0 && lane < nR) bad_rows += (u32)(sLr[lane] >> 8);
                if (warp == 0 && !cand_counted) {
                    Lc_mine = lane < nC ? sLc[lane] : 0;
                    bad_rows += (u32)(Lc_mine >> 8);
                }
                cand_counted = true;
                // ---- the references' tokens -> dictionary; position bit -> mask[slot][reference].  Half a warp
                // per row (captions are mostly shorter than 16 tokens); rows beyond that get a second pass
                // for their upper halves ----
#pragma unroll 1
                for (int half = 0; half < (any_long ? 2 : 1); ++half) {
#pragma unroll 1
                    for (int t = (int)threadIdx.x; t < nR * 16; t += MT_THREADS) {
                        const int r = t >> 4, j = (t & 15) + 16 * half;
                        const int L = sLr[r] & 0xff;
                        if (j < (L > 0 ? L : 1)) {
                            const u32 code = L > 0 ? (u32)ctokR[r * CST2_TOK_STRIDE + j] : 0u;   // (0: the empty token)
                            u32 idx = rg_home(code, C);
#pragma unroll 1
                            for (;;) {
                                const u32 cur = *reinterpret_cast<volatile u32*>(&dict[idx]);
                                if (cur == code) break;
                                if (cur == 0xffffffffu) {
                                    const u32 old = atomicCAS(&dict[idx], 0xffffffffu, code);
                                    if (old == 0xffffffffu || old == code) break;
                                }
                                idx = (idx + 1u == C) ? 0u : idx + 1u;
                            }
                            atomicOr(&mask[idx * (u32)nR + (u32)r], 1u << j);
                        }
                    }
                }
                __syncthreads();
                // ---- a warp per candidate: lane i looks token i up ONCE (-> its slot, or "in no reference of
                // the batch"); then lane r runs my_lcs against reference r as the bit-vector recurrence, the
                // slots going round by shuffle ----
#pragma unroll 1
                for (int h = warp; h < nC; h += MT_WARPS) {
                    const int Lh = sLc[h] & 0xff;
                    const int steps = Lh > 0 ? Lh : 1;
                    u32 off = RG_NONE;               // word offset of the token's row of masks, if any reference holds it
                    if (lane < steps) {
                        const u32 code = Lh > 0 ? (u32)ctokC[h * CST2_TOK_STRIDE + lane] : 0u;
                        u32 idx = rg_home(code, C);
#pragma unroll 1
                        for (;;) {
                            const u32 cur = dict[idx];
                            if (cur == code) { off = idx * (u32)nR; break; }
                            if (cur == 0xffffffffu) break;
                            idx = (idx + 1u == C) ? 0u : idx + 1u;
                        }
                    }
                    const bool on = lane < nR;
                    const u32* mr = mask + lane;
                    u32 V = 0xffffffffu;
                    // (a token no reference of the batch holds leaves V as it is: only the others are stepped through)
                    u32 present = __ballot_sync(CST_FULL, off != RG_NONE);
#pragma unroll 1
                    while (present) {
                        const int i = __ffs(present) - 1;
                        present &= present - 1u;
                        const u32 d = __shfl_sync(CST_FULL, off, i);
                        const u32 U = on ? (V & mr[d]) : 0u;
                        V = (V + U) | (V - U);
                    }
                    // prec = lcs / len(candidate): the largest lcs; rec = lcs / len(reference): fractions with
                    // denominators <= 32 differ by >= 2^-10, so floor(4096 lcs / len) orders them, and equal
                    // fractions are equal doubles whichever (lcs, len) pair speaks for them
                    u32 lcs = 0u, key = 0u;
                    if (on) {
                        const int Lr = sLr[lane] & 0xff;
                        const u32 lenr = (u32)(Lr > 0 ? Lr : 1);
                        lcs = (u32)__popc(~V & (lenr >= 32u ? 0xffffffffu : ((1u << lenr) - 1u)));
                        // (the key only has to ORDER distinct fractions: they lie >= 4 apart on this scale, the fast
                        //  float quotient is off by less than 1; equal fractions may get keys one apart, and either
                        //  one's (lcs, len) gives the same double)
                        key = ((u32)(__fdividef((float)(lcs << 12), (float)lenr)) << 12) | (lcs << 6) | lenr;
                    }
                    lcs = __reduce_max_sync(CST_FULL, lcs);
                    key = __reduce_max_sync(CST_FULL, key);
                    if (lane == 0) {                 // (one warp owns a candidate: no atomics)
                        if ((int)lcs > s_lcs[h]) s_lcs[h] = (int)lcs;
                        if (key > s_key[h]) s_key[h] = key;
                    }
                }
                rb += nR;
                __syncthreads();
            }
            if (r0 >= r1) {                          // (no references: the candidates' lengths still have to land)
                __syncthreads();
                if (warp == 0) {
                    Lc_mine = lane < nC ? sLc[lane] : 0;
                    bad_rows += (u32)(Lc_mine >> 8);
                }
            }
            // Rouge.calc_score: F-measure of the two maxima, in the scorer's operation order.  Warp 0 alone;
            // the others go on to the next batch (nothing they touch before its first barrier is read here).
            if (warp == 0) {
                const int lbest = s_lcs[lane];
                const u32 kbest = s_key[lane];
                s_lcs[lane] = 0;
                s_key[lane] = 0u;
                if (lane < nC) {
                    const int c = cb + lane;
                    const int64_t out = (c >= S) ? p.n_cand + video * S + (c - S) : video * S + c;
                    const int Lh = Lc_mine & 0xff;
                    const double prec = (double)lbest / (double)(Lh > 0 ? Lh : 1);
                    const double rec = kbest ? (double)((kbest >> 6) & 63u) / (double)(kbest & 63u) : 0.0;
                    const double beta2 = 1.2 * 1.2;
                    double score = 0.0;
                    if (prec != 0.0 && rec != 0.0)
                        score = __ddiv_rn(__dmul_rn(__dmul_rn(__dadd_rn(1.0, beta2), prec), rec),
                                          __dadd_rn(rec, __dmul_rn(beta2, prec)));
                    scores[out] = score;
                }
            }
        }
        if (bad_rows && p.bad_out) atomicAdd(p.bad_out, bad_rows);
    }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
static int mt_dtype_bytes(int dtype) {
    return dtype == CST_DTYPE_I32 ? 4 : dtype == CST_DTYPE_I64 ? 8 : dtype == CST_DTYPE_I16 ? 2 : 0;
}

static int mt_view(const char* who, RowsView& p, const void* cand, const void* greedy, int32_t cand_dtype,
                   int64_t n_cand, int32_t cand_width, int64_t cand_stride, const void* refs, int32_t ref_dtype,
                   int32_t ref_width, int64_t ref_stride, const int64_t* ref_offsets, int32_t refs_per_video,
                   int64_t n_videos, int32_t cands_per_video, int32_t use_eos, uint32_t* bad_rows_out) {
    if (n_cand < 0 || n_videos < 0 || cands_per_video < 1) return cst_fail(CST_E_ARG, "%s: negative size", who);
    if (n_cand != n_videos * (int64_t)cands_per_video)
        return cst_fail(CST_E_ARG, "%s: n_cand (%lld) != n_videos (%lld) * cands_per_video (%d)", who, (long long)n_cand,
                        (long long)n_videos, cands_per_video);
    if (cand_width < 1 || cand_width > CST_MAX_ROW_WIDTH || ref_width < 1 || ref_width > CST_MAX_ROW_WIDTH)
        return cst_fail(CST_E_RANGE, "%s: row widths %d/%d outside [1, %d]", who, cand_width, ref_width, CST_MAX_ROW_WIDTH);
    if (cand_stride < cand_width || ref_stride < ref_width) return cst_fail(CST_E_ARG, "%s: stride < width", who);
    const int ce = mt_dtype_bytes(cand_dtype), re = mt_dtype_bytes(ref_dtype);
    if (!ce || !re) return cst_fail(CST_E_ARG, "%s: bad dtype", who);
    if (refs_per_video < 0 || (refs_per_video == 0 && !ref_offsets))
        return cst_fail(CST_E_ARG, "%s: refs_per_video must be > 0", who);
    if (cands_per_video > 4096) return cst_fail(CST_E_RANGE, "%s: cands_per_video %d > 4096", who, cands_per_video);
    if ((uint64_t)cand_stride * (uint64_t)ce > 0xffffffffull || (uint64_t)ref_stride * (uint64_t)re > 0xffffffffull)
        return cst_fail(CST_E_RANGE, "%s: row pitch beyond 4 GiB", who);
    if (n_cand > 0 && (!cand || !refs)) return cst_fail(CST_E_ARG, "%s: NULL rows", who);
    memset(&p, 0, sizeof(p));
    p.cand = cand;
    p.greedy = greedy;
    p.refs = refs;
    p.ref_offsets = ref_offsets;
    p.n_cand = n_cand;
    p.n_videos = n_videos;
    p.cand_esz = ce;
    p.ref_esz = re;
    p.cand_width = cand_width;
    p.ref_width = ref_width;
    p.pitch_c = (u32)((uint64_t)cand_stride * (uint64_t)ce);
    p.pitch_r = (u32)((uint64_t)ref_stride * (uint64_t)re);
    p.refs_per_video = refs_per_video;
    p.cands_per_video = cands_per_video;
    p.use_eos = use_eos ? 1 : 0;
    p.bad_out = bad_rows_out;
    return CST_OK;
}

static unsigned mt_grid(int device, int64_t n_videos, int ctas_per_sm) {
    int sms = 148;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || sms < 1) sms = 148;
    // (the error, if any, is sticky-free: a failed attribute query leaves no pending error)
    const int64_t full = (int64_t)sms * ctas_per_sm * 8;      // several waves: CTAs of uneven videos even out
    return (unsigned)(n_videos < full ? n_videos : full);
}

extern "C" int cst_bleu_scores(const void* cand, const void* greedy, int32_t cand_dtype, int64_t n_cand,
                               int32_t cand_width, int64_t cand_stride, const void* refs, int32_t ref_dtype,
                               int32_t ref_width, int64_t ref_stride, const int64_t* ref_offsets,
                               int32_t refs_per_video, int64_t n_videos, int32_t cands_per_video, int32_t use_eos,
                               double* bleu_out, int32_t* comps_out, uint32_t* bad_rows_out, int32_t device,
                               void* stream) {
    RowsView p;
    int rc = mt_view("cst_bleu_scores", p, cand, greedy, cand_dtype, n_cand, cand_width, cand_stride, refs, ref_dtype,
                     ref_width, ref_stride, ref_offsets, refs_per_video, n_videos, cands_per_video, use_eos, bad_rows_out);
    if (rc != CST_OK) return rc;
    if (n_cand == 0) return CST_OK;
    if (!bleu_out) return cst_fail(CST_E_ARG, "cst_bleu_scores: NULL output");
    CstDeviceGuard guard(device);
    if (!guard.ok) return cst_fail(CST_E_CUDA, "cst_bleu_scores: cannot select device %d", device);
    static bool attr_set[64] = {false};
    if (device < 0 || device >= 64 || !attr_set[device]) {
        CST_CUDA_TRY(cudaFuncSetAttribute(k_bleu<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BL_SMEM));
        CST_CUDA_TRY(cudaFuncSetAttribute(k_bleu<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BL_SMEM));
        if (device >= 0 && device < 64) attr_set[device] = true;     // (a constant: racing threads set the same value)
    }
    // (int32 rows on both sides -- what the callers of this package keep on the device -- get the element size compiled in)
    if (p.cand_esz == 4 && p.ref_esz == 4)
        k_bleu<4><<<mt_grid(device, n_videos, 6), MT_THREADS, BL_SMEM, (cudaStream_t)stream>>>(p, bleu_out, comps_out);
    else
        k_bleu<0><<<mt_grid(device, n_videos, 6), MT_THREADS, BL_SMEM, (cudaStream_t)stream>>>(p, bleu_out, comps_out);
    cst_count_launch();
    CST_CUDA_TRY(cudaGetLastError());
    return CST_OK;
}

extern "C" int cst_rouge_scores(const void* cand, const void* greedy, int32_t cand_dtype, int64_t n_cand,
                                int32_t cand_width, int64_t cand_stride, const void* refs, int32_t ref_dtype,
                                int32_t ref_width, int64_t ref_stride, const int64_t* ref_offsets,
                                int32_t refs_per_video, int64_t n_videos, int32_t cands_per_video, int32_t use_eos,
                                double* scores_out, uint32_t* bad_rows_out, int32_t device, void* stream) {
    RowsView p;
    int rc = mt_view("cst_rouge_scores", p, cand, greedy, cand_dtype, n_cand, cand_width, cand_stride, refs, ref_dtype,
                     ref_width, ref_stride, ref_offsets, refs_per_video, n_videos, cands_per_video, use_eos, bad_rows_out);
    if (rc != CST_OK) return rc;
    if (n_cand == 0) return CST_OK;
    if (!scores_out) return cst_fail(CST_E_ARG, "cst_rouge_scores: NULL output");
    CstDeviceGuard guard(device);
    if (!guard.ok) return cst_fail(CST_E_CUDA, "cst_rouge_scores: cannot select device %d", device);
    if (p.cand_esz == 4 && p.ref_esz == 4)
        k_rouge<4><<<mt_grid(device, n_videos, 6), MT_THREADS, 0, (cudaStream_t)stream>>>(p, scores_out);
    else
        k_rouge<0><<<mt_grid(device, n_videos, 6), MT_THREADS, 0, (cudaStream_t)stream>>>(p, scores_out);
    cst_count_launch();
    CST_CUDA_TRY(cudaGetLastError());
    return CST_OK;
}
