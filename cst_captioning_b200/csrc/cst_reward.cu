// CIDEr-D reward path for B200 (sm_100a): kernels + C ABI (include/cst_reward.h).
//
// This file holds the df table (host build, device build, probes), the baseline-only kernel, the
// fold of the per-video statistics and the host side of every entry point; the scoring kernel
// itself (raw rows, staged references, cooking, leave-one-out counts) is csrc/cst_score2.cuh.
//
// No tensor cores: the work is sparse hashing, not a dense contraction.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <new>
#include <utility>
#include <vector>

#include "cst_common.cuh"
#include "cst_internal.h"
#include "cst_reward.h"

// ---------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------
static thread_local char g_err[512] = "";

#define fail cst_fail
#define CUDA_TRY CST_CUDA_TRY

extern "C" const char* cst_last_error(void) { return g_err; }
extern "C" int cst_abi_version(void) { return 1; }

// every kernel this library launches is counted here (bench.py reports it)
static std::atomic<long long> g_launches{0};
#define COUNT_LAUNCH() g_launches.fetch_add(1, std::memory_order_relaxed)
extern "C" int64_t cst_kernel_launches(void) { return (int64_t)g_launches.load(std::memory_order_relaxed); }

// entry points of cst_internal.h for the other translation units
int cst_fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
void cst_count_launch(void) { COUNT_LAUNCH(); }

typedef CstDeviceGuard DeviceGuard;

// bytes per token id of a CST_DTYPE_*, 0 if unknown
static int dtype_bytes(int32_t dtype) {
    return dtype == CST_DTYPE_I32 ? 4 : dtype == CST_DTYPE_I64 ? 8 : dtype == CST_DTYPE_I16 ? 2 : 0;
}

// ---------------------------------------------------------------------------
// df table handle (layout: DfView in cst_common.cuh)
// ---------------------------------------------------------------------------
struct cst_df_table {
    int device;
    u32 n_buckets;
    int64_t size;
    int64_t n_documents;
    double idf_unseen;
    u64* keys;       // [n_buckets * 4]
    double* idfv;    // [n_buckets * 4]
    u32* dfc;        // [n_buckets * 4]
    double* lut;     // [n_documents + 1]
    double* uni;     // [CST_UNI_N]
    double* gauss;   // [CST_GAUSS_N] e^{-d^2 / (2 sigma^2)}, d = index - CST_GAUSS_HALF
};

static DfView view_of(const cst_df_table* t) {
    DfView v;
    v.keys = t->keys;
    v.idfv = t->idfv;
    v.dfc = t->dfc;
    v.lut = t->lut;
    v.uni = t->uni;
    v.n_buckets = t->n_buckets;
    v.unseen = t->idf_unseen;
    return v;
}

static u32 df_buckets_for(int64_t n) {
    int64_t nb = (2 * n + CST_BUCKET - 1) / CST_BUCKET + 4;   // load factor <= 0.5
#ifdef CST_EXPERIMENT
    if (const char* e = getenv("CST_EXP_LOAD_PCT")) nb = (100 * n / atoi(e) + CST_BUCKET - 1) / CST_BUCKET + 4;
#endif
    if (nb < 256) nb = 256;
    return (u32)nb;
}

// Gaussian length penalty exactly as the scorer evaluates it:
// np.e ** (-(delta ** 2) / (2 * sigma ** 2)) on Python floats, i.e. libm pow.
static void fill_gauss(double* g) {
    const double e = 2.718281828459045;
    for (int i = 0; i < CST_GAUSS_N; ++i) {
        double delta = (double)(i - CST_GAUSS_HALF);
        g[i] = pow(e, -(delta * delta) / (2.0 * (CST_SIGMA * CST_SIGMA)));
    }
}

// Allocate a zeroed table and upload the idf LUT (idf_by_df: n_documents + 1 values).
static int table_alloc(int device, u32 n_buckets, const double* idf_by_df, int64_t n_documents,
                       cst_df_table** out) {
    cst_df_table* t = new (std::nothrow) cst_df_table();
    if (!t) return fail(CST_E_NOMEM, "out of host memory");
    memset(t, 0, sizeof(*t));
    t->device = device;
    t->n_buckets = n_buckets;
    t->n_documents = n_documents;
    t->idf_unseen = idf_by_df[0];
    double g[CST_GAUSS_N];
    fill_gauss(g);
    const size_t slots = (size_t)n_buckets * CST_BUCKET;
    cudaError_t e = cudaMalloc(&t->keys, slots * sizeof(u64));
    if (e == cudaSuccess) e = cudaMalloc(&t->dfc, slots * sizeof(u32));
    if (e == cudaSuccess) e = cudaMalloc(&t->idfv, slots * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&t->lut, (size_t)(n_documents + 1) * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&t->uni, (size_t)CST_UNI_N * sizeof(double));
    if (e == cudaSuccess) e = cudaMalloc(&t->gauss, sizeof(g));
    if (e == cudaSuccess) e = cudaMemset(t->keys, 0, slots * sizeof(u64));
    if (e == cudaSuccess) e = cudaMemset(t->dfc, 0, slots * sizeof(u32));
    if (e == cudaSuccess) e = cudaMemcpy(t->lut, idf_by_df, (size_t)(n_documents + 1) * sizeof(double), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(t->gauss, g, sizeof(g), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(t->keys);
        cudaFree(t->dfc);
        cudaFree(t->idfv);
        cudaFree(t->lut);
        cudaFree(t->uni);
        cudaFree(t->gauss);
        delete t;
        return fail(e == cudaErrorMemoryAllocation ? CST_E_NOMEM : CST_E_CUDA, "df table allocation: %s",
                    cudaGetErrorString(e));
    }
    *out = t;
    return CST_OK;
}

// idfv[slot] = lut[dfc[slot]] for every stored key
__global__ void k_fill_idf(const u64* __restrict__ keys, const u32* __restrict__ dfc, const double* __restrict__ lut,
                           double* idfv, size_t slots) {
    const size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= slots) return;
    idfv[s] = keys[s] ? lut[dfc[s]] : 0.0;
}

// uni[token] = idf of the unigram `token` (direct-indexed shortcut of the hash table)
__global__ void k_fill_uni(DfView t, double* uni) {
    const u32 tok = blockIdx.x * blockDim.x + threadIdx.x;
    if (tok >= CST_UNI_N) return;
    const u64 key = (u64)tok + 1ull;
    u64 k[4];
    const u32 b = cst_slot(cst_hash(key), t.n_buckets);
    ld_bucket(t.keys + (size_t)b * CST_BUCKET, k);
    const int slot = bucket_resolve(t.keys, t.n_buckets, b, k, key);
    uni[tok] = (slot >= 0) ? t.lut[t.dfc[slot]] : t.unseen;
}

static int table_finish(cst_df_table* t) {
    const size_t slots = (size_t)t->n_buckets * CST_BUCKET;
    k_fill_idf<<<(unsigned)((slots + 255) / 256), 256>>>(t->keys, t->dfc, t->lut, t->idfv, slots);
    COUNT_LAUNCH();
    CUDA_TRY(cudaGetLastError());
    k_fill_uni<<<(CST_UNI_N + 255) / 256, 256>>>(view_of(t), t->uni);
    COUNT_LAUNCH();
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaDeviceSynchronize());
    return CST_OK;
}

static int check_lut(const char* who, const double* idf_by_df, int64_t n_documents) {
    if (!idf_by_df || n_documents < 1) return fail(CST_E_ARG, "%s: idf_by_df / n_documents missing", who);
    if (n_documents > (int64_t)1 << 28) return fail(CST_E_RANGE, "%s: n_documents too large", who);
    for (int64_t d = 0; d <= n_documents; ++d)
        if (!(idf_by_df[d] >= 0.0)) return fail(CST_E_ARG, "%s: idf_by_df[%lld] is negative or NaN", who, (long long)d);
    return CST_OK;
}

extern "C" int cst_df_create(const uint64_t* keys, const double* df, int64_t n, const double* idf_by_df,
                             int64_t n_documents, int device, cst_df_table** out) {
    if (!out) return fail(CST_E_ARG, "cst_df_create: out is NULL");
    *out = nullptr;
    if (n < 0 || (n > 0 && (!keys || !df))) return fail(CST_E_ARG, "cst_df_create: NULL input");
    if (n > (int64_t)1 << 30) return fail(CST_E_RANGE, "cst_df_create: %lld n-grams exceed 2^30", (long long)n);
    int rc = check_lut("cst_df_create", idf_by_df, n_documents);
    if (rc != CST_OK) return rc;
    DeviceGuard guard(device);
    if (!guard.ok) return fail(CST_E_CUDA, "cst_df_create: cannot select CUDA device %d", device);

    const u32 nb = df_buckets_for(n);
    const size_t slots = (size_t)nb * CST_BUCKET;
    std::vector<u64> h_keys;
    std::vector<u32> h_dfc;
    std::vector<unsigned char> h_over;        // bucket b was full when an insertion came by
    try {
        h_keys.assign(slots, 0ull);
        h_dfc.assign(slots, 0u);
        h_over.assign(nb, 0);
    } catch (const std::bad_alloc&) {
        return fail(CST_E_NOMEM, "cst_df_create: out of host memory for %zu slots", slots);
    }
    for (int64_t i = 0; i < n; ++i) {
        const u64 k = keys[i];
        if (k == 0ull || (k & 0xffffull) == 0ull)
            return fail(CST_E_ARG, "cst_df_create: key %lld is not a packed n-gram", (long long)i);
        const double d = df[i];
        if (!(d >= 0.0) || d > (double)n_documents || d != floor(d))
            return fail(CST_E_ARG, "cst_df_create: df[%lld] = %g is not an integer in [0, n_documents]", (long long)i, d);
        u32 b = cst_slot(cst_hash(k), nb);
        for (;;) {
            int j = 0;
            for (; j < CST_BUCKET; ++j) {
                u64 cur = h_keys[(size_t)b * CST_BUCKET + j];
                if (cur == k) return fail(CST_E_ARG, "cst_df_create: duplicate key at %lld", (long long)i);
                if (cur == 0ull) break;
            }
            if (j < CST_BUCKET) {
                h_keys[(size_t)b * CST_BUCKET + j] = k;
                h_dfc[(size_t)b * CST_BUCKET + j] = (u32)d;
                break;
            }
            h_over[b] = 1;
            b = (b + 1 == nb) ? 0u : b + 1;
        }
    }
    // the overflow mark of a full bucket: the order of its first two keys (bucket_resolve)
    for (u32 b = 0; b < nb; ++b) {
        u64* k = &h_keys[(size_t)b * CST_BUCKET];
        if (k[CST_BUCKET - 1] == 0ull) continue;
        if ((k[0] > k[1]) != (h_over[b] != 0)) {
            std::swap(k[0], k[1]);
            std::swap(h_dfc[(size_t)b * CST_BUCKET], h_dfc[(size_t)b * CST_BUCKET + 1]);
        }
    }
    cst_df_table* t = nullptr;
    rc = table_alloc(device, nb, idf_by_df, n_documents, &t);
    if (rc != CST_OK) return rc;
    cudaError_t e = cudaMemcpy(t->keys, h_keys.data(), slots * sizeof(u64), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(t->dfc, h_dfc.data(), slots * sizeof(u32), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cst_df_destroy(t);
        return fail(CST_E_CUDA, "cst_df_create: upload failed: %s", cudaGetErrorString(e));
    }
    t->size = n;
    rc = table_finish(t);
    if (rc != CST_OK) {
        cst_df_destroy(t);
        return rc;
    }
    *out = t;
    return CST_OK;
}

extern "C" int cst_df_destroy(cst_df_table* t) {
    if (!t) return CST_OK;
    DeviceGuard guard(t->device);
    cudaFree(t->keys);
    cudaFree(t->dfc);
    cudaFree(t->idfv);
    cudaFree(t->lut);
    cudaFree(t->uni);
    cudaFree(t->gauss);
    delete t;
    return CST_OK;
}

extern "C" int64_t cst_df_size(const cst_df_table* t) { return t ? t->size : -1; }
extern "C" int64_t cst_df_capacity(const cst_df_table* t) { return t ? (int64_t)t->n_buckets * CST_BUCKET : -1; }
extern "C" int cst_df_device(const cst_df_table* t) { return t ? t->device : -1; }

extern "C" int cst_df_export(const cst_df_table* t, uint64_t* keys_out, double* df_out) {
    if (!t || (t->size > 0 && (!keys_out || !df_out))) return fail(CST_E_ARG, "cst_df_export: NULL argument");
    DeviceGuard guard(t->device);
    if (!guard.ok) return fail(CST_E_CUDA, "cst_df_export: cannot select device %d", t->device);
    const size_t slots = (size_t)t->n_buckets * CST_BUCKET;
    std::vector<u64> h_keys(slots);
    std::vector<u32> h_dfc(slots);
    CUDA_TRY(cudaMemcpy(h_keys.data(), t->keys, slots * sizeof(u64), cudaMemcpyDeviceToHost));
    CUDA_TRY(cudaMemcpy(h_dfc.data(), t->dfc, slots * sizeof(u32), cudaMemcpyDeviceToHost));
    int64_t at = 0;
    for (size_t s = 0; s < slots; ++s) {
        if (h_keys[s] == 0ull) continue;
        if (at >= t->size) return fail(CST_E_CUDA, "cst_df_export: table holds more keys than its size");
        keys_out[at] = h_keys[s];
        df_out[at] = (double)h_dfc[s];
        ++at;
    }
    if (at != t->size) return fail(CST_E_CUDA, "cst_df_export: found %lld keys, expected %lld", (long long)at, (long long)t->size);
    return CST_OK;
}

// ---------------------------------------------------------------------------
// parity probes
// ---------------------------------------------------------------------------
// Looks keys up exactly the way the scoring kernels do: unigram idf through the
// direct array, everything else through the bucketed table + LUT.
__global__ void k_df_lookup(DfView t, const u64* __restrict__ keys, int64_t n, double* df_out, double* idf_out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u64 key = keys[i];
    double d = 0.0, w = t.unseen;
    if (key != 0ull) {
        u64 k[4];
        const u32 b = cst_slot(cst_hash(key), t.n_buckets);
        ld_bucket(t.keys + (size_t)b * CST_BUCKET, k);
        const int slot = bucket_resolve(t.keys, t.n_buckets, b, k, key);
        if (slot >= 0) {
            const u32 c = t.dfc[slot];
            d = (double)c;
            w = t.idfv[slot];
        }
        if ((key >> 16) == 0ull && (u32)key - 1u < CST_UNI_N) w = t.uni[(u32)key - 1u];
    }
    df_out[i] = d;
    if (idf_out) idf_out[i] = w;
}

extern "C" int cst_df_lookup(const cst_df_table* t, const uint64_t* keys, int64_t n, double* df_out,
                             double* idf_out, void* stream) {
    if (!t || n < 0 || (n > 0 && (!keys || !df_out))) return fail(CST_E_ARG, "cst_df_lookup: bad argument");
    if (n == 0) return CST_OK;
    DeviceGuard guard(t->device);
    if (!guard.ok) return fail(CST_E_CUDA, "cst_df_lookup: cannot select device %d", t->device);
    const int threads = 256;
    const int64_t blocks = (n + threads - 1) / threads;
    k_df_lookup<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(view_of(t), (const u64*)keys, n, df_out, idf_out);
    COUNT_LAUNCH();
    CUDA_TRY(cudaGetLastError());
    return CST_OK;
}

__global__ void k_ngram_counts(const void* __restrict__ rows, int esz, int64_t n_rows, int width, int64_t stride,
                               int use_eos, u64* keys_out, int* counts_out, int* n_out, int* len_out) {
    __shared__ u32 tokbuf[8][32];
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int64_t row = (int64_t)blockIdx.x * 8 + warp;
    if (row >= n_rows) return;
    u64 key[4];
    int L;
    const char* base = reinterpret_cast<const char*>(rows) + (size_t)row * (size_t)stride * (size_t)esz;
    canon_row(base, esz, width, use_eos, lane, 32, tokbuf[warp], key, L);
    int tf[4];
    bool lead[4];
    row_tf(key, lane, 32, -1, tf, lead);
    int at0 = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const u32 lm = __ballot_sync(CST_FULL, lead[k]);
        if (lead[k]) {
            const int at = at0 + __popc(lm & lanemask_lt(lane));
            keys_out[row * CST_NGRAM_SLOTS_PER_ROW + at] = key[k];
            counts_out[row * CST_NGRAM_SLOTS_PER_ROW + at] = tf[k];
        }
        at0 += __popc(lm);
    }
    if (lane == 0) {
        n_out[row] = at0;
        len_out[row] = L;
    }
}

extern "C" int cst_ngram_counts(const void* rows, int32_t dtype, int64_t n_rows, int32_t width,
                                int64_t stride, int32_t use_eos, uint64_t* keys_out, int32_t* counts_out,
                                int32_t* n_out, int32_t* len_out, int device, void* stream) {
    if (n_rows < 0 || (n_rows > 0 && (!rows || !keys_out || !counts_out || !n_out || !len_out)))
        return fail(CST_E_ARG, "cst_ngram_counts: NULL argument");
    if (width < 1 || width > CST_MAX_ROW_WIDTH)
        return fail(CST_E_RANGE, "cst_ngram_counts: row width %d outside [1, %d]", width, CST_MAX_ROW_WIDTH);
    if (stride < width) return fail(CST_E_ARG, "cst_ngram_counts: stride < width");
    if (!dtype_bytes(dtype)) return fail(CST_E_ARG, "cst_ngram_counts: bad dtype");
    if (n_rows == 0) return CST_OK;
    DeviceGuard guard(device);
    if (!guard.ok) return fail(CST_E_CUDA, "cst_ngram_counts: cannot select device %d", device);
    const unsigned blocks = (unsigned)((n_rows + 7) / 8);
    k_ngram_counts<<<blocks, 256, 0, (cudaStream_t)stream>>>(rows, dtype_bytes(dtype), n_rows, width, stride,
                                                           use_eos, (u64*)keys_out, counts_out, n_out, len_out);
    COUNT_LAUNCH();
    CUDA_TRY(cudaGetLastError());
    return CST_OK;
}

// ---------------------------------------------------------------------------
// baseline (phase B): utils.py:292-322 and 219-224 on one group of S scores
// ---------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(CST_FULL, v, o);
    return v;
}

// Sum of the k lowest of x[0..S) by one warp: every lane ranks its elements by
// counting (ties broken by index, like a stable sort) and the elements of rank < k are
// summed with a fixed shuffle tree (deterministic).
__device__ __forceinline__ double k_lowest_sum(const double* x, int S, int k, int lane) {
    double acc = 0.0;
    for (int i = lane; i < S; i += 32) {
        const double xi = x[i];
        int rank = 0;
        for (int j = 0; j < S; ++j) {
            const double xj = x[j];
            rank += (xj < xi || (xj == xi && j < i)) ? 1 : 0;
        }
        if (rank < k) acc += xi;
    }
    return warp_sum(acc);
}

// One warp turns the S scores of a group into rewards and the group's partial
// statistics.  sc: S sampled scores (shared or global); other: greedy scores
// (GREEDY) or bcmrscores row (SCB_GT); reward: global float out.
__device__ __forceinline__ void baseline_group(const double* sc, const double* other, int S, int mode, int k,
                                               int lane, float* reward, double& sum_m,
                                               double& sum_b) {
    double base = 0.0;
    double sb = 0.0;
    if (mode == CST_BASE_SCB_GT) {
        base = k_lowest_sum(other, S, k, lane) / (double)k;
        double a = 0.0;
        for (int i = lane; i < S; i += 32) a += other[i];
        sb = warp_sum(a);
    } else if (mode == CST_BASE_SCB_SAMPLE) {
        const double low = k_lowest_sum(sc, S, k, lane);
        base = low / (double)k;
        sb = low;
    }
    double am = 0.0, ag = 0.0;
    for (int i = lane; i < S; i += 32) {
        const double s = sc[i];
        double r = s - base;
        if (mode == CST_BASE_GREEDY) {
            const double g = other[i];
            r = s - g;
            ag += g;
        }
        reward[i] = (float)r;
        am += s;
    }
    sum_m = warp_sum(am);
    if (mode == CST_BASE_GREEDY) sb = warp_sum(ag);
    sum_b = sb;
}

// The fold itself: thread t adds the groups t, t + blockDim, ...; warps by shuffle tree; the warps'
// sums in order.  The bits depend on blockDim only, so the in-kernel fold and k_fold_stats agree.
__device__ __forceinline__ void fold_stats_body(u32* ctl, const double* partials, int64_t n_groups, double n_m,
                                                double n_b, double* stats, double* red /*[2*nwarps]*/) {
    double am = 0.0, ab = 0.0;
    for (int64_t g = threadIdx.x; g < n_groups; g += blockDim.x) {

        am += __ldcg(partials + 2 * g);
        ab += __ldcg(partials + 2 * g + 1);
    }
    am = warp_sum(am);
    ab = warp_sum(ab);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    if (lane == 0) { red[2 * warp] = am; red[2 * warp + 1] = ab; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double tm = 0.0, tb = 0.0;
        for (int w = 0; w < nw; ++w) { tm += red[2 * w]; tb += red[2 * w + 1]; }
        stats[0] = tm;
        stats[1] = tb;
        stats[2] = n_m;
        stats[3] = n_b;
        stats[4] = (double)atomicExch(&ctl[1], 0u);
        stats[5] = tm / n_m;      // m_score and b_score of THIS launch, so a single-GPU caller
        stats[6] = tb / n_b;      // needs no follow-up kernel (ranks re-derive them from the sums)
        __threadfence();
        ctl[0] = 0u;
    }
}

// Deterministic fold of per-group partials by the last CTA to finish.
// ctl[0] = CTA arrival counter, ctl[1] = bad-row counter; both return to 0.
__device__ __forceinline__ void fold_stats_if_last(u32* ctl, const double* partials, int64_t n_groups,
                                                   double n_m, double n_b, double* stats, double* red /*[2*nwarps]*/) {
    __shared__ int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const u32 done = atomicAdd(&ctl[0], 1u);
        s_last = (done == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    fold_stats_body(ctl, partials, n_groups, n_m, n_b, stats, red);
}

// ---------------------------------------------------------------------------
// bulk asynchronous copy global -> shared (the TMA engine's 1-D form) completing on an mbarrier
// ---------------------------------------------------------------------------
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, u32 arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// one thread: expect `bytes`, then start the copy; earlier generic-proxy accesses of the
// destination are ordered before the engine's writes by the proxy fence
__device__ __forceinline__ void bulk_fetch(void* dst, const void* src, u32 bytes, unsigned long long* bar) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    // the image streams through once: evict-first in L2, so it does not push out the df table
    unsigned long long policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long* bar, u32 parity) {
    u32 ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

// Large batches: the per-CTA arrival protocol (fence, atomic, two barriers) costs every CTA more than one
// tiny follow-up launch costs the batch, so there the fold is its own kernel (same summation order).
__global__ void k_fold_stats(u32* ctl, const double* partials, int64_t n_groups, double n_m, double n_b,
                             double* stats) {
    extern __shared__ __align__(16) unsigned char fold_smem[];
    fold_stats_body(ctl, partials, n_groups, n_m, n_b, stats, reinterpret_cast<double*>(fold_smem));
}

// ---------------------------------------------------------------------------
// fused scoring kernel
// ---------------------------------------------------------------------------
struct ScoreParams {
    DfView df;
    const double* gauss;
    const void* cand;
    const void* greedy;
    int cand_esz;
    int64_t n_cand;
    int cand_width;
    int64_t cand_stride;
    const void* refs;
    int ref_esz;
    int ref_width;
    int64_t ref_stride;
    const int64_t* ref_offsets;
    int refs_per_video;   // exact count (ref_offsets == NULL) or declared maximum
    int chunk_refs;       // refs staged per pass, <= 32
    int64_t n_videos;
    int cands_per_video;
    int use_eos;
    int mode;             // CST_BASE_* or -1: scores only
    int scb_k;
    const double* bcmr;
    float* reward;
    double* scores;
    double* stats;
    u32* ctl;
    u32* bad_out;         // += rows holding ids outside [-1, 65534] (= ctl + 1 when there is scratch)
    double* partials;
    double n_m, n_b;
    u32 tab_slots;        // capacity of the per-video shared table
    // staged references (cst_refs_cook): the result of phase R kept in HBM, one image per video
    const unsigned char* cooked;    // scoring from images: phase R becomes a copy
    const int64_t* cooked_index;    // image of video v = cooked_index[v] (NULL: v)
    int64_t cooked_stride;          // bytes per image
    unsigned char* cook_out;        // cooking: no candidates, phase R's result is written out
    int fold_apart;                 // the statistics are folded by k_fold_stats after this launch
    u32 list_cap;                   // k_score2: n-gram instances per batch (shared-memory list)
    int prefetch_ahead;             // k_score2: how many videos ahead a CTA prefetches rows into L2
    u32* excl;                      // leave-one-out counts [caption][df slot] (CST2_MODE_EXCL)
    size_t excl_stride;             // = df slots
    int n_local;                    // k_score2: candidate rows per video (cands_per_video, twice with greedy rows)
    int cand_rows;                  // k_score2: candidate rows per batch
};

#ifndef CST_FOLD_APART_CTAS
#define CST_FOLD_APART_CTAS 16   /* batches of at least this many CTAs per SM fold their statistics in k_fold_stats */
#endif

#include "cst_score2.cuh"

// Standalone baseline over (n_groups, S) scores: one warp per group.  Groups of up to 32 scores (every
// training shape) stay in registers, lane = caption: ranks by shuffle, no second pass over memory.
// fold_apart: the statistics are folded by k_fold_stats after the launch (large batches).
__global__ void __launch_bounds__(128) k_scb(const double* __restrict__ scores, const double* __restrict__ other,
                                             int64_t n_groups, int S, int mode, int k, float* reward,
                                             double* stats, u32* ctl, double* partials, double n_m, double n_b,
                                             int fold_apart) {
    extern __shared__ __align__(16) unsigned char smem[];
    double* red = reinterpret_cast<double*>(smem);   // [8]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t g = (int64_t)blockIdx.x * 4 + warp; g < n_groups; g += (int64_t)gridDim.x * 4) {
        double sm, sb;
        if (S <= 32) {
            const bool on = lane < S;
            const double s = on ? scores[g * S + lane] : 0.0;
            const double o = (on && other) ? other[g * S + lane] : 0.0;
            const bool scb = (mode == CST_BASE_SCB_GT || mode == CST_BASE_SCB_SAMPLE);
            const double x = (mode == CST_BASE_SCB_GT) ? o : s;
            double low = 0.0;
            if (scb) {
                int rank = 0;
                if (k < S) {                         // k lowest (utils.py:294-312): ties by index = stable sort
#pragma unroll 1
                    for (int j = 0; j < S; ++j) {
                        const double xj = __shfl_sync(CST_FULL, x, j);
                        rank += (xj < x || (xj == x && j < lane)) ? 1 : 0;
                    }
                }
                low = warp_sum((on && rank < k) ? x : 0.0);
            }
            const double base = scb ? low / (double)k : 0.0;
            if (on) reward[g * S + lane] = (float)((mode == CST_BASE_GREEDY) ? s - o : s - base);
            sm = warp_sum(s);
            sb = (mode == CST_BASE_SCB_SAMPLE) ? low : (mode == CST_BASE_NONE) ? 0.0 : warp_sum(o);
        } else {
            baseline_group(scores + g * S, other ? other + g * S : nullptr, S, mode, k, lane, reward + g * S, sm, sb);
        }
        if (lane == 0) {
            partials[2 * g] = sm;
            partials[2 * g + 1] = sb;
        }
    }
    if (!fold_apart) fold_stats_if_last(ctl, partials, n_groups, n_m, n_b, stats, red);
}

// ---------------------------------------------------------------------------
// host side of scoring
// ---------------------------------------------------------------------------
static int grams_per_row(int width) {
    int w = width < 32 ? width : 32, g = 0;
    for (int k = 1; k <= 4; ++k) g += (w - k + 1 > 0) ? (w - k + 1) : 0;
    return g;
}

static u32 next_pow2(u32 x) {
    u32 p = 1;
    while (p < x) p <<= 1;
    return p;
}

extern "C" int64_t cst_reward_scratch_bytes(int64_t n_groups) {
    if (n_groups < 0) return -1;
    return 256 + 16 * n_groups;
}

static void stat_counts(int mode, int64_t n_cand, int seq_per_img, int k, double& n_m, double& n_b) {
    n_m = (double)n_cand;
    switch (mode) {
        case CST_BASE_GREEDY: n_b = (double)n_cand; break;
        case CST_BASE_SCB_GT: n_b = (double)n_cand; break;
        case CST_BASE_SCB_SAMPLE: n_b = (double)(n_cand / seq_per_img) * (double)k; break;
        default: n_b = 1.0; break;   // b_score = 0 / 1
    }
}

extern "C" int64_t cst_refs_cooked_bytes(int32_t ref_width, int32_t refs_per_video) {
    if (ref_width < 1 || ref_width > CST_MAX_ROW_WIDTH || refs_per_video < 1 || refs_per_video > 32) return -1;
    return ((int64_t)CST2_IMG_BYTES + 127) / 128 * 128;       // one size: the table capacity is fixed
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is per function and device, shared by all host threads:
// remember the largest value set so far process-wide and only ever raise it.
static std::mutex g_smem_mutex;
static int g_sm_count[64] = {0};

static int device_sms(int dev, int* sms) {
    *sms = 148;
    if (dev < 0 || dev >= 64) return CST_OK;
    std::lock_guard<std::mutex> lock(g_smem_mutex);
    if (!g_sm_count[dev]) CUDA_TRY(cudaDeviceGetAttribute(&g_sm_count[dev], cudaDevAttrMultiProcessorCount, dev));
    *sms = g_sm_count[dev];
    return CST_OK;
}

template <int MODE, int ESZ, int NW, int CTAS>
static int launch_mode(const ScoreParams& p, unsigned grid, size_t smem, int dev, cudaStream_t st) {
    {
        static size_t set[64] = {0};                  // (one array per instantiation)
        std::lock_guard<std::mutex> lock(g_smem_mutex);
        if (dev < 0 || dev >= 64 || set[dev] < smem) {
            CUDA_TRY(cudaFuncSetAttribute(k_score2<MODE, NW, CTAS, ESZ>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)smem));
            if (dev >= 0 && dev < 64) set[dev] = smem;
        }
    }
    k_score2<MODE, NW, CTAS, ESZ><<<grid, NW * 32, smem, st>>>(p);
    COUNT_LAUNCH();
    CUDA_TRY(cudaGetLastError());
    return CST_OK;
}

// Batches that two CTAs per SM hold at once (BASELINE configs 1 and 3: 64 videos) are latency-bound: one
// video's chain of stages is the whole call.  CTAs of 16 warps halve the rounds of every stage
// (config 3 from a CUDA graph: 44.5 -> 34.3 us); at full occupancy the same CTAs would be 75 % slower.
#define CST2_SMALL_WARPS 16
#define CST2_SMALL_CTAS 2
template <int MODE, int ESZ>
static int launch_sized(const ScoreParams& p, unsigned grid, size_t smem_big, size_t smem_small, bool small, int dev,
                        cudaStream_t st) {
    return small ? launch_mode<MODE, ESZ, CST2_SMALL_WARPS, CST2_SMALL_CTAS>(p, grid, smem_small, dev, st)
                 : launch_mode<MODE, ESZ, CST2_WARPS, CST2_CTAS>(p, grid, smem_big, dev, st);
}

// One launch of the scoring kernel (raw rows, from staged images, or cooking images), plus the fold of
// the statistics on large batches.
static int launch_score2(const ScoreParams& p_in, cudaStream_t st) {
    ScoreParams p = p_in;
    p.tab_slots = CST2_TAB_SLOTS;
    p.list_cap = CST2_LIST_CAP;
    const int n_local = p.cands_per_video * (p.greedy ? 2 : 1);
    const int n_batches = (n_local + CST2_ROWS - 1) / CST2_ROWS;
    p.n_local = n_local;
    p.cand_rows = n_batches ? (n_local + n_batches - 1) / n_batches : 1;
    const size_t smem = score2_smem(n_local, p.cands_per_video, CST2_WARPS);
    const size_t smem_small = score2_smem(n_local, p.cands_per_video, CST2_SMALL_WARPS);
    if (smem_small > 227 * 1024) return fail(CST_E_RANGE, "scoring needs %zu bytes of shared memory per CTA", smem_small);
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    int sms = 148;
    int rc = device_sms(dev, &sms);
    if (rc != CST_OK) return rc;
    // one CTA per video: the hardware's CTA scheduler balances better than a persistent grid with a
    // static stride (measured again this round: 1 / 2 / 4 / 8 / 16 CTAs per resident slot 2.25 / 2.23 /
    // 2.17 / 2.11 / 2.06 ms against 2.03)
    const unsigned grid = (unsigned)(p.n_videos < 0x7fffffffLL ? p.n_videos : 0x7fffffffLL);
    const bool small = p.n_videos <= (int64_t)CST2_SMALL_CTAS * sms;
    p.fold_apart = (p.mode >= 0 && !p.cook_out && !p.excl && p.n_videos >= (int64_t)CST_FOLD_APART_CTAS * sms) ? 1 : 0;
    p.prefetch_ahead = sms * CST2_CTAS;
    // int32 rows on both sides (what the device-resident training loop passes) take an instantiation with
    // the element size compiled in; int16 / int64 / mixed rows the run-time one
    const bool i32 = p.cand_esz == 4 && (p.cooked || p.ref_esz == 4);
    if (p.excl)
        rc = launch_mode<CST2_MODE_EXCL, 0, CST2_WARPS, CST2_CTAS>(p, grid, smem, dev, st);
    else if (p.cook_out)
        rc = launch_mode<CST2_MODE_COOK, 0, CST2_WARPS, CST2_CTAS>(p, grid, smem, dev, st);
    else if (p.cooked)
        rc = i32 ? launch_sized<CST2_MODE_STAGED, 4>(p, grid, smem, smem_small, small, dev, st)
                 : launch_sized<CST2_MODE_STAGED, 0>(p, grid, smem, smem_small, small, dev, st);
    else
        rc = i32 ? launch_sized<CST2_MODE_RAW, 4>(p, grid, smem, smem_small, small, dev, st)
                 : launch_sized<CST2_MODE_RAW, 0>(p, grid, smem, smem_small, small, dev, st);
    if (rc != CST_OK) return rc;
    if (p.fold_apart) {
        k_fold_stats<<<1, 1024, 2 * 32 * sizeof(double), st>>>(p.ctl, p.partials, p.n_videos, p.n_m, p.n_b, p.stats);
        COUNT_LAUNCH();
        CUDA_TRY(cudaGetLastError());
    }
    return CST_OK;
}

static int score_common(const char* who, const cst_df_table* t, const void* cand, const void* greedy,
                        int32_t cand_dtype, int64_t n_cand, int32_t cand_width, int64_t cand_stride,
                        const void* refs, int32_t ref_dtype, int32_t ref_width, int64_t ref_stride,
                        const int64_t* ref_offsets, int32_t refs_per_video, int64_t n_videos,
                        int32_t cands_per_video, int32_t use_eos, int mode, int scb_k, const double* bcmr,
                        float* reward, double* scores, double* stats, void* scratch, int64_t scratch_bytes,
                        cudaStream_t st, const void* cooked = nullptr, const int64_t* cooked_index = nullptr,
                        int64_t cooked_stride = 0) {
    if (!t) return fail(CST_E_ARG, "%s: table is NULL", who);
    if (n_cand < 0 || n_videos < 0 || cands_per_video < 1) return fail(CST_E_ARG, "%s: negative size", who);
    if (n_cand != n_videos * (int64_t)cands_per_video)
        return fail(CST_E_ARG, "%s: n_cand (%lld) != n_videos (%lld) * cands_per_video (%d)", who,
                    (long long)n_cand, (long long)n_videos, cands_per_video);
    if (cand_width < 1 || cand_width > CST_MAX_ROW_WIDTH || ref_width < 1 || ref_width > CST_MAX_ROW_WIDTH)
        return fail(CST_E_RANGE, "%s: row widths %d/%d outside [1, %d]", who, cand_width, ref_width,
                    CST_MAX_ROW_WIDTH);
    if (cand_stride < cand_width || ref_stride < ref_width) return fail(CST_E_ARG, "%s: stride < width", who);
    if (!dtype_bytes(cand_dtype) || !dtype_bytes(ref_dtype))
        return fail(CST_E_ARG, "%s: bad dtype", who);
    if (refs_per_video < 0 || (refs_per_video == 0 && !ref_offsets))
        return fail(CST_E_ARG, "%s: refs_per_video must be > 0", who);
    if (cands_per_video > 4096) return fail(CST_E_RANGE, "%s: cands_per_video %d > 4096", who, cands_per_video);
    if (n_cand == 0) {
        return CST_OK;
    }
    if (!cand || (!refs && !cooked)) return fail(CST_E_ARG, "%s: NULL rows", who);
    if (cooked) {
        if (refs_per_video < 1 || refs_per_video > 32)
            return fail(CST_E_RANGE, "%s: staged references hold at most 32 refs per video (got %d)", who,
                        refs_per_video);
        if (cooked_stride != cst_refs_cooked_bytes(ref_width, refs_per_video))
            return fail(CST_E_ARG, "%s: image stride %lld does not match ref_width %d / refs_per_video %d", who,
                        (long long)cooked_stride, ref_width, refs_per_video);
        if (reinterpret_cast<uintptr_t>(cooked) & 15u) return fail(CST_E_ARG, "%s: images must be 16-byte aligned", who);
    }
    if (mode == CST_BASE_GREEDY && !greedy) return fail(CST_E_ARG, "%s: greedy rows required", who);

    DeviceGuard guard(t->device);
    if (!guard.ok) return fail(CST_E_CUDA, "%s: cannot select device %d", who, t->device);

    ScoreParams p;
    memset(&p, 0, sizeof(p));
    p.df = view_of(t);
    p.gauss = t->gauss;
    p.cand = cand;
    p.greedy = greedy;
    p.cand_esz = dtype_bytes(cand_dtype);
    p.ref_esz = dtype_bytes(ref_dtype);
    p.n_cand = n_cand;
    p.cand_width = cand_width;
    p.cand_stride = cand_stride;
    p.refs = refs;
    p.ref_width = ref_width;
    p.ref_stride = ref_stride;
    p.ref_offsets = cooked ? nullptr : ref_offsets;
    p.refs_per_video = refs_per_video;
    p.chunk_refs = refs_per_video < 32 ? (refs_per_video < 1 ? 1 : refs_per_video) : 32;
    p.cooked = reinterpret_cast<const unsigned char*>(cooked);
    p.cooked_index = cooked_index;
    p.cooked_stride = cooked_stride;
    p.n_videos = n_videos;
    p.cands_per_video = cands_per_video;
    p.use_eos = use_eos ? 1 : 0;
    p.mode = mode;
    p.scb_k = scb_k;
    p.bcmr = bcmr;
    p.reward = reward;
    p.scores = scores;
    p.stats = stats;
    if (mode >= 0 || scratch) {
        const int64_t need = (mode >= 0) ? cst_reward_scratch_bytes(n_videos) : 256;
        if (!scratch || scratch_bytes < need)
            return fail(CST_E_ARG, "%s: scratch too small (%lld < %lld bytes)", who, (long long)scratch_bytes,
                        (long long)need);
        p.ctl = reinterpret_cast<u32*>(scratch);   // [1] also collects bad rows in scores-only mode
        p.bad_out = p.ctl + 1;
        p.partials = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(scratch) + 256);
        if (mode >= 0) stat_counts(mode, n_cand, cands_per_video, scb_k, p.n_m, p.n_b);
    }
    // (the kernel addresses rows with 32-bit pitches)
    if ((uint64_t)cand_stride * 8u >= (1ull << 32) || (uint64_t)ref_stride * 8u >= (1ull << 32))
        return fail(CST_E_RANGE, "%s: rows more than 4 GB apart", who);
    return launch_score2(p, st);
}

extern "C" int cst_cider_scores(const cst_df_table* table, const void* cand, const void* greedy,
                                int32_t cand_dtype, int64_t n_cand, int32_t cand_width, int64_t cand_stride,
                                const void* refs, int32_t ref_dtype, int32_t ref_width, int64_t ref_stride,
                                const int64_t* ref_offsets, int32_t refs_per_video, int64_t n_videos,
                                int32_t cands_per_video, int32_t use_eos, double* scores_out, void* stream) {
    if (n_cand > 0 && !scores_out) return fail(CST_E_ARG, "cst_cider_scores: scores_out is NULL");
    return score_common("cst_cider_scores", table, cand, greedy, cand_dtype, n_cand, cand_width, cand_stride, refs,
                        ref_dtype, ref_width, ref_stride, ref_offsets, refs_per_video, n_videos, cands_per_video,
                        use_eos, -1, 0, nullptr, nullptr, scores_out, nullptr, nullptr, 0, (cudaStream_t)stream);
}

static int check_baseline_args(const char* who, int mode, int seq_per_img, int scb_k, const void* other) {
    if (mode < CST_BASE_NONE || mode > CST_BASE_SCB_SAMPLE) return fail(CST_E_ARG, "%s: unknown baseline mode %d", who, mode);
    if (seq_per_img < 1) return fail(CST_E_ARG, "%s: seq_per_img < 1", who);
    if ((mode == CST_BASE_SCB_GT || mode == CST_BASE_SCB_SAMPLE) && (scb_k < 1 || scb_k > seq_per_img))
        return fail(CST_E_ARG, "%s: scb_k %d outside [1, %d]", who, scb_k, seq_per_img);
    if (mode == CST_BASE_SCB_GT && !other) return fail(CST_E_ARG, "%s: bcmrscores required", who);
    return CST_OK;
}

extern "C" int cst_scb_apply(const double* scores, const double* other, int64_t n_rows, int32_t seq_per_img,
                             int32_t mode, int32_t scb_k, float* reward_out, double* stats_out, void* scratch,
                             int64_t scratch_bytes, int device, void* stream) {
    int rc = check_baseline_args("cst_scb_apply", mode, seq_per_img, scb_k, other);
    if (rc != CST_OK) return rc;
    if (n_rows < 0) return fail(CST_E_ARG, "cst_scb_apply: n_rows < 0");
    if (n_rows == 0) return CST_OK;
    if (!scores || !reward_out || !stats_out) return fail(CST_E_ARG, "cst_scb_apply: NULL argument");
    if (mode == CST_BASE_GREEDY && !other) return fail(CST_E_ARG, "cst_scb_apply: greedy scores required");
    if (seq_per_img > 4096) return fail(CST_E_RANGE, "cst_scb_apply: seq_per_img %d > 4096", seq_per_img);
    if (!scratch || scratch_bytes < cst_reward_scratch_bytes(n_rows))
        return fail(CST_E_ARG, "cst_scb_apply: scratch too small");
    DeviceGuard guard(device);
    if (!guard.ok) return fail(CST_E_CUDA, "cst_scb_apply: cannot select device %d", device);
    double n_m, n_b;
    stat_counts(mode, n_rows * seq_per_img, seq_per_img, scb_k, n_m, n_b);
    const size_t smem = 8 * sizeof(double);
    int64_t blocks = (n_rows + 3) / 4;
    if (blocks > 148 * 16) blocks = 148 * 16;
    // a 128-thread CTA folding tens of thousands of partials is most of such a launch: large batches fold apart
    const int fold_apart = n_rows >= 4096;
    u32* ctl = reinterpret_cast<u32*>(scratch);
    double* partials = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(scratch) + 256);
    k_scb<<<(unsigned)blocks, 128, smem, (cudaStream_t)stream>>>(scores, other, n_rows, seq_per_img, mode, scb_k,
                                                               reward_out, stats_out, ctl, partials, n_m, n_b, fold_apart);
    COUNT_LAUNCH();
    CUDA_TRY(cudaGetLastError());
    if (fold_apart) {
        k_fold_stats<<<1, 1024, 2 * 32 * sizeof(double), (cudaStream_t)stream>>>(ctl, partials, n_rows, n_m, n_b, stats_out);
        COUNT_LAUNCH();
        CUDA_TRY(cudaGetLastError());
    }
    return CST_OK;
}

static int reward_impl(const char* who, const cst_df_table* table, const void* cand, const void* greedy,
                       int32_t cand_dtype, int64_t n_cand, int32_t cand_width, int64_t cand_stride, const void* refs,
                       int32_t ref_dtype, int32_t ref_width, int64_t ref_stride, const int64_t* ref_offsets,
                       int32_t refs_per_video, int64_t n_videos, int32_t cands_per_video, int32_t use_eos,
                       int32_t mode, int32_t seq_per_img, int32_t scb_k, const double* bcmrscores, float* reward_out,
                       double* scores_out, double* stats_out, void* scratch, int64_t scratch_bytes, void* stream,
                       const void* cooked, const int64_t* cooked_index, int64_t cooked_stride) {
    int rc = check_baseline_args(who, mode, seq_per_img, scb_k, bcmrscores);
    if (rc != CST_OK) return rc;
    if (n_cand > 0 && (!reward_out || !stats_out)) return fail(CST_E_ARG, "%s: NULL output", who);
    if (n_cand % seq_per_img != 0)
        return fail(CST_E_ARG, "%s: n_cand %lld is not a multiple of seq_per_img %d", who, (long long)n_cand,
                    seq_per_img);
    if (cands_per_video == seq_per_img) {
        return score_common(who, table, cand, greedy, cand_dtype, n_cand, cand_width, cand_stride, refs,
                            ref_dtype, ref_width, ref_stride, ref_offsets, refs_per_video, n_videos,
                            cands_per_video, use_eos, mode, scb_k, bcmrscores, reward_out, scores_out, stats_out,
                            scratch, scratch_bytes, (cudaStream_t)stream, cooked, cooked_index, cooked_stride);
    }
    // baseline rows do not coincide with videos (expand_feat == 0): score, then baseline
    if (!scores_out) return fail(CST_E_ARG, "%s: scores_out is required when cands_per_video != seq_per_img", who);
    rc = score_common(who, table, cand, greedy, cand_dtype, n_cand, cand_width, cand_stride, refs, ref_dtype,
                      ref_width, ref_stride, ref_offsets, refs_per_video, n_videos, cands_per_video, use_eos, -1, 0,
                      nullptr, nullptr, scores_out, nullptr, scratch, scratch_bytes, (cudaStream_t)stream, cooked,
                      cooked_index, cooked_stride);
    if (rc != CST_OK) return rc;
    const double* other = (mode == CST_BASE_GREEDY) ? scores_out + n_cand : bcmrscores;
    return cst_scb_apply(scores_out, other, n_cand / seq_per_img, seq_per_img, mode, scb_k, reward_out, stats_out,
                         scratch, scratch_bytes, table->device, stream);
}

extern "C" int cst_reward(const cst_df_table* table, const void* cand, const void* greedy, int32_t cand_dtype,
                          int64_t n_cand, int32_t cand_width, int64_t cand_stride, const void* refs,
                          int32_t ref_dtype, int32_t ref_width, int64_t ref_stride, const int64_t* ref_offsets,
                          int32_t refs_per_video, int64_t n_videos, int32_t cands_per_video, int32_t use_eos,
                          int32_t mode, int32_t seq_per_img, int32_t scb_k, const double* bcmrscores,
                          float* reward_out, double* scores_out, double* stats_out, void* scratch,
                          int64_t scratch_bytes, void* stream) {
    return reward_impl("cst_reward", table, cand, greedy, cand_dtype, n_cand, cand_width, cand_stride, refs, ref_dtype,
                       ref_width, ref_stride, ref_offsets, refs_per_video, n_videos, cands_per_video, use_eos, mode,
                       seq_per_img, scb_k, bcmrscores, reward_out, scores_out, stats_out, scratch, scratch_bytes,
                       stream, nullptr, nullptr, 0);
}

extern "C" int cst_reward_cooked(const cst_df_table* table, const void* cand, const void* greedy, int32_t cand_dtype,
                                 int64_t n_cand, int32_t cand_width, int64_t cand_stride, const void* images,
                                 int64_t image_stride, const int64_t* image_index, int32_t ref_width,
                                 int32_t refs_per_video, int64_t n_videos, int32_t cands_per_video, int32_t use_eos,
                                 int32_t mode, int32_t seq_per_img, int32_t scb_k, const double* bcmrscores,
                                 float* reward_out, double* scores_out, double* stats_out, void* scratch,
                                 int64_t scratch_bytes, void* stream) {
    if (n_cand > 0 && !images) return fail(CST_E_ARG, "cst_reward_cooked: images is NULL");
    return reward_impl("cst_reward_cooked", table, cand, greedy, cand_dtype, n_cand, cand_width, cand_stride, nullptr,
                       CST_DTYPE_I32, ref_width, ref_width, nullptr, refs_per_video, n_videos, cands_per_video,
                       use_eos, mode, seq_per_img, scb_k, bcmrscores, reward_out, scores_out, stats_out, scratch,
                       scratch_bytes, stream, images, image_index, image_stride);
}

// The reference stages of k_score2 run once per video and kept: one image per video at images_out + v * image_stride
// (cst_refs_cooked_bytes).  bad_rows_out (device, optional) receives the number of reference rows that
// held ids outside [-1, 65534], plus one per video beyond one image (more than 32 refs or 1,100 n-gram
// instances: its image is marked invalid).
extern "C" int cst_refs_cook(const cst_df_table* t, const void* refs, int32_t ref_dtype, int64_t n_rows,
                             int32_t ref_width, int64_t ref_stride, const int64_t* ref_offsets,
                             int32_t refs_per_video, int64_t n_videos, int32_t use_eos, void* images_out,
                             int64_t image_stride, uint32_t* bad_rows_out, void* stream) {
    const char* who = "cst_refs_cook";
    if (!t) return fail(CST_E_ARG, "%s: table is NULL", who);
    if (n_videos < 0 || n_rows < 0) return fail(CST_E_ARG, "%s: negative size", who);
    if (ref_width < 1 || ref_width > CST_MAX_ROW_WIDTH)
        return fail(CST_E_RANGE, "%s: row width %d outside [1, %d]", who, ref_width, CST_MAX_ROW_WIDTH);
    if (ref_stride < ref_width) return fail(CST_E_ARG, "%s: stride < width", who);
    if (!dtype_bytes(ref_dtype)) return fail(CST_E_ARG, "%s: bad dtype", who);
    if (refs_per_video < 1 || refs_per_video > 32)
        return fail(CST_E_RANGE, "%s: at most 32 refs per video can be staged (got %d)", who, refs_per_video);
    if (!ref_offsets && n_rows != n_videos * (int64_t)refs_per_video)
        return fail(CST_E_ARG, "%s: n_rows (%lld) != n_videos (%lld) * refs_per_video (%d)", who, (long long)n_rows,
                    (long long)n_videos, refs_per_video);
    if (image_stride != cst_refs_cooked_bytes(ref_width, refs_per_video))
        return fail(CST_E_ARG, "%s: image stride %lld, expected %lld", who, (long long)image_stride,
                    (long long)cst_refs_cooked_bytes(ref_width, refs_per_video));
    if (n_videos == 0) return CST_OK;
    if (!refs || !images_out) return fail(CST_E_ARG, "%s: NULL argument", who);
    if (reinterpret_cast<uintptr_t>(images_out) & 15u) return fail(CST_E_ARG, "%s: images must be 16-byte aligned", who);
    DeviceGuard guard(t->device);
    if (!guard.ok) return fail(CST_E_CUDA, "%s: cannot select device %d", who, t->device);

    ScoreParams p;
    memset(&p, 0, sizeof(p));
    p.df = view_of(t);
    p.gauss = t->gauss;
    p.refs = refs;
    p.ref_esz = dtype_bytes(ref_dtype);
    p.cand_esz = p.ref_esz;
    p.ref_width = ref_width;
    p.cand_width = 1;
    p.ref_stride = ref_stride;
    p.ref_offsets = ref_offsets;
    p.refs_per_video = refs_per_video;
    p.chunk_refs = refs_per_video;
    p.n_videos = n_videos;
    p.cands_per_video = 0;
    p.use_eos = use_eos ? 1 : 0;
    p.mode = -1;
    p.bad_out = bad_rows_out;
    p.cook_out = reinterpret_cast<unsigned char*>(images_out);
    p.cooked_stride = image_stride;
    if ((uint64_t)ref_stride * 8u >= (1ull << 32)) return fail(CST_E_RANGE, "%s: rows more than 4 GB apart", who);
    return launch_score2(p, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------
// leave-one-out document frequencies (the corpus-mode df of compute_scores.py:81-92)
// ---------------------------------------------------------------------------
// excl_out[i * slots + s] += number of videos in which caption i is the only one holding the n-gram of
// df slot s (slots = cst_df_capacity).  The table must have been built from these very rows.
extern "C" int cst_df_sole_holders(const cst_df_table* t, const void* rows, int32_t dtype, int64_t n_rows,
                                   int32_t width, int64_t stride, int32_t per_video, int64_t n_videos,
                                   int32_t use_eos, uint32_t* excl_out, uint32_t* too_large_out, void* stream) {
    const char* who = "cst_df_sole_holders";
    if (!t) return fail(CST_E_ARG, "%s: table is NULL", who);
    if (n_videos < 0 || n_rows != n_videos * (int64_t)per_video) return fail(CST_E_ARG, "%s: bad sizes", who);
    if (per_video < 1 || per_video > 32) return fail(CST_E_RANGE, "%s: 1..32 captions per video (got %d)", who, per_video);
    if (width < 1 || width > CST_MAX_ROW_WIDTH) return fail(CST_E_RANGE, "%s: row width %d outside [1, %d]", who, width, CST_MAX_ROW_WIDTH);
    if (stride < width || (uint64_t)stride * 8u >= (1ull << 32)) return fail(CST_E_ARG, "%s: bad stride", who);
    if (!dtype_bytes(dtype)) return fail(CST_E_ARG, "%s: bad dtype", who);
    if (n_videos == 0) return CST_OK;
    if (!rows || !excl_out) return fail(CST_E_ARG, "%s: NULL argument", who);
    DeviceGuard guard(t->device);
    if (!guard.ok) return fail(CST_E_CUDA, "%s: cannot select device %d", who, t->device);
    ScoreParams p;
    memset(&p, 0, sizeof(p));
    p.df = view_of(t);
    p.gauss = t->gauss;
    p.refs = rows;
    p.ref_esz = dtype_bytes(dtype);
    p.cand_esz = p.ref_esz;
    p.ref_width = width;
    p.cand_width = 1;
    p.ref_stride = stride;
    p.refs_per_video = per_video;
    p.chunk_refs = per_video;
    p.n_videos = n_videos;
    p.use_eos = use_eos ? 1 : 0;
    p.mode = -1;
    p.bad_out = too_large_out;
    p.excl = excl_out;
    p.excl_stride = (size_t)t->n_buckets * CST_BUCKET;
    return launch_score2(p, (cudaStream_t)stream);
}

// idf of every stored n-gram for the pass that leaves caption `caption` out:
// idfv[s] = lut[df[s] - excl[caption][s]], and the unigram shortcut array follows
__global__ void k_fill_idf_loo(const u64* __restrict__ keys, const u32* __restrict__ dfc, const u32* __restrict__ excl,
                               const double* __restrict__ lut, double* idfv, double* uni, size_t slots) {
    const size_t s = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= slots) return;
    const u64 key = keys[s];
    if (key == 0ull) return;
    const u32 d = dfc[s] - excl[s];
    const double w = lut[d];
    idfv[s] = w;
    if ((key >> 16) == 0ull && (u32)key - 1u < CST_UNI_N) uni[(u32)key - 1u] = w;
}

extern "C" int cst_df_leave_one_out(cst_df_table* t, const uint32_t* excl, int32_t caption, void* stream) {
    if (!t || !excl || caption < 0) return fail(CST_E_ARG, "cst_df_leave_one_out: bad argument");
    DeviceGuard guard(t->device);
    if (!guard.ok) return fail(CST_E_CUDA, "cst_df_leave_one_out: cannot select device %d", t->device);
    const size_t slots = (size_t)t->n_buckets * CST_BUCKET;
    k_fill_idf_loo<<<(unsigned)((slots + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        t->keys, t->dfc, excl + (size_t)caption * slots, t->lut, t->idfv, t->uni, slots);
    COUNT_LAUNCH();
    CUDA_TRY(cudaGetLastError());
    return CST_OK;
}

// ---------------------------------------------------------------------------
// device-side document frequency (compute_ciderdf.py:56-70 on packed keys)
// ---------------------------------------------------------------------------
// Insert into a shared-memory key set; true iff this call created the entry.
__device__ __forceinline__ bool set_insert_new(u64* keys, u32 mask, u32 idx, u64 key) {
    for (;;) {
        u64 cur = *reinterpret_cast<volatile u64*>(&keys[idx]);
        if (cur == key) return false;
        if (cur == 0ull) {
            u64 old = atomicCAS(&keys[idx], 0ull, key);
            if (old == 0ull) return true;
            if (old == key) return false;
        }
        idx = (idx + 1) & mask;
    }
}

// One CTA per video: the set union of its refs' n-grams is formed in shared memory
// (compute_ciderdf.py:66-67); each member then adds 1 to its global counter (:68).
__global__ void __launch_bounds__(256) k_df_build(const void* __restrict__ rows, int esz, int width, int64_t stride,
                                                  int use_eos, const int64_t* __restrict__ video_offsets,
                                                  int64_t n_videos, u32 set_slots, int set_shift, u64* g_keys,
                                                  u32* g_cnt, u32 g_cap, u32* n_unique, u32* n_bad) {
    extern __shared__ __align__(16) unsigned char smem[];
    u64* set = reinterpret_cast<u64*>(smem);
    u32* tok_all = reinterpret_cast<u32*>(set + set_slots);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    u32* tokbuf = tok_all + warp * 32;
    for (int64_t video = blockIdx.x; video < n_videos; video += gridDim.x) {
        for (u32 i = threadIdx.x; i < set_slots; i += blockDim.x) set[i] = 0ull;
        __syncthreads();
        const int64_t r0 = video_offsets[video], r1 = video_offsets[video + 1];
        for (int64_t r = r0 + warp; r < r1; r += nwarps) {
            u64 key[4];
            int L;
            const char* row = reinterpret_cast<const char*>(rows) + (size_t)r * (size_t)stride * (size_t)esz;
            const bool bad = canon_row(row, esz, width, use_eos, lane, 32, tokbuf, key, L);
            if (bad && lane == 0) atomicAdd(n_bad, 1u);
            int tf[4];
            bool lead[4];
            row_tf(key, lane, 32, -1, tf, lead);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (!lead[k]) continue;
                const u32 h = cst_hash(key[k]);
                if (!set_insert_new(set, set_slots - 1u, h >> set_shift, key[k])) continue;
                u32 s = cst_slot(h, g_cap);
                for (;;) {
                    u64 cur = *reinterpret_cast<volatile u64*>(&g_keys[s]);
                    if (cur == key[k]) break;
                    if (cur == 0ull) {
                        u64 old = atomicCAS(&g_keys[s], 0ull, key[k]);
                        if (old == 0ull) { atomicAdd(n_unique, 1u); break; }
                        if (old == key[k]) break;
                    }
                    s = (s + 1 == g_cap) ? 0u : s + 1;
                }
                atomicAdd(&g_cnt[s], 1u);
            }
        }
        __syncthreads();
    }
}

// Move the counted keys into the final bucketed table.  Slots of a bucket fill front
// to back (the lookup relies on it: an empty last slot ends a probe).
__global__ void k_df_rehash(const u64* __restrict__ g_keys, const u32* __restrict__ g_cnt, u32 g_cap, u64* keys,
                            u32* dfc, u32 n_buckets, unsigned long long* over /* [n_buckets * 4], zeroed */) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g_cap) return;
    const u64 key = g_keys[i];
    if (key == 0ull) return;
    u32 b = cst_slot(cst_hash(key), n_buckets);
    for (;;) {
        for (int j = 0; j < CST_BUCKET; ++j) {
            const size_t s = (size_t)b * CST_BUCKET + j;
            if (atomicCAS(&keys[s], 0ull, key) == 0ull) {
                dfc[s] = g_cnt[i];
                return;
            }
        }
        over[(size_t)b * CST_BUCKET] = 1ull;      // full when this key came by (every writer stores the same)
        b = (b + 1 == n_buckets) ? 0u : b + 1;
    }
}

// The overflow mark of a full bucket is the order of its first two keys (bucket_resolve).
__global__ void k_df_order(u64* keys, u32* dfc, const unsigned long long* __restrict__ over, u32 n_buckets) {
    const u32 b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_buckets) return;
    u64* k = keys + (size_t)b * CST_BUCKET;
    if (k[CST_BUCKET - 1] == 0ull) return;
    const u64 k0 = k[0], k1 = k[1];
    if ((k0 > k1) != (over[(size_t)b * CST_BUCKET] != 0ull)) {
        u32* d = dfc + (size_t)b * CST_BUCKET;
        const u32 d0 = d[0];
        k[0] = k1;
        k[1] = k0;
        d[0] = d[1];
        d[1] = d0;
    }
}

extern "C" int cst_df_build_from_rows(const void* rows, int32_t dtype, int64_t n_rows, int32_t width,
                                      int64_t stride, const int64_t* video_offsets, int64_t n_videos,
                                      int32_t use_eos, const double* idf_by_df, int device, cst_df_table** out) {
    if (!out) return fail(CST_E_ARG, "cst_df_build_from_rows: out is NULL");
    *out = nullptr;
    if (n_rows < 0 || n_videos < 1 || !video_offsets || (n_rows > 0 && !rows))
        return fail(CST_E_ARG, "cst_df_build_from_rows: bad argument");
    {
        int lrc = check_lut("cst_df_build_from_rows", idf_by_df, n_videos);
        if (lrc != CST_OK) return lrc;
    }
    if (width < 1 || width > CST_MAX_ROW_WIDTH)
        return fail(CST_E_RANGE, "cst_df_build_from_rows: row width %d outside [1, %d]", width, CST_MAX_ROW_WIDTH);
    if (stride < width) return fail(CST_E_ARG, "cst_df_build_from_rows: stride < width");
    if (!dtype_bytes(dtype)) return fail(CST_E_ARG, "cst_df_build_from_rows: bad dtype");
    if (n_videos > 0x7fffffffLL) return fail(CST_E_RANGE, "cst_df_build_from_rows: too many videos");
    int64_t max_refs = 0;
    if (video_offsets[0] != 0 || video_offsets[n_videos] != n_rows)
        return fail(CST_E_ARG, "cst_df_build_from_rows: video_offsets must span [0, n_rows]");
    for (int64_t v = 0; v < n_videos; ++v) {
        const int64_t c = video_offsets[v + 1] - video_offsets[v];
        if (c < 0) return fail(CST_E_ARG, "cst_df_build_from_rows: video_offsets not monotone at %lld", (long long)v);
        if (c > max_refs) max_refs = c;
    }
    const int g_row = grams_per_row(width);
    const int64_t per_video = max_refs * g_row;
    if (per_video + per_video / 4 + 8 > (1 << 14))
        return fail(CST_E_RANGE, "cst_df_build_from_rows: %lld refs in one video exceed the shared-memory set",
                    (long long)max_refs);
    const int64_t instances = n_rows * (int64_t)g_row;
    if (2 * instances + 16 > 0xffffffffLL) return fail(CST_E_RANGE, "cst_df_build_from_rows: corpus too large");

    DeviceGuard guard(device);
    if (!guard.ok) return fail(CST_E_CUDA, "cst_df_build_from_rows: cannot select device %d", device);

    const u32 want = (u32)(per_video + per_video / 4 + 8);
    const u32 set_slots = next_pow2(want < 64 ? 64 : want);
    const int set_shift = 32 - __builtin_ctz(set_slots);
    const u32 g_cap = (u32)(2 * instances + 16);
    u64* g_keys = nullptr;
    u32* g_cnt = nullptr;
    u32* counters = nullptr;      // [0] unique keys, [1] bad rows
    int64_t* d_off = nullptr;
    cst_df_table* t = nullptr;
    int rc = CST_OK;
    cudaError_t e = cudaSuccess;
    u32 h_counters[2] = {0, 0};
    const size_t smem = (size_t)set_slots * sizeof(u64) + 8 * 32 * sizeof(u32);
    const unsigned grid = (unsigned)(n_videos < 148 * 32 ? n_videos : 148 * 32);
#define BUILD_TRY(expr)                                                                              \
    do {                                                                                             \
        e = (expr);                                                                                  \
        if (e != cudaSuccess) {                                                                      \
            rc = fail(e == cudaErrorMemoryAllocation ? CST_E_NOMEM : CST_E_CUDA, "%s failed: %s", #expr, \
                      cudaGetErrorString(e));                                                        \
            goto done;                                                                               \
        }                                                                                            \
    } while (0)
    BUILD_TRY(cudaMalloc(&g_keys, (size_t)g_cap * sizeof(u64)));
    BUILD_TRY(cudaMalloc(&g_cnt, (size_t)g_cap * sizeof(u32)));
    BUILD_TRY(cudaMalloc(&counters, 2 * sizeof(u32)));
    BUILD_TRY(cudaMalloc(&d_off, (size_t)(n_videos + 1) * sizeof(int64_t)));
    BUILD_TRY(cudaMemset(g_keys, 0, (size_t)g_cap * sizeof(u64)));
    BUILD_TRY(cudaMemset(g_cnt, 0, (size_t)g_cap * sizeof(u32)));
    BUILD_TRY(cudaMemset(counters, 0, 2 * sizeof(u32)));
    BUILD_TRY(cudaMemcpy(d_off, video_offsets, (size_t)(n_videos + 1) * sizeof(int64_t), cudaMemcpyHostToDevice));
    BUILD_TRY(cudaFuncSetAttribute(k_df_build, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_df_build<<<grid, 256, smem>>>(rows, dtype_bytes(dtype), width, stride, use_eos ? 1 : 0, d_off, n_videos,
                                    set_slots, set_shift, g_keys, g_cnt, g_cap, counters, counters + 1);
    COUNT_LAUNCH();
    BUILD_TRY(cudaGetLastError());
    BUILD_TRY(cudaMemcpy(h_counters, counters, sizeof(h_counters), cudaMemcpyDeviceToHost));
    if (h_counters[1] != 0) {
        rc = fail(CST_E_RANGE, "cst_df_build_from_rows: %u rows hold token ids outside [0, %d]", h_counters[1],
                  CST_MAX_TOKEN_ID);
        goto done;
    }
    rc = table_alloc(device, df_buckets_for((int64_t)h_counters[0]), idf_by_df, n_videos, &t);
    if (rc != CST_OK) goto done;
    // the idf array is filled last (table_finish): until then it holds the overflow marks
    BUILD_TRY(cudaMemset(t->idfv, 0, (size_t)t->n_buckets * CST_BUCKET * sizeof(double)));
    k_df_rehash<<<(g_cap + 255) / 256, 256>>>(g_keys, g_cnt, g_cap, t->keys, t->dfc, t->n_buckets,
                                              reinterpret_cast<unsigned long long*>(t->idfv));
    COUNT_LAUNCH();
    BUILD_TRY(cudaGetLastError());
    k_df_order<<<(t->n_buckets + 255) / 256, 256>>>(t->keys, t->dfc, reinterpret_cast<const unsigned long long*>(t->idfv),
                                                    t->n_buckets);
    COUNT_LAUNCH();
    BUILD_TRY(cudaGetLastError());
    BUILD_TRY(cudaDeviceSynchronize());
    t->size = (int64_t)h_counters[0];
    rc = table_finish(t);
    if (rc != CST_OK) goto done;
done:
#undef BUILD_TRY
    cudaFree(g_keys);
    cudaFree(g_cnt);
    cudaFree(counters);
    cudaFree(d_off);
    if (rc != CST_OK) {
        if (t) cst_df_destroy(t);
        return rc;
    }
    *out = t;
    return CST_OK;
}

// ---------------------------------------------------------------------------
// cst_reward_noalloc: the reward launch for hosts that cannot allocate
// ---------------------------------------------------------------------------
// Some hosts give every byte of HBM to their own memory pool when they start, so a
// later cudaMalloc from a plugin simply fails; others ban allocation inside their
// step loop because cudaMalloc / cudaFree stall the whole device.  For them the
// statistics scratch that cst_reward takes from the caller lives here instead, in a
// fixed array that is part of the library image.  Every GPU that loads the library
// holds these CST_NOALLOC_SCRATCH_BYTES (about 256 KiB) until the process exits.
#define CST_NOALLOC_SCRATCH_BYTES (256 + 16 * CST_NOALLOC_MAX_GROUPS)
static __device__ __align__(256) unsigned char g_noalloc_scratch[CST_NOALLOC_SCRATCH_BYTES];

static std::mutex g_noalloc_mutex;
static void* g_noalloc_ptr[64] = {nullptr};

// Device address of g_noalloc_scratch on `device` (which must be current).  In host
// code the symbol itself is only a stand-in, so the address is asked for once per
// device and remembered.
static int noalloc_scratch(int device, void** out) {
    if (device < 0 || device >= 64) return fail(CST_E_RANGE, "cst_reward_noalloc: device %d outside [0, 63]", device);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr)
            return fail(CST_E_CUDA, "cst_reward_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
        g_noalloc_ptr[device] = ptr;
    }
    *out = g_noalloc_ptr[device];
    return CST_OK;
}

extern "C" int cst_reward_noalloc(const cst_df_table* table, const void* cand, const void* greedy,
                                  int32_t cand_dtype, int64_t n_cand, int32_t cand_width, int64_t cand_stride,
                                  const void* refs, int32_t ref_dtype, int32_t ref_width, int64_t ref_stride,
                                  const int64_t* ref_offsets, int32_t refs_per_video, int64_t n_videos,
                                  int32_t cands_per_video, int32_t use_eos, int32_t mode, int32_t seq_per_img,
                                  int32_t scb_k, const double* bcmrscores, float* reward_out, double* scores_out,
                                  double* stats_out, void* stream) {
    if (!table) return fail(CST_E_ARG, "cst_reward_noalloc: table is NULL");
    if (seq_per_img < 1) return fail(CST_E_ARG, "cst_reward_noalloc: seq_per_img < 1");
    const int64_t groups = (n_videos > n_cand / seq_per_img) ? n_videos : n_cand / seq_per_img;
    if (groups > CST_NOALLOC_MAX_GROUPS)
        return fail(CST_E_RANGE,
                    "cst_reward_noalloc: %lld videos exceed the built-in scratch (CST_NOALLOC_MAX_GROUPS = %d); "
                    "split the batch or use cst_reward with caller-provided scratch",
                    (long long)groups, CST_NOALLOC_MAX_GROUPS);
    void* scratch = nullptr;
    {
        DeviceGuard guard(table->device);
        if (!guard.ok) return fail(CST_E_CUDA, "cst_reward_noalloc: cannot select device %d", table->device);
        int rc = noalloc_scratch(table->device, &scratch);
        if (rc != CST_OK) return rc;
        // the array keeps whatever the previous call (possibly an aborted one) left:
        // the arrival and bad-row counters must start at zero
        CUDA_TRY(cudaMemsetAsync(scratch, 0, 256, (cudaStream_t)stream));
    }
    return cst_reward(table, cand, greedy, cand_dtype, n_cand, cand_width, cand_stride, refs, ref_dtype, ref_width,
                      ref_stride, ref_offsets, refs_per_video, n_videos, cands_per_video, use_eos, mode, seq_per_img,
                      scb_k, bcmrscores, reward_out, scores_out, stats_out, scratch, CST_NOALLOC_SCRATCH_BYTES, stream);
}
