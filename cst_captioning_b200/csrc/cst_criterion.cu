// Fused consumer of the reward: the policy-gradient loss of model.py:7-23
// (RewardCriterion) and its gradient, taking the per-caption reward column directly
// so the (N, T) broadcast of utils.py:224,322 is never materialised.
//
//   mask[i][0] = 1, mask[i][t] = (seq[i][t-1] > 0)          (model.py:15-18)
//   loss = sum(-logp * reward * mask) / sum(mask)            (model.py:20-21)
//   dloss/dlogp[i][t] = -reward[i] * mask[i][t] / sum(mask)
//
// Both kernels stream their inputs once: HBM-bound by construction (per token 8 B of
// int64 seq + 4 B of logp read; the backward also writes 4 B).
#include <stdint.h>

#include "cst_common.cuh"
#include "cst_internal.h"
#include "cst_reward.h"

#define CRIT_THREADS 256
#ifndef CRIT_ROWS_PER_ITER
#define CRIT_ROWS_PER_ITER 2      /* rows in flight per warp; measured: 2 rows at 8 CTAs/SM beats 4 rows (spills) and 1 */
#endif
#ifndef CRIT_MIN_CTAS
#define CRIT_MIN_CTAS 8        /* 32 registers: 64 warps per SM */
#endif

__device__ __forceinline__ double crit_warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(CST_FULL, v, o);
    return v;
}

// One warp per row, lane = column (T <= 32 per pass, wider rows loop).  Per-CTA partial
// sums go to scratch; the last CTA adds them in a fixed order (deterministic loss).
template <typename ST>
__global__ void __launch_bounds__(CRIT_THREADS, CRIT_MIN_CTAS) k_criterion_fwd(
    const ST* __restrict__ seq, int64_t seq_stride, const float* __restrict__ logp, int64_t logp_stride,
    const float* __restrict__ reward, int64_t reward_row_stride, int64_t reward_col_stride, int64_t n_rows, int width,
    float* loss_out, float* den_out, u32* ctl, double* partials) {
    __shared__ double s_red[2 * (CRIT_THREADS / 32)];
    __shared__ int s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t warps_total = (int64_t)gridDim.x * (CRIT_THREADS / 32);
    const int64_t gw = (int64_t)blockIdx.x * (CRIT_THREADS / 32) + warp;
    double num = 0.0, den = 0.0;
    for (int64_t r0 = gw * CRIT_ROWS_PER_ITER; r0 < n_rows; r0 += warps_total * CRIT_ROWS_PER_ITER) {
        for (int t = lane; t < width; t += 32) {
            // all loads of the CRIT_ROWS_PER_ITER rows first, then the arithmetic: the kernel
            // is a pure stream and lives on the number of loads in flight
            ST prev[CRIT_ROWS_PER_ITER];
            float lp[CRIT_ROWS_PER_ITER], rw[CRIT_ROWS_PER_ITER];
#pragma unroll
            for (int q = 0; q < CRIT_ROWS_PER_ITER; ++q) {
                const int64_t i = r0 + q;
                const bool in = i < n_rows;
                prev[q] = (in && t > 0) ? __ldcs(seq + i * seq_stride + (t - 1)) : (ST)(in ? 1 : 0);
                lp[q] = in ? __ldcs(logp + i * logp_stride + t) : 0.0f;
                rw[q] = in ? __ldg(reward + i * reward_row_stride + t * reward_col_stride) : 0.0f;
            }
#pragma unroll
            for (int q = 0; q < CRIT_ROWS_PER_ITER; ++q) {
                if (prev[q] > 0) {                         // mask: first column, or previous token > 0
                    num += (double)(-lp[q] * rw[q]);
                    den += 1.0;
                }
            }
        }
    }
    num = crit_warp_sum(num);
    den = crit_warp_sum(den);
    if (lane == 0) {
        s_red[2 * warp] = num;
        s_red[2 * warp + 1] = den;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < CRIT_THREADS / 32; ++w) {
            a += s_red[2 * w];
            b += s_red[2 * w + 1];
        }
        partials[2 * blockIdx.x] = a;
        partials[2 * blockIdx.x + 1] = b;
        __threadfence();
        s_last = (atomicAdd(&ctl[0], 1u) == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double a = 0.0, b = 0.0;
    for (int g = threadIdx.x; g < (int)gridDim.x; g += CRIT_THREADS) {
        a += __ldcg(partials + 2 * g);
        b += __ldcg(partials + 2 * g + 1);
    }
    a = crit_warp_sum(a);
    b = crit_warp_sum(b);
    __syncthreads();
    if (lane == 0) {
        s_red[2 * warp] = a;
        s_red[2 * warp + 1] = b;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ta = 0.0, tb = 0.0;
        for (int w = 0; w < CRIT_THREADS / 32; ++w) {
            ta += s_red[2 * w];
            tb += s_red[2 * w + 1];
        }
        *loss_out = (float)(ta / tb);
        *den_out = (float)tb;
        ctl[0] = 0u;
    }
}

template <typename ST>
__global__ void __launch_bounds__(CRIT_THREADS, CRIT_MIN_CTAS) k_criterion_bwd(
    const ST* __restrict__ seq, int64_t seq_stride, const float* __restrict__ reward, int64_t reward_row_stride,
    int64_t reward_col_stride, const float* __restrict__ den, const float* __restrict__ grad_out, int64_t n_rows,
    int width, float* grad_logp, int64_t grad_stride) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t warps_total = (int64_t)gridDim.x * (CRIT_THREADS / 32);
    const int64_t gw = (int64_t)blockIdx.x * (CRIT_THREADS / 32) + warp;
    const float scale = -(*grad_out) / (*den);
    for (int64_t r0 = gw * CRIT_ROWS_PER_ITER; r0 < n_rows; r0 += warps_total * CRIT_ROWS_PER_ITER) {
        for (int t = lane; t < width; t += 32) {
            ST prev[CRIT_ROWS_PER_ITER];
            float rw[CRIT_ROWS_PER_ITER];
#pragma unroll
            for (int q = 0; q < CRIT_ROWS_PER_ITER; ++q) {
                const int64_t i = r0 + q;
                const bool in = i < n_rows;
                prev[q] = (in && t > 0) ? __ldcs(seq + i * seq_stride + (t - 1)) : (ST)1;
                rw[q] = in ? __ldg(reward + i * reward_row_stride + t * reward_col_stride) : 0.0f;
            }
#pragma unroll
            for (int q = 0; q < CRIT_ROWS_PER_ITER; ++q) {
                const int64_t i = r0 + q;
                if (i < n_rows) __stcs(grad_logp + i * grad_stride + t, prev[q] > 0 ? scale * rw[q] : 0.0f);
            }
        }
    }
}

// ---------------------------------------------------------------------------
// Dense (N, T) inputs -- what the training loop passes -- as ONE element stream: a thread
// takes 4 consecutive elements per slot (16-byte loads of logp, 16/32 bytes of seq) and
// keeps CRIT_SLOTS slots in flight, so every lane moves full vectors whatever T is (the
// row-per-warp kernels above leave 4 of 32 lanes idle at T = 28 and load 4/8-byte scalars).
// (row, t) of a thread's elements advance by a fixed step: no division inside the loop.
// The mask of an element needs the token BEFORE it: inside a group that is the thread's
// own data, for the group's first element it comes from the lane to the left (lane 0:
// one extra scalar load of a sector its neighbour warp streams anyway).
// ---------------------------------------------------------------------------
#ifndef CRIT_SLOTS
#define CRIT_SLOTS 2
#endif
#ifndef CRIT_DENSE_CTAS
#define CRIT_DENSE_CTAS 4
#endif

// bit k = (token k of the group > 0)
__device__ __forceinline__ unsigned crit_pos4(const int64_t* p) {
    const longlong2 a = __ldcs(reinterpret_cast<const longlong2*>(p));
    const longlong2 b = __ldcs(reinterpret_cast<const longlong2*>(p) + 1);
    return (a.x > 0 ? 1u : 0u) | (a.y > 0 ? 2u : 0u) | (b.x > 0 ? 4u : 0u) | (b.y > 0 ? 8u : 0u);
}
__device__ __forceinline__ unsigned crit_pos4(const int32_t* p) {
    const int4 a = __ldcs(reinterpret_cast<const int4*>(p));
    return (a.x > 0 ? 1u : 0u) | (a.y > 0 ? 2u : 0u) | (a.z > 0 ? 4u : 0u) | (a.w > 0 ? 8u : 0u);
}

struct CritCursor {       // (row, t) of the first element of a thread's group, and how it advances
    int64_t row;
    int t;
    __device__ __forceinline__ void advance(int64_t drow, int dt, int width) {
        row += drow;
        t += dt;
        if (t >= width) {
            t -= width;
            ++row;
        }
    }
};

template <typename ST>
__global__ void __launch_bounds__(CRIT_THREADS, CRIT_DENSE_CTAS) k_criterion_fwd_dense(
    const ST* __restrict__ seq, const float* __restrict__ logp, const float* __restrict__ reward,
    int64_t reward_row_stride, int64_t n_rows, int width, float* loss_out, float* den_out, u32* ctl,
    double* partials) {
    __shared__ double s_red[2 * (CRIT_THREADS / 32)];
    __shared__ int s_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t n_elem = n_rows * (int64_t)width;
    const int64_t n_groups = n_elem >> 2;
    const int64_t nt = (int64_t)gridDim.x * CRIT_THREADS;
    const int64_t g0 = (int64_t)blockIdx.x * CRIT_THREADS + threadIdx.x;
    const int64_t slot_drow = (4 * nt) / width;                 // one slot further = 4 * nt elements further
    const int slot_dt = (int)((4 * nt) - slot_drow * width);
    CritCursor cur;
    cur.row = (4 * g0) / width;
    cur.t = (int)(4 * g0 - cur.row * width);
    double num = 0.0;
    unsigned den = 0;
#pragma unroll 1
    for (int64_t gb = g0 - lane; gb < n_groups; gb += nt * CRIT_SLOTS) {      // warp-uniform trip count
        unsigned pos[CRIT_SLOTS], left[CRIT_SLOTS];
        float4 lp[CRIT_SLOTS];
        float rw_a[CRIT_SLOTS], rw_b[CRIT_SLOTS];
        CritCursor c[CRIT_SLOTS];
#pragma unroll
        for (int u = 0; u < CRIT_SLOTS; ++u) {                  // every load of every slot first
            const int64_t g = gb + lane + (int64_t)u * nt;
            const bool in = g < n_groups;
            c[u] = cur;
            cur.advance(slot_drow, slot_dt, width);
            pos[u] = in ? crit_pos4(seq + 4 * g) : 0u;
            lp[u] = in ? __ldcs(reinterpret_cast<const float4*>(logp) + g) : make_float4(0.f, 0.f, 0.f, 0.f);
            left[u] = (in && lane == 0 && g > 0) ? (__ldg(seq + 4 * g - 1) > 0 ? 1u : 0u) : 0u;
            rw_a[u] = in ? __ldg(reward + c[u].row * reward_row_stride) : 0.0f;
            rw_b[u] = (in && c[u].t + 3 >= width && c[u].row + 1 < n_rows)
                          ? __ldg(reward + (c[u].row + 1) * reward_row_stride) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < CRIT_SLOTS; ++u) {
            const unsigned up = __shfl_up_sync(CST_FULL, pos[u] >> 3, 1);
            const unsigned prev = ((pos[u] << 1) | (lane == 0 ? left[u] : up)) & 0xfu;   // bit k = token before element k > 0
            const bool in = gb + lane + (int64_t)u * nt < n_groups;
            const float l4[4] = {lp[u].x, lp[u].y, lp[u].z, lp[u].w};
            float acc = 0.0f;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int tk = c[u].t + k;
                const bool wrap = tk >= width;
                if (wrap) tk -= width;
                const bool m = in && (tk == 0 || ((prev >> k) & 1u));
                if (m) {
                    acc += -l4[k] * (wrap ? rw_b[u] : rw_a[u]);
                    ++den;
                }
            }
            num += (double)acc;
        }
    }
    if (g0 == 0) {                                              // the n_elem % 4 elements after the last group
        for (int64_t e = 4 * n_groups; e < n_elem; ++e) {
            const int64_t row = e / width;
            const int t = (int)(e - row * width);
            if (t == 0 || seq[e - 1] > 0) {
                num += (double)(-logp[e] * reward[row * reward_row_stride]);
                ++den;
            }
        }
    }
    num = crit_warp_sum(num);
    double dden = crit_warp_sum((double)den);
    if (lane == 0) {
        s_red[2 * warp] = num;
        s_red[2 * warp + 1] = dden;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < CRIT_THREADS / 32; ++w) {
            a += s_red[2 * w];
            b += s_red[2 * w + 1];
        }
        partials[2 * blockIdx.x] = a;
        partials[2 * blockIdx.x + 1] = b;
        __threadfence();
        s_last = (atomicAdd(&ctl[0], 1u) == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double a = 0.0, b = 0.0;
    for (int g = threadIdx.x; g < (int)gridDim.x; g += CRIT_THREADS) {
        a += __ldcg(partials + 2 * g);
        b += __ldcg(partials + 2 * g + 1);
    }
    a = crit_warp_sum(a);
    b = crit_warp_sum(b);
    __syncthreads();
    if (lane == 0) {
        s_red[2 * warp] = a;
        s_red[2 * warp + 1] = b;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ta = 0.0, tb = 0.0;
        for (int w = 0; w < CRIT_THREADS / 32; ++w) {
            ta += s_red[2 * w];
            tb += s_red[2 * w + 1];
        }
        *loss_out = (float)(ta / tb);
        *den_out = (float)tb;
        ctl[0] = 0u;
    }
}

template <typename ST>
__global__ void __launch_bounds__(CRIT_THREADS, CRIT_DENSE_CTAS) k_criterion_bwd_dense(
    const ST* __restrict__ seq, const float* __restrict__ reward, int64_t reward_row_stride,
    const float* __restrict__ den, const float* __restrict__ grad_out, int64_t n_rows, int width,
    float* __restrict__ grad_logp) {
    const int lane = threadIdx.x & 31;
    const int64_t n_elem = n_rows * (int64_t)width;
    const int64_t n_groups = n_elem >> 2;
    const int64_t nt = (int64_t)gridDim.x * CRIT_THREADS;
    const int64_t g0 = (int64_t)blockIdx.x * CRIT_THREADS + threadIdx.x;
    const int64_t slot_drow = (4 * nt) / width;
    const int slot_dt = (int)((4 * nt) - slot_drow * width);
    CritCursor cur;
    cur.row = (4 * g0) / width;
    cur.t = (int)(4 * g0 - cur.row * width);
    const float scale = -(*grad_out) / (*den);
#pragma unroll 1
    for (int64_t gb = g0 - lane; gb < n_groups; gb += nt * CRIT_SLOTS) {
        unsigned pos[CRIT_SLOTS], left[CRIT_SLOTS];
        float rw_a[CRIT_SLOTS], rw_b[CRIT_SLOTS];
        CritCursor c[CRIT_SLOTS];
#pragma unroll
        for (int u = 0; u < CRIT_SLOTS; ++u) {
            const int64_t g = gb + lane + (int64_t)u * nt;
            const bool in = g < n_groups;
            c[u] = cur;
            cur.advance(slot_drow, slot_dt, width);
            pos[u] = in ? crit_pos4(seq + 4 * g) : 0u;
            left[u] = (in && lane == 0 && g > 0) ? (__ldg(seq + 4 * g - 1) > 0 ? 1u : 0u) : 0u;
            rw_a[u] = in ? __ldg(reward + c[u].row * reward_row_stride) : 0.0f;
            rw_b[u] = (in && c[u].t + 3 >= width && c[u].row + 1 < n_rows)
                          ? __ldg(reward + (c[u].row + 1) * reward_row_stride) : 0.0f;
        }
#pragma unroll
        for (int u = 0; u < CRIT_SLOTS; ++u) {
            const unsigned up = __shfl_up_sync(CST_FULL, pos[u] >> 3, 1);
            const unsigned prev = ((pos[u] << 1) | (lane == 0 ? left[u] : up)) & 0xfu;
            const int64_t g = gb + lane + (int64_t)u * nt;
            float o[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int tk = c[u].t + k;
                const bool wrap = tk >= width;
                if (wrap) tk -= width;
                const bool m = tk == 0 || ((prev >> k) & 1u);
                o[k] = m ? scale * (wrap ? rw_b[u] : rw_a[u]) : 0.0f;
            }
            if (g < n_groups) __stcs(reinterpret_cast<float4*>(grad_logp) + g, make_float4(o[0], o[1], o[2], o[3]));
        }
    }
    if (g0 == 0) {
        for (int64_t e = 4 * n_groups; e < n_elem; ++e) {
            const int64_t row = e / width;
            const int t = (int)(e - row * width);
            grad_logp[e] = (t == 0 || seq[e - 1] > 0) ? scale * reward[row * reward_row_stride] : 0.0f;
        }
    }
}

static bool crit_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int crit_dense_grid(int64_t n_rows, int width) {
    const int64_t groups = (n_rows * (int64_t)width) >> 2;
    int64_t want = (groups + (int64_t)CRIT_THREADS * CRIT_SLOTS - 1) / ((int64_t)CRIT_THREADS * CRIT_SLOTS);
    const int64_t cap = 148 * CRIT_DENSE_CTAS;
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    return (int)want;
}

static int crit_grid(int64_t n_rows) {
    int64_t want = (n_rows + (CRIT_THREADS / 32) * CRIT_ROWS_PER_ITER - 1) / ((CRIT_THREADS / 32) * CRIT_ROWS_PER_ITER);
    const int64_t cap = 148 * 8;      // 8 resident CTAs of 256 threads per SM
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    return (int)want;
}

extern "C" int64_t cst_criterion_scratch_bytes(void) { return 256 + 16 * 148 * 8; }

static int crit_check(const char* who, const void* seq, int32_t seq_dtype, const void* reward, int64_t n_rows,
                      int32_t width, int64_t seq_stride) {
    if (n_rows < 0 || width < 1) return cst_fail(CST_E_ARG, "%s: bad shape", who);
    if (seq_dtype != CST_DTYPE_I32 && seq_dtype != CST_DTYPE_I64) return cst_fail(CST_E_ARG, "%s: bad seq dtype", who);
    if (n_rows > 0 && (!seq || !reward)) return cst_fail(CST_E_ARG, "%s: NULL input", who);
    if (seq_stride < width) return cst_fail(CST_E_ARG, "%s: seq stride < width", who);
    return CST_OK;
}

extern "C" int cst_reward_criterion_fwd(const void* seq, int32_t seq_dtype, int64_t seq_stride, const float* logprobs,
                                        int64_t logp_stride, const float* reward, int64_t reward_row_stride,
                                        int64_t reward_col_stride, int64_t n_rows, int32_t width, float* loss_out,
                                        float* den_out, void* scratch, int64_t scratch_bytes, int device,
                                        void* stream) {
    int rc = crit_check("cst_reward_criterion_fwd", seq, seq_dtype, reward, n_rows, width, seq_stride);
    if (rc != CST_OK) return rc;
    if (n_rows == 0) return cst_fail(CST_E_ARG, "cst_reward_criterion_fwd: empty batch (the loss would be 0/0)");
    if (!logprobs || !loss_out || !den_out) return cst_fail(CST_E_ARG, "cst_reward_criterion_fwd: NULL argument");
    if (!scratch || scratch_bytes < cst_criterion_scratch_bytes())
        return cst_fail(CST_E_ARG, "cst_reward_criterion_fwd: scratch too small");
    CstDeviceGuard guard(device);
    if (!guard.ok) return cst_fail(CST_E_CUDA, "cst_reward_criterion_fwd: cannot select device %d", device);
    u32* ctl = reinterpret_cast<u32*>(scratch);
    double* partials = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(scratch) + 256);
    cudaStream_t st = (cudaStream_t)stream;
    const bool dense = seq_stride == width && logp_stride == width && reward_col_stride == 0 && width >= 4 &&
                       crit_aligned16(seq) && crit_aligned16(logprobs);
    if (dense) {
        const int dgrid = crit_dense_grid(n_rows, width);
        if (seq_dtype == CST_DTYPE_I64)
            k_criterion_fwd_dense<int64_t><<<dgrid, CRIT_THREADS, 0, st>>>(
                (const int64_t*)seq, logprobs, reward, reward_row_stride, n_rows, width, loss_out, den_out, ctl, partials);
        else
            k_criterion_fwd_dense<int32_t><<<dgrid, CRIT_THREADS, 0, st>>>(
                (const int32_t*)seq, logprobs, reward, reward_row_stride, n_rows, width, loss_out, den_out, ctl, partials);
        cst_count_launch();
        CST_CUDA_TRY(cudaGetLastError());
        return CST_OK;
    }
    const int grid = crit_grid(n_rows);
    if (seq_dtype == CST_DTYPE_I64)
        k_criterion_fwd<int64_t><<<grid, CRIT_THREADS, 0, st>>>((const int64_t*)seq, seq_stride, logprobs, logp_stride,
                                                                 reward, reward_row_stride, reward_col_stride, n_rows,
                                                                 width, loss_out, den_out, ctl, partials);
    else
        k_criterion_fwd<int32_t><<<grid, CRIT_THREADS, 0, st>>>((const int32_t*)seq, seq_stride, logprobs, logp_stride,
                                                                 reward, reward_row_stride, reward_col_stride, n_rows,
                                                                 width, loss_out, den_out, ctl, partials);
    cst_count_launch();
    CST_CUDA_TRY(cudaGetLastError());
    return CST_OK;
}

extern "C" int cst_reward_criterion_bwd(const void* seq, int32_t seq_dtype, int64_t seq_stride, const float* reward,
                                        int64_t reward_row_stride, int64_t reward_col_stride, const float* den,
                                        const float* grad_out, int64_t n_rows, int32_t width, float* grad_logprobs,
                                        int64_t grad_stride, int device, void* stream) {
    int rc = crit_check("cst_reward_criterion_bwd", seq, seq_dtype, reward, n_rows, width, seq_stride);
    if (rc != CST_OK) return rc;
    if (n_rows == 0) return CST_OK;
    if (!den || !grad_out || !grad_logprobs) return cst_fail(CST_E_ARG, "cst_reward_criterion_bwd: NULL argument");
    if (grad_stride < width) return cst_fail(CST_E_ARG, "cst_reward_criterion_bwd: grad stride < width");
    CstDeviceGuard guard(device);
    if (!guard.ok) return cst_fail(CST_E_CUDA, "cst_reward_criterion_bwd: cannot select device %d", device);
    cudaStream_t st = (cudaStream_t)stream;
    const bool dense = seq_stride == width && grad_stride == width && reward_col_stride == 0 && width >= 4 &&
                       crit_aligned16(seq) && crit_aligned16(grad_logprobs);
    if (dense) {
        const int dgrid = crit_dense_grid(n_rows, width);
        if (seq_dtype == CST_DTYPE_I64)
            k_criterion_bwd_dense<int64_t><<<dgrid, CRIT_THREADS, 0, st>>>(
                (const int64_t*)seq, reward, reward_row_stride, den, grad_out, n_rows, width, grad_logprobs);
        else
            k_criterion_bwd_dense<int32_t><<<dgrid, CRIT_THREADS, 0, st>>>(
                (const int32_t*)seq, reward, reward_row_stride, den, grad_out, n_rows, width, grad_logprobs);
        cst_count_launch();
        CST_CUDA_TRY(cudaGetLastError());
        return CST_OK;
    }
    const int grid = crit_grid(n_rows);
    if (seq_dtype == CST_DTYPE_I64)
        k_criterion_bwd<int64_t><<<grid, CRIT_THREADS, 0, st>>>((const int64_t*)seq, seq_stride, reward,
                                                                 reward_row_stride, reward_col_stride, den, grad_out,
                                                                 n_rows, width, grad_logprobs, grad_stride);
    else
        k_criterion_bwd<int32_t><<<grid, CRIT_THREADS, 0, st>>>((const int32_t*)seq, seq_stride, reward,
                                                                 reward_row_stride, reward_col_stride, den, grad_out,
                                                                 n_rows, width, grad_logprobs, grad_stride);
    cst_count_launch();
    CST_CUDA_TRY(cudaGetLastError());
    return CST_OK;
}
