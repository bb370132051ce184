// kernels_topk.cuh -- per-query streaming top-k over a database (`similarity(idxs)` over Q x N_t instead of
// N_t x N_t, algorithm_template.py:121-140, with the ranking getEvalStatistics applies, :235).  No reference counterpart as a function:
// the reference materialises the whole matrix on disk and argsorts it.
//
// Per query row the device keeps TK_SLOTS 64-bit keys: slots [0, k) = the best k so far, slots [k, TK_SLOTS) = the scores of the
// database chunk just computed.  key = (order-preserving image of the normalised score) << 32 | database position, so an ascending
// sort ranks by similarity with ties towards the lowest position (what a stable argsort of the full row gives).
#pragma once
#include "common.cuh"

#define TK_SLOTS 4096
#define TK_THREADS 512
#define TK_WORST 0xFFFFFFFFFFFFFFFFull

__device__ __forceinline__ uint32_t tk_ord(float f)          // monotone float -> uint32 (ascending)
{
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float tk_ord_inv(uint32_t k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void tk_init_kernel(unsigned long long* __restrict__ cand, int64_t n)
{
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) cand[e] = TK_WORST;
}

// scores of one (rows x cols) chunk, row-major, -> slots [k, k + cols) of every row; the rest of the chunk area -> worst.
// norm -1: raw score, descending.  0: score / sqrt(len_db), descending (Serra09.normalize_by_length, rqa_serra09.py:71-83).
// 1: sqrt(len_db) / score, ascending (ChenFusion.normalize_by_length, latefusion_chen.py:75-85).  float64 arithmetic, float32 store,
// like the reference's memmap update.  A query that is its own database entry is left out (the diagonal is never ranked, :233).
__global__ void tk_scatter_kernel(const float* __restrict__ scores, int rows, int cols, int k, int norm,
                                  const int32_t* __restrict__ q_tracks, const int32_t* __restrict__ db_tracks, int db_pos0,
                                  const int64_t* __restrict__ offs, unsigned long long* __restrict__ cand)
{
    const int area = TK_SLOTS - k;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < (int64_t)rows * area; e += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(e / area), c = (int)(e - (int64_t)r * area);
        unsigned long long key = TK_WORST;
        if (c < cols) {
            const int t = db_tracks[c];
            if (t != q_tracks[r]) {
                const float s = scores[(int64_t)r * cols + c];
                float v = s;
                if (norm >= 0) {
                    const double len = sqrt((double)(offs[t + 1] - offs[t]));
                    v = (float)(norm == 0 ? (double)s / len : len / (double)s);
                }
                const uint32_t o = (norm == 1) ? tk_ord(v) : ~tk_ord(v);
                key = ((unsigned long long)o << 32) | (uint32_t)(db_pos0 + c);
            }
        }
        cand[(int64_t)r * TK_SLOTS + k + c] = key;
    }
}

// one block per row: bitonic sort of the TK_SLOTS keys in shared memory, the best k written back to slots [0, k)
__global__ void __launch_bounds__(TK_THREADS) tk_sort_kernel(unsigned long long* __restrict__ cand, int k)
{
    __shared__ unsigned long long s[TK_SLOTS];
    unsigned long long* row = cand + (int64_t)blockIdx.x * TK_SLOTS;
    for (int i = threadIdx.x; i < TK_SLOTS; i += TK_THREADS) s[i] = row[i];
    __syncthreads();
    for (int size = 2; size <= TK_SLOTS; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = threadIdx.x; t < TK_SLOTS / 2; t += TK_THREADS) {
                const int i = 2 * t - (t & (stride - 1));           // lower index of the compare-exchange
                const int j = i + stride;
                const bool up = (i & size) == 0;
                const unsigned long long a = s[i], b = s[j];
                if ((a > b) == up) { s[i] = b; s[j] = a; }
            }
            __syncthreads();
        }
    for (int i = threadIdx.x; i < k; i += TK_THREADS) row[i] = s[i];
}

// keys -> (database position | -1, normalised score | -inf / +inf) per query
__global__ void tk_emit_kernel(const unsigned long long* __restrict__ cand, int rows, int k, int norm,
                               int32_t* __restrict__ out_idx, float* __restrict__ out_score)
{
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < (int64_t)rows * k; e += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(e / k), c = (int)(e - (int64_t)r * k);
        const unsigned long long key = cand[(int64_t)r * TK_SLOTS + c];
        if (key == TK_WORST) { out_idx[e] = -1; out_score[e] = (norm == 1) ? INFINITY : -INFINITY; continue; }
        const uint32_t o = (uint32_t)(key >> 32);
        out_idx[e] = (int32_t)(uint32_t)key;
        out_score[e] = tk_ord_inv(norm == 1 ? o : ~o);
    }
}
