// kernels_eval.cuh -- N x N score-matrix stage: scatter/mirror, length normalisation,
// rank counting for MR/MRR/MDR/MAP/Top-k (sm_100a).
//
// Reference: CoverAlgorithm.all_pairwise (acoss/algorithms/algorithm_template.py:168-190),
// Serra09.normalize_by_length (rqa_serra09.py:71-83), ChenFusion.normalize_by_length
// (latefusion_chen.py:75-85), CoverAlgorithm.getEvalStatistics (algorithm_template.py:206-291).
#pragma once
#include "common.cuh"

// D[i, j] = score (and D[j, i] = score when symmetric: "Ds += Ds.T" on a matrix
// whose lower triangle and diagonal are zero is exactly this assignment).
__global__ void scatter_scores_kernel(const int32_t* __restrict__ pairs,
                                      const float* __restrict__ scores, int64_t n_pairs,
                                      int n, int symmetric, float* __restrict__ D)
{
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    int i = pairs[2 * p], j = pairs[2 * p + 1];
    float s = scores[p];
    D[(int64_t)i * n + j] = s;
    if (symmetric) D[(int64_t)j * n + i] = s;
}

// mode 0: D[i,j] = f32( f64(D[i,j]) / sqrt(f64(len_j)) )     (Serra09)
// mode 1: D[i,j] = f32( sqrt(f64(len_j)) / f64(D[i,j]) )     (ChenFusion; 0 -> inf)
__global__ void normalize_by_length_kernel(float* __restrict__ D, int n,
                                           const int32_t* __restrict__ lengths, int mode)
{
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)n * n) return;
    int j = (int)(idx % n);
    double nf = sqrt((double)lengths[j]);
    double v = (double)D[idx];
    D[idx] = (float)(mode == 0 ? v / nf : nf / v);
}

// One CTA per position p of the clique-contiguous order.  For every clique mate c
// (positions cstart .. cstart+K-1, the song itself included with value -inf) the
// rank of c in row p's descending, position-stable ordering:
//   rank_c = 1 + #{x : val(x) > val(c)  or  (val(x) == val(c) and x < c)}
// with val(x) = -inf for x == p else D[order[p], order[x]].  The K ranks are
// written unsorted to mate_ranks[rank_off[p] + (c - cstart)]; cliques of size < 2
// are skipped (the reference stops at the first one; they come last).
__global__ void __launch_bounds__(256)
eval_rank_kernel(const float* __restrict__ D, int n,
                 const int32_t* __restrict__ order, const int32_t* __restrict__ csize,
                 const int32_t* __restrict__ cstart, const int64_t* __restrict__ rank_off,
                 int32_t* __restrict__ mate_ranks)
{
    extern __shared__ float row[];     // n values of row p in position order
    const int p = blockIdx.x;
    const int K = csize[p];
    if (K < 2) return;
    const int c0 = cstart[p];
    const float* Drow = D + (int64_t)order[p] * n;
    for (int x = threadIdx.x; x < n; x += blockDim.x)
        row[x] = (x == p) ? -INFINITY : Drow[order[x]];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int k = warp; k < K; k += nw) {
        const int c = c0 + k;
        const float vc = row[c];
        int cnt = 0;
        for (int x = lane; x < n; x += 32) {
            float vx = row[x];
            cnt += (vx > vc) || (vx == vc && x < c);
        }
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if (lane == 0) mate_ranks[rank_off[p] + k] = cnt + 1;
    }
}
