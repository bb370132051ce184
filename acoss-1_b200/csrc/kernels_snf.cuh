// kernels_snf.cuh -- Similarity Network Fusion on N x N matrices (sm_100a).
//
// Replaces acoss/algorithms/utils/similarity_fusion.py: getW (:15-36), getP (:101-122, the
// diagRegularize=False branch that doSimilarityFusionWs uses), getS (:124-144),
// doSimilarityFusionWs (:146-186) and doSimilarityFusion (:188-196), as called by
// ChenFusion.do_late_fusion (latefusion_chen.py:87-91) and EarlyFusion.do_late_fusion
// (earlyfusion_traile.py:200-206).
//
// Numeric types follow numpy's promotion in the reference: the affinity stage (DSym, kNN radii,
// W, P, S) runs in the input dtype T (float32 matrices stay float32), the cross-diffusion runs in
// float64 (nextPts = np.zeros(...) is float64).  The iteration order reproduces the reference's
// list aliasing: iteration 0 is a Jacobi sweep from the P matrices, every later
// iteration is Gauss-Seidel (matrix i sees the already-updated matrices k < i).
//
// Bound: all kernels are streaming passes over N x N matrices (HBM / L2); the two K-sparse x dense
// products per update read K rows of the dense operand per output row (2*K*8*N^2 bytes, L2-resident
// at N_t = 1000).
#pragma once
#include "common.cuh"

#define SNF_THREADS 256

template <typename T> struct SnfKey;
template <> struct SnfKey<float> {
    typedef uint32_t type;
    static constexpr int BITS = 32;
    __device__ static uint32_t enc(float f) { uint32_t b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
    __device__ static float dec(uint32_t k) { return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k); }
};
template <> struct SnfKey<double> {
    typedef unsigned long long type;
    static constexpr int BITS = 64;
    __device__ static unsigned long long enc(double f)
    {
        unsigned long long b = (unsigned long long)__double_as_longlong(f);
        return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
    }
    __device__ static double dec(unsigned long long k)
    {
        return __longlong_as_double((long long)((k & 0x8000000000000000ull) ? (k & 0x7fffffffffffffffull) : ~k));
    }
};

__device__ __forceinline__ float snf_exp(float x) { return expf(x); }
__device__ __forceinline__ double snf_exp(double x) { return exp(x); }

// block-wide integer sum (all threads get the result); red = 33 ints of shared memory
__device__ __forceinline__ int block_sum_int(int v, int* red)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = __reduce_add_sync(0xffffffffu, v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    int t = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
    return t;
}

template <typename T>
__device__ __forceinline__ T block_sum(T v, T* red)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    T t = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
    return t;
}

// key of the k-th smallest (0-based) among keys[0..n) in shared memory: bitwise bisection
template <typename KT, int BITS>
__device__ KT block_select_key(const KT* keys, int n, int k, int* red)
{
    KT ans = 0;
    for (int bit = BITS - 1; bit >= 0; --bit) {
        const KT cand = ans | ((KT)1 << bit);
        int cnt = 0;
        for (int j = threadIdx.x; j < n; j += blockDim.x) cnt += (keys[j] < cand);
        cnt = block_sum_int(cnt, red);
        if (cnt <= k) ans = cand;      // at most k keys are < cand  =>  key_(k) >= cand
    }
    return ans;
}

// DSym = 0.5 * (D + D^T), diagonal 0   (similarity_fusion.py:23-24)
template <typename T>
__global__ void snf_sym_kernel(const T* __restrict__ D, int n, T* __restrict__ S)
{
    __shared__ T tile[32][33];
    const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {          // tile[r][c] = D[bx + r][by + c]  (the transposed block)
        const int i = bx + r, j = by + threadIdx.x;
        tile[r][threadIdx.x] = (i < n && j < n) ? D[(int64_t)i * n + j] : (T)0;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int i = by + r, j = bx + threadIdx.x;
        if (i < n && j < n) {
            const T a = D[(int64_t)i * n + j], b = tile[threadIdx.x][r];
            S[(int64_t)i * n + j] = (i == j) ? (T)0 : (T)0.5 * (a + b);
        }
    }
}

// MeanDist[i] = mean(K+1 smallest of row i) * (K+1) / K      (similarity_fusion.py:26-27)
template <typename T>
__global__ void __launch_bounds__(SNF_THREADS)
snf_radius_kernel(const T* __restrict__ S, int n, int K, T* __restrict__ mean_dist)
{
    typedef typename SnfKey<T>::type KT;
    extern __shared__ __align__(16) unsigned char snf_smem[];
    KT* keys = reinterpret_cast<KT*>(snf_smem);
    __shared__ int red[33];
    __shared__ T redT[32];
    const int i = blockIdx.x;
    const T* row = S + (int64_t)i * n;
    for (int j = threadIdx.x; j < n; j += blockDim.x) keys[j] = SnfKey<T>::enc(row[j]);
    __syncthreads();
    const KT kk = block_select_key<KT, SnfKey<T>::BITS>(keys, n, K, red);       // (K+1)-th smallest
    T s = 0;
    int less = 0;
    for (int j = threadIdx.x; j < n; j += blockDim.x)
        if (keys[j] < kk) { s += row[j]; ++less; }
    less = block_sum_int(less, red);
    s = block_sum<T>(s, redT);
    const T vshare = SnfKey<T>::dec(kk);          // the (K+1)-th smallest value itself (ties counted with it)
    if (threadIdx.x == 0) {
        const T total = s + (T)(K + 1 - less) * vshare;
        const T mean = total / (T)(K + 1);
        mean_dist[i] = mean * (T)(K + 1) / (T)K;
    }
}

// W = exp(-DSym^2 / Denom), Denom = 2 (Mu Eps)^2, Eps = (md_i + md_j + DSym) / 3, in place on S
// (similarity_fusion.py:31-36; Mu = 0.5)
template <typename T>
__global__ void snf_affinity_kernel(T* __restrict__ S, int n, const T* __restrict__ md)
{
    const int64_t tot = (int64_t)n * n;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < tot; e += (int64_t)gridDim.x * blockDim.x) {
        const int i = (int)(e / n), j = (int)(e - (int64_t)i * n);
        const T d = S[e];
        T eps = (md[i] + md[j]) + d;
        eps = eps / (T)3;
        const T me = (T)0.5 * eps;
        T den = (T)2 * (me * me);
        if (den == (T)0) den = (T)1;
        S[e] = snf_exp(-(d * d) / den);
    }
}

// P = W / rowsum (rowsum 0 -> 1), written as float64            (similarity_fusion.py:119-122)
template <typename T>
__global__ void __launch_bounds__(SNF_THREADS)
snf_rownorm_kernel(const T* __restrict__ W, int n, double* __restrict__ P)
{
    __shared__ T redT[32];
    const int i = blockIdx.x;
    const T* row = W + (int64_t)i * n;
    T s = 0;
    for (int j = threadIdx.x; j < n; j += blockDim.x) s += row[j];
    s = block_sum<T>(s, redT);
    if (s == (T)0) s = (T)1;
    for (int j = threadIdx.x; j < n; j += blockDim.x) P[(int64_t)i * n + j] = (double)(row[j] / s);
}

// S = K largest entries of each row of W, L1-normalised (similarity_fusion.py:133-144).
// Output in CSR order (ascending column, what coo.tocsr() canonicalises to): J[n][K], V[n][K] (float64).
// Ties at the K-th value are resolved towards the lowest column index.
template <typename T>
__global__ void __launch_bounds__(SNF_THREADS)
snf_topk_kernel(const T* __restrict__ W, int n, int K, int32_t* __restrict__ J, double* __restrict__ V)
{
    typedef typename SnfKey<T>::type KT;
    extern __shared__ __align__(16) unsigned char snf_smem[];
    KT* keys = reinterpret_cast<KT*>(snf_smem);
    __shared__ int red[33];
    const int i = blockIdx.x;
    const T* row = W + (int64_t)i * n;
    for (int j = threadIdx.x; j < n; j += blockDim.x) keys[j] = ~SnfKey<T>::enc(row[j]);    // descending order
    __syncthreads();
    const KT kk = block_select_key<KT, SnfKey<T>::BITS>(keys, n, K - 1, red);               // K-th largest
    int greater = 0;
    for (int j = threadIdx.x; j < n; j += blockDim.x) greater += (keys[j] < kk);
    greater = block_sum_int(greater, red);
    const int ties_wanted = K - greater;
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        int pos = 0, ties = 0;
        T sum = 0;
        for (int j0 = 0; j0 < n; j0 += 32) {
            const int j = j0 + lane;
            const KT key = (j < n) ? keys[j] : ~(KT)0;
            const bool gt = (j < n) && key < kk;
            const bool eq = (j < n) && key == kk;
            const unsigned meq = __ballot_sync(0xffffffffu, eq);
            const int my_tie = ties + __popc(meq & ((1u << lane) - 1u));
            const bool take = gt || (eq && my_tie < ties_wanted);
            const unsigned mt = __ballot_sync(0xffffffffu, take);
            if (take) {
                const int o = pos + __popc(mt & ((1u << lane) - 1u));
                J[(int64_t)i * K + o] = j;
            }
            pos += __popc(mt);
            ties += __popc(meq);
        }
        __syncwarp();
        // L1 norm over the K kept values in CSR order, then normalise
        for (int o = 0; o < K; ++o) sum += row[J[(int64_t)i * K + o]];
        if (sum == (T)0) sum = (T)1;
        for (int o = lane; o < K; o += 32) V[(int64_t)i * K + o] = (double)(row[J[(int64_t)i * K + o]] / sum);
    }
}

// dense image of the K-per-row matrix S (getS returns it sparse; the free function hands back scipy CSR built from this)
__global__ void snf_dense_S_kernel(const int32_t* __restrict__ J, const double* __restrict__ V, int n, int K, double* __restrict__ S)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (int64_t)n * K) return;
    const int i = (int)(e / K);
    S[(int64_t)i * n + J[e]] = V[e];
}

// Xt = transpose( (sum_{k != skip} P_k) / (n_mats - 1) )        (similarity_fusion.py:168-174)
// Ps: n_mats consecutive n x n float64 matrices.
__global__ void snf_mean_others_T_kernel(const double* __restrict__ Ps, int n_mats, int skip, int n, double* __restrict__ Xt)
{
    __shared__ double tile[32][33];
    const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
    const double inv_den = (double)(n_mats - 1);
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int i = by + r, j = bx + threadIdx.x;
        double acc = 0.0;
        if (i < n && j < n) {
            for (int k = 0; k < n_mats; ++k)
                if (k != skip) acc = __dadd_rn(acc, Ps[(int64_t)k * n * n + (int64_t)i * n + j]);
            acc = acc / inv_den;
        }
        tile[r][threadIdx.x] = acc;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int i = bx + r, j = by + threadIdx.x;          // Xt[i][j] = X[j][i]
        if (i < n && j < n) Xt[(int64_t)i * n + j] = tile[threadIdx.x][r];
    }
}

// out = S . in  (S: K entries per row, CSR order; sequential y += a * x like scipy's csr_matvecs),
// TRANSPOSE_OUT: store out^T instead.  add_diag is added to out[r][r].
template <bool TRANSPOSE_OUT>
__global__ void snf_spmm_kernel(const int32_t* __restrict__ J, const double* __restrict__ V, int K,
                                const double* __restrict__ in, int n, double add_diag, double* __restrict__ out)
{
    __shared__ double tile[32][33];
    const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
    const int c = c0 + threadIdx.x;
    for (int rr = threadIdx.y; rr < 32; rr += blockDim.y) {
        const int r = r0 + rr;
        double acc = 0.0;
        if (r < n && c < n) {
            for (int k = 0; k < K; ++k) {
                const int j = J[(int64_t)r * K + k];
                acc = __dadd_rn(acc, __dmul_rn(V[(int64_t)r * K + k], in[(int64_t)j * n + c]));
            }
            if (r == c) acc = __dadd_rn(acc, add_diag);
        }
        if (TRANSPOSE_OUT) tile[rr][threadIdx.x] = acc;
        else if (r < n && c < n) out[(int64_t)r * n + c] = acc;
    }
    if (TRANSPOSE_OUT) {
        __syncthreads();
        for (int rr = threadIdx.y; rr < 32; rr += blockDim.y) {
            const int i = c0 + rr, j = r0 + threadIdx.x;       // out^T[i][j] = out[j][i]
            if (i < n && j < n) out[(int64_t)i * n + j] = tile[threadIdx.x][rr];
        }
    }
}

// fused = (sum_k P_k) / n_mats                                   (similarity_fusion.py:183-186)
__global__ void snf_mean_kernel(const double* __restrict__ Ps, int n_mats, int64_t tot, double* __restrict__ out)
{
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < tot; e += (int64_t)gridDim.x * blockDim.x) {
        double acc = 0.0;
        for (int k = 0; k < n_mats; ++k) acc = __dadd_rn(acc, Ps[(int64_t)k * tot + e]);
        out[e] = acc / (double)n_mats;
    }
}
