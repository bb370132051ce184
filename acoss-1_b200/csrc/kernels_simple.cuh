// kernels_simple.cuh -- SiMPle matrix-profile similarity and FTM2D shingle distance (sm_100a).
//
// SiMPle (acoss/algorithms/simple_silva.py): Simple.oti (:45-54), Simple.simple_sim (:68-118),
// Simple.similarity (:120-126).  For the ordered pair (i, j):
//     B' = roll(B_j, oti, axis=0),  oti = argsort(<sum_t A_i, roll(sum_t B_j, k)>)[-1]
//     MP[p] = min_q  sum_{k < SSLEN} sum_{c < 12} (A[c][p+k] - B'[c][q+k])^2,   p <= La - SSLEN, q <= Lb - SSLEN
//     score = -median(MP)                                   (float64 throughout, like the reference)
// The reference evaluates the same quantity through FFT-seeded sliding dot products (STOMP); here
// each diagonal of the (La x Lb) frame-distance matrix is walked once with a running SSLEN-window sum,
// i.e. 12 subtract + 12 FMA per cell and no dependence on the reference's rounding path.
// One CTA per ordered pair; both sequences and the profile live in shared memory.  Compute-bound on
// the FP64 pipe (~26 flop per cell, 8*12*(La+Lb) bytes in, 8 bytes out).
//
// FTM2D (acoss/algorithms/ftm2d.py:85-97): sim(i, j) = exp(-sum((s_i - s_j)^2)) on 900-d float64
// shingles: one tiled all-pairs kernel (4*N_t^2 output bytes; 2*900 flop per pair).
#pragma once
#include "common.cuh"

#define SIMPLE_THREADS 256
#define SIMPLE_SEG 49            // profile rows per work item (plus SSLEN-1 warm-up cells); ODD, see SIMPLE_FS
#define SIMPLE_FS 14             // doubles per frame in shared memory (12 bins + 2 pad).  The lanes of a warp walk segments whose
                                 // first rows are SIMPLE_SEG apart: with 12-double frames and 48-row segments every lane hit the
                                 // same bank (48 x 24 words = 0 mod 32: ncu showed the shared-memory pipe 97 % busy on conflicts);
                                 // 49 x 28 words = 28 mod 32 puts the eight lanes of a 16-byte-load phase on eight different bank quads

// feats: frame-major (L_t, 12) float64 per track; prof: (n_tracks, 12) sums over time
__global__ void simple_prepare_kernel(const double* __restrict__ rowmajor /* (12, L) per track */,
                                      const int64_t* __restrict__ offs, int n_tracks,
                                      double* __restrict__ framemajor, double* __restrict__ prof)
{
    const int t = blockIdx.x;
    if (t >= n_tracks) return;
    const int64_t o = offs[t];
    const int L = (int)(offs[t + 1] - o);
    const double* src = rowmajor + o * ACOSS_NBINS;
    double* dst = framemajor + o * ACOSS_NBINS;
    for (int e = threadIdx.x; e < L * ACOSS_NBINS; e += blockDim.x) {
        const int p = e / ACOSS_NBINS, c = e - p * ACOSS_NBINS;
        dst[e] = src[(int64_t)c * L + p];
    }
    if (threadIdx.x < ACOSS_NBINS) {
        double s = 0.0;
        for (int p = 0; p < L; ++p) s += src[(int64_t)threadIdx.x * L + p];
        prof[t * ACOSS_NBINS + threadIdx.x] = s;
    }
}

template <int SSLEN>
__device__ __forceinline__ void simple_walk(const double* __restrict__ As, const double* __restrict__ Bs,
                                            int p0, int q0, int n_out, unsigned long long* mp)
{
    // cells (p0 + t, q0 + t), t = 0 .. n_out + SSLEN - 2; output row p0 + t - (SSLEN - 1) for t >= SSLEN - 1
    double ring[SSLEN];
#pragma unroll
    for (int k = 0; k < SSLEN; ++k) ring[k] = 0.0;
    double s = 0.0;
    const int total = n_out + SSLEN - 1;
    for (int base = 0; base < total; base += SSLEN) {
#pragma unroll
        for (int k = 0; k < SSLEN; ++k) {
            const int t = base + k;
            if (t < total) {
                const double* a = As + (p0 + t) * SIMPLE_FS;
                const double* b = Bs + (q0 + t) * SIMPLE_FS;
                double g = 0.0;
#pragma unroll
                for (int c = 0; c < ACOSS_NBINS; ++c) { const double d = a[c] - b[c]; g = fma(d, d, g); }
                s += g - ring[k];
                ring[k] = g;
                if (t >= SSLEN - 1) {
                    // re-sum the window exactly every SSLEN steps is unnecessary in fp64 (drift ~1e-16 per step)
                    const double v = s < 0.0 ? 0.0 : s;
                    const int row = p0 + t - (SSLEN - 1);
                    const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
                    if (bits < mp[row]) atomicMin(&mp[row], bits);
                }
            }
        }
    }
}

// generic window length: the outgoing cell is recomputed instead of kept in a register ring
__device__ __forceinline__ double simple_cell(const double* a, const double* b)
{
    double g = 0.0;
#pragma unroll
    for (int c = 0; c < ACOSS_NBINS; ++c) { const double d = a[c] - b[c]; g = fma(d, d, g); }
    return g;
}

template <int SSLEN_T>
__global__ void __launch_bounds__(SIMPLE_THREADS)
simple_pair_kernel(const int32_t* __restrict__ pairs, int64_t n_pairs, const double* __restrict__ feats,
                   const int64_t* __restrict__ offs, const double* __restrict__ prof, int sslen, int maxL,
                   double* __restrict__ out)
{
    extern __shared__ __align__(16) unsigned char simple_smem[];
    double* As = reinterpret_cast<double*>(simple_smem);
    double* Bs = As + (size_t)maxL * SIMPLE_FS;
    unsigned long long* mp = reinterpret_cast<unsigned long long*>(Bs + (size_t)maxL * SIMPLE_FS);
    int* segpre = reinterpret_cast<int*>(mp + maxL);            // [2 maxL + 2] segments in the diagonals before d
    __shared__ double otiv[ACOSS_NBINS];
    __shared__ int s_oti;
    __shared__ double s_med[2];
    const int tid = threadIdx.x;
    for (int64_t pi = blockIdx.x; pi < n_pairs; pi += gridDim.x) {
        const int ta = pairs[2 * pi], tb = pairs[2 * pi + 1];
        const int64_t oa = offs[ta], ob = offs[tb];
        const int La = (int)(offs[ta + 1] - oa), Lb = (int)(offs[tb + 1] - ob);
        __syncthreads();
        if (tid < ACOSS_NBINS) {
            double v = 0.0;                    // np.dot(profile_a, np.roll(profile_b, tid))
            for (int c = 0; c < ACOSS_NBINS; ++c) v += prof[ta * ACOSS_NBINS + c] * prof[tb * ACOSS_NBINS + rot_src(c, tid)];
            otiv[tid] = v;
        }
        __syncthreads();
        if (tid == 0) {
            int best = 0;                      // argsort(...)[-1]: the last (largest index) among equal maxima
            for (int k = 1; k < ACOSS_NBINS; ++k) if (otiv[k] >= otiv[best]) best = k;
            s_oti = best;
        }
        __syncthreads();
        const int rot = s_oti;
        for (int e = tid; e < La * ACOSS_NBINS; e += SIMPLE_THREADS) {
            const int p = e / ACOSS_NBINS, c = e - p * ACOSS_NBINS;
            As[p * SIMPLE_FS + c] = feats[oa * ACOSS_NBINS + e];
        }
        for (int e = tid; e < Lb * ACOSS_NBINS; e += SIMPLE_THREADS) {
            const int q = e / ACOSS_NBINS, c = e - q * ACOSS_NBINS;
            Bs[q * SIMPLE_FS + c] = feats[(ob + q) * ACOSS_NBINS + rot_src(c, rot)];
        }
        const int n_mp = La - sslen + 1, n_q = Lb - sslen + 1;
        for (int e = tid; e < n_mp; e += SIMPLE_THREADS) mp[e] = 0x7FF0000000000000ull;     // +inf
        __syncthreads();
        // work items: (diagonal delta = q - p, segment of SIMPLE_SEG profile rows).  The item -> (diagonal, segment) map goes
        // through a prefix table of segments per diagonal, so that thread t takes items t, t + 256, ... and ALL lanes of a
        // warp walk their segments at the same time.  (Dealing the items out inside the enumeration loop left ONE lane of a
        // warp active at a time: ncu showed ~30x the warp-level instructions the cells need.)
        const int nd = n_mp + n_q - 1;                         // diagonals, index d = delta + n_mp - 1
        for (int d = tid; d < nd; d += SIMPLE_THREADS) {
            const int delta = d - (n_mp - 1);
            const int p_lo = delta < 0 ? -delta : 0, p_hi = min(n_mp - 1, n_q - 1 - delta);
            segpre[d + 1] = (p_hi - p_lo + SIMPLE_SEG) / SIMPLE_SEG;
        }
        if (tid == 0) segpre[0] = 0;
        __syncthreads();
        if (tid < 32) {                                        // inclusive scan of segpre[1 .. nd] by warp 0
            const int per = (nd + 31) / 32;
            const int lo = 1 + tid * per, hi = min(nd + 1, lo + per);
            int sum = 0;
            for (int i = lo; i < hi; ++i) sum += segpre[i];
            int incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t_ = __shfl_up_sync(0xffffffffu, incl, o); if (tid >= o) incl += t_; }
            int run = incl - sum;
            for (int i = lo; i < hi; ++i) { run += segpre[i]; segpre[i] = run; }
        }
        __syncthreads();
        const int n_items = segpre[nd];
        for (int item = tid; item < n_items; item += SIMPLE_THREADS) {
            int lo = 0, hi = nd;                               // segpre[lo] <= item < segpre[hi]
            while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (segpre[mid] <= item) lo = mid; else hi = mid; }
            const int delta = lo - (n_mp - 1);
            const int p_lo = delta < 0 ? -delta : 0;                  // first profile row on this diagonal
            const int p_hi = min(n_mp - 1, n_q - 1 - delta);          // last
            const int n_rows = p_hi - p_lo + 1;
            const int r0 = (item - segpre[lo]) * SIMPLE_SEG;
            {
                const int n_out = min(SIMPLE_SEG, n_rows - r0);
                const int p0 = p_lo + r0, q0 = p0 + delta;
                if (SSLEN_T > 0) simple_walk<(SSLEN_T > 0 ? SSLEN_T : 1)>(As, Bs, p0, q0, n_out, mp);
                else {
                    double s = 0.0;
                    for (int t = 0; t < n_out + sslen - 1; ++t) {
                        s += simple_cell(As + (p0 + t) * SIMPLE_FS, Bs + (q0 + t) * SIMPLE_FS);
                        if (t >= sslen) s -= simple_cell(As + (p0 + t - sslen) * SIMPLE_FS, Bs + (q0 + t - sslen) * SIMPLE_FS);
                        if (t >= sslen - 1) {
                            const double v = s < 0.0 ? 0.0 : s;
                            const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
                            const int row = p0 + t - (sslen - 1);
                            if (bits < mp[row]) atomicMin(&mp[row], bits);
                        }
                    }
                }
            }
        }
        __syncthreads();
        // np.median(matrix_profile): rank by counting (non-negative doubles order like their bit patterns)
        const int k_lo = (n_mp - 1) / 2, k_hi = n_mp / 2;
        for (int e = tid; e < n_mp; e += SIMPLE_THREADS) {
            const unsigned long long v = mp[e];
            int rank = 0;
            for (int j = 0; j < n_mp; ++j) { const unsigned long long w = mp[j]; rank += (w < v) || (w == v && j < e); }
            if (rank == k_lo) s_med[0] = __longlong_as_double((long long)v);
            if (rank == k_hi) s_med[1] = __longlong_as_double((long long)v);
        }
        __syncthreads();
        if (tid == 0) out[pi] = -((k_lo == k_hi) ? s_med[0] : (s_med[0] + s_med[1]) / 2.0);
    }
}

// FTM2D: D[i][j] = exp(-sum_k (S[i][k] - S[j][k])^2); 64 x 64 tile per CTA, 4 x 4 per thread
#define FTM_TILE 64
#define FTM_KC 16
__global__ void __launch_bounds__(256)
ftm2d_allpairs_kernel(const double* __restrict__ S, int n, int d, double* __restrict__ D)
{
    __shared__ double sa[FTM_KC][FTM_TILE + 1];
    __shared__ double sb[FTM_KC][FTM_TILE + 1];
    const int i0 = blockIdx.y * FTM_TILE, j0 = blockIdx.x * FTM_TILE;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    double acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
    for (int k0 = 0; k0 < d; k0 += FTM_KC) {
        for (int e = threadIdx.x; e < FTM_TILE * FTM_KC; e += 256) {
            const int r = e / FTM_KC, k = e - r * FTM_KC;
            sa[k][r] = (i0 + r < n && k0 + k < d) ? S[(int64_t)(i0 + r) * d + k0 + k] : 0.0;
            sb[k][r] = (j0 + r < n && k0 + k < d) ? S[(int64_t)(j0 + r) * d + k0 + k] : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < FTM_KC; ++k) {
            double av[4], bv[4];
#pragma unroll
            for (int a = 0; a < 4; ++a) { av[a] = sa[k][ty * 4 + a]; bv[a] = sb[k][tx * 4 + a]; }
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) { const double df = av[a] - bv[b]; acc[a][b] = fma(df, df, acc[a][b]); }
        }
        __syncthreads();
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int i = i0 + ty * 4 + a, j = j0 + tx * 4 + b;
            if (i < n && j < n) D[(int64_t)i * n + j] = exp(-acc[a][b]);
        }
}

// ---- SiMPle per-track preprocessing (Simple.load_features + Simple.smooth, acoss/algorithms/simple_silva.py:34-43, 56-66) ----
// Block per track.  (1) column k of the reduced feature = mean over raw frames k SKIP .. k SKIP + WIN - 1 (clipped at the end
// of the track) of every chroma bin: numpy reduces the transposed float32 view sequentially over time in float32 and divides
// in float32 (checked against numpy in tests/); (2) 'same'-aligned convolution with the normalised Hann window (float64, taps
// computed on the host); (3) every column divided by its L2 norm over the 12 bins (summed bin-ascending), norms below the
// smallest normal double left alone.  `stage` holds the (12, L) float64 output of step (1).
__global__ void simple_preprocess_kernel(const float* __restrict__ raw, const int64_t* __restrict__ raw_offs,
                                         const int64_t* __restrict__ col_offs, int win, int skip,
                                         const double* __restrict__ taps, int n_taps,
                                         double* __restrict__ stage, double* __restrict__ out)
{
    const int t = blockIdx.x;
    const int64_t T = raw_offs[t + 1] - raw_offs[t];
    const int L = (int)(col_offs[t + 1] - col_offs[t]);
    const float* x = raw + raw_offs[t] * ACOSS_NBINS;
    double* st = stage + col_offs[t] * ACOSS_NBINS;
    double* o = out + col_offs[t] * ACOSS_NBINS;
    for (int e = threadIdx.x; e < L * ACOSS_NBINS; e += blockDim.x) {
        const int k = e / ACOSS_NBINS, c = e - k * ACOSS_NBINS;        // consecutive threads: consecutive bins of one frame
        const int64_t j0 = (int64_t)k * skip;
        const int cnt = (int)min((int64_t)win, T - j0);
        float acc = 0.0f;
        for (int j = 0; j < cnt; ++j) acc = __fadd_rn(acc, x[(j0 + j) * ACOSS_NBINS + c]);
        st[(int64_t)c * L + k] = (double)__fdiv_rn(acc, (float)cnt);
    }
    __syncthreads();
    const int start = (n_taps - 1) / 2;
    for (int e = threadIdx.x; e < L * ACOSS_NBINS; e += blockDim.x) {
        const int c = e / L, i = e - c * L;
        const int nfull = i + start;                                   // index into the 'full' convolution
        double acc = 0.0;
        for (int k = 0; k < n_taps; ++k) {
            const int j = nfull - k;
            if (j >= 0 && j < L) acc = __dadd_rn(acc, __dmul_rn(st[(int64_t)c * L + j], taps[k]));
        }
        o[(int64_t)c * L + i] = acc;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < L; i += blockDim.x) {
        double ss = 0.0;
        for (int c = 0; c < ACOSS_NBINS; ++c) { const double v = fabs(o[(int64_t)c * L + i]); ss = __dadd_rn(ss, __dmul_rn(v, v)); }
        double len = sqrt(ss);
        if (len < 2.2250738585072014e-308) len = 1.0;
        for (int c = 0; c < ACOSS_NBINS; ++c) o[(int64_t)c * L + i] = o[(int64_t)c * L + i] / len;
    }
}
