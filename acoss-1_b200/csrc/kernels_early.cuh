// kernels_early.cuh -- EarlyFusionTralie per-pair pipeline (sm_100a).
//
// Replaces EarlyFusion.similarity (acoss/algorithms/earlyfusion_traile.py:157-198) and the free
// functions it calls: get_csm (cross_recurrence.py:30-48), get_csm_cosine (:53-73), get_oti (:75-103),
// get_csm_blocked_oti (:105-134), csm_to_binary (:136-161), getWCSM (similarity_fusion.py:38-54) and
// smith_waterman_constrained (alignment_tools.py:25-46; kernel in kernels_dp.cuh).
//
// Per pair (M, N = blocks of the two tracks), all float32 like the reference's numpy path:
//   CSM_mfccs, CSM_ssms = sqrt(max(xx_i + yy_j - 2 <x_i, y_j>, 0))        (GEMM, d = 50*n_mfcc and 1225)
//   CSM_chromas         = 1 - <roll(x_i, oti) / |x_i|, y_j / |y_j|>        (GEMM, d = 480)
//   B_s   = csm_to_binary(CSM_s, kappa): the round(kappa*N) smallest of every row -> 1
//   W_s   = exp(-C^2 / (2 (0.5 eps)^2)),  eps = (rowmean_K(i) + colmean_K(j) + C) / 3
//   early = exp(-(W_mfccs + W_ssms + W_chromas));  B_early = csm_to_binary(early, kappa)
//   scores = smith_waterman_constrained(B_*)                                 (4 per pair)
// Layout: a batch of pairs owns, per pair, four float32 M x N matrices (3 CSMs + early) and four uint8
// M x N binary matrices in HBM scratch; tracks' block features are uploaded once.
#pragma once
#include "common.cuh"

struct EfPair {
    int32_t ta, tb;          // track indices (query i, reference j)
    int32_t M, N;            // number of blocks
    int32_t oti;             // get_oti(chroma_med_i, chroma_med_j), device-written
    int32_t k_bin;           // int(np.round(kappa * N))
    int32_t pitch;           // row pitch (bytes) of the binary matrices: N rounded up to the Smith-Waterman chunk, padding = 0
    int32_t pad_;
    int64_t blk_a, blk_b;    // first block row of the two tracks in the block store
    int64_t csm_off;         // float offset of 4 consecutive M x N matrices (mfccs, ssms, chromas, early)
    int64_t bin_off;         // byte offset of 4 consecutive M x N binary matrices
    int64_t row_off;         // float offset of rowmean[3][M]
    int64_t col_off;         // float offset of colmean[3][N]
    int64_t out_idx;
};

struct EfTile { int32_t pair, feat, m0, n0; };

#define EF_TM 64
#define EF_TN 64
#define EF_TK 16

// per-row squared norms of one feature: np.sum(X**2, 1) in float32
__global__ void ef_sqnorm_kernel(const float* __restrict__ X, int64_t rows, int d, float* __restrict__ out)
{
    const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (r >= rows) return;
    float s = 0.f;
    for (int c = lane; c < d; c += 32) { const float v = X[r * d + c]; s = fmaf(v, v, s); }
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) out[r] = s;
}

// get_oti(C1, C2): argmax_i sum(roll(C1, i) * C2), first maximum
__global__ void ef_oti_kernel(EfPair* __restrict__ pairs, int64_t n, const float* __restrict__ cmed)
{
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const float* c1 = cmed + (int64_t)pairs[p].ta * ACOSS_NBINS;
    const float* c2 = cmed + (int64_t)pairs[p].tb * ACOSS_NBINS;
    int best = 0;
    float bv = -INFINITY;
    for (int i = 0; i < ACOSS_NBINS; ++i) {
        float s = 0.f;
        for (int c = 0; c < ACOSS_NBINS; ++c) s = __fadd_rn(s, __fmul_rn(c1[rot_src(c, i)], c2[c]));
        if (s > bv) { bv = s; best = i; }
    }
    pairs[p].oti = best;
}

// One 64 x 64 tile of one CSM of one pair.  256 threads, 4 x 4 outputs each, K staged 16 at a time.
// feat 0/1: Euclidean; feat 2: cosine on OTI-rolled, row-normalised chroma blocks.
__global__ void __launch_bounds__(256)
ef_csm_kernel(const EfPair* __restrict__ pairs, const EfTile* __restrict__ tiles, int n_tiles,
              const float* __restrict__ X0, const float* __restrict__ X1, const float* __restrict__ X2,
              const float* __restrict__ sq0, const float* __restrict__ sq1, const float* __restrict__ sq2,
              int d0, int d1, int d2, float* __restrict__ csm)
{
    __shared__ __align__(16) float As[EF_TK][EF_TM + 4];
    __shared__ __align__(16) float Bs[EF_TK][EF_TN + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const EfTile tl = tiles[t];
        const EfPair pd = pairs[tl.pair];
        const float* X = tl.feat == 0 ? X0 : (tl.feat == 1 ? X1 : X2);
        const float* sq = tl.feat == 0 ? sq0 : (tl.feat == 1 ? sq1 : sq2);
        const int d = tl.feat == 0 ? d0 : (tl.feat == 1 ? d1 : d2);
        const bool cosine = (tl.feat == 2);
        const int rot = cosine ? pd.oti : 0;
        // this thread stages row (tid >> 2) of A and of B, 4 consecutive k at (tid & 3) * 4
        const int lr = tid >> 2, lk = (tid & 3) * 4;
        const int ra = tl.m0 + lr, rb = tl.n0 + lr;
        const bool va = ra < pd.M, vb = rb < pd.N;
        const float* pa = X + (pd.blk_a + (va ? ra : 0)) * d;
        const float* pb = X + (pd.blk_b + (vb ? rb : 0)) * d;
        float ia = 1.f, ib = 1.f;
        if (cosine) {
            float na = va ? sqrtf(sq[pd.blk_a + ra]) : 1.f, nb = vb ? sqrtf(sq[pd.blk_b + rb]) : 1.f;
            if (na == 0.f) na = 1.f;
            if (nb == 0.f) nb = 1.f;
            ia = na; ib = nb;
        }
        float acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
        for (int k0 = 0; k0 < d; k0 += EF_TK) {
            __syncthreads();
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int k = k0 + lk + q;
                float av = 0.f, bv = 0.f;
                if (k < d) {
                    if (va) {
                        int ks = k;
                        if (cosine) { const int g = k / ACOSS_NBINS, c = k - g * ACOSS_NBINS; ks = g * ACOSS_NBINS + rot_src(c, rot); }
                        av = pa[ks];
                        if (cosine) av = av / ia;
                    }
                    if (vb) { bv = pb[k]; if (cosine) bv = bv / ib; }
                }
                As[lk + q][lr] = av;
                Bs[lk + q][lr] = bv;
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < EF_TK; ++k) {
                const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
                const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
                const float av[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
            }
        }
        float* out = csm + pd.csm_off + (int64_t)tl.feat * pd.M * pd.N;
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int i = tl.m0 + ty * 4 + a;
            if (i >= pd.M) continue;
            const float xx = cosine ? 0.f : sq[pd.blk_a + i];
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int j = tl.n0 + tx * 4 + b;
                if (j >= pd.N) continue;
                float v;
                if (cosine) v = 1.f - acc[a][b];
                else {
                    v = (xx + sq[pd.blk_b + j]) - 2.f * acc[a][b];
                    if (v < 0.f) v = 0.f;
                    v = sqrtf(v);
                }
                out[(int64_t)i * pd.N + j] = v;
            }
        }
    }
}

__device__ __forceinline__ uint32_t ef_key(float f)
{
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ef_key_inv(uint32_t k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// warp-cooperative: key of the k-th smallest (0-based) of n keys held as keys[q] = element q*32 + lane
template <int R>
__device__ __forceinline__ uint32_t ef_warp_select(const uint32_t (&keys)[R], int k)
{
    uint32_t ans = 0;
    for (int bit = 31; bit >= 0; --bit) {
        const uint32_t cand = ans | (1u << bit);
        int cnt = 0;
#pragma unroll
        for (int q = 0; q < R; ++q) cnt += (keys[q] < cand);
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if (cnt <= k) ans = cand;
    }
    return ans;
}

// R-loops over a row's keys (element q * 32 + lane): blocks of four q behind ONE warp-uniform branch, so that a row shorter than
// 32 R pays for the blocks it has (a per-element guard is if-converted and the padding then costs as much as the data).
// Needs `N` (the row length) in scope.
#define EF_Q(q, ...)                                                             \
    _Pragma("unroll") for (int qb_ = 0; qb_ < R; qb_ += 4)                         \
        if (qb_ * 32 < N) { _Pragma("unroll") for (int q = qb_; q < qb_ + 4 && q < R; ++q) { __VA_ARGS__ } }

// Two order statistics of one row at once (ranks kA and, if wantB, kB; 0-based) without the 32-step bitwise descent (2 x 32 x R
// compare-and-count instructions per row were 49 % of an EarlyFusion step): one 256-bin histogram over the row's key range in
// shared memory (bins are linear in KEY space: (key - min) >> shift, exact and monotone), a scan that locates the bin of each
// rank, and an exact ranking of that bin's members (usually 1 - 4 of them).  A bin with more than 32 members (long runs of equal
// distances) falls back to the bitwise descent.  hist: 256 ints, cand: 64 keys, ccnt: 2 ints, all private to the warp.
template <int R>
__device__ __forceinline__ void ef_warp_select2(const uint32_t (&keys)[R], int N, int kA, int kB, bool wantB,
                                                int* hist, uint32_t* cand, int* ccnt, uint32_t& outA, uint32_t& outB,
                                                int (&less)[2], int (&equal)[2])
{
    // less[w] = number of keys below the w-th result, equal[w] = number of keys equal to it (-1: unknown, the caller counts)
    const int lane = threadIdx.x & 31;
    uint32_t mn = 0xffffffffu, mx = 0u;
    EF_Q(q,
        mn = min(mn, keys[q]);                                   // padding keys are 0xffffffff
        if (q * 32 + lane < N) mx = max(mx, keys[q]);)
    mn = __reduce_min_sync(0xffffffffu, mn);
    mx = __reduce_max_sync(0xffffffffu, mx);
    const uint32_t range = mx - mn;
    const int shift = range >= 256u ? (32 - __clz(range)) - 8 : 0;       // (range >> shift) < 256
    int4* h4 = reinterpret_cast<int4*>(hist);
    h4[lane] = make_int4(0, 0, 0, 0); h4[lane + 32] = make_int4(0, 0, 0, 0);
    if (lane < 2) ccnt[lane] = 0;
    __syncwarp();
    EF_Q(q, if (q * 32 + lane < N) atomicAdd(&hist[(keys[q] - mn) >> shift], 1);)
    __syncwarp();
    const int4 ca = h4[2 * lane], cb = h4[2 * lane + 1];             // bins 8 lane .. 8 lane + 7
    const int c[8] = {ca.x, ca.y, ca.z, ca.w, cb.x, cb.y, cb.z, cb.w};
    int tot = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) tot += c[j];
    int incl = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t_ = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t_; }
    const int excl = incl - tot;
    int bt[2] = {0, 0}, below[2] = {0, 0};
    int cum[8];                                   // inclusive prefix of this lane's eight bins, from the start of the row
    {
        int run = excl;
#pragma unroll
        for (int j = 0; j < 8; ++j) { run += c[j]; cum[j] = run; }
    }
#pragma unroll
    for (int w = 0; w < 2; ++w) {
        const int k = w == 0 ? kA : kB;
        const bool mine = k >= excl && k < incl;
        int nb_ = 0, bl = excl;                   // bins of this lane that end at or before rank k, and how many keys they hold
#pragma unroll
        for (int j = 0; j < 7; ++j) if (k >= cum[j]) { ++nb_; bl = cum[j]; }
        const unsigned m = __ballot_sync(0xffffffffu, mine);
        const int src = m ? __ffs(m) - 1 : 0;
        bt[w] = __shfl_sync(0xffffffffu, 8 * lane + nb_, src);
        below[w] = __shfl_sync(0xffffffffu, bl, src);
    }
    EF_Q(q,
        if (q * 32 + lane < N) {
            const int b = (int)((keys[q] - mn) >> shift);
            if (b == bt[0]) { const int p = atomicAdd(&ccnt[0], 1); if (p < 32) cand[p] = keys[q]; }
            if (wantB && b == bt[1]) { const int p = atomicAdd(&ccnt[1], 1); if (p < 32) cand[32 + p] = keys[q]; }
        })
    __syncwarp();
#pragma unroll
    for (int w = 0; w < 2; ++w) {
        if (w == 1 && !wantB) { outB = 0; less[1] = -1; equal[1] = -1; break; }
        const int n_c = ccnt[w], want = (w == 0 ? kA : kB) - below[w];
        uint32_t res;
        if (n_c <= 32) {
            const uint32_t mykey = lane < n_c ? cand[32 * w + lane] : 0xffffffffu;
            int rank = 0;
            for (int j = 0; j < n_c; ++j) {
                const uint32_t o = __shfl_sync(0xffffffffu, mykey, j);
                rank += (o < mykey) || (o == mykey && j < lane);
            }
            const unsigned m = __ballot_sync(0xffffffffu, lane < n_c && rank == want);
            res = __shfl_sync(0xffffffffu, mykey, m ? __ffs(m) - 1 : 0);
            less[w] = below[w] + __popc(__ballot_sync(0xffffffffu, lane < n_c && mykey < res));
            equal[w] = __popc(__ballot_sync(0xffffffffu, lane < n_c && mykey == res));
        } else {
            res = ef_warp_select<R>(keys, w == 0 ? kA : kB);
            less[w] = -1; equal[w] = -1;
        }
        if (w == 0) outA = res; else outB = res;
    }
    __syncwarp();
}

// One warp per (pair, matrix, row): binarise the k_bin smallest (ties -> lowest column) and, for the three
// CSMs (want_mean), the mean of the K smallest.  Rows up to 32*R columns.  which = matrix index 0..3.
template <int R>
__global__ void __launch_bounds__(256, R >= 64 ? 2 : 3)
ef_row_kernel(const EfPair* __restrict__ pairs, const int64_t* __restrict__ row_task_off, int n_pairs, int64_t n_rows_total,
              int first_mat, int n_mats, int K, const float* __restrict__ csm, uint8_t* __restrict__ bin,
              float* __restrict__ rowmean)
{
    __shared__ __align__(16) int s_hist[8 * 256];
    __shared__ uint32_t s_cand[8 * 64];
    __shared__ int s_ccnt[8 * 2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    (void)row_task_off; (void)n_rows_total;
    // block = (pair, matrix) [x] x row slice [y]: no search for the pair that owns a row (a 12-step binary search through global
    // memory per row was 7 % of the kernel's instructions and a chain of dependent loads)
    for (int64_t bp = blockIdx.x; bp < (int64_t)n_pairs * n_mats; bp += gridDim.x) {
      const int mat = first_mat + (int)(bp % n_mats);
      const EfPair pd = pairs[bp / n_mats];
      for (int i = blockIdx.y * 8 + warp; i < pd.M; i += 8 * gridDim.y) {
        const int N = pd.N;
        const float* row = csm + pd.csm_off + ((int64_t)mat * pd.M + i) * N;
        uint32_t keys[R];
#pragma unroll
        for (int q = 0; q < R; ++q) keys[q] = 0xffffffffu;
        EF_Q(q, const int j = q * 32 + lane; if (j < N) keys[q] = ef_key(row[j]);)
        uint8_t* brow = bin + pd.bin_off + ((int64_t)mat * pd.M + i) * pd.pitch;
        for (int j = N + lane; j < pd.pitch; j += 32) brow[j] = 0;
        // both order statistics of the row in one go: the kb-th smallest (binarisation) and the K-th smallest (neighbourhood mean)
        const bool want_bin = pd.k_bin > 0 && pd.k_bin < N;
        const bool want_mean = mat < 3 && rowmean != nullptr;
        const int Ke = K < N ? K : N;
        uint32_t selA = 0, selB = 0;
        int sel_less[2] = {-1, -1}, sel_eq[2] = {-1, -1};       // [0] binarisation rank, [1] mean rank
        {
            int* wh = s_hist + (threadIdx.x >> 5) * 256;
            uint32_t* wc = s_cand + (threadIdx.x >> 5) * 64;
            int* wn = s_ccnt + (threadIdx.x >> 5) * 2;
            if (want_bin) ef_warp_select2<R>(keys, N, pd.k_bin - 1, Ke - 1, want_mean, wh, wc, wn, selA, selB, sel_less, sel_eq);
            else if (want_mean) {
                int l2[2], e2[2];
                ef_warp_select2<R>(keys, N, Ke - 1, 0, false, wh, wc, wn, selB, selA, l2, e2);
                sel_less[1] = l2[0]; sel_eq[1] = e2[0];
            }
        }
        const int kb = pd.k_bin;
        if (kb <= 0) {
            EF_Q(q, const int j = q * 32 + lane; if (j < N) brow[j] = 0;)
        } else if (kb >= N) {
            EF_Q(q, const int j = q * 32 + lane; if (j < N) brow[j] = 1;)
        } else {
            const uint32_t kk = selA;                                     // kb-th smallest
            if (sel_eq[0] == 1) {
                // the kb-th smallest is unique (the rule for distances in general position): the kb smallest are the keys <= it
                EF_Q(q, const int j = q * 32 + lane; if (j < N) brow[j] = keys[q] <= kk ? 1 : 0;)
            } else {
                int less = 0;
                EF_Q(q, less += (keys[q] < kk);)
                less = __reduce_add_sync(0xffffffffu, less);
                int ties_left = kb - less;
                EF_Q(q,
                    const int j = q * 32 + lane;
                    const bool eq = (j < N) && keys[q] == kk;
                    const unsigned meq = __ballot_sync(0xffffffffu, eq);
                    const bool one = (j < N) && (keys[q] < kk || (eq && __popc(meq & ((1u << lane) - 1u)) < ties_left));
                    ties_left -= __popc(meq);
                    if (j < N) brow[j] = one ? 1 : 0;)
            }
        }
        if (mat < 3 && rowmean) {
            // np.mean(np.partition(C, K, 1)[:, 0:K], 1): mean of the K smallest (float32)
            const uint32_t kk = selB;
            float s = 0.f;
            int less = 0;
            EF_Q(q, if (keys[q] < kk) { s += ef_key_inv(keys[q]); ++less; })
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            less = __reduce_add_sync(0xffffffffu, less);
            s += (float)(Ke - less) * ef_key_inv(kk);
            if (lane == 0) rowmean[pd.row_off + (int64_t)mat * pd.M + i] = s / (float)Ke;
        }
      }
    }
}

// One thread per (pair, CSM, column): mean of the K (<= 16) smallest over rows, coalesced across columns.
__global__ void __launch_bounds__(128)
ef_col_kernel(const EfPair* __restrict__ pairs, const int64_t* __restrict__ col_task_off, int n_pairs, int64_t n_cols_total,
              int K, const float* __restrict__ csm, float* __restrict__ colmean)
{
    const int64_t task = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (task >= n_cols_total * 3) return;
    const int mat = (int)(task / n_cols_total);
    const int64_t cc = task - (int64_t)mat * n_cols_total;
    int lo = 0, hi = n_pairs - 1;
    while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (col_task_off[mid] <= cc) lo = mid; else hi = mid - 1; }
    const EfPair pd = pairs[lo];
    const int j = (int)(cc - col_task_off[lo]);
    const float* col = csm + pd.csm_off + (int64_t)mat * pd.M * pd.N + j;
    float top[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) top[q] = INFINITY;
    const int Ke = K < pd.M ? K : pd.M;
    for (int i = 0; i < pd.M; ++i) {
        float x = col[(int64_t)i * pd.N];
        if (x < top[15]) {
            // insert into the ascending list (static indices: bubble the new value down)
#pragma unroll
            for (int q = 0; q < 16; ++q) { const float lo_v = fminf(top[q], x); x = fmaxf(top[q], x); top[q] = lo_v; }
        }
    }
    // only the Ke smallest matter; entries beyond Ke are ignored
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < 16; ++q) if (q < Ke) s += top[q];
    colmean[pd.col_off + (int64_t)mat * pd.N + j] = s / (float)Ke;
}

// early = exp(-(W_0 + W_1 + W_2)), W_s = exp(-C^2 / (2 (0.5 eps)^2)), eps = (r_i + c_j + C) / 3
__global__ void __launch_bounds__(256)
ef_early_kernel(const EfPair* __restrict__ pairs, int n_pairs, int chunks_per_pair,
                float* __restrict__ csm, const float* __restrict__ rowmean, const float* __restrict__ colmean)
{
    const int p = blockIdx.x / chunks_per_pair, ch = blockIdx.x - p * chunks_per_pair;
    if (p >= n_pairs) return;
    const EfPair pd = pairs[p];
    const int64_t MN = (int64_t)pd.M * pd.N;
    for (int64_t c = (int64_t)ch * blockDim.x + threadIdx.x; c < MN; c += (int64_t)chunks_per_pair * blockDim.x) {
        const int i = (int)(c / pd.N), j = (int)(c - (int64_t)i * pd.N);
        float sum = 0.f;
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            const float C = csm[pd.csm_off + s * MN + c];
            float eps = (rowmean[pd.row_off + (int64_t)s * pd.M + i] + colmean[pd.col_off + (int64_t)s * pd.N + j]) + C;
            eps = eps / 3.0f;
            const float me = 0.5f * eps;
            const float den = 2.0f * (me * me);
            sum += expf(-(C * C) / den);
        }
        csm[pd.csm_off + 3 * MN + c] = expf(-sum);
    }
}
