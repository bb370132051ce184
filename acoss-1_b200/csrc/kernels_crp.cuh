// kernels_crp.cuh -- binary cross-recurrence plot of a batch of pairs (sm_100a).
//
// Replaces essentia ChromaCrossSimilarity(frameStackSize=m, frameStackStride=tau,
// binarizePercentile=kappa, oti=True)(query, reference) as called at
// acoss/algorithms/rqa_serra09.py:60-66 (semantics: SURVEY Appendix A D1-D7,
// restated in oracle/acoss_oracle.c:oracle_crp).
//
// Structure ("strip" kernels): the thresholds of D7 are percentiles of whole rows
// (query axis) and whole columns (reference axis) of the M x N distance matrix.
// A CTA owns a strip of R_S consecutive rows of "side A" and ALL columns of
// "side B"; the strip's distances live in shared memory (never in HBM), the
// per-row order statistics are selected there, and
//   pass 1 (side A = reference): writes thr_r[N]           (column thresholds)
//   pass 2 (side A = query):     computes thr_q[M] and binarises against both,
//                                writing the CRP to HBM in the DP kernel's
//                                "delta" byte code (0x02 = recurrence, 0xFF = none).
// HBM traffic per pair: the two tracks' frames (L2-resident), (M+N) thresholds
// and the M x pitch CRP bytes.  The distance matrix is recomputed once per pass
// instead of being written (4 B/cell) and read back twice.
#pragma once
#include "common.cuh"

struct StripTask {
    int32_t pair;
    int32_t row0;     // first side-A row of this strip
    int32_t nrows;    // <= R_S
    int32_t pad;
};

#define CRP_DELTA_ONE  0x02u   // R = 1: score delta +2 half-units (+1.0)
#define CRP_DELTA_ZERO 0xFFu   // R = 0: score delta -1 half-unit  (-0.5), sign-extends to 0xFFFF

#define STRIP_THREADS 256

// D6: percentile by linear interpolation between two adjacent order statistics.
// n values per row; qf = (kappa*100)/100 in fp32 (host-computed the oracle's way).
struct PercentileRank {
    int lo, hi;       // order statistics needed (0-based)
    float w_lo, w_hi; // interpolation weights; hi == lo -> take v[lo]
};
__device__ __forceinline__ PercentileRank percentile_rank(int n, float qf)
{
    PercentileRank pr;
    if (n == 1) { pr.lo = pr.hi = 0; pr.w_lo = 1.f; pr.w_hi = 0.f; return pr; }
    float k = xmul((float)(n - 1), qf);
    float fl = floorf(k), ce = ceilf(k);
    pr.lo = (int)fl; pr.hi = (int)ce;
    pr.w_lo = xsub(ce, k);
    pr.w_hi = xsub(k, fl);
    return pr;
}
__device__ __forceinline__ float percentile_combine(const PercentileRank& pr, float v_lo, float v_hi)
{
    if (pr.lo == pr.hi) return v_lo;
    return xadd(xmul(v_lo, pr.w_lo), xmul(v_hi, pr.w_hi));
}

// Warp-cooperative exact selection of order statistics `k` and `k+1` (0-based)
// among n non-negative floats in shared memory.  Bitwise bisection on the IEEE
// bit pattern (monotone for non-negative floats): 31 counting passes, then one
// pass for the successor.  Returns v_k; v_k1 by reference (== v_k when k+1 >= n).
__device__ __forceinline__ float warp_select_pair(const float* __restrict__ v, int n, int k,
                                                  int lane, float& v_k1)
{
    uint32_t ans = 0;
    for (int bit = 30; bit >= 0; --bit) {
        uint32_t cand = ans | (1u << bit);
        int cnt = 0;
        for (int j = lane; j < n; j += 32) cnt += (__float_as_uint(v[j]) < cand);
        cnt = __reduce_add_sync(0xffffffffu, cnt);
        if (cnt <= k) ans = cand;   // at most k elements are < cand  =>  v_(k) >= cand
    }
    int cle = 0;
    uint32_t mgt = 0xffffffffu;
    for (int j = lane; j < n; j += 32) {
        uint32_t key = __float_as_uint(v[j]);
        cle += (key <= ans);
        if (key > ans && key < mgt) mgt = key;
    }
    cle = __reduce_add_sync(0xffffffffu, cle);
    mgt = __reduce_min_sync(0xffffffffu, mgt);
    v_k1 = (cle >= k + 2 || mgt == 0xffffffffu) ? __uint_as_float(ans) : __uint_as_float(mgt);
    return __uint_as_float(ans);
}

// Exact strip kernel.  Dynamic shared memory layout (floats):
//   As  [12*m][R_S]   embedded side-A rows, element-major so one LDS.128 feeds 4 rows
//   xxs [R_S]         exact self dots of the side-A rows
//   thA [R_S]         thresholds of the strip rows
//   Bt  [12][TBF]     side-B frames of the current column tile, bin-major
//   V   [R_S][NBpad]  distances of the strip
// side_a_is_query = 0: pass 1 (thr_r);  1: pass 2 (thr_q + CRP).
__global__ void __launch_bounds__(STRIP_THREADS, 1)
crp_strip_exact_kernel(const PairDesc* __restrict__ descs,
                       const StripTask* __restrict__ tasks, int n_tasks,
                       const float* __restrict__ frames,
                       const float* __restrict__ selfdot,
                       float* __restrict__ thr,          // batch threshold buffer
                       uint8_t* __restrict__ crp,        // batch CRP buffer (pass 2)
                       int side_a_is_query, int m, int tau, float qf,
                       int R_S, int NBpad, int TBF)
{
    extern __shared__ __align__(16) float smem[];
    const int d = ACOSS_NBINS * m;
    float* As = smem;
    float* xxs = As + d * R_S;
    float* thA = xxs + R_S;
    float* Bt = thA + R_S;
    float* V = Bt + ACOSS_NBINS * TBF;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int step = (m == 1) ? 1 : tau;   // frame step between embedded rows

    for (int task = blockIdx.x; task < n_tasks; task += gridDim.x) {
        const StripTask tk = tasks[task];
        const PairDesc pd = descs[tk.pair];
        const int rot = pd.oti % ACOSS_NBINS;
        // side A / side B geometry
        const int MA = side_a_is_query ? pd.M : pd.N;
        const int NB = side_a_is_query ? pd.N : pd.M;
        const int64_t a_f0 = side_a_is_query ? pd.q_frame_off : pd.r_frame_off;
        const int64_t b_f0 = side_a_is_query ? pd.r_frame_off : pd.q_frame_off;
        const int nb_frames = side_a_is_query ? pd.nr : pd.nq;
        const int rot_a = side_a_is_query ? 0 : rot;     // rotation applies to the reference
        const int rot_b = side_a_is_query ? rot : 0;
        const float* sdA = selfdot + (side_a_is_query ? pd.q_sd_off : pd.r_sd_off) + (int64_t)rot_a * MA;
        const float* sdB = selfdot + (side_a_is_query ? pd.r_sd_off : pd.q_sd_off) + (int64_t)rot_b * NB;
        (void)MA;

        __syncthreads();   // previous task's readers are done with smem
        // ---- stage side-A rows: As[e][r] = x_{row0+r}[e], zero for r >= nrows
        for (int idx = tid; idx < d * R_S; idx += STRIP_THREADS) {
            int e = idx / R_S, r = idx - e * R_S;
            float v = 0.0f;
            if (r < tk.nrows) {
                int s = e / ACOSS_NBINS, c = e - s * ACOSS_NBINS;
                int64_t f = a_f0 + (int64_t)(tk.row0 + r) * step + (int64_t)s * tau;
                v = frames[f * ACOSS_NBINS + rot_src(c, rot_a)];
            }
            As[idx] = v;
        }
        if (tid < R_S) xxs[tid] = (tid < tk.nrows) ? sdA[tk.row0 + tid] : 0.0f;

        // ---- loop over side-B column tiles
        for (int j0 = 0; j0 < NB; j0 += STRIP_THREADS) {
            __syncthreads();
            // frames j0*step .. (j0+255)*step + (m-1)*tau, clipped to the track
            const int fb0 = j0 * step;
            for (int idx = tid; idx < ACOSS_NBINS * TBF; idx += STRIP_THREADS) {
                int c = idx / TBF, fl = idx - c * TBF;
                int f = fb0 + fl;
                float v = 0.0f;
                if (f < nb_frames) v = frames[(b_f0 + f) * ACOSS_NBINS + rot_src(c, rot_b)];
                Bt[idx] = v;
            }
            __syncthreads();
            const int j = j0 + tid;
            if (j < NB) {
                const float sb = sdB[j];
                const float* yb = Bt + tid * step;
                for (int r = 0; r < R_S; r += 4) {
                    if (r >= tk.nrows) break;
                    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
                    for (int s = 0; s < m; ++s) {
                        const float* ys = yb + s * tau;
                        const float4* xs = reinterpret_cast<const float4*>(As + (s * ACOSS_NBINS) * R_S + r);
#pragma unroll
                        for (int c = 0; c < ACOSS_NBINS; ++c) {
                            float y = ys[c * TBF];
                            float4 x = xs[c * (R_S / 4)];
                            a0 = xadd(a0, xmul(x.x, y));
                            a1 = xadd(a1, xmul(x.y, y));
                            a2 = xadd(a2, xmul(x.z, y));
                            a3 = xadd(a3, xmul(x.w, y));
                        }
                    }
                    float acc[4] = {a0, a1, a2, a3};
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        float sa = xxs[r + k];
                        float xx = side_a_is_query ? sa : sb;   // x = query, y = reference (D5)
                        float yy = side_a_is_query ? sb : sa;
                        float item = xadd(xsub(xx, xmul(2.0f, acc[k])), yy);
                        if (item < 0.0f) item = 0.0f;
                        V[(r + k) * NBpad + j] = __fsqrt_rn(item);
                    }
                }
            }
        }
        __syncthreads();

        // ---- per-row percentile thresholds (one warp per row)
        const PercentileRank pr = percentile_rank(NB, qf);
        float* thr_out = thr + (side_a_is_query ? pd.thr_q_off : pd.thr_r_off);
        for (int r = warp; r < tk.nrows; r += STRIP_THREADS / 32) {
            float v_hi;
            float v_lo = warp_select_pair(V + r * NBpad, NB, pr.lo, lane, v_hi);
            float t = percentile_combine(pr, v_lo, v_hi);
            if (lane == 0) { thA[r] = t; thr_out[tk.row0 + r] = t; }
        }
        if (!side_a_is_query) continue;   // pass 1 ends here (uniform per launch)
        __syncthreads();

        // ---- pass 2: binarise against both thresholds, write delta-coded CRP
        const float* thB = thr + pd.thr_r_off;
        const int words = pd.pitch >> 2;
        for (int r = 0; r < tk.nrows; ++r) {
            uint32_t* out = reinterpret_cast<uint32_t*>(crp + pd.crp_off + (int64_t)(tk.row0 + r) * pd.pitch);
            const float ta = thA[r];
            const float* vr = V + r * NBpad;
            for (int w = tid; w < words; w += STRIP_THREADS) {
                uint32_t word = 0;
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    int j = w * 4 + b;
                    uint32_t code = CRP_DELTA_ZERO;
                    if (j < NB) {
                        float dv = vr[j];
                        if (dv <= ta && dv <= thB[j]) code = CRP_DELTA_ONE;
                    }
                    word |= code << (8 * b);
                }
                out[w] = word;
            }
        }
    }
}
