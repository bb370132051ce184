// kernels_ftm.cuh -- FTM2D per-track shingles on the device (sm_100a).
//
// Replaces the body of FTM2D.load_features (acoss/algorithms/ftm2d.py:52-66) and the helpers it calls, chrompwr (:100-117) and
// btchroma_to_fftmat (:120-138):
//   hpcp   = librosa.util.sync(chroma.T, onsets, aggregate=np.median)     beat-synchronous median (boundaries {0} u onsets u {T})
//   chroma = chrompwr(hpcp, PWR)                                          per beat: x / |x| to the power PWR, renormalised to |x| (float32)
//   patch  = fftshift(|fft2(chroma[:, i : i + WIN])|).flatten()            one 12 x WIN 2-D DFT magnitude per beat position i
//   s_i    = log(C * patch / |patch| + 1)                                 (float64)
//   shingle = median_i s_i, normalised to unit length                      (float64, 12 * WIN values)
//
// The 2-D DFT is separable: the 12-point DFT over the chroma axis is computed ONCE per beat (ftm_coldft_kernel), the WIN-point
// DFT over time per window position directly from those columns (ftm_window_kernel, twiddles in shared memory).  Arithmetic:
// the reference's scipy.fftpack.fft2 runs in complex64 on the float32 chroma, with whatever summation order its mixed-radix
// FFT has; here every DFT sum is accumulated in float64 from the same float32 inputs and the magnitude is rounded to float32 as
// the reference's is -- the two agree to float32 rounding of the magnitudes (tests: 2e-5 of the largest shingle entry).
#pragma once
#include "common.cuh"

#define FTM_MAXSEG 512          // longest beat segment (frames) the median kernel sorts in local memory

// beat-synchronous median: thread per (segment, bin).  bounds = concatenated segment boundaries of all tracks (per track
// b_0 = 0 < b_1 < ... < b_S = T), seg_off[t] = first segment of track t, out[(seg_off[t] + s) * 12 + c].
__global__ void ftm_sync_median_kernel(const float* __restrict__ raw, const int64_t* __restrict__ raw_offs,
                                       const int32_t* __restrict__ bounds, const int64_t* __restrict__ bound_off,
                                       const int64_t* __restrict__ seg_off, int track0, float* __restrict__ out)
{
    const int track = track0 + blockIdx.y;
    const int64_t nseg = seg_off[track + 1] - seg_off[track];
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nseg * ACOSS_NBINS) return;
    const int64_t seg = idx / ACOSS_NBINS;
    const int c = (int)(idx - seg * ACOSS_NBINS);
    const int32_t* b = bounds + bound_off[track] + seg;
    const int b0 = b[0], b1 = b[1];
    const int cnt = min(b1 - b0, FTM_MAXSEG);          // the host rejects longer segments
    float v[FTM_MAXSEG];
    const float* src = raw + (raw_offs[track] + b0) * ACOSS_NBINS + c;
    for (int k = 0; k < cnt; ++k) {
        const float x = src[(int64_t)k * ACOSS_NBINS];
        int j = k;
        while (j > 0 && v[j - 1] > x) { v[j] = v[j - 1]; --j; }
        v[j] = x;
    }
    float med;
    if (cnt & 1) med = v[cnt >> 1];
    else med = __fdiv_rn(__fadd_rn(v[(cnt >> 1) - 1], v[cnt >> 1]), 2.0f);      // np.median of an even count: float32 mean of the two
    out[(seg_off[track] + seg) * ACOSS_NBINS + c] = med;
}

// chrompwr (float32, numpy's op order: sums over the 12 bins sequentially) + the 12-point DFT over the chroma axis (float64)
__global__ void ftm_coldft_kernel(const float* __restrict__ hpcp, int64_t n_beats, float pwr, double2* __restrict__ coldft)
{
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_beats) return;
    float x[ACOSS_NBINS], p[ACOSS_NBINS];
    float nn = 0.0f;
#pragma unroll
    for (int c = 0; c < ACOSS_NBINS; ++c) { x[c] = hpcp[t * ACOSS_NBINS + c]; nn = __fadd_rn(nn, __fmul_rn(x[c], x[c])); }
    float cmn = __fsqrt_rn(nn);
    if (cmn == 0.0f) cmn = 1.0f;
    float pn = 0.0f;
#pragma unroll
    for (int c = 0; c < ACOSS_NBINS; ++c) { p[c] = powf(__fdiv_rn(x[c], cmn), pwr); pn = __fadd_rn(pn, __fmul_rn(p[c], p[c])); }
    float cmpn = __fsqrt_rn(pn);
    if (cmpn == 0.0f) cmpn = 1.0f;
    double y[ACOSS_NBINS];
#pragma unroll
    for (int c = 0; c < ACOSS_NBINS; ++c) y[c] = (double)__fmul_rn(cmn, __fdiv_rn(p[c], cmpn));
    // exact twelfth roots of unity: cos / sin of multiples of 30 degrees
    const double h = 0.5, s3 = 0.86602540378443864676;
    const double co[12] = {1.0, s3, h, 0.0, -h, -s3, -1.0, -s3, -h, 0.0, h, s3};
    const double si[12] = {0.0, h, s3, 1.0, s3, h, 0.0, -h, -s3, -1.0, -s3, -h};
#pragma unroll
    for (int k = 0; k < ACOSS_NBINS; ++k) {
        double re = 0.0, im = 0.0;
#pragma unroll
        for (int a = 0; a < ACOSS_NBINS; ++a) { const int m = (k * a) % 12; re += y[a] * co[m]; im -= y[a] * si[m]; }
        coldft[t * ACOSS_NBINS + k] = make_double2(re, im);
    }
}

// one block per (window position, track): F[k][l] = sum_t C[k][i + t] e^{-2 pi i l t / WIN}, |F| rounded to float32, fft-shifted
// and flattened (12 x WIN, row-major), s = log(C * |F| / norm + 1) with norm = sqrt(sum |F|^2) (0 -> 1), float64
__global__ void __launch_bounds__(256)
ftm_window_kernel(const double2* __restrict__ coldft, const int64_t* __restrict__ beat_off, const int64_t* __restrict__ win_off,
                  int win, double Cc, double* __restrict__ svals)
{
    extern __shared__ double2 fsm[];
    double2* tw = fsm;                       // [win] twiddles
    double2* cs = fsm + win;                 // [12][win] column DFTs of this window
    __shared__ double red[8];
    const int track = blockIdx.y;
    const int64_t nwin = win_off[track + 1] - win_off[track];
    const int i = blockIdx.x;
    if (i >= nwin) return;
    const int tid = threadIdx.x;
    for (int t = tid; t < win; t += blockDim.x) {
        double s, c;
        sincospi(2.0 * (double)t / (double)win, &s, &c);
        tw[t] = make_double2(c, -s);
    }
    for (int e = tid; e < ACOSS_NBINS * win; e += blockDim.x) {
        const int t = e / ACOSS_NBINS, k = e - t * ACOSS_NBINS;
        cs[k * win + t] = coldft[(beat_off[track] + i + t) * ACOSS_NBINS + k];
    }
    __syncthreads();
    const int total = ACOSS_NBINS * win;
    double sq = 0.0;
    double* out = svals + (win_off[track] + i) * total;
    for (int e = tid; e < total; e += blockDim.x) {
        const int k = e / win, l = e - k * win;
        double re = 0.0, im = 0.0;
        int m = 0;                           // (l * t) mod win
        for (int t = 0; t < win; ++t) {
            const double2 c = cs[k * win + t], w = tw[m];
            re += c.x * w.x - c.y * w.y;
            im += c.x * w.y + c.y * w.x;
            m += l; if (m >= win) m -= win;
        }
        const float mag = __fsqrt_rn(__fadd_rn(__fmul_rn((float)re, (float)re), __fmul_rn((float)im, (float)im)));
        // fftshift: out[(k + 6) % 12][(l + win / 2) % win] = F[k][l]
        const int ks = (k + ACOSS_NBINS / 2) % ACOSS_NBINS, ls = (l + win / 2) % win;
        out[ks * win + ls] = (double)mag;
        sq += (double)mag * (double)mag;
    }
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    if ((tid & 31) == 0) red[tid >> 5] = sq;
    __syncthreads();
    double norm = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) norm += red[w];
    norm = sqrt(norm);
    if (norm == 0.0) norm = 1.0;
    for (int e = tid; e < total; e += blockDim.x) out[e] = log(Cc * out[e] / norm + 1.0);
}

// median over the window positions of every one of the 12 * WIN entries (np.median(shingles, 0), float64): warp per (track, entry),
// bitwise bisection on the order-preserving 64-bit key; even counts average the two middle values
__device__ __forceinline__ unsigned long long ftm_key(double d)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(d);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double ftm_key_inv(unsigned long long k)
{
    const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}
__global__ void __launch_bounds__(256)
ftm_median_kernel(const double* __restrict__ svals, const int64_t* __restrict__ win_off, int n_tracks, int total, double* __restrict__ shingle)
{
    const int lane = threadIdx.x & 31;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (wid >= (int64_t)n_tracks * total) return;
    const int track = (int)(wid / total), e = (int)(wid - (int64_t)track * total);
    const int n = (int)(win_off[track + 1] - win_off[track]);
    const double* col = svals + win_off[track] * total + e;
    auto kth = [&](int k) {
        unsigned long long ans = 0;
        for (int bit = 63; bit >= 0; --bit) {
            const unsigned long long cand = ans | (1ull << bit);
            int cnt = 0;
            for (int j = lane; j < n; j += 32) cnt += (ftm_key(col[(int64_t)j * total]) < cand);
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if (cnt <= k) ans = cand;
        }
        return ftm_key_inv(ans);
    };
    double med;
    if (n & 1) med = kth(n >> 1);
    else med = (kth((n >> 1) - 1) + kth(n >> 1)) / 2.0;
    if (lane == 0) shingle[(int64_t)track * total + e] = med;
}

// shingle / sqrt(sum(shingle ** 2)): block per track
__global__ void __launch_bounds__(256)
ftm_normalize_kernel(double* __restrict__ shingle, int total)
{
    __shared__ double red[8];
    double* s = shingle + (int64_t)blockIdx.x * total;
    double sq = 0.0;
    for (int e = threadIdx.x; e < total; e += blockDim.x) sq += s[e] * s[e];
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sq;
    __syncthreads();
    double norm = 0.0;
    for (int w = 0; w < 8; ++w) norm += red[w];
    norm = sqrt(norm);
    for (int e = threadIdx.x; e < total; e += blockDim.x) s[e] = s[e] / norm;
}
