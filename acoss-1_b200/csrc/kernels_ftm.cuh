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

// =====================================================================================================================
// EarlyFusionTralie block features (acoss/algorithms/earlyfusion_traile.py:104-150, resize_block :214-247).
//
// resize_block(X, i1, i2, n_out) = skimage.transform.resize(X[i1:i2].astype(float64), (n_out, n_feats), anti_aliasing=True,
// mode='constant').  skimage is an unpinned dependency that is absent here (parity of this stage is UNPINNED); the restatement
// (oracle/ports.py:resize_block, which uses the real scipy.ndimage.gaussian_filter for step 1) follows skimage 0.15 / 0.16:
// with f = (i2 - i1) / n_out,
//   1. Gaussian filter along time only (the feature axis keeps its size: sigma = 0): sigma = max(0, (f - 1) / 2), taps
//      exp(-k^2 / (2 sigma^2)) for |k| <= int(4 sigma + 0.5), normalised, samples outside the block = 0 (mode='constant');
//   2. order-1 warp: output frame o samples row r = f (o + 0.5) - 0.5 by linear interpolation between floor(r) and ceil(r),
//      rows outside [0, n_in - 1] count as 0 (cval).
// Both steps are fused: an output needs the filtered values of two rows only.  float64 like the reference.
// =====================================================================================================================
#define EF_MAXLW 1024           // Gaussian half-width limit (frames)

// normalised Gaussian taps of one resize in shared memory, w[|k|]; returns the half-width.  All threads of the block call it.
__device__ __forceinline__ int ef_resize_taps(double* w, int n_in, int n_out, int tid, int nthreads)
{
    const double f = (double)n_in / (double)n_out;
    const double sigma = fmax(0.0, (f - 1.0) / 2.0);
    const int lw = sigma > 1e-15 ? min((int)(4.0 * sigma + 0.5), EF_MAXLW) : 0;
    __syncthreads();
    for (int k = tid; k <= lw; k += nthreads) w[k] = lw ? exp(-0.5 / (sigma * sigma) * (double)k * (double)k) : 1.0;
    __syncthreads();
    if (tid == 0) {
        double sum = 0.0;
        for (int k = -lw; k <= lw; ++k) sum += w[k < 0 ? -k : k];
        w[EF_MAXLW + 1] = sum;
    }
    __syncthreads();
    const double sum = w[EF_MAXLW + 1];
    __syncthreads();
    for (int k = tid; k <= lw; k += nthreads) w[k] = w[k] / sum;
    __syncthreads();
    return lw;
}

__device__ __forceinline__ double ef_resized(const float* __restrict__ X, int64_t row0, int n_in, int n_feat, int c, int n_out, int o,
                                             const double* __restrict__ w, int lw)
{
    const double f = (double)n_in / (double)n_out;
    const double r = f * ((double)o + 0.5) - 0.5;
    const double rf = floor(r), rc = ceil(r);
    const double dr = r - rf;
    double out = 0.0;
#pragma unroll
    for (int side = 0; side < 2; ++side) {
        if (side == 1 && rc == rf) break;               // r is an integer: one sample, weight 1
        const int rr = (int)(side ? rc : rf);
        const double wgt = side ? dr : 1.0 - dr;
        if (rr < 0 || rr > n_in - 1) continue;
        double g = 0.0;
        for (int k = -lw; k <= lw; ++k) {
            const int q = rr + k;
            if (q < 0 || q >= n_in) continue;
            double v = (double)X[(row0 + q) * n_feat + c];
            if (isnan(v)) v = 0.0;                       // mfcc[np.isnan(mfcc)] = 0 (earlyfusion_traile.py:107)
            g += w[k < 0 ? -k : k] * v;
        }
        out += wgt * g;
    }
    if (isinf(out) || isnan(out)) out = 0.0;            // ret[np.isinf(ret)] = 0; ret[np.isnan(ret)] = 0
    return out;
}

struct EfBlockTask { int64_t frame0_m, frame0_c, out_blk; int32_t i1, i2m, i2c, pad; };   // MFCC frames [i1, i2m), chroma [i1, i2c)

// one CTA per block: resized + z-normalised MFCC block (n_mf frames x n_mfcc), its SSM lower triangle, resized chroma block
__global__ void __launch_bounds__(256)
ef_block_features_kernel(const float* __restrict__ mfcc, const float* __restrict__ chroma, const EfBlockTask* __restrict__ tasks, int n_tasks,
                         int n_mfcc, int n_mf, int n_ch, float* __restrict__ out_mfccs, float* __restrict__ out_ssms, float* __restrict__ out_chromas)
{
    extern __shared__ double bsm[];
    double* x = bsm;                         // [n_mf][n_mfcc]
    double* colmean = x + n_mf * n_mfcc;     // [n_mfcc]
    double* rnorm = colmean + n_mfcc;        // [n_mf] row norms, then squared norms of the normalised rows
    double* taps = rnorm + n_mf;             // [EF_MAXLW + 2]
    const int tid = threadIdx.x;
    for (int t = blockIdx.x; t < n_tasks; t += gridDim.x) {
        const EfBlockTask tk = tasks[t];
        int lw = ef_resize_taps(taps, tk.i2m - tk.i1, n_mf, tid, blockDim.x);
        for (int e = tid; e < n_mf * n_mfcc; e += blockDim.x) {
            const int o = e / n_mfcc, c = e - o * n_mfcc;
            x[e] = ef_resized(mfcc, tk.frame0_m + tk.i1, tk.i2m - tk.i1, n_mfcc, c, n_mf, o, taps, lw);
        }
        __syncthreads();
        for (int c = tid; c < n_mfcc; c += blockDim.x) {            // x -= np.mean(x, 0)
            double s = 0.0;
            for (int o = 0; o < n_mf; ++o) s += x[o * n_mfcc + c];
            colmean[c] = s / (double)n_mf;
        }
        __syncthreads();
        for (int e = tid; e < n_mf * n_mfcc; e += blockDim.x) x[e] -= colmean[e % n_mfcc];
        __syncthreads();
        for (int o = tid; o < n_mf; o += blockDim.x) {              // xn = x / sqrt(sum(x ** 2, 1)), zero norms -> 1
            double s = 0.0;
            for (int c = 0; c < n_mfcc; ++c) s += x[o * n_mfcc + c] * x[o * n_mfcc + c];
            s = sqrt(s);
            rnorm[o] = s == 0.0 ? 1.0 : s;
        }
        __syncthreads();
        float* om = out_mfccs + tk.out_blk * (int64_t)(n_mf * n_mfcc);
        for (int e = tid; e < n_mf * n_mfcc; e += blockDim.x) { x[e] = x[e] / rnorm[e / n_mfcc]; om[e] = (float)x[e]; }
        __syncthreads();
        for (int o = tid; o < n_mf; o += blockDim.x) {
            double s = 0.0;
            for (int c = 0; c < n_mfcc; ++c) s += x[o * n_mfcc + c] * x[o * n_mfcc + c];
            rnorm[o] = s;
        }
        __syncthreads();
        // get_ssm(xn)[I < J] with I, J = np.meshgrid(pix, pix): entries (r, c), c < r, in row-major order
        float* os = out_ssms + tk.out_blk * (int64_t)(n_mf * (n_mf - 1) / 2);
        for (int e = tid; e < n_mf * (n_mf - 1) / 2; e += blockDim.x) {
            int r = (int)((1.0 + sqrt(1.0 + 8.0 * (double)e)) / 2.0);
            while (r * (r - 1) / 2 > e) --r;
            while ((r + 1) * r / 2 <= e) ++r;
            const int c = e - r * (r - 1) / 2;
            double dot = 0.0;
            for (int k = 0; k < n_mfcc; ++k) dot += x[r * n_mfcc + k] * x[c * n_mfcc + k];
            double d2 = rnorm[r] + rnorm[c] - 2.0 * dot;
            if (d2 < 0.0) d2 = 0.0;
            os[e] = (float)sqrt(d2);
        }
        float* oc = out_chromas + tk.out_blk * (int64_t)(n_ch * ACOSS_NBINS);
        lw = ef_resize_taps(taps, tk.i2c - tk.i1, n_ch, tid, blockDim.x);
        for (int e = tid; e < n_ch * ACOSS_NBINS; e += blockDim.x) {
            const int o = e / ACOSS_NBINS, c = e - o * ACOSS_NBINS;
            oc[e] = (float)ef_resized(chroma, tk.frame0_c + tk.i1, tk.i2c - tk.i1, ACOSS_NBINS, c, n_ch, o, taps, lw);
        }
    }
}

// np.median(chroma, axis=0) per track: warp per (track, bin), bisection on the float32 key
__global__ void __launch_bounds__(128)
ef_chroma_median_kernel(const float* __restrict__ chroma, const int64_t* __restrict__ frame_off, int n_tracks, float* __restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (wid >= n_tracks * ACOSS_NBINS) return;
    const int track = wid / ACOSS_NBINS, c = wid - track * ACOSS_NBINS;
    const int n = (int)(frame_off[track + 1] - frame_off[track]);
    const float* col = chroma + frame_off[track] * ACOSS_NBINS + c;
    auto key = [](float f) { const uint32_t b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); };
    auto kth = [&](int k) {
        uint32_t ans = 0;
        for (int bit = 31; bit >= 0; --bit) {
            const uint32_t cand = ans | (1u << bit);
            int cnt = 0;
            for (int j = lane; j < n; j += 32) cnt += (key(col[(int64_t)j * ACOSS_NBINS]) < cand);
            cnt = __reduce_add_sync(0xffffffffu, cnt);
            if (cnt <= k) ans = cand;
        }
        return __uint_as_float((ans & 0x80000000u) ? (ans & 0x7fffffffu) : ~ans);
    };
    float med = 0.0f;
    if (n > 0) med = (n & 1) ? kth(n >> 1) : __fdiv_rn(__fadd_rn(kth((n >> 1) - 1), kth(n >> 1)), 2.0f);
    if (lane == 0) out[track * ACOSS_NBINS + c] = med;
}
