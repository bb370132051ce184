// kernels_prep.cuh -- per-track preparation and per-pair OTI (sm_100a).
//
// Reference semantics (essentia, restated in oracle/acoss_oracle.c, SURVEY
// Appendix A D1): g = sum of frames / max(g); oti = first argmax over k = 0..12
// of <g_q, roll(g_r, k)>.  All sums sequential fp32 without FMA so the device
// result is bit-identical to the oracle's.
#pragma once
#include "common.cuh"

// One warp per track.  Lanes 0..11 own one chroma bin each and walk the frames
// sequentially (the order of the oracle's sum); the max is a 12-lane shuffle.
__global__ void global_chroma_kernel(const float* __restrict__ frames,
                                     const int64_t* __restrict__ offs,
                                     int n_tracks, float* __restrict__ gchroma)
{
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (warp >= n_tracks) return;
    int64_t f0 = offs[warp], f1 = offs[warp + 1];
    float acc = 0.0f;
    if (lane < ACOSS_NBINS)
        for (int64_t t = f0; t < f1; ++t) acc = xadd(acc, frames[t * ACOSS_NBINS + lane]);
    float mx = acc;
    // max over lanes 0..11 (lanes >= 12 hold 0, sums of non-negative chroma are >= 0;
    // to be safe for arbitrary input they are excluded explicitly)
    float v = lane < ACOSS_NBINS ? acc : -INFINITY;
    mx = v;
    for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
    if (lane < ACOSS_NBINS) {
        float g = (mx != 0.0f) ? __fdiv_rn(acc, mx) : acc;
        gchroma[(int64_t)warp * ACOSS_NBINS + lane] = g;
    }
}

// Exact self dot products of the delay-embedded rows, for all 12 rotations:
// sd[sd_off[track] + rot*M + t] = dot(x_t, x_t), x_t = concat_s roll(frame[t*tau + s*tau], rot),
// summed in the oracle's order (s major, bin minor).  Rotation changes the
// order in which the bins are visited, hence the fp32 rounding, hence one table
// per rotation.  Thread per (rot, t), grid.y = track.
__global__ void selfdot_kernel(const float* __restrict__ frames,
                               const int64_t* __restrict__ offs,
                               const int32_t* __restrict__ emb_len,
                               const int64_t* __restrict__ sd_off,
                               int m, int tau, float* __restrict__ sd, int track0)
{
    int track = track0 + blockIdx.y;
    int M = emb_len[track];
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= M * ACOSS_NBINS) return;
    int rot = idx / M, t = idx - rot * M;
    const float* f = frames + (offs[track] + (int64_t)(m == 1 ? t : t * tau)) * ACOSS_NBINS;
    float acc = 0.0f;
    for (int s = 0; s < m; ++s) {
        const float* fs = f + (int64_t)s * tau * ACOSS_NBINS;
#pragma unroll
        for (int c = 0; c < ACOSS_NBINS; ++c) {
            float v = fs[rot_src(c, rot)];
            acc = xadd(acc, xmul(v, v));
        }
    }
    sd[sd_off[track] + (int64_t)rot * M + t] = acc;
}

// Thread per pair: 13 shifts x 12 sequential products.
__global__ void oti_kernel(PairDesc* __restrict__ descs, int64_t n_pairs,
                           const float* __restrict__ gchroma, int use_oti)
{
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    if (!use_oti) { descs[p].oti = 0; return; }
    float gq[ACOSS_NBINS], gr[ACOSS_NBINS];
    const float* a = gchroma + (int64_t)descs[p].q * ACOSS_NBINS;
    const float* b = gchroma + (int64_t)descs[p].r * ACOSS_NBINS;
#pragma unroll
    for (int c = 0; c < ACOSS_NBINS; ++c) { gq[c] = a[c]; gr[c] = b[c]; }
    int best = 0;
    float bestv = 0.0f;
#pragma unroll
    for (int k = 0; k <= ACOSS_NBINS; ++k) {
        float acc = 0.0f;
#pragma unroll
        for (int c = 0; c < ACOSS_NBINS; ++c) {
            int src = (c - k + 2 * ACOSS_NBINS) % ACOSS_NBINS;
            acc = xadd(acc, xmul(gq[c], gr[src]));
        }
        if (k == 0 || acc > bestv) { bestv = acc; best = k; }
    }
    descs[p].oti = best;
}

// librosa.util.sync(chroma.T, arange(0, T, fac), aggregate=np.median).T for all
// tracks: one thread per (output frame, bin); the segment (<= fac values) is
// sorted in local memory by insertion; np.median = mean of the two middle
// values for even counts, computed as numpy does (float32 mean of two: (a+b)/2
// via np.mean -> add in fp32, divide by 2).
template <int MAXFAC>
__global__ void median_downsample_kernel(const float* __restrict__ raw,
                                         const int64_t* __restrict__ raw_offs,
                                         const int64_t* __restrict__ out_offs,
                                         int n_tracks, int fac, float* __restrict__ out, int track0)
{
    int track = track0 + blockIdx.y;
    int64_t T = raw_offs[track + 1] - raw_offs[track];
    int64_t nout = out_offs[track + 1] - out_offs[track];
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nout * ACOSS_NBINS) return;
    int64_t seg = idx / ACOSS_NBINS;
    int c = (int)(idx - seg * ACOSS_NBINS);
    int64_t b0 = seg * fac, b1 = b0 + fac;
    if (b1 > T) b1 = T;
    int cnt = (int)(b1 - b0);
    float v[MAXFAC];
    const float* src = raw + (raw_offs[track] + b0) * ACOSS_NBINS + c;
    for (int k = 0; k < cnt; ++k) {
        float x = src[(int64_t)k * ACOSS_NBINS];
        int j = k;
        while (j > 0 && v[j - 1] > x) { v[j] = v[j - 1]; --j; }
        v[j] = x;
    }
    float med;
    if (cnt & 1) med = v[cnt >> 1];
    else med = __fdiv_rn(xadd(v[(cnt >> 1) - 1], v[cnt >> 1]), 2.0f);
    out[(out_offs[track] + seg) * ACOSS_NBINS + c] = med;
}
