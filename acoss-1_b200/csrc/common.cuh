// common.cuh -- shared device/host definitions of the acoss_b200 hot path (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#define ACOSS_NBINS 12
#define ACOSS_N_COUNTERS 48        // device profiling counters per context (debug ABI exposes them)
#define ACOSS_NUM_SMS 148

// One (query, reference) pair of a batch, filled by the host planner except `oti`.
struct PairDesc {
    int32_t q, r;            // track indices
    int32_t nq, nr;          // frames per track (after downsampling)
    int32_t M, N;            // delay-embedded rows (query / reference)
    int32_t oti;             // optimal transposition index 0..12 (device-written)
    int32_t pitch;           // CRP row pitch in bytes (multiple of the DP chunk)
    int64_t q_frame_off;     // first frame of q / r in the frame store
    int64_t r_frame_off;
    int64_t q_sd_off;        // offset of the self-dot table of q / r (floats)
    int64_t r_sd_off;
    int64_t crp_off;         // byte offset of this pair's CRP in the batch CRP buffer
    int64_t thr_q_off;       // float offset of thr_q[M] / thr_r[N] in the batch threshold buffer
    int64_t thr_r_off;
    int64_t out_idx;         // index of this pair in the caller's output arrays
    int64_t edge_off;        // float offset of this pair's column band edges (phase-split CRP pipeline, pass 2)
};

// exact fp32 arithmetic helpers: the oracle's op order, never contracted to FMA
// ("fp32, sequential sums, no FMA").
__device__ __forceinline__ float xmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float xadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float xsub(float a, float b) { return __fsub_rn(a, b); }

// rotated bin index: roll(frame, rot)[c] = frame[(c - rot) mod 12], rot in 0..11
__device__ __forceinline__ int rot_src(int c, int rot)
{
    int s = c - rot;
    return s < 0 ? s + ACOSS_NBINS : s;
}

#define ACOSS_CUDA_CHECK_RET(ctx, expr)                                           \
    do {                                                                          \
        cudaError_t _e = (expr);                                                  \
        if (_e != cudaSuccess) {                                                  \
            (ctx)->set_error(std::string(#expr) + ": " + cudaGetErrorString(_e)); \
            return ACOSS_ERR_CUDA;                                                \
        }                                                                         \
    } while (0)
