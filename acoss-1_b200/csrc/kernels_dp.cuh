// kernels_dp.cuh -- Qmax / Dmax / constrained Smith-Waterman recurrences (sm_100a).
//
// Replaces essentia CoverSongSimilarity(alignmentType='serra09'|'chen17',
// distanceType='symmetric')(csm)[1] (acoss/algorithms/rqa_serra09.py:64,67,
// latefusion_chen.py:67-71; semantics SURVEY Appendix A D8/D9, restated in
// oracle/acoss_oracle.c:oracle_cover_similarity) and
// smith_waterman_constrained (acoss/algorithms/utils/alignment_tools.py:25-46).
//
// Key observation: cell (i, j) depends on (i-1, j-1), (i-2, j-1), (i-1, j-2) only,
// i.e. on the two previous ROWS.  A row is therefore fully data-parallel; the
// only serial dimension is i.  One warp owns one pair and keeps rows i-1 and
// i-2 of the score matrix entirely in registers:
//   * scores are exact integers in half units (gamma_o = gamma_e = 0.5 =>
//     Q in (1/2)Z): R=1 -> +2, R=0 -> max(0, . - 1); two cells per 32-bit
//     register as int16x2, updated with the DPX instructions
//     (__vimax3_s16x2 / __viaddmax_s16x2);
//   * a lane owns "groups" of 2G consecutive cells; inside a group register p
//     holds cells (p, p+G), so the "column j-1" and "column j-2" views of a row
//     are register renames except at p = 0/1, where one PRMT + one SHFL fetch
//     the neighbour group's edge;
//   * the CRP row arrives as one coalesced 2G-byte load per lane and chunk
//     (chunk = 32 lanes x 2G cells, 512 B contiguous for G = 8), in the "delta"
//     byte code (0x02 / 0xFF): one PRMT with sign replication expands two bytes
//     into the packed int16x2 increment.
// HBM traffic = the CRP bytes, read exactly once, fully coalesced: the kernel is
// bound by HBM bandwidth at 1 B/cell.
#pragma once
#include "common.cuh"

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

// ---- coalesced chunk loads of 2G delta bytes per lane ---------------------------
template <int G> struct RowBytes;
// plain 0 / 1 bytes -> 0x00 (one) / 0xFF (zero): the sign-replicating PRMT of delta<P>() then yields 0x0000 / 0xFFFF halves
__device__ __forceinline__ uint32_t sw_code(uint32_t w) { return ~(w * 0xFFu); }

template <> struct RowBytes<8> {
    uint4 v;
    __device__ __forceinline__ void load(const uint8_t* p) { v = __ldg(reinterpret_cast<const uint4*>(p)); }
    __device__ __forceinline__ void to_sw() { v = make_uint4(sw_code(v.x), sw_code(v.y), sw_code(v.z), sw_code(v.w)); }
    __device__ __forceinline__ void fill(uint32_t w) { v = make_uint4(w, w, w, w); }
    template <int P> __device__ __forceinline__ uint32_t delta() const {
        const uint32_t x = (P >> 2) == 0 ? v.x : v.y;
        const uint32_t y = (P >> 2) == 0 ? v.z : v.w;
        constexpr uint32_t b = P & 3;
        return prmt(x, y, b | ((b | 8) << 4) | ((4 + b) << 8) | (((4 + b) | 8) << 12));
    }
};
template <> struct RowBytes<4> {
    uint2 v;
    __device__ __forceinline__ void load(const uint8_t* p) { v = __ldg(reinterpret_cast<const uint2*>(p)); }
    __device__ __forceinline__ void to_sw() { v = make_uint2(sw_code(v.x), sw_code(v.y)); }
    __device__ __forceinline__ void fill(uint32_t w) { v = make_uint2(w, w); }
    template <int P> __device__ __forceinline__ uint32_t delta() const {
        constexpr uint32_t b = P;
        return prmt(v.x, v.y, b | ((b | 8) << 4) | ((4 + b) << 8) | (((4 + b) | 8) << 12));
    }
};
template <> struct RowBytes<2> {
    uint32_t v;
    __device__ __forceinline__ void load(const uint8_t* p) { v = __ldg(reinterpret_cast<const uint32_t*>(p)); }
    __device__ __forceinline__ void to_sw() { v = sw_code(v); }
    __device__ __forceinline__ void fill(uint32_t w) { v = w; }
    template <int P> __device__ __forceinline__ uint32_t delta() const {
        constexpr uint32_t b = P;
        return prmt(v, v, b | ((b | 8) << 4) | ((2 + b) << 8) | (((2 + b) | 8) << 12));
    }
};

// bonus of the chen17 recurrence: +R (= +2 half units) from the packed delta
// (0x0002 -> 2, 0xFFFF -> 0):  delta & ~(delta >> 1) & 0x00020002
__device__ __forceinline__ uint32_t delta_to_bonus(uint32_t dl)
{
    return dl & ~(dl >> 1) & 0x00020002u;
}

template <int G, int P> struct StepUnroll;

// One row of the recurrence for one warp: row i-1 is in Aprev, row i-2 in Bold; the new row i is written IN PLACE
// over Bold (register P of the new row needs Bold[P-1] only, so descending P never clobbers a live value).
// Calling it with the two arrays swapped on alternate rows removes every row-to-row register copy.
template <int G, int NCH, int ALIGN>
__device__ __forceinline__ void dp_row_step(uint32_t (&Aprev)[NCH][G], uint32_t (&Bold)[NCH][G], uint32_t (&carry)[NCH],
                                            uint32_t (&Bon)[NCH][G], const RowBytes<G> (&cur)[NCH],
                                            const uint32_t (&vmask)[ALIGN ? G : 1], int nch, int lane, uint32_t zmask,
                                            uint32_t& rm)
{
    const int src = (lane + 31) & 31;
    const bool edge = (lane == 31);
    // chunks in DESCENDING order so that Aprev[ch-1] still holds row i-1
#pragma unroll
    for (int ch = NCH - 1; ch >= 0; --ch) {
        if (ch < nch) {
            // left neighbour group's edge registers of row i-1: lane 31 of chunk ch-1 is the left neighbour of lane 0
            // of chunk ch; for ch == 0 lane 0 has no left neighbour (columns < 0 are zero)
            const uint32_t e1 = (ch > 0) ? Aprev[ch > 0 ? ch - 1 : 0][G - 1] : 0u;
            const uint32_t e2 = (ch > 0) ? Aprev[ch > 0 ? ch - 1 : 0][G - 2] : 0u;
            const uint32_t g1 = edge ? e1 : Aprev[ch][G - 1];
            const uint32_t g2 = edge ? e2 : Aprev[ch][G - 2];
            const uint32_t n1 = __shfl_sync(0xffffffffu, g1, src);
            const uint32_t n2 = __shfl_sync(0xffffffffu, g2, src);
            const uint32_t s1_0 = prmt(n1, Aprev[ch][G - 1], 0x5432);   // (nbr cell 2G-1, own cell G-1)
            const uint32_t s2_0 = prmt(n2, Aprev[ch][G - 2], 0x5432);   // (nbr cell 2G-2, own cell G-2)
            uint32_t dsh = 0;   // chen17: packed delta of (i, j-1) for p = 0
            if (ALIGN == 1) {
                const uint32_t dl_last = cur[ch].template delta<G - 1>();
                const uint32_t dprev = (ch > 0) ? cur[ch > 0 ? ch - 1 : 0].template delta<G - 1>() : 0xFFFFFFFFu;
                const uint32_t nd = __shfl_sync(0xffffffffu, edge ? dprev : dl_last, src);
                dsh = prmt(nd, dl_last, 0x5432);
            }
            StepUnroll<G, G - 1>::template run<ALIGN>(Aprev[ch], Bold[ch], cur[ch], s1_0, s2_0, carry[ch], Bon[ch], dsh);
            if (ch == 0) { Bold[0][0] &= zmask; Bold[0][1] &= zmask; }
            carry[ch] = s1_0;
            if (ALIGN == 2) {
                // constrained Smith-Waterman: Bold holds S of the new row; the maximum is taken over S (last chunk: without
                // the excluded last column and the padding), then the row is stored as U = S + dl(B) for its successors
                if (ch == nch - 1) {
#pragma unroll
                    for (int p = 0; p < G; p += 2)
                        rm = __vimax3_s16x2(rm, Bold[ch][p] & vmask[p], Bold[ch][p + 1] & vmask[p + 1]);
                } else {
#pragma unroll
                    for (int p = 0; p < G; p += 2) rm = __vimax3_s16x2(rm, Bold[ch][p], Bold[ch][p + 1]);
                }
                StepUnroll<G, G - 1>::sw_penalty(cur[ch], Bold[ch]);
                continue;
            }
            if (ALIGN == 1) {
                StepUnroll<G, G - 1>::bonus(cur[ch], Bon[ch]);
                if (ch == nch - 1) {
                    // a padded column can receive the +R bonus of the last valid column: keep it out of the maximum
#pragma unroll
                    for (int p = 0; p < G; p += 2)
                        rm = __vimax3_s16x2(rm, Bold[ch][p] & vmask[p], Bold[ch][p + 1] & vmask[p + 1]);
                    continue;
                }
            }
#pragma unroll
            for (int p = 0; p < G; p += 2) rm = __vimax3_s16x2(rm, Bold[ch][p], Bold[ch][p + 1]);
        }
    }
}

// One warp per pair.  ALIGN: 0 = serra09 (Qmax), 1 = chen17 (Dmax).
// Requirements: pitch % 16 == 0 (16-byte aligned rows), N <= pitch, N <= NCH*64*G, min(M,N) small enough for
// int16 half units (host checks: 4*min(M,N) < 32000).
template <int G, int NCH, int ALIGN>
__global__ void __launch_bounds__(128)
dp_rowpar_kernel(const PairDesc* __restrict__ descs, int64_t n_pairs,
                 const uint8_t* __restrict__ crp, float* __restrict__ out,
                 const int32_t* __restrict__ order, unsigned* __restrict__ counter)
{
    // order / counter (both or neither): pairs are handed out dynamically, largest M x N first (the host sorts `order`),
    // so that a batch of several waves of unequal pairs ends with a short tail; null: static round-robin.
    constexpr int CHUNK = 64 * G;       // cells (= bytes) per chunk
    constexpr int LB = 2 * G;           // bytes per lane and chunk
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const uint32_t zmask = (lane == 0) ? 0xFFFF0000u : 0xFFFFFFFFu;   // clears cells j = 0, 1

    for (int64_t it = warp0;; it += nwarps) {
        int64_t k = it;
        if (counter) {
            unsigned kk = 0;
            if (lane == 0) kk = atomicAdd(counter, 1u);
            k = __shfl_sync(0xffffffffu, kk, 0);
        }
        if (k >= n_pairs) break;
        const PairDesc pd = descs[order ? order[k] : k];
        const int M = pd.M, N = pd.N, pitch = pd.pitch;
        const int nch = (N + CHUNK - 1) / CHUNK;
        const uint8_t* base = crp + pd.crp_off + (int64_t)lane * LB;
        // the pitch is a multiple of 16 B (128 B from the pair pipeline), not of the chunk: in the last chunk the lanes past the
        // pitch load nothing and keep the "no recurrence" fill, so no padding beyond the 128-byte line is ever streamed
        const bool ok_last = (nch - 1) * CHUNK + (lane + 1) * LB <= pitch;
#define DP_LD_OK(ch) ((ch) < nch - 1 || ((ch) == nch - 1 && ok_last))

        uint32_t A[NCH][G], B[NCH][G], carry[NCH];     // rows i-1, i-2 (roles alternate); carried (i-2) edge
        uint32_t Bon[NCH][G];                          // chen17: +R of row i-1 (unshifted)
        uint32_t rm = 0;
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
            carry[ch] = 0;
#pragma unroll
            for (int p = 0; p < G; ++p) { A[ch][p] = 0; B[ch][p] = 0; Bon[ch][p] = 0; }
        }
        // valid-column masks of the last chunk (chen17 only)
        uint32_t vmask[ALIGN ? G : 1];
        if (ALIGN) {
            const int c0 = (nch - 1) * CHUNK + lane * LB;
#pragma unroll
            for (int p = 0; p < G; ++p)
                vmask[p] = ((c0 + p < N) ? 0x0000FFFFu : 0u) | ((c0 + p + G < N) ? 0xFFFF0000u : 0u);
        } else vmask[0] = 0;

        RowBytes<G> r0[NCH], r1[NCH];
        // chunks beyond the pair's width are never loaded: they stay "no recurrence" for the whole pair
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) { r0[ch].fill(0xFFFFFFFFu); r1[ch].fill(0xFFFFFFFFu); }
        if (M > 2) {
            if (ALIGN) {
                // bonus of row 1
#pragma unroll
                for (int ch = 0; ch < NCH; ++ch)
                    if (ch < nch) {
                        RowBytes<G> rb;
                        rb.fill(0xFFFFFFFFu);
                        if (DP_LD_OK(ch)) rb.load(base + (int64_t)1 * pitch + ch * CHUNK);
                        StepUnroll<G, G - 1>::bonus(rb, Bon[ch]);
                    }
            }
#pragma unroll
            for (int ch = 0; ch < NCH; ++ch)
                if (DP_LD_OK(ch)) r0[ch].load(base + (int64_t)2 * pitch + ch * CHUNK);
        }
        const int64_t pitch64 = pitch;
        const uint8_t* rp = base + 2 * pitch64;             // this lane's bytes of row i
        const bool pf_lane = ((lane * LB) & 127) == 0;      // one L2 prefetch per 128-byte line
        for (int i = 2; i < M; i += 2, rp += 2 * pitch64) {
            // row i + 1 -> r1 while row i is processed
            const bool more1 = (i + 1 < M);
            if (more1) {
                const uint8_t* p1 = rp + pitch64;
#pragma unroll
                for (int ch = 0; ch < NCH; ++ch)
                    if (DP_LD_OK(ch)) r1[ch].load(p1 + ch * CHUNK);
            }
            if (pf_lane && i + 7 < M) {
                const uint8_t* p6 = rp + 4 * pitch64;
                const uint8_t* p7 = p6 + pitch64;
#pragma unroll
                for (int ch = 0; ch < NCH; ++ch)
                    if (DP_LD_OK(ch)) {
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(p6 + ch * CHUNK));
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(p7 + ch * CHUNK));
                    }
            }
            dp_row_step<G, NCH, ALIGN>(A, B, carry, Bon, r0, vmask, nch, lane, zmask, rm);      // row i   -> B
            if (more1) {
                if (i + 2 < M) {
                    const uint8_t* p2 = rp + 2 * pitch64;
#pragma unroll
                    for (int ch = 0; ch < NCH; ++ch)
                        if (DP_LD_OK(ch)) r0[ch].load(p2 + ch * CHUNK);
                }
                dp_row_step<G, NCH, ALIGN>(B, A, carry, Bon, r1, vmask, nch, lane, zmask, rm);  // row i+1 -> A
            }
        }
        int best = max((int)(rm & 0xFFFFu), (int)(rm >> 16));
        best = __reduce_max_sync(0xffffffffu, best);
        if (lane == 0) out[pd.out_idx] = 0.5f * (float)best;
    }
#undef DP_LD_OK
}

// ---- constrained Smith-Waterman on the same row-parallel DPX scheme ---------------------------------------------------
// smith_waterman_constrained(B) (acoss/algorithms/utils/alignment_tools.py:25-46), exact in integer tenths (increments
// {+1, +0.3, -1, -1.7} x 10).  With the score matrix shifted by one row and one column, S'(i, j) = S(i + 1, j + 1),
//     S'(i, j) = max(0, mv(B(i, j)) + max(U(i-1, j-1), U(i-2, j-1), U(i-1, j-2))),   U(i, j) = S'(i, j) + dl(B(i, j)),
// mv = +10 / -10, dl = 0 / -7 by B(i, j): every term is indexed like B itself and the dependency shape is Qmax's (two previous
// rows), so dp_row_step<G, NCH, 2> applies: rows of U in registers as int16x2, VIMNMX3 + 2 x VIADDMNMX per register.
// S' = 0 for i < 2 or j < 2 (U = dl(B) there); the reference's loops stop at i = M - 1, j = N - 1, i.e. the last row and the
// last column of B are never scored: the row loop ends at M - 2 and the maximum masks columns >= N - 1.
// Input: plain 0 / 1 bytes, row pitch a multiple of the chunk (64 G bytes), padding bytes 0.  Output: the maximum in tenths.
struct SwMat { int64_t off; int32_t M, N, pitch, out_idx; };

template <int G, int NCH>
__global__ void __launch_bounds__(128)
sw_rowpar_kernel(const SwMat* __restrict__ mats, int64_t n_mats, const uint8_t* __restrict__ bin, float* __restrict__ out_tenths,
                 unsigned* __restrict__ counter)
{
    constexpr int CHUNK = 64 * G, LB = 2 * G;
    const int lane = threadIdx.x & 31;
    const uint32_t zmask = (lane == 0) ? 0xFFFF0000u : 0xFFFFFFFFu;   // clears cells j = 0, 1
    for (;;) {
        unsigned kk = 0;
        if (lane == 0) kk = atomicAdd(counter, 1u);
        const int64_t k = __shfl_sync(0xffffffffu, kk, 0);
        if (k >= n_mats) break;
        const SwMat mt = mats[k];
        const int M = mt.M, N = mt.N, pitch = mt.pitch;
        const int nch = (N + CHUNK - 1) / CHUNK;
        const uint8_t* base = bin + mt.off + (int64_t)lane * LB;
        uint32_t A[NCH][G], B[NCH][G], carry[NCH], Bon[NCH][G];
        uint32_t rm = 0;
        uint32_t vmask[G];
        {
            const int c0 = (nch - 1) * CHUNK + lane * LB;
#pragma unroll
            for (int p = 0; p < G; ++p)
                vmask[p] = ((c0 + p < N - 1) ? 0x0000FFFFu : 0u) | ((c0 + p + G < N - 1) ? 0xFFFF0000u : 0u);
        }
        RowBytes<G> r0[NCH], r1[NCH];
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
            carry[ch] = 0;
            r0[ch].fill(0xFFFFFFFFu); r1[ch].fill(0xFFFFFFFFu);       // chunks beyond the matrix: B = 0
#pragma unroll
            for (int p = 0; p < G; ++p) { A[ch][p] = 0; B[ch][p] = 0; Bon[ch][p] = 0; }
        }
        if (M >= 4 && N >= 4) {
            // rows 0 and 1: S' = 0, U = dl(B).  Row i-1 must sit in A and row i-2 in B for i = 2: row 0 -> B, row 1 -> A; the
            // carried (i-2, j-1) edge of row 0 is rebuilt from B the way dp_row_step builds it
#pragma unroll
            for (int ch = 0; ch < NCH; ++ch)
                if (ch < nch) {
                    RowBytes<G> t0, t1;
                    t0.load(base + ch * CHUNK); t0.to_sw();
                    t1.load(base + (int64_t)pitch + ch * CHUNK); t1.to_sw();
                    StepUnroll<G, G - 1>::sw_penalty(t0, B[ch]);
                    StepUnroll<G, G - 1>::sw_penalty(t1, A[ch]);
                }
            {
                const int src = (lane + 31) & 31;
                const bool edge = (lane == 31);
#pragma unroll
                for (int ch = 0; ch < NCH; ++ch)
                    if (ch < nch) {
                        const uint32_t e1 = (ch > 0) ? B[ch > 0 ? ch - 1 : 0][G - 1] : 0u;
                        const uint32_t g1 = edge ? e1 : B[ch][G - 1];
                        const uint32_t n1 = __shfl_sync(0xffffffffu, g1, src);
                        carry[ch] = prmt(n1, B[ch][G - 1], 0x5432);
                    }
            }
#pragma unroll
            for (int ch = 0; ch < NCH; ++ch)
                if (ch < nch) { r0[ch].load(base + (int64_t)2 * pitch + ch * CHUNK); r0[ch].to_sw(); }
            const int64_t pitch64 = pitch;
            const uint8_t* rp = base + 2 * pitch64;
            const bool pf_lane = ((lane * LB) & 127) == 0;
            const int Mend = M - 1;                              // the last row of B is never scored
            for (int i = 2; i < Mend; i += 2, rp += 2 * pitch64) {
                const bool more1 = (i + 1 < Mend);
                if (more1) {
                    const uint8_t* p1 = rp + pitch64;
#pragma unroll
                    for (int ch = 0; ch < NCH; ++ch)
                        if (ch < nch) { r1[ch].load(p1 + ch * CHUNK); r1[ch].to_sw(); }
                }
                if (pf_lane && i + 7 < Mend) {
                    const uint8_t* p6 = rp + 4 * pitch64;
                    const uint8_t* p7 = p6 + pitch64;
#pragma unroll
                    for (int ch = 0; ch < NCH; ++ch)
                        if (ch < nch) {
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(p6 + ch * CHUNK));
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(p7 + ch * CHUNK));
                        }
                }
                dp_row_step<G, NCH, 2>(A, B, carry, Bon, r0, vmask, nch, lane, zmask, rm);      // row i   -> B
                if (more1) {
                    if (i + 2 < Mend) {
                        const uint8_t* p2 = rp + 2 * pitch64;
#pragma unroll
                        for (int ch = 0; ch < NCH; ++ch)
                            if (ch < nch) { r0[ch].load(p2 + ch * CHUNK); r0[ch].to_sw(); }
                    }
                    dp_row_step<G, NCH, 2>(B, A, carry, Bon, r1, vmask, nch, lane, zmask, rm);  // row i+1 -> A
                }
            }
        }
        int best = max((int)(rm & 0xFFFFu), (int)(rm >> 16));
        best = __reduce_max_sync(0xffffffffu, best);
        if (lane == 0) out_tenths[mt.out_idx] = (float)best;
    }
}

// Fully unrolled per-register update, P = G-1 .. 0 (recursion keeps every register index a compile-time
// constant).  Writes the new row over B in place.
template <int G, int P> struct StepUnroll {
    template <int ALIGN>
    __device__ __forceinline__ static void run(const uint32_t (&A)[G], uint32_t (&B)[G],
                                               const RowBytes<G>& row, uint32_t s1_0, uint32_t s2_0,
                                               uint32_t carry, const uint32_t (&Bon)[G], uint32_t dsh)
    {
        const uint32_t a1 = (P >= 1) ? A[P >= 1 ? P - 1 : 0] : s1_0;                 // row i-1, col j-1
        const uint32_t b1 = (P >= 1) ? B[P >= 1 ? P - 1 : 0] : carry;                // row i-2, col j-1
        const uint32_t a2 = (P >= 2) ? A[P >= 2 ? P - 2 : 0] : (P == 1 ? s1_0 : s2_0); // row i-1, col j-2
        const uint32_t dl = row.template delta<P>();
        uint32_t m;
        if (ALIGN == 2) {
            // constrained Smith-Waterman in integer tenths (alignment_tools.py:25-46), rows hold U = S + dl(B):
            //   S = max(0, mv + max(U(i-1,j-1), U(i-2,j-1), U(i-1,j-2))),  mv = +10 / -10 by B(i,j)   (dl: 0x0000 = one, 0xFFFF = zero)
            m = __vimax3_s16x2(a1, b1, a2);
            B[P] = __viaddmax_s16x2(m, 0x000A000Au ^ (dl & 0xFFFCFFFCu), 0u);
            StepUnroll<G, P - 1>::template run<ALIGN>(A, B, row, s1_0, s2_0, carry, Bon, dsh);
            return;
        }
        if (ALIGN) {
            // chen17: (i-2,j-1) + R(i-1,j)  and  (i-1,j-2) + R(i,j-1)
            const uint32_t dprev = (P >= 1) ? row.template delta<(P >= 1 ? P - 1 : 0)>() : dsh;
            const uint32_t b1b = b1 + Bon[P];
            const uint32_t a2b = a2 + delta_to_bonus(dprev);
            m = __vimax3_s16x2(a1, b1b, a2b);
        } else {
            m = __vimax3_s16x2(a1, b1, a2);
        }
        B[P] = __viaddmax_s16x2(m, dl, 0u);    // R=1: m+2 ; R=0: max(m-1, 0)
        StepUnroll<G, P - 1>::template run<ALIGN>(A, B, row, s1_0, s2_0, carry, Bon, dsh);
    }
    __device__ __forceinline__ static void bonus(const RowBytes<G>& row, uint32_t (&Bon)[G])
    {
        Bon[P] = delta_to_bonus(row.template delta<P>());
        StepUnroll<G, P - 1>::bonus(row, Bon);
    }
    // U = S + dl(B): -7 tenths where B = 0 (packed add without carry between the halves: viaddmax against -32768)
    __device__ __forceinline__ static void sw_penalty(const RowBytes<G>& row, uint32_t (&B)[G])
    {
        B[P] = __viaddmax_s16x2(B[P], row.template delta<P>() & 0xFFF9FFF9u, 0x80008000u);
        StepUnroll<G, P - 1>::sw_penalty(row, B);
    }
};
template <int G> struct StepUnroll<G, -1> {
    template <int ALIGN>
    __device__ __forceinline__ static void run(const uint32_t (&)[G], uint32_t (&)[G], const RowBytes<G>&,
                                               uint32_t, uint32_t, uint32_t, const uint32_t (&)[G], uint32_t) {}
    __device__ __forceinline__ static void bonus(const RowBytes<G>&, uint32_t (&)[G]) {}
    __device__ __forceinline__ static void sw_penalty(const RowBytes<G>&, uint32_t (&)[G]) {}
};

// ---- generic fallback: any N, any gammas, fp32 like the oracle -------------------
// One warp per pair; rows i-1 / i-2 of the score matrix and the needed CRP rows in
// shared memory.  Used for tracks longer than the register kernel supports and for
// disOnset != disExtension or values other than 0.5.  CRP in delta code.
__global__ void __launch_bounds__(32)
dp_generic_kernel(const PairDesc* __restrict__ descs, int64_t n_pairs,
                  const uint8_t* __restrict__ crp, int alignment,
                  float g_onset, float g_ext, int maxN, float* __restrict__ out)
{
    extern __shared__ float sm[];
    float* S1 = sm;                 // row i-1
    float* S2 = S1 + maxN;          // row i-2
    float* S0 = S2 + maxN;          // row i
    const int lane = threadIdx.x;
    for (int64_t pidx = blockIdx.x; pidx < n_pairs; pidx += gridDim.x) {
        const PairDesc pd = descs[pidx];
        const int M = pd.M, N = pd.N;
        const uint8_t* R = crp + pd.crp_off;
        const int64_t pitch = pd.pitch;
        for (int j = lane; j < N; j += 32) { S1[j] = 0.f; S2[j] = 0.f; S0[j] = 0.f; }
        __syncwarp();
        float best = 0.f;
        for (int i = 2; i < M; ++i) {
            const uint8_t* r0 = R + (int64_t)i * pitch;
            const uint8_t* r1 = r0 - pitch;
            const uint8_t* r2 = r1 - pitch;
            for (int j = 2 + lane; j < N; j += 32) {
                float p1 = S1[j - 1], p2 = S2[j - 1], p3 = S1[j - 2];
                if (alignment == 1) {
                    p2 += (r1[j] == CRP_DELTA_ONE) ? 1.f : 0.f;
                    p3 += (r0[j - 1] == CRP_DELTA_ONE) ? 1.f : 0.f;
                }
                float v;
                if (r0[j] == CRP_DELTA_ONE) {
                    v = fmaxf(p1, fmaxf(p2, p3)) + 1.0f;
                } else {
                    float c1 = p1 - ((r1[j - 1] == CRP_DELTA_ONE) ? g_onset : g_ext);
                    float c2 = p2 - ((r2[j - 1] == CRP_DELTA_ONE) ? g_onset : g_ext);
                    float c3 = p3 - ((r1[j - 2] == CRP_DELTA_ONE) ? g_onset : g_ext);
                    v = fmaxf(fmaxf(0.f, c1), fmaxf(c2, c3));
                }
                S0[j] = v;
                best = fmaxf(best, v);
            }
            __syncwarp();
            float* t = S2; S2 = S1; S1 = S0; S0 = t;
            // S0 (old i-2 row) keeps stale values at j >= 2, all overwritten next row; j < 2 stay 0
        }
        for (int o = 16; o > 0; o >>= 1) best = fmaxf(best, __shfl_xor_sync(0xffffffffu, best, o));
        if (lane == 0) out[pd.out_idx] = best;
        __syncwarp();
    }
}

// ---- constrained Smith-Waterman (alignment_tools.py:25-46), integer tenths ----------
// S[i][j] = max(0, S[i-1][j-1]+mv+dl(B[i-2][j-2]), S[i-2][j-1]+mv+dl(B[i-3][j-2]),
//               S[i-1][j-2]+mv+dl(B[i-2][j-3])),  mv = +1/-1 by B[i-1][j-1],
// dl(v) = 0 if v > 0 else -0.7 (Q7: the gap_opening branch is unreachable).
// All increments are multiples of 0.1: {+1, +0.3, -1, -1.7} -> exact int32 tenths.
// Same dependency shape as Qmax (two previous rows) => same row-parallel scheme;
// one warp per matrix, rows in shared memory.  Input: plain 0/1 bytes.
__global__ void __launch_bounds__(32)
sw_constrained_kernel(const uint8_t* __restrict__ mats, const int64_t* __restrict__ mat_off,
                      const int32_t* __restrict__ Ms, const int32_t* __restrict__ Ns,
                      int n, int maxN, double* __restrict__ out)
{
    extern __shared__ int smi[];
    int* S1 = smi;
    int* S2 = S1 + maxN;
    int* S0 = S2 + maxN;
    const int lane = threadIdx.x;
    for (int k = blockIdx.x; k < n; k += gridDim.x) {
        const int M = Ms[k], N = Ns[k];
        const uint8_t* Bm = mats + mat_off[k];
        int best = 0;
        if (M >= 4 && N >= 4) {
            for (int j = lane; j < N; j += 32) { S1[j] = 0; S2[j] = 0; S0[j] = 0; }
            __syncwarp();
            for (int i = 3; i < M; ++i) {
                const uint8_t* b1 = Bm + (int64_t)(i - 1) * N;
                const uint8_t* b2 = b1 - N;
                const uint8_t* b3 = b2 - N;
                for (int j = 3 + lane; j < N; j += 32) {
                    int mv = b1[j - 1] ? 10 : -10;
                    int d1 = S1[j - 1] + mv + (b2[j - 2] ? 0 : -7);
                    int d2 = S2[j - 1] + mv + (b3[j - 2] ? 0 : -7);
                    int d3 = S1[j - 2] + mv + (b2[j - 3] ? 0 : -7);
                    int v = max(max(d1, d2), max(d3, 0));
                    S0[j] = v;
                    best = max(best, v);
                }
                __syncwarp();
                int* t = S2; S2 = S1; S1 = S0; S0 = t;
            }
            best = __reduce_max_sync(0xffffffffu, best);
        }
        if (lane == 0) out[k] = (double)best / 10.0;
        __syncwarp();
    }
}
