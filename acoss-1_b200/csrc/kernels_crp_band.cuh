// kernels_crp_band.cuh -- fast bit-exact CRP: approximate cross-similarity + rigorous
// error band + exact re-check (sm_100a).  Same contract as kernels_crp.cuh (which stays
// as the always-exact reference path and as the fallback for m != 9 or tau != 1).
//
// Why it is exact.  Let rho_o be the oracle's radicand (fp32, sequential 108-term sums,
// no FMA) and rho~ the value computed here.  Forward error
// analysis (u = 2^-24; all terms relative to xx + yy, the squared norms of the two
// embedded rows; |x.y| <= (xx+yy)/2):
//     oracle:  |rho_o - rho*| <= 220 u (xx + yy)   (three 108-term chains + 2 roundings)
//     here:    |rho~  - rho*| <=  42 u (xx + yy)   (12-term FMA chains, 9-term tree)
//     =>  |rho~ - rho_o| <= 262 u (xx + yy)  <  EPS_REL (xx + yy),   EPS_REL = 2^-15.
// D = sqrt_rn(max(rho_o, 0)) is monotone in rho_o, so
//   * order statistics: the exact k-th smallest lies within E of the approximate k-th
//     smallest; every element whose rho~ is within 2E of it is recomputed exactly (the
//     oracle's op order) and the order statistic is selected among those exact values;
//   * binarisation: D <= eps is decided by rho~ against T = eps*eps when
//     |rho~ - T| > E + 8uT, otherwise the cell is recomputed exactly.
// Everything that is written out (thresholds, CRP bits) therefore equals the oracle's.
//
// How the approximation is computed ("sliding Gram").  With m = 9, tau = 1 the squared
// embedded distance is a diagonal sum of squared 12-d frame distances,
//     rho*(i, j) = sum_{k<9} G2[i+k][j+k],   G2[a][b] = |f_a|^2 + |f_b|^2 - 2 <f_a, f_b>,
// i.e. 12 FMA per G2 entry shared by 9 cells instead of 108 MACs per cell.
//   phase 1: a thread owns one G2 column b (B frame in registers) and streams the strip's
//            24 A frames from shared memory (broadcast LDS.128, pre-scaled by -2); G2 goes to
//            shared memory in a skewed layout Gs[a][(b - a) mod TW] so that
//   phase 2: a thread owns a diagonal delta = b - a, reads its 24 G2 values conflict-free,
//            forms the 16 window sums with a 4-level tree in registers and writes rho~ into
//            the strip buffer V (fp32, shared memory).
#pragma once
#include "common.cuh"
#include "kernels_crp.cuh"
#include "kernels_crp_tc.cuh"

#define BAND_RS 16                    // strip rows
#define BAND_GR (BAND_RS + 8)         // G rows per strip
#define BAND_TW 512                   // G columns per tile
#define BAND_ND (BAND_TW - BAND_GR + 1)   // complete diagonals per tile (489)
#define BAND_THREADS 512
#define BAND_WARPS (BAND_THREADS / 32)
#define BAND_LMAX 256                 // per-warp candidate list capacity
#define BAND_LPL (BAND_LMAX / 32)     // list entries per lane
#define BAND_EPS_REL 3.0517578125e-05f    // 2^-15
#define BAND_SAMPLE 512
#define BAND_SPL (BAND_SAMPLE / 32)
#define BAND_MIN_NB 128               // below this every row takes the exact-row path
#define BAND_AFS 16                   // floats per staged A frame: 12 bins (x -2), |f|^2, pad
#define BAND_WSCR 768                 // per-warp scratch words inside Gs: lidx[256] lval[256] stage[256]
#define BAND_SMEM_FIXED (BAND_GR * BAND_AFS + 3 * BAND_RS + 32 + BAND_GR * BAND_TW)   // + 17 * NBpad floats
#define BAND_RGN_TC (TC_OPB_BYTES / 4 + BAND_GR * TC_TW)   // tensor-core mode: OpB (32 KB) + skewed tile (24 KB) in place of Gs

// exact distance of (query embedded row u, reference embedded row v): the oracle's op order
__device__ __noinline__ float exact_dist(const float* __restrict__ frames, const PairDesc& pd, int rot,
                                         const float* __restrict__ selfdot, int u, int v)
{
    const float* qf = frames + (pd.q_frame_off + u) * ACOSS_NBINS;
    const float* rf = frames + (pd.r_frame_off + v) * ACOSS_NBINS;
    float acc = 0.0f;
    for (int s = 0; s < 9; ++s) {
#pragma unroll
        for (int c = 0; c < ACOSS_NBINS; ++c)
            acc = xadd(acc, xmul(qf[s * ACOSS_NBINS + c], rf[s * ACOSS_NBINS + rot_src(c, rot)]));
    }
    const float xx = selfdot[pd.q_sd_off + u];
    const float yy = selfdot[pd.r_sd_off + (int64_t)rot * pd.N + v];
    float item = xadd(xsub(xx, xmul(2.0f, acc)), yy);
    if (item < 0.0f) item = 0.0f;
    return __fsqrt_rn(item);
}

// order-preserving map float -> uint32 for values of either sign (rho~ can be slightly negative)
__device__ __forceinline__ uint32_t fkey(float f)
{
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float fkey_inv(uint32_t k)
{
    uint32_t b = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
    return __uint_as_float(b);
}

// counters: [0] exact cells  [1] fallback rows  [2..5] cycles (thread 0): stage, tiles, select, binarise
//           [7] tasks  [8..12] cycles (warp 0): pivots, count+list, list select, exact, rank  [13] rows profiled
//
// Shared-memory carve-up (floats):
//   Afr [24*16]  A frames of the strip: -2 * bins (pre-rotated), |f|^2, pad
//   xxA [16] thA [16] flag[16] red[32]
//   yyB [NBpad]  exact self dots of all side-B rows; later: lower band edges of the column thresholds
//   Gs  [24*512] skewed G2 tile; later: per-warp scratch; later: upper band edges + ambiguity list
//   V   [16*NBpad]
//
// TC = 1: the G2 tiles come from the tensor cores (kernels_crp_tc.cuh); shared memory additionally holds the
// strip operand OpA, one mbarrier and the TMEM base address behind V; the Gs region doubles as the side-B
// operand buffer OpB.  `fnorm` = per-frame |f|^2 split (tc_frame_norm_kernel).
#define TC_ND (TC_TW - BAND_GR + 1)      // complete diagonals per tensor-core tile (233)
template <int TC>
__global__ void __launch_bounds__(BAND_THREADS, 1)
crp_strip_band_kernel(const PairDesc* __restrict__ descs,
                      const StripTask* __restrict__ tasks, int n_tasks,
                      const float* __restrict__ frames,
                      const float* __restrict__ selfdot,
                      float* __restrict__ thr, uint8_t* __restrict__ crp,
                      int side_a_is_query, float qf, int NBpad,
                      unsigned long long* __restrict__ counters, const float2* __restrict__ fnorm)
{
    extern __shared__ __align__(16) float smem[];
    float* Afr = smem;
    float* xxA = Afr + BAND_GR * BAND_AFS;
    float* thA = xxA + BAND_RS;
    int* flag = reinterpret_cast<int*>(thA + BAND_RS);
    float* red = reinterpret_cast<float*>(flag + BAND_RS);          // 32 words
    float* yyB = red + 32;
    float* Gs = yyB + NBpad;
    float* V = Gs + (TC ? BAND_RGN_TC : BAND_GR * BAND_TW);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long n_exact = 0;
    // tensor-core state
    unsigned char* OpB = reinterpret_cast<unsigned char*>(Gs);
    unsigned char* OpA = reinterpret_cast<unsigned char*>(V + (size_t)BAND_RS * NBpad);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(OpA + TC_OPA_BYTES);
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(mbar + 1);
    uint32_t tmem_base = 0, mma_phase = 0;
    if (TC) {
        if (warp == 0) tmem_alloc(tmem_holder);
        if (tid == 0) { mbar_init(mbar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
        tmem_base = *tmem_holder;
    }

    for (int task = blockIdx.x; task < n_tasks; task += gridDim.x) {
        const StripTask tk = tasks[task];
        const PairDesc pd = descs[tk.pair];
        const int rot = pd.oti % ACOSS_NBINS;
        const int NB = side_a_is_query ? pd.N : pd.M;
        const int na_frames = side_a_is_query ? pd.nq : pd.nr;
        const int nb_frames = side_a_is_query ? pd.nr : pd.nq;
        const int64_t a_f0 = side_a_is_query ? pd.q_frame_off : pd.r_frame_off;
        const int64_t b_f0 = side_a_is_query ? pd.r_frame_off : pd.q_frame_off;
        // <roll(ref, rot), query>: the B frames stay unrotated in registers; the A frames are staged
        // pre-rotated: by +rot when A is the reference, by -rot when A is the query.
        const int rot_a = side_a_is_query ? (ACOSS_NBINS - rot) % ACOSS_NBINS : rot;
        const float* sdA = selfdot + (side_a_is_query ? pd.q_sd_off : pd.r_sd_off + (int64_t)rot * pd.N);
        const float* sdB = selfdot + (side_a_is_query ? pd.r_sd_off + (int64_t)rot * pd.N : pd.q_sd_off);

        __syncthreads();
        long long ph0 = 0, ph1 = 0, ph2 = 0, ph3 = 0;
        if (tid == 0) ph0 = clock64();
        // ---- stage A frames (x -2, with |f|^2), self dots, flags
        if (tid < BAND_GR * BAND_AFS) {
            const int a = tid / BAND_AFS, c = tid - a * BAND_AFS;
            const int f = tk.row0 + a;
            float v = 0.0f;
            if (f < na_frames) {
                const float* fp = frames + (a_f0 + f) * ACOSS_NBINS;
                if (c < ACOSS_NBINS) v = -2.0f * fp[rot_src(c, rot_a)];
                else if (c == ACOSS_NBINS) {
#pragma unroll
                    for (int q = 0; q < ACOSS_NBINS; ++q) v = fmaf(fp[q], fp[q], v);
                }
            }
            Afr[tid] = v;
        }
        if (tid < BAND_RS) { xxA[tid] = (tid < tk.nrows) ? sdA[tk.row0 + tid] : 0.0f; flag[tid] = 0; }
        float ymax = 0.0f;
        for (int j = tid; j < NB; j += BAND_THREADS) { float y = sdB[j]; yyB[j] = y; ymax = fmaxf(ymax, y); }
        for (int o = 16; o > 0; o >>= 1) ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
        if (lane == 0) red[warp] = ymax;
        __syncthreads();
        ymax = red[0];
#pragma unroll
        for (int w = 1; w < BAND_WARPS; ++w) ymax = fmaxf(ymax, red[w]);
        float xmax = 0.0f;
#pragma unroll
        for (int i = 0; i < BAND_RS; ++i) xmax = fmaxf(xmax, xxA[i]);
        const float E = BAND_EPS_REL * (xmax + ymax);     // one bound for the whole strip (conservative)

        const bool tiny = NB < BAND_MIN_NB;
        if (tid == 0) ph1 = clock64();
        if (!tiny) {
            // ---- approximate radicands of the strip: sliding Gram over column tiles
            const bool full_rows = (tk.nrows == BAND_RS);
            if (TC) {
                // ---- strip operand OpA[slice][chunk][row][4]: built once per task
                for (int e = tid; e < TC_SLICES * 2 * TC_NROWS; e += BAND_THREADS) {
                    const int n = e & (TC_NROWS - 1), ch = (e >> 5) & 1, sl = e >> 6;
                    const bool valid = n < BAND_GR && (tk.row0 + n) < na_frames;
                    const float* ar = Afr + (n < BAND_GR ? n : 0) * BAND_AFS;
                    float4 o;
                    o.x = tc_opa_elem(ar, sl, 4 * ch + 0, valid);
                    o.y = tc_opa_elem(ar, sl, 4 * ch + 1, valid);
                    o.z = tc_opa_elem(ar, sl, 4 * ch + 2, valid);
                    o.w = tc_opa_elem(ar, sl, 4 * ch + 3, valid);
                    *reinterpret_cast<float4*>(OpA + (size_t)e * 16) = o;
                }
                const float* fr_b = frames + b_f0 * ACOSS_NBINS;
                const float2* fn_b = fnorm + b_f0;
                float* GsT = Gs + TC_OPB_BYTES / 4;          // skewed G2 tile, behind the side-B operand buffer
                // thread -> (frame fl = tid >> 1, bins 6 (tid & 1) .. + 5) of a tile; threads 0..255 also own the norm chunks
                const int ld_fl = tid >> 1, ld_h = tid & 1;
                float* ob_row = reinterpret_cast<float*>(OpB) + (size_t)(ld_fl >> 7) * (TC_CHUNKS * TC_BLK * 4) + (size_t)(ld_fl & (TC_BLK - 1)) * 4;
                float4* on_row = reinterpret_cast<float4*>(OpB) + (size_t)(tid >> 7) * (TC_CHUNKS * TC_BLK) + (tid & (TC_BLK - 1));
                float2 px0, px1, px2, pn;
                auto tc_fetch = [&](int b0_) {      // raw bins + norm split of tile b0_ -> registers
                    const int col = b0_ + ld_fl;
                    px0 = px1 = px2 = make_float2(0.f, 0.f);
                    if (col >= 0 && col < nb_frames) {
                        const float2* p = reinterpret_cast<const float2*>(fr_b + (int64_t)col * ACOSS_NBINS + 6 * ld_h);
                        px0 = __ldg(p); px1 = __ldg(p + 1); px2 = __ldg(p + 2);
                    }
                    pn = make_float2(0.f, 0.f);
                    if (tid < TC_TW) { const int cn = b0_ + tid; if (cn >= 0 && cn < nb_frames) pn = __ldg(fn_b + cn); }
                };
                auto tc_store = [&]() {             // hi / lo split -> OpB (chunks 0-2 hi, 3-5 lo, 6 = [1 1 nh nl], 7 = 0)
                    float h0, l0, h1, l1, h2, l2, h3, l3, h4, l4, h5, l5;
                    tf32_split(px0.x, h0, l0); tf32_split(px0.y, h1, l1); tf32_split(px1.x, h2, l2);
                    tf32_split(px1.y, h3, l3); tf32_split(px2.x, h4, l4); tf32_split(px2.y, h5, l5);
                    if (ld_h == 0) {                // bins 0..5: chunk 0 whole, chunk 1 first half
                        *reinterpret_cast<float4*>(ob_row) = make_float4(h0, h1, h2, h3);
                        *reinterpret_cast<float2*>(ob_row + 1 * TC_BLK * 4) = make_float2(h4, h5);
                        *reinterpret_cast<float4*>(ob_row + 3 * TC_BLK * 4) = make_float4(l0, l1, l2, l3);
                        *reinterpret_cast<float2*>(ob_row + 4 * TC_BLK * 4) = make_float2(l4, l5);
                    } else {                        // bins 6..11: chunk 1 second half, chunk 2 whole
                        *reinterpret_cast<float2*>(ob_row + 1 * TC_BLK * 4 + 2) = make_float2(h0, h1);
                        *reinterpret_cast<float4*>(ob_row + 2 * TC_BLK * 4) = make_float4(h2, h3, h4, h5);
                        *reinterpret_cast<float2*>(ob_row + 4 * TC_BLK * 4 + 2) = make_float2(l0, l1);
                        *reinterpret_cast<float4*>(ob_row + 5 * TC_BLK * 4) = make_float4(l2, l3, l4, l5);
                    }
                    if (tid < TC_TW) {
                        on_row[6 * TC_BLK] = make_float4(1.f, 1.f, pn.x, pn.y);
                        on_row[7 * TC_BLK] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                    fence_proxy_async_smem();
                };
                auto tc_mma = [&]() {               // one thread: 2 blocks x 6 k-slices, then commit to the mbarrier
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(OpB), b_base = smem_u32(OpA);
#pragma unroll
                    for (int q = 0; q < 2; ++q)
#pragma unroll
                        for (int sl = 0; sl < TC_SLICES; ++sl) {
                            const int c0 = (sl < 4) ? 2 * sl : 2 * (sl - 4);      // A-operand chunk pair of this slice
                            const uint64_t ad = umma_desc(a_base + (uint32_t)((q * TC_CHUNKS + c0) * TC_BLK * 16), TC_BLK * 16, 128);
                            const uint64_t bd = umma_desc(b_base + (uint32_t)(sl * 2 * TC_NROWS * 16), TC_NROWS * 16, 128);
                            umma_tf32(tmem_base + (uint32_t)(q * TC_NROWS), ad, bd, sl > 0 ? 1u : 0u);
                        }
                    umma_commit(mbar);
                };
                // prologue: tile 0 -> OpB -> MMA
                const int b_first = -(BAND_RS - 1);
                tc_fetch(b_first);
                tc_store();
                tc_fence_before();
                __syncthreads();
                if (tid == 0) tc_mma();
                for (int b0 = b_first; b0 < NB; b0 += TC_ND) {
                    const bool more = (b0 + TC_ND < NB);
                    if (more) tc_fetch(b0 + TC_ND);                 // global loads of the next tile fly during the wait below
                    mbar_wait(mbar, mma_phase);
                    mma_phase ^= 1u;
                    tc_fence_after();
                    // TMEM -> skewed G2 tile: warp w reads lane quarter w & 3 of block (w >> 2) & 1, strip rows 12 (w >> 3) ..
                    {
                        const int quarter = warp & 3, q = (warp >> 2) & 1, half = warp >> 3;
                        float v[16];
                        tmem_ld16(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(q * TC_NROWS + half * 12), v);
                        // unskewed tile GsT[a][column]: lanes = consecutive columns, constant offsets per strip row
                        float* gp = GsT + (half * 12) * TC_TW + q * TC_BLK + quarter * 32 + lane;
#pragma unroll
                        for (int aa = 0; aa < 12; ++aa) gp[aa * TC_TW] = v[aa];
                    }
                    if (more) tc_store();                           // OpB is free: its MMAs have completed
                    tc_fence_before();
                    __syncthreads();
                    if (more && tid == 0) tc_mma();                 // next tile's MMAs run under this tile's window sums
                    // diagonal window sums: thread (delta, h) -> rows 8 h .. 8 h + 7 of diagonal delta
                    if (tid < 2 * TC_ND) {
                        const int h = tid >= TC_ND ? 1 : 0, dl = tid - h * TC_ND;
                        float g[16];
                        // diagonal delta of the tile: G2[8h + a][dl + 8h + a]  (lanes = consecutive diagonals)
                        const float* gq = GsT + (8 * h) * (TC_TW + 1) + dl;
#pragma unroll
                        for (int a = 0; a < 16; ++a) g[a] = gq[a * (TC_TW + 1)];
                        float t2[15], t4[13], t8[9];
#pragma unroll
                        for (int a = 0; a < 15; ++a) t2[a] = g[a] + g[a + 1];
#pragma unroll
                        for (int a = 0; a < 13; ++a) t4[a] = t2[a] + t2[a + 2];
#pragma unroll
                        for (int a = 0; a < 9; ++a) t8[a] = t4[a] + t4[a + 4];
                        const int jb = b0 + dl;
                        float* vp = V + jb + (8 * h) * (NBpad + 1);
                        if (full_rows && jb >= 0 && jb + BAND_RS <= NB) {
#pragma unroll
                            for (int ii = 0; ii < 8; ++ii) { *vp = t8[ii] + g[ii + 8]; vp += NBpad + 1; }
                        } else {
#pragma unroll
                            for (int ii = 0; ii < 8; ++ii) {
                                const int i = 8 * h + ii, j = jb + i;
                                if (i < tk.nrows && j >= 0 && j < NB) *vp = t8[ii] + g[ii + 8];
                                vp += NBpad + 1;
                            }
                        }
                    }
                    __syncthreads();
                }
            } else {
                for (int b0 = -(BAND_RS - 1); b0 < NB; b0 += BAND_ND) {
                    // phase 1: thread (h, t) = (tid >> 8, tid & 255) owns G2 columns b0 + t and b0 + t + 256 for the
                    // A rows [12 h, 12 h + 12): every broadcast A load feeds two columns
                    {
                        const int t = tid & 255, a_base = (tid >> 8) * (BAND_GR / 2);
                        float B0[ACOSS_NBINS], B1[ACOSS_NBINS];
                        const int c0 = b0 + t, c1 = c0 + 256;
                        const bool v0 = (c0 >= 0 && c0 < nb_frames), v1 = (c1 >= 0 && c1 < nb_frames);
                        const float4* p0 = reinterpret_cast<const float4*>(frames + (b_f0 + (v0 ? c0 : 0)) * ACOSS_NBINS);
                        const float4* p1 = reinterpret_cast<const float4*>(frames + (b_f0 + (v1 ? c1 : 0)) * ACOSS_NBINS);
    #pragma unroll
                        for (int q = 0; q < 3; ++q) {
                            const float4 x = v0 ? __ldg(p0 + q) : make_float4(0.f, 0.f, 0.f, 0.f);
                            const float4 y = v1 ? __ldg(p1 + q) : make_float4(0.f, 0.f, 0.f, 0.f);
                            B0[4 * q] = x.x; B0[4 * q + 1] = x.y; B0[4 * q + 2] = x.z; B0[4 * q + 3] = x.w;
                            B1[4 * q] = y.x; B1[4 * q + 1] = y.y; B1[4 * q + 2] = y.z; B1[4 * q + 3] = y.w;
                        }
                        float ff0 = 0.f, ff1 = 0.f;
    #pragma unroll
                        for (int q = 0; q < ACOSS_NBINS; ++q) { ff0 = fmaf(B0[q], B0[q], ff0); ff1 = fmaf(B1[q], B1[q], ff1); }
    #pragma unroll
                        for (int aa = 0; aa < BAND_GR / 2; ++aa) {
                            const int a = a_base + aa;
                            const float4* ap = reinterpret_cast<const float4*>(Afr + a * BAND_AFS);
                            const float4 x0 = ap[0], x1 = ap[1], x2 = ap[2], x3 = ap[3];
                            float g0 = x3.x + ff0, g1 = x3.x + ff1;       // |f_a|^2 + |f_b|^2
                            g0 = fmaf(x0.x, B0[0], g0); g1 = fmaf(x0.x, B1[0], g1);
                            g0 = fmaf(x0.y, B0[1], g0); g1 = fmaf(x0.y, B1[1], g1);
                            g0 = fmaf(x0.z, B0[2], g0); g1 = fmaf(x0.z, B1[2], g1);
                            g0 = fmaf(x0.w, B0[3], g0); g1 = fmaf(x0.w, B1[3], g1);
                            g0 = fmaf(x1.x, B0[4], g0); g1 = fmaf(x1.x, B1[4], g1);
                            g0 = fmaf(x1.y, B0[5], g0); g1 = fmaf(x1.y, B1[5], g1);
                            g0 = fmaf(x1.z, B0[6], g0); g1 = fmaf(x1.z, B1[6], g1);
                            g0 = fmaf(x1.w, B0[7], g0); g1 = fmaf(x1.w, B1[7], g1);
                            g0 = fmaf(x2.x, B0[8], g0); g1 = fmaf(x2.x, B1[8], g1);
                            g0 = fmaf(x2.y, B0[9], g0); g1 = fmaf(x2.y, B1[9], g1);
                            g0 = fmaf(x2.z, B0[10], g0); g1 = fmaf(x2.z, B1[10], g1);
                            g0 = fmaf(x2.w, B0[11], g0); g1 = fmaf(x2.w, B1[11], g1);
                            // Gs[a][(col - a) mod 512], col = t and t + 256
                            Gs[a * BAND_TW + ((t - a) & (BAND_TW - 1))] = g0;
                            Gs[a * BAND_TW + ((t + 256 - a) & (BAND_TW - 1))] = g1;
                        }
                    }
                    __syncthreads();
                    // phase 2: diagonal delta = tid
                    if (tid < BAND_ND) {
                        float g[BAND_GR];
    #pragma unroll
                        for (int a = 0; a < BAND_GR; ++a) g[a] = Gs[a * BAND_TW + tid];
                        float t2[BAND_GR - 1], t4[BAND_GR - 3], t8[BAND_GR - 7];
    #pragma unroll
                        for (int a = 0; a < BAND_GR - 1; ++a) t2[a] = g[a] + g[a + 1];
    #pragma unroll
                        for (int a = 0; a < BAND_GR - 3; ++a) t4[a] = t2[a] + t2[a + 2];
    #pragma unroll
                        for (int a = 0; a < BAND_GR - 7; ++a) t8[a] = t4[a] + t4[a + 4];
                        const int jb = b0 + tid;                       // column of row 0 on this diagonal
                        float* vp = V + jb;
                        if (full_rows && jb >= 0 && jb + BAND_RS <= NB) {
    #pragma unroll
                            for (int i = 0; i < BAND_RS; ++i) vp[i * (NBpad + 1)] = t8[i] + g[i + 8];
                        } else {
    #pragma unroll
                            for (int i = 0; i < BAND_RS; ++i) {
                                const int j = jb + i;
                                if (i < tk.nrows && j >= 0 && j < NB) vp[i * (NBpad + 1)] = t8[i] + g[i + 8];
                            }
                        }
                    }
                    __syncthreads();
                }
            }
        }
        if (tid == 0) ph2 = clock64();

        // ---- per-row percentile thresholds: one warp per row.  Per-warp scratch inside Gs.
        const PercentileRank pr = percentile_rank(NB, qf);
        float* thr_out = thr + (side_a_is_query ? pd.thr_q_off : pd.thr_r_off);
        if (warp < tk.nrows) {
            const int r = warp;
            int* lidx = reinterpret_cast<int*>(Gs) + warp * BAND_WSCR;
            float* lval = reinterpret_cast<float*>(lidx + BAND_LMAX);
            float* stage = lval + BAND_LMAX;
            float* vr = V + r * NBpad;
            const int u_or_v = tk.row0 + r;     // the strip row's index on side A
            bool exact_row = tiny;
            float v_lo = 0.f, v_hi = 0.f;
            const bool prof = (warp == 0 && lane == 0);
            long long tc0 = 0, tc1 = 0, tc2 = 0, tc3 = 0, tc4 = 0;
            if (!exact_row) {
                if (prof) tc0 = clock64();
                // (a) sample (registers): 16 values per lane, evenly spread over the row
                float sm[BAND_SPL];
                float smin = INFINITY, smax = -INFINITY;
#pragma unroll
                for (int q = 0; q < BAND_SPL; ++q) {
                    const int pos = (int)(((long long)(q * 32 + lane) * NB) >> 9);     // / BAND_SAMPLE
                    sm[q] = vr[pos];
                    smin = fminf(smin, sm[q]); smax = fmaxf(smax, sm[q]);
                }
                for (int o = 16; o > 0; o >>= 1) {
                    smin = fminf(smin, __shfl_xor_sync(0xffffffffu, smin, o));
                    smax = fmaxf(smax, __shfl_xor_sync(0xffffffffu, smax, o));
                }
                // (b) 512-bin histogram of the sample (per-warp shared scratch, two 16-bit counters per word) ->
                //     pivots at bin edges.  Pivots only need to be roughly right: pass A below verifies the bracket.
                unsigned* hist = reinterpret_cast<unsigned*>(stage);
                const float hscale = (smax > smin) ? 512.0f / (smax - smin) : 0.0f;
                const float hwidth = (smax - smin) * (1.0f / 512.0f);
#pragma unroll
                for (int q = 0; q < 8; ++q) hist[q * 32 + lane] = 0u;
                __syncwarp();
#pragma unroll
                for (int q = 0; q < BAND_SPL; ++q) {
                    const int bin = min(511, (int)((sm[q] - smin) * hscale));
                    atomicAdd(&hist[bin >> 1], 1u << (16 * (bin & 1)));
                }
                __syncwarp();
                int hc[16], hpre;                        // lane owns bins 16 lane .. 16 lane + 15
                {
                    const uint4 h0 = *reinterpret_cast<const uint4*>(hist + 8 * lane);
                    const uint4 h1 = *reinterpret_cast<const uint4*>(hist + 8 * lane + 4);
                    const unsigned hw[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
                    int tot = 0;
#pragma unroll
                    for (int q = 0; q < 8; ++q) { hc[2 * q] = hw[q] & 0xffffu; hc[2 * q + 1] = hw[q] >> 16; tot += hc[2 * q] + hc[2 * q + 1]; }
                    int incl = tot;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const int t_ = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t_; }
                    hpre = incl - tot;                   // samples in bins < 16 lane
                }
                const float frac = ((float)pr.lo + 0.5f) / (float)NB;
                const float rs = frac * (float)BAND_SAMPLE;
                const float sg = sqrtf(fmaxf(rs * (1.0f - frac), 1.0f));
                bool ok = false;
                int c_a = 0, c_ab = 0;
                float pa = -INFINITY, pb = INFINITY;
                for (int attempt = 0; attempt < 2 && !ok; ++attempt) {
                    const float ns = attempt == 0 ? 2.0f : 4.0f;
                    const int r_lo = (int)floorf(rs - ns * sg) - 1;       // sample ranks the bracket must enclose
                    const int r_hi = (int)ceilf(rs + ns * sg) + 1;
                    // pa = lower edge of the last bin b with #(samples in bins < b) <= r_lo;
                    // pb = upper edge of the first bin b with #(samples in bins <= b) > r_hi
                    int ba = -1, bb = 512;
                    {
                        int run = hpre;
#pragma unroll
                        for (int q = 0; q < 16; ++q) {
                            if (run <= r_lo) ba = 16 * lane + q;
                            run += hc[q];
                            if (run > r_hi && bb == 512) bb = 16 * lane + q;
                        }
                        ba = __reduce_max_sync(0xffffffffu, ba);
                        bb = __reduce_min_sync(0xffffffffu, bb);
                    }
                    pa = (r_lo >= 0 && ba > 0) ? smin + (float)ba * hwidth : -INFINITY;
                    pb = (r_hi < BAND_SAMPLE && bb < 511) ? smin + (float)(bb + 1) * hwidth : INFINITY;
                    if (prof && attempt == 0) tc1 = clock64();
                    // (c) pass A: per-lane counts.  A lane owns 4-column groups (one LDS.128 each); rho~ may be slightly
                    //     negative, so the comparisons stay in floating point; counts accumulate from the predicates.
                    const int n_full = NB >> 2;                       // complete 4-column groups
                    const float4* vr4 = reinterpret_cast<const float4*>(vr);
                    int ca_l = 0, cb_l = 0;
                    int g = lane;
                    for (; g + 32 < n_full; g += 64) {
                        const float4 x0 = vr4[g], x1 = vr4[g + 32];
                        ca_l += (x0.x < pa) + (x0.y < pa) + (x0.z < pa) + (x0.w < pa) + (x1.x < pa) + (x1.y < pa) + (x1.z < pa) + (x1.w < pa);
                        cb_l += (x0.x < pb) + (x0.y < pb) + (x0.z < pb) + (x0.w < pb) + (x1.x < pb) + (x1.y < pb) + (x1.z < pb) + (x1.w < pb);
                    }
                    for (; g < n_full; g += 32) {
                        const float4 x0 = vr4[g];
                        ca_l += (x0.x < pa) + (x0.y < pa) + (x0.z < pa) + (x0.w < pa);
                        cb_l += (x0.x < pb) + (x0.y < pb) + (x0.z < pb) + (x0.w < pb);
                    }
                    if (lane < (NB & 3)) { const float x = vr[4 * n_full + lane]; ca_l += (x < pa); cb_l += (x < pb); }
                    const int cab_l = cb_l - ca_l;
                    const int pk = __reduce_add_sync(0xffffffffu, ca_l | (cab_l << 16));
                    c_a = pk & 0xffff; c_ab = pk >> 16;
                    ok = (c_a <= pr.lo) && (c_a + c_ab >= pr.hi + 1) && (c_ab <= BAND_LMAX);
                    if (!ok) continue;
                    int incl = cab_l;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const int t_ = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t_; }
                    // pass B: every lane writes its in-band elements at its private offset
                    int off = incl - cab_l;
                    for (g = lane; g < n_full; g += 32) {
                        const float4 x0 = vr4[g];
                        const float xs[4] = {x0.x, x0.y, x0.z, x0.w};
                        if (fminf(fminf(x0.x, x0.y), fminf(x0.z, x0.w)) < pb) {
#pragma unroll
                            for (int q = 0; q < 4; ++q)
                                if (xs[q] >= pa && xs[q] < pb) { lidx[off] = 4 * g + q; lval[off] = xs[q]; ++off; }
                        }
                    }
                    if (lane < (NB & 3)) {
                        const float x = vr[4 * n_full + lane];
                        if (x >= pa && x < pb) { lidx[off] = 4 * n_full + lane; lval[off] = x; ++off; }
                    }
                    __syncwarp();
                }
                if (prof) tc2 = clock64();
                float w_lo = 0.f, w_hi = 0.f;
                float lv[BAND_LPL];
                if (ok) {
                    // (d) shrink [w_lo, w_hi) by value bisection inside the list while it keeps both wanted order
                    //     statistics (ranks kk and kk + need2 - 1 of the list), until it holds few elements
                    float vmin = INFINITY, vmax = -INFINITY;
#pragma unroll
                    for (int q = 0; q < BAND_LPL; ++q) {
                        const int e = q * 32 + lane;
                        lv[q] = (e < c_ab) ? lval[e] : INFINITY;
                        if (e < c_ab) { vmin = fminf(vmin, lv[q]); vmax = fmaxf(vmax, lv[q]); }
                    }
                    for (int o = 16; o > 0; o >>= 1) {
                        vmin = fminf(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
                        vmax = fmaxf(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
                    }
                    const int kk = pr.lo - c_a;                     // rank of the lower wanted order statistic inside the list
                    w_lo = vmin; w_hi = INFINITY;                   // invariant: #(list < w_lo) <= kk < #(list < w_hi)
                    int c_lo = 0, c_hi = c_ab;
                    float b_hi = vmax;                              // finite upper end used for the midpoint
                    for (int it = 0; it < 24 && c_hi - c_lo > 3; ++it) {
                        const float mid = 0.5f * (w_lo + b_hi);
                        if (!(mid > w_lo) || !(mid < b_hi)) break;
                        int cm = 0;
#pragma unroll
                        for (int q = 0; q < BAND_LPL; ++q) cm += (lv[q] < mid);
                        cm = __reduce_add_sync(0xffffffffu, cm);
                        if (cm <= kk) { w_lo = mid; c_lo = cm; }
                        else { w_hi = mid; b_hi = mid; c_hi = cm; }
                    }
                    // the upper wanted order statistic (rank kk + 1) is inside the window or is the smallest element >= w_hi
                    if (pr.hi != pr.lo && c_hi < kk + 2) {
                        float nx = INFINITY;
#pragma unroll
                        for (int q = 0; q < BAND_LPL; ++q) if (lv[q] >= w_hi) nx = fminf(nx, lv[q]);
                        for (int o = 16; o > 0; o >>= 1) nx = fminf(nx, __shfl_xor_sync(0xffffffffu, nx, o));
                        w_hi = nx;                                  // candidates up to and including nx (+ 2E)
                    }
                    // the candidate window (+- 2E) must lie strictly inside the list's value range [pa, pb)
                    const float hi_edge = (w_hi == INFINITY) ? vmax : w_hi;
                    ok = (w_lo - 2.0f * E > pa || pa == -INFINITY) && (hi_edge + 2.0f * E < pb || pb == INFINITY);
                }
                if (prof) tc3 = clock64();
                if (ok) {
                    // (e) candidates = list elements in [w_lo - 2E, w_hi + 2E]; every exact order statistic is among them
                    const float c_min = w_lo - 2.0f * E;
                    const float c_max = (w_hi == INFINITY) ? INFINITY : w_hi + 2.0f * E;
                    int n_below = c_a, n_c = 0;
#pragma unroll
                    for (int q = 0; q < BAND_LPL; ++q) {
                        const int e = q * 32 + lane;
                        const bool valid = e < c_ab;
                        const bool below = valid && lv[q] < c_min;
                        const bool cand = valid && !below && lv[q] <= c_max;
                        const int jj = valid ? lidx[e] : 0;
                        n_below += __popc(__ballot_sync(0xffffffffu, below));
                        const unsigned mc = __ballot_sync(0xffffffffu, cand);
                        __syncwarp();
                        // compact candidate columns to the front of lidx (slot <= e: nothing unread is overwritten)
                        if (cand) { const int slot = n_c + __popc(mc & ((1u << lane) - 1u)); if (slot < 32) lidx[slot] = jj; }
                        n_c += __popc(mc);
                        __syncwarp();
                    }
                    const int kx = pr.lo - n_below;
                    if (n_c > 32 || kx < 0 || kx + (pr.hi - pr.lo) >= n_c) ok = false;
                    else {
                        if (lane < n_c) {           // pull the candidates' 9 frames (432 B) towards L1 while the A row is staged
                            const float* bp = frames + (b_f0 + lidx[lane]) * ACOSS_NBINS;
#pragma unroll
                            for (int q = 0; q < 4; ++q) asm volatile("prefetch.global.L1 [%0];" ::"l"(bp + q * 32));
                        }
                        // exact values of all candidates at once, one per lane, in the ORACLE's element order: the A row
                        // goes to lval (108 floats, un-scaled, oracle order); bin c of the oracle is bin (st + c) mod 12 of
                        // both staged operands (see the staging rotation above); lane l walks the 108-term chain of its
                        // candidate straight from the frame store.
                        const int st = side_a_is_query ? rot_a : 0;
                        for (int e = lane; e < 108; e += 32) {
                            const int s_ = e / ACOSS_NBINS, c = e - s_ * ACOSS_NBINS;
                            int cb = st + c; if (cb >= ACOSS_NBINS) cb -= ACOSS_NBINS;
                            lval[e] = -0.5f * Afr[(r + s_) * BAND_AFS + cb];       // exact: staged as -2 * a
                        }
                        __syncwarp();
                        int cbs[ACOSS_NBINS];
#pragma unroll
                        for (int c = 0; c < ACOSS_NBINS; ++c) { int cb = st + c; if (cb >= ACOSS_NBINS) cb -= ACOSS_NBINS; cbs[c] = cb; }
                        float dex = INFINITY;
                        if (lane < n_c) {
                            const int jj = lidx[lane];
                            const float* bp = frames + (b_f0 + jj) * ACOSS_NBINS;
                            const float4* as = reinterpret_cast<const float4*>(lval);
                            float acc = 0.0f;
#pragma unroll 3
                            for (int s_ = 0; s_ < 9; ++s_) {
                                const float4 a0 = as[3 * s_], a1 = as[3 * s_ + 1], a2 = as[3 * s_ + 2];
                                const float* bs = bp + s_ * ACOSS_NBINS;
                                acc = xadd(acc, xmul(a0.x, __ldg(bs + cbs[0])));
                                acc = xadd(acc, xmul(a0.y, __ldg(bs + cbs[1])));
                                acc = xadd(acc, xmul(a0.z, __ldg(bs + cbs[2])));
                                acc = xadd(acc, xmul(a0.w, __ldg(bs + cbs[3])));
                                acc = xadd(acc, xmul(a1.x, __ldg(bs + cbs[4])));
                                acc = xadd(acc, xmul(a1.y, __ldg(bs + cbs[5])));
                                acc = xadd(acc, xmul(a1.z, __ldg(bs + cbs[6])));
                                acc = xadd(acc, xmul(a1.w, __ldg(bs + cbs[7])));
                                acc = xadd(acc, xmul(a2.x, __ldg(bs + cbs[8])));
                                acc = xadd(acc, xmul(a2.y, __ldg(bs + cbs[9])));
                                acc = xadd(acc, xmul(a2.z, __ldg(bs + cbs[10])));
                                acc = xadd(acc, xmul(a2.w, __ldg(bs + cbs[11])));
                            }
                            const float sa = xxA[r], sb = yyB[jj];
                            const float xx = side_a_is_query ? sa : sb, yy = side_a_is_query ? sb : sa;
                            float item = xadd(xsub(xx, xmul(2.0f, acc)), yy);
                            if (item < 0.0f) item = 0.0f;
                            dex = __fsqrt_rn(item);
                            ++n_exact;
                        }
                        __syncwarp();
                        if (prof) tc4 = clock64();
                        // rank by comparison through shuffles (ties by lane), pick ranks kx and kx+1
                        int rank = 0;
                        for (int b = 0; b < n_c; ++b) {
                            const float db = __shfl_sync(0xffffffffu, dex, b);
                            rank += (db < dex) || (db == dex && b < lane);
                        }
                        const unsigned m0 = __ballot_sync(0xffffffffu, lane < n_c && rank == kx);
                        const unsigned m1 = __ballot_sync(0xffffffffu, lane < n_c && rank == kx + 1);
                        v_lo = __shfl_sync(0xffffffffu, dex, __ffs(m0) - 1);
                        v_hi = m1 ? __shfl_sync(0xffffffffu, dex, __ffs(m1) - 1) : v_lo;
                        if (prof) {
                            const long long tc5 = clock64();
                            atomicAdd(&counters[8], (unsigned long long)(tc1 - tc0));
                            atomicAdd(&counters[9], (unsigned long long)(tc2 - tc1));
                            atomicAdd(&counters[10], (unsigned long long)(tc3 - tc2));
                            atomicAdd(&counters[11], (unsigned long long)(tc4 - tc3));
                            atomicAdd(&counters[12], (unsigned long long)(tc5 - tc4));
                            atomicAdd(&counters[13], 1ull);
                        }
                    }
                }
                if (!ok) exact_row = true;
            }
            if (exact_row) {
                // fallback (tiny rows, heavy ties, list overflow): the whole row exactly, as in kernels_crp.cuh
                for (int j = lane; j < NB; j += 32) {
                    vr[j] = side_a_is_query ? exact_dist(frames, pd, rot, selfdot, u_or_v, j)
                                            : exact_dist(frames, pd, rot, selfdot, j, u_or_v);
                }
                n_exact += (NB + 31 - lane) / 32;
                __syncwarp();
                v_lo = warp_select_pair(vr, NB, pr.lo, lane, v_hi);
                if (lane == 0) { flag[r] = 1; if (!tiny) atomicAdd(&counters[1], 1ull); }
            }
            const float t = percentile_combine(pr, v_lo, v_hi);
            if (lane == 0) { thA[r] = t; thr_out[tk.row0 + r] = t; }
        }
        __syncthreads();
        if (tid == 0) {
            ph3 = clock64();
            atomicAdd(&counters[2], (unsigned long long)(ph1 - ph0));
            atomicAdd(&counters[3], (unsigned long long)(ph2 - ph1));
            atomicAdd(&counters[4], (unsigned long long)(ph3 - ph2));
            atomicAdd(&counters[7], 1ull);
        }
        if (!side_a_is_query) continue;

        // ---- pass 2: binarise.  Per column the band edges of T_r = thr_r^2 go to shared memory once
        //      (lo in yyB, hi in Gs); per cell: sure-in  <=> rho~ < min(lo_q, lo_r[j]),
        //      sure-out <=> rho~ > min(hi_q, hi_r[j]); anything else is decided exactly afterwards.
        const float* thB = thr + pd.thr_r_off;
        float* lo_r = yyB;
        float* hi_r = Gs;
        int* amb = reinterpret_cast<int*>(Gs + NBpad);
        int* amb_cnt = reinterpret_cast<int*>(red + 24);
        const int amb_cap = BAND_GR * BAND_TW - NBpad;
        if (tid == 0) *amb_cnt = 0;
        for (int j = tid; j < NBpad; j += BAND_THREADS) {
            float lo = -INFINITY, hi = -INFINITY;          // padding columns: never "in", always "out"
            if (j < NB) {
                const float tb = __ldg(thB + j);
                const float Tr = xmul(tb, tb);
                const float mr = E + Tr * 4.76837158203125e-07f;        // E + 8uT
                lo = Tr - mr; hi = Tr + mr;
            }
            lo_r[j] = lo; hi_r[j] = hi;
        }
        __syncthreads();
        const int words = pd.pitch >> 2;
        // per-row band edges of T_q = thr_q^2 (lo in red[0..15], hi in thA-adjacent scratch: reuse xxA, no longer needed)
        if (tid < BAND_RS) {
            const float ta = thA[tid];
            const float Tq = xmul(ta, ta);
            const float mq = E + Tq * 4.76837158203125e-07f;
            red[tid] = Tq - mq;
            xxA[tid] = Tq + mq;
        }
        __syncthreads();
        // thread = one 4-column group (one output word per row); the column band edges stay in registers while
        // the thread walks down the strip's rows; consecutive threads write consecutive words of a CRP row
        int any_flag = 0;
#pragma unroll
        for (int r = 0; r < BAND_RS; ++r) any_flag |= flag[r];
        const uint32_t zero_word = (CRP_DELTA_ZERO) | (CRP_DELTA_ZERO << 8) | (CRP_DELTA_ZERO << 16) | (CRP_DELTA_ZERO << 24);
        const uint32_t flip = CRP_DELTA_ZERO ^ CRP_DELTA_ONE;
        for (int w = tid; w < words; w += BAND_THREADS) {
            const int j0 = w * 4;
            uint32_t* out = reinterpret_cast<uint32_t*>(crp + pd.crp_off + (int64_t)tk.row0 * pd.pitch) + w;
            const int wstride = pd.pitch >> 2;
            if (j0 >= NBpad) {
                for (int r = 0; r < tk.nrows; ++r) out[(int64_t)r * wstride] = zero_word;
                continue;
            }
            const float4 l4 = *reinterpret_cast<const float4*>(lo_r + j0);
            const float4 h4 = *reinterpret_cast<const float4*>(hi_r + j0);
            const bool tail = (j0 + 3 >= NB);
            for (int r = 0; r < tk.nrows; ++r) {
                const float4 x4 = *reinterpret_cast<const float4*>(V + r * NBpad + j0);
                uint32_t word = zero_word;
                if (any_flag && flag[r] != 0) {
                    const float ta = thA[r];
                    const float xs[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        const int j = j0 + b;
                        if (j < NB && xs[b] <= ta && xs[b] <= __ldg(thB + j)) word ^= flip << (8 * b);
                    }
                } else {
                    const float lo_q = red[r], hi_q = xxA[r];
                    const bool i0 = x4.x < fminf(lo_q, l4.x), i1 = x4.y < fminf(lo_q, l4.y);
                    const bool i2 = x4.z < fminf(lo_q, l4.z), i3 = x4.w < fminf(lo_q, l4.w);
                    const bool o0 = x4.x > fminf(hi_q, h4.x), o1 = x4.y > fminf(hi_q, h4.y);
                    const bool o2 = x4.z > fminf(hi_q, h4.z), o3 = x4.w > fminf(hi_q, h4.w);
                    word ^= (i0 ? flip : 0u) | (i1 ? flip << 8 : 0u) | (i2 ? flip << 16 : 0u) | (i3 ? flip << 24 : 0u);
                    if (!((i0 | o0) & (i1 | o1) & (i2 | o2) & (i3 | o3))) {
                        // at least one cell is neither surely in nor surely out (padding columns are always "out")
                        const bool am[4] = {!(i0 | o0), !(i1 | o1), !(i2 | o2), !(i3 | o3)};
#pragma unroll
                        for (int b = 0; b < 4; ++b)
                            if (am[b] && (!tail || j0 + b < NB)) {
                                const int slot = atomicAdd(amb_cnt, 1);
                                if (slot < amb_cap) amb[slot] = (r << 24) | (j0 + b);
                            }
                    }
                }
                out[(int64_t)r * wstride] = word;
            }
        }
        __syncthreads();
        {
            const int na = *amb_cnt;
            if (na > amb_cap) {
                // pathological (massive exact ties): decide every non-exact row of the strip exactly
                for (int r = 0; r < tk.nrows; ++r) {
                    if (flag[r]) continue;
                    uint8_t* out = crp + pd.crp_off + (int64_t)(tk.row0 + r) * pd.pitch;
                    for (int j = tid; j < NB; j += BAND_THREADS) {
                        const float dv = exact_dist(frames, pd, rot, selfdot, tk.row0 + r, j);
                        ++n_exact;
                        out[j] = (dv <= thA[r] && dv <= thB[j]) ? (uint8_t)CRP_DELTA_ONE : (uint8_t)CRP_DELTA_ZERO;
                    }
                }
            } else {
                for (int e = tid; e < na; e += BAND_THREADS) {
                    const int code = amb[e];
                    const int r = code >> 24, j = code & 0xFFFFFF;
                    const float dv = exact_dist(frames, pd, rot, selfdot, tk.row0 + r, j);
                    ++n_exact;
                    if (dv <= thA[r] && dv <= thB[j])
                        crp[pd.crp_off + (int64_t)(tk.row0 + r) * pd.pitch + j] = (uint8_t)CRP_DELTA_ONE;
                }
            }
        }
        __syncthreads();
        if (tid == 0) atomicAdd(&counters[5], (unsigned long long)(clock64() - ph3));
    }
    for (int o = 16; o > 0; o >>= 1) n_exact += __shfl_xor_sync(0xffffffffu, n_exact, o);
    if (lane == 0 && n_exact) atomicAdd(&counters[0], n_exact);
    if (TC) {
        tc_fence_before();
        __syncthreads();
        if (warp == 0) tmem_free(tmem_base);
    }
}
