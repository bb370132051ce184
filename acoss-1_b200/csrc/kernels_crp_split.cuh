// kernels_crp_split.cuh -- the tensor-core CRP as a PHASE-SPLIT pipeline (sm_100a, crp_mode 5; the default for 512..3072-frame
// tracks).  Same arithmetic, the same error bound and the same bit-exact result as crp_strip_gemm_kernel
// (kernels_crp_gemm.cuh), but every phase of a strip is its own kernel running over ALL strips of a chunk:
//
//   crp_pivot_kernel    sample GEMM (512 spread columns, TMA-fetched per-track operand) -> per-row pivots pa <= pb
//   crp_stream_kernel<0> count pass: streams every 128 x 256 tile once; c_a = #(rho~ < pa), elements of [pa, pb) -> row lists
//   crp_select_kernel   lists -> candidates within 2E of the wanted order statistics -> exact 108-term chains -> thresholds
//   crp_edges_kernel    (pass 2) per-column band edges of the column thresholds
//   crp_stream_kernel<1> binarise pass: streams every tile again, writes the CRP bytes, decides ambiguous cells exactly
//
// Why split: in the one-kernel form 48 % of a strip's time was threshold selection with no MMA in flight (one 640-thread
// CTA per SM, 192 KB of shared memory shared by all phases).  Here the two streaming kernels do nothing but stream: a
// dedicated TMA warp fetches the side-B operand planes -- PRECOMPUTED per track at upload in their final shared-memory
// layout (hi / lo TF32 planes + norm planes, sp_operand kernels) -- with cp.async.bulk onto mbarriers, a dedicated warp
// issues the tcgen05 MMAs, the 16 epilogue warps only read TMEM; the side-A planes are double-buffered so that the MMA
// stream runs across strip boundaries without a bubble.  The selection kernel runs at four CTAs per SM (latency-bound list
// and chain loads get 2x the warps), the pivot kernel keeps the whole SM for its histogram.
//
// Scratch per strip of a chunk (global memory, indexed by the strip's position in the chunk):
//   piv   float2[128]            (pa, pb) per strip row
//   cnt4  int[128][4]            #(rho~ < pa) counted by each of the row's four owner threads
//   lcnt  int[128][4]            list entries written by each owner
//   lists float2[128][4][SUBCAP] (rho~, column) -- a row's four sub-lists are contiguous: the selection kernel reads them
//                                with one coalesced warp load per 32 slots
#pragma once
#include "common.cuh"
#include "kernels_crp.cuh"
#include "kernels_crp_tc.cuh"
#include "kernels_crp_band.cuh"
#include "kernels_crp_gemm.cuh"

#define SP_BLOCK (GM_THREADS + 128)          // 16 epilogue warps + MMA warp + TMA warp (+ 2 idle warps of the quad)
#define SP_NB_RING 3                         // side-B operand buffers
#define SP_EDGE_SLOTS 6                      // column-edge ring (see crp_stream_kernel: >= 5 slots make the reuse safe)
#define SP_AMB_CAP 8192                      // ambiguous cells per strip before the exact-everything path
#define SP_LIST_STRIDE (4 * GM_SUBCAP)       // float2 entries per row
// crp_stream_kernel<2>: count pass on the hi.hi product alone (14 MMAs per tile instead of 41), see SP_EPS_REL_HH; context option "count_hh"
// Error bound of the hi.hi-only radicand against the oracle's, relative to xx + yy: TF32 rounding of both operands (unit roundoff
// 2^-11 each) gives |x_h.y_h - x.y| <= (2^-10 + 2^-22) sum |x_k y_k| <= 2^-11 (xx + yy); the radicand carries -2 x.y: 2^-10 (xx + yy);
// plus the three-product bound GM_EPS_REL = 2^-13 for everything else (norms, accumulation, the oracle's own rounding).  1.25 x 2^-10.
#define SP_EPS_REL_HH 1.220703125e-03f
#define SP_SEL_THREADS 256
#define SP_SEL_WARPS (SP_SEL_THREADS / 32)
#define SP_SEL_LPL (SP_LIST_STRIDE / 32)     // list entries per lane in the selection kernel

// ---- per-track operand store (built once at upload, m = 9, tau = 1) -------------------------------------------------
// planes:  [6][P][16 B], P = 256 * n_tiles + 8 entries, entry e = frame e, zero past the end of the track:
//          planes 3 4 5 = H0 H1 H2: bins 4 c3 .. 4 c3 + 3 as TF32 hi parts (the hi.hi product, 14 tf32 MMAs per tile)
//          planes 0 1 2 = the CROSS-PRODUCT operands in bf16 (x_hi y_lo + x_lo y_hi needs 8 bits of either factor: the error
//          this adds is 2^-18 (xx + yy), 3 % of E), laid out so that 14 kind::f16 MMAs (K = 16) replace 27 tf32 ones:
//            plane 0 = bf16(y_lo) bins 0..7, plane 1 = bf16(y_hi) bins 0..7      (K = [x_hi 0..7 | x_lo 0..7] . [y_lo | y_hi], one MMA per delay)
//            plane 2 = [bf16(y_lo) bins 8..11 | bf16(y_hi) bins 8..11]           (against [x_hi 8..11 | x_lo 8..11], two delays per MMA)
// norms:   [12 rotations][256 * n_tiles][16 B] = (1, 1, |y|^2 hi, |y|^2 lo) from the exact self-dot table of that rotation;
//          padding columns get (1, 1, GM_BIG, 0): rho~ huge, never below a pivot or an edge
// sdmax:   [12] max of the self-dot table per rotation (the strip kernels' error bound uses a per-pair E)
__host__ __device__ __forceinline__ int sp_tiles(int emb) { return emb > 0 ? (emb + GM_TN - 1) / GM_TN : 0; }

// eight floats -> eight bf16 (round to nearest even) in one 16-byte operand chunk, K order = argument order
__device__ __forceinline__ uint4 sp_bf16x8(float a0, float a1, float a2, float a3, float a4, float a5, float a6, float a7)
{
    const __nv_bfloat162 p0 = __floats2bfloat162_rn(a0, a1), p1 = __floats2bfloat162_rn(a2, a3);
    const __nv_bfloat162 p2 = __floats2bfloat162_rn(a4, a5), p3 = __floats2bfloat162_rn(a6, a7);
    uint4 r;
    r.x = *reinterpret_cast<const uint32_t*>(&p0); r.y = *reinterpret_cast<const uint32_t*>(&p1);
    r.z = *reinterpret_cast<const uint32_t*>(&p2); r.w = *reinterpret_cast<const uint32_t*>(&p3);
    return r;
}

__global__ void sp_planes_kernel(const float* __restrict__ frames, const int64_t* __restrict__ offs, const int32_t* __restrict__ emb_len,
                                 const int64_t* __restrict__ op_off, unsigned char* __restrict__ opb, int track0)
{
    const int track = track0 + blockIdx.y;
    const int nt = sp_tiles(emb_len[track]);
    if (nt == 0) return;
    const int P = GM_TN * nt + 8;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= P) return;
    const int n = (int)(offs[track + 1] - offs[track]);
    float4 x[3];
    if (e < n) {
        const float4* fp = reinterpret_cast<const float4*>(frames + (offs[track] + e) * ACOSS_NBINS);
        x[0] = fp[0]; x[1] = fp[1]; x[2] = fp[2];
    } else {
        x[0] = x[1] = x[2] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    unsigned char* base = opb + op_off[2 * track];
    float h[12], l[12];
#pragma unroll
    for (int c3 = 0; c3 < 3; ++c3) {
        float4 hi, lo;
        tf32_split(x[c3].x, hi.x, lo.x); tf32_split(x[c3].y, hi.y, lo.y);
        tf32_split(x[c3].z, hi.z, lo.z); tf32_split(x[c3].w, hi.w, lo.w);
        *reinterpret_cast<float4*>(base + ((size_t)(3 + c3) * P + e) * 16) = hi;
        h[4 * c3] = hi.x; h[4 * c3 + 1] = hi.y; h[4 * c3 + 2] = hi.z; h[4 * c3 + 3] = hi.w;
        l[4 * c3] = lo.x; l[4 * c3 + 1] = lo.y; l[4 * c3 + 2] = lo.z; l[4 * c3 + 3] = lo.w;
    }
    *reinterpret_cast<uint4*>(base + ((size_t)0 * P + e) * 16) = sp_bf16x8(l[0], l[1], l[2], l[3], l[4], l[5], l[6], l[7]);
    *reinterpret_cast<uint4*>(base + ((size_t)1 * P + e) * 16) = sp_bf16x8(h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
    *reinterpret_cast<uint4*>(base + ((size_t)2 * P + e) * 16) = sp_bf16x8(l[8], l[9], l[10], l[11], h[8], h[9], h[10], h[11]);
}

__global__ void sp_norms_kernel(const int32_t* __restrict__ emb_len, const int64_t* __restrict__ sd_off, const float* __restrict__ selfdot,
                                const int64_t* __restrict__ op_off, unsigned char* __restrict__ opn, int track0)
{
    const int track = track0 + blockIdx.y;
    const int M = emb_len[track];
    const int nt = sp_tiles(M);
    if (nt == 0) return;
    const int W = GM_TN * nt;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= ACOSS_NBINS * W) return;
    const int rot = idx / W, j = idx - rot * W;
    float4 nv = make_float4(1.0f, 1.0f, GM_BIG, 0.f);
    if (j < M) tf32_split(selfdot[sd_off[track] + (int64_t)rot * M + j], nv.z, nv.w);
    *reinterpret_cast<float4*>(opn + op_off[2 * track + 1] + ((size_t)rot * W + j) * 16) = nv;
}

__global__ void sp_sdmax_kernel(const int32_t* __restrict__ emb_len, const int64_t* __restrict__ sd_off, const float* __restrict__ selfdot,
                                float* __restrict__ sdmax)
{
    const int track = blockIdx.x, M = emb_len[track];
    __shared__ float red[8];
    for (int rot = 0; rot < ACOSS_NBINS; ++rot) {
        float mx = 0.0f;
        for (int j = threadIdx.x; j < M; j += blockDim.x) mx = fmaxf(mx, selfdot[sd_off[track] + (int64_t)rot * M + j]);
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int w = 1; w < (int)(blockDim.x >> 5); ++w) mx = fmaxf(mx, red[w]);
            sdmax[track * ACOSS_NBINS + rot] = mx;
        }
        __syncthreads();
    }
}

// per-pair error bound and "unbounded below" pivot: functions of the two tracks only, so that all four kernels agree
struct SpPairBound { float E, plow, E_hh; };      // E: three-product streams; E_hh: the hi.hi-only count pass
__device__ __forceinline__ SpPairBound sp_pair_bound(const PairDesc& pd, const float* __restrict__ sdmax)
{
    const int rot = pd.oti % ACOSS_NBINS;
    const float s = __ldg(sdmax + pd.q * ACOSS_NBINS) + __ldg(sdmax + pd.r * ACOSS_NBINS + rot);
    SpPairBound b;
    b.E = GM_EPS_REL * s;
    b.E_hh = SP_EPS_REL_HH * s;
    b.plow = -(4.0f * s + 1.0e-30f);
    return b;
}

// stage the side-A planes of a strip (entry e = frame row0 + e of -2 x the rotated bins: planes 0 1 2 = TF32 hi parts per chunk of
// four bins; planes 3 4 5 = the bf16 cross-product operands [hi 0..7], [lo 0..7], [hi 8..11 | lo 8..11], see the operand store;
// norm plane), optionally only the hi planes (the sample GEMM)
__device__ __forceinline__ void sp_stage_a(unsigned char* Apl, const float* __restrict__ fr_a, const float* __restrict__ sdA,
                                           int row0, int nrows, int na_frames, int rot_a, int tid, bool hi_only)
{
    if (hi_only) {
        for (int e = tid; e < (GM_ROWS + 8) * 3; e += GM_THREADS) {
            const int fe = e / 3, c3 = e - 3 * fe;
            const int f = row0 + fe;
            float4 hi = make_float4(0.f, 0.f, 0.f, 0.f), lo = hi;
            if (f < na_frames) {
                const float* fp = fr_a + (int64_t)f * ACOSS_NBINS;
                tf32_split(-2.0f * fp[rot_src(4 * c3 + 0, rot_a)], hi.x, lo.x);
                tf32_split(-2.0f * fp[rot_src(4 * c3 + 1, rot_a)], hi.y, lo.y);
                tf32_split(-2.0f * fp[rot_src(4 * c3 + 2, rot_a)], hi.z, lo.z);
                tf32_split(-2.0f * fp[rot_src(4 * c3 + 3, rot_a)], hi.w, lo.w);
            }
            *reinterpret_cast<float4*>(Apl + c3 * GM_APLANE + fe * 16) = hi;
        }
    } else {
        for (int fe = tid; fe < GM_ROWS + 8; fe += GM_THREADS) {
            const int f = row0 + fe;
            float h[12], l[12];
#pragma unroll
            for (int c = 0; c < 12; ++c) { h[c] = 0.f; l[c] = 0.f; }
            if (f < na_frames) {
                const float* fp = fr_a + (int64_t)f * ACOSS_NBINS;
#pragma unroll
                for (int c = 0; c < 12; ++c) tf32_split(-2.0f * fp[rot_src(c, rot_a)], h[c], l[c]);
            }
#pragma unroll
            for (int c3 = 0; c3 < 3; ++c3)
                *reinterpret_cast<float4*>(Apl + c3 * GM_APLANE + fe * 16) = make_float4(h[4 * c3], h[4 * c3 + 1], h[4 * c3 + 2], h[4 * c3 + 3]);
            *reinterpret_cast<uint4*>(Apl + 3 * GM_APLANE + fe * 16) = sp_bf16x8(h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
            *reinterpret_cast<uint4*>(Apl + 4 * GM_APLANE + fe * 16) = sp_bf16x8(l[0], l[1], l[2], l[3], l[4], l[5], l[6], l[7]);
            *reinterpret_cast<uint4*>(Apl + 5 * GM_APLANE + fe * 16) = sp_bf16x8(h[8], h[9], h[10], h[11], l[8], l[9], l[10], l[11]);
        }
    }
    if (tid < GM_ROWS) {
        float4 nv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tid < nrows) { tf32_split(sdA[row0 + tid], nv.x, nv.y); nv.z = 1.0f; nv.w = 1.0f; }
        *reinterpret_cast<float4*>(Apl + 6 * GM_APLANE + tid * 16) = nv;
    }
}

// the 14 MMAs of the hi.hi product of one 128 x 256 tile (operand layouts of gm_issue_tile: A planes H0 H1 H2 L0 L1 L2 NRM,
// B planes L0 L1 L2 H0 H1 H2 NRM)
__device__ __forceinline__ void sp_issue_tile_hh(uint32_t d_tmem, uint32_t a0, uint32_t b0)
{
    const uint32_t al = (a0 & 0x3FFFFu) >> 4, bl = (b0 & 0x3FFFFu) >> 4;
    uint32_t acc = 0;
#pragma unroll
    for (int c3 = 0; c3 < 3; ++c3)
#pragma unroll
        for (int kp = 0; kp < 4; ++kp) {
            gm_mma(d_tmem, GM_DLO(al, c3 * GM_APLANE + 32 * kp, 16), GM_DLO(bl, (3 + c3) * GM_BPLANE + 32 * kp, 16), acc);
            acc = 1;
        }
    gm_mma(d_tmem, GM_DLO(al, 128, GM_APLANE), GM_DLO(bl, 3 * GM_BPLANE + 128, GM_BPLANE), 1);                                        // (H0,H0) (H1,H1), offset 8
    gm_mma(d_tmem, GM_DLO(al, 2 * GM_APLANE + 128, 4 * GM_APLANE - 128), GM_DLO(bl, 5 * GM_BPLANE + 128, GM_BPLANE - 128), 1);        // (H2,H2) norms
}

// The 28 MMAs of one 128 x 256 tile of the streaming passes: the hi.hi product in TF32 (14) + both cross products in bf16 (14).
// Measured: an MMA whose two K chunks are ADJACENT entries of one plane (LBO = 16 B: the Hankel case, the chunks of neighbouring
// rows overlap in shared memory) takes ~153 clk, one whose chunks come from two planes ~260 clk -- the operand fetch, not the
// arithmetic, sets the pace.  So the cross products pair consecutive delays wherever they can:
//   bins 0..7:  x_hi . y_lo (A plane 3, B plane 0) and x_lo . y_hi (A plane 4, B plane 1): two delays per MMA, 4 + 4;
//               delay 8 of both in ONE MMA, K = [x_hi | x_lo] . [y_lo | y_hi] (LBO = one plane)
//   bins 8..11: the [hi | lo] chunk of two consecutive delays per MMA (A plane 5 against B plane 2), 4; delay 8 alone pairs
//               with a chunk of zeros on the A side (`zero_saddr`, 128 x 16 bytes) so that whatever B holds there drops out
__device__ __forceinline__ void sp_issue_tile_mixed(uint32_t d_tmem, uint32_t a0, uint32_t b0, uint32_t zero_saddr)
{
    sp_issue_tile_hh(d_tmem, a0, b0);
    const uint32_t al = (a0 & 0x3FFFFu) >> 4, bl = (b0 & 0x3FFFFu) >> 4;
#pragma unroll
    for (int pl = 0; pl < 3; ++pl)
#pragma unroll
        for (int kp = 0; kp < 4; ++kp)
            gm_mma_bf16(d_tmem, GM_DLO(al, (3 + pl) * GM_APLANE + 32 * kp, 16), GM_DLO(bl, pl * GM_BPLANE + 32 * kp, 16), 1);
    gm_mma_bf16(d_tmem, GM_DLO(al, 3 * GM_APLANE + 128, GM_APLANE), GM_DLO(bl, 128, GM_BPLANE), 1);
    const uint32_t a8 = a0 + 5 * GM_APLANE + 128;                          // delay 8 of plane 5; its K partner is the zero chunk
    gm_mma_bf16(d_tmem, ((a8 & 0x3FFFFu) >> 4) + (((zero_saddr - a8) >> 4) << 16), GM_DLO(bl, 2 * GM_BPLANE + 128, 16), 1);
}

// geometry of one strip task, shared by all kernels
struct SpGeom {
    int rot, NB, na_frames, rot_a, trackB, rotB, n_tiles;
    const float* fr_a; const float* fr_b; const float* sdA; const float* sdB;
};
__device__ __forceinline__ SpGeom sp_geom(const PairDesc& pd, int side_a_is_query, const float* __restrict__ frames,
                                          const float* __restrict__ selfdot)
{
    SpGeom g;
    g.rot = pd.oti % ACOSS_NBINS;
    g.NB = side_a_is_query ? pd.N : pd.M;
    g.na_frames = side_a_is_query ? pd.nq : pd.nr;
    // <roll(ref, rot), query>: the B frames stay unrotated; the A frames are staged pre-rotated: by +rot when A is the
    // reference, by -rot when A is the query
    g.rot_a = side_a_is_query ? (ACOSS_NBINS - g.rot) % ACOSS_NBINS : g.rot;
    g.trackB = side_a_is_query ? pd.r : pd.q;
    g.rotB = side_a_is_query ? g.rot : 0;
    g.n_tiles = (g.NB + GM_TN - 1) / GM_TN;
    g.fr_a = frames + (side_a_is_query ? pd.q_frame_off : pd.r_frame_off) * ACOSS_NBINS;
    g.fr_b = frames + (side_a_is_query ? pd.r_frame_off : pd.q_frame_off) * ACOSS_NBINS;
    g.sdA = selfdot + (side_a_is_query ? pd.q_sd_off : pd.r_sd_off + (int64_t)g.rot * pd.N);
    g.sdB = selfdot + (side_a_is_query ? pd.r_sd_off + (int64_t)g.rot * pd.N : pd.q_sd_off);
    return g;
}

// =====================================================================================================================
// K1: sample GEMM + per-row pivots.  One 512-thread CTA per SM.  Per strip: the A hi planes are staged while the first half
// of the side-B track's precomputed sample operand (115 KB) arrives by one bulk copy; 14 MMAs per half into TMEM; the four
// owner threads of a row bin their 128 samples into private byte histograms (no atomics); a scan over the owners' counters
// gives pa / pb at bin edges (see crp_strip_gemm_kernel for the derivation of every step; the code is the same).
// =====================================================================================================================
#define SP_PIV_SMEM (GM_A_BYTES + GM_R1_BYTES + 2 * GM_THREADS * 4 + 64)

__global__ void __launch_bounds__(GM_THREADS, 1)
crp_pivot_kernel(const PairDesc* __restrict__ descs, const StripTask* __restrict__ tasks, int n_tasks,
                 const float* __restrict__ frames, const float* __restrict__ selfdot, const float* __restrict__ sdmax,
                 const unsigned char* __restrict__ smpop, int side_a_is_query, float qf, float nsigma,
                 float2* __restrict__ piv, unsigned long long* __restrict__ counters)
{
    extern __shared__ __align__(128) unsigned char gsm[];
    unsigned char* Apl = gsm;
    unsigned char* R1 = gsm + GM_A_BYTES;
    float* xf = reinterpret_cast<float*>(R1 + GM_R1_BYTES);      // [2][GM_THREADS] exchange
    int* xc = reinterpret_cast<int*>(xf);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(xf + 2 * GM_THREADS);
    uint64_t* mb_smp = mbar, *mb_ld = mbar + 1;
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(mbar + 2);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int lq = warp & 3, cg = warp >> 2;
    const int myrow = 32 * lq + lane;
    uint32_t psmp = 0, pld = 0;

    if (warp == 0) gm_tmem_alloc(tmem_holder);
    if (tid == 0) {
        mbar_init(mb_smp, 1); mbar_init(mb_ld, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
    const uint32_t my_taddr = tmem_base + ((uint32_t)(32 * lq) << 16) + 16 * cg;
    const uint32_t a_saddr = smem_u32(Apl), r1_saddr = smem_u32(R1);

    for (int task = blockIdx.x; task < n_tasks; task += gridDim.x) {
        const StripTask tk = tasks[task];
        const PairDesc pd = descs[tk.pair];
        const SpGeom g = sp_geom(pd, side_a_is_query, frames, selfdot);
        float2* piv_out = piv + (size_t)task * GM_ROWS;
        if (g.NB < GM_MIN_NB) {          // tiny side B: the selection kernel takes every row exactly
            if (tid < GM_ROWS) piv_out[tid] = make_float2(0.f, 0.f);
            continue;
        }
        const PercentileRank pr = percentile_rank(g.NB, qf);
        const SpPairBound pbd = sp_pair_bound(pd, sdmax);
        const unsigned char* smp_src = smpop + (size_t)g.trackB * GM_SMP_OP_BYTES;
        fence_proxy_async_smem();        // the histogram's generic-proxy accesses of R1 before the bulk copy's async-proxy writes
        __syncthreads();                 // R1 / A planes / TMEM of the previous strip are free
        if (tid == 0) gm_bulk_load(R1, smp_src, GM_SMP_HALF_BYTES, mb_ld);
        sp_stage_a(Apl, g.fr_a, g.sdA, tk.row0, tk.nrows, g.na_frames, g.rot_a, tid, true);
        fence_proxy_async_smem();
        for (int half = 0; half < GM_SHALVES; ++half) {
            if (half > 0 && tid == 0) gm_bulk_load(R1, smp_src + (size_t)half * GM_SMP_HALF_BYTES, GM_SMP_HALF_BYTES, mb_ld);
            mbar_wait(mb_ld, pld); pld ^= 1;
            tc_fence_before();
            __syncthreads();
            if (tid == 0) {
                tc_fence_after();
                gm_issue_sample(tmem_base + half * GM_TN, a_saddr, r1_saddr);
                umma_commit(mb_smp);
            }
            mbar_wait(mb_smp, psmp); psmp ^= 1;     // the operand buffer is rebuilt for the second half
            tc_fence_after();
        }
        // ---- pivots
        static_assert(GM_S == 512 && GM_THREADS == 512, "four owners x 128 samples per row");
        unsigned* hall = reinterpret_cast<unsigned*>(R1);                       // [16 warps][64 words][32 lanes] = 128 KB
        const int xi = myrow * 4;
#pragma unroll
        for (int i = 0; i < 16; ++i) reinterpret_cast<uint4*>(hall + warp * 2048)[i * 32 + lane] = make_uint4(0u, 0u, 0u, 0u);
        float smin = INFINITY, smax = -INFINITY;
#pragma unroll
        for (int b = 0; b < GM_SHALVES; ++b)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float v[32];
                gm_tmem_ld32(my_taddr + 48 * cg + b * GM_TN + 32 * h, v);
#pragma unroll
                for (int q = 0; q < 32; ++q) { smin = fminf(smin, v[q]); smax = fmaxf(smax, v[q]); }
            }
        xf[xi + cg] = smin; xf[GM_THREADS + xi + cg] = smax;
        __syncthreads();
        {
            const float4 a4 = *reinterpret_cast<const float4*>(xf + xi), b4 = *reinterpret_cast<const float4*>(xf + GM_THREADS + xi);
            smin = fminf(fminf(a4.x, a4.y), fminf(a4.z, a4.w));
            smax = fmaxf(fmaxf(b4.x, b4.y), fmaxf(b4.z, b4.w));
        }
        const float hscale = (smax > smin) ? 256.0f / (smax - smin) : 0.0f;
        const float hwidth = (smax - smin) * (1.0f / 256.0f);
        const float hoff = -smin * hscale;
        unsigned* hw = hall + warp * 2048 + lane;                                                // bin b: byte b % 4 of word b / 4
#pragma unroll
        for (int b = 0; b < GM_SHALVES; ++b)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float v[32];
                gm_tmem_ld32(my_taddr + 48 * cg + b * GM_TN + 32 * h, v);
#pragma unroll
                for (int q = 0; q < 32; ++q) {
                    const int bin = min(255, (int)fmaf(v[q], hscale, hoff));
                    // the histogram is private to the thread (bank = lane): the reduction instruction is used for its fire-and-forget
                    // issue, not for atomicity -- a byte read-modify-write put a load-add-store chain on every one of the 128 samples
                    atomicAdd(hw + (bin >> 2) * 32, 1u << (8 * (bin & 3)));
                }
            }
        tc_fence_before();
        __syncthreads();
        const float frac = ((float)pr.lo + 0.5f) / (float)g.NB;
        const float rs = frac * (float)GM_S;
        const float sg = sqrtf(fmaxf(rs * (1.0f - frac), 1.0f));
        const int r_lo = (int)floorf(rs - nsigma * sg) - 1;          // sample ranks the bracket must enclose
        const int r_hi = (int)ceilf(rs + nsigma * sg) + 1;
        unsigned lo[16], hi[16];
        int tot = 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            unsigned l = 0, h = 0;
#pragma unroll
            for (int o = 0; o < 4; ++o) {
                const unsigned x = hall[(4 * o + lq) * 2048 + (16 * cg + k) * 32 + lane];
                l += x & 0x00ff00ffu; h += (x >> 8) & 0x00ff00ffu;
            }
            lo[k] = l; hi[k] = h;
            tot += (int)((l & 0xffffu) + (l >> 16) + (h & 0xffffu) + (h >> 16));
        }
        xc[xi + cg] = tot;           // the min / max exchange was read before the barrier above
        __syncthreads();
        int ba = -1, bb = 256;
        {
            const int4 c = *reinterpret_cast<const int4*>(xc + xi);
            const int cc[4] = {c.x, c.y, c.z, c.w};
            int run = 0;
#pragma unroll
            for (int q = 0; q < 4; ++q) if (q < cg) run += cc[q];
            // pa = lower edge of the last bin b with #(samples in bins < b) <= r_lo;
            // pb = upper edge of the first bin b with #(samples in bins <= b) > r_hi
#pragma unroll
            for (int k = 0; k < 16; ++k)
#pragma unroll
                for (int f = 0; f < 4; ++f) {
                    const int bin = 64 * cg + 4 * k + f;
                    const int cnt = (int)(f == 0 ? (lo[k] & 0xffffu) : f == 1 ? (hi[k] & 0xffffu) : f == 2 ? (lo[k] >> 16) : (hi[k] >> 16));
                    if (run <= r_lo) ba = bin;
                    run += cnt;
                    if (run > r_hi && bb == 256) bb = bin;
                }
        }
        __syncthreads();                  // xc is rewritten
        xc[xi + cg] = ba; xc[GM_THREADS + xi + cg] = bb;
        __syncthreads();
        if (cg == 0) {
            const int4 a4 = *reinterpret_cast<const int4*>(xc + xi), b4 = *reinterpret_cast<const int4*>(xc + GM_THREADS + xi);
            ba = max(max(a4.x, a4.y), max(a4.z, a4.w));
            bb = min(min(b4.x, b4.y), min(b4.z, b4.w));
            const float pa = (r_lo >= 0 && ba > 0) ? fmaxf(smin + (float)ba * hwidth, pbd.plow) : pbd.plow;
            const float pb = (r_hi < GM_S && bb < 255) ? fminf(smin + (float)(bb + 1) * hwidth, GM_PMAX) : GM_PMAX;
            piv_out[myrow] = make_float2(pa, pb);
        }
    }
    if (tid == 0 && counters) atomicAdd(&counters[7], 0ull);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) gm_tmem_free(tmem_base);
}

// =====================================================================================================================
// K2 / K4: the streaming kernels.  MODE 0 = count pass, MODE 2 = count pass on hi.hi alone, MODE 1 = binarise pass (pass 2 only).
//   warps 0..15  epilogue: thread (row = 32 lq + lane, column group cg) reads its 4 x 16 accumulator columns of every tile
//   warp 16      MMA issuer: 41 tcgen05.mma per tile from one elected lane, commits to tmem_full[] and b_empty[]
//   warp 17      TMA producer: 7 (8) cp.async.bulk per tile -- six operand planes, the norm plane (and the column edges) --
//                onto b_full[] with expect_tx
// Rings: side-B buffers g % 3, TMEM accumulators g % 2, A planes (strip index) % 2, with g the CTA's running tile count.
// The epilogue warps stage the A planes of strip i + 1 at the start of strip i (its buffer was last read by the MMAs of
// strip i - 1, which every epilogue warp has seen complete), so the MMA stream crosses strip boundaries without waiting.
// Column edges: slot g % 6.  The copy for tile g + 6 is issued after b_empty of tile g + 3, i.e. after the MMAs of tile g + 3
// have completed; those could only start once every warp had released TMEM buffer (g + 1) % 2, i.e. finished tile g + 1,
// hence tile g: the slot is free.
// =====================================================================================================================
template <int MODE> struct SpStreamSmem {
    static constexpr int A_OFF = 0;
    static constexpr int B_OFF = 2 * GM_A_BYTES;
    static constexpr int EDGE_OFF = B_OFF + SP_NB_RING * GM_B_BYTES;
    static constexpr int ARAW_OFF = EDGE_OFF + (MODE == 1 ? SP_EDGE_SLOTS * GM_TN * 4 : 0);
    static constexpr int AMB_OFF = ARAW_OFF + (MODE == 1 ? GM_ARAW_BYTES : 0);
    static constexpr int ROWV_OFF = AMB_OFF + (MODE == 1 ? SP_AMB_CAP * 4 : 0);
    static constexpr int ZERO_OFF = ROWV_OFF + 4 * GM_ROWV_BYTES;           // s_xx s_th s_loq + spare; then 128 x 16 bytes of zeros
    static constexpr int MBAR_OFF = ZERO_OFF + GM_ROWS * 16;
    static constexpr int BYTES = MBAR_OFF + 16 * 8 + 64;
};

template <int MODE>
__global__ void __launch_bounds__(SP_BLOCK, 1)
crp_stream_kernel(const PairDesc* __restrict__ descs, const StripTask* __restrict__ tasks, int n_tasks,
                  const float* __restrict__ frames, const float* __restrict__ selfdot, const float* __restrict__ sdmax,
                  const int32_t* __restrict__ emb_len, const int64_t* __restrict__ op_off,
                  const unsigned char* __restrict__ opb, const unsigned char* __restrict__ opn,
                  const float* __restrict__ thr, const float* __restrict__ edges, uint8_t* __restrict__ crp,
                  int side_a_is_query,
                  const float2* __restrict__ piv, int* __restrict__ cnt4, int* __restrict__ lcnt, float2* __restrict__ lists,
                  unsigned long long* __restrict__ counters)
{
    using L = SpStreamSmem<MODE>;
    extern __shared__ __align__(128) unsigned char gsm[];
    unsigned char* Apl = gsm + L::A_OFF;
    unsigned char* Bbuf = gsm + L::B_OFF;
    float* loC = reinterpret_cast<float*>(gsm + L::EDGE_OFF);
    float* Araw = reinterpret_cast<float*>(gsm + L::ARAW_OFF);
    int* amb = reinterpret_cast<int*>(gsm + L::AMB_OFF);
    float* s_xx = reinterpret_cast<float*>(gsm + L::ROWV_OFF);
    float* s_th = s_xx + GM_ROWS;
    float* s_loq = s_th + GM_ROWS;
    int* amb_cnt = reinterpret_cast<int*>(s_loq + GM_ROWS);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(gsm + L::MBAR_OFF);
    uint64_t* mb_tfull = mbar, *mb_tempty = mbar + 2, *mb_bfull = mbar + 4, *mb_bempty = mbar + 7, *mb_afull = mbar + 10;
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(mbar + 12);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int lq = warp & 3, cg = (warp >> 2) & 3;
    const int myrow = 32 * lq + lane;
    unsigned long long n_exact = 0;

    if (warp == 0) gm_tmem_alloc(tmem_holder);
    for (int e = tid; e < GM_ROWS * 4; e += SP_BLOCK) reinterpret_cast<uint32_t*>(gsm + L::ZERO_OFF)[e] = 0u;
    fence_proxy_async_smem();              // the zero chunk is read by the tensor core (async proxy)
    if (tid == 0) {
        mbar_init(mb_tfull, 1); mbar_init(mb_tfull + 1, 1);
        mbar_init(mb_tempty, GM_WARPS); mbar_init(mb_tempty + 1, GM_WARPS);
        for (int k = 0; k < SP_NB_RING; ++k) { mbar_init(mb_bfull + k, 1); mbar_init(mb_bempty + k, 1); }
        mbar_init(mb_afull, GM_WARPS); mbar_init(mb_afull + 1, GM_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
    const uint32_t a_saddr = smem_u32(Apl), b_saddr = smem_u32(Bbuf);

    // tiny strips (NB < GM_MIN_NB) have no count pass: every role skips them the same way
    auto skip = [&](const PairDesc& pd) { return MODE != 1 && (side_a_is_query ? pd.N : pd.M) < GM_MIN_NB; };

    if (warp == GM_WARPS + 1) {
        // ================= TMA producer =================
        if (lane == 0) {
            uint32_t g = 0;
            for (int task = blockIdx.x; task < n_tasks; task += gridDim.x) {
                const StripTask tk = tasks[task];
                const PairDesc pd = descs[tk.pair];
                if (skip(pd)) continue;
                const int rot = pd.oti % ACOSS_NBINS;
                const int NB = side_a_is_query ? pd.N : pd.M;
                const int trackB = side_a_is_query ? pd.r : pd.q, rotB = side_a_is_query ? rot : 0;
                const int n_tiles = (NB + GM_TN - 1) / GM_TN;
                const int ntb = sp_tiles(emb_len[trackB]);           // == n_tiles
                const size_t P = (size_t)GM_TN * ntb + 8;
                const unsigned char* pl = opb + op_off[2 * trackB];
                const unsigned char* nr = opn + op_off[2 * trackB + 1] + (size_t)rotB * GM_TN * ntb * 16;
                const float* ed = MODE == 1 ? edges + pd.edge_off : nullptr;
                for (int t = 0; t < n_tiles; ++t, ++g) {
                    const uint32_t sb = g % SP_NB_RING;
                    if (g >= SP_NB_RING) mbar_wait(mb_bempty + sb, ((g / SP_NB_RING) - 1) & 1u);
                    unsigned char* dst = Bbuf + sb * GM_B_BYTES;
                    constexpr int P0 = (MODE == 2) ? 3 : 0;          // hi.hi count pass: the H planes only
                    const uint32_t total = (6 - P0) * GM_BPLANE + GM_TN * 16 + (MODE == 1 ? GM_TN * 4 : 0);
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(mb_bfull + sb)), "r"(total) : "memory");
#pragma unroll
                    for (int p = P0; p < 6; ++p)
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                     ::"r"(smem_u32(dst + p * GM_BPLANE)), "l"(pl + ((size_t)p * P + (size_t)GM_TN * t) * 16),
                                       "r"((uint32_t)GM_BPLANE), "r"(smem_u32(mb_bfull + sb)) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(smem_u32(dst + 6 * GM_BPLANE)), "l"(nr + (size_t)GM_TN * t * 16),
                                   "r"((uint32_t)(GM_TN * 16)), "r"(smem_u32(mb_bfull + sb)) : "memory");
                    if (MODE == 1)
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                     ::"r"(smem_u32(loC + (g % SP_EDGE_SLOTS) * GM_TN)), "l"(ed + (size_t)GM_TN * t),
                                       "r"((uint32_t)(GM_TN * 4)), "r"(smem_u32(mb_bfull + sb)) : "memory");
                }
            }
        }
    } else if (warp == GM_WARPS) {
        // ================= MMA issuer =================
        uint32_t g = 0, i = 0;
        for (int task = blockIdx.x; task < n_tasks; task += gridDim.x) {
            const StripTask tk = tasks[task];
            const PairDesc pd = descs[tk.pair];
            if (skip(pd)) continue;
            const int NB = side_a_is_query ? pd.N : pd.M;
            const int n_tiles = (NB + GM_TN - 1) / GM_TN;
            const uint32_t ab = i & 1;
            mbar_wait(mb_afull + ab, (i >> 1) & 1u);
            for (int t = 0; t < n_tiles; ++t, ++g) {
                const uint32_t sb = g % SP_NB_RING, buf = g & 1;
                mbar_wait(mb_bfull + sb, (g / SP_NB_RING) & 1u);
                if (g >= 2) mbar_wait(mb_tempty + buf, ((g >> 1) - 1) & 1u);
                tc_fence_after();
                if (gm_elect_one()) {
                    if (MODE == 2) sp_issue_tile_hh(tmem_base + buf * GM_TN, a_saddr + ab * GM_A_BYTES, b_saddr + sb * GM_B_BYTES);
                    else sp_issue_tile_mixed(tmem_base + buf * GM_TN, a_saddr + ab * GM_A_BYTES, b_saddr + sb * GM_B_BYTES, smem_u32(gsm + L::ZERO_OFF));
                    umma_commit(mb_tfull + buf);
                    umma_commit(mb_bempty + sb);
                }
                __syncwarp();
            }
            ++i;
        }
    } else if (warp < GM_WARPS) {
        // ================= epilogue warps =================
        const uint32_t my_taddr = tmem_base + ((uint32_t)(32 * lq) << 16) + 16 * cg;
        uint32_t g = 0, i = 0;
        // next non-skipped task at or after `from`
        auto next_task = [&](int from) {
            while (from < n_tasks && skip(descs[tasks[from].pair])) from += gridDim.x;
            return from;
        };
        auto stage = [&](int task, uint32_t ab) {
            const StripTask tk = tasks[task];
            const PairDesc pd = descs[tk.pair];
            const SpGeom ge = sp_geom(pd, side_a_is_query, frames, selfdot);
            sp_stage_a(Apl + ab * GM_A_BYTES, ge.fr_a, ge.sdA, tk.row0, tk.nrows, ge.na_frames, ge.rot_a, tid, false);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(mb_afull + ab);
        };
        int task = next_task(blockIdx.x);
        if (task < n_tasks) stage(task, 0);
        while (task < n_tasks) {
            const StripTask tk = tasks[task];
            const PairDesc pd = descs[tk.pair];
            const SpGeom ge = sp_geom(pd, side_a_is_query, frames, selfdot);
            const int NB = ge.NB, n_tiles = ge.n_tiles;
            const bool rowvalid = myrow < tk.nrows;
            const int nxt = next_task(task + gridDim.x);
            float E = 0.0f;
            if (MODE == 1) {
                E = sp_pair_bound(pd, sdmax).E;
                gm_csync();                          // the previous strip's exact chains are done with Araw / amb / s_*
                for (int e = tid; e < (GM_ROWS + 8) * ACOSS_NBINS; e += GM_THREADS) {
                    const int fe = e / ACOSS_NBINS, c = e - fe * ACOSS_NBINS;
                    const int f = tk.row0 + fe;
                    Araw[e] = (f < ge.na_frames) ? ge.fr_a[(int64_t)f * ACOSS_NBINS + c] : 0.0f;       // pass 2: A = query, unrotated
                }
                if (tid < GM_ROWS) {
                    const float xx = (tid < tk.nrows) ? ge.sdA[tk.row0 + tid] : 0.0f;
                    const float ta = (tid < tk.nrows) ? thr[pd.thr_q_off + tk.row0 + tid] : 0.0f;
                    const float Tq = xmul(ta, ta);
                    s_xx[tid] = xx; s_th[tid] = ta;
                    s_loq[tid] = Tq - (E + Tq * 4.76837158203125e-07f);          // E + 8uT
                }
                if (tid == 0) *amb_cnt = 0;
            }
            if (nxt < n_tasks) stage(nxt, (i + 1) & 1);
            if (MODE == 1) gm_csync();

            if (MODE != 1) {
                // ---- count pass: c_a = #(rho~ < pa); elements of [pa, pb) go to the row's list.  d = rho~ - pa: "below" is the
                //      sign bit of d, "inside [pa, pb)" is one unsigned compare of d's bit pattern against that of pb - pa.
                const size_t srow = (size_t)task * GM_ROWS + myrow;
                const float2 pv = rowvalid ? piv[srow] : make_float2(0.f, 0.f);
                const float pa = pv.x;
                const uint32_t wbits = __float_as_uint(pv.y - pv.x);
                float2* ml = lists + (srow * 4 + cg) * GM_SUBCAP;
                int ca = 0, ln = 0;
                for (int t = 0; t < n_tiles; ++t, ++g) {
                    const uint32_t buf = g & 1;
                    mbar_wait(mb_tfull + buf, (g >> 1) & 1u);
                    tc_fence_after();
                    uint32_t rr[2][16];                       // two batches of 16 columns: one in flight while the other is processed
                    tmem_ld16_issue(my_taddr + buf * GM_TN, rr[0]);
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        tmem_ld_wait16(rr[q4 & 1]);
                        if (q4 < 3) tmem_ld16_issue(my_taddr + buf * GM_TN + 64 * (q4 + 1), rr[(q4 + 1) & 1]);
                        if (rowvalid) {
                            const int jbase = t * GM_TN + 64 * q4 + 16 * cg;
                            if (ln + 16 <= GM_SUBCAP) {
                                // room for the whole batch: the write pointer just advances (no clamp in the dependency chain)
                                float2* wp = ml + ln;
#pragma unroll
                                for (int q = 0; q < 16; ++q) {
                                    const float vq = __uint_as_float(rr[q4 & 1][q]);
                                    const uint32_t u = __float_as_uint(vq - pa);
                                    ca += u >> 31;
                                    if (u < wbits) { *wp = make_float2(vq, __int_as_float(jbase + q)); ++wp; }
                                }
                                ln = (int)(wp - ml);
                            } else {
#pragma unroll
                                for (int q = 0; q < 16; ++q) {
                                    const float vq = __uint_as_float(rr[q4 & 1][q]);
                                    const uint32_t u = __float_as_uint(vq - pa);
                                    ca += u >> 31;
                                    if (u < wbits) {
                                        ml[min(ln, GM_SUBCAP - 1)] = make_float2(vq, __int_as_float(jbase + q));     // overflow is detected from ln
                                        ++ln;
                                    }
                                }
                            }
                        }
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(mb_tempty + buf);
                }
                cnt4[srow * 4 + cg] = ca;
                lcnt[srow * 4 + cg] = ln;
            } else {
                // ---- binarise pass.  Per cell: sure-in <=> rho~ < min(lo_q, lo_r[j]); 0 <= rho~ - min(...) <= 2.2 E is ambiguous
                //      and decided exactly after the strip; anything above is out.  32 CRP bytes per tcgen05.ld go straight
                //      from registers to HBM.
                const float lo_q = s_loq[myrow];
                const uint32_t zero_word = (CRP_DELTA_ZERO) | (CRP_DELTA_ZERO << 8) | (CRP_DELTA_ZERO << 16) | (CRP_DELTA_ZERO << 24);
                const uint32_t flip = CRP_DELTA_ZERO ^ CRP_DELTA_ONE;
                uint8_t* orow = crp + pd.crp_off + (int64_t)(tk.row0 + myrow) * pd.pitch;
                const int pitch = pd.pitch;
                const uint32_t wamb = __float_as_uint(2.2f * E);
                for (int t = 0; t < n_tiles; ++t, ++g) {
                    const uint32_t buf = g & 1;
                    mbar_wait(mb_tfull + buf, (g >> 1) & 1u);
                    tc_fence_after();
                    const float* lcs = loC + (g % SP_EDGE_SLOTS) * GM_TN;
                    uint32_t rr[2][16];
                    tmem_ld16_issue(my_taddr + buf * GM_TN, rr[0]);
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        tmem_ld_wait16(rr[q4 & 1]);
                        if (q4 < 3) tmem_ld16_issue(my_taddr + buf * GM_TN + 64 * (q4 + 1), rr[(q4 + 1) & 1]);
                        float v[16];
#pragma unroll
                        for (int q = 0; q < 16; ++q) v[q] = __uint_as_float(rr[q4 & 1][q]);
                        const int cbase = 64 * q4 + 16 * cg, jbase = t * GM_TN + cbase;
                        const float4* lc = reinterpret_cast<const float4*>(lcs + cbase);
                        uint32_t w[4], amask = 0;
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            const float4 l4 = lc[q];
                            const uint32_t u0 = __float_as_uint(v[4 * q] - fminf(lo_q, l4.x)), u1 = __float_as_uint(v[4 * q + 1] - fminf(lo_q, l4.y));
                            const uint32_t u2 = __float_as_uint(v[4 * q + 2] - fminf(lo_q, l4.z)), u3 = __float_as_uint(v[4 * q + 3] - fminf(lo_q, l4.w));
                            uint32_t s01, s23, sg;
                            asm("prmt.b32 %0, %1, %2, 0x00fb;" : "=r"(s01) : "r"(u0), "r"(u1));      // sign-replicated top bytes of u0, u1
                            asm("prmt.b32 %0, %1, %2, 0x00fb;" : "=r"(s23) : "r"(u2), "r"(u3));
                            asm("prmt.b32 %0, %1, %2, 0x5410;" : "=r"(sg) : "r"(s01), "r"(s23));
                            w[q] = zero_word ^ (sg & (flip * 0x01010101u));
                            if (min(min(u0, u1), min(u2, u3)) <= wamb)
                                amask |= ((uint32_t)(u0 <= wamb) | ((uint32_t)(u1 <= wamb) << 1) | ((uint32_t)(u2 <= wamb) << 2) | ((uint32_t)(u3 <= wamb) << 3)) << (4 * q);
                        }
                        if (rowvalid && jbase + 16 <= pitch) *reinterpret_cast<uint4*>(orow + jbase) = make_uint4(w[0], w[1], w[2], w[3]);
                        if (amask && rowvalid) {
                            int at = atomicAdd(amb_cnt, __popc(amask));
                            while (amask) {
                                const int b = __ffs(amask) - 1;
                                amask &= amask - 1;
                                if (at < SP_AMB_CAP) amb[at] = (myrow << 24) | (jbase + b);
                                ++at;
                            }
                        }
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(mb_tempty + buf);
                }
                // columns past the last tile up to the pitch
                const int done = n_tiles * GM_TN;
                if (done < pitch) {
                    const int wrem = (pitch - done) >> 2;
                    for (int idx = tid; idx < tk.nrows * wrem; idx += GM_THREADS) {
                        const int r = idx / wrem, w = idx - r * wrem;
                        reinterpret_cast<uint32_t*>(crp + pd.crp_off + (int64_t)(tk.row0 + r) * pitch + done)[w] = zero_word;
                    }
                }
                __threadfence_block();
                gm_csync();
                const int na = *amb_cnt;
                const float* thB = thr + pd.thr_r_off;
                const int rot = ge.rot;
                if (na > SP_AMB_CAP) {
                    // pathological (massive exact ties): decide every row of the strip exactly
                    for (int r = 0; r < tk.nrows; ++r) {
                        uint8_t* out = crp + pd.crp_off + (int64_t)(tk.row0 + r) * pd.pitch;
                        for (int j = tid; j < NB; j += GM_THREADS) {
                            const float dv = exact_dist(frames, pd, rot, selfdot, tk.row0 + r, j);
                            ++n_exact;
                            out[j] = (dv <= s_th[r] && dv <= thB[j]) ? (uint8_t)CRP_DELTA_ONE : (uint8_t)CRP_DELTA_ZERO;
                        }
                    }
                } else {
                    // side A = query (rows staged in Araw in the oracle's order), side B = reference, rotated by boff
                    const int boff = (ACOSS_NBINS - rot) % ACOSS_NBINS;
                    for (int e = tid; e < na; e += GM_THREADS) {
                        const int code = amb[e];
                        const int r = (code >> 24) & 0xFF, j = code & 0xFFFFFF;
                        const float4* bp = reinterpret_cast<const float4*>(ge.fr_b + (int64_t)j * ACOSS_NBINS);
                        const float4* as = reinterpret_cast<const float4*>(Araw + r * ACOSS_NBINS);
                        float acc = 0.0f;
#pragma unroll 3
                        for (int s_ = 0; s_ < 9; ++s_) {
                            const float4 a0 = as[3 * s_], a1 = as[3 * s_ + 1], a2 = as[3 * s_ + 2];
                            const float4 b0 = __ldg(bp + 3 * s_), b1 = __ldg(bp + 3 * s_ + 1), b2 = __ldg(bp + 3 * s_ + 2);
                            const float av[12] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w, a2.x, a2.y, a2.z, a2.w};
                            const float bv[12] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w, b2.x, b2.y, b2.z, b2.w};
                            acc = gm_chain12_rot(acc, av, bv, boff);
                        }
                        float item = xadd(xsub(s_xx[r], xmul(2.0f, acc)), __ldg(ge.sdB + j));
                        if (item < 0.0f) item = 0.0f;
                        const float dv = __fsqrt_rn(item);
                        ++n_exact;
                        if (dv <= s_th[r] && dv <= thB[j])
                            crp[pd.crp_off + (int64_t)(tk.row0 + r) * pd.pitch + j] = (uint8_t)CRP_DELTA_ONE;
                    }
                }
            }
            task = nxt;
            ++i;
        }
        for (int o = 16; o > 0; o >>= 1) n_exact += __shfl_xor_sync(0xffffffffu, n_exact, o);
        if (lane == 0 && n_exact) atomicAdd(&counters[0], n_exact);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) gm_tmem_free(tmem_base);
}

// per-column lower band edges of the column thresholds of every pair of a batch (pass 2): lo_r[j] = T - (E + 8uT),
// T = thr_r[j]^2, E the per-pair bound; -inf for the padding columns up to a multiple of 256 ("never in")
__global__ void crp_edges_kernel(const PairDesc* __restrict__ descs, int64_t n_pairs, const float* __restrict__ sdmax,
                                 const float* __restrict__ thr, float* __restrict__ edges)
{
    for (int64_t p = blockIdx.x; p < n_pairs; p += gridDim.x) {
        const PairDesc pd = descs[p];
        const float E = sp_pair_bound(pd, sdmax).E;
        const int W = sp_tiles(pd.N) * GM_TN;
        for (int j = threadIdx.x; j < W; j += blockDim.x) {
            float lo = -INFINITY;
            if (j < pd.N) {
                const float tb = thr[pd.thr_r_off + j];
                const float Tr = xmul(tb, tb);
                lo = Tr - (E + Tr * 4.76837158203125e-07f);
            }
            edges[pd.edge_off + j] = lo;
        }
    }
}

// =====================================================================================================================
// K3: selection.  256 threads per strip, four CTAs per SM.
//   (A) warp per row: the row's list (<= 512 entries, 16 per lane, one coalesced load per 32 slots) -> value bisection until
//       the window around the wanted rank holds at most three elements -> candidates = everything within 2E of the window
//   (B) thread per candidate (all rows of the strip pooled): the exact 108-term chain in the oracle's op order
//   (C) warp per row: exact order statistics among the candidates -> percentile -> threshold
//   rows that could not be bracketed (bracket miss, list overflow, tiny side B): the whole row exactly
// =====================================================================================================================
#define SP_POOL 3072                         // candidate entries per strip (flat: a row takes what it needs, at most SP_CAND_MAX)
#define SP_CAND_MAX 96
#define SP_SEL_SMEM (GM_ARAW_BYTES + 6 * GM_ROWS * 4 + 8 + 2 * SP_POOL * 4 + SP_SEL_WARPS * 64 * 4)

__global__ void __launch_bounds__(SP_SEL_THREADS, 4)
crp_select_kernel(const PairDesc* __restrict__ descs, const StripTask* __restrict__ tasks, int n_tasks,
                  const float* __restrict__ frames, const float* __restrict__ selfdot, const float* __restrict__ sdmax,
                  float* __restrict__ thr, int side_a_is_query, float qf,
                  const float2* __restrict__ piv, const int* __restrict__ cnt4, const int* __restrict__ lcnt,
                  const float2* __restrict__ lists, float* __restrict__ fb_scratch, int fb_stride, int count_hh,
                  unsigned long long* __restrict__ counters)
{
    extern __shared__ __align__(16) unsigned char ssm[];
    float* Araw = reinterpret_cast<float*>(ssm);                            // [GM_ROWS + 8][12] strip frames in the oracle's bin order
    float* s_xx = Araw + (GM_ROWS + 8) * ACOSS_NBINS;
    int* s_nc = reinterpret_cast<int*>(s_xx + GM_ROWS);
    int* s_kx = s_nc + GM_ROWS;
    int* s_flag = s_kx + GM_ROWS;
    int* s_off = s_flag + GM_ROWS;                                          // [GM_ROWS + 1]
    int* s_used = s_off + GM_ROWS;                                          // pool entries handed out
    int* pool_col = s_off + GM_ROWS + 2;                                    // [SP_POOL] (row << 24) | column
    float* pool_dex = reinterpret_cast<float*>(pool_col + SP_POOL);         // [SP_POOL] approximate radicand, then the exact distance
    int* s_hist = reinterpret_cast<int*>(pool_dex + SP_POOL);               // [SP_SEL_WARPS][64]
    int* s_pk = s_hist + SP_SEL_WARPS * 64;                                 // [GM_ROWS] sub-list boundaries of the row's compacted list
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned long long n_exact = 0;
    float max_err = 0.0f;

    for (int task = blockIdx.x; task < n_tasks; task += gridDim.x) {
        const StripTask tk = tasks[task];
        const PairDesc pd = descs[tk.pair];
        const SpGeom g = sp_geom(pd, side_a_is_query, frames, selfdot);
        const int rot = g.rot, NB = g.NB;
        const PercentileRank pr = percentile_rank(NB, qf);
        const SpPairBound pbd = sp_pair_bound(pd, sdmax);
        const float E = count_hh ? pbd.E_hh : pbd.E, plow = pbd.plow;
        float* thr_out = thr + (side_a_is_query ? pd.thr_q_off : pd.thr_r_off);
        const bool tiny = NB < GM_MIN_NB;
        __syncthreads();
        for (int e = tid; e < (GM_ROWS + 8) * ACOSS_NBINS; e += SP_SEL_THREADS) {
            const int fe = e / ACOSS_NBINS, c = e - fe * ACOSS_NBINS;
            const int f = tk.row0 + fe;
            Araw[e] = (f < g.na_frames) ? g.fr_a[(int64_t)f * ACOSS_NBINS + (side_a_is_query ? c : rot_src(c, rot))] : 0.0f;
        }
        if (tid < GM_ROWS) {
            s_xx[tid] = (tid < tk.nrows) ? g.sdA[tk.row0 + tid] : 0.0f;
            s_nc[tid] = 0; s_kx[tid] = 0; s_off[tid] = 0;
            if (tid == 0) *s_used = 0;
            s_flag[tid] = (tid < tk.nrows && tiny) ? 1 : 0;
        }
        __syncthreads();
        // ---- (A)
        if (!tiny)
            for (int r = warp; r < tk.nrows; r += SP_SEL_WARPS) {
                const size_t srow = (size_t)task * GM_ROWS + r;
                const int4 c4r = *reinterpret_cast<const int4*>(lcnt + srow * 4);
                const int4 a4r = *reinterpret_cast<const int4*>(cnt4 + srow * 4);
                const int cseg[4] = {c4r.x, c4r.y, c4r.z, c4r.w};
                const int c_ar = a4r.x + a4r.y + a4r.z + a4r.w, c_abr = c4r.x + c4r.y + c4r.z + c4r.w;
                const float2 pv = piv[srow];
                const float par = pv.x, pbr = pv.y;
                const bool ovf = max(max(c4r.x, c4r.y), max(c4r.z, c4r.w)) > GM_SUBCAP;
                bool okr = (c_ar <= pr.lo) && (c_ar + c_abr >= pr.hi + 1) && !ovf;
                if (!okr) {
                    if (lane == 0) { s_flag[r] = 1; atomicAdd(&counters[1], 1ull); if (ovf) atomicAdd(&counters[15], 1ull); else atomicAdd(&counters[13], 1ull); }
                    continue;
                }
                // The row's list (four sub-lists, <= 512 entries of [pa, pb) in all, ~200 typically) is read COMPACTED: entry
                // e = 32 i + lane for i < ceil(T / 32), so every later pass costs per entry in use, not per slot of the 16 x 32
                // layout (ncu, source view: the slot-wise passes cost 1600 warp instructions per row, nine tenths on empty slots).
                // One 64-bin histogram over the bracket (shared-memory atomics) + a prefix scan find the bin(s) holding the wanted
                // ranks; the window shrinks to the extremes of those bins (again if it still holds more than 8 entries); everything
                // within 2E of the window is a candidate for the exact chains.
                const float2* lst = lists + srow * SP_LIST_STRIDE;
                int* h = s_hist + warp * 64;
                const int T = c_abr, p1 = c4r.x, p2 = p1 + c4r.y, p3 = p2 + c4r.z;
                if (lane == 0) s_pk[r] = p1 | (p2 << 10) | (p3 << 20);
                const int nI = (T + 31) >> 5;                                       // warp-uniform, <= SP_SEL_LPL
                // every pass below walks the entries in blocks of four behind ONE warp-uniform branch per block (a per-entry guard is
                // if-converted by the compiler and the empty entries then cost as much as the live ones); entries past T hold +inf
#define SP_ENTRIES(i, ...)                                                                      \
                _Pragma("unroll") for (int ib_ = 0; ib_ < SP_SEL_LPL; ib_ += 4)                 \
                    if (ib_ < nI) { _Pragma("unroll") for (int i = ib_; i < ib_ + 4; ++i) { __VA_ARGS__ } }
                float lx[SP_SEL_LPL];
#pragma unroll
                for (int i = 0; i < SP_SEL_LPL; ++i) lx[i] = INFINITY;
                SP_ENTRIES(i,
                    const int e = 32 * i + lane;
                    if (e < T) {
                        const int seg = (e >= p1) + (e >= p2) + (e >= p3);
                        const int sl = e - (seg == 0 ? 0 : seg == 1 ? p1 : seg == 2 ? p2 : p3);
                        const float2 v = lst[seg * GM_SUBCAP + sl];
                        lx[i] = v.x;
                    })
                float base = par, top = pbr;
                if (par == plow || pbr == GM_PMAX) {                // open-ended bracket: the list's own value range
                    float vmn = INFINITY, vmx = -INFINITY;
                    SP_ENTRIES(i, if (lx[i] != INFINITY) { vmn = fminf(vmn, lx[i]); vmx = fmaxf(vmx, lx[i]); })
                    for (int o = 16; o > 0; o >>= 1) {
                        vmn = fminf(vmn, __shfl_xor_sync(0xffffffffu, vmn, o));
                        vmx = fmaxf(vmx, __shfl_xor_sync(0xffffffffu, vmx, o));
                    }
                    base = vmn; top = vmx + fabsf(vmx) * 1.0e-6f + 1.0e-30f;
                }
                int klo = pr.lo - c_ar, khi = klo + (pr.hi - pr.lo);                // ranks wanted among the entries in the window
                float w_lo = base, w_hi = top;
                uint32_t binw[SP_SEL_LPL / 4];                                      // bin of entry i: byte i % 4 of word i / 4; 0xFF = outside
#pragma unroll
                for (int q = 0; q < SP_SEL_LPL / 4; ++q) binw[q] = 0xFFFFFFFFu;
                okr = true;
                for (int round = 0; round < 3; ++round) {
                    const float bw = (top - base) * (1.0f / 64.0f);
                    const float bsc = (bw > 0.0f) ? 1.0f / bw : 0.0f;
                    h[lane] = 0; h[lane + 32] = 0;
                    __syncwarp();
                    SP_ENTRIES(i,
                        const uint32_t ob = (binw[i >> 2] >> (8 * (i & 3))) & 0xFFu;
                        const bool in = round == 0 ? (lx[i] != INFINITY) : (ob != 0xFFu);
                        uint32_t b = 0xFFu;
                        if (in) {
                            b = (uint32_t)min(63, max(0, (int)((lx[i] - base) * bsc)));
                            atomicAdd(&h[b], 1);
                        }
                        binw[i >> 2] = (binw[i >> 2] & ~(0xFFu << (8 * (i & 3)))) | (b << (8 * (i & 3)));)
                    __syncwarp();
                    const int h0 = h[2 * lane], h1 = h[2 * lane + 1];
                    __syncwarp();
                    int incl = h0 + h1;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const int t_ = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t_; }
                    const int excl = incl - h0 - h1;
                    const int my_lo = (klo >= excl && klo < incl) ? (klo < excl + h0 ? 2 * lane : 2 * lane + 1) : -1;
                    const int my_hi = (khi >= excl && khi < incl) ? (khi < excl + h0 ? 2 * lane : 2 * lane + 1) : -1;
                    const unsigned m_lo = __ballot_sync(0xffffffffu, my_lo >= 0), m_hi = __ballot_sync(0xffffffffu, my_hi >= 0);
                    if (m_lo == 0 || m_hi == 0) { okr = false; break; }
                    const int l_lo = __ffs(m_lo) - 1, l_hi = __ffs(m_hi) - 1;
                    const uint32_t b_lo = (uint32_t)__shfl_sync(0xffffffffu, my_lo, l_lo), b_hi = (uint32_t)__shfl_sync(0xffffffffu, my_hi, l_hi);
                    const int below_pair = __shfl_sync(0xffffffffu, excl, l_lo), h0_lo = __shfl_sync(0xffffffffu, h0, l_lo);
                    const int n_below_bin = below_pair + ((b_lo & 1) ? h0_lo : 0);
                    const int incl_hi = __shfl_sync(0xffffffffu, incl, l_hi), h1_hi = __shfl_sync(0xffffffffu, h1, l_hi);
                    const int n_upto_bin = incl_hi - ((b_hi & 1) ? 0 : h1_hi);
                    klo -= n_below_bin; khi -= n_below_bin;
                    float vmn = INFINITY, vmx = -INFINITY;
                    SP_ENTRIES(i,
                        const uint32_t b = (binw[i >> 2] >> (8 * (i & 3))) & 0xFFu;
                        if (b - b_lo <= b_hi - b_lo) { vmn = fminf(vmn, lx[i]); vmx = fmaxf(vmx, lx[i]); }     // 0xFF is never inside
                        else binw[i >> 2] |= 0xFFu << (8 * (i & 3));)
                    for (int o = 16; o > 0; o >>= 1) {
                        vmn = fminf(vmn, __shfl_xor_sync(0xffffffffu, vmn, o));
                        vmx = fmaxf(vmx, __shfl_xor_sync(0xffffffffu, vmx, o));
                    }
                    w_lo = vmn; w_hi = vmx;
                    if (n_upto_bin - n_below_bin <= 8 || !(vmx > vmn)) break;
                    base = vmn; top = vmx + fabsf(vmx) * 1.0e-6f + 1.0e-30f;
                }
                int n_c = 0, kx = 0;
                if (okr) {
                    const float c_min = w_lo - 2.0f * E, c_max = w_hi + 2.0f * E;
                    // the candidate window must lie strictly inside the list's value range [pa, pb)
                    okr = (c_min > par || par == plow) && (c_max < pbr || pbr == GM_PMAX);
                    int my_c = 0;
                    if (okr) {
                        int my_below = 0;
                        SP_ENTRIES(i,                                   // empty entries hold +inf: neither below nor candidates
                            my_below += lx[i] < c_min;
                            my_c += (lx[i] >= c_min) && (lx[i] <= c_max);)
                        const int n_below = c_ar + __reduce_add_sync(0xffffffffu, my_below);
                        n_c = __reduce_add_sync(0xffffffffu, my_c);
                        kx = pr.lo - n_below;
                        if (n_c > SP_CAND_MAX || kx < 0 || kx + (pr.hi - pr.lo) >= n_c) okr = false;
                    }
                    if (okr) {
                        int incl = my_c;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) { const int t_ = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t_; }
                        int off = 0;
                        if (lane == 0) off = atomicAdd(s_used, n_c);
                        off = __shfl_sync(0xffffffffu, off, 0);
                        if (off + n_c > SP_POOL) {                  // pool exhausted: the row goes the exact way; its slots stay marked empty
                            okr = false;
                            for (int sl = off + lane; sl < min(off + n_c, SP_POOL); sl += 32) pool_col[sl] = -1;
                        } else {
                            int at = off + incl - my_c;
                            SP_ENTRIES(i,
                                if ((lx[i] >= c_min) && (lx[i] <= c_max)) { pool_col[at] = (r << 24) | (32 * i + lane); pool_dex[at] = lx[i]; ++at; })
                            if (lane == 0) s_off[r] = off;
                        }
                    }
                }
#undef SP_ENTRIES
                if (lane == 0) {
                    if (okr) { s_nc[r] = n_c; s_kx[r] = kx; }
                    else { s_flag[r] = 1; atomicAdd(&counters[1], 1ull); atomicAdd(&counters[12], 1ull); }
                }
            }
        __syncthreads();
        // ---- (B)
        {
            const int boff = side_a_is_query ? (ACOSS_NBINS - rot) % ACOSS_NBINS : 0;     // rot_src(c, rot) = (c + boff) % 12
            const int n_tasks_c = min(*s_used, SP_POOL);             // entries of rows that did not fit are never written: their rows are flagged
            for (int idx = tid; idx < n_tasks_c; idx += SP_SEL_THREADS) {
                const int code = pool_col[idx];
                if (code < 0) continue;
                const int r = (code >> 24) & 0xFF, e = code & 0xFFFFFF;
                int jj;
                {   // entry e of the row's compacted list -> its column (second word of the list entry; the selection pass keeps
                    // only the values in registers: 16 registers less per thread = four CTAs per SM instead of three)
                    const int pk = s_pk[r], q1 = pk & 1023, q2 = (pk >> 10) & 1023, q3 = (pk >> 20) & 1023;
                    const int seg = (e >= q1) + (e >= q2) + (e >= q3);
                    const int sl = e - (seg == 0 ? 0 : seg == 1 ? q1 : seg == 2 ? q2 : q3);
                    jj = __float_as_int(lists[((size_t)task * GM_ROWS + r) * SP_LIST_STRIDE + seg * GM_SUBCAP + sl].y);
                }
                const float4* bp = reinterpret_cast<const float4*>(g.fr_b + (int64_t)jj * ACOSS_NBINS);
                const float4* as = reinterpret_cast<const float4*>(Araw + r * ACOSS_NBINS);
                float acc = 0.0f;
#pragma unroll 3
                for (int s_ = 0; s_ < 9; ++s_) {
                    const float4 a0 = as[3 * s_], a1 = as[3 * s_ + 1], a2 = as[3 * s_ + 2];
                    const float4 b0 = __ldg(bp + 3 * s_), b1 = __ldg(bp + 3 * s_ + 1), b2 = __ldg(bp + 3 * s_ + 2);
                    const float av[12] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w, a2.x, a2.y, a2.z, a2.w};
                    const float bv[12] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w, b2.x, b2.y, b2.z, b2.w};
                    acc = gm_chain12_rot(acc, av, bv, boff);
                }
                const float sa = s_xx[r], sb = __ldg(g.sdB + jj);
                const float xx = side_a_is_query ? sa : sb, yy = side_a_is_query ? sb : sa;
                float item = xadd(xsub(xx, xmul(2.0f, acc)), yy);
                max_err = fmaxf(max_err, fabsf(item - pool_dex[idx]) / (sa + sb));
                if (item < 0.0f) item = 0.0f;
                pool_dex[idx] = __fsqrt_rn(item);
                ++n_exact;
            }
        }
        __syncthreads();
        // ---- (C) warp per row: ranks kx and kx + 1 among the row's exact candidate distances (ties by slot); up to three
        //      candidates per lane, the others read as shared-memory broadcasts
        for (int r = warp; r < tk.nrows; r += SP_SEL_WARPS) {
            const int n_c = s_nc[r], kx = s_kx[r];
            if (n_c == 0) continue;
            const float* dex = pool_dex + s_off[r];
            float da[3];
            int rank[3] = {0, 0, 0};
#pragma unroll
            for (int c = 0; c < 3; ++c) da[c] = (32 * c + lane < n_c) ? dex[32 * c + lane] : INFINITY;
            if (n_c <= 32) {
                for (int b_ = 0; b_ < n_c; ++b_) {
                    const float db = dex[b_];
                    rank[0] += (db < da[0]) || (db == da[0] && b_ < lane);
                }
            } else {
                for (int b_ = 0; b_ < n_c; ++b_) {
                    const float db = dex[b_];
#pragma unroll
                    for (int c = 0; c < 3; ++c) rank[c] += (db < da[c]) || (db == da[c] && b_ < 32 * c + lane);
                }
            }
            float v_lo = 0.0f, v_hi = 0.0f;
            bool have_hi = false;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                if (32 * c >= n_c) break;
                const bool in = 32 * c + lane < n_c;
                const unsigned m_lo = __ballot_sync(0xffffffffu, in && rank[c] == kx);
                const unsigned m_hi = __ballot_sync(0xffffffffu, in && rank[c] == kx + 1);
                if (m_lo) v_lo = __shfl_sync(0xffffffffu, da[c], __ffs(m_lo) - 1);
                if (m_hi) { v_hi = __shfl_sync(0xffffffffu, da[c], __ffs(m_hi) - 1); have_hi = true; }
            }
            if (!have_hi) v_hi = v_lo;
            if (lane == 0) thr_out[tk.row0 + r] = percentile_combine(pr, v_lo, v_hi);
        }
        // ---- rows that could not be bracketed (and tiny side B): the whole row exactly, one row at a time
        __syncthreads();
        {
            int any = 0;
            if (tid < GM_ROWS) any = s_flag[tid];
            any = __syncthreads_or(any);
            float* vr = fb_scratch + (size_t)blockIdx.x * fb_stride;
            for (int r = 0; any && r < tk.nrows; ++r) {
                if (!s_flag[r]) continue;
                for (int j = tid; j < NB; j += SP_SEL_THREADS) {
                    vr[j] = side_a_is_query ? exact_dist(frames, pd, rot, selfdot, tk.row0 + r, j)
                                            : exact_dist(frames, pd, rot, selfdot, j, tk.row0 + r);
                    ++n_exact;
                }
                __threadfence_block();
                __syncthreads();
                if (warp == 0) {
                    float v_hi;
                    const float v_lo = warp_select_pair(vr, NB, pr.lo, lane, v_hi);
                    if (lane == 0) thr_out[tk.row0 + r] = percentile_combine(pr, v_lo, v_hi);
                }
                __syncthreads();
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        n_exact += __shfl_xor_sync(0xffffffffu, n_exact, o);
        max_err = fmaxf(max_err, __shfl_xor_sync(0xffffffffu, max_err, o));
    }
    if (lane == 0 && n_exact) atomicAdd(&counters[0], n_exact);
    if (lane == 0 && max_err > 0.0f) atomicMax(&counters[14], (unsigned long long)__float_as_uint(max_err));
}

// ---- measurement aid: clocks per MMA of the instruction mixes the streaming kernels issue (one CTA, operands = zeros) --------------
// out[v] = clock64 ticks of `reps` back-to-back tiles of variant v: 0 = 14 TF32 (hi.hi), 1 = 41 TF32 (three-product tile),
// 2 = 28 mixed (14 TF32 + 14 bf16), 3 = 14 bf16 with both K chunks adjacent (LBO 16), 4 = 14 bf16 with the K chunks in two planes.
__global__ void __launch_bounds__(128, 1) sp_mma_probe_kernel(long long* __restrict__ out, int reps)
{
    extern __shared__ __align__(128) unsigned char psm[];
    uint64_t* mbar = reinterpret_cast<uint64_t*>(psm + GM_A_BYTES + GM_B_BYTES + GM_ROWS * 16);
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(mbar + 1);
    for (int e = threadIdx.x; e < (GM_A_BYTES + GM_B_BYTES + GM_ROWS * 16) / 4; e += blockDim.x) reinterpret_cast<uint32_t*>(psm)[e] = 0u;
    fence_proxy_async_smem();
    if (threadIdx.x < 32) gm_tmem_alloc(tmem_holder);
    if (threadIdx.x == 0) { mbar_init(mbar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
    const uint32_t a0 = smem_u32(psm), b0 = a0 + GM_A_BYTES, z0 = b0 + GM_B_BYTES;
    if (threadIdx.x < 32) {
        uint32_t ph = 0;
        for (int v = 0; v < 5; ++v) {
            long long t0 = 0;
            for (int warm = 0; warm < 2; ++warm) {
                t0 = clock64();
                if (gm_elect_one()) {
                    for (int r = 0; r < reps; ++r) {
                        const uint32_t d = tmem_base + (r & 1) * GM_TN;
                        if (v == 0) sp_issue_tile_hh(d, a0, b0);
                        else if (v == 1) gm_issue_tile(d, a0, b0);
                        else if (v == 2) sp_issue_tile_mixed(d, a0, b0, z0);
                        else {
                            const uint32_t al = (a0 & 0x3FFFFu) >> 4, bl = (b0 & 0x3FFFFu) >> 4;
                            for (int k = 0; k < 14; ++k) {
                                if (v == 3) gm_mma_bf16(d, GM_DLO(al, 3 * GM_APLANE + 16 * (k % 8), 16), GM_DLO(bl, 16 * (k % 8), 16), k ? 1u : 0u);
                                else gm_mma_bf16(d, GM_DLO(al, 3 * GM_APLANE + 16 * (k % 8), GM_APLANE), GM_DLO(bl, 16 * (k % 8), GM_BPLANE), k ? 1u : 0u);
                            }
                        }
                    }
                    umma_commit(mbar);
                }
                __syncwarp();
                mbar_wait(mbar, ph); ph ^= 1;
                tc_fence_after();
            }
            if (threadIdx.x == 0) out[v] = clock64() - t0;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) gm_tmem_free(tmem_base);
}
