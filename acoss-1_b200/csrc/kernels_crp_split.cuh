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
#define SP_LIDX(e) ((((uint32_t)(e) >> 2) << 11) | ((uint32_t)(e) & 3u))     // entry e of a sub-list, relative to the sub-list's first sector
#define SP_LMID ((4u * GM_ROWS - 1u) << 2)       // the sub-list field of a list index (bits 2..10)
static_assert(GM_ROWS == 128 && GM_SUBCAP == 128, "SP_LIDX: 2 + 9 + 5 index bits");
#ifdef SP_TIMING        // debug build (ACOSS_NVCC_EXTRA=-DSP_TIMING): clocks of the count pass summed over the CTAs -- MMA issuer: counters[2..5] =
                        // wait a_full, b_full, tmem_empty, total; epilogue warp 0: [6] = wait tmem_full, [8] = tcgen05.wait::ld, [9] = batch bodies, [10] = total
#define SP_TDECL(x) long long x = 0
#define SP_T0(x) const long long x = clock64()
#define SP_T1(x, acc) acc += clock64() - x
#else
#define SP_TDECL(x)
#define SP_T0(x)
#define SP_T1(x, acc)
#endif
#ifdef SP_TIMING_BIN    // debug build: clocks of the binarise pass's epilogue warp 0 -- counters[2..5] = before the tiles, tiles, after the tiles, total; [6] = ambiguous cells
#define SB_T(x) const long long x = clock64()
#else
#define SB_T(x)
#endif
#define SP_NO_ENTRY __int_as_float(0x7fc00000)   // filler of a batched list walk past the end of a sub-list: a NaN is neither below nor inside any window
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
    // binarise pass: `lists` / `lcnt`, unused by this pass, carry the batch's list of ambiguous cells (task, row << 24 | column) and
    // its fill count when the host found room for it (crp_fix_kernel then decides the cells); null: the strip decides its own
    int2* amb_g = reinterpret_cast<int2*>(lists);
    unsigned long long* amb_gcnt = reinterpret_cast<unsigned long long*>(lcnt);
    const bool deferred = MODE == 1 && lists != nullptr;
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
        SP_TDECL(tm_a); SP_TDECL(tm_b); SP_TDECL(tm_e);
        SP_T0(tm_all);
        for (int task = blockIdx.x; task < n_tasks; task += gridDim.x) {
            const StripTask tk = tasks[task];
            const PairDesc pd = descs[tk.pair];
            if (skip(pd)) continue;
            const int NB = side_a_is_query ? pd.N : pd.M;
            const int n_tiles = (NB + GM_TN - 1) / GM_TN;
            const uint32_t ab = i & 1;
            SP_T0(ta);
            mbar_wait(mb_afull + ab, (i >> 1) & 1u);
            SP_T1(ta, tm_a);
            for (int t = 0; t < n_tiles; ++t, ++g) {
                const uint32_t sb = g % SP_NB_RING, buf = g & 1;
                SP_T0(tb);
                mbar_wait(mb_bfull + sb, (g / SP_NB_RING) & 1u);
                SP_T1(tb, tm_b);
                SP_T0(te);
                if (g >= 2) mbar_wait(mb_tempty + buf, ((g >> 1) - 1) & 1u);
                SP_T1(te, tm_e);
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
#ifdef SP_TIMING
        if (lane == 0 && counters) {
            if (MODE != 1) {
                atomicAdd(&counters[2], (unsigned long long)tm_a); atomicAdd(&counters[3], (unsigned long long)tm_b);
                atomicAdd(&counters[4], (unsigned long long)tm_e); atomicAdd(&counters[5], (unsigned long long)(clock64() - tm_all));
            }
        }
#endif
    } else if (warp < GM_WARPS) {
        // ================= epilogue warps =================
        const uint32_t my_taddr = tmem_base + ((uint32_t)(32 * lq) << 16) + 16 * cg;
        uint32_t g = 0, i = 0;
        SP_TDECL(tm_f); SP_TDECL(tm_w); SP_TDECL(tm_p);
        SP_T0(tm_epi_all);
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
            SB_T(sb0);
            if (MODE == 1) {
                E = sp_pair_bound(pd, sdmax).E;
                gm_csync();                          // the previous strip's exact chains are done with Araw / amb / s_*
                if (!deferred) for (int e = tid; e < (GM_ROWS + 8) * ACOSS_NBINS; e += GM_THREADS) {
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
                // sector-grouped lists: entry e of sub-list (row, cg) at SP_LIDX(e) of the strip's region, i.e. the four entries
                // e & ~3 .. | 3 of a sub-list fill one 32-byte sector and the sectors of the 512 sub-lists of a group lie side by
                // side (the selection kernel's lanes -- one sub-list each -- read 1 KB contiguous per step).  Entry-major order
                // (one 8-byte entry per sub-list side by side) made the four warps of a row complete each sector at different
                // times: ncu showed a DRAM read per sector (1.4 GB per launch, the list volume) and the kernel ran 30 % longer.
                // The write cursor keeps the sub-list field of the index at all ones, so that + 1 carries from the entry's low two
                // bits straight into the group number; the sub-list's own field is folded into the base pointer.
                float2* mc = lists + (size_t)task * (GM_SUBCAP * 4 * GM_ROWS) + (myrow * 4 + cg) * 4 - SP_LMID;
                int ca = 0;
                uint32_t c = SP_LMID;                         // write cursor: entry 0
                constexpr uint32_t C_FULL = (uint32_t)((GM_SUBCAP - 16) >> 2) << 11;      // fewer than 16 free slots from this group on
                // one batch of 16 columns.  A sub-list with fewer than 16 free slots counts as overflowed from there on (its row is
                // then taken exactly by the selection kernel), so the epilogue carries one store sequence without a clamp.
                auto batch = [&](const uint32_t (&r)[16], int jbase) {
                    if (rowvalid && c < C_FULL) {
#pragma unroll
                        for (int q = 0; q < 16; ++q) {
                            const float vq = __uint_as_float(r[q]);
                            const uint32_t u = __float_as_uint(vq - pa);
                            ca += u >> 31;
                            if (u < wbits) { mc[c] = make_float2(vq, __int_as_float(jbase + q)); c = (c + 1) | SP_LMID; }
                        }
                    }
                };
                for (int t = 0; t < n_tiles; ++t, ++g) {
                    const uint32_t buf = g & 1;
                    SP_T0(tf);
                    mbar_wait(mb_tfull + buf, (g >> 1) & 1u);
                    SP_T1(tf, tm_f);
                    tc_fence_after();
                    uint32_t r0[16], r1[16];                  // two batches of 16 columns: one in flight while the other is processed
                    const uint32_t ta = my_taddr + buf * GM_TN;
                    const int jb = t * GM_TN + 16 * cg;
                    tmem_ld16_issue(ta, r0);
                    // two rounds of (even batch, odd batch): half the code of the four-way unrolled loop -- the three warp roles of
                    // this kernel run different code and ncu showed 10 % of the stall samples waiting for instructions
#pragma unroll 1
                    for (int h = 0; h < 2; ++h) {
                        SP_T0(tw0);
                        tmem_ld_wait16(r0);
                        SP_T1(tw0, tm_w);
                        tmem_ld16_issue(ta + 128 * h + 64, r1);
                        SP_T0(tp0);
                        batch(r0, jb + 128 * h);
                        SP_T1(tp0, tm_p);
                        SP_T0(tw1);
                        tmem_ld_wait16(r1);
                        SP_T1(tw1, tm_w);
                        if (h == 0) tmem_ld16_issue(ta + 128, r0);
                        SP_T0(tp1);
                        batch(r1, jb + 128 * h + 64);
                        SP_T1(tp1, tm_p);
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(mb_tempty + buf);
                }
                const int ln = (c >= C_FULL) ? GM_SUBCAP + 1 : (int)((c & 3u) | ((c >> 11) << 2));
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
                SB_T(sb1);
                for (int t = 0; t < n_tiles; ++t, ++g) {
                    const uint32_t buf = g & 1;
                    SP_T0(tf);
                    mbar_wait(mb_tfull + buf, (g >> 1) & 1u);
                    SP_T1(tf, tm_f);
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
                SB_T(sb2);
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
                } else if (deferred) {
                    // the strip's ambiguous cells go to the batch's list (crp_fix_kernel decides them after this pass): resolving
                    // them here kept all 16 epilogue warps of the CTA off the next strip's tiles for ~10 000 clocks per strip
                    // (clock64 probes: a fifth of the pass), for ~140 cells.  Warp 0 copies while the others go on to the next strip
                    // (they meet at its first barrier).  A strip that finds the list full has warp 0 decide its cells here,
                    // straight from global memory, and void the slots it drew.
                    if (warp == 0 && na > 0) {
                        unsigned long long base = 0;
                        if (lane == 0) base = atomicAdd(amb_gcnt, (unsigned long long)na);
                        base = __shfl_sync(0xffffffffu, base, 0);
                        const unsigned long long cap = amb_gcnt[1];
                        const bool ok = base + (unsigned long long)na <= cap;
                        for (int e = lane; e < na; e += 32) {
                            const int code = amb[e];
                            if (base + e < cap) amb_g[base + e] = ok ? make_int2(task, code) : make_int2(-1, 0);
                            if (!ok) {
                                const int r = (code >> 24) & 0xFF, j = code & 0xFFFFFF;
                                const float dv = exact_dist(frames, pd, rot, selfdot, tk.row0 + r, j);
                                ++n_exact;
                                if (dv <= s_th[r] && dv <= thB[j]) crp[pd.crp_off + (int64_t)(tk.row0 + r) * pd.pitch + j] = (uint8_t)CRP_DELTA_ONE;
                            }
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
#ifdef SP_TIMING_BIN
                if (tid == 0 && counters) {
                    const long long sb3 = clock64();
                    atomicAdd(&counters[2], (unsigned long long)(sb1 - sb0)); atomicAdd(&counters[3], (unsigned long long)(sb2 - sb1));
                    atomicAdd(&counters[4], (unsigned long long)(sb3 - sb2)); atomicAdd(&counters[5], (unsigned long long)(sb3 - sb0));
                    atomicAdd(&counters[6], (unsigned long long)na);
                }
#endif
            }
            task = nxt;
            ++i;
        }
        for (int o = 16; o > 0; o >>= 1) n_exact += __shfl_xor_sync(0xffffffffu, n_exact, o);
        if (lane == 0 && n_exact) atomicAdd(&counters[0], n_exact);
#ifdef SP_TIMING
        if (tid == 0 && counters) {
            atomicAdd(&counters[MODE == 1 ? 7 : 6], (unsigned long long)tm_f);
            if (MODE != 1) {            // count pass, epilogue warp 0: clocks in tcgen05.wait::ld, in the batch bodies, in total
                atomicAdd(&counters[8], (unsigned long long)tm_w); atomicAdd(&counters[9], (unsigned long long)tm_p);
                atomicAdd(&counters[10], (unsigned long long)(clock64() - tm_epi_all));
            }
        }
#endif
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
// K4b: the ambiguous cells of the binarise pass, one thread per cell: the exact 108-term chain in the oracle's op order against both
// thresholds (the cell's CRP byte was written as "out" by the pass).
// =====================================================================================================================
__global__ void sp_amb_init_kernel(unsigned long long* amb_gcnt, unsigned long long cap) { amb_gcnt[0] = 0ull; amb_gcnt[1] = cap; }

__global__ void __launch_bounds__(256)
crp_fix_kernel(const PairDesc* __restrict__ descs, const StripTask* __restrict__ tasks, const float* __restrict__ frames,
               const float* __restrict__ selfdot, const float* __restrict__ thr, uint8_t* __restrict__ crp,
               const int2* __restrict__ amb_g, const unsigned long long* __restrict__ amb_gcnt, unsigned long long* __restrict__ counters)
{
    const unsigned long long n = min(amb_gcnt[0], amb_gcnt[1]);            // fill count, capacity
    unsigned long long n_exact = 0;
    for (unsigned long long e = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (unsigned long long)gridDim.x * blockDim.x) {
        const int2 ent = amb_g[e];
        if (ent.x < 0) continue;                                            // voided by a strip that found the list full
        const StripTask tk = tasks[ent.x];
        const PairDesc pd = descs[tk.pair];
        const SpGeom ge = sp_geom(pd, 1, frames, selfdot);                  // side A = query (unrotated), side B = reference
        const int r = (ent.y >> 24) & 0xFF, j = ent.y & 0xFFFFFF, row = tk.row0 + r;
        const int boff = (ACOSS_NBINS - ge.rot) % ACOSS_NBINS;
        const float4* ap = reinterpret_cast<const float4*>(ge.fr_a + (int64_t)row * ACOSS_NBINS);
        const float4* bp = reinterpret_cast<const float4*>(ge.fr_b + (int64_t)j * ACOSS_NBINS);
        float acc = 0.0f;
#pragma unroll 3
        for (int s_ = 0; s_ < 9; ++s_) {
            const float4 a0 = __ldg(ap + 3 * s_), a1 = __ldg(ap + 3 * s_ + 1), a2 = __ldg(ap + 3 * s_ + 2);
            const float4 b0 = __ldg(bp + 3 * s_), b1 = __ldg(bp + 3 * s_ + 1), b2 = __ldg(bp + 3 * s_ + 2);
            const float av[12] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w, a2.x, a2.y, a2.z, a2.w};
            const float bv[12] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w, b2.x, b2.y, b2.z, b2.w};
            acc = gm_chain12_rot(acc, av, bv, boff);
        }
        float item = xadd(xsub(__ldg(ge.sdA + row), xmul(2.0f, acc)), __ldg(ge.sdB + j));
        if (item < 0.0f) item = 0.0f;
        const float dv = __fsqrt_rn(item);
        ++n_exact;
        if (dv <= __ldg(thr + pd.thr_q_off + row) && dv <= __ldg(thr + pd.thr_r_off + j))
            crp[pd.crp_off + (int64_t)row * pd.pitch + j] = (uint8_t)CRP_DELTA_ONE;
    }
    for (int o = 16; o > 0; o >>= 1) n_exact += __shfl_xor_sync(0xffffffffu, n_exact, o);
    if ((threadIdx.x & 31) == 0 && n_exact && counters) atomicAdd(&counters[0], n_exact);
}

// =====================================================================================================================
#define SP_SEL_SMEM (GM_ARAW_BYTES + 5 * GM_ROWS * 4 + 16 + 2 * GM_ROWS * 32 * 4 + 16 * SP_SEL_THREADS * 4)

// =====================================================================================================================
// K3: selection.  256 threads per strip, four CTAs per SM.
//   (A) a quad of lanes per row (one lane per sub-list): 64-bin histogram of the list over the bracket, refined by value interval
//       until the bins of the wanted ranks hold at most 12 entries -> candidates = everything within 2E of that window
//   (B) thread per candidate (all rows of the strip pooled): the exact 108-term chain in the oracle's op order
//   (C) warp per row: exact order statistics among the candidates -> percentile -> threshold
//   rows that could not be bracketed (bracket miss, list overflow, tiny side B): the whole row exactly
// =====================================================================================================================
__global__ void __launch_bounds__(SP_SEL_THREADS, 3)
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
    int* pool_col = s_off + GM_ROWS + 4;                                    // [GM_ROWS][32] candidate columns
    float* pool_dex = reinterpret_cast<float*>(pool_col + GM_ROWS * 32);    // [GM_ROWS][32] approximate radicand, then the exact distance
    unsigned* s_hq = reinterpret_cast<unsigned*>(pool_dex + GM_ROWS * 32);  // [16 words][SP_SEL_THREADS] private 64-bin byte histograms
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
        // the strip's frames and self-dots are only needed by part (B): their loads are issued here and land in shared memory after
        // part (A), whose walks hide the latency (they used to sit in front of a barrier: 6 % of the kernel's stall samples)
        constexpr int SP_ARAW_PER_THREAD = ((GM_ROWS + 8) * ACOSS_NBINS + SP_SEL_THREADS - 1) / SP_SEL_THREADS;
        float areg[SP_ARAW_PER_THREAD], xreg = 0.0f;
#pragma unroll
        for (int k = 0; k < SP_ARAW_PER_THREAD; ++k) {
            const int e = tid + k * SP_SEL_THREADS;
            const int fe = e / ACOSS_NBINS, c = e - fe * ACOSS_NBINS;
            const int f = tk.row0 + fe;
            areg[k] = (e < (GM_ROWS + 8) * ACOSS_NBINS && f < g.na_frames)
                          ? g.fr_a[(int64_t)f * ACOSS_NBINS + (side_a_is_query ? c : rot_src(c, rot))] : 0.0f;
        }
        if (tid < GM_ROWS) {
            xreg = (tid < tk.nrows) ? g.sdA[tk.row0 + tid] : 0.0f;
            s_nc[tid] = 0; s_kx[tid] = 0;
            s_flag[tid] = (tid < tk.nrows && tiny) ? 1 : 0;
        }
        __syncthreads();
        // ---- (A) a QUAD of lanes per row -- one lane per sub-list -- eight rows per warp.  With a warp per row every pass paid ~30
        //      warp-wide instructions for 5 list entries per lane plus scans, ballots and reductions per row (ncu: 770 warp
        //      instructions per row in this part, 39 % of the CRP stage).  Here a lane walks its own sub-list (~42 entries; the lists
        //      are sector-grouped, so the 32 lanes of a warp read 1 KB contiguous per pair of loads), bins it into a private 64-bin byte
        //      histogram (shared memory, bank = thread), the quad adds its four histograms, finds the bin(s) of the wanted ranks,
        //      and a second walk counts what lies below the candidate window and appends the candidates to the row's pool.
        if (!tiny) {
            const int sub = lane & 3, rw = lane >> 2, qb = lane & ~3;
            unsigned* hq = s_hq + tid;                                      // word w of this thread's histogram: hq[w * SP_SEL_THREADS]
            for (int r0 = 8 * warp; r0 < tk.nrows; r0 += 8 * SP_SEL_WARPS) {
                const int r = r0 + rw;
                const bool live = r < tk.nrows;
                const size_t srow = (size_t)task * GM_ROWS + (live ? r : r0);
                const int n_s = live ? lcnt[srow * 4 + sub] : 0, a_s = live ? cnt4[srow * 4 + sub] : 0;
                int c_ar = a_s, c_abr = n_s, n_mx = n_s;
                c_ar += __shfl_xor_sync(0xffffffffu, c_ar, 1); c_ar += __shfl_xor_sync(0xffffffffu, c_ar, 2);
                c_abr += __shfl_xor_sync(0xffffffffu, c_abr, 1); c_abr += __shfl_xor_sync(0xffffffffu, c_abr, 2);
                n_mx = max(n_mx, __shfl_xor_sync(0xffffffffu, n_mx, 1)); n_mx = max(n_mx, __shfl_xor_sync(0xffffffffu, n_mx, 2));
                const float2 pv = piv[srow];
                const float par = pv.x, pbr = pv.y;
                const bool ovf = n_mx > GM_SUBCAP;
                bool okr = live && (c_ar <= pr.lo) && (c_ar + c_abr >= pr.hi + 1) && !ovf;
                if (live && !okr && sub == 0) {
                    s_flag[r] = 1; atomicAdd(&counters[1], 1ull);
                    if (ovf) atomicAdd(&counters[15], 1ull); else atomicAdd(&counters[13], 1ull);
                }
                if (!__any_sync(0xffffffffu, okr)) continue;
                // entry e of this lane's sub-list: lists[(task * GM_SUBCAP + e) * (4 GM_ROWS) + 4 row + sub]
                const float2* ml = lists + (size_t)task * GM_SUBCAP * (4 * GM_ROWS) + ((live ? r : r0) * 4 + sub) * 4;
                const int n_e = okr ? n_s : 0;
                const int n_w = __reduce_max_sync(0xffffffffu, n_e);          // walk length of the warp
                float base = par, top = pbr;
                const bool open_ended = okr && (par == plow || pbr == GM_PMAX);
                if (__any_sync(0xffffffffu, open_ended)) {                    // open-ended bracket: the list's own value range
                    float vmn = INFINITY, vmx = -INFINITY;
#pragma unroll 8
                    for (int e = 0; e < n_w; ++e)
                        if (e < n_e) { const float v = ml[SP_LIDX(e)].x; vmn = fminf(vmn, v); vmx = fmaxf(vmx, v); }
                    vmn = fminf(vmn, __shfl_xor_sync(0xffffffffu, vmn, 1)); vmn = fminf(vmn, __shfl_xor_sync(0xffffffffu, vmn, 2));
                    vmx = fmaxf(vmx, __shfl_xor_sync(0xffffffffu, vmx, 1)); vmx = fmaxf(vmx, __shfl_xor_sync(0xffffffffu, vmx, 2));
                    if (open_ended) { base = vmn; top = vmx + fabsf(vmx) * 1.0e-6f + 1.0e-30f; }
                }
                // walk 1 (up to three rounds): histogram of the entries inside the current value interval [lo_v, hi_v] over [base, top);
                // the bin(s) of the wanted ranks become the next interval while they still hold more than 12 entries (runs of
                // near-equal distances).  Membership is by VALUE and the entries below the interval are re-counted in every walk,
                // so the rounds need no per-entry state.
                float lo_v = -INFINITY, hi_v = INFINITY;
                bool refining = okr;
                for (int round = 0; round < 3; ++round) {
                    if (!__any_sync(0xffffffffu, refining)) break;
                    const float bw = (top - base) * (1.0f / 64.0f);
                    const float bsc = (bw > 0.0f) ? 1.0f / bw : 0.0f;
#pragma unroll
                    for (int w = 0; w < 16; ++w) hq[w * SP_SEL_THREADS] = 0u;
                    int below = 0;
                    // eight entries in flight per lane (one load per step left every step waiting for the L2 round trip)
                    for (int e0 = 0; e0 < n_w; e0 += 8) {
                        float v8[8];
#pragma unroll
                        for (int h = 0; h < 4; ++h) {                                   // two sectors of the sub-list, two entries per load
                            const bool in = (e0 + 2 * h < n_e) && refining;
                            const float4 t4 = in ? *reinterpret_cast<const float4*>(ml + SP_LIDX(e0 + 2 * h)) : make_float4(0.f, 0.f, 0.f, 0.f);
                            v8[2 * h] = in ? t4.x : SP_NO_ENTRY;
                            v8[2 * h + 1] = (in && e0 + 2 * h + 1 < n_e) ? t4.z : SP_NO_ENTRY;
                        }
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            // straight-line body (a branch per entry cost two reconvergence points in a divergent warp)
                            const float v = v8[u];
                            below += (v < lo_v);
                            const unsigned b = min(63u, __float2uint_rz((v - base) * bsc));             // negative and NaN -> 0
                            // private counter: the reduction is fire-and-forget.  Entries outside the interval (and the NaN of
                            // "no entry", which fails the compares) add zero: the compiler keeps a conditional atomic behind a branch
                            atomicAdd(hq + (b >> 2) * SP_SEL_THREADS, (v >= lo_v && v <= hi_v) ? 1u << ((8u * b) & 31u) : 0u);
                        }
                    }
                    __syncwarp();
                    below += __shfl_xor_sync(0xffffffffu, below, 1); below += __shfl_xor_sync(0xffffffffu, below, 2);
                    // the quad's histogram: lane `sub` adds bins 16 sub .. 16 sub + 15 of the four lanes (16-bit pairs: bins 0 | 2 and 1 | 3 of a word)
                    unsigned lo2[4], hi2[4];
                    int tot = 0;
#pragma unroll
                    for (int w = 0; w < 4; ++w) {
                        unsigned l = 0, h = 0;
#pragma unroll
                        for (int o = 0; o < 4; ++o) {
                            const unsigned x = s_hq[(4 * sub + w) * SP_SEL_THREADS + (tid & ~3) + o];
                            l += x & 0x00ff00ffu; h += (x >> 8) & 0x00ff00ffu;
                        }
                        lo2[w] = l; hi2[w] = h;
                        tot += (int)((l & 0xffffu) + (l >> 16) + (h & 0xffffu) + (h >> 16));
                    }
                    __syncwarp();
                    int excl = 0;
#pragma unroll
                    for (int o = 0; o < 3; ++o) { const int t_ = __shfl_sync(0xffffffffu, tot, qb + o); if (o < sub) excl += t_; }
                    const int klo = pr.lo - c_ar - below, khi = klo + (pr.hi - pr.lo);      // ranks wanted among the entries of the interval
                    int b_lo = -1, b_hi = -1, n_lo = 0, n_hi = 0;                            // bins of the ranks, entries before b_lo / through b_hi
                    {
                        int run = excl;
#pragma unroll
                        for (int w = 0; w < 4; ++w)
#pragma unroll
                            for (int f = 0; f < 4; ++f) {
                                const int cnt = (int)(f == 0 ? (lo2[w] & 0xffffu) : f == 1 ? (hi2[w] & 0xffffu) : f == 2 ? (lo2[w] >> 16) : (hi2[w] >> 16));
                                if (klo >= run && klo < run + cnt) { b_lo = 16 * sub + 4 * w + f; n_lo = run; }
                                if (khi >= run && khi < run + cnt) { b_hi = 16 * sub + 4 * w + f; n_hi = run + cnt; }
                                run += cnt;
                            }
                    }
                    b_lo = max(b_lo, __shfl_xor_sync(0xffffffffu, b_lo, 1)); b_lo = max(b_lo, __shfl_xor_sync(0xffffffffu, b_lo, 2));
                    b_hi = max(b_hi, __shfl_xor_sync(0xffffffffu, b_hi, 1)); b_hi = max(b_hi, __shfl_xor_sync(0xffffffffu, b_hi, 2));
                    n_lo = max(n_lo, __shfl_xor_sync(0xffffffffu, n_lo, 1)); n_lo = max(n_lo, __shfl_xor_sync(0xffffffffu, n_lo, 2));
                    n_hi = max(n_hi, __shfl_xor_sync(0xffffffffu, n_hi, 1)); n_hi = max(n_hi, __shfl_xor_sync(0xffffffffu, n_hi, 2));
                    if (refining) {
                        if (b_lo < 0 || b_hi < 0) { okr = false; refining = false; }
                        else {
                            // the next interval: the bins of the ranks, 2 % of a bin wider for the rounding of the binning map
                            lo_v = fmaxf(lo_v, base + ((float)b_lo - 0.02f) * bw);
                            hi_v = fminf(hi_v, base + ((float)b_hi + 1.02f) * bw);
                            if (n_hi - n_lo <= 12 || !(hi_v > lo_v) || bw <= 0.0f) refining = false;
                            else { base = lo_v; top = hi_v; }
                        }
                    }
                }
                float c_min = lo_v - 2.0f * E, c_max = hi_v + 2.0f * E;
                {
                    // a window that reaches over an end of the bracket by its BIN EDGE may still lie inside by its entries (the wanted
                    // rank sits in the first or the last bin for ~3 % of the rows): those rows take the extremes of the entries
                    const bool e_lo = okr && !(c_min > par || par == plow), e_hi = okr && !(c_max < pbr || pbr == GM_PMAX);
                    if (__any_sync(0xffffffffu, e_lo || e_hi)) {
                        float vmn = INFINITY, vmx = -INFINITY;
                        for (int e = 0; e < n_w; ++e)
                            if (e < n_e && (e_lo || e_hi)) {
                                const float v = ml[SP_LIDX(e)].x;
                                if (v >= lo_v && v <= hi_v) { vmn = fminf(vmn, v); vmx = fmaxf(vmx, v); }
                            }
                        vmn = fminf(vmn, __shfl_xor_sync(0xffffffffu, vmn, 1)); vmn = fminf(vmn, __shfl_xor_sync(0xffffffffu, vmn, 2));
                        vmx = fmaxf(vmx, __shfl_xor_sync(0xffffffffu, vmx, 1)); vmx = fmaxf(vmx, __shfl_xor_sync(0xffffffffu, vmx, 2));
                        if (e_lo) c_min = vmn - 2.0f * E;
                        if (e_hi) c_max = vmx + 2.0f * E;
                    }
                }
                okr = okr && (c_min > par || par == plow) && (c_max < pbr || pbr == GM_PMAX);
                // walk 2: entries below the window, candidates appended to the row's pool
                int my_below = 0, my_c = 0;
                for (int e0 = 0; e0 < n_w; e0 += 8) {
                    float2 e8[8];
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        const bool in = (e0 + 2 * h < n_e) && okr;
                        const float4 t4 = in ? *reinterpret_cast<const float4*>(ml + SP_LIDX(e0 + 2 * h)) : make_float4(0.f, 0.f, 0.f, 0.f);
                        e8[2 * h] = make_float2(in ? t4.x : SP_NO_ENTRY, t4.y);
                        e8[2 * h + 1] = make_float2((in && e0 + 2 * h + 1 < n_e) ? t4.z : SP_NO_ENTRY, t4.w);
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const float2 ent = e8[u];
                        if (ent.x < c_min) ++my_below;
                        else if (ent.x <= c_max) {                                  // a NaN (no entry) fails both compares
                            const int slot = atomicAdd(&s_nc[r], 1);
                            if (slot < 32) { pool_col[r * 32 + slot] = __float_as_int(ent.y); pool_dex[r * 32 + slot] = ent.x; }
                            ++my_c;
                        }
                    }
                }
                my_below += __shfl_xor_sync(0xffffffffu, my_below, 1); my_below += __shfl_xor_sync(0xffffffffu, my_below, 2);
                my_c += __shfl_xor_sync(0xffffffffu, my_c, 1); my_c += __shfl_xor_sync(0xffffffffu, my_c, 2);
                const int kx = pr.lo - (c_ar + my_below);
                if (my_c > 32 || kx < 0 || kx + (pr.hi - pr.lo) >= my_c) okr = false;
                __syncwarp();
                if (live && sub == 0 && !s_flag[r]) {
                    if (okr) { s_nc[r] = my_c; s_kx[r] = kx; }
                    else { s_nc[r] = 0; s_flag[r] = 1; atomicAdd(&counters[1], 1ull); atomicAdd(&counters[12], 1ull); }
                }
            }
        }
#pragma unroll
        for (int k = 0; k < SP_ARAW_PER_THREAD; ++k) {
            const int e = tid + k * SP_SEL_THREADS;
            if (e < (GM_ROWS + 8) * ACOSS_NBINS) Araw[e] = areg[k];
        }
        if (tid < GM_ROWS) s_xx[tid] = xreg;
        __syncthreads();
        // offsets of the rows' candidates in the flat task list
        if (warp == 0) {
            int c[4], tot = 0;
#pragma unroll
            for (int q = 0; q < 4; ++q) { const int r = 4 * lane + q; c[q] = (r < tk.nrows) ? s_nc[r] : 0; tot += c[q]; }
            int incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t_ = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t_; }
            int run = incl - tot;
#pragma unroll
            for (int q = 0; q < 4; ++q) { s_off[4 * lane + q] = run; run += c[q]; }
            if (lane == 31) s_off[GM_ROWS] = run;
        }
        __syncthreads();
        // ---- (B) thread per candidate (all rows of the strip pooled): the exact 108-term chain in the oracle's op order
        {
            const int boff = side_a_is_query ? (ACOSS_NBINS - rot) % ACOSS_NBINS : 0;     // rot_src(c, rot) = (c + boff) % 12
            const int n_tasks_c = s_off[GM_ROWS];
            for (int idx = tid; idx < n_tasks_c; idx += SP_SEL_THREADS) {
                int r = 0;                                            // last row with s_off[r] <= idx
#pragma unroll
                for (int st = 64; st > 0; st >>= 1) if (s_off[r + st] <= idx) r += st;
                const int slot = idx - s_off[r];
                const int jj = pool_col[r * 32 + slot];
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
                max_err = fmaxf(max_err, fabsf(item - pool_dex[r * 32 + slot]) / (sa + sb));
                if (item < 0.0f) item = 0.0f;
                pool_dex[r * 32 + slot] = __fsqrt_rn(item);
                ++n_exact;
            }
        }
        __syncthreads();
        // ---- (C) warp per row: ranks kx and kx + 1 among the row's exact candidate distances (ties by slot)
        for (int r = warp; r < tk.nrows; r += SP_SEL_WARPS) {
            const int n_c = s_nc[r], kx = s_kx[r];
            if (n_c == 0) continue;
            const float da = (lane < n_c) ? pool_dex[r * 32 + lane] : INFINITY;
            int rank = 0;
            for (int b_ = 0; b_ < n_c; ++b_) {
                const float db = __shfl_sync(0xffffffffu, da, b_);
                rank += (db < da) || (db == da && b_ < lane);
            }
            const unsigned m_lo = __ballot_sync(0xffffffffu, lane < n_c && rank == kx);
            const unsigned m_hi = __ballot_sync(0xffffffffu, lane < n_c && rank == kx + 1);
            const float v_lo = __shfl_sync(0xffffffffu, da, __ffs(m_lo) - 1);
            const float v_hi = m_hi ? __shfl_sync(0xffffffffu, da, __ffs(m_hi) - 1) : v_lo;
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
