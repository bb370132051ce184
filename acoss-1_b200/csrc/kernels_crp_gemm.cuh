// kernels_crp_gemm.cuh -- bit-exact CRP with the cross-similarity on the tensor cores (sm_100a, crp_mode 4).
//
// Same contract and the same exactness argument as kernels_crp_band.cuh (approximate radicand rho~,
// rigorous error band E, exact re-check of everything inside the band), but rho~ of a whole
// 128 x 256 block of cells is ONE accumulated tcgen05 GEMM:
//     rho~(i, j) = |x_i|^2 + |y_j|^2 - 2 <x_i, y_j>,     x_i, y_j = 108-d delay embeddings (m = 9, tau = 1)
// in TF32 with the split x = x_hi + x_lo (three products hh + hl + lh, the lo.lo term is inside E).
//
// Hankel operands.  Embedded row r is frames r .. r+8, so the K-major operand "row r, 16-byte chunk
// (k, c3)" (time offset k < 9, bins 4 c3 .. 4 c3 + 3) is chunk c3 of frame r + k.  Shared memory holds
// every frame ONCE, as six planes (hi / lo x c3) of 16-byte entries; the UMMA descriptor of the
// k-slice {(k, c3), (k + 1, c3)} simply starts at plane + 16 k with LBO = 16: rows are 16 B apart
// (the canonical no-swizzle core matrix), 8-row groups 128 B apart (SBO).  Nothing is expanded nine
// times, neither in shared memory nor on the way from L2.  Per 128 x 256 tile: 36 MMAs (3 products x
// 3 planes x 4 offset pairs) + 5 MMAs for offset 8 and the norm chunk = 41 x (M 128, N 256, K 8).
//
// Rows of the strip are TMEM lanes, so after tcgen05.ld a thread owns ONE strip row: counting against
// its row's pivots, list insertion and binarisation are thread-local, no shuffles, no shared-memory
// strip buffer.  Accumulators are double-buffered in TMEM (2 x 256 columns): the epilogue of tile t
// overlaps the MMAs of tile t + 1.
//
// Per strip (128 rows of side A x all NB rows of side B):
//   sample    256 spread columns, hi.hi product only (14 MMAs on an explicitly gathered operand)
//             -> per-row pivots pa <= pb bracketing the wanted order statistics (sample histogram)
//   count     stream all tiles: c_a = #(rho~ < pa), elements in [pa, pb) go to the row's list (global
//             scratch, L2-resident)
//   select    warp per row on its list: value bisection, candidates within 2E, exact 108-term chains
//             (the oracle's op order), exact order statistics -> threshold
//   binarise  (pass 2 only) stream all tiles again: rho~ against the band edges of both thresholds,
//             ambiguous cells decided exactly afterwards; CRP bytes go straight from registers to HBM.
#pragma once
#include "common.cuh"
#include "kernels_crp.cuh"
#include "kernels_crp_tc.cuh"
#include "kernels_crp_band.cuh"

#define GM_ROWS 128                   // strip rows = MMA M = TMEM lanes
#define GM_TN 256                     // side-B embedded rows per tile = MMA N = TMEM columns per buffer
#define GM_THREADS 512
#define GM_WARPS (GM_THREADS / 32)
#define GM_BLOCK (GM_THREADS + 128)   // + one MMA-issuing warp (tcgen05.mma issue blocks ~100 clk per instruction while the pipe is
                                      // busy); warps are allocated in fours, the other three exit at once
#ifndef GM_S
#define GM_S 512                      // sample columns per row (MMA blocks of 256); 256 overflows the per-owner lists
#endif
#define GM_SHALVES (GM_S / GM_TN)
#define GM_SROUND (32768 / GM_S)       // strip rows whose samples fit the 128 KB dump buffer at once
#define GM_SPL (GM_S / 32)
#define GM_SUBCAP 128                 // list capacity per (row, column group): private to the owning thread, no atomics
#define GM_LCAP (4 * GM_SUBCAP)        // list slots per row
#define GM_LPL (GM_LCAP / 32)
#define GM_EPS_REL 1.220703125e-04f   // 2^-13, see the error budget below
#define GM_NSIGMA 2.5f                   // bracket half-width in binomial sigmas.  Measured (L = 2057): on music-like frames the spread sample brackets far tighter than binomial (1.0 sigma missed 64 of 8.3 M rows), on white-noise frames 2.0 sigma misses 1 % of the rows and each miss costs an exact row; 2.5 is within 2 % of the best on both
#define GM_MIN_NB 128                 // below this every row takes the exact-row path
#define GM_MAX_NB 16384
#define GM_APLANE ((GM_ROWS + 8) * 16)
#define GM_BPLANE ((GM_TN + 8) * 16)
#define GM_A_BYTES (6 * GM_APLANE + GM_ROWS * 16)       // 15 104
#define GM_B_BYTES (6 * GM_BPLANE + GM_TN * 16)         // 29 440
#define GM_SMP_CHUNKS 28
#define GM_SMP_PSTRIDE (GM_TN * 16 + 16)   // plane stride of the gathered sample operand: skewed by one 16-byte entry, so that a warp
                                           // writing the 28 chunk planes of ONE sample spreads over all bank groups
#define GM_R1_BYTES 131072                     // 131 072: sample operand / sample dump (64 rows) / 3 B buffers + edges + ambiguity list / scratch
#define GM_EDGE_OFF (3 * GM_B_BYTES)                    // column band edges lo[4][256] (floats), slot = tile % 4
#define GM_AMB_OFF 93184                                // ambiguity list (ints)
#define GM_AMB_CAP ((GM_R1_BYTES - GM_AMB_OFF) / 4)
#define GM_WSCR_WORDS 512                               // per-warp scratch: hist[256] | arow[108] cand[32]
#define GM_ROWV_BYTES (GM_ROWS * 4)
#define GM_ARAW_BYTES ((GM_ROWS + 8) * ACOSS_NBINS * 4)  // strip frames in the oracle's bin order, for the exact chains
// shared memory: A planes | R1 | per-warp scratch | per-row arrays (pa pb thA xx loq hiq: float; cnta lcnt flag: int) | 2 mbarriers | tmem holder
#define GM_SMP_HALF_BYTES (GM_SMP_CHUNKS * GM_SMP_PSTRIDE)          // 115 136: one half (256 samples) of a track's sample operand
#define GM_SMP_OP_BYTES (GM_SHALVES * GM_SMP_HALF_BYTES)           // per track, precomputed at upload (gm_sample_operand_kernel)
#define GM_SMEM_BYTES (GM_A_BYTES + GM_R1_BYTES + GM_WARPS * GM_WSCR_WORDS * 4 + 12 * GM_ROWV_BYTES + GM_ARAW_BYTES + 128)
#define GM_TMEM_COLS 512
// list entry (value, column) of slot s of a sub-list, relative to the owner's base (region + 2 * row): slots are stored in pairs,
// [slot / 2][row][slot % 2], so that a thread reads two consecutive slots with one 16-byte load and a warp 512 contiguous bytes
#define GM_LIDX(s) ((((s) >> 1) * (2 * GM_ROWS)) + ((s) & 1))
#define GM_POOL(r, slot) ((r) * 32 + (((slot) + (r)) & 31))   // candidate pool index: thread-per-row access is bank-conflict-free
#define GM_BIG 1.0e30f                 // norm entry of padding columns: rho~ = huge, never below a pivot or an edge
#define GM_PMAX 1.0e29f                // "unbounded" pivots
#define GM_IDESC ((1u << 4) | (2u << 7) | (2u << 10) | ((GM_TN >> 3) << 17) | ((GM_ROWS >> 4) << 24))

// Error budget of rho~ against the oracle's radicand, in units of u (xx + yy), u = 2^-24:
//   oracle's own rounding (kernels_crp_band.cuh)                                       220
//   norms taken from the oracle's self-dot tables (108-term chains) + hi/lo residual    112
//   dropped lo.lo and residual cross terms of the three-product TF32 split               12
//   fp32 accumulation inside the tensor core, assuming no better than truncation of
//   every addend to the largest exponent: 41 MMAs x 9 addends x 2u x |partial| <= 2(xx+yy)   1476
//   total 1820 u < 2048 u = 2^-13.   counters[14] records the largest error actually observed.

__device__ __forceinline__ void gm_tmem_alloc(uint32_t* holder_smem)
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(holder_smem)), "n"(GM_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void gm_tmem_free(uint32_t taddr)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(GM_TMEM_COLS) : "memory");
}
// One MMA.  The two shared-memory descriptors differ from call to call only in their low words (start address >> 4 in
// bits 0-13, LBO >> 4 in bits 16-29); the high word (SBO = 128 B, descriptor version 1, no swizzle) is common.
#define GM_DESC_HI (((128u >> 4) & 0x3FFFu) | (1u << 14))
__device__ __forceinline__ void gm_mma(uint32_t d_tmem, uint32_t adesc_lo, uint32_t bdesc_lo, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 da, {%1, %5};\n\t"
        "mov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %3, p;\n\t"
        "}\n" ::"r"(d_tmem), "r"(adesc_lo), "r"(bdesc_lo), "r"((uint32_t)GM_IDESC), "r"(accumulate), "r"((uint32_t)GM_DESC_HI) : "memory");
}
// Same for bf16 operands (kind::f16, K = 16 per instruction = two 16-byte chunks of eight values), fp32 accumulation in the same TMEM tile
#define GM_IDESC_BF16 ((1u << 4) | (1u << 7) | (1u << 10) | ((GM_TN >> 3) << 17) | ((GM_ROWS >> 4) << 24))
__device__ __forceinline__ void gm_mma_bf16(uint32_t d_tmem, uint32_t adesc_lo, uint32_t bdesc_lo, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "mov.b64 da, {%1, %5};\n\t"
        "mov.b64 db, {%2, %5};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t"
        "}\n" ::"r"(d_tmem), "r"(adesc_lo), "r"(bdesc_lo), "r"((uint32_t)GM_IDESC_BF16), "r"(accumulate), "r"((uint32_t)GM_DESC_HI) : "memory");
}
// low descriptor word of (base + byte offset, LBO): additive in the offset because shared memory addresses stay below 256 KB
#define GM_DLO(base_lo, off, lbo) ((base_lo) + (((uint32_t)(off)) >> 4) + ((((uint32_t)(lbo)) >> 4) << 16))
// barrier over the 16 consumer warps only (the producer warp never joins it)
__device__ __forceinline__ void gm_csync() { asm volatile("bar.sync 1, 512;" ::: "memory"); }
__device__ __forceinline__ int gm_csync_or(int pred)
{
    int r;
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "setp.ne.b32 p, %1, 0;\n\t"
        "bar.red.or.pred q, 1, 512, p;\n\t"
        "selp.b32 %0, 1, 0, q;\n\t"
        "}\n" : "=r"(r) : "r"(pred) : "memory");
    return r;
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool gm_elect_one()
{
    uint32_t pred;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}\n" : "=r"(pred));
    return pred != 0;
}

// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void gm_tmem_ld32(uint32_t taddr, float (&v)[32])
{
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int q = 0; q < 32; ++q) v[q] = __uint_as_float(r[q]);
}

// the 41 MMAs of one 128 x 256 tile.  a0 / b0: shared-memory byte addresses of the A planes / of the B buffer.
// A plane order: H0 H1 H2 L0 L1 L2 NRM;  B plane order: L0 L1 L2 H0 H1 H2 NRM (so that every paired
// chunk of the offset-8 leftovers has a non-negative LBO on both operands).
__device__ __forceinline__ void gm_issue_tile(uint32_t d_tmem, uint32_t a0, uint32_t b0)
{
    const uint32_t al = (a0 & 0x3FFFFu) >> 4, bl = (b0 & 0x3FFFFu) >> 4;
    uint32_t acc = 0;
#pragma unroll
    for (int prod = 0; prod < 3; ++prod) {              // 0: hi.hi  1: hi.lo  2: lo.hi
#pragma unroll
        for (int c3 = 0; c3 < 3; ++c3) {
            const int ap = ((prod == 2 ? 3 : 0) + c3) * GM_APLANE;
            const int bp = ((prod == 1 ? 0 : 3) + c3) * GM_BPLANE;
#pragma unroll
            for (int kp = 0; kp < 4; ++kp) {
                gm_mma(d_tmem, GM_DLO(al, ap + 32 * kp, 16), GM_DLO(bl, bp + 32 * kp, 16), acc);
                acc = 1;
            }
        }
    }
    const int aH0 = 128, aH2 = 2 * GM_APLANE + 128, aL0 = 3 * GM_APLANE + 128;
    const int bL0 = 128, bL2 = 2 * GM_BPLANE + 128, bH0 = 3 * GM_BPLANE + 128, bH2 = 5 * GM_BPLANE + 128;
    gm_mma(d_tmem, GM_DLO(al, aH0, GM_APLANE), GM_DLO(bl, bH0, GM_BPLANE), 1);                  // (H0,H0) (H1,H1)
    gm_mma(d_tmem, GM_DLO(al, aH0, GM_APLANE), GM_DLO(bl, bL0, GM_BPLANE), 1);                  // (H0,L0) (H1,L1)
    gm_mma(d_tmem, GM_DLO(al, aL0, GM_APLANE), GM_DLO(bl, bH0, GM_BPLANE), 1);                  // (L0,H0) (L1,H1)
    gm_mma(d_tmem, GM_DLO(al, aH2, 3 * GM_APLANE), GM_DLO(bl, bL2, 3 * GM_BPLANE), 1);          // (H2,L2) (L2,H2)
    gm_mma(d_tmem, GM_DLO(al, aH2, 4 * GM_APLANE - 128), GM_DLO(bl, bH2, GM_BPLANE - 128), 1);  // (H2,H2) norms
}

// the 14 MMAs of a sample block: A = hi planes, B = gathered sample operand, chunk column q at s0 + q * 4096:
// q = 8 c3 + k (k < 8), 24..26 = (c3, 8), 27 = norms
__device__ __forceinline__ void gm_issue_sample(uint32_t d_tmem, uint32_t a0, uint32_t s0)
{
    const uint32_t al = (a0 & 0x3FFFFu) >> 4, sl = (s0 & 0x3FFFFu) >> 4;
    uint32_t acc = 0;
#pragma unroll
    for (int c3 = 0; c3 < 3; ++c3) {
#pragma unroll
        for (int kp = 0; kp < 4; ++kp) {
            gm_mma(d_tmem, GM_DLO(al, c3 * GM_APLANE + 32 * kp, 16), GM_DLO(sl, (8 * c3 + 2 * kp) * GM_SMP_PSTRIDE, GM_SMP_PSTRIDE), acc);
            acc = 1;
        }
    }
    gm_mma(d_tmem, GM_DLO(al, 128, GM_APLANE), GM_DLO(sl, 24 * GM_SMP_PSTRIDE, GM_SMP_PSTRIDE), 1);
    gm_mma(d_tmem, GM_DLO(al, 2 * GM_APLANE + 128, 4 * GM_APLANE - 128), GM_DLO(sl, 26 * GM_SMP_PSTRIDE, GM_SMP_PSTRIDE), 1);
}

// One frame (12 terms) of the oracle's sequential chain: acc += a[c] * b[(c + OFF) % 12], c ascending, no FMA.
template <int OFF>
__device__ __forceinline__ float gm_chain12(float acc, const float (&a)[12], const float (&b)[12])
{
#pragma unroll
    for (int c = 0; c < 12; ++c) acc = xadd(acc, xmul(a[c], b[(c + OFF) % 12]));
    return acc;
}
__device__ __forceinline__ float gm_chain12_rot(float acc, const float (&a)[12], const float (&b)[12], int off)
{
    switch (off) {          // uniform per task
    case 0: return gm_chain12<0>(acc, a, b);
    case 1: return gm_chain12<1>(acc, a, b);
    case 2: return gm_chain12<2>(acc, a, b);
    case 3: return gm_chain12<3>(acc, a, b);
    case 4: return gm_chain12<4>(acc, a, b);
    case 5: return gm_chain12<5>(acc, a, b);
    case 6: return gm_chain12<6>(acc, a, b);
    case 7: return gm_chain12<7>(acc, a, b);
    case 8: return gm_chain12<8>(acc, a, b);
    case 9: return gm_chain12<9>(acc, a, b);
    case 10: return gm_chain12<10>(acc, a, b);
    default: return gm_chain12<11>(acc, a, b);
    }
}

// ---- per-track sample operand, built once at upload ---------------------------------------------------------------
// The sample columns of a strip depend on side B only (positions spread over its NB embedded rows), so the gathered,
// TF32-rounded K-major operand of every track is precomputed in its final shared-memory layout and a strip fetches it
// with one bulk copy per half (TMA engine, no LSU requests: the per-SM LSU path from L2 sustains only ~15 B/clk here).
// Norm entries use the unrotated self-dot table: the sample only steers the pivots, which the count pass verifies.
__global__ void gm_sample_operand_kernel(const float* __restrict__ frames, const int64_t* __restrict__ offs,
                                         const int32_t* __restrict__ emb_len, const int64_t* __restrict__ sd_off,
                                         const float* __restrict__ selfdot, unsigned char* __restrict__ out)
{
    const int track = blockIdx.x;
    const int NB = emb_len[track];
    if (NB <= 0) return;
    const float* fr = frames + offs[track] * ACOSS_NBINS;
    const float* sd = selfdot + sd_off[track];
    unsigned char* o = out + (size_t)track * GM_SMP_OP_BYTES;
    for (int e = threadIdx.x; e < GM_S * GM_SMP_CHUNKS; e += blockDim.x) {
        const int idx = e / GM_SMP_CHUNKS, q = e - idx * GM_SMP_CHUNKS;      // sample, chunk plane
        const int half = idx / GM_TN, s = idx - half * GM_TN;
        const int pos = (int)(((long long)idx * NB) / GM_S);
        float4 v;
        if (q < 27) {
            const int c3 = q < 24 ? (q >> 3) : q - 24, k = q < 24 ? (q & 7) : 8;
            const float4 x = *reinterpret_cast<const float4*>(fr + (int64_t)(pos + k) * ACOSS_NBINS + 4 * c3);
            v = make_float4(tf32_rna(x.x), tf32_rna(x.y), tf32_rna(x.z), tf32_rna(x.w));
        } else {
            v.x = 1.0f; v.y = 1.0f;
            tf32_split(sd[pos], v.z, v.w);
        }
        *reinterpret_cast<float4*>(o + (size_t)half * GM_SMP_HALF_BYTES + q * GM_SMP_PSTRIDE + s * 16) = v;
    }
}
__device__ __forceinline__ void gm_bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// counters: [0] exact cells  [1] fallback rows  [2..6] cycles (thread 0): stage+sample, pivots, count pass, select, binarise
//           [7] tasks  [11] too many candidates  [12] window not inside the list  [13] bracket miss  [15] list overflow
//           [14] max observed |rho~ - rho_oracle| / (xx + yy) (float bits)
__global__ void __launch_bounds__(GM_BLOCK, 1)
crp_strip_gemm_kernel(const PairDesc* __restrict__ descs,
                      const StripTask* __restrict__ tasks, int n_tasks,
                      const float* __restrict__ frames,
                      const float* __restrict__ selfdot,
                      float* __restrict__ thr, uint8_t* __restrict__ crp,
                      int side_a_is_query, float qf,
                      float2* __restrict__ lists,            // [gridDim.x][4][GM_SUBCAP][GM_ROWS] (rho~, column) of the listed elements
                      const unsigned char* __restrict__ smpop, // per-track sample operands (gm_sample_operand_kernel) or null: gather in the kernel
                      unsigned long long* __restrict__ counters, int dbg_arg)
{
#ifdef ACOSS_PROFILE_BUILD
    const int dbg = dbg_arg;            // kernel-timing experiments (skip phases; results are wrong): profile builds only
#else
    (void)dbg_arg;
    constexpr int dbg = 0;              // release: every `dbg &` branch folds away
#endif

    extern __shared__ __align__(128) unsigned char gsm[];
    unsigned char* Apl = gsm;
    unsigned char* R1 = gsm + GM_A_BYTES;
    float* rowv = reinterpret_cast<float*>(R1 + GM_R1_BYTES + GM_WARPS * GM_WSCR_WORDS * 4);
    float* s_pa = rowv;
    float* s_pb = rowv + GM_ROWS;
    float* s_th = rowv + 2 * GM_ROWS;
    float* s_xx = rowv + 3 * GM_ROWS;
    float* s_loq = rowv + 4 * GM_ROWS;
    float* s_hiq = rowv + 5 * GM_ROWS;
    int* s_cnta = reinterpret_cast<int*>(rowv + 6 * GM_ROWS);
    int* s_lcnt = s_cnta + GM_ROWS;
    int* s_nc = reinterpret_cast<int*>(s_loq);                           // candidates per row / rank among them: live before the
    int* s_kx = reinterpret_cast<int*>(s_hiq);                           // binarise pass sets the band edges
    int* s_flag = s_lcnt + 4 * GM_ROWS;                                  // s_lcnt: [row][column group]
    float* Araw = reinterpret_cast<float*>(s_flag + GM_ROWS);            // [GM_ROWS + 8][12]
    // mbarriers: full[2] (MMAs of a tile done, by tcgen05.commit), empty[2] (a TMEM buffer has been read; one arrival per
    // warp), built[3] (a side-B operand buffer has been written; one arrival per warp), smp (sample MMAs done)
    uint64_t* mbar = reinterpret_cast<uint64_t*>(Araw + (GM_ROWS + 8) * ACOSS_NBINS);
    uint64_t* mb_full = mbar, *mb_empty = mbar + 2, *mb_built = mbar + 4, *mb_smp = mbar + 7, *mb_ld = mbar + 8;
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(mbar + 9);
    float* s_red = reinterpret_cast<float*>(tmem_holder + 2);            // 8 words: ambiguity counter, strip error bound for the producer
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int lq = warp & 3, cg = warp >> 2;          // TMEM lane quarter (strip rows 32 lq ..), column group (64 cg ..)
    const int myrow = 32 * lq + lane;
    unsigned long long n_exact = 0;
    uint32_t pf0 = 0, pf1 = 0, psmp = 0, pld = 0;     // mbarrier phase parities: full (every thread), sample, sample-operand bulk load
    float max_err = 0.0f;

    if (warp == 0) gm_tmem_alloc(tmem_holder);
    if (tid == 0) {
        mbar_init(mb_full, 1); mbar_init(mb_full + 1, 1); mbar_init(mb_empty, GM_WARPS); mbar_init(mb_empty + 1, GM_WARPS);
        mbar_init(mb_built, GM_WARPS); mbar_init(mb_built + 1, GM_WARPS); mbar_init(mb_built + 2, GM_WARPS); mbar_init(mb_smp, 1); mbar_init(mb_ld, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
    const uint32_t my_taddr = tmem_base + ((uint32_t)(32 * lq) << 16) + 16 * cg;
    const uint32_t a_saddr = smem_u32(Apl), r1_saddr = smem_u32(R1);
    float2* my_lists = lists + (size_t)blockIdx.x * GM_ROWS * GM_LCAP;

    if (warp > GM_WARPS) {
        // spare warps of the allocation quad: nothing to do
    } else if (warp == GM_WARPS) {
        // ================= MMA-issuing warp =================
        // Mirrors the consumers' sequence of streaming passes (count pass unless the pair is tiny, binarise pass in pass 2).
        // Per tile: wait until all 16 consumer warps have written their share of the side-B planes (built) and have read
        // the TMEM buffer's previous contents (empty), then issue the 41 MMAs from one elected lane and commit to full.
        uint32_t pbi = 0, pe0 = 1, pe1 = 1;               // a fresh mbarrier passes a wait on parity 1
        for (int task = blockIdx.x; task < n_tasks; task += gridDim.x) {
            const StripTask tk = tasks[task];
            const PairDesc pd = descs[tk.pair];
            const int NB = side_a_is_query ? pd.N : pd.M;
            const int n_tiles = (NB + GM_TN - 1) / GM_TN;
            const bool tiny = NB < GM_MIN_NB;
            for (int pass = 0; pass < 2; ++pass) {
                if (pass == 0 ? tiny : !side_a_is_query) continue;
                for (int t = 0; t < n_tiles; ++t) {
                    const int buf = t & 1, sb = t % 3;
                    mbar_wait(mb_built + sb, (pbi >> sb) & 1u); pbi ^= 1u << sb;
                    if (buf == 0) { mbar_wait(mb_empty, pe0); pe0 ^= 1; } else { mbar_wait(mb_empty + 1, pe1); pe1 ^= 1; }
                    tc_fence_after();
                    if (gm_elect_one()) {
                        if (!(dbg & 2)) gm_issue_tile(tmem_base + buf * GM_TN, a_saddr, r1_saddr + sb * GM_B_BYTES);
                        umma_commit(mb_full + buf);
                    }
                    __syncwarp();
                }
            }
        }
    } else {
    for (int task = blockIdx.x; task < n_tasks; task += gridDim.x) {
        const StripTask tk = tasks[task];
        const PairDesc pd = descs[tk.pair];
        const int rot = pd.oti % ACOSS_NBINS;
        const int NB = side_a_is_query ? pd.N : pd.M;
        const int na_frames = side_a_is_query ? pd.nq : pd.nr;
        const int nb_frames = side_a_is_query ? pd.nr : pd.nq;
        const int64_t a_f0 = side_a_is_query ? pd.q_frame_off : pd.r_frame_off;
        const int64_t b_f0 = side_a_is_query ? pd.r_frame_off : pd.q_frame_off;
        // <roll(ref, rot), query>: the B frames stay unrotated; the A frames are staged pre-rotated:
        // by +rot when A is the reference, by -rot when A is the query.
        const int rot_a = side_a_is_query ? (ACOSS_NBINS - rot) % ACOSS_NBINS : rot;
        const float* sdA = selfdot + (side_a_is_query ? pd.q_sd_off : pd.r_sd_off + (int64_t)rot * pd.N);
        const float* sdB = selfdot + (side_a_is_query ? pd.r_sd_off + (int64_t)rot * pd.N : pd.q_sd_off);
        const float* fr_a = frames + a_f0 * ACOSS_NBINS;
        const float* fr_b = frames + b_f0 * ACOSS_NBINS;
        const int n_tiles = (NB + GM_TN - 1) / GM_TN;
        const bool rowvalid = myrow < tk.nrows;
        const PercentileRank pr = percentile_rank(NB, qf);
        float* thr_out = thr + (side_a_is_query ? pd.thr_q_off : pd.thr_r_off);
        const bool tiny = NB < GM_MIN_NB;

        gm_csync();
        long long ph0 = 0, ph1 = 0, ph2 = 0, ph3 = 0, ph4 = 0;
        long long prb = 0;
        if (tid == 0) { ph0 = clock64(); ph1 = ph0; ph2 = ph0; prb = ph0; }
#define GM_PROBE(i) do { if (tid == 0) { const long long c_ = clock64(); atomicAdd(&counters[16 + (i)], (unsigned long long)(c_ - prb)); prb = c_; } } while (0)
        // ---- stage the A planes: entry e = frame row0 + e (e < 136), chunk c3: -2 * rotated bins, hi / lo; norms
        for (int e = tid; e < (GM_ROWS + 8) * 3; e += GM_THREADS) {
            const int fe = e / 3, c3 = e - 3 * fe;
            const int f = tk.row0 + fe;
            float4 hi = make_float4(0.f, 0.f, 0.f, 0.f), lo = hi;
            if (f < na_frames) {
                const float* fp = fr_a + (int64_t)f * ACOSS_NBINS;
                tf32_split(-2.0f * fp[rot_src(4 * c3 + 0, rot_a)], hi.x, lo.x);
                tf32_split(-2.0f * fp[rot_src(4 * c3 + 1, rot_a)], hi.y, lo.y);
                tf32_split(-2.0f * fp[rot_src(4 * c3 + 2, rot_a)], hi.z, lo.z);
                tf32_split(-2.0f * fp[rot_src(4 * c3 + 3, rot_a)], hi.w, lo.w);
            }
            *reinterpret_cast<float4*>(Apl + c3 * GM_APLANE + fe * 16) = hi;
            *reinterpret_cast<float4*>(Apl + (3 + c3) * GM_APLANE + fe * 16) = lo;
        }
        for (int e = tid; e < (GM_ROWS + 8) * ACOSS_NBINS; e += GM_THREADS) {
            const int fe = e / ACOSS_NBINS, c = e - fe * ACOSS_NBINS;
            const int f = tk.row0 + fe;
            Araw[e] = (f < na_frames) ? fr_a[(int64_t)f * ACOSS_NBINS + (side_a_is_query ? c : rot_src(c, rot))] : 0.0f;
        }
        if (tid < GM_ROWS) {
            float4 nv = make_float4(0.f, 0.f, 0.f, 0.f);
            float xx = 0.0f;
            if (tid < tk.nrows) { xx = sdA[tk.row0 + tid]; tf32_split(xx, nv.x, nv.y); nv.z = 1.0f; nv.w = 1.0f; }
            *reinterpret_cast<float4*>(Apl + 6 * GM_APLANE + tid * 16) = nv;
            s_xx[tid] = xx; s_flag[tid] = 0; s_cnta[tid] = 0;
            s_pa[tid] = 0.0f; s_pb[tid] = 0.0f;
        }
        float ymax = 0.0f, xmax = 0.0f;
        for (int j = tid; j < NB; j += GM_THREADS) ymax = fmaxf(ymax, __ldg(sdB + j));
        if (tid < tk.nrows) xmax = sdA[tk.row0 + tid];
        for (int o = 16; o > 0; o >>= 1) {
            ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
            xmax = fmaxf(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
        }
        float* red = reinterpret_cast<float*>(R1);       // R1 is free here
        if (lane == 0) { red[warp] = ymax; red[GM_WARPS + warp] = xmax; }
        gm_csync();
        ymax = red[0]; xmax = red[GM_WARPS];
#pragma unroll
        for (int w = 1; w < GM_WARPS; ++w) { ymax = fmaxf(ymax, red[w]); xmax = fmaxf(xmax, red[GM_WARPS + w]); }
        const float E = GM_EPS_REL * (xmax + ymax);       // one bound for the whole strip (conservative)
        const float plow = -(4.0f * (xmax + ymax) + 1.0e-30f);    // "unbounded below" pivot: under every rho~, small enough for rho~ - plow to be meaningful
        gm_csync();
        GM_PROBE(0);

        if (!tiny) {
            // ---- sample: gather GM_S spread columns (blocks of 256) into an explicit K-major operand (hi parts only; the pivots only
            //      have to be roughly right, the count pass verifies the bracket), 14 MMAs per half into TMEM buffer 0 / 1
            const unsigned char* smp_src = smpop ? smpop + (size_t)(side_a_is_query ? pd.r : pd.q) * GM_SMP_OP_BYTES : nullptr;
            for (int half = 0; half < GM_SHALVES; ++half) {
                if (smp_src) {
                    // precomputed operand of the side-B track: one bulk copy (R1 is free: the barrier above / the wait on the
                    // previous half's MMAs), completion on an mbarrier
                    if (tid == 0) gm_bulk_load(R1, smp_src + (size_t)half * GM_SMP_HALF_BYTES, GM_SMP_HALF_BYTES, mb_ld);
                    mbar_wait(mb_ld, pld); pld ^= 1;
                } else {
                    // A warp gathers 16 samples of the half, one sample per load instruction: lanes 0 .. 26 read the 27
                    // consecutive 16-byte chunks of frames pos .. pos + 8 (432 contiguous bytes = 4 or 5 L1 tag lookups; one
                    // chunk per thread at scattered addresses cost 27, and the L1 tag stage was the bound of this phase), lane
                    // 27 the norm.  Memory chunk m = 3 k + c3 goes to operand plane q = 8 c3 + k (k < 8) or 24 + c3 (k = 8).
                    static_assert(GM_THREADS == 2 * GM_TN && GM_SMP_CHUNKS == 28, "sample gather mapping");
                    static_assert(GM_SMP_CHUNKS * GM_SMP_PSTRIDE <= GM_R1_BYTES, "sample operand");
                    const int mk = lane / 3, mc3 = lane - 3 * mk;
                    const int q = lane < 24 ? 8 * mc3 + mk : lane;          // lanes 24 .. 26: k = 8, q = 24 + c3 = lane; lane 27: norms
#pragma unroll
                    for (int i0 = 0; i0 < 16; i0 += 8) {
                        // every lane issues the same unconditional loads (lanes 27 .. 31 re-read chunk 26, the norm is a
                        // broadcast load), so that all 16 are in flight at once; the lane's role is resolved by selects
                        float4 x[8];
                        float nrm[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int s = 16 * warp + i0 + i;
                            const int pos = (int)(((long long)(half * GM_TN + s) * NB) / GM_S);
                            x[i] = __ldg(reinterpret_cast<const float4*>(fr_b + (int64_t)pos * ACOSS_NBINS) + min(lane, 26));
                            nrm[i] = __ldg(sdB + pos);
                        }
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int s = 16 * warp + i0 + i;
                            float nh, nl;
                            tf32_split(nrm[i], nh, nl);
                            const bool isn = lane >= 27;
                            const float4 o = make_float4(isn ? 1.0f : tf32_rna(x[i].x), isn ? 1.0f : tf32_rna(x[i].y),
                                                         isn ? nh : tf32_rna(x[i].z), isn ? nl : tf32_rna(x[i].w));
                            if (lane < 28) *reinterpret_cast<float4*>(R1 + q * GM_SMP_PSTRIDE + s * 16) = o;
                        }
                    }
                }
                fence_proxy_async_smem();
                tc_fence_before();
                gm_csync();
                GM_PROBE(1);
                if (tid == 0) {
                    tc_fence_after();
                    gm_issue_sample(tmem_base + half * GM_TN, a_saddr, r1_saddr);
                    umma_commit(mb_smp);
                }
                mbar_wait(mb_smp, psmp); psmp ^= 1;     // the operand buffer is rebuilt for the second half
                tc_fence_after();
                GM_PROBE(2);
            }
            if (tid == 0) ph1 = clock64();
            // ---- pivots.  Per row the 256-bin histogram of its GM_S samples over [min, max] -> pa / pb at bin edges.  The four
            //      owner threads of a row bin their own 128 samples straight out of TMEM into PRIVATE histograms with 8-bit
            //      counters (64 words per thread, word w of lane l at hw[w * 32 + l]: bank = lane, conflict-free, no atomics --
            //      shared-memory atomics were the whole cost of this phase); the scan adds the four owners' counters.
            {
                static_assert(GM_S == 512 && GM_THREADS == 512, "four owners x 128 samples per row");
                unsigned* hall = reinterpret_cast<unsigned*>(R1);                       // [16 warps][64 words][32 lanes] = 128 KB
                float* xf = reinterpret_cast<float*>(R1 + GM_R1_BYTES);                 // [2][GM_THREADS] exchange (the per-warp scratch is free here)
                int* xc = reinterpret_cast<int*>(xf);
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
                gm_csync();
                {
                    const float4 a4 = *reinterpret_cast<const float4*>(xf + xi), b4 = *reinterpret_cast<const float4*>(xf + GM_THREADS + xi);
                    smin = fminf(fminf(a4.x, a4.y), fminf(a4.z, a4.w));
                    smax = fmaxf(fmaxf(b4.x, b4.y), fmaxf(b4.z, b4.w));
                }
                const float hscale = (smax > smin) ? 256.0f / (smax - smin) : 0.0f;
                const float hwidth = (smax - smin) * (1.0f / 256.0f);
                const float hoff = -smin * hscale;
                unsigned char* hb = reinterpret_cast<unsigned char*>(hall + warp * 2048) + 4 * lane;     // bin b: byte b % 4 of word b / 4
#pragma unroll
                for (int b = 0; b < GM_SHALVES; ++b)
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        float v[32];
                        gm_tmem_ld32(my_taddr + 48 * cg + b * GM_TN + 32 * h, v);
#pragma unroll
                        for (int q = 0; q < 32; ++q) {
                            // these phases are issue-bound (65 536 samples per strip): a byte-wide read-modify-write keeps
                            // the instruction count per sample at its minimum
                            const int bin = min(255, (int)fmaf(v[q], hscale, hoff));
                            hb[(bin >> 2) * 128 + (bin & 3)] += 1;
                        }
                    }
                tc_fence_before();
                gm_csync();
                GM_PROBE(3);
                const float frac = ((float)pr.lo + 0.5f) / (float)NB;
                const float rs = frac * (float)GM_S;
                const float sg = sqrtf(fmaxf(rs * (1.0f - frac), 1.0f));
                const int r_lo = (int)floorf(rs - GM_NSIGMA * sg) - 1;       // sample ranks the bracket must enclose
                const int r_hi = (int)ceilf(rs + GM_NSIGMA * sg) + 1;
                // owner cg scans bins 64 cg .. 64 cg + 63 (words 16 cg .. 16 cg + 15 of the four owners' histograms)
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
                gm_csync();
                int ba = -1, bb = 256;
                {
                    const int4 c = *reinterpret_cast<const int4*>(xc + xi);
                    const int cc[4] = {c.x, c.y, c.z, c.w};
                    int run = 0;
#pragma unroll
                    for (int g = 0; g < 4; ++g) if (g < cg) run += cc[g];
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
                gm_csync();                  // xc is rewritten
                xc[xi + cg] = ba; xc[GM_THREADS + xi + cg] = bb;
                gm_csync();
                if (cg == 0) {
                    const int4 a4 = *reinterpret_cast<const int4*>(xc + xi), b4 = *reinterpret_cast<const int4*>(xc + GM_THREADS + xi);
                    ba = max(max(a4.x, a4.y), max(a4.z, a4.w));
                    bb = min(min(b4.x, b4.y), min(b4.z, b4.w));
                    s_pa[myrow] = (r_lo >= 0 && ba > 0) ? fmaxf(smin + (float)ba * hwidth, plow) : plow;
                    s_pb[myrow] = (r_hi < GM_S && bb < 255) ? fminf(smin + (float)(bb + 1) * hwidth, GM_PMAX) : GM_PMAX;
                }
                gm_csync();
                GM_PROBE(4);
            }
        }
        if (tid == 0) ph2 = clock64();

        // ---- tile streaming.  Everything inside a pass is ordered by mbarriers, not CTA barriers: a warp runs the epilogue of
        //      tile t out of TMEM buffer t % 2, releases it (empty), writes its share of the side-B planes of tile t + 2 into
        //      shared-memory buffer (t + 2) % 3 (free: the MMAs of tile t - 1 are done) and signals built; the issuing warp
        //      waits for all shares and for the TMEM buffer and issues the MMAs of tile t + 2.  The MMAs of tile t + 1 run
        //      meanwhile.
        const float* thB = thr + pd.thr_r_off;            // column thresholds (pass 2)
        float* loC = reinterpret_cast<float*>(R1 + GM_EDGE_OFF);      // [4][256] lower band edges of the column thresholds, slot = tile % 4:
                                                                      // a warp writing tile t + 2 may run beside warps still reading tiles t - 1 and t
        // side-B planes of tile t (frames 256 t .. 256 t + 263, hi / lo split in registers), norms, column band edges
        auto build_b = [&](int t, bool edges) {
            const int sb = t % 3;
            unsigned char* Bb = R1 + sb * GM_B_BYTES;
            const int j0 = t * GM_TN;
            for (int e = tid; e < (GM_TN + 8) + GM_TN; e += GM_THREADS) {
                if (e < GM_TN + 8) {
                    const int f = j0 + e;
                    float4 x[3];
                    if (f < nb_frames && !(dbg & 16)) {
                        const float4* fp = reinterpret_cast<const float4*>(fr_b + (int64_t)f * ACOSS_NBINS);
                        x[0] = __ldg(fp); x[1] = __ldg(fp + 1); x[2] = __ldg(fp + 2);
                    } else {
                        x[0] = x[1] = x[2] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int c3 = 0; c3 < 3; ++c3) {
                        float4 hi, lo;
                        tf32_split(x[c3].x, hi.x, lo.x); tf32_split(x[c3].y, hi.y, lo.y);
                        tf32_split(x[c3].z, hi.z, lo.z); tf32_split(x[c3].w, hi.w, lo.w);
                        *reinterpret_cast<float4*>(Bb + c3 * GM_BPLANE + e * 16) = lo;
                        *reinterpret_cast<float4*>(Bb + (3 + c3) * GM_BPLANE + e * 16) = hi;
                    }
                } else {
                    const int c = e - (GM_TN + 8), j = j0 + c;
                    float4 nv = make_float4(1.0f, 1.0f, GM_BIG, 0.f);     // padding columns: rho~ huge
                    float lo = -INFINITY;                          // ... and never "in", always "out"
                    if (j < NB) {
                        tf32_split(__ldg(sdB + j), nv.z, nv.w);
                        if (edges) {
                            const float tb = __ldg(thB + j);
                            const float Tr = xmul(tb, tb);
                            lo = Tr - (E + Tr * 4.76837158203125e-07f);        // E + 8uT
                        }
                    }
                    *reinterpret_cast<float4*>(Bb + 6 * GM_BPLANE + c * 16) = nv;
                    if (edges) loC[(t & 3) * GM_TN + c] = lo;
                }
            }
            fence_proxy_async_smem();
        };
        auto built = [&](int t) { __syncwarp(); if (lane == 0) mbar_arrive(mb_built + t % 3); };
        auto stream = [&](auto&& epi, bool edges) {
            gm_csync();                                   // every warp is done with R1 and with TMEM
            build_b(0, edges); built(0);
            if (n_tiles > 1) { build_b(1, edges); built(1); }
            for (int t = 0; t < n_tiles; ++t) {
                const int buf = t & 1;
                if (buf == 0) { mbar_wait(mb_full, pf0); pf0 ^= 1; } else { mbar_wait(mb_full + 1, pf1); pf1 ^= 1; }
                tc_fence_after();
                // column ownership interleaved in runs of 16: thread (row, cg) reads columns 16 (4 i + cg) .. + 15, so that a
                // cluster of consecutive in-band columns spreads over the four owners of the row
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float v[16];
                    tmem_ld16(my_taddr + buf * GM_TN + 64 * i, v);
                    if (!(dbg & 1)) epi(t * GM_TN + 64 * i + 16 * cg, t & 3, 64 * i + 16 * cg, v);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(mb_empty + buf);
                if (t + 2 < n_tiles) { build_b(t + 2, edges); built(t + 2); }
            }
            gm_csync();
        };

        int ca = 0, ln = 0;      // count pass results of this thread: #(rho~ < pa), listed elements
        if (!tiny) {
            // ---- count pass: c_a = #(rho~ < pa); elements of [pa, pb) go to the row's list
            const float pa = s_pa[myrow], pb = s_pb[myrow];
            float2* ml = my_lists + cg * GM_SUBCAP * GM_ROWS + 2 * myrow;    // slot s at ml[GM_LIDX(s)]: private to this thread, no atomics;
                                                                              // rows are adjacent, so reading a slot is coalesced across a warp
            // d = rho~ - pa: "below" is the sign bit of d, "inside [pa, pb)" is one unsigned compare of d's bit pattern against
            // that of pb - pa (a negative d has the top bit set).  Elements within a rounding error of pb may be left out;
            // the selection only trusts the list at least 2E away from pb.
            const uint32_t wbits = __float_as_uint(pb - pa);
            stream([&](int jbase, int, int, const float (&v)[16]) {
                if (!rowvalid) return;
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const uint32_t u = __float_as_uint(v[q] - pa);
                    ca += u >> 31;
                    if (u < wbits) {
                        if (!(dbg & 8)) ml[GM_LIDX(min(ln, GM_SUBCAP - 1))] = make_float2(v[q], __int_as_float(jbase + q));     // overflow is detected from ln
                        ++ln;
                    }
                }
            }, false);
            if (rowvalid) s_lcnt[myrow * 4 + cg] = ln;
            if (rowvalid) atomicAdd(&s_cnta[myrow], ca);
            __threadfence_block();
            gm_csync();
        }
        if (tid == 0) ph3 = clock64();
        GM_PROBE(5);

        // ---- select.  (A) warp per row: value bisection on the row's list, candidates (everything within 2E of the
        //      wanted order statistics) go to a pool in shared memory;  (B) thread per candidate: the exact 108-term
        //      chain in the oracle's op order;  (C) warp per row: exact order statistics among its candidates.
        int* pool_col = reinterpret_cast<int*>(R1 + 98304);               // [GM_ROWS][32] candidate columns
        float* pool_dex = reinterpret_cast<float*>(R1 + 114688);          // [GM_ROWS][32] rho~ of the candidate, then its exact distance
        int* s_off = reinterpret_cast<int*>(R1 + GM_R1_BYTES + 4096);     // [GM_ROWS + 1] candidate offsets (behind the exchange buffers)
        if (!tiny) {
            // (A) four owner threads per row, each on its own sub-list (the elements it listed in the count pass; slots are
            //     row-adjacent in memory, so a warp reads them coalesced).  One binning pass refines the bracket: 128
            //     equal-width buckets over the list's value range (the list is close to uniform inside the narrow bracket, so
            //     a bucket holds about two elements) into PRIVATE histograms with 8-bit counters (a sub-list holds at most 128
            //     elements; word w of lane l of warp W at ph[(W * 32 + w) * 32 + l]: bank = lane, conflict-free, no atomics); a
            //     scan over the four owners' counters finds the bucket(s) of the wanted ranks; a second pass over the values
            //     (still in registers) collects the candidates: everything within 2E of those buckets.
            int* xc = reinterpret_cast<int*>(R1 + GM_R1_BYTES);           // [2][GM_THREADS] exchange buffers (the per-warp scratch is free here)
            float* xf = reinterpret_cast<float*>(xc);
            unsigned* hall = reinterpret_cast<unsigned*>(R1);             // [16 warps][32 words][32 lanes] = 64 KB
            unsigned char* pbyte = reinterpret_cast<unsigned char*>(hall + warp * 1024) + 4 * lane;      // bucket b: byte b % 4 of this thread's word b / 4
            const int xi = myrow * 4;
            int4 c4 = make_int4(0, 0, 0, 0);
            if (rowvalid) c4 = *reinterpret_cast<const int4*>(s_lcnt + 4 * myrow);
            const int c_a = s_cnta[myrow], c_ab = c4.x + c4.y + c4.z + c4.w;
            const bool ovf = max(max(c4.x, c4.y), max(c4.z, c4.w)) > GM_SUBCAP;
            bool ok = rowvalid && (c_a <= pr.lo) && (c_a + c_ab >= pr.hi + 1) && !ovf;
            if (rowvalid && !ok && cg == 0 && ovf) atomicAdd(&counters[15], 1ull);
            const float2* ml = my_lists + cg * GM_SUBCAP * GM_ROWS + 2 * myrow;
            const int n = ok ? ln : 0;
            const float pa = s_pa[myrow], pb = s_pb[myrow];
#pragma unroll
            for (int i = 0; i < 8; ++i) reinterpret_cast<uint4*>(hall + warp * 1024)[i * 32 + lane] = make_uint4(0u, 0u, 0u, 0u);
            __syncwarp();
            // The sub-list is streamed from L2 twice, in chunks of 16 loads in flight (binning pass, then candidate pass):
            // holding it in "registers" put all 64 values of every thread in local memory (128 KB per CTA, more than the
            // L1 left beside 200 KB of shared memory), which is where the time of this phase went.
            float2* cstage = reinterpret_cast<float2*>(R1 + 65536);       // [8][GM_THREADS] (value, column) of a thread's first candidates
            // value range of the list: the pivots, unless a pivot is the "unbounded" sentinel
            float vmin = pa, vmax = pb;
            {
                float lmin = INFINITY, lmax = -INFINITY;
                const bool need = ok && (pa == plow || pb == GM_PMAX);
                if (need)
                    for (int q = 0; q < n; ++q) { const float x = ml[GM_LIDX(q)].x; lmin = fminf(lmin, x); lmax = fmaxf(lmax, x); }
                xf[xi + cg] = lmin; xf[GM_THREADS + xi + cg] = lmax;
                gm_csync();
                if (need) {
                    const float4 a = *reinterpret_cast<const float4*>(xf + xi), b = *reinterpret_cast<const float4*>(xf + GM_THREADS + xi);
                    vmin = fminf(fminf(a.x, a.y), fminf(a.z, a.w));
                    vmax = fmaxf(fmaxf(b.x, b.y), fmaxf(b.z, b.w));
                    vmax = vmax + fabsf(vmax) * 1.0e-6f + 1.0e-30f;          // strictly above every element
                }
            }
            GM_PROBE(6);
            // binning pass (monotone map value -> bucket)
            const float bw = (vmax - vmin) * (1.0f / 128.0f);
            const float bscale = (bw > 0.0f) ? 1.0f / bw : 0.0f;
            for (int q0 = 0; q0 < n; q0 += 32) {
                // unconditional loads (index clamped to the last element) so that all 32 are in flight at once; the
                // increment carries the predicate
                float t[32];
#pragma unroll
                for (int q = 0; q < 32; q += 2) {
                    const float4 e2 = *reinterpret_cast<const float4*>(ml + GM_LIDX(min(q0 + q, (n - 1) & ~1)));
                    t[q] = e2.x; t[q + 1] = e2.z;
                }
#pragma unroll
                for (int q = 0; q < 32; ++q) {
                    const int b = min(127, max(0, (int)((t[q] - vmin) * bscale)));
                    pbyte[(b >> 2) * 128 + (b & 3)] += (unsigned char)(q0 + q < n);
                }
            }
            gm_csync();
            GM_PROBE(7);
            // scan: owner cg takes buckets 32 cg .. 32 cg + 31; bucket b_lo holds rank kk, b_hi rank kk + (pr.hi - pr.lo)
            const int kk = pr.lo - c_a, kk2 = kk + (pr.hi - pr.lo);
            int hc[32], mysum = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                unsigned l = 0, h = 0;
#pragma unroll
                for (int o = 0; o < 4; ++o) {
                    const unsigned x = hall[((4 * o + lq) * 32 + 8 * cg + k) * 32 + lane];
                    l += x & 0x00ff00ffu; h += (x >> 8) & 0x00ff00ffu;
                }
                hc[4 * k] = (int)(l & 0xffffu); hc[4 * k + 1] = (int)(h & 0xffffu); hc[4 * k + 2] = (int)(l >> 16); hc[4 * k + 3] = (int)(h >> 16);
                mysum += hc[4 * k] + hc[4 * k + 1] + hc[4 * k + 2] + hc[4 * k + 3];
            }
            xc[xi + cg] = mysum;
            gm_csync();
            int b_lo = -1, b_hi = -1;
            {
                const int4 c = *reinterpret_cast<const int4*>(xc + xi);
                const int cc[4] = {c.x, c.y, c.z, c.w};
                int run = 0;
#pragma unroll
                for (int g = 0; g < 4; ++g) if (g < cg) run += cc[g];
#pragma unroll
                for (int b = 0; b < 32; ++b) {
                    const int nxt = run + hc[b];
                    if (run <= kk && kk < nxt) b_lo = 32 * cg + b;
                    if (run <= kk2 && kk2 < nxt) b_hi = 32 * cg + b;
                    run = nxt;
                }
            }
            gm_csync();                 // xc is rewritten
            xc[xi + cg] = b_lo; xc[GM_THREADS + xi + cg] = b_hi;
            gm_csync();
            GM_PROBE(8);
            float c_min = 0.0f, c_max = 0.0f;
            if (ok) {
                const int4 a = *reinterpret_cast<const int4*>(xc + xi), b = *reinterpret_cast<const int4*>(xc + GM_THREADS + xi);
                b_lo = max(max(a.x, a.y), max(a.z, a.w));
                b_hi = max(max(b.x, b.y), max(b.z, b.w));
                // bucket edges with 1 % slack for the rounding of the binning map, then the 2E band
                const float w_lo = vmin + ((float)b_lo - 0.01f) * bw, w_hi = vmin + ((float)b_hi + 1.01f) * bw;
                c_min = w_lo - 2.0f * E; c_max = w_hi + 2.0f * E;
                // the candidate window must lie strictly inside the list's value range [pa, pb)
                const bool inside = b_lo >= 0 && b_hi >= 0 && (c_min > pa || pa == plow) && (c_max < pb || pb == GM_PMAX);
                if (!inside) { ok = false;  }
            }
            // candidates = list elements in [c_min, c_max]; every exact order statistic is among them.  Second pass over the
            // sub-list: count the elements below the window, stage the first eight candidates of this thread in shared memory
            {
                int nb_l = 0, nc_l = 0;
                if (ok)
                    for (int q0 = 0; q0 < n; q0 += 16) {
                        float2 t[16];
#pragma unroll
                        for (int q = 0; q < 16; q += 2) {
                            const float4 e2 = *reinterpret_cast<const float4*>(ml + GM_LIDX(min(q0 + q, (n - 1) & ~1)));
                            t[q] = make_float2(e2.x, e2.y); t[q + 1] = make_float2(e2.z, e2.w);
                        }
#pragma unroll
                        for (int q = 0; q < 16; ++q) {           // branch-free: predicated store into the staging slot
                            const bool live = q0 + q < n;
                            const bool cand = live && t[q].x >= c_min && t[q].x <= c_max;
                            nb_l += (live && t[q].x < c_min);
                            if (cand) cstage[min(nc_l, 7) * GM_THREADS + tid] = t[q];
                            nc_l += cand;
                        }
                    }
                gm_csync();             // xc is rewritten
                xc[xi + cg] = nb_l | (nc_l << 16);
                gm_csync();
                if (ok) {
                    const int4 c = *reinterpret_cast<const int4*>(xc + xi);
                    const int cc[4] = {c.x, c.y, c.z, c.w};
                    int n_c = 0, n_b = 0, slot = 0;
#pragma unroll
                    for (int g = 0; g < 4; ++g) { if (g < cg) slot += cc[g] >> 16; n_c += cc[g] >> 16; n_b += cc[g] & 0xffff; }
                    const int kx = pr.lo - (c_a + n_b);
                    if (kx < 0 || kx + (pr.hi - pr.lo) >= n_c) { ok = false;  }
                    else if (n_c > 32) { if (cg == 0) { s_nc[myrow] = 0; s_flag[myrow] = 2; } }      // non-uniform list: bisection below
                    else {
                        if (nc_l <= 8) {
                            for (int k = 0; k < nc_l; ++k) {
                                const float2 ent = cstage[k * GM_THREADS + tid];
                                pool_col[GM_POOL(myrow, slot)] = __float_as_int(ent.y); pool_dex[GM_POOL(myrow, slot)] = ent.x; ++slot;
                            }
                        } else {
                            // more candidates than staging slots (plateaus of near-equal distances put them in one owner's
                            // sub-list): read the sub-list again, 16 loads in flight
                            for (int q0 = 0; q0 < n; q0 += 16) {
                                float2 t[16];
#pragma unroll
                                for (int q = 0; q < 16; ++q) t[q] = ml[GM_LIDX(min(q0 + q, n - 1))];
#pragma unroll
                                for (int q = 0; q < 16; ++q)
                                    if (q0 + q < n && t[q].x >= c_min && t[q].x <= c_max) {
                                        pool_col[GM_POOL(myrow, slot)] = __float_as_int(t[q].y); pool_dex[GM_POOL(myrow, slot)] = t[q].x; ++slot;
                                    }
                            }
                        }
                        if (cg == 0) { s_nc[myrow] = n_c; s_kx[myrow] = kx; }
                    }
                }
                if (rowvalid && !ok && cg == 0) { s_nc[myrow] = 0; s_flag[myrow] = 1; atomicAdd(&counters[1], 1ull); }
            }
            gm_csync();
            GM_PROBE(9);
            // (A') rows whose bucket window held too many elements (the list is far from uniform): warp per row, value
            //      bisection on the whole list until the window holds at most three elements
            for (int r = warp; r < tk.nrows; r += GM_WARPS) {
                if (s_flag[r] != 2) continue;
                const int4 c4r = *reinterpret_cast<const int4*>(s_lcnt + 4 * r);
                const int cseg[4] = {c4r.x, c4r.y, c4r.z, c4r.w};
                const int c_ar = s_cnta[r], c_abr = c4r.x + c4r.y + c4r.z + c4r.w;
                const float par = s_pa[r], pbr = s_pb[r];
                const float2* lst = my_lists + 2 * r;
                float l2[GM_LPL];
                float vmn = INFINITY, vmx = -INFINITY;
#pragma unroll
                for (int q = 0; q < GM_LPL; ++q) {              // slot (q % 4) * 32 + lane of segment q / 4
                    const int sl = (q & 3) * 32 + lane;
                    l2[q] = INFINITY;
                    if (sl < cseg[q >> 2]) { l2[q] = lst[(q >> 2) * GM_SUBCAP * GM_ROWS + GM_LIDX(sl)].x; vmn = fminf(vmn, l2[q]); vmx = fmaxf(vmx, l2[q]); }
                }
                for (int o = 16; o > 0; o >>= 1) {
                    vmn = fminf(vmn, __shfl_xor_sync(0xffffffffu, vmn, o));
                    vmx = fmaxf(vmx, __shfl_xor_sync(0xffffffffu, vmx, o));
                }
                const int kkr = pr.lo - c_ar;                   // invariant: #(list < w_lo) <= kkr < #(list < w_hi)
                float w_lo = vmn, w_hi = INFINITY, b_hi = vmx;
                int c_lo = 0, c_hi = c_abr;
                for (int it = 0; it < 32 && c_hi - c_lo > 3; ++it) {
                    const float mid = 0.5f * (w_lo + b_hi);
                    if (!(mid > w_lo) || !(mid < b_hi)) break;
                    int cm = 0;
#pragma unroll
                    for (int q = 0; q < GM_LPL; ++q) cm += (l2[q] < mid);
                    cm = __reduce_add_sync(0xffffffffu, cm);
                    if (cm <= kkr) { w_lo = mid; c_lo = cm; }
                    else { w_hi = mid; b_hi = mid; c_hi = cm; }
                }
                if (pr.hi != pr.lo && c_hi < kkr + 2) {         // rank kkr + 1 is the smallest element >= w_hi
                    float nx = INFINITY;
#pragma unroll
                    for (int q = 0; q < GM_LPL; ++q) if (l2[q] >= w_hi) nx = fminf(nx, l2[q]);
                    for (int o = 16; o > 0; o >>= 1) nx = fminf(nx, __shfl_xor_sync(0xffffffffu, nx, o));
                    w_hi = nx;
                }
                const float hi_edge = (w_hi == INFINITY) ? vmx : w_hi;
                bool okr = (w_lo - 2.0f * E > par || par == plow) && (hi_edge + 2.0f * E < pbr || pbr == GM_PMAX);
                int n_c = 0, kx = 0;
                if (okr) {
                    const float c_min = w_lo - 2.0f * E;
                    const float c_max = (w_hi == INFINITY) ? INFINITY : w_hi + 2.0f * E;
                    int n_below = c_ar;
#pragma unroll
                    for (int q = 0; q < GM_LPL; ++q) {
                        const bool valid = l2[q] != INFINITY;
                        const bool below = valid && l2[q] < c_min;
                        const bool cand = valid && !below && l2[q] <= c_max;
                        n_below += __popc(__ballot_sync(0xffffffffu, below));
                        const unsigned mc = __ballot_sync(0xffffffffu, cand);
                        if (cand) {
                            const int slot = n_c + __popc(mc & ((1u << lane) - 1u));
                            if (slot < 32) {
                                pool_col[GM_POOL(r, slot)] = __float_as_int(lst[(q >> 2) * GM_SUBCAP * GM_ROWS + GM_LIDX((q & 3) * 32 + lane)].y);
                                pool_dex[GM_POOL(r, slot)] = l2[q];
                            }
                        }
                        n_c += __popc(mc);
                    }
                    kx = pr.lo - n_below;
                    if (n_c > 32 || kx < 0 || kx + (pr.hi - pr.lo) >= n_c) okr = false;
                }
                if (lane == 0) {
                    if (okr) { s_nc[r] = n_c; s_kx[r] = kx; s_flag[r] = 0; }
                    else { s_nc[r] = 0; s_flag[r] = 1; atomicAdd(&counters[1], 1ull); }
                }
            }
            gm_csync();
            GM_PROBE(10);
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
            gm_csync();
            GM_PROBE(11);
            {
                const int boff = side_a_is_query ? (ACOSS_NBINS - rot) % ACOSS_NBINS : 0;     // rot_src(c, rot) = (c + boff) % 12
                const int n_tasks_c = s_off[GM_ROWS];
                for (int idx = tid; idx < n_tasks_c; idx += GM_THREADS) {
                    int r = 0;                                            // last row with s_off[r] <= idx
#pragma unroll
                    for (int st = 64; st > 0; st >>= 1) if (s_off[r + st] <= idx) r += st;
                    const int slot = idx - s_off[r];
                    const int jj = pool_col[GM_POOL(r, slot)];
                    const float4* bp = reinterpret_cast<const float4*>(fr_b + (int64_t)jj * ACOSS_NBINS);
                    const float4* as = reinterpret_cast<const float4*>(Araw + r * ACOSS_NBINS);
                    float acc = 0.0f;
                    // oracle bin c of the side-B frame is its bin cbs[c] = (c + boff) % 12: whole frames are loaded as three
                    // 16-byte vectors and the rotation is resolved at compile time inside a uniform switch
#pragma unroll 3
                    for (int s_ = 0; s_ < 9; ++s_) {
                        const float4 a0 = as[3 * s_], a1 = as[3 * s_ + 1], a2 = as[3 * s_ + 2];
                        const float4 b0 = __ldg(bp + 3 * s_), b1 = __ldg(bp + 3 * s_ + 1), b2 = __ldg(bp + 3 * s_ + 2);
                        const float av[12] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w, a2.x, a2.y, a2.z, a2.w};
                        const float bv[12] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w, b2.x, b2.y, b2.z, b2.w};
                        acc = gm_chain12_rot(acc, av, bv, boff);
                    }
                    const float sa = s_xx[r], sb = __ldg(sdB + jj);
                    const float xx = side_a_is_query ? sa : sb, yy = side_a_is_query ? sb : sa;
                    float item = xadd(xsub(xx, xmul(2.0f, acc)), yy);
                    max_err = fmaxf(max_err, fabsf(item - pool_dex[GM_POOL(r, slot)]) / (sa + sb));
                    if (item < 0.0f) item = 0.0f;
                    pool_dex[GM_POOL(r, slot)] = __fsqrt_rn(item);
                    ++n_exact;
                }
            }
            gm_csync();
            GM_PROBE(12);
            // (C) four threads per row: ranks kx and kx + 1 among the row's exact candidate distances (ties by slot); the row's
            //     distances sit in registers, thread cg ranks candidates cg, cg + 4, ...
            {
                float* xlo = reinterpret_cast<float*>(R1 + GM_R1_BYTES);          // [GM_ROWS] (the exchange buffers are free here)
                float* xhi = xlo + GM_THREADS;
                const int n_c = rowvalid ? s_nc[myrow] : 0, kx = s_kx[myrow];
                if (n_c > 0) {
                    float d[32];
#pragma unroll
                    for (int b_ = 0; b_ < 32; ++b_) d[b_] = (b_ < n_c) ? pool_dex[GM_POOL(myrow, b_)] : INFINITY;
                    for (int a_ = cg; a_ < n_c; a_ += 4) {
                        const float da = pool_dex[GM_POOL(myrow, a_)];
                        int rank = 0;
#pragma unroll
                        for (int b_ = 0; b_ < 32; ++b_) rank += (d[b_] < da) || (d[b_] == da && b_ < a_);
                        if (rank == kx) { xlo[myrow] = da; if (kx + 1 >= n_c) xhi[myrow] = da; }
                        if (rank == kx + 1) xhi[myrow] = da;
                    }
                }
                gm_csync();
                if (cg == 0 && n_c > 0) {
                    const float t = percentile_combine(pr, xlo[myrow], xhi[myrow]);
                    s_th[myrow] = t; thr_out[tk.row0 + myrow] = t;
                }
            }
        } else if (tid < GM_ROWS) {
            s_flag[tid] = 1;
        }
        gm_csync();
        GM_PROBE(13);
        // ---- rows that could not be bracketed (and tiny pairs): the whole row exactly, one row at a time
        const int any_flag = gm_csync_or(tid < tk.nrows && s_flag[tid]);
        for (int r = 0; any_flag && r < tk.nrows; ++r) {
            if (!s_flag[r]) continue;
            float* vr = reinterpret_cast<float*>(R1);
            for (int j = tid; j < NB; j += GM_THREADS) {
                vr[j] = side_a_is_query ? exact_dist(frames, pd, rot, selfdot, tk.row0 + r, j)
                                        : exact_dist(frames, pd, rot, selfdot, j, tk.row0 + r);
                ++n_exact;
            }
            gm_csync();
            if (warp == 0) {
                float v_hi;
                const float v_lo = warp_select_pair(vr, NB, pr.lo, lane, v_hi);
                const float t = percentile_combine(pr, v_lo, v_hi);
                if (lane == 0) { s_th[r] = t; thr_out[tk.row0 + r] = t; }
            }
            gm_csync();
        }
        GM_PROBE(14);
        if (tid == 0) {
            ph4 = clock64();
            atomicAdd(&counters[2], (unsigned long long)(ph1 - ph0));
            atomicAdd(&counters[3], (unsigned long long)(ph2 - ph1));
            atomicAdd(&counters[4], (unsigned long long)(ph3 - ph2));
            atomicAdd(&counters[5], (unsigned long long)(ph4 - ph3));
            atomicAdd(&counters[7], 1ull);
        }
        if (!side_a_is_query) continue;

        // ---- pass 2: binarise.  Per cell: sure-in <=> rho~ < min(lo_q, lo_r[j]), sure-out <=> rho~ > min(hi_q, hi_r[j]);
        //      anything else is decided exactly afterwards.  A thread owns one CRP row: 32 bytes per tcgen05.ld go
        //      straight from registers to HBM.
        int* amb = reinterpret_cast<int*>(R1 + GM_AMB_OFF);
        int* amb_cnt = reinterpret_cast<int*>(s_red);
        if (tid == 0) *amb_cnt = 0;
        if (tid < GM_ROWS) {
            const float ta = (tid < tk.nrows) ? s_th[tid] : 0.0f;
            const float Tq = xmul(ta, ta);
            const float mq = E + Tq * 4.76837158203125e-07f;
            s_loq[tid] = Tq - mq;
        }
        gm_csync();
        {
            const float lo_q = s_loq[myrow];
            const uint32_t zero_word = (CRP_DELTA_ZERO) | (CRP_DELTA_ZERO << 8) | (CRP_DELTA_ZERO << 16) | (CRP_DELTA_ZERO << 24);
            const uint32_t flip = CRP_DELTA_ZERO ^ CRP_DELTA_ONE;
            uint8_t* orow = crp + pd.crp_off + (int64_t)(tk.row0 + myrow) * pd.pitch;
            const int pitch = pd.pitch;
            // d = rho~ - min(lo_q, lo_r[j]): the sign bit of d is "surely in"; 0 <= d <= W (one unsigned compare on the bit
            // pattern) is "ambiguous", where W bounds the width of every band (hi - lo <= 2 (E + 8u T) < 2.2 E).  The four sign
            // bits of a group become the four CRP bytes with three PRMTs (sign replication) and one LOP3.
            const uint32_t wamb = __float_as_uint(2.2f * E);
            stream([&](int jbase, int slot, int cbase, const float (&v)[16]) {
                const float4* lc = reinterpret_cast<const float4*>(loC + slot * GM_TN + cbase);
                uint32_t w[4], amask = 0;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const float4 l4 = lc[g];
                    const uint32_t u0 = __float_as_uint(v[4 * g] - fminf(lo_q, l4.x)), u1 = __float_as_uint(v[4 * g + 1] - fminf(lo_q, l4.y));
                    const uint32_t u2 = __float_as_uint(v[4 * g + 2] - fminf(lo_q, l4.z)), u3 = __float_as_uint(v[4 * g + 3] - fminf(lo_q, l4.w));
                    uint32_t s01, s23, sg;
                    asm("prmt.b32 %0, %1, %2, 0x00fb;" : "=r"(s01) : "r"(u0), "r"(u1));      // sign-replicated top bytes of u0, u1
                    asm("prmt.b32 %0, %1, %2, 0x00fb;" : "=r"(s23) : "r"(u2), "r"(u3));
                    asm("prmt.b32 %0, %1, %2, 0x5410;" : "=r"(sg) : "r"(s01), "r"(s23));
                    w[g] = zero_word ^ (sg & (flip * 0x01010101u));
                    if (min(min(u0, u1), min(u2, u3)) <= wamb)
                        amask |= ((uint32_t)(u0 <= wamb) | ((uint32_t)(u1 <= wamb) << 1) | ((uint32_t)(u2 <= wamb) << 2) | ((uint32_t)(u3 <= wamb) << 3)) << (4 * g);
                }
                if (rowvalid && jbase + 16 <= pitch) *reinterpret_cast<uint4*>(orow + jbase) = make_uint4(w[0], w[1], w[2], w[3]);
                if (amask && rowvalid) {
                    int at = atomicAdd(amb_cnt, __popc(amask));
                    while (amask) {
                        const int b = __ffs(amask) - 1;
                        amask &= amask - 1;
                        if (at < GM_AMB_CAP) amb[at] = (myrow << 24) | (jbase + b);
                        ++at;
                    }
                }
            }, true);
            // columns past the last tile up to the pitch
            const int done = n_tiles * GM_TN;
            if (done < pitch) {
                const int wrem = (pitch - done) >> 2;
                for (int idx = tid; idx < tk.nrows * wrem; idx += GM_THREADS) {
                    const int r = idx / wrem, w = idx - r * wrem;
                    reinterpret_cast<uint32_t*>(crp + pd.crp_off + (int64_t)(tk.row0 + r) * pitch + done)[w] = zero_word;
                }
            }
        }
        gm_csync();
        GM_PROBE(15);
        {
            const int na = *amb_cnt;
            if (na > GM_AMB_CAP) {
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
                // pass 2: side A = query (rows staged in Araw in the oracle's order), side B = reference, rotated by boff
                const int boff = (ACOSS_NBINS - rot) % ACOSS_NBINS;
                for (int e = tid; e < na; e += GM_THREADS) {
                    const int code = amb[e];
                    const int r = code >> 24, j = code & 0xFFFFFF;
                    const float4* bp = reinterpret_cast<const float4*>(fr_b + (int64_t)j * ACOSS_NBINS);
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
                    float item = xadd(xsub(s_xx[r], xmul(2.0f, acc)), __ldg(sdB + j));
                    if (item < 0.0f) item = 0.0f;
                    const float dv = __fsqrt_rn(item);
                    ++n_exact;
                    if (dv <= s_th[r] && dv <= thB[j])
                        crp[pd.crp_off + (int64_t)(tk.row0 + r) * pd.pitch + j] = (uint8_t)CRP_DELTA_ONE;
                }
            }
        }
        gm_csync();
        GM_PROBE(16);
        if (tid == 0) atomicAdd(&counters[6], (unsigned long long)(clock64() - ph4));
    }
    for (int o = 16; o > 0; o >>= 1) {
        n_exact += __shfl_xor_sync(0xffffffffu, n_exact, o);
        max_err = fmaxf(max_err, __shfl_xor_sync(0xffffffffu, max_err, o));
    }
    if (lane == 0 && n_exact) atomicAdd(&counters[0], n_exact);
    if (lane == 0 && max_err > 0.0f) atomicMax(&counters[14], (unsigned long long)__float_as_uint(max_err));
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) gm_tmem_free(tmem_base);
}
