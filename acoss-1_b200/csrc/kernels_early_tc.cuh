// kernels_early_tc.cuh -- the three EarlyFusionTralie cross-similarity matrices of a pair on the tensor cores (sm_100a).
//
// Replaces get_csm (acoss/algorithms/utils/cross_recurrence.py:30-48: Gram-expansion Euclid, d = 50 * n_mfcc and 1225) and
// get_csm_blocked_oti(get_csm_cosine) (:53-73, :105-134: 1 - cosine after rolling the query's chroma bins by the OTI,
// d = 480), called at earlyfusion_traile.py:167-174.  The contraction <x_i, y_j> is a tcgen05 GEMM in TF32 with the split
// x = x_hi + x_lo (three products hh + hl + lh): the dropped lo.lo term is 2^-22 relative, i.e. the result is as close to the
// exact product as the reference's float32 BLAS (whose summation order nothing can reproduce).
//
// Operand store (ef_planes_kernel, once per upload).  Per track and feature the rows are kept CHUNK-MAJOR as TF32 hi and lo
// planes, [variant][hi | lo][chunk kc][row][16 B] (chunk = 4 consecutive floats along d, d zero-padded to a multiple of 8):
// for any tile the 16-byte chunks kc of 128 / 256 consecutive rows are one contiguous 2 / 4 KB run, which is exactly one
// no-swizzle K-major UMMA core-matrix column (rows 16 B apart, SBO = 128 B) -- a stage of the pipeline is filled by 16
// cp.async.bulk copies (TMA engine) straight into its final layout, no LSU work, no conversion in the kernel.
//   * chroma rows are stored normalised (x / |x|, zero norms -> 1 like the reference) in FOUR variants r = 0..3 with the 12 bins
//     of every chroma frame rolled by r; a roll by oti = 4 q + r is then variant r with the three 4-bin chunks of every frame
//     permuted by q -- a pure re-addressing of the bulk copies of the query side.
//
// Kernel (ef_gemm_kernel): persistent CTA, tile = 128 query rows x 256 reference rows x one feature; warp 8 = TMA producer
// (4-stage ring, 4 chunks = 16 floats of K per stage, 48 KB), warp 9 = MMA issuer (6 MMAs of 128 x 256 x 8 per stage), warps
// 0..7 = epilogue (thread = TMEM lane = query row x column half).  MEASURED: the tensor core adds into its fp32 accumulator
// with truncation towards zero -- a 1000-term contraction accumulated in one TMEM accumulator came out 1e-5 low in magnitude
// (round 2, first version of this kernel).  The contraction is therefore cut into SPANS of 64 floats of K, each accumulated in
// a fresh TMEM accumulator (two alternate), and the spans are summed by the epilogue warps in float32 round-to-nearest
// registers; then xx + yy - 2 acc, clip, sqrt (or 1 - acc) and 16-byte row segments to HBM.
#pragma once
#include "common.cuh"
#include "kernels_crp_tc.cuh"
#include "kernels_crp_gemm.cuh"
#include "kernels_early.cuh"

#define EG_TM 128
#define EG_TN 256
#define EG_STAGE_CH 4                                   // 16-byte K chunks per stage
#define EG_STAGES 4
#define EG_A_CH (EG_TM * 16)                            // bytes of one chunk column of the A tile
#define EG_B_CH (EG_TN * 16)
#define EG_STAGE_BYTES (EG_STAGE_CH * 2 * (EG_A_CH + EG_B_CH))     // 49 152: [A hi][A lo][B hi][B lo] x 4 chunks
#define EG_EPI_WARPS 8
#define EG_SPAN 4                                       // stages (= 8 K-steps of 8 floats) accumulated in one TMEM accumulator
#define EG_THREADS (32 * (EG_EPI_WARPS + 2))
#define EG_SMEM (EG_STAGES * EG_STAGE_BYTES + EG_TN * 4 + 256)

__host__ __device__ __forceinline__ int eg_chunks(int d) { return ((d + 7) / 8) * 2; }     // 16-byte chunks, padded to whole K = 8 steps

struct EgTile { int32_t pair, feat, m0, n0; };      // m0 < 0: no tile (the odd partner of a cluster pair, see ef_gemm_kernel<2>)

// ---- cluster helpers (ef_gemm_kernel<2>: two CTAs share every reference-side operand stage through TMA multicast) ----
__device__ __forceinline__ uint32_t eg_cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void eg_cluster_sync()
{
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the mbarrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void eg_mbar_arrive_remote(uint64_t* bar, uint32_t rank)
{
    asm volatile(
        "{\n\t"
        ".reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t"
        "}\n" ::"r"(smem_u32(bar)), "r"(rank) : "memory");
}
// wait with cluster-scope acquire (the data guarded by the barrier was written from the partner CTA)
__device__ __forceinline__ void eg_mbar_wait_cluster(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "EGW_LOOP:\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra EGW_DONE;\n\t"
        "bra EGW_LOOP;\n\t"
        "EGW_DONE:\n\t"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// store one word at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void eg_st_remote(int* p, uint32_t rank, int v)
{
    asm volatile(
        "{\n\t"
        ".reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "st.shared::cluster.b32 [ra], %2;\n\t"
        "}\n" ::"r"(smem_u32(p)), "r"(rank), "r"(v) : "memory");
}
// MMA completion -> one arrival on the mbarrier at this offset in every CTA of `mask`
__device__ __forceinline__ void eg_umma_commit_mc(uint64_t* bar, uint16_t mask)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}

// planes of one feature of all tracks: thread per (row, chunk); variant v rolls the 12 bins of every chroma frame by v and is
// only built for the cosine feature (n_var = 4), whose rows are also normalised
__global__ void ef_planes_kernel(const float* __restrict__ X, const float* __restrict__ sq, const int64_t* __restrict__ blk_offs,
                                 const int64_t* __restrict__ plane_off, int n_tracks, int d, int nkc, int n_var, int cosine,
                                 unsigned char* __restrict__ planes)
{
    const int track = blockIdx.y;
    const int64_t r0 = blk_offs[track];
    const int rows = (int)(blk_offs[track + 1] - r0);
    const int64_t total = (int64_t)rows * nkc;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int kc = (int)(idx / rows), row = (int)(idx - (int64_t)kc * rows);      // rows fastest: coalesced stores
        const float* xr = X + (r0 + row) * d;
        float nrm = 1.0f;
        if (cosine) { nrm = sqrtf(sq[r0 + row]); if (nrm == 0.0f) nrm = 1.0f; }
        for (int v = 0; v < n_var; ++v) {
            float x[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int k = 4 * kc + e;
                float val = 0.0f;
                if (k < d) {
                    int ks = k;
                    if (v) { const int g = k / ACOSS_NBINS, c = k - g * ACOSS_NBINS; ks = g * ACOSS_NBINS + rot_src(c, v); }
                    val = xr[ks];
                    if (cosine) val = val / nrm;
                }
                x[e] = val;
            }
            float4 hi, lo;
            tf32_split(x[0], hi.x, lo.x); tf32_split(x[1], hi.y, lo.y); tf32_split(x[2], hi.z, lo.z); tf32_split(x[3], hi.w, lo.w);
            unsigned char* base = planes + plane_off[track] + (size_t)v * 2 * nkc * rows * 16;
            *reinterpret_cast<float4*>(base + ((size_t)kc * rows + row) * 16) = hi;
            *reinterpret_cast<float4*>(base + ((size_t)(nkc + kc) * rows + row) * 16) = lo;
        }
    }
}

struct EgFeat {
    const unsigned char* planes;      // plane store of this feature
    const int64_t* plane_off;         // per track byte offset
    const float* sq;                  // squared row norms (all tracks)
    int nkc;                          // chunks per row (even)
};

// CL = 1: one CTA per tile.  CL = 2 (launched as clusters of two CTAs): tiles[2 c] and tiles[2 c + 1] are two 128-row query
// tiles against the SAME 256 reference rows; each CTA fetches its own query operand and HALF of the reference operand, the
// latter multicast into both CTAs' shared memory, so a stage costs 32 KB instead of 48 KB of L2 reads per CTA (the GEMM is bound
// by L2 -> SM operand delivery, not by the tensor pipe).  A stage buffer is refilled once BOTH CTAs' MMAs have consumed it.
template <int CL>
__global__ void __launch_bounds__(EG_THREADS, 1)
ef_gemm_kernel(const EfPair* __restrict__ pairs, const EgTile* __restrict__ tiles, int n_tiles,
               EgFeat f0, EgFeat f1, EgFeat f2, float* __restrict__ csm, int* __restrict__ tile_counter)
{
    const uint32_t crank = CL == 2 ? eg_cluster_rank() : 0u;
    const int t_first = CL == 2 ? 2 * (int)(blockIdx.x >> 1) + (int)crank : (int)blockIdx.x;
    const int t_step = CL == 2 ? 2 * (int)(gridDim.x >> 1) : (int)gridDim.x;
    extern __shared__ __align__(128) unsigned char esm[];
    unsigned char* stage0 = esm;
    float* s_yy = reinterpret_cast<float*>(esm + EG_STAGES * EG_STAGE_BYTES);          // [EG_TN] norms of the tile's reference rows
    uint64_t* mbar = reinterpret_cast<uint64_t*>(s_yy + EG_TN);
    uint64_t* mb_full = mbar, *mb_empty = mbar + EG_STAGES, *mb_tfull = mbar + 2 * EG_STAGES, *mb_tempty = mbar + 2 * EG_STAGES + 2;
    // Tiles are handed out in list order from a global counter (CL = 1): the tiles in flight are then a contiguous window of the
    // list, so the eight tiles of one (pair, feature) start within microseconds of each other and find each other's operand
    // rows in L2.  With the static round-robin the CTAs drifted apart (tile times differ 2.5 x between the features) and 71 % of
    // all operand reads came from DRAM (ncu: 236 GB per 4995 pairs).  The producer warp draws the index and passes it to the
    // MMA and epilogue warps through a four-slot ring.
    uint64_t* mb_tq_full = mbar + 2 * EG_STAGES + 4, *mb_tq_empty = mbar + 2 * EG_STAGES + 8;
    uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(mbar + 2 * EG_STAGES + 12);
    int* s_tile = reinterpret_cast<int*>(tmem_holder + 2);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // CL = 2: the counter hands out tile PAIRS; the leader CTA (rank 0) draws, writes the index into both CTAs' rings and signals
    // both; every reader of either CTA releases the slot on the LEADER's empty barrier.
    const bool dynamic = tile_counter != nullptr;
    auto next_tile = [&](uint32_t& it) -> int {              // MMA and epilogue warps (and the producer of the non-leader CTA)
        int t;
        if (!dynamic) { t = t_first + (int)it * t_step; if (t >= n_tiles) t = -1; }
        else {
            const uint32_t q = it & 3u;
            if (CL == 2) eg_mbar_wait_cluster(mb_tq_full + q, (it >> 2) & 1u); else mbar_wait(mb_tq_full + q, (it >> 2) & 1u);
            t = s_tile[q];
            __syncwarp();
            if (lane == 0) { if (CL == 2 && crank != 0) eg_mbar_arrive_remote(mb_tq_empty + q, 0u); else mbar_arrive(mb_tq_empty + q); }
            if (CL == 2 && t >= 0) t += (int)crank;
        }
        ++it;
        return t;
    };

    if (warp == 0) gm_tmem_alloc(tmem_holder);
    if (tid == 0) {
        for (int s = 0; s < EG_STAGES; ++s) { mbar_init(mb_full + s, 1); mbar_init(mb_empty + s, CL); }
        mbar_init(mb_tfull, 1); mbar_init(mb_tfull + 1, 1); mbar_init(mb_tempty, EG_EPI_WARPS); mbar_init(mb_tempty + 1, EG_EPI_WARPS);
        for (int q = 0; q < 4; ++q) { mbar_init(mb_tq_full + q, 1); mbar_init(mb_tq_empty + q, CL == 2 ? 2 * (EG_EPI_WARPS + 1) + 1 : EG_EPI_WARPS + 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (CL == 2) eg_cluster_sync();              // the partner's barriers exist before anything is sent to them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_holder;
    const uint32_t st_saddr = smem_u32(stage0);

    if (warp == EG_EPI_WARPS) {
        // ================= TMA producer =================
        // The 16 bulk copies of a stage (4 K chunks x hi / lo x query / reference side) are issued by 16 LANES at once: one
        // elected lane issuing them in turn took longer per stage than the stage's six MMAs.
        {
            uint32_t g = 0;                       // running stage count
            const int c = lane >> 2, h = (lane >> 1) & 1, side_b = lane & 1;      // this lane's copy of a stage (lanes 0..15)
            // the operand planes are re-read by the other tiles of the pair (and the query side by the next pairs): keep them in L2
            // ahead of the CSM stores streaming out (ncu: 236 GB of DRAM reads per 4995 pairs, 4x what one pass over the operands needs)
            uint64_t pol;
            asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
            for (uint32_t it = 0;;) {
                int t;
                if (!dynamic) { t = t_first + (int)it * t_step; if (t >= n_tiles) t = -1; }
                else if (CL == 2 && crank != 0) {
                    t = next_tile(it);          // advances it
                    --it;
                } else {
                    t = 0;
                    if (lane == 0) {
                        const uint32_t q = it & 3u;
                        if (it >= 4) mbar_wait(mb_tq_empty + q, ((it >> 2) - 1) & 1u);
                        t = atomicAdd(tile_counter, 1) * CL;
                        if (t >= n_tiles) t = -1;
                        s_tile[q] = t;
                        if (CL == 2) { eg_st_remote(s_tile + q, 1u, t); eg_mbar_arrive_remote(mb_tq_full + q, 1u); }
                        mbar_arrive(mb_tq_full + q);
                    }
                    t = __shfl_sync(0xffffffffu, t, 0);
                }
                ++it;
                if (t < 0) break;
                const EgTile tl = tiles[t];
                const EfPair pd = pairs[tl.pair];
                const EgFeat F = tl.feat == 0 ? f0 : (tl.feat == 1 ? f1 : f2);
                const int nkc = F.nkc;
                const int ra = tl.m0 >= 0 ? min(EG_TM, pd.M - tl.m0) : 0, rb = min(EG_TN, pd.N - tl.n0);      // valid rows: only those are copied
                // reference rows this CTA fetches (CL = 2: its half, multicast to both CTAs)
                const int b_lo = CL == 2 ? (int)crank * (EG_TN / 2) : 0;
                const int b_cnt = CL == 2 ? max(0, min(EG_TN / 2, rb - b_lo)) : rb;
                const int oti = tl.feat == 2 ? pd.oti % ACOSS_NBINS : 0;
                const int var = oti & 3, qrot = oti >> 2;
                const unsigned char* pa = F.planes + F.plane_off[pd.ta] + (size_t)var * 2 * nkc * pd.M * 16;
                const unsigned char* pb = F.planes + F.plane_off[pd.tb];
                for (int k0 = 0; k0 < nkc; k0 += EG_STAGE_CH, ++g) {
                    const uint32_t s = g % EG_STAGES;
                    if (g >= EG_STAGES) mbar_wait(mb_empty + s, ((g / EG_STAGES) - 1) & 1u);
                    const int nch = min(EG_STAGE_CH, nkc - k0);
                    unsigned char* dst = stage0 + s * EG_STAGE_BYTES;
                    const uint32_t bar = smem_u32(mb_full + s);
                    if (lane == 0)
                        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)(nch * 2 * 16 * (ra + rb))) : "memory");
                    __syncwarp();
                    if (lane < 16 && c < nch) {
                        const int kc = k0 + c;
                        if (!side_b) {
                            // query side: a roll by 4 q permutes the three chunks of a 12-bin frame (chunk c3 of the rolled row is
                            // chunk (c3 - q) mod 3 of the stored variant)
                            int kca = kc;
                            if (qrot) { const int gfr = kc / 3, c3 = kc - 3 * gfr; kca = 3 * gfr + (c3 + 3 - qrot) % 3; }
                            if (ra > 0)
                                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                                             ::"r"(smem_u32(dst + (h * EG_STAGE_CH + c) * EG_A_CH)),
                                               "l"(pa + ((size_t)(h * nkc + kca) * pd.M + tl.m0) * 16), "r"((uint32_t)(ra * 16)), "r"(bar), "l"(pol) : "memory");
                        } else if (CL == 2) {
                            if (b_cnt > 0)
                                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
                                             ::"r"(smem_u32(dst + 2 * EG_STAGE_CH * EG_A_CH + (h * EG_STAGE_CH + c) * EG_B_CH + b_lo * 16)),
                                               "l"(pb + ((size_t)(h * nkc + kc) * pd.N + tl.n0 + b_lo) * 16), "r"((uint32_t)(b_cnt * 16)), "r"(bar),
                                               "h"((uint16_t)3) : "memory");
                        } else {
                            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                                         ::"r"(smem_u32(dst + 2 * EG_STAGE_CH * EG_A_CH + (h * EG_STAGE_CH + c) * EG_B_CH)),
                                           "l"(pb + ((size_t)(h * nkc + kc) * pd.N + tl.n0) * 16), "r"((uint32_t)(rb * 16)), "r"(bar), "l"(pol) : "memory");
                        }
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp == EG_EPI_WARPS + 1) {
        // ================= MMA issuer =================
        uint32_t g = 0, u = 0;                   // running stage / span counts
        for (uint32_t it = 0;;) {
            const int t = next_tile(it);
            if (t < 0) break;
            const EgTile tl = tiles[t];
            const EgFeat F = tl.feat == 0 ? f0 : (tl.feat == 1 ? f1 : f2);
            const int nkc = F.nkc;
            int sidx = 0;
            if (CL == 2 && tl.m0 < 0) {
                // no tile on this side of the pair: the stages still arrive here (the partner's half is multicast into this CTA
                // as well); hand every one straight back to both producers
                for (int k0 = 0; k0 < nkc; k0 += EG_STAGE_CH, ++g) {
                    const uint32_t s = g % EG_STAGES;
                    mbar_wait(mb_full + s, (g / EG_STAGES) & 1u);
                    if (lane == 0) { mbar_arrive(mb_empty + s); eg_mbar_arrive_remote(mb_empty + s, crank ^ 1u); }
                    __syncwarp();
                }
                continue;
            }
            for (int k0 = 0; k0 < nkc; k0 += EG_STAGE_CH, ++g, ++sidx) {
                const uint32_t s = g % EG_STAGES, buf = u & 1;
                const bool first = (sidx % EG_SPAN) == 0, last = (sidx % EG_SPAN) == EG_SPAN - 1 || k0 + EG_STAGE_CH >= nkc;
                if (first && u >= 2) mbar_wait(mb_tempty + buf, ((u >> 1) - 1) & 1u);
                mbar_wait(mb_full + s, (g / EG_STAGES) & 1u);
                tc_fence_after();
                if (gm_elect_one()) {
                    const int nch = min(EG_STAGE_CH, nkc - k0);
                    const uint32_t sa = st_saddr + s * EG_STAGE_BYTES;
                    const uint32_t a_hi = (sa & 0x3FFFFu) >> 4, a_lo = ((sa + EG_STAGE_CH * EG_A_CH) & 0x3FFFFu) >> 4;
                    const uint32_t b_hi = ((sa + 2 * EG_STAGE_CH * EG_A_CH) & 0x3FFFFu) >> 4;
                    const uint32_t b_lo = ((sa + 2 * EG_STAGE_CH * EG_A_CH + EG_STAGE_CH * EG_B_CH) & 0x3FFFFu) >> 4;
                    uint32_t acc = first ? 0u : 1u;
                    for (int c = 0; c < nch; c += 2) {         // one K = 8 step = chunks c, c + 1 (LBO = one chunk column)
                        // the small cross products first: they are added while the accumulator is still small
                        gm_mma(tmem_base + buf * EG_TN, GM_DLO(a_hi, c * EG_A_CH, EG_A_CH), GM_DLO(b_lo, c * EG_B_CH, EG_B_CH), acc);
                        gm_mma(tmem_base + buf * EG_TN, GM_DLO(a_lo, c * EG_A_CH, EG_A_CH), GM_DLO(b_hi, c * EG_B_CH, EG_B_CH), 1);
                        gm_mma(tmem_base + buf * EG_TN, GM_DLO(a_hi, c * EG_A_CH, EG_A_CH), GM_DLO(b_hi, c * EG_B_CH, EG_B_CH), 1);
                        acc = 1;
                    }
                    if (CL == 2) eg_umma_commit_mc(mb_empty + s, (uint16_t)3); else umma_commit(mb_empty + s);
                    if (last) umma_commit(mb_tfull + buf);
                }
                __syncwarp();
                if (last) ++u;
            }
        }
    } else {
        // ================= epilogue: warps 0..7, thread = (TMEM lane = query row, column half) =================
        uint32_t u = 0;
        const int lq = warp & 3, half = warp >> 2;
        const int row = 32 * lq + lane;
        for (uint32_t it = 0;;) {
            const int t = next_tile(it);
            if (t < 0) break;
            const EgTile tl = tiles[t];
            if (CL == 2 && tl.m0 < 0) continue;
            const EfPair pd = pairs[tl.pair];
            const EgFeat F = tl.feat == 0 ? f0 : (tl.feat == 1 ? f1 : f2);
            const bool cosine = tl.feat == 2;
            const int i = tl.m0 + row;
            const bool rv = i < pd.M;
            const int n_stage = (F.nkc + EG_STAGE_CH - 1) / EG_STAGE_CH, n_span = (n_stage + EG_SPAN - 1) / EG_SPAN;
            // reference-row norms of this tile (the previous tile's readers are done: named barrier of the epilogue warps)
            asm volatile("bar.sync 1, 256;" ::: "memory");
            for (int c = tid; c < EG_TN; c += 32 * EG_EPI_WARPS) s_yy[c] = (tl.n0 + c < pd.N && !cosine) ? F.sq[pd.blk_b + tl.n0 + c] : 0.0f;
            asm volatile("bar.sync 1, 256;" ::: "memory");
            const float xx = (rv && !cosine) ? F.sq[pd.blk_a + i] : 0.0f;
            float acc[EG_TN / 2];
#pragma unroll
            for (int q = 0; q < EG_TN / 2; ++q) acc[q] = 0.0f;
            for (int sp = 0; sp < n_span; ++sp, ++u) {
                const uint32_t buf = u & 1;
                mbar_wait(mb_tfull + buf, (u >> 1) & 1u);
                tc_fence_after();
                const uint32_t taddr = tmem_base + ((uint32_t)(32 * lq) << 16) + buf * EG_TN + half * (EG_TN / 2);
#pragma unroll
                for (int c0 = 0; c0 < EG_TN / 2; c0 += 16) {
                    float v[16];
                    tmem_ld16(taddr + c0, v);
#pragma unroll
                    for (int q = 0; q < 16; ++q) acc[c0 + q] = __fadd_rn(acc[c0 + q], v[q]);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(mb_tempty + buf);
            }
            float* out = csm + pd.csm_off + (int64_t)tl.feat * pd.M * pd.N + (int64_t)i * pd.N + tl.n0 + half * (EG_TN / 2);
            const int ncols = min(EG_TN / 2, pd.N - tl.n0 - half * (EG_TN / 2));
            const bool vec_ok = ((pd.N & 3) == 0) && ((pd.csm_off & 3) == 0) && ((((int64_t)tl.feat * pd.M * pd.N) & 3) == 0);
            if (rv) {
#pragma unroll
                for (int c0 = 0; c0 < EG_TN / 2; c0 += 4) {
                    float r4[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        float r_;
                        if (cosine) r_ = 1.0f - acc[c0 + q];
                        else {
                            r_ = (xx + s_yy[half * (EG_TN / 2) + c0 + q]) - 2.0f * acc[c0 + q];
                            if (r_ < 0.0f) r_ = 0.0f;
                            r_ = sqrtf(r_);
                        }
                        r4[q] = r_;
                    }
                    if (c0 + 4 <= ncols && vec_ok) __stcs(reinterpret_cast<float4*>(out + c0), make_float4(r4[0], r4[1], r4[2], r4[3]));
                    else {
#pragma unroll
                        for (int q = 0; q < 4; ++q) if (c0 + q < ncols) __stcs(out + c0 + q, r4[q]);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CL == 2) eg_cluster_sync();              // nothing of the partner is still on its way into this CTA
    if (warp == 0) gm_tmem_free(tmem_base);
}
