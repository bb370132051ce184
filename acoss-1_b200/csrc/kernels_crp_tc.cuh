// kernels_crp_tc.cuh -- tcgen05 (5th-gen tensor core) producer of the sliding-Gram tiles used by
// crp_strip_band_kernel<1> (sm_100a).
//
// The approximate radicand is rho~(i, j) = sum_{k<9} G2[i+k][j+k] with
//     G2[a][b] = |f_a|^2 + |f_b|^2 - 2 <f_a, f_b>          (12-d chroma frames)
// (see kernels_crp_band.cuh).  G2 is one small GEMM per tile: 128 side-B frames (MMA M) x 32 strip
// frames (MMA N, 24 used), in TF32 with the split  x = x_hi + x_lo  (both TF32-exact):
//     side-B row (A operand, 8 chunks of 4 floats):  [ b_hi(12) | b_lo(12) | 1 1 nb_hi nb_lo | 0 0 0 0 ]
//     strip row  (B operand, 6 k-slices of 8):       S1 = -2a_hi[0..7]   S2 = [-2a_hi[8..11], -2a_hi[0..3]]
//                                                    S3 = -2a_hi[4..11]  S4 = [na_hi na_lo 1 1 0 0 0 0]
//                                                    T1 = -2a_lo[0..7]   T2 = [-2a_lo[8..11], 0 0 0 0]
// Six accumulating MMAs with the A-operand k-slices (chunks 0-1, 2-3, 4-5, 6-7, 0-1, 2-3) give
//     -2(a_hi.b_hi + a_hi.b_lo + a_lo.b_hi) + |a|^2 + |b|^2
// without duplicating b_hi in shared memory (a k-slice is any two 16-byte chunks one LBO apart).
// All products are exact in fp32 (11-bit x 11-bit significands); the dropped a_lo.b_lo term, the
// rounding of the lo parts and the fp32 accumulation inside the tensor core are covered by the error
// bound E of the band kernel.
//
// Data path per tile of 256 side-B frames: raw frames (48 B) + per-frame |f|^2 split (8 B) from L2 ->
// hi/lo split in registers -> canonical no-swizzle K-major UMMA layout in shared memory ->
// 12 x tcgen05.mma.kind::tf32 (M=128, N=32, K=8) issued by one thread, accumulators in TMEM ->
// tcgen05.commit on an mbarrier -> tcgen05.ld (32x32b) by all 16 warps -> G2 tile in shared
// memory for the diagonal window sums.
#pragma once
#include "common.cuh"

#define TC_CHUNKS 8              // 16-byte chunks per side-B operand row
#define TC_SLICES 6              // k-slices (MMAs) per block
#define TC_BLK 128               // side-B frames per MMA (M)
#define TC_NROWS 32              // strip frames per MMA (N); 24 used
#define TC_TW 256                // side-B frames per tile (2 MMA blocks)
#define TC_TMEM_COLS 64          // 2 blocks x 32 accumulator columns

__device__ __forceinline__ float tf32_rna(float x)
{
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
// x = hi + lo with hi, lo exactly representable in TF32 up to the final rounding of lo
__device__ __forceinline__ void tf32_split(float x, float& hi, float& lo)
{
    hi = tf32_rna(x);
    lo = tf32_rna(__fsub_rn(x, hi));
}

// per-frame |f|^2 (12-term fmaf chain) split into TF32 hi / lo
__global__ void tc_frame_norm_kernel(const float* __restrict__ frames, int64_t n_frames, float2* __restrict__ fnorm)
{
    const int64_t f = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= n_frames) return;
    const float* fp = frames + f * ACOSS_NBINS;
    float nn = 0.f;
#pragma unroll
    for (int c = 0; c < ACOSS_NBINS; ++c) nn = fmaf(fp[c], fp[c], nn);
    float nh, nl;
    tf32_split(nn, nh, nl);
    fnorm[f] = make_float2(nh, nl);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid)
{
    const int sz = valid ? 16 : 0;      // src-size 0 -> 16 bytes of zeros
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra WAIT_DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* holder_smem)      // one full warp
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(holder_smem)), "n"(TC_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free(uint32_t taddr)              // the same warp
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(TC_TMEM_COLS) : "memory");
}

// shared-memory matrix descriptor, K-major, no swizzle: core matrix = 8 rows x 16 B (rows 16 B apart);
// lbo = bytes between the two 16-byte chunks of a k-slice, sbo = bytes between 8-row groups
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;                                   // descriptor version (sm_100)
    return d;                                                 // base_offset 0, lbo_mode 0, layout SWIZZLE_NONE
}

// instruction descriptor: D = F32, A = B = TF32, both K-major, M = 128, N = 32, dense
#define TC_IDESC ((1u << 4) | (2u << 7) | (2u << 10) | ((TC_NROWS >> 3) << 17) | ((TC_BLK >> 4) << 24))

__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"((uint32_t)TC_IDESC), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// 16 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16])
{
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int q = 0; q < 16; ++q) v[q] = __uint_as_float(r[q]);
}

// The same load split in two, so that the next 16 columns are in flight while the previous ones are processed: the registers
// of `r` are written asynchronously and may only be read after tmem_ld_wait16(r) (tcgen05.wait::ld waits for EVERY load of the
// thread still in flight, so the wait for one batch goes before the issue of the next).
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[16])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait16(uint32_t (&r)[16])
{
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                   "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :
                 : "memory");
}

// shared-memory operand layouts (bytes), rows 16 B apart inside a chunk column
//   OpB (side-B frames, MMA A operand): [block q (2)][chunk c (8)][row m (128)][16 B]          = 32 768 B
//   OpA (strip frames,  MMA B operand): [slice s (6)][chunk (2)][row n (32)][16 B]             =  6 144 B
#define TC_OPB_BYTES (2 * TC_CHUNKS * TC_BLK * 16)
#define TC_OPA_BYTES (TC_SLICES * 2 * TC_NROWS * 16)

// element e (0..7) of k-slice s of the strip-frame operand row, from the staged frame
// (afr_row: -2 * bins, already rotated; |f|^2 at index 12).  -2a is exact, so split(-2a) = -2 split(a).
__device__ __forceinline__ float tc_opa_elem(const float* __restrict__ afr_row, int s, int e, bool valid)
{
    if (!valid) return 0.f;
    int bin = -1;
    bool want_lo = false;
    switch (s) {
    case 0: bin = e; break;                                    // S1
    case 1: bin = e < 4 ? 8 + e : e - 4; break;                // S2
    case 2: bin = 4 + e; break;                                // S3
    case 4: bin = e; want_lo = true; break;                    // T1
    case 5: bin = e < 4 ? 8 + e : -1; want_lo = true; break;   // T2
    default: break;                                            // S4
    }
    if (s == 3) {
        float nh, nl;
        tf32_split(afr_row[ACOSS_NBINS], nh, nl);
        return e == 0 ? nh : (e == 1 ? nl : (e < 4 ? 1.f : 0.f));
    }
    if (bin < 0) return 0.f;
    float hi, lo;
    tf32_split(afr_row[bin], hi, lo);
    return want_lo ? lo : hi;
}
