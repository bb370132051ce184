/*
 * acoss_b200.h -- C-ABI of the B200-native all-pairs cover-song-similarity hot path.
 *
 * This is the drop-in boundary.  The reference (acoss,
 * pure Python) reaches its per-pair arithmetic through two essentia Algorithm
 * objects and a handful of numpy/numba free functions; a maintainer binds this
 * library with ctypes at exactly those call sites.  Every
 * entry point cites the reference interface it replaces (paths relative to the
 * reference repository root).
 *
 * Conventions: extern "C"; plain pointers and sizes; int return (0 = ok, <0 =
 * ACOSS_ERR_*; text via acoss_last_error); caller-owned input/output buffers;
 * the context owns all device memory; one context per GPU; no exceptions cross
 * the ABI; no CPU fallback -- every compute entry point fails with
 * ACOSS_ERR_CUDA if the device is unusable.
 *
 * Pointer residency is part of each signature: "_host" parameters are ordinary
 * host memory (numpy buffers); "_dev" parameters are device pointers valid on
 * the context's device (e.g. torch tensors' data_ptr()).
 */
#ifndef ACOSS_B200_H
#define ACOSS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct acoss_ctx acoss_ctx;

enum {
    ACOSS_OK = 0,
    ACOSS_ERR_ARG = -1,         /* bad argument (null pointer, bad size, unknown enum) */
    ACOSS_ERR_CUDA = -2,        /* CUDA runtime / launch failure, no usable device     */
    ACOSS_ERR_TOO_SHORT = -3,   /* a track is too short to stack m frames (essentia raises) */
    ACOSS_ERR_TOO_LONG = -4,    /* a track exceeds the on-chip strip budget            */
    ACOSS_ERR_NOMEM = -5,       /* device allocation failed                            */
    ACOSS_ERR_STATE = -6        /* call order (e.g. pairs before upload)               */
};

/* Parameters of the Serra09 / LateFusionChen pair kernel.  Replaces the keyword
 * arguments of essentia ChromaCrossSimilarity(frameStackSize, frameStackStride,
 * binarizePercentile, oti) and CoverSongSimilarity(disOnset, disExtension) as
 * used at acoss/algorithms/rqa_serra09.py:60-64 and latefusion_chen.py:63-68. */
typedef struct acoss_crp_params {
    int32_t m;              /* frameStackSize, default 9                      */
    int32_t tau;            /* frameStackStride, default 1                    */
    float   kappa;          /* binarizePercentile, default 0.095              */
    int32_t oti;            /* 1 = transpose reference by OTI (default)       */
    float   dis_onset;      /* CoverSongSimilarity disOnset, default 0.5      */
    float   dis_extension;  /* CoverSongSimilarity disExtension, default 0.5  */
    int32_t crp_mode;       /* 0 = auto, 1 = exact-everywhere strips,
                               2 = approximate (FMA sliding Gram) + rigorous band + exact recheck,
                               3 = same with tcgen05 G2 tiles, 4 = tcgen05 GEMM cross-similarity
                               (128 x 256 tiles, Hankel operands) + band + exact recheck in ONE kernel,
                               5 = the same arithmetic as a phase-split pipeline (pivot / count / select /
                               binarise kernels, TMA-staged precomputed operand planes).
                               Every mode returns the same bits.                 */
} acoss_crp_params;

/* Timing / accounting of the last pair call (device time by CUDA events on the
 * context's stream).  bench.py's roofline block is computed from these. */
typedef struct acoss_pair_stats {
    double  ms_total;        /* whole call, device events                      */
    double  ms_prep;         /* OTI + descriptors                              */
    double  ms_crp;          /* cross-similarity + thresholds + binarise       */
    double  ms_dp;           /* Qmax/Dmax kernels only                         */
    int64_t dp_cells;        /* sum (M-2)*(N-2) per DP run                     */
    int64_t crp_cells;       /* sum M*N                                        */
    int64_t crp_bytes;       /* bytes of CRP written to / read from HBM        */
    int64_t n_launches;      /* kernels launched by this call                  */
    int64_t n_dp_launches;
    int64_t n_batches;
    int64_t n_exact_cells;   /* cells computed with the oracle's exact op order */
    int64_t n_fallback_rows; /* band mode: rows that fell back to the exact-row path */
    int64_t crp_kernel_mask; /* which CRP kernels ran (one bit per batch decision), ACOSS_CRP_KERNEL_*:
                                auto mode picks per batch from the track lengths, so a downgrade is visible here */
    int64_t dp_kernel_mask;  /* which alignment kernels ran, ACOSS_DP_KERNEL_*              */
} acoss_pair_stats;

enum {
    ACOSS_CRP_KERNEL_EXACT = 1,    /* crp_strip_exact_kernel: every cell in the oracle's op order      */
    ACOSS_CRP_KERNEL_BAND = 2,     /* crp_strip_band_kernel<0>: sliding Gram on the fp32 pipes + band  */
    ACOSS_CRP_KERNEL_BAND_TC = 4,  /* crp_strip_band_kernel<1>: tcgen05 G2 tiles                       */
    ACOSS_CRP_KERNEL_GEMM = 8,     /* crp_strip_gemm_kernel: tcgen05 cross-similarity GEMM, one kernel */
    ACOSS_CRP_KERNEL_SPLIT = 16,   /* crp_sample/count/select/binarise kernels (phase-split pipeline)  */
    ACOSS_DP_KERNEL_ROWPAR = 1,    /* dp_rowpar_kernel (DPX int16x2, HBM-bound)                        */
    ACOSS_DP_KERNEL_GENERIC = 2    /* dp_generic_kernel (fp32, any gammas / very long rows)            */
};

/* ---- context ------------------------------------------------------------- */

/* Creates a context on CUDA device `device` (cudaSetDevice + stream).  */
int acoss_ctx_create(int device, acoss_ctx **out);
void acoss_ctx_destroy(acoss_ctx *ctx);
/* Last error text of this context (or of the failed create when ctx == NULL). */
const char *acoss_last_error(const acoss_ctx *ctx);
/* Library identification: "acoss_b200 <version> sm_100a". */
const char *acoss_version(void);
/* The CUDA stream all work of this context is ordered on (cudaStream_t). */
void *acoss_ctx_stream(acoss_ctx *ctx);
/* Blocks until all work queued on the context's stream has finished. */
int acoss_ctx_synchronize(acoss_ctx *ctx);
/* Orders every later call of this context on the caller's stream (a cudaStream_t of the context's device, e.g.
 * torch.cuda.current_stream().cuda_stream); NULL returns to the context's own stream.  The stream the context leaves is
 * synchronised first.  The `_dev` entry points then need no extra synchronisation against the caller's own kernels. */
int acoss_ctx_set_stream(acoss_ctx *ctx, void *stream);

/* Tuning / test knobs of a context.  Keys:
 *   "crp_budget_bytes"       per-batch CRP working set of acoss_serra09_pairs* (default: a quarter of the free HBM,
 *                            at most 32 GiB; also settable by the environment variable ACOSS_CRP_BUDGET_MB at creation)
 *   "auto_split"             1 (default): crp_mode 0 (auto) picks the phase-split CRP pipeline (crp_mode 5) for batches whose reference
 *                            tracks have >= 1024 embedded rows; 0: auto never uses it
 *   "piv_nsigma_x10"         bracket half-width of the phase-split pivots in tenths of a binomial sigma (default 25)
 *   "split_overlap"          1: selection kernels on a second stream beside the next chunk's streaming kernels (default 0: no gain measured)
 *   "count_hh"               1: the phase-split count pass multiplies the TF32 hi parts only (half of the tensor work, a ten times
 *                            wider candidate band in the selection); results are bit-identical either way.  An experiment, not a
 *                            setting: the band overflows the selection's 32-candidate pool on most rows, which then take the exact
 *                            whole-row path (about 20 x slower at 2000-frame tracks)
 *   "ef_dynamic_tiles"       1 (default): the EarlyFusion CSM GEMM hands its tiles out in list order from a global counter
 *   "ef_cluster"             1: the EarlyFusion CSM GEMM runs as clusters of two CTAs that share the reference-side operand stages
 *                            by TMA multicast (bit-identical, measured slower: default 0)
 *   "ef_simt_csm"            1: acoss_earlyfusion_upload skips the tensor-core operand planes (the CSMs then run on the
 *                            fp32 SIMT tile kernel); set before the upload
 *   "inject_alloc_failures"  test hook: the next `value` device allocations of the pair pipeline's batch buffers
 *                            report ACOSS_ERR_NOMEM (exercises the halve-the-budget retry)
 * Returns ACOSS_ERR_ARG for an unknown key or a non-positive budget. */
int acoss_ctx_set_option(acoss_ctx *ctx, const char *key, int64_t value);

/* ---- track store --------------------------------------------------------- */

/* Uploads the (already median-downsampled) 12-bin chroma of all tracks:
 * frames_host = concatenation of (n_t, 12) float32 arrays, offsets_host[n_tracks+1]
 * = frame offsets.  Replaces the per-object feature cache Serra09.all_feats
 * (rqa_serra09.py:44-53, latefusion_chen.py:46-56).  Also computes, on the
 * device, each track's global chroma (essentia globalAverageChroma) and the
 * exact self dot products of its delay-embedded rows for all 12 rotations. */
int acoss_upload_tracks(acoss_ctx *ctx, const float *frames_host,
                        const int64_t *offsets_host, int32_t n_tracks,
                        int32_t m, int32_t tau);

/* Median downsampling on the device: librosa.util.sync(chroma.T,
 * arange(0, T, fac), aggregate=np.median).T for every track
 * (rqa_serra09.py:51).  raw offsets in raw frames; out_offsets_host[n_tracks+1]
 * and out_frames_host (sum ceil(T/fac), 12) are written. */
int acoss_median_downsample(acoss_ctx *ctx, const float *raw_frames_host,
                            const int64_t *raw_offsets_host, int32_t n_tracks,
                            int32_t fac, float *out_frames_host,
                            int64_t *out_offsets_host);

/* ---- Serra09 / LateFusionChen pairs -------------------------------------- */

/* For every pair (query, reference) in pairs_host[(P,2) int32]:
 *   csm = ChromaCrossSimilarity(m, tau, kappa, oti)(feat[query], feat[reference])
 *   qmax = CoverSongSimilarity('serra09','symmetric')(csm)[1]
 *   dmax = CoverSongSimilarity('chen17','symmetric')(csm)[1]     (if out_dmax != NULL)
 * i.e. the body of Serra09.similarity (rqa_serra09.py:55-69) and
 * ChenFusion.similarity (latefusion_chen.py:58-73), batched.  Outputs are the
 * raw maxima of the score matrices (float32, P each).  Host buffers. */
int acoss_serra09_pairs(acoss_ctx *ctx, const int32_t *pairs_host, int64_t n_pairs,
                        const acoss_crp_params *params,
                        float *out_qmax_host, float *out_dmax_host);

/* Same, outputs written to device memory (asynchronously on the context's
 * stream; inputs still host). */
int acoss_serra09_pairs_dev(acoss_ctx *ctx, const int32_t *pairs_host, int64_t n_pairs,
                            const acoss_crp_params *params,
                            float *out_qmax_dev, float *out_dmax_dev);

/* Query-versus-database mode: the scores of every (queries[a], db[b]) pair -- the body of
 * `similarity(idxs)` (algorithm_template.py:121-140) over a Q x N_db rectangle instead of the upper triangle -- ranked per query
 * the way getEvalStatistics ranks a row (:233-235, argsort of the similarity, the query itself left out), keeping only the best k.
 * The Q x N_db matrix exists nowhere: the device holds k + one database chunk of keys per query and merges chunk by chunk.
 * which: 0 Qmax, 1 Dmax.  norm_mode: -1 raw score (higher first); 0 score / sqrt(len_db) (Serra09.normalize_by_length,
 * rqa_serra09.py:71-83; higher first); 1 sqrt(len_db) / score (ChenFusion.normalize_by_length, latefusion_chen.py:75-85, a distance:
 * lower first).  Ties go to the lower database position.  out_idx_host (Q, k) int32: positions in db_host (-1 where the database has
 * fewer than k other tracks); out_score_host (Q, k) float32: the normalised scores.  1 <= k <= ACOSS_TOPK_MAX.
 * Shard over GPUs by splitting queries_host; there is no exchange between ranks. */
#define ACOSS_TOPK_MAX 1000
int acoss_serra09_topk(acoss_ctx *ctx, const int32_t *queries_host, int32_t n_queries, const int32_t *db_host, int32_t n_db,
                       const acoss_crp_params *params, int32_t which, int32_t norm_mode, int32_t k,
                       int32_t *out_idx_host, float *out_score_host);

/* Accounting of the last acoss_serra09_pairs* call (synchronises the stream). */
int acoss_get_pair_stats(acoss_ctx *ctx, acoss_pair_stats *out);

/* Profiling aid: raw device counters of the last pair call (clock64 cycles per phase of the
 * CRP strip kernel summed over CTAs; layout documented in kernels_crp_band.cuh). */
int acoss_debug_counters(acoss_ctx *ctx, uint64_t *out16);
/* Same, the first n of the extended set (slots 16.. hold the sub-phase cycle probes of
 * crp_strip_gemm_kernel; layout in kernels_crp_gemm.cuh). */
int acoss_debug_counters_ext(acoss_ctx *ctx, uint64_t *out, int32_t n);
/* Measurement aid: clock ticks of `reps` back-to-back 128 x 256 tiles of the streaming kernels' MMA mixes on one SM
 * (0: 14 TF32, 1: 41 TF32, 2: 14 TF32 + 14 bf16, 3: 14 bf16 adjacent K chunks, 4: 14 bf16 K chunks in two planes). */
int acoss_debug_mma_probe(acoss_ctx *ctx, int32_t reps, int64_t *out5);

/* Test hook: the binary cross-recurrence plot of ONE pair, (M, N) uint8 0/1,
 * rows = query frames (ChromaCrossSimilarity output, rqa_serra09.py:66).
 * Optional (may be NULL): thresholds thr_q[M], thr_r[N], oti. */
int acoss_crp_pair(acoss_ctx *ctx, int32_t query, int32_t reference,
                   const acoss_crp_params *params, uint8_t *out_crp_host,
                   int32_t *out_M, int32_t *out_N,
                   float *out_thr_q_host, float *out_thr_r_host, int32_t *out_oti);

/* Test hook / first kernel: CoverSongSimilarity on caller-supplied binary
 * matrices.  crps_host = concatenation of n (M_k, N_k) uint8 0/1 row-major
 * matrices; alignment 0 = 'serra09' (Qmax), 1 = 'chen17' (Dmax).
 * Replaces alignment_algo(csm)[1] (rqa_serra09.py:67). */
int acoss_cover_similarity_from_crp(acoss_ctx *ctx, const uint8_t *crps_host,
                                    const int32_t *Ms_host, const int32_t *Ns_host,
                                    int32_t n, int32_t alignment,
                                    float dis_onset, float dis_extension,
                                    float *out_scores_host);

/* Allocation-free variant of acoss_cover_similarity_from_crp.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid allocation).  The call performs no cudaMalloc / cudaFree and uses no allocating library routine: the
 * kernel's scratch (pair descriptors and the matrices re-coded into the DP kernel's pitched byte layout) lives in a static
 * device array of ACOSS_NOALLOC_MAX_CRP_BYTES + 65 KiB that every GPU loading the library holds until exit.
 *   crps_dev        DEVICE pointer: concatenation of n (M_k, N_k) uint8 0/1 row-major matrices
 *   Ms_host/Ns_host HOST arrays of n shapes;  alignment 0 = 'serra09' (Qmax), 1 = 'chen17' (Dmax); disOnset = disExtension = 0.5
 *   out_scores_dev  DEVICE pointer, n float32 (valid when the call returns: it synchronises the context's stream)
 * Limits (ACOSS_ERR_TOO_LONG beyond them): n <= ACOSS_NOALLOC_MAX_MATRICES, N_k <= 2560, sum_k M_k * pitch_k <=
 * ACOSS_NOALLOC_MAX_CRP_BYTES with pitch_k = N_k rounded up to 128 / 256 / 512 (by the widest matrix: <= 512 / <= 1280 / wider).
 * All calls on a device share the array and it keeps the last call's contents (what must start at zero is cleared by the
 * call): calls that overlap on the GPU (two contexts / streams on one device) race -- serialise them. */
#define ACOSS_NOALLOC_MAX_MATRICES 1024
#define ACOSS_NOALLOC_MAX_CRP_BYTES (16 << 20)
int acoss_cover_similarity_from_crp_noalloc(acoss_ctx *ctx, const uint8_t *crps_dev,
                                            const int32_t *Ms_host, const int32_t *Ns_host,
                                            int32_t n, int32_t alignment, float *out_scores_dev);

/* ---- N x N score matrix: scatter, mirror, length normalisation, evaluation -- */

/* D (n x n float32, device) := 0, then D[i,j] = score for each pair, then
 * D += D.T when symmetric != 0  (CoverAlgorithm.all_pairwise,
 * algorithm_template.py:168-190). */
int acoss_scores_to_matrix_dev(acoss_ctx *ctx, const int32_t *pairs_dev,
                               const float *scores_dev, int64_t n_pairs,
                               int32_t n, int32_t symmetric, float *D_dev);

/* mode 0: D[i,j] /= sqrt(len_j)        (Serra09.normalize_by_length, rqa_serra09.py:71-83)
 * mode 1: D[i,j] = sqrt(len_j)/D[i,j]  (ChenFusion.normalize_by_length, latefusion_chen.py:75-85)
 * lengths = frames per track after downsampling (int32, device). float64
 * arithmetic, float32 store, like the reference's memmap update. */
int acoss_normalize_by_length_dev(acoss_ctx *ctx, float *D_dev, int32_t n,
                                  const int32_t *lengths_dev, int32_t mode);

/* CoverAlgorithm.getEvalStatistics (algorithm_template.py:206-291) on the device.
 * D_dev: (n x n) float32, higher = more similar.  order_host[n]: the track
 * indices in clique-contiguous order (cliques sorted by size, descending);
 * clique_size_host[n]: for position p in that order the size of its clique;
 * clique_start_host[n]: position where its clique starts.
 * out_host[4 + n_tops] = MR, MRR, MDR, MAP, Top-k counts (float64).
 * out_ranks_host (may be NULL): first-hit rank per position (NaN where the
 * reference leaves NaN).  Ties are ranked by position (stable), SURVEY H5. */
int acoss_eval_statistics(acoss_ctx *ctx, const float *D_dev, int32_t n,
                          const int32_t *order_host, const int32_t *clique_size_host,
                          const int32_t *clique_start_host,
                          const int32_t *tops_host, int32_t n_tops,
                          double *out_host, double *out_ranks_host);

/* numpy's pairwise float64 summation, used by acoss_eval_statistics so that MR/MRR/MAP
 * round exactly like np.mean/np.nanmean (algorithm_template.py:262-267).  Pure host
 * arithmetic, exported so a CPU-only test can pin it against numpy. */
double acoss_np_pairwise_sum(const double *a, int64_t n);

/* ---- constrained Smith-Waterman (EarlyFusionTralie) ----------------------- */

/* smith_waterman_constrained(B) for n binary matrices
 * (acoss/algorithms/utils/alignment_tools.py:25-46).  Same packing as
 * acoss_cover_similarity_from_crp.  Scores float64. */
int acoss_sw_constrained_from_binary(acoss_ctx *ctx, const uint8_t *mats_host,
                                     const int32_t *Ms_host, const int32_t *Ns_host,
                                     int32_t n, double *out_scores_host);

/* ---- similarity network fusion (LateFusionChen, EarlyFusionTralie) -------------- */

/* doSimilarityFusion(Scores, K, niters, reg_diag) (acoss/algorithms/utils/similarity_fusion.py:188-196;
 * getW :15-36, getP :101-122, getS :124-144, doSimilarityFusionWs :146-186) on n_mats
 * (n x n) distance matrices, concatenated in mats_host.  dtype 0 = float32, 1 = float64: the
 * affinity stage runs in that type, the diffusion in float64, as numpy promotes in the reference.
 * out_fused_host: (n x n) float64.  out_Ws_host (may be NULL): the n_mats affinity matrices in the
 * input dtype (first element of the tuple the reference returns).  The reference's list aliasing
 * (iteration 0 Jacobi, then Gauss-Seidel) is reproduced. */
int acoss_snf(acoss_ctx *ctx, const void *mats_host, int32_t dtype, int32_t n_mats, int32_t n,
              int32_t K, int32_t niters, double reg_diag, double *out_fused_host, void *out_Ws_host);

/* The stages of doSimilarityFusion one at a time, for callers of the reference's free functions (getW :15-36, getP :101-122,
 * getS :124-144, doSimilarityFusionWs :146-186).  input_kind 0: mats are distances (as acoss_snf); 1: mats are the affinities Ws
 * themselves.  n_mats >= 1; the diffusion (niters > 0 or out_fused_host) needs >= 2.  Optional outputs (NULL to skip):
 * out_fused_host (n, n) float64; out_Ws_host n_mats x (n, n) in the input dtype; out_P_host n_mats x (n, n) float64 = getP(W);
 * out_S_host n_mats x (n, n) float64 = getS(W, K) as a dense matrix (K non-zeros per row). */
int acoss_snf_ex(acoss_ctx *ctx, const void *mats_host, int32_t dtype, int32_t input_kind, int32_t n_mats, int32_t n,
                 int32_t K, int32_t niters, double reg_diag, double *out_fused_host, void *out_Ws_host,
                 double *out_P_host, double *out_S_host);

/* ---- free functions of acoss/algorithms/utils on caller-supplied matrices ------------------------------------------ */

/* get_csm(X, Y) (kind 0, cross_recurrence.py:31-48) / get_csm_cosine(X, Y) (kind 1, :54-73) of two row-major float32 point
 * clouds (M, d), (N, d) -> (M, N) float32.  roll in 1..11 (kind 1, d a multiple of 12): the 12-bin frames of X are rolled
 * by `roll` first, which is what get_csm_blocked_oti (:106-134) does with roll = get_oti(C1, C2). */
int acoss_csm(acoss_ctx *ctx, const float *X_host, int32_t M, const float *Y_host, int32_t N, int32_t d,
              int32_t kind, int32_t roll, float *out_host);

/* get_oti(C1, C2) (cross_recurrence.py:76-103) for n pairs of 12-bin float32 chroma vectors, (n, 12) each: out_host[p] =
 * argmax_i sum(roll(C1_p, i) * C2_p), first maximum. */
int acoss_oti(acoss_ctx *ctx, const float *C1_host, const float *C2_host, int32_t n, int32_t *out_host);

/* csm_to_binary(D, kappa) (cross_recurrence.py:137-161): (M, N) float32 -> (M, N) uint8, the kappa-nearest columns of every
 * row (kappa = 0: all ones; kappa < 1: round(kappa * N), half to even; else int(kappa)); ties towards the lowest column.
 * N <= 2048.  ACOSS_ERR_ARG where np.argpartition raises (kth >= N). */
int acoss_csm_to_binary(acoss_ctx *ctx, const float *D_host, int32_t M, int32_t N, double kappa, uint8_t *out_bin_host);

/* getWCSM(CSMAB, k1, k2) (similarity_fusion.py:38-54, Mu = 0.5): (M, N) float32 -> (M, N) float32.  k1 <= 16, N <= 2048. */
int acoss_wcsm(acoss_ctx *ctx, const float *C_host, int32_t M, int32_t N, int32_t k1, int32_t k2, float *out_W_host);

/* ---- SiMPle (matrix profile) and FTM2D ------------------------------------------- */

/* Uploads the SiMPle features of all tracks: feats_host = concatenation of (12, L_t) row-major
 * float64 arrays (what Simple.load_features returns, acoss/algorithms/simple_silva.py:34-43),
 * col_offsets_host[n_tracks+1] = column offsets.  Also computes np.sum(seq, 1) per track (:46-47). */
int acoss_simple_upload(acoss_ctx *ctx, const double *feats_host, const int64_t *col_offsets_host,
                        int32_t n_tracks);

/* Per-track SiMPle preprocessing on the device (replaces the Python loops of
 * Simple.load_features and Simple.smooth, acoss/algorithms/simple_silva.py:34-43, 56-66): raw (T_t, 12)
 * float32 chroma -> (12, int(T_t / SKIP)) float64: mean over WIN-frame windows every SKIP frames, 'same'
 * convolution with the taps (the normalised Hann window of length win_len_smooth + 2, computed by the
 * caller), columns L2-normalised.  out_col_offsets_host[n_tracks + 1] receives the column offsets;
 * out_feats_host must hold 12 * sum_t int(T_t / SKIP) doubles (track t at 12 * offset, row-major (12, L_t)). */
int acoss_simple_preprocess(acoss_ctx *ctx, const float *raw_frames_host, const int64_t *raw_offsets_host,
                            int32_t n_tracks, int32_t win, int32_t skip, const double *taps_host,
                            int32_t n_taps, double *out_feats_host, int64_t *out_col_offsets_host);

/* For every ORDERED pair (i, j): Sj = Simple.oti(Si, Sj)[0]; out = -Simple.simple_sim(Si, Sj)
 * (simple_silva.py:45-54, 68-118, 120-126); sslen = SSLEN (default 10).  float64 scores. */
int acoss_simple_pairs(acoss_ctx *ctx, const int32_t *pairs_host, int64_t n_pairs, int32_t sslen,
                       double *out_scores_host);

/* FTM2D.similarity for all pairs at once (acoss/algorithms/ftm2d.py:85-97):
 * out_D_host[i*n + j] = exp(-sum((s_i - s_j)**2)), shingles_host = (n, d) float64 (d = 900). */
int acoss_ftm2d_allpairs(acoss_ctx *ctx, const double *shingles_host, int32_t n, int32_t d,
                         double *out_D_host);

/* Per-track FTM2D shingles on the device (replaces the body of FTM2D.load_features,
 * acoss/algorithms/ftm2d.py:52-66, with chrompwr :100-117 and btchroma_to_fftmat :120-138): raw (T_t, 12) float32 chroma and the
 * beat onsets (frame indices, concatenated; onset_offsets_host[n_tracks + 1]) of every track -> beat-synchronous median
 * (librosa.util.sync, boundaries {0} u onsets u {T}), chrompwr(PWR), |fft2| of every 12 x WIN patch (fft-shifted, flattened),
 * log(C * patch / |patch| + 1), median over the patches, unit norm.  out_shingles_host: (n_tracks, 12 * WIN) float64.
 * ACOSS_ERR_TOO_SHORT when a track has fewer than WIN beats (the reference fails there). */
int acoss_ftm2d_shingles(acoss_ctx *ctx, const float *raw_frames_host, const int64_t *raw_offsets_host,
                         const int64_t *onsets_host, const int64_t *onset_offsets_host, int32_t n_tracks,
                         double pwr, int32_t win, double C, double *out_shingles_host);

/* ---- EarlyFusionTralie ------------------------------------------------------------ */

/* Uploads the blocked features of all tracks (what EarlyFusion.load_features returns,
 * acoss/algorithms/earlyfusion_traile.py:67-154): 'mfccs' (n_blocks, d_mfcc), 'ssms' (n_blocks, d_ssm),
 * 'chromas' (n_blocks, d_chroma = 12 * chromas_per_block), all float32 and concatenated over tracks;
 * 'chroma_med' (n_tracks, 12) float32; block_offsets_host[n_tracks+1]. */
int acoss_earlyfusion_upload(acoss_ctx *ctx, const float *mfccs_host, int32_t d_mfcc,
                             const float *ssms_host, int32_t d_ssm,
                             const float *chromas_host, int32_t d_chroma,
                             const float *chroma_med_host, const int64_t *block_offsets_host,
                             int32_t n_tracks);

/* Blocked features of all tracks on the device (replaces the body of EarlyFusion.load_features,
 * acoss/algorithms/earlyfusion_traile.py:104-150, and resize_block :214-247 = skimage.transform.resize(anti_aliasing=True,
 * mode='constant'), restated: Gaussian anti-alias filter sigma = (f - 1) / 2 along time + order-1 interpolation; parity of the
 * resize is UNPINNED, skimage is absent).  Inputs, concatenated over tracks: mfcc (T_t, n_mfcc) float32 (feats['mfcc_htk'].T, NaN -> 0),
 * chroma (T_t, 12) float32, beat onsets (frame indices).  Per track n_blocks = n_beats - blocksize; per block: MFCC frames
 * [onset_b, onset_{b+blocksize-1}) resized to mfccs_per_block frames, column-centred, rows L2-normalised -> 'mfccs'; the lower
 * triangle of their Euclidean SSM -> 'ssms'; chroma frames [onset_b, onset_{b+blocksize}) resized to chromas_per_block -> 'chromas';
 * np.median(chroma, 0) -> 'chroma_med'.  out_block_offsets_host[n_tracks + 1] is always written; pass NULL for the four feature
 * outputs to size them first ((n_blocks_total, mfccs_per_block * n_mfcc), (.., mpb (mpb - 1) / 2), (.., 12 cpb), (n_tracks, 12)). */
int acoss_earlyfusion_block_features(acoss_ctx *ctx, const float *mfcc_host, int32_t n_mfcc, const int64_t *mfcc_offsets_host,
                                     const float *chroma_host, const int64_t *chroma_offsets_host,
                                     const int64_t *onsets_host, const int64_t *onset_offsets_host, int32_t n_tracks,
                                     int32_t blocksize, int32_t mfccs_per_block, int32_t chromas_per_block,
                                     float *out_mfccs_host, float *out_ssms_host, float *out_chromas_host,
                                     float *out_chroma_med_host, int64_t *out_block_offsets_host);

/* The body of EarlyFusion.similarity (earlyfusion_traile.py:157-198) for every pair (i, j):
 * get_csm x2, get_csm_blocked_oti(get_csm_cosine), csm_to_binary(kappa), getWCSM(K, K), the early
 * fusion exp(-sum W) and four smith_waterman_constrained scores.
 * out_scores_host: (n_pairs, 4) float64 in the order mfccs, ssms, chromas, early. */
int acoss_earlyfusion_pairs(acoss_ctx *ctx, const int32_t *pairs_host, int64_t n_pairs,
                            double kappa, int32_t K, double *out_scores_host);

/* Test hook: the 3 CSMs + the early matrix (4, M, N) float32 and the 4 binary matrices (4, M, N)
 * uint8 of one pair, plus its 4 scores. */
int acoss_earlyfusion_pair_matrices(acoss_ctx *ctx, int32_t i, int32_t j, double kappa, int32_t K,
                                    float *out_csm_host, uint8_t *out_bin_host,
                                    double *out_scores_host);

/* Device time (CUDA events) of the last acoss_earlyfusion_pairs call: out_ms3 = {CSM GEMMs,
 * selection + W + early, Smith-Waterman}; GEMM flop; sum of M*N. */
int acoss_earlyfusion_stats(acoss_ctx *ctx, double *out_ms3, double *out_flop, int64_t *out_cells);

/* ---- plain device memory helpers for host languages without a CUDA binding -- */
int acoss_dev_alloc(acoss_ctx *ctx, int64_t bytes, void **out_dev);
int acoss_dev_free(acoss_ctx *ctx, void *dev);
int acoss_memcpy_h2d(acoss_ctx *ctx, void *dst_dev, const void *src_host, int64_t bytes);
int acoss_memcpy_d2h(acoss_ctx *ctx, void *dst_host, const void *src_dev, int64_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* ACOSS_B200_H */
