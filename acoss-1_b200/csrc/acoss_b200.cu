// acoss_b200.cu -- context + C-ABI of the B200-native acoss hot path (sm_100a).
// Interface contract: include/acoss_b200.h.  No CPU fallback anywhere in this
// file: every compute entry point launches the CUDA kernels of kernels_*.cuh.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <string>
#include <vector>

#include "../../include/acoss_b200.h"
#include "common.cuh"
#include "kernels_prep.cuh"
#include "kernels_crp.cuh"
#include "kernels_crp_band.cuh"
#include "kernels_crp_gemm.cuh"
#include "kernels_crp_split.cuh"
#include "kernels_dp.cuh"
#include "kernels_eval.cuh"
#include "kernels_snf.cuh"
#include "kernels_simple.cuh"
#include "kernels_early.cuh"
#include "kernels_early_tc.cuh"
#include "kernels_ftm.cuh"
#include "kernels_topk.cuh"

static std::string g_create_error;

// growable device buffer owned by the context
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { e = cudaMalloc(&p, bytes); want = bytes; }
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct acoss_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;    // the stream every call of this context is ordered on (own_stream, or the caller's: acoss_ctx_set_stream)
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream2 = nullptr;   // phase-split pipeline: the selection kernels run beside the next chunk's streaming kernels
    cudaEvent_t ev_cnt[2] = {}, ev_sel[2] = {};
    // pair pipeline: the per-batch descriptors / strip tasks / DP order go up on a copy stream from pinned staging, into the set
    // (batch parity) the running batch does not use, so that the device never waits for the host between batches
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_up[2] = {}, ev_done[2] = {};
    void* h_pin[2] = {nullptr, nullptr};
    size_t h_pin_cap[2] = {0, 0};
    DevBuf d_descs_alt, d_tasks_alt, d_dp_order_alt;
    float piv_nsigma = 2.5f;          // bracket half-width of the phase-split pivots in binomial sigmas ("piv_nsigma_x10")
    int split_overlap = 0;            // "split_overlap": 0 = all phase-split kernels on one stream
    std::string err;
    // track store
    int n_tracks = 0, m = 0, tau = 0;
    std::vector<int64_t> h_offs;      // frame offsets
    std::vector<int32_t> h_emb;       // embedded rows per track
    std::vector<int64_t> h_sd_off;    // self-dot table offsets
    DevBuf d_frames, d_offs, d_emb, d_sd_off, d_selfdot, d_gchroma, d_fnorm;
    DevBuf d_dp_order;                // per batch: DP hand-out order (largest pair first) + two launch counters
    DevBuf d_smpop;                   // per-track sample operands of the tensor-core CRP kernel (m = 9, tau = 1 only)
    bool have_smpop = false;
    // phase-split CRP pipeline: per-track side-B operand planes + norm planes in their shared-memory layout, byte offsets
    // (2 per track), max of the self-dot table per rotation; per-chunk scratch (pivots, counts, lists), column edges
    DevBuf d_opb, d_opn, d_op_off, d_sdmax, d_piv, d_cnt4, d_lcnt, d_slists, d_edges, d_fb;
    bool have_split = false;
    int64_t split_launches = 0;
    int count_hh = 0;                 // phase-split count pass on the hi.hi product alone ("count_hh")
    int auto_split = 1;               // crp_mode 0 picks the phase-split pipeline (acoss_ctx_set_option "auto_split")
    // SiMPle feature store (frame-major float64) + per-track chroma profiles
    int sp_tracks = 0;
    std::vector<int64_t> sp_offs;
    DevBuf d_sp_feats, d_sp_offs, d_sp_prof;
    // EarlyFusion block-feature store (mfccs, ssms, chromas), squared row norms, chroma medians; pair scratch
    int ef_tracks = 0, ef_d[3] = {0, 0, 0};
    std::vector<int64_t> ef_offs;
    DevBuf d_ef_x[3], d_ef_sq[3], d_ef_cmed, d_ef_csm, d_ef_bin;
    DevBuf d_ef_planes[3], d_ef_poff[3];   // tensor-core operand planes of the three features + per-track byte offsets
    int ef_nkc[3] = {0, 0, 0};
    bool have_ef_planes = false;
    int ef_dynamic_tiles = 1;             // CSM GEMM tiles handed out from a global counter ("ef_dynamic_tiles")
    DevBuf d_ef_cnt;
    int ef_cluster = 0;                   // "ef_cluster": CSM GEMM as clusters of two CTAs sharing the reference operand by TMA multicast
    int ef_force_simt = 0;                // acoss_ctx_set_option "ef_simt_csm": keep the fp32 SIMT tile kernel (comparison / fallback)
    int ef_csm_kernel_used = 0;           // 1 = SIMT tiles, 2 = tcgen05 GEMM (last acoss_earlyfusion_pairs call)
    double ef_ms[3] = {0, 0, 0}, ef_flop = 0;
    int64_t ef_cells = 0;
    // per-call scratch
    DevBuf d_descs, d_tasks, d_thr, d_crp, d_scores, d_misc, d_misc2, d_counters, d_lists;
    cudaEvent_t ev[8] = {};
    acoss_pair_stats stats = {};
    unsigned long long h_counters[ACOSS_N_COUNTERS] = {};
    size_t crp_budget = (size_t)8 << 30;
    int inject_alloc_failures = 0;    // test hook (acoss_ctx_set_option)
    int max_smem = 0;
    int num_sms = ACOSS_NUM_SMS;

    void set_error(const std::string& s) { err = s; }
};

#define CK(expr) ACOSS_CUDA_CHECK_RET(ctx, expr)
#define CK_LAUNCH() CK(cudaGetLastError())

extern "C" const char* acoss_version(void) { return "acoss_b200 0.1 sm_100a"; }

extern "C" const char* acoss_last_error(const acoss_ctx* ctx)
{
    return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

extern "C" int acoss_ctx_create(int device, acoss_ctx** out)
{
    if (!out) return ACOSS_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_error = std::string("no usable CUDA device: ") + cudaGetErrorString(e);
        return ACOSS_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) { g_create_error = "device index out of range"; return ACOSS_ERR_ARG; }
    e = cudaSetDevice(device);
    if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); return ACOSS_ERR_CUDA; }
    acoss_ctx* ctx = new acoss_ctx();
    ctx->device = device;
    e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); delete ctx; return ACOSS_ERR_CUDA; }
    ctx->own_stream = ctx->stream;
    for (auto& ev : ctx->ev) cudaEventCreate(&ev);
    if (cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking) != cudaSuccess) ctx->stream2 = nullptr;
    if (cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) ctx->copy_stream = nullptr;
    for (int i = 0; i < 2; ++i) {
        cudaEventCreateWithFlags(&ctx->ev_cnt[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ctx->ev_sel[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ctx->ev_up[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming);
    }
    cudaDeviceGetAttribute(&ctx->max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
    cudaDeviceGetAttribute(&ctx->num_sms, cudaDevAttrMultiProcessorCount, device);
    {   // CRP working set per batch: a quarter of the free HBM, at most 32 GiB (several DP waves per batch: with the
        // largest-first hand-out the tail of a batch is short), at least the old 8 GiB default when that much is free
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && free_b > 0)
            ctx->crp_budget = std::max<size_t>((size_t)1 << 30, std::min<size_t>((size_t)32 << 30, free_b / 4));
    }
    if (const char* s = getenv("ACOSS_EF_CLUSTER")) ctx->ef_cluster = atoi(s) != 0;
    if (const char* s = getenv("ACOSS_EF_DYNAMIC")) ctx->ef_dynamic_tiles = atoi(s) != 0;      // measurement switch, same as option "ef_cluster"
    if (const char* s = getenv("ACOSS_CRP_BUDGET_MB")) {
        long mb = atol(s);
        if (mb > 0) ctx->crp_budget = (size_t)mb << 20;
    }
    *out = ctx;
    return ACOSS_OK;
}

extern "C" void acoss_ctx_destroy(acoss_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    DevBuf* bufs[] = {&ctx->d_frames, &ctx->d_offs, &ctx->d_emb, &ctx->d_sd_off, &ctx->d_selfdot,
                      &ctx->d_gchroma, &ctx->d_fnorm, &ctx->d_smpop, &ctx->d_dp_order, &ctx->d_descs, &ctx->d_tasks, &ctx->d_descs_alt, &ctx->d_tasks_alt, &ctx->d_dp_order_alt, &ctx->d_thr, &ctx->d_crp,
                      &ctx->d_scores, &ctx->d_misc, &ctx->d_misc2, &ctx->d_counters, &ctx->d_lists,
                      &ctx->d_opb, &ctx->d_opn, &ctx->d_op_off, &ctx->d_sdmax, &ctx->d_piv, &ctx->d_cnt4, &ctx->d_lcnt, &ctx->d_slists,
                      &ctx->d_edges, &ctx->d_fb,
                      &ctx->d_sp_feats, &ctx->d_sp_offs, &ctx->d_sp_prof,
                      &ctx->d_ef_x[0], &ctx->d_ef_x[1], &ctx->d_ef_x[2], &ctx->d_ef_sq[0], &ctx->d_ef_sq[1], &ctx->d_ef_sq[2],
                      &ctx->d_ef_cmed, &ctx->d_ef_csm, &ctx->d_ef_bin, &ctx->d_ef_cnt,
                      &ctx->d_ef_planes[0], &ctx->d_ef_planes[1], &ctx->d_ef_planes[2], &ctx->d_ef_poff[0], &ctx->d_ef_poff[1], &ctx->d_ef_poff[2]};
    for (DevBuf* b : bufs) b->release();
    for (auto& ev : ctx->ev) if (ev) cudaEventDestroy(ev);
    for (int i = 0; i < 2; ++i) {
        if (ctx->ev_cnt[i]) cudaEventDestroy(ctx->ev_cnt[i]);
        if (ctx->ev_sel[i]) cudaEventDestroy(ctx->ev_sel[i]);
        if (ctx->ev_up[i]) cudaEventDestroy(ctx->ev_up[i]);
        if (ctx->ev_done[i]) cudaEventDestroy(ctx->ev_done[i]);
        if (ctx->h_pin[i]) cudaFreeHost(ctx->h_pin[i]);
    }
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

extern "C" int acoss_ctx_set_option(acoss_ctx* ctx, const char* key, int64_t value)
{
    if (!ctx || !key) return ACOSS_ERR_ARG;
    const std::string k(key);
    if (k == "crp_budget_bytes") {
        if (value <= 0) { ctx->set_error("crp_budget_bytes must be positive"); return ACOSS_ERR_ARG; }
        ctx->crp_budget = (size_t)value;
        return ACOSS_OK;
    }
    if (k == "piv_nsigma_x10") { if (value < 1 || value > 100) return ACOSS_ERR_ARG; ctx->piv_nsigma = 0.1f * (float)value; return ACOSS_OK; }
    if (k == "split_overlap") { ctx->split_overlap = value != 0; return ACOSS_OK; }
    if (k == "ef_dynamic_tiles") { ctx->ef_dynamic_tiles = value != 0; return ACOSS_OK; }
    if (k == "ef_cluster") { ctx->ef_cluster = value != 0; return ACOSS_OK; }
    if (k == "count_hh") { ctx->count_hh = value != 0; return ACOSS_OK; }
    if (k == "auto_split") { ctx->auto_split = value != 0; return ACOSS_OK; }
    if (k == "ef_simt_csm") { ctx->ef_force_simt = value != 0; return ACOSS_OK; }
    if (k == "inject_alloc_failures") {
        ctx->inject_alloc_failures = (int)std::max<int64_t>(0, value);
        return ACOSS_OK;
    }
    ctx->set_error("acoss_ctx_set_option: unknown key " + k);
    return ACOSS_ERR_ARG;
}

extern "C" void* acoss_ctx_stream(acoss_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

extern "C" int acoss_ctx_set_stream(acoss_ctx* ctx, void* stream)
{
    if (!ctx) return ACOSS_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));                 // nothing of this context is left in flight on the stream it leaves
    ctx->stream = stream ? (cudaStream_t)stream : ctx->own_stream;
    return ACOSS_OK;
}

extern "C" int acoss_ctx_synchronize(acoss_ctx* ctx)
{
    if (!ctx) return ACOSS_ERR_ARG;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    return ACOSS_OK;
}

static int embedded_len(int n, int m, int tau)
{
    if (m == 1) return n;
    int inc = m * tau;
    if (n < inc + 1) return 0;
    return (n - inc + tau - 1) / tau;
}

static int reserve(acoss_ctx* ctx, DevBuf& b, size_t bytes)
{
    cudaError_t e = b.reserve(bytes);
    if (e != cudaSuccess) {
        ctx->set_error(std::string("device allocation of ") + std::to_string(bytes) + " bytes failed: " + cudaGetErrorString(e));
        cudaGetLastError();
        return ACOSS_ERR_NOMEM;
    }
    return ACOSS_OK;
}
#define RESERVE(buf, bytes) do { int _r = reserve(ctx, buf, bytes); if (_r) return _r; } while (0)

// ---------------------------------------------------------------------------
// track store
// ---------------------------------------------------------------------------
extern "C" int acoss_upload_tracks(acoss_ctx* ctx, const float* frames_host, const int64_t* offsets_host,
                                   int32_t n_tracks, int32_t m, int32_t tau)
{
    if (!ctx) return ACOSS_ERR_ARG;
    if (!frames_host || !offsets_host || n_tracks <= 0 || m < 1 || m > 32 || tau < 1) {
        ctx->set_error("acoss_upload_tracks: bad argument");
        return ACOSS_ERR_ARG;
    }
    CK(cudaSetDevice(ctx->device));
    ctx->n_tracks = n_tracks; ctx->m = m; ctx->tau = tau;
    ctx->h_offs.assign(offsets_host, offsets_host + n_tracks + 1);
    ctx->h_emb.resize(n_tracks);
    ctx->h_sd_off.resize(n_tracks + 1);
    int64_t sd = 0;
    int max_emb = 0;
    for (int t = 0; t < n_tracks; ++t) {
        int64_t n = ctx->h_offs[t + 1] - ctx->h_offs[t];
        if (n < 0 || n > (int64_t)1 << 30) { ctx->set_error("acoss_upload_tracks: bad offsets"); return ACOSS_ERR_ARG; }
        int M = embedded_len((int)n, m, tau);
        ctx->h_emb[t] = M;       // 0 = too short; reported when such a track is used in a pair
        ctx->h_sd_off[t] = sd;
        sd += (int64_t)ACOSS_NBINS * M;
        max_emb = std::max(max_emb, M);
    }
    ctx->h_sd_off[n_tracks] = sd;
    const int64_t total = ctx->h_offs[n_tracks];
    RESERVE(ctx->d_frames, sizeof(float) * ACOSS_NBINS * (size_t)std::max<int64_t>(total, 1));
    RESERVE(ctx->d_offs, sizeof(int64_t) * (n_tracks + 1));
    RESERVE(ctx->d_emb, sizeof(int32_t) * n_tracks);
    RESERVE(ctx->d_sd_off, sizeof(int64_t) * (n_tracks + 1));
    RESERVE(ctx->d_selfdot, sizeof(float) * (size_t)std::max<int64_t>(sd, 1));
    RESERVE(ctx->d_gchroma, sizeof(float) * ACOSS_NBINS * n_tracks);
    RESERVE(ctx->d_fnorm, sizeof(float2) * (size_t)std::max<int64_t>(total, 1));
    cudaStream_t st = ctx->stream;
    CK(cudaMemcpyAsync(ctx->d_frames.p, frames_host, sizeof(float) * ACOSS_NBINS * (size_t)total, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ctx->d_offs.p, ctx->h_offs.data(), sizeof(int64_t) * (n_tracks + 1), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ctx->d_emb.p, ctx->h_emb.data(), sizeof(int32_t) * n_tracks, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(ctx->d_sd_off.p, ctx->h_sd_off.data(), sizeof(int64_t) * (n_tracks + 1), cudaMemcpyHostToDevice, st));
    {
        int threads = 128, warps_per_block = threads / 32;
        int blocks = (n_tracks + warps_per_block - 1) / warps_per_block;
        global_chroma_kernel<<<blocks, threads, 0, st>>>(ctx->d_frames.as<float>(), ctx->d_offs.as<int64_t>(),
                                                         n_tracks, ctx->d_gchroma.as<float>());
        CK_LAUNCH();
    }
    if (total > 0) {
        tc_frame_norm_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(ctx->d_frames.as<float>(), total, ctx->d_fnorm.as<float2>());
        CK_LAUNCH();
    }
    if (max_emb > 0) {
        // grid.y is limited to 65535: tracks go in y-chunks
        for (int t0 = 0; t0 < n_tracks; t0 += 65535) {
            dim3 grid((max_emb * ACOSS_NBINS + 127) / 128, std::min(65535, n_tracks - t0));
            selfdot_kernel<<<grid, 128, 0, st>>>(ctx->d_frames.as<float>(), ctx->d_offs.as<int64_t>(),
                                                 ctx->d_emb.as<int32_t>(), ctx->d_sd_off.as<int64_t>(), m, tau,
                                                 ctx->d_selfdot.as<float>(), t0);
            CK_LAUNCH();
        }
    }
    // sample operands of the tensor-core CRP kernel: 225 KB per track, built only when they fit a 6 GiB budget (the kernel
    // gathers the samples itself otherwise)
    ctx->have_smpop = false;
    if (m == 9 && tau == 1 && max_emb > 0 && (size_t)n_tracks * GM_SMP_OP_BYTES <= ((size_t)6 << 30)) {
        RESERVE(ctx->d_smpop, (size_t)n_tracks * GM_SMP_OP_BYTES);
        gm_sample_operand_kernel<<<n_tracks, 256, 0, st>>>(ctx->d_frames.as<float>(), ctx->d_offs.as<int64_t>(), ctx->d_emb.as<int32_t>(),
                                                           ctx->d_sd_off.as<int64_t>(), ctx->d_selfdot.as<float>(), ctx->d_smpop.as<unsigned char>());
        CK_LAUNCH();
        ctx->have_smpop = true;
    }
    // operand store of the phase-split CRP pipeline (kernels_crp_split.cuh): 288 B per frame, built when it fits 8 GiB
    ctx->have_split = false;
    if (ctx->have_smpop) {
        std::vector<int64_t> op_off(2 * (size_t)n_tracks);
        size_t pl_bytes = 0, nr_bytes = 0;
        int max_tiles = 0;
        for (int t = 0; t < n_tracks; ++t) {
            const int nt = sp_tiles(ctx->h_emb[t]);
            op_off[2 * t] = (int64_t)pl_bytes; op_off[2 * t + 1] = (int64_t)nr_bytes;
            if (nt > 0) { pl_bytes += (size_t)6 * (GM_TN * nt + 8) * 16; nr_bytes += (size_t)ACOSS_NBINS * GM_TN * nt * 16; }
            max_tiles = std::max(max_tiles, nt);
        }
        if (max_tiles > 0 && pl_bytes + nr_bytes <= ((size_t)8 << 30)) {
            RESERVE(ctx->d_opb, pl_bytes + 64);
            RESERVE(ctx->d_opn, nr_bytes + 64);
            RESERVE(ctx->d_op_off, sizeof(int64_t) * op_off.size());
            RESERVE(ctx->d_sdmax, sizeof(float) * ACOSS_NBINS * n_tracks);
            CK(cudaMemcpyAsync(ctx->d_op_off.p, op_off.data(), sizeof(int64_t) * op_off.size(), cudaMemcpyHostToDevice, st));
            const int P = GM_TN * max_tiles + 8, W = GM_TN * max_tiles;
            for (int t0 = 0; t0 < n_tracks; t0 += 65535) {
                const int ny = std::min(65535, n_tracks - t0);
                sp_planes_kernel<<<dim3((P + 127) / 128, ny), 128, 0, st>>>(ctx->d_frames.as<float>(), ctx->d_offs.as<int64_t>(), ctx->d_emb.as<int32_t>(),
                                                                            ctx->d_op_off.as<int64_t>(), ctx->d_opb.as<unsigned char>(), t0);
                CK_LAUNCH();
                sp_norms_kernel<<<dim3((ACOSS_NBINS * W + 127) / 128, ny), 128, 0, st>>>(ctx->d_emb.as<int32_t>(), ctx->d_sd_off.as<int64_t>(),
                                                                                         ctx->d_selfdot.as<float>(), ctx->d_op_off.as<int64_t>(),
                                                                                         ctx->d_opn.as<unsigned char>(), t0);
                CK_LAUNCH();
            }
            sp_sdmax_kernel<<<n_tracks, 256, 0, st>>>(ctx->d_emb.as<int32_t>(), ctx->d_sd_off.as<int64_t>(), ctx->d_selfdot.as<float>(),
                                                      ctx->d_sdmax.as<float>());
            CK_LAUNCH();
            ctx->have_split = true;
        }
    }
    CK(cudaStreamSynchronize(st));   // host buffers may be released by the caller
    return ACOSS_OK;
}

extern "C" int acoss_median_downsample(acoss_ctx* ctx, const float* raw_frames_host, const int64_t* raw_offsets_host,
                                       int32_t n_tracks, int32_t fac, float* out_frames_host, int64_t* out_offsets_host)
{
    if (!ctx) return ACOSS_ERR_ARG;
    if (!raw_frames_host || !raw_offsets_host || !out_frames_host || !out_offsets_host || n_tracks <= 0 || fac < 1 || fac > 256) {
        ctx->set_error("acoss_median_downsample: bad argument (need non-null buffers, n_tracks > 0 and 1 <= fac <= 256: a segment is sorted in a 256-entry per-thread array)");
        return ACOSS_ERR_ARG;
    }
    CK(cudaSetDevice(ctx->device));
    std::vector<int64_t> oo(n_tracks + 1, 0);
    int64_t max_out = 0;
    for (int t = 0; t < n_tracks; ++t) {
        int64_t T = raw_offsets_host[t + 1] - raw_offsets_host[t];
        int64_t no = (T + fac - 1) / fac;
        oo[t + 1] = oo[t] + no;
        max_out = std::max(max_out, no);
    }
    const int64_t total_raw = raw_offsets_host[n_tracks], total_out = oo[n_tracks];
    DevBuf raw, roff, ooff, outb;
    auto cleanup = [&]() { raw.release(); roff.release(); ooff.release(); outb.release(); };
    if (raw.reserve(sizeof(float) * ACOSS_NBINS * (size_t)std::max<int64_t>(total_raw, 1)) != cudaSuccess ||
        roff.reserve(sizeof(int64_t) * (n_tracks + 1)) != cudaSuccess ||
        ooff.reserve(sizeof(int64_t) * (n_tracks + 1)) != cudaSuccess ||
        outb.reserve(sizeof(float) * ACOSS_NBINS * (size_t)std::max<int64_t>(total_out, 1)) != cudaSuccess) {
        cleanup(); cudaGetLastError();
        ctx->set_error("acoss_median_downsample: device allocation failed");
        return ACOSS_ERR_NOMEM;
    }
    cudaStream_t st = ctx->stream;
    cudaError_t e = cudaMemcpyAsync(raw.p, raw_frames_host, sizeof(float) * ACOSS_NBINS * (size_t)total_raw, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(roff.p, raw_offsets_host, sizeof(int64_t) * (n_tracks + 1), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(ooff.p, oo.data(), sizeof(int64_t) * (n_tracks + 1), cudaMemcpyHostToDevice, st);
    for (int t0 = 0; e == cudaSuccess && max_out > 0 && t0 < n_tracks; t0 += 65535) {      // grid.y <= 65535
        dim3 grid((unsigned)((max_out * ACOSS_NBINS + 127) / 128), std::min(65535, n_tracks - t0));
        median_downsample_kernel<256><<<grid, 128, 0, st>>>(raw.as<float>(), roff.as<int64_t>(), ooff.as<int64_t>(),
                                                            n_tracks, fac, outb.as<float>(), t0);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_frames_host, outb.p, sizeof(float) * ACOSS_NBINS * (size_t)total_out, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    cleanup();
    if (e != cudaSuccess) { ctx->set_error(std::string("acoss_median_downsample: ") + cudaGetErrorString(e)); return ACOSS_ERR_CUDA; }
    memcpy(out_offsets_host, oo.data(), sizeof(int64_t) * (n_tracks + 1));
    return ACOSS_OK;
}

#include "api_pairs.inc"
#include "api_matrix.inc"
#include "api_snf.inc"
#include "api_simple.inc"
#include "api_early.inc"
#include "api_features.inc"
#include "api_topk.inc"

// ---------------------------------------------------------------------------
// Allocation-free variant of acoss_cover_similarity_from_crp (no counterpart in the reference: added
// for callers that cannot call cudaMalloc -- CUDA graph capture, tight serving loops, harnesses that forbid it).
// The scratch the Qmax / Dmax kernel needs (pair descriptors + the CRPs re-coded into the pitched "delta" byte layout) lives in
// ONE file-scope device array: every GPU that loads the library holds ACOSS_NOALLOC_SCRATCH_BYTES of it until exit, every call
// on a device shares it (calls that overlap on the GPU race), and it keeps the previous call's contents (the error flag is
// cleared at the start of every call; everything else is overwritten before it is read).
// In host code `g_noalloc_scratch` is only a stand-in for the symbol: its device address comes from cudaGetSymbolAddress, once
// per device with that device current.
// ---------------------------------------------------------------------------
#define NOALLOC_DESC_BYTES (ACOSS_NOALLOC_MAX_MATRICES * sizeof(PairDesc))      // 64 KB
#define NOALLOC_FLAG_OFF NOALLOC_DESC_BYTES                                     // int error flag (+ padding to 256 B)
#define NOALLOC_CRP_OFF (NOALLOC_DESC_BYTES + 256)
#define ACOSS_NOALLOC_SCRATCH_BYTES (NOALLOC_CRP_OFF + ACOSS_NOALLOC_MAX_CRP_BYTES + 256)
static __device__ __align__(256) unsigned char g_noalloc_scratch[ACOSS_NOALLOC_SCRATCH_BYTES];
static void* g_noalloc_ptr[64] = {};          // device address of the array, per device ordinal

// plain 0 / 1 bytes -> delta code (0x02 / 0xFF) in the pitched layout of the DP kernel; padding columns = "no recurrence"
__global__ void noalloc_recode_kernel(const PairDesc* __restrict__ descs, int n, const uint8_t* __restrict__ src,
                                      uint8_t* __restrict__ dst, int* __restrict__ flag)
{
    for (int k = blockIdx.y; k < n; k += gridDim.y) {
        const PairDesc pd = descs[k];
        const int64_t total = (int64_t)pd.M * pd.pitch;
        const uint8_t* s = src + pd.q_frame_off;            // q_frame_off: byte offset of this matrix in the caller's packed input
        for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
            const int i = (int)(e / pd.pitch), j = (int)(e - (int64_t)i * pd.pitch);
            uint8_t code = (uint8_t)CRP_DELTA_ZERO;
            if (j < pd.N) {
                const uint8_t v = s[(int64_t)i * pd.N + j];
                if (v > 1) atomicExch(flag, 1);
                if (v == 1) code = (uint8_t)CRP_DELTA_ONE;
            }
            dst[pd.crp_off + e] = code;
        }
    }
}

extern "C" int acoss_cover_similarity_from_crp_noalloc(acoss_ctx* ctx, const uint8_t* crps_dev, const int32_t* Ms, const int32_t* Ns,
                                                       int32_t n, int32_t alignment, float* out_scores_dev)
{
    if (!ctx) return ACOSS_ERR_ARG;
    if (!crps_dev || !Ms || !Ns || n < 0 || !out_scores_dev || alignment < 0 || alignment > 1) { ctx->set_error("bad argument"); return ACOSS_ERR_ARG; }
    if (n > ACOSS_NOALLOC_MAX_MATRICES) {
        ctx->set_error("acoss_cover_similarity_from_crp_noalloc: more than ACOSS_NOALLOC_MAX_MATRICES (1024) matrices in one call");
        return ACOSS_ERR_TOO_LONG;
    }
    CK(cudaSetDevice(ctx->device));
    if (n == 0) return ACOSS_OK;
    if (ctx->device >= 64) { ctx->set_error("acoss_cover_similarity_from_crp_noalloc: device ordinal >= 64"); return ACOSS_ERR_ARG; }
    if (!g_noalloc_ptr[ctx->device]) {          // set-up, once per device, with that device current
        void* p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_noalloc_ptr[ctx->device] = p;
    }
    unsigned char* scratch = static_cast<unsigned char*>(g_noalloc_ptr[ctx->device]);
    int maxN = 0, max_min = 0;
    for (int k = 0; k < n; ++k) {
        if (Ms[k] < 1 || Ns[k] < 1) { ctx->set_error("empty matrix"); return ACOSS_ERR_ARG; }
        maxN = std::max(maxN, Ns[k]); max_min = std::max(max_min, std::min(Ms[k], Ns[k]));
    }
    const DpConfig cfg = pick_dp_config(maxN, max_min, true);
    if (cfg.G == 0) {
        ctx->set_error("acoss_cover_similarity_from_crp_noalloc: matrices wider than 2560 columns (or scores beyond int16 half units) need the allocating entry point");
        return ACOSS_ERR_TOO_LONG;
    }
    const int gran = 64 * cfg.G;
    static thread_local PairDesc descs[ACOSS_NOALLOC_MAX_MATRICES];      // host staging: no heap allocation either
    size_t bytes = 0, src = 0;
    for (int k = 0; k < n; ++k) {
        PairDesc pd = {};
        pd.M = Ms[k]; pd.N = Ns[k]; pd.pitch = (Ns[k] + gran - 1) / gran * gran;
        pd.crp_off = (int64_t)bytes; pd.out_idx = k; pd.q_frame_off = (int64_t)src;
        bytes += (size_t)pd.M * pd.pitch; src += (size_t)pd.M * pd.N;
        descs[k] = pd;
    }
    if (bytes > (size_t)ACOSS_NOALLOC_MAX_CRP_BYTES) {
        ctx->set_error("acoss_cover_similarity_from_crp_noalloc: the matrices need " + std::to_string(bytes) +
                       " bytes of scratch, the static array holds ACOSS_NOALLOC_MAX_CRP_BYTES (16 MiB)");
        return ACOSS_ERR_TOO_LONG;
    }
    cudaStream_t st = ctx->stream;
    PairDesc* d_descs = reinterpret_cast<PairDesc*>(scratch);
    int* d_flag = reinterpret_cast<int*>(scratch + NOALLOC_FLAG_OFF);
    uint8_t* d_crp = scratch + NOALLOC_CRP_OFF;
    CK(cudaMemsetAsync(d_flag, 0, sizeof(int), st));
    CK(cudaMemcpyAsync(d_descs, descs, sizeof(PairDesc) * n, cudaMemcpyHostToDevice, st));
    noalloc_recode_kernel<<<dim3(64, std::min(n, 1024)), 256, 0, st>>>(d_descs, n, crps_dev, d_crp, d_flag);
    CK_LAUNCH();
    CK(launch_dp_any(ctx, cfg, alignment, d_descs, n, d_crp, out_scores_dev, 0.5f, 0.5f, maxN));
    int flag = 0;
    CK(cudaMemcpyAsync(&flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (flag) { ctx->set_error("CoverSongSimilarity: non-binary element in input"); return ACOSS_ERR_ARG; }
    return ACOSS_OK;
}
