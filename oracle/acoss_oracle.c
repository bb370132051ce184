/*
 * acoss_oracle.c -- CPU ORACLE.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product path (acoss-1_b200/)
 * never links, imports or calls anything in oracle/.
 *
 * What it is: a plain-C restatement of the arithmetic behind the reference's
 * Serra09 / LateFusionChen per-pair call
 *     csm = ChromaCrossSimilarity(frameStackSize=m, frameStackStride=tau,
 *                                 binarizePercentile=kappa, oti=True)(query, reference)
 *     _, score = CoverSongSimilarity(alignmentType='serra09'|'chen17',
 *                                    distanceType='symmetric')(csm)
 * (reference call sites: acoss/algorithms/rqa_serra09.py:60-67,
 *  acoss/algorithms/latefusion_chen.py:63-71) and of the reference's in-tree
 * constrained Smith-Waterman (acoss/algorithms/utils/alignment_tools.py:7-46).
 *
 * PARITY UNPINNED for the essentia part: the two Algorithm objects live in the
 * third-party C++ library essentia (MTG/essentia, PyPI "essentia"), an
 * UNPINNED optional dependency of the reference (setup.py:53, install.sh:32)
 * whose source is not under /root/reference and which cannot be installed in
 * this image.  The reference's own tests pin no numeric result at that
 * boundary (test/basetest.py:1-21 is import-only).  This file therefore
 * restates essentia's published algorithm under the decisions D1..D9 of
 * SURVEY.md Appendix A; each function cites the decision it follows.  The
 * Smith-Waterman part IS pinned: it is checked against the reference's own
 * numba code run verbatim (tests/golden/, oracle/make_golden.py).
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC (oracle/Makefile).
 * -ffp-contract=off matters: D5 fixes "fp32, sequential sums, no FMA".
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define NBINS 12

/* ---- D1: global chroma + optimal transposition index --------------------- */

/* g = sum over frames (sequential, fp32), divided by its max element unless the
 * max is zero (essentia globalAverageChroma + normalize). */
void oracle_global_chroma(const float *x, int n, float *g)
{
    for (int c = 0; c < NBINS; ++c) g[c] = 0.0f;
    for (int t = 0; t < n; ++t)
        for (int c = 0; c < NBINS; ++c) g[c] = g[c] + x[(size_t)t * NBINS + c];
    float mx = g[0];
    for (int c = 1; c < NBINS; ++c) if (g[c] > mx) mx = g[c];
    if (mx != 0.0f)
        for (int c = 0; c < NBINS; ++c) g[c] = g[c] / mx;
}

/* D1: v_k = <g_q, roll(g_r, k)> for k = 0..12 (13 shifts), roll = right
 * rotation (np.roll); oti = FIRST argmax.  Each dot is a sequential fp32 sum. */
int oracle_oti_from_global(const float *gq, const float *gr)
{
    int best = 0;
    float bestv = 0.0f;
    for (int k = 0; k <= NBINS; ++k) {
        float acc = 0.0f;
        for (int c = 0; c < NBINS; ++c) {
            /* roll(g_r, k)[c] = g_r[(c - k) mod 12] */
            int src = (c - k) % NBINS; if (src < 0) src += NBINS;
            acc = acc + gq[c] * gr[src];
        }
        if (k == 0 || acc > bestv) { bestv = acc; best = k; }
    }
    return best;
}

int oracle_oti(const float *q, int nq, const float *r, int nr)
{
    float gq[NBINS], gr[NBINS];
    oracle_global_chroma(q, nq, gq);
    oracle_global_chroma(r, nr, gr);
    return oracle_oti_from_global(gq, gr);
}

/* ---- D3/D4: delay embedding ---------------------------------------------- */

/* number of stacked rows: loop "for (i = 0; i < n - m*tau; i += tau)";
 * m == 1 returns the frames unchanged.  <= 0 means "too short" (essentia throws). */
int oracle_embedded_len(int n, int m, int tau)
{
    if (m == 1) return n;
    int inc = m * tau;
    if (n < inc + 1) return 0;
    return (n - inc + tau - 1) / tau;
}

/* out: (M, 12*m); row t = concat(frame[t*tau + s*tau], s = 0..m-1); the frames
 * are rotated right by `rot` bins first (rot = oti for the reference track, 0
 * for the query; D1 "every reference frame is rotated right by oti"). */
static void embed(const float *x, int n, int m, int tau, int rot, float *out, int M)
{
    (void)n;
    for (int t = 0; t < M; ++t) {
        int i = (m == 1) ? t : t * tau;
        for (int s = 0; s < m; ++s) {
            const float *f = x + (size_t)(i + s * tau) * NBINS;
            float *o = out + ((size_t)t * m + s) * NBINS;
            for (int c = 0; c < NBINS; ++c) {
                int src = (c - rot) % NBINS; if (src < 0) src += NBINS;
                o[c] = f[src];
            }
        }
    }
}

/* ---- D5: distance ---------------------------------------------------------- */

static float dotf(const float *a, const float *b, int d)
{
    float acc = 0.0f;
    for (int c = 0; c < d; ++c) acc = acc + a[c] * b[c];
    return acc;
}

/* ---- D6: percentile -------------------------------------------------------- */

static int cmp_float(const void *a, const void *b)
{
    float x = *(const float *)a, y = *(const float *)b;
    return (x > y) - (x < y);
}

/* sorted ascending input.  q in percent. */
static float percentile_sorted(const float *v, int n, float qpercent)
{
    if (n == 1) return v[0];
    float q = qpercent / 100.0f;
    float k = (float)(n - 1) * q;
    float fl = floorf(k), ce = ceilf(k);
    if (fl == ce) return v[(int)fl];
    float d0 = v[(int)fl] * (ce - k);
    float d1 = v[(int)ce] * (k - fl);
    return d0 + d1;
}

/* exported for tests */
float oracle_percentile(const float *v_unsorted, int n, float kappa)
{
    float *tmp = (float *)malloc(sizeof(float) * (size_t)n);
    memcpy(tmp, v_unsorted, sizeof(float) * (size_t)n);
    qsort(tmp, (size_t)n, sizeof(float), cmp_float);
    float p = percentile_sorted(tmp, n, kappa * 100.0f);
    free(tmp);
    return p;
}

/* ---- a3: ChromaCrossSimilarity -------------------------------------------- */

/* Returns 0 on success, -1 if a track is too short to stack (essentia throws).
 * crp: (M, N) uint8 0/1, row-major, rows = query.  Optional outputs (may be
 * NULL): dist (M,N) f32, thr_q (M), thr_r (N), n_clamped (# negative radicands).
 */
int oracle_crp(const float *q, int nq, const float *r, int nr,
               int m, int tau, float kappa, int use_oti,
               uint8_t *crp, float *dist_out, float *thr_q_out, float *thr_r_out,
               int *oti_out, int64_t *n_clamped_out)
{
    int M = oracle_embedded_len(nq, m, tau);
    int N = oracle_embedded_len(nr, m, tau);
    if (M <= 0 || N <= 0) return -1;
    int d = NBINS * m;
    int oti = use_oti ? oracle_oti(q, nq, r, nr) : 0;
    if (oti_out) *oti_out = oti;

    float *X = (float *)malloc(sizeof(float) * (size_t)M * d);
    float *Y = (float *)malloc(sizeof(float) * (size_t)N * d);
    float *D = (float *)malloc(sizeof(float) * (size_t)M * N);
    float *thr_q = (float *)malloc(sizeof(float) * (size_t)M);
    float *thr_r = (float *)malloc(sizeof(float) * (size_t)N);
    float *tmp = (float *)malloc(sizeof(float) * (size_t)(M > N ? M : N));
    embed(q, nq, m, tau, 0, X, M);
    embed(r, nr, m, tau, oti % NBINS, Y, N);

    int64_t clamped = 0;
    for (int i = 0; i < M; ++i) {
        const float *xi = X + (size_t)i * d;
        for (int j = 0; j < N; ++j) {
            const float *yj = Y + (size_t)j * d;
            /* D5: item = dot(x,x) - 2*dot(x,y) + dot(y,y), left to right */
            float item = dotf(xi, xi, d) - 2.0f * dotf(xi, yj, d) + dotf(yj, yj, d);
            if (item < 0.0f) { item = 0.0f; ++clamped; }
            D[(size_t)i * N + j] = sqrtf(item);
        }
    }
    float qp = kappa * 100.0f;
    for (int j = 0; j < N; ++j) {                   /* reference-axis thresholds */
        for (int i = 0; i < M; ++i) tmp[i] = D[(size_t)i * N + j];
        qsort(tmp, (size_t)M, sizeof(float), cmp_float);
        thr_r[j] = percentile_sorted(tmp, M, qp);
    }
    for (int i = 0; i < M; ++i) {                   /* query-axis thresholds */
        memcpy(tmp, D + (size_t)i * N, sizeof(float) * (size_t)N);
        qsort(tmp, (size_t)N, sizeof(float), cmp_float);
        thr_q[i] = percentile_sorted(tmp, N, qp);
    }
    /* D7: R_ij = [D_ij <= thr_q[i]] * [D_ij <= thr_r[j]]  (Heaviside(0) = 1) */
    for (int i = 0; i < M; ++i)
        for (int j = 0; j < N; ++j) {
            float v = D[(size_t)i * N + j];
            crp[(size_t)i * N + j] = (uint8_t)((v <= thr_q[i]) && (v <= thr_r[j]));
        }
    if (dist_out) memcpy(dist_out, D, sizeof(float) * (size_t)M * N);
    if (thr_q_out) memcpy(thr_q_out, thr_q, sizeof(float) * (size_t)M);
    if (thr_r_out) memcpy(thr_r_out, thr_r, sizeof(float) * (size_t)N);
    if (n_clamped_out) *n_clamped_out = clamped;
    free(X); free(Y); free(D); free(thr_q); free(thr_r); free(tmp);
    return 0;
}

/* ---- a4/a5, D8/D9: CoverSongSimilarity ------------------------------------ */

static inline float gamma_state(uint8_t v, float dis_onset, float dis_ext)
{
    return v == 1 ? dis_onset : dis_ext;
}
static inline float max4f(float a, float b, float c, float d)
{
    float m = a; if (b > m) m = b; if (c > m) m = c; if (d > m) m = d; return m;
}
static inline float max3f(float a, float b, float c)
{
    float m = a; if (b > m) m = b; if (c > m) m = c; return m;
}

/* alignment: 0 = 'serra09' (Qmax), 1 = 'chen17' (Dmax).  Returns the raw
 * maximum of the score matrix (D9).  score_out (M,N) f32 optional. */
float oracle_cover_similarity(const uint8_t *R, int M, int N, int alignment,
                              float dis_onset, float dis_ext, float *score_out)
{
    float *S = (float *)calloc((size_t)M * (size_t)N, sizeof(float));
    float best = 0.0f;
#define S_(i, j) S[(size_t)(i) * N + (j)]
#define R_(i, j) R[(size_t)(i) * N + (j)]
    for (int i = 2; i < M; ++i) {
        for (int j = 2; j < N; ++j) {
            float c1, c2, c3, v;
            if (alignment == 0) {
                if (R_(i, j) == 1) {
                    v = max3f(S_(i-1, j-1), S_(i-2, j-1), S_(i-1, j-2)) + 1.0f;
                } else {
                    c1 = S_(i-1, j-1) - gamma_state(R_(i-1, j-1), dis_onset, dis_ext);
                    c2 = S_(i-2, j-1) - gamma_state(R_(i-2, j-1), dis_onset, dis_ext);
                    c3 = S_(i-1, j-2) - gamma_state(R_(i-1, j-2), dis_onset, dis_ext);
                    v = max4f(0.0f, c1, c2, c3);
                }
            } else {
                float p2 = S_(i-2, j-1) + (float)R_(i-1, j);
                float p3 = S_(i-1, j-2) + (float)R_(i, j-1);
                if (R_(i, j) == 1) {
                    v = max3f(S_(i-1, j-1), p2, p3) + 1.0f;
                } else {
                    c1 = S_(i-1, j-1) - gamma_state(R_(i-1, j-1), dis_onset, dis_ext);
                    c2 = p2 - gamma_state(R_(i-2, j-1), dis_onset, dis_ext);
                    c3 = p3 - gamma_state(R_(i-1, j-2), dis_onset, dis_ext);
                    v = max4f(0.0f, c1, c2, c3);
                }
            }
            S_(i, j) = v;
            if (v > best) best = v;
        }
    }
#undef S_
#undef R_
    if (score_out) memcpy(score_out, S, sizeof(float) * (size_t)M * N);
    free(S);
    return best;
}

/* ---- a9: constrained Smith-Waterman (alignment_tools.py:7-46) ------------- */

/* delta_func as written (Q7): 0 if value_a > 0 else gap_extension (-0.7);
 * the gap_opening branch is unreachable.  float64 like the numba code. */
double oracle_sw_constrained(const uint8_t *B, int M, int N)
{
    double best = 0.0;
    if (N < 4 || M < 4) return best;
    double *S = (double *)calloc((size_t)M * (size_t)N, sizeof(double));
#define S_(i, j) S[(size_t)(i) * N + (j)]
#define B_(i, j) B[(size_t)(i) * N + (j)]
    for (int i = 3; i < M; ++i) {
        for (int j = 3; j < N; ++j) {
            double mv = B_(i-1, j-1) == 0 ? -1.0 : 1.0;
            double d1 = S_(i-1, j-1) + mv + (B_(i-2, j-2) > 0 ? 0.0 : -0.7);
            double d2 = S_(i-2, j-1) + mv + (B_(i-3, j-2) > 0 ? 0.0 : -0.7);
            double d3 = S_(i-1, j-2) + mv + (B_(i-2, j-3) > 0 ? 0.0 : -0.7);
            double v = d1; if (d2 > v) v = d2; if (d3 > v) v = d3; if (0.0 > v) v = 0.0;
            S_(i, j) = v;
            if (v > best) best = v;
        }
    }
#undef S_
#undef B_
    free(S);
    return best;
}

/* ---- batch driver used by the CPU baseline timing -------------------------- */

/* frames: concatenated (sum n_t, 12) f32; offs[n_tracks+1] frame offsets;
 * pairs (P,2) int32 (query, reference).  out_q / out_d: raw Qmax / Dmax maxima
 * (out_d may be NULL).  Single-threaded per call, like essentia; the harness
 * fans out over threads.  cells_out: sum of (M-2)*(N-2) (may be NULL).
 * Returns 0, or -(k+1) if pair k had a too-short track. */
int oracle_serra09_pairs(const float *frames, const int64_t *offs,
                         const int32_t *pairs, int64_t P,
                         int m, int tau, float kappa, int use_oti,
                         float dis_onset, float dis_ext,
                         float *out_q, float *out_d, int64_t *cells_out)
{
    int64_t cells = 0;
    for (int64_t k = 0; k < P; ++k) {
        int a = pairs[2 * k], b = pairs[2 * k + 1];
        const float *q = frames + (size_t)offs[a] * NBINS;
        const float *r = frames + (size_t)offs[b] * NBINS;
        int nq = (int)(offs[a + 1] - offs[a]), nr = (int)(offs[b + 1] - offs[b]);
        int M = oracle_embedded_len(nq, m, tau), N = oracle_embedded_len(nr, m, tau);
        if (M <= 0 || N <= 0) return (int)-(k + 1);
        uint8_t *crp = (uint8_t *)malloc((size_t)M * N);
        if (oracle_crp(q, nq, r, nr, m, tau, kappa, use_oti, crp, 0, 0, 0, 0, 0) != 0) {
            free(crp); return (int)-(k + 1);
        }
        if (out_q) out_q[k] = oracle_cover_similarity(crp, M, N, 0, dis_onset, dis_ext, 0);
        if (out_d) out_d[k] = oracle_cover_similarity(crp, M, N, 1, dis_onset, dis_ext, 0);
        if (M > 2 && N > 2) cells += (int64_t)(M - 2) * (N - 2);
        free(crp);
    }
    if (cells_out) *cells_out = cells;
    return 0;
}
