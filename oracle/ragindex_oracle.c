/*
 * ragindex_oracle.c -- CPU restatement of the reference's retrieval hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing on the product path may link, import or
 * call this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker or the
 * timed CPU baseline.
 *
 * PARITY UNPINNED: the reference ships no tests, no golden vectors and no
 * stored embeddings for this path, and its own implementation (C#, net9.0)
 * cannot run in this environment (no dotnet/mono).  Equivalence to the C# is
 * by inspection of the ~20 lines restated here; the oracle is pinned only
 * against hand-derivable known answers (tests/golden/kat_*.json) and against
 * an independent numpy restatement (oracle/oracle_np.py).
 *
 * Reference (paths relative to /root/reference/Minimal_Agent_RAG/Minimal_RAG):
 *   Helpers/RagIndex.cs:124-135   static double Cosine(float[] a, float[] b)
 *   Helpers/RagIndex.cs:59-67     QueryAsync: Select(Cosine) ->
 *                                 OrderByDescending(Score) -> Take(max(1,k))
 *   Program.cs:326-327            RagChunk / RagHit records (float[] in,
 *                                 double Score out)
 *
 * Arithmetic contract that must hold for this file to be the oracle
 * (build with: gcc -O2 -ffp-contract=off, no -ffast-math, x86-64 SSE2):
 *   - a[i]*b[i] is an fp32 multiply rounded to fp32 (C# float*float);
 *   - the fp32 product is widened to fp64 and added to an fp64 accumulator,
 *     strictly in index order, three accumulators starting at +0.0;
 *   - dot / (sqrt(na) * sqrt(nb) + 1e-9) entirely in fp64;
 *   - length mismatch returns exactly -1.0;
 *   - ordering is a STABLE descending sort on the fp64 score with
 *     double.CompareTo semantics (NaN is smaller than every number, NaN==NaN,
 *     -0.0 == +0.0), so equal scores keep insertion order (lower row first).
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <float.h>
#if defined(__FAST_MATH__)
#error "the oracle must not be built with -ffast-math"
#endif
#if FLT_EVAL_METHOD != 0
#error "the oracle needs float*float evaluated in fp32 (x86-64 SSE2, FLT_EVAL_METHOD == 0)"
#endif

/* RagIndex.cs:124-135 */
double rag_oracle_cosine(const float *a, int64_t a_len, const float *b, int64_t b_len)
{
    if (a_len != b_len) return -1.0;                 /* :126 */
    double dot = 0.0, na = 0.0, nb = 0.0;            /* :127 */
    for (int64_t i = 0; i < a_len; i++) {            /* :128 */
        float p0 = a[i] * b[i];                      /* fp32 product, rounded   :130 */
        float p1 = a[i] * a[i];                      /*                          :131 */
        float p2 = b[i] * b[i];                      /*                          :132 */
        dot += (double)p0;
        na += (double)p1;
        nb += (double)p2;
    }
    return dot / (sqrt(na) * sqrt(nb) + 1e-9);       /* :134 */
}

/* double.CompareTo as used by Comparer<double>.Default: <0, 0, >0 */
static int cmp_double_dotnet(double x, double y)
{
    if (x < y) return -1;
    if (x > y) return 1;
    if (x == y) return 0;
    /* at least one NaN: NaN is less than any number, equal to NaN */
    int xn = isnan(x), yn = isnan(y);
    if (xn) return yn ? 0 : -1;
    return 1;
}

/* "a ranks strictly before b" under OrderByDescending + stable tie-break */
static int ranks_before(double sa, int64_t ia, double sb, int64_t ib)
{
    int c = cmp_double_dotnet(sa, sb);
    if (c != 0) return c > 0;
    return ia < ib;
}

/* Score every row.  rows: n x d fp32 row-major; row_dim is the stored row
 * length (the reference never checks it against the query at insert time). */
void rag_oracle_score_all(const float *rows, int64_t n, int64_t row_dim,
                          const float *q, int64_t q_dim, double *out_scores)
{
    for (int64_t r = 0; r < n; r++)
        out_scores[r] = rag_oracle_cosine(q, q_dim, rows + r * row_dim, row_dim);
}

typedef struct {
    const float *rows; int64_t r0, r1, row_dim; const float *q; int64_t q_dim; double *out;
} score_job;

static void *score_worker(void *p)
{
    score_job *j = (score_job *)p;
    for (int64_t r = j->r0; r < j->r1; r++)
        j->out[r] = rag_oracle_cosine(j->q, j->q_dim, j->rows + r * j->row_dim, j->row_dim);
    return NULL;
}

/* Row-parallel variant: per-row arithmetic unchanged (rows are independent). */
void rag_oracle_score_all_mt(const float *rows, int64_t n, int64_t row_dim,
                             const float *q, int64_t q_dim, double *out_scores, int n_threads)
{
    if (n_threads <= 1 || n < 2 * (int64_t)n_threads) {
        rag_oracle_score_all(rows, n, row_dim, q, q_dim, out_scores);
        return;
    }
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)n_threads);
    score_job *jobs = (score_job *)malloc(sizeof(score_job) * (size_t)n_threads);
    int64_t per = (n + n_threads - 1) / n_threads;
    for (int t = 0; t < n_threads; t++) {
        int64_t r0 = per * t, r1 = r0 + per;
        if (r0 > n) r0 = n;
        if (r1 > n) r1 = n;
        jobs[t] = (score_job){rows, r0, r1, row_dim, q, q_dim, out_scores};
        pthread_create(&th[t], NULL, score_worker, &jobs[t]);
    }
    for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
}

/* Stable descending selection of the first `take` entries (RagIndex.cs:64-66).
 * A bounded insertion list is exactly what "stable sort then Take" returns. */
int64_t rag_oracle_select(const double *scores, int64_t n, int64_t top_k,
                          int64_t *out_idx, double *out_score)
{
    int64_t take = top_k < 1 ? 1 : top_k;            /* Math.Max(1, topK)  :65 */
    if (take > n) take = n;
    int64_t have = 0;
    for (int64_t r = 0; r < n; r++) {
        double s = scores[r];
        if (have == take && !ranks_before(s, r, out_score[have - 1], out_idx[have - 1]))
            continue;
        int64_t pos = have < take ? have : take - 1;
        while (pos > 0 && ranks_before(s, r, out_score[pos - 1], out_idx[pos - 1])) {
            out_score[pos] = out_score[pos - 1];
            out_idx[pos] = out_idx[pos - 1];
            pos--;
        }
        out_score[pos] = s;
        out_idx[pos] = r;
        if (have < take) have++;
    }
    return have;
}

/* The whole path: QueryAsync after the embedding call (RagIndex.cs:62-66).
 * Returns the number of hits written = min(n, max(1, top_k)). */
int64_t rag_oracle_query(const float *rows, int64_t n, int64_t row_dim,
                         const float *q, int64_t q_dim, int64_t top_k,
                         int64_t *out_idx, double *out_score, int n_threads)
{
    if (n <= 0) return 0;
    double *scores = (double *)malloc(sizeof(double) * (size_t)n);
    rag_oracle_score_all_mt(rows, n, row_dim, q, q_dim, scores, n_threads);
    int64_t got = rag_oracle_select(scores, n, top_k, out_idx, out_score);
    free(scores);
    return got;
}

/* ------------------------------------------------------------------------
 * Synthetic corpus generator shared (bit for bit) with the CUDA generator in
 * csrc/synth.cu.  Integer-only sampling (Irwin-Hall sum of four 16-bit
 * uniforms, centred) so that no transcendental enters; the row is normalised
 * with an exact integer sum of squares, one IEEE sqrt and IEEE divisions in
 * fp64, then rounded once to fp32 (and optionally once more, RNE, to bf16).
 * Not part of the reference: SURVEY.md section 8(d) "Synthetic inputs".
 * ---------------------------------------------------------------------- */
static inline uint64_t splitmix64(uint64_t z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static inline int32_t synth_sample(uint64_t row_key, int64_t col)
{
    uint64_t u = splitmix64(row_key ^ ((uint64_t)col * 0xD1B54A32D192ED03ull + 0x8CB92BA72F3D8DD7ull));
    int32_t s = (int32_t)(u & 0xFFFF) + (int32_t)((u >> 16) & 0xFFFF) +
                (int32_t)((u >> 32) & 0xFFFF) + (int32_t)((u >> 48) & 0xFFFF);
    return s - 131070;
}

static inline float bf16_round_f32(float x)
{
    uint32_t u;
    memcpy(&u, &x, 4);
    if ((u & 0x7F800000u) != 0x7F800000u) {          /* finite: round to nearest even */
        u += 0x7FFFu + ((u >> 16) & 1u);
    }
    u &= 0xFFFF0000u;
    memcpy(&x, &u, 4);
    return x;
}

void rag_synth_rows(float *out, int64_t row0, int64_t n_rows, int64_t dim, int64_t ld,
                    uint64_t seed, int round_bf16)
{
    /* the samples of a row are drawn once and kept (same values as drawing them again for the second pass) */
    int32_t *smp = (int32_t *)malloc(sizeof(int32_t) * (size_t)(dim > 0 ? dim : 1));
    for (int64_t r = 0; r < n_rows; r++) {
        uint64_t row_key = splitmix64(seed + 0x632BE59BD9B4E019ull * (uint64_t)(row0 + r + 1));
        int64_t n2 = 0;
        for (int64_t c = 0; c < dim; c++) {
            int64_t s = smp[c] = synth_sample(row_key, c);
            n2 += s * s;
        }
        double norm = sqrt((double)n2);
        float *o = out + r * ld;
        for (int64_t c = 0; c < dim; c++) {
            double v = n2 ? (double)smp[c] / norm : 0.0;
            float f = (float)v;
            o[c] = round_bf16 ? bf16_round_f32(f) : f;
        }
        for (int64_t c = dim; c < ld; c++) o[c] = 0.0f;
    }
    free(smp);
}

typedef struct { float *out; int64_t row0, n, dim, ld; uint64_t seed; int bf; } synth_job;
static void *synth_worker(void *p)
{
    synth_job *j = (synth_job *)p;
    rag_synth_rows(j->out, j->row0, j->n, j->dim, j->ld, j->seed, j->bf);
    return NULL;
}

void rag_synth_rows_mt(float *out, int64_t row0, int64_t n_rows, int64_t dim, int64_t ld,
                       uint64_t seed, int round_bf16, int n_threads)
{
    if (n_threads <= 1 || n_rows < 4 * (int64_t)n_threads) {
        rag_synth_rows(out, row0, n_rows, dim, ld, seed, round_bf16);
        return;
    }
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)n_threads);
    synth_job *jobs = (synth_job *)malloc(sizeof(synth_job) * (size_t)n_threads);
    int64_t per = (n_rows + n_threads - 1) / n_threads;
    for (int t = 0; t < n_threads; t++) {
        int64_t a = per * t, b = a + per;
        if (a > n_rows) a = n_rows;
        if (b > n_rows) b = n_rows;
        jobs[t] = (synth_job){out + a * ld, row0 + a, b - a, dim, ld, seed, round_bf16};
        pthread_create(&th[t], NULL, synth_worker, &jobs[t]);
    }
    for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
}
