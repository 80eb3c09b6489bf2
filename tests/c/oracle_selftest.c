/* oracle_selftest.c -- the CPU oracle (oracle/ragindex_oracle.c) exercised from plain C so that it can run under
 * AddressSanitizer + UndefinedBehaviorSanitizer (tests/test_sanitizers.py): generator, multi-threaded scoring, selection
 * with ties / NaN / k <= 0 / k > n, empty index, dimension mismatch.  Known answers as in tests/golden/kat_v1.json. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

double rag_oracle_cosine(const float*, int64_t, const float*, int64_t);
int64_t rag_oracle_query(const float*, int64_t, int64_t, const float*, int64_t, int64_t, int64_t*, double*, int);
void rag_synth_rows_mt(float*, int64_t, int64_t, int64_t, int64_t, uint64_t, int, int);

static int fails = 0;
#define CHECK(c) do { if (!(c)) { fprintf(stderr, "FAIL line %d: %s\n", __LINE__, #c); fails++; } } while (0)

int main(void) {
    const float rows[4][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {0, 2, 0}};
    const float q[3] = {0, 1, 0};
    int64_t idx[8];
    double sc[8];
    CHECK(rag_oracle_query(&rows[0][0], 4, 3, q, 3, 4, idx, sc, 1) == 4);
    CHECK(idx[0] == 3 && idx[1] == 1 && idx[2] == 0 && idx[3] == 2);
    CHECK(sc[0] == 2.0 / (1.0 * 2.0 + 1e-9) && sc[1] == 1.0 / (1.0 + 1e-9) && sc[2] == 0.0);
    CHECK(rag_oracle_query(&rows[0][0], 4, 3, q, 3, 0, idx, sc, 1) == 1 && idx[0] == 3);       /* Math.Max(1, topK) */
    CHECK(rag_oracle_query(&rows[0][0], 4, 3, q, 3, -5, idx, sc, 1) == 1);
    CHECK(rag_oracle_query(&rows[0][0], 4, 3, q, 2, 3, idx, sc, 1) == 3 && sc[0] == -1.0 && idx[2] == 2);   /* length mismatch */
    CHECK(rag_oracle_query(&rows[0][0], 0, 3, q, 3, 5, idx, sc, 1) == 0);                        /* empty index */
    const float nanrows[3][2] = {{NAN, 1}, {0, 1}, {1, 0}};
    const float q2[2] = {0, 1};
    CHECK(rag_oracle_query(&nanrows[0][0], 3, 2, q2, 2, 3, idx, sc, 1) == 3 && idx[0] == 1 && idx[1] == 2 && idx[2] == 0 && isnan(sc[2]));
    /* a seeded corpus through every thread count: same answer, and rows of odd sizes at the buffer's edge */
    const int64_t n = 5003, d = 67;
    float* corpus = (float*)malloc(sizeof(float) * (size_t)(n * d));
    float* query = (float*)malloc(sizeof(float) * (size_t)d);
    rag_synth_rows_mt(corpus, 0, n, d, d, 1234, 0, 5);
    rag_synth_rows_mt(query, 0, 1, d, d, 5678, 0, 1);
    int64_t i1[100], i7[100];
    double s1[100], s7[100];
    CHECK(rag_oracle_query(corpus, n, d, query, d, 100, i1, s1, 1) == 100);
    CHECK(rag_oracle_query(corpus, n, d, query, d, 100, i7, s7, 7) == 100);
    for (int j = 0; j < 100; j++) CHECK(i1[j] == i7[j] && s1[j] == s7[j]);
    for (int j = 1; j < 100; j++) CHECK(s1[j - 1] > s1[j] || (s1[j - 1] == s1[j] && i1[j - 1] < i1[j]));
    double n2 = 0;
    for (int c = 0; c < d; c++) n2 += (double)corpus[17 * d + c] * corpus[17 * d + c];
    CHECK(fabs(n2 - 1.0) < 1e-6);
    free(corpus);
    free(query);
    printf("oracle_selftest %s\n", fails ? "FAILED" : "ok");
    return fails ? 1 : 0;
}
