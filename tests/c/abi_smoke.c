/* abi_smoke.c -- the C ABI used from plain C, the way a P/Invoke or cgo binding would use it: no Python, no
 * C++, no CUDA headers.  Stores a tiny corpus, searches it, checks the hand-derivable answer
 * (tests/golden/kat_v1.json "onehot_orthogonal" and "ties_lower_index_first").
 * Exit codes: 0 = pass, 3 = no GPU (library refused cleanly, RAGSTORE_ERR_NO_DEVICE), other = failure. */
#include <stdio.h>
#include <string.h>

#include "../../include/ragstore.h"

static int fail(const char* what) {
    char buf[512];
    ragstore_last_error(buf, sizeof buf);
    fprintf(stderr, "FAIL %s: %s\n", what, buf);
    return 1;
}

int main(void) {
    ragstore_t* st = NULL;
    int rc = ragstore_create(3, RAGSTORE_F32, 16, 0, RAGSTORE_CREATE_DEFAULT, &st);
    if (rc == RAGSTORE_ERR_NO_DEVICE) {
        printf("no device: library refused (there is no CPU path)\n");
        return 3;
    }
    if (rc) return fail("create");
    const float rows[4][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}, {0, 2, 0}};
    int64_t first = -1, count = 0;
    if (ragstore_add(st, &rows[0][0], 4, 3, RAGSTORE_ADD_HOST, &first) || first != 0) return fail("add");
    if (ragstore_count(st, &count) || count != 4) return fail("count");
    const float q[3] = {0, 1, 0};
    int64_t idx[4];
    double score[4];
    int32_t found = 0, flags = 0;
    if (ragstore_search(st, q, 1, 3, 4, idx, score, &found, &flags)) return fail("search");
    /* rows 1 and 3 are parallel to q; row 3 is longer, so the reference's 1e-9 in the denominator ranks it first */
    const double s3 = 2.0 / (1.0 * 2.0 + 1e-9), s1 = 1.0 / (1.0 * 1.0 + 1e-9);
    if (found != 4 || idx[0] != 3 || idx[1] != 1 || idx[2] != 0 || idx[3] != 2) return fail("order");
    if (score[0] != s3 || score[1] != s1 || score[2] != 0.0 || score[3] != 0.0) return fail("scores");
    if (!(flags & RAGSTORE_HIT_CERTIFIED) || !(flags & RAGSTORE_HIT_EXACT_ROWS)) return fail("flags");
    uint8_t cited[4];
    double best;
    int32_t low;
    if (ragstore_hit_filters(score, found, 0.35, 0.6, 0.20, cited, &best, &low)) return fail("filters");
    if (best != s3 || low || !cited[0] || !cited[1] || cited[2] || cited[3]) return fail("filter values");
    if (ragstore_search(st, q, 1, 2, 3, idx, score, &found, &flags)) return fail("search dim mismatch");
    if (found != 3 || idx[0] != 0 || idx[2] != 2 || score[0] != -1.0) return fail("dim mismatch semantics");
    if (ragstore_destroy(st)) return fail("destroy");
    /* the same corpus behind ONE handle over two shards (the same GPU listed twice when the box has only one):
     * blocks of 2 rows are dealt round-robin, ids stay insertion order, the answer is the same */
    {
        const int devices[2] = {0, 0};
        if (ragstore_create_sharded(3, RAGSTORE_F32, 16, devices, 2, 2, RAGSTORE_CREATE_DEFAULT, &st)) return fail("create_sharded");
        if (ragstore_add(st, &rows[0][0], 3, 3, RAGSTORE_ADD_HOST, &first) || first != 0) return fail("sharded add 1");
        if (ragstore_add(st, &rows[3][0], 1, 3, RAGSTORE_ADD_HOST, &first) || first != 3) return fail("sharded add 2");
        if (ragstore_count(st, &count) || count != 4) return fail("sharded count");
        if (ragstore_search(st, q, 1, 3, 4, idx, score, &found, &flags)) return fail("sharded search");
        if (found != 4 || idx[0] != 3 || idx[1] != 1 || idx[2] != 0 || idx[3] != 2) return fail("sharded order");
        if (score[0] != s3 || score[1] != s1 || score[2] != 0.0 || score[3] != 0.0) return fail("sharded scores");
        if (ragstore_destroy(st)) return fail("sharded destroy");
    }
    printf("abi_smoke ok\n");
    return 0;
}
