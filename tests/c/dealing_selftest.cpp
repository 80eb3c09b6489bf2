// dealing_selftest.cpp -- the block-cyclic dealing of a multi-device handle (csrc/dealing.h), checked exhaustively on small
// cases and under UBSan: every append is cut into pieces that tile it, every piece lands at the tail of its shard, ids are the
// insertion order, and locate / global id are inverse to each other.
#include <cstdio>
#include <vector>

#include "../../llm-kubernetes-minikube-sharp4dev_b200/csrc/dealing.h"

int main() {
    int fails = 0;
#define CHECK(c) do { if (!(c)) { if (fails++ < 10) std::fprintf(stderr, "FAIL line %d: %s (B=%lld G=%d)\n", __LINE__, #c, (long long)B, G); } } while (0)
    for (int64_t B : {1, 2, 7, 64}) {
        for (int G : {1, 2, 3, 8}) {
            std::vector<std::vector<int64_t>> shard(G);       // per shard: the global ids of its rows, in local order
            int64_t count = 0;
            const int64_t appends[] = {0, 1, B - 1 > 0 ? B - 1 : 1, B + 3, 5 * B * G + 2, 1, 3 * B};
            for (int64_t n : appends) {
                int64_t covered = 0;
                rag::deal_append(count, n, B, G, [&](int s, int64_t off, int64_t rows) {
                    CHECK(off == covered && rows > 0);                                    // pieces tile the append, in order
                    CHECK(static_cast<int64_t>(shard[s].size()) == rag::shard_rows(count + off, s, B, G));   // lands at the shard's tail
                    for (int64_t r = 0; r < rows; ++r) shard[s].push_back(count + off + r);
                    covered += rows;
                });
                CHECK(covered == n);
                count += n;
                int64_t sum = 0;
                for (int s = 0; s < G; ++s) {
                    CHECK(static_cast<int64_t>(shard[s].size()) == rag::shard_rows(count, s, B, G));
                    sum += static_cast<int64_t>(shard[s].size());
                }
                CHECK(sum == count);
            }
            for (int s = 0; s < G; ++s)
                for (int64_t l = 0; l < static_cast<int64_t>(shard[s].size()); ++l) {
                    CHECK(rag::dealt_global_id(l, B, G, s) == shard[s][l]);              // id = place in the insertion order
                    int sh = -1;
                    int64_t loc = -1;
                    rag::dealt_locate(shard[s][l], B, G, &sh, &loc);
                    CHECK(sh == s && loc == l);
                    if (l) CHECK(shard[s][l] > shard[s][l - 1]);                          // local order = global order within a shard
                }
            // a prefix of the insertion order is a prefix of every shard
            for (int64_t n = 0; n <= count; n += (count / 17 + 1))
                for (int s = 0; s < G; ++s) {
                    const int64_t lim = rag::shard_rows(n, s, B, G);
                    if (lim > 0) CHECK(shard[s][lim - 1] < n);
                    if (lim < static_cast<int64_t>(shard[s].size())) CHECK(shard[s][lim] >= n);
                }
        }
    }
    std::printf("dealing_selftest %s\n", fails ? "FAILED" : "ok");
    return fails ? 1 : 0;
}
