// dealing.h -- how a multi-device handle deals rows to its shards (host-only arithmetic, no CUDA: tested on its own by
// tests/c/dealing_selftest.cpp).  Rows are appended in insertion order (RagIndex.cs:48); blocks of B rows go to the G shards
// round-robin; a row's id stays its place in the insertion order, so (score desc, id asc) over the shards' hits is the order
// one List<RagChunk> would give (RagIndex.cs:64-65).
#pragma once
#include <cstdint>

namespace rag {

// rows of shard s among the first `total` rows of the global insertion order
inline int64_t shard_rows(int64_t total, int s, int64_t B, int G) {
    const int64_t full = total / B, rem = total % B;
    return (full / G + (s < full % G ? 1 : 0)) * B + (s == full % G ? rem : 0);
}

// global id of local row `local` of shard s (the device-side twin is global_row_id() in topk_device.cuh)
inline int64_t dealt_global_id(int64_t local, int64_t B, int G, int s, int64_t base = 0) {
    if (B <= 0) return base + local;
    return base + ((local / B) * G + s) * B + local % B;
}

// where global row g lives
inline void dealt_locate(int64_t g, int64_t B, int G, int* shard, int64_t* local) {
    const int64_t blk = g / B;
    *shard = static_cast<int>(blk % G);
    *local = (blk / G) * B + g % B;
}

// Cuts the append [first, first + n) at block boundaries and calls seg(shard, offset_in_append, rows) for every piece, in
// insertion order.  Each piece lands at the tail of its shard: shard_rows(first + offset, shard) rows are there before it.
template <class Seg>
inline void deal_append(int64_t first, int64_t n, int64_t B, int G, Seg seg) {
    for (int64_t g = first; g < first + n;) {
        const int64_t blk = g / B;
        const int64_t upto = (blk + 1) * B < first + n ? (blk + 1) * B : first + n;
        seg(static_cast<int>(blk % G), g - first, upto - g);
        g = upto;
    }
}

}  // namespace rag
