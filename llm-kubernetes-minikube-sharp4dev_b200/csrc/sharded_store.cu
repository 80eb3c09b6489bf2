// sharded_store.cu -- ONE store handle over several GPUs of one process.
//
// The reference's index is one object in one process: a DI singleton (Program.cs:35,43) whose QueryAsync every handler
// calls (Program.cs:72,116).  For corpora larger than one GPU the drop-in must therefore keep that
// shape: ragstore_create_sharded returns the same opaque handle as ragstore_create, and ragstore_add / ragstore_search /
// ragstore_count / ragstore_search_coalesced work on it unchanged.
//
//   store   rows are dealt to the shards in blocks of `block_rows`, round-robin, in insertion order; a row's id is its
//           place in that order (the position the reference's List<RagChunk>.Add would have given it, RagIndex.cs:48);
//   search  ONE host thread launches every shard's fused scan + top-k + exact-rescoring kernel on that shard's own
//           stream (the query rides in the kernel parameters).  Each shard's last CTA stores its k rescored records
//           STRAIGHT into a gather buffer on the first shard's GPU -- plain stores over NVLink through peer access, no
//           IPC, no NCCL, no host hop -- and a merge kernel there, ordered behind the shard kernels by events, ranks
//           the G x k records (score desc, id asc: RagIndex.cs:64-65) and writes the hits into page-locked host memory;
//   batches take the tensor-core path on every shard (each shard's TMA loads read that shard's own memory only).
// A search works on a prefix of the insertion order: the published global row count is translated into per-shard row
// limits, so rows of an append still in flight on some shard are never mixed into a result.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <deque>
#include <functional>
#include <memory>
#include <thread>

#include "dealing.h"
#include "store_impl.h"

namespace rag {

struct ShardedCtx {
    cudaStream_t st0 = nullptr;            // merge stream on the first shard's device
    ragstore_cand* d_gather = nullptr;     // first device: [shard][query][req] records
    int32_t* d_flags = nullptr;            // first device: [shard][query]
    std::vector<ragstore_cand*> d_local;   // per shard, on its own device: staging when it cannot store into d_gather
    std::vector<int32_t*> d_lflags;
    std::vector<cudaEvent_t> done;         // per shard: its launches of this call have been issued and will complete
    uint8_t* h_out = nullptr;              // page-locked: [idx][score][found] of a query block, written by the merge kernel
    int32_t* h_flags = nullptr;            // page-locked: [shard][query]
    float* h_q = nullptr;                  // page-locked copy of the query block, read in place by every shard's kernels
};

// One host thread per shard, parked on its device: a batch of queries costs every shard a query upload, two tensor-map
// encodings and five launches (~0.1 ms of host time), which one caller thread would pay G times in a row per 256-query
// block; the workers pay it side by side.  Single queries stay on the caller's thread (one launch per shard).
struct ShardWorker {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    std::deque<std::function<void()>> jobs;
    bool stop = false;
};

struct ShardedState {
    std::vector<std::unique_ptr<ShardWorker>> workers;
    int n = 0;
    int64_t block_rows = 0;
    std::vector<int> devices;
    std::vector<char> direct;              // shard s stores straight into the first device's gather buffer
    std::mutex add_mu;
    std::vector<ShardedCtx> ctxs;
    std::vector<int> free_ctx;
    std::mutex ctx_mu;
    std::condition_variable ctx_cv;
};

namespace {

constexpr size_t kGatherRecords = size_t(kCtxQueries) * kMaxKp;   // per shard and context

#define SH_TRY(expr)                                                                             \
    do {                                                                                         \
        cudaError_t e__ = (expr);                                                                \
        if (e__ != cudaSuccess) return api_fail(RAGSTORE_ERR_CUDA, #expr, cudaGetErrorString(e__)); \
    } while (0)

// double.CompareTo order of OrderByDescending (NaN lowest, -0 == +0), then lower id first
uint64_t ord_key(double s) {
    if (s != s) return 0ull;
    s = s + 0.0;
    uint64_t u;
    std::memcpy(&u, &s, 8);
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}

void worker_loop(ShardWorker* w, int device) {
    cudaSetDevice(device);
    for (;;) {
        std::function<void()> job;
        {
            std::unique_lock<std::mutex> lk(w->mu);
            w->cv.wait(lk, [&] { return w->stop || !w->jobs.empty(); });
            if (w->jobs.empty()) return;   // stop requested and nothing left to do
            job = std::move(w->jobs.front());
            w->jobs.pop_front();
        }
        job();
    }
}

void post(ShardWorker* w, std::function<void()> job) {
    {
        std::lock_guard<std::mutex> lk(w->mu);
        w->jobs.push_back(std::move(job));
    }
    w->cv.notify_one();
}

void stop_workers(ShardedState* ss) {
    for (auto& w : ss->workers) {
        {
            std::lock_guard<std::mutex> lk(w->mu);
            w->stop = true;
        }
        w->cv.notify_one();
        if (w->th.joinable()) w->th.join();
    }
    ss->workers.clear();
}

void free_sharded_ctx(ShardedState* ss, ShardedCtx& c) {
    {
        DeviceGuard g(ss->devices[0]);
        if (c.st0) cudaStreamDestroy(c.st0);
        cudaFree(c.d_gather);
        cudaFree(c.d_flags);
    }
    for (int s = 0; s < ss->n; ++s) {
        DeviceGuard g(ss->devices[s]);
        if (s < static_cast<int>(c.d_local.size())) cudaFree(c.d_local[s]);
        if (s < static_cast<int>(c.d_lflags.size())) cudaFree(c.d_lflags[s]);
        if (s < static_cast<int>(c.done.size()) && c.done[s]) cudaEventDestroy(c.done[s]);
    }
    if (c.h_out) cudaFreeHost(c.h_out);
    if (c.h_flags) cudaFreeHost(c.h_flags);
    if (c.h_q) cudaFreeHost(c.h_q);
    c = ShardedCtx{};
}

int alloc_sharded_ctx(ragstore* p, ShardedCtx& c) {
    ShardedState* ss = p->sharded;
    const int G = ss->n;
    c.d_local.assign(G, nullptr);
    c.d_lflags.assign(G, nullptr);
    c.done.assign(G, nullptr);
    {
        DeviceGuard g(ss->devices[0]);
        SH_TRY(cudaStreamCreateWithFlags(&c.st0, cudaStreamNonBlocking));
        SH_TRY(cudaMalloc(&c.d_gather, sizeof(ragstore_cand) * kGatherRecords * G));
        SH_TRY(cudaMalloc(&c.d_flags, sizeof(int32_t) * kCtxQueries * G));
    }
    for (int s = 0; s < G; ++s) {
        DeviceGuard g(ss->devices[s]);
        SH_TRY(cudaEventCreateWithFlags(&c.done[s], cudaEventDisableTiming));
        if (!ss->direct[s]) {
            SH_TRY(cudaMalloc(&c.d_local[s], sizeof(ragstore_cand) * kGatherRecords));
            SH_TRY(cudaMalloc(&c.d_lflags[s], sizeof(int32_t) * kCtxQueries));
        }
    }
    // portable: every device may write into it (the merge kernel does) and read from it (the kernels read query blocks in place)
    SH_TRY(cudaHostAlloc(reinterpret_cast<void**>(&c.h_out), kGatherRecords * 16 + kCtxQueries * 8, cudaHostAllocPortable));
    SH_TRY(cudaHostAlloc(reinterpret_cast<void**>(&c.h_flags), sizeof(int32_t) * kCtxQueries * G, cudaHostAllocPortable));
    SH_TRY(cudaHostAlloc(reinterpret_cast<void**>(&c.h_q), sizeof(float) * size_t(kCtxQueries) * p->ld, cudaHostAllocPortable));
    return RAGSTORE_OK;
}

// Enqueue shard s's part of a query block on its own stream: records for queries `which` (all nq when null) into
// out_cand[q * req ...], flags into out_flags[q].  kp_force > 0: fp32 scans with that list length (the redo of queries the
// first pass could not certify).
// h_pinned: the same queries in page-locked memory that every device can read in place, or null.
int launch_shard(ragstore* sh, SearchCtx& c, int64_t n, int req, const float* h_queries, const float* h_pinned, int nq, int kp_force,
                 const std::vector<int>* which, ragstore_cand* out_cand, int32_t* out_flags) {
    const int take = static_cast<int>(std::min<int64_t>(n, req));
    if (n == 0) {
        // an empty shard contributes only padding; 0xFF bytes give id = -1 and a NaN score (ranked last, skipped)
        SH_TRY(cudaMemsetAsync(out_cand, 0xFF, sizeof(ragstore_cand) * static_cast<size_t>(req) * nq, c.st));
        // nothing to certify and nothing inexact: both flags, for every query (the context's pinned query block is idle)
        int32_t* h = reinterpret_cast<int32_t*>(c.h_q);
        for (int q = 0; q < nq; ++q) h[q] = RAGSTORE_HIT_CERTIFIED | RAGSTORE_HIT_EXACT_ROWS;
        SH_TRY(cudaMemcpyAsync(out_flags, h, sizeof(int32_t) * nq, cudaMemcpyHostToDevice, c.st));
        return RAGSTORE_OK;
    }
    const int kp = kp_force > 0 ? kp_force : choose_kp(take);
    const float* h2d_src = h_pinned;
    auto stage_queries = [&]() {   // without a shared page-locked copy, use this context's own
        if (h2d_src) return;
        std::memcpy(c.h_q, h_queries, sizeof(float) * static_cast<size_t>(nq) * sh->dim);
        h2d_src = c.h_q;
    };
    // The kernels read the queries STRAIGHT from page-locked host memory (device-addressable under unified addressing):
    // a few hundred KB of loads over PCIe instead of a copy command that would queue behind a 32 MB ingest chunk on the
    // copy engine while rows are being appended.
    if (kp_force <= 0 && !which && batch_eligible(sh, n, nq, kp)) {
        stage_queries();
        BatchSearchArgs b{};
        fill_batch_args(sh, n, take, kp, b);
        b.scan.out_flags = out_flags;
        b.scan.out_cand = out_cand;
        b.scan.cand_slots = req;
        b.d_queries = h2d_src;
        b.nq = nq;
        b.out_stride = req;
        b.scratch = c.batch_scratch;
        sh->n_batch_launches.fetch_add(1, std::memory_order_relaxed);
        const int rc = launch_batch_search(b, c.st);
        if (rc) return api_fail(RAGSTORE_ERR_CUDA, "batched search launch (shard)", cudaGetErrorString(static_cast<cudaError_t>(rc)));
        return RAGSTORE_OK;
    }
    const bool in_params = sh->dim <= kParamQueryMax;
    const int count = which ? static_cast<int>(which->size()) : nq;
    if (!in_params) stage_queries();
    for (int i = 0; i < count; ++i) {
        const int q = which ? (*which)[i] : i;
        // a shadowed fp32 shard answers from its bf16 image first; the redo pass (kp_force) is the fp32 scan
        const bool from_shadow = kp_force <= 0 && shadow_scan_pays(sh, n);
        ScanArgs a = from_shadow ? shadow_scan_args(sh, n, take, shadow_scan_kp(take)) : base_args(sh, n, take, kp);
        a.d_query = in_params ? nullptr : h2d_src + static_cast<size_t>(q) * sh->dim;
        a.scratch = c.scratch;
        a.out_flags = out_flags + q;
        a.out_cand = out_cand + static_cast<size_t>(q) * req;
        a.cand_slots = req;
        sh->n_scan_launches.fetch_add(1, std::memory_order_relaxed);
        const int rc = launch_scan(a, c.st, in_params ? h_queries + static_cast<size_t>(q) * sh->dim : nullptr);
        if (rc) return api_fail(RAGSTORE_ERR_CUDA, "scan kernel launch (shard)", cudaGetErrorString(static_cast<cudaError_t>(rc)));
    }
    return RAGSTORE_OK;
}

}  // namespace

int sharded_destroy(ragstore* p) {
    ShardedState* ss = p->sharded;
    if (ss) {
        stop_workers(ss);
        for (int s = 0; s < ss->n; ++s) {
            DeviceGuard g(ss->devices[s]);
            cudaDeviceSynchronize();
        }
        for (auto& c : ss->ctxs) free_sharded_ctx(ss, c);
    }
    for (ragstore* sh : p->shards) ragstore_destroy(sh);
    delete ss;
    delete p;
    return RAGSTORE_OK;
}

int sharded_reserve_rows(ragstore* p, int64_t n_rows) {
    ShardedState* ss = p->sharded;
    for (int s = 0; s < ss->n; ++s) {
        const int rc = ragstore_reserve_rows(p->shards[s], shard_rows(n_rows, s, ss->block_rows, ss->n));
        if (rc) return rc;
    }
    return RAGSTORE_OK;
}

int sharded_add(ragstore* p, const float* rows, int64_t n_rows, uint32_t flags, int64_t* out_first_row) {
    NvtxRange nvtx("ragstore_add (multi-device handle): deal blocks, upload + normalise per shard");
    ShardedState* ss = p->sharded;
    std::lock_guard<std::mutex> lk(ss->add_mu);
    const int64_t first = p->count.load(std::memory_order_acquire);
    if (out_first_row) *out_first_row = first;
    if (n_rows == 0) return RAGSTORE_OK;
    if (first + n_rows > p->capacity) return api_fail(RAGSTORE_ERR_CAPACITY, "ragstore_add: capacity exceeded");
    const int G = ss->n;
    const int64_t B = ss->block_rows;
    // cut [first, first + n_rows) at block boundaries; block b belongs to shard b % G and lands at that shard's tail
    std::vector<std::vector<const float*>> seg_ptr(G);
    std::vector<std::vector<int64_t>> seg_rows(G);
    deal_append(first, n_rows, B, G, [&](int s, int64_t off, int64_t n) {
        seg_ptr[s].push_back(rows + off * p->dim);
        seg_rows[s].push_back(n);
    });
    if (flags & RAGSTORE_ADD_DEVICE) {
        // a device-resident source is read in place by every shard's upload kernel: that needs the source on the shard's own
        // GPU, or on the first GPU with peer access from the shard's (what the handle set up at create)
        cudaPointerAttributes pa{};
        const bool known = cudaPointerGetAttributes(&pa, rows) == cudaSuccess && pa.type == cudaMemoryTypeDevice;
        cudaGetLastError();
        for (int s = 0; s < G; ++s) {
            if (seg_ptr[s].empty()) continue;
            const bool reachable = known && (pa.device == ss->devices[s] || (pa.device == ss->devices[0] && ss->direct[s]));
            if (!reachable)
                return api_fail(RAGSTORE_ERR_UNSUPPORTED, "ragstore_add: device-resident rows must live on the handle's first GPU (with peer access from every shard's GPU) "
                                                          "or on the GPU of the shard that receives them; pass host rows otherwise");
        }
    }
    std::vector<int> rcs(G, RAGSTORE_OK);
    std::vector<std::string> errs(G);
    auto feed = [&](int s) {
        if (seg_ptr[s].empty()) return;
        cudaSetDevice(ss->devices[s]);   // a fresh thread starts on device 0
        rcs[s] = ragstore_add_many(p->shards[s], seg_ptr[s].data(), seg_rows[s].data(), static_cast<int>(seg_ptr[s].size()), p->dim, flags, nullptr);
        if (rcs[s]) {
            char buf[512];
            ragstore_last_error(buf, sizeof buf);
            errs[s] = buf;
        }
    };
    int busy = 0;
    for (int s = 0; s < G; ++s) busy += !seg_ptr[s].empty();
    if (busy > 1 && n_rows * static_cast<int64_t>(p->dim) >= (1 << 20)) {
        // every shard's upload pipeline (PCIe copy + upload/normalise kernel) runs at the same time, one host thread each
        std::vector<std::thread> th;
        for (int s = 0; s < G; ++s) th.emplace_back(feed, s);
        for (auto& t : th) t.join();
    } else {
        int prev = -1;
        cudaGetDevice(&prev);
        for (int s = 0; s < G; ++s) feed(s);
        if (prev >= 0) cudaSetDevice(prev);
    }
    for (int s = 0; s < G; ++s)
        if (rcs[s]) return api_fail(rcs[s], "ragstore_add (shard)", errs[s].c_str());
    p->count.store(first + n_rows, std::memory_order_release);   // publish: searches work on a prefix of the insertion order
    return RAGSTORE_OK;
}

int sharded_info(ragstore* p, int64_t* out, int n) {
    int64_t v[16] = {p->dim, p->ld, p->dtype, p->count.load(), p->capacity, 0, p->device, p->sm_count, 0, 0, 0, 1, 0, 0, 0, 0};
    for (ragstore* sh : p->shards) {
        int64_t w[16] = {0};
        const int rc = ragstore_info(sh, w, 16);
        if (rc) return rc;
        v[5] += w[5];
        v[8] = w[8];
        for (int i : {9, 10, 12, 13, 14, 15}) v[i] += w[i];
    }
    for (int i = 0; i < n && i < 16; ++i) out[i] = v[i];
    return RAGSTORE_OK;
}

int sharded_search(ragstore* p, const float* queries, int n_queries, int top_k, int64_t* out_idx, double* out_score,
                   int32_t* out_found, int32_t* out_flags) {
    NvtxRange nvtx("ragstore_search (multi-device handle): per-shard scans -> peer stores -> merge");
    ShardedState* ss = p->sharded;
    const int G = ss->n;
    const int64_t N = p->count.load(std::memory_order_acquire);   // the prefix this search works on
    const int req = top_k < 1 ? 1 : top_k;
    const int take = static_cast<int>(std::min<int64_t>(N, req));
    std::vector<int64_t> n_s(G);
    for (int s = 0; s < G; ++s) n_s[s] = shard_rows(N, s, ss->block_rows, G);

    if (take > kMaxKp) {
        // More hits than one launch collects: every shard answers through its own host search (repeated scans) and the
        // G lists are merged here.  G x take records: microseconds of host work next to G x passes corpus sweeps.
        std::vector<int64_t> idx(static_cast<size_t>(G) * req);
        std::vector<double> sc(static_cast<size_t>(G) * req);
        std::vector<int32_t> fnd(G), flg(G);
        struct Rec {
            uint64_t key;
            int64_t id;
            double score;
        };
        std::vector<Rec> recs;
        for (int q = 0; q < n_queries; ++q) {
            recs.clear();
            int32_t flags_all = RAGSTORE_HIT_CERTIFIED | RAGSTORE_HIT_EXACT_ROWS;
            for (int s = 0; s < G; ++s) {
                if (n_s[s] == 0) continue;
                const int rc = search_host_limited(p->shards[s], n_s[s], queries + static_cast<size_t>(q) * p->dim, 1, p->dim, top_k,
                                                   idx.data() + static_cast<size_t>(s) * req, sc.data() + static_cast<size_t>(s) * req, &fnd[s], &flg[s]);
                if (rc) return rc;
                flags_all &= flg[s];
                for (int j = 0; j < fnd[s]; ++j)
                    recs.push_back({ord_key(sc[static_cast<size_t>(s) * req + j]), idx[static_cast<size_t>(s) * req + j], sc[static_cast<size_t>(s) * req + j]});
            }
            std::sort(recs.begin(), recs.end(), [](const Rec& a, const Rec& b) { return a.key != b.key ? a.key > b.key : a.id < b.id; });
            const int found = static_cast<int>(std::min<size_t>(recs.size(), static_cast<size_t>(take)));
            for (int j = 0; j < found; ++j) {
                out_idx[static_cast<size_t>(q) * req + j] = recs[j].id;
                out_score[static_cast<size_t>(q) * req + j] = recs[j].score;
            }
            out_found[q] = found;
            if (out_flags) out_flags[q] = flags_all;
        }
        return RAGSTORE_OK;
    }

    // ---- one context of the parent + one of every shard for the duration of the call
    int pci;
    {
        std::unique_lock<std::mutex> lk(ss->ctx_mu);
        ss->ctx_cv.wait(lk, [&] { return !ss->free_ctx.empty(); });
        pci = ss->free_ctx.back();
        ss->free_ctx.pop_back();
    }
    std::vector<int> cis(G, -1);
    for (int s = 0; s < G; ++s) cis[s] = acquire_ctx(p->shards[s]);
    struct Release {
        ragstore* p;
        ShardedState* ss;
        int pci;
        std::vector<int>& cis;
        ~Release() {
            for (size_t s = 0; s < cis.size(); ++s)
                if (cis[s] >= 0) release_ctx(p->shards[s], cis[s]);
            {
                std::lock_guard<std::mutex> lk(ss->ctx_mu);
                ss->free_ctx.push_back(pci);
            }
            ss->ctx_cv.notify_one();
        }
    } rel{p, ss, pci, cis};
    ShardedCtx& pc = ss->ctxs[pci];
    int prev_dev = -1;
    cudaGetDevice(&prev_dev);
    struct RestoreDev {
        int d;
        ~RestoreDev() {
            if (d >= 0) cudaSetDevice(d);
        }
    } restore{prev_dev};

    for (int q0 = 0; q0 < n_queries; q0 += kCtxQueries) {
        const int nq = std::min(kCtxQueries, n_queries - q0);
        const float* hq = queries + static_cast<size_t>(q0) * p->dim;
        const size_t nh = static_cast<size_t>(nq) * req;
        int64_t* h_idx = reinterpret_cast<int64_t*>(pc.h_out);
        double* h_score = reinterpret_cast<double*>(h_idx + nh);
        int32_t* h_found = reinterpret_cast<int32_t*>(h_score + nh);
        // Several queries: ONE page-locked copy of the block that every shard's host->device copy reads (a copy per shard
        // cost 8 x 0.8 MB of memcpy per 256-query block on 8 GPUs -- more host time than the sweep takes on the devices)
        const float* h_pinned = nullptr;
        if (nq > 1 || p->dim > kParamQueryMax) {
            std::memcpy(pc.h_q, hq, sizeof(float) * static_cast<size_t>(nq) * p->dim);
            h_pinned = pc.h_q;
        }
        std::vector<std::vector<int>> redo(G);   // per shard: queries to run again with the widest candidate lists
        for (int pass = 0; pass < 2; ++pass) {
            // shard s's part of this block: launches on its own stream, records (and flags) into the first GPU's memory
            auto shard_part = [&](int s) -> int {
                ragstore* sh = p->shards[s];
                SearchCtx& c = sh->ctxs[cis[s]];
                SH_TRY(cudaSetDevice(ss->devices[s]));
                ragstore_cand* dst = pc.d_gather + static_cast<size_t>(s) * nh;
                int32_t* fdst = pc.d_flags + static_cast<size_t>(s) * nq;
                ragstore_cand* oc = ss->direct[s] ? dst : pc.d_local[s];
                int32_t* of = ss->direct[s] ? fdst : pc.d_lflags[s];
                const int rc = launch_shard(sh, c, n_s[s], req, hq, h_pinned, nq, pass == 1 ? kMaxKp : 0, pass == 1 ? &redo[s] : nullptr, oc, of);
                if (rc) return rc;
                if (!ss->direct[s]) {   // no peer access from this device: the copy engine carries the records over
                    SH_TRY(cudaMemcpyPeerAsync(dst, ss->devices[0], oc, ss->devices[s], sizeof(ragstore_cand) * nh, c.st));
                    SH_TRY(cudaMemcpyPeerAsync(fdst, ss->devices[0], of, ss->devices[s], sizeof(int32_t) * nq, c.st));
                }
                SH_TRY(cudaEventRecord(pc.done[s], c.st));
                return RAGSTORE_OK;
            };
            if (pass == 0 && nq > 1 && G > 1 && !ss->workers.empty()) {
                // a block of queries: every shard's worker thread issues its part at the same time
                std::mutex mu;
                std::condition_variable cv;
                int left = G, first_rc = RAGSTORE_OK;
                std::string first_err;
                for (int s = 0; s < G; ++s)
                    post(ss->workers[s].get(), [&, s] {
                        const int rc = shard_part(s);
                        std::string err;
                        if (rc) {
                            char buf[512];
                            ragstore_last_error(buf, sizeof buf);   // the worker thread's own error text
                            err = buf;
                        }
                        std::lock_guard<std::mutex> lk(mu);
                        if (rc && !first_rc) first_rc = rc, first_err = err;
                        if (--left == 0) cv.notify_one();
                    });
                {
                    std::unique_lock<std::mutex> lk(mu);
                    cv.wait(lk, [&] { return left == 0; });
                }
                if (first_rc) return api_fail(first_rc, "ragstore_search (shard)", first_err.c_str());
            } else {
                for (int s = 0; s < G; ++s) {
                    if (pass == 1 && redo[s].empty()) continue;
                    const int rc = shard_part(s);
                    if (rc) return rc;
                }
            }
            SH_TRY(cudaSetDevice(ss->devices[0]));
            for (int s = 0; s < G; ++s)
                if (pass == 0 || !redo[s].empty()) SH_TRY(cudaStreamWaitEvent(pc.st0, pc.done[s], 0));
            // rank the G x req records of every query; the hits go straight into page-locked host memory
            const int rc = launch_merge_shards(pc.d_gather, G, nq, req, req, h_idx, h_score, h_found, pc.st0);
            if (rc) return api_fail(RAGSTORE_ERR_CUDA, "merge kernel launch", cudaGetErrorString(static_cast<cudaError_t>(rc)));
            SH_TRY(cudaMemcpyAsync(pc.h_flags, pc.d_flags, sizeof(int32_t) * static_cast<size_t>(G) * nq, cudaMemcpyDeviceToHost, pc.st0));
            SH_TRY(cudaStreamSynchronize(pc.st0));
            if (pass == 1) break;
            bool any = false;
            for (int s = 0; s < G; ++s)
                for (int q = 0; q < nq; ++q)
                    if (!(pc.h_flags[static_cast<size_t>(s) * nq + q] & RAGSTORE_HIT_CERTIFIED) && n_s[s] > 0) {
                        redo[s].push_back(q);
                        any = true;
                    }
            if (!any) break;
        }
        for (int q = 0; q < nq; ++q) {
            const int found = h_found[q];
            const size_t o = static_cast<size_t>(q0 + q) * req;
            std::memcpy(out_idx + o, h_idx + static_cast<size_t>(q) * req, sizeof(int64_t) * found);
            std::memcpy(out_score + o, h_score + static_cast<size_t>(q) * req, sizeof(double) * found);
            out_found[q0 + q] = found;
            if (out_flags) {
                int32_t f = RAGSTORE_HIT_CERTIFIED | RAGSTORE_HIT_EXACT_ROWS;
                for (int s = 0; s < G; ++s) f &= pc.h_flags[static_cast<size_t>(s) * nq + q];
                out_flags[q0 + q] = f;
            }
        }
    }
    return RAGSTORE_OK;
}

}  // namespace rag

namespace rag {

std::mutex* sharded_add_mutex(ragstore* p, int64_t* block_rows) {
    if (block_rows) *block_rows = p->sharded->block_rows;
    return &p->sharded->add_mu;
}

// the parent object around existing shard stores: peer access towards the first device, search contexts
static int sharded_wrap(ragstore* p, const int* devices, int n_devices, int64_t block_rows) {
    ShardedState* ss = p->sharded;
    ss->n = n_devices;
    ss->block_rows = block_rows;
    ss->devices.assign(devices, devices + n_devices);
    ss->direct.assign(n_devices, 1);
    p->device = devices[0];
    p->ld = p->shards[0]->ld;
    p->sm_count = p->shards[0]->sm_count;
    // a shard on another GPU stores its hit records straight into the first GPU's gather buffer when peer access exists
    for (int s = 1; s < n_devices; ++s) {
        if (devices[s] == devices[0]) continue;
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, devices[s], devices[0]) != cudaSuccess) can = 0;
        if (can) {
            DeviceGuard g(devices[s]);
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[0], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) can = 0;
        }
        cudaGetLastError();
        ss->direct[s] = can ? 1 : 0;
    }
    ss->ctxs.resize(kNumCtx);
    for (int i = 0; i < kNumCtx; ++i) {
        const int rc = alloc_sharded_ctx(p, ss->ctxs[i]);
        if (rc) return rc;
        ss->free_ctx.push_back(i);
    }
    if (n_devices > 1)
        for (int s = 0; s < n_devices; ++s) {
            ss->workers.emplace_back(new ShardWorker());
            ShardWorker* w = ss->workers.back().get();
            w->th = std::thread(worker_loop, w, devices[s]);
        }
    return RAGSTORE_OK;
}

int sharded_adopt(std::vector<ragstore*>& shards, const int* devices, int dim, int dtype, int64_t capacity_rows, int64_t block_rows,
                  int64_t count, ragstore** out) {
    const int G = static_cast<int>(shards.size());
    for (int s = 0; s < G; ++s) {
        if (shards[s]->dim != dim || shards[s]->dtype != dtype || shards[s]->count.load() != shard_rows(count, s, block_rows, G))
            return api_fail(RAGSTORE_ERR_INVALID, "ragstore_load_sharded: a shard's image does not belong to this corpus (dim, dtype or row count)");
        const int rc = ragstore_set_row_blocking(shards[s], block_rows, G, s);
        if (rc) return rc;
    }
    ragstore* p = new ragstore();
    p->sharded = new ShardedState();
    p->dim = dim;
    p->dtype = dtype;
    p->capacity = capacity_rows;
    p->shards = shards;
    const int rc = sharded_wrap(p, devices, G, block_rows);
    if (rc) {
        char buf[512];
        ragstore_last_error(buf, sizeof buf);
        const std::string why = buf;
        p->shards.clear();   // the caller keeps (and destroys) the shards
        sharded_destroy(p);
        return api_fail(rc, why.c_str());
    }
    p->count.store(count, std::memory_order_release);
    *out = p;
    return RAGSTORE_OK;
}

}  // namespace rag

using namespace rag;

extern "C" int ragstore_create_sharded(int dim, int dtype, int64_t capacity_rows, const int* devices, int n_devices,
                                       int64_t block_rows, uint32_t flags, ragstore_t** out) {
    if (!out) return api_fail(RAGSTORE_ERR_INVALID, "ragstore_create_sharded: out is null");
    *out = nullptr;
    if (!devices || n_devices < 1 || n_devices > 64 || dim <= 0 || capacity_rows <= 0 || block_rows < 0)
        return api_fail(RAGSTORE_ERR_INVALID, "ragstore_create_sharded: need 1..64 devices, positive dim and capacity, block_rows >= 0");
    if (block_rows == 0) block_rows = 65536;
    ragstore* p = new ragstore();
    p->sharded = new ShardedState();
    p->sharded->n = n_devices;
    p->sharded->devices.assign(devices, devices + n_devices);
    p->dim = dim;
    p->dtype = dtype;
    p->capacity = capacity_rows;
    auto bail = [&](int rc) {
        char buf[512];
        ragstore_last_error(buf, sizeof buf);   // keep the cause across the clean-up calls
        const std::string why = buf;
        sharded_destroy(p);
        return api_fail(rc, why.c_str());
    };
    for (int s = 0; s < n_devices; ++s) {
        ragstore* sh = nullptr;
        const int64_t cap = std::max<int64_t>(shard_rows(capacity_rows, s, block_rows, n_devices), block_rows);
        int rc = ragstore_create(dim, dtype, cap, devices[s], flags, &sh);
        if (rc) return bail(rc);
        p->shards.push_back(sh);
        rc = ragstore_set_row_blocking(sh, block_rows, n_devices, s);
        if (rc) return bail(rc);
    }
    const int rc = sharded_wrap(p, devices, n_devices, block_rows);
    if (rc) return bail(rc);
    *out = p;
    return RAGSTORE_OK;
}
