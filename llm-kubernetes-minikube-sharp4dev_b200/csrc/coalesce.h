// coalesce.h -- the queue behind ragstore_search_coalesced, host-only and free of CUDA so that it can be
// stress-tested on its own under ThreadSanitizer (tests/c/coalesce_stress.cpp).
//
// The reference serves /rag/search and /agent_rag concurrently from one singleton index (Program.cs:43,62,106); every
// request scans the whole corpus on its own thread.  Here concurrent single-query callers are answered by one sweep:
// whoever finds no batch running while its own request is pending runs the next batch -- the oldest request's KIND
// (class_of), at most max_batch of them -- then steps down and wakes the others.  Leadership is held for ONE batch, so a
// caller returns as soon as its own answer is in, however many clients keep re-entering, and a request is only batched
// with its own kind, so one caller's unusual top_k neither changes the path nor the error of anybody else.
#pragma once
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

namespace rag {

struct CoPending {
    const float* q;
    int top_k;
    int64_t* idx;
    double* score;
    int32_t* found;
    int32_t* flags;
    int status = 0;
    bool done = false;
    std::string err;
};

struct Coalescer {
    std::mutex mu;
    std::condition_variable cv;
    std::vector<CoPending*> queue;
    bool leader = false;
    int window_us = 0, max_batch = 256;
    // what the last batch looked like (guarded by mu): drives the settle wait below
    double last_batch_us = 0.0;
    int last_batch_n = 0;

    // class_of(top_k) -> int: requests of one class may share a batch.
    // run(batch) answers every request of the batch (outputs, status, err) -- called WITHOUT the lock held.
    // Returns me.status once me has been answered.
    template <class ClassOf, class Run>
    int submit(CoPending& me, ClassOf class_of, Run run) {
        std::unique_lock<std::mutex> lk(mu);
        queue.push_back(&me);
        cv.notify_all();  // lets a leader that waits for company see that its batch has grown
        while (!me.done) {
            if (leader) {
                cv.wait(lk, [&] { return me.done || !leader; });
                continue;
            }
            leader = true;
            // Batches form by themselves: whoever arrives while a batch is executing is queued and taken, all together,
            // by the next one, so by default nobody waits (a lone caller on an idle store runs at once).  A window makes
            // every batch's first caller wait that long for company: larger batches, at the price of that much latency
            // (measured with 16 closed-loop clients on 10k x 768: 2.9 -> 10.3 queries per sweep).
            if (window_us > 0) {
                const auto deadline = std::chrono::steady_clock::now() + std::chrono::microseconds(window_us);
                cv.wait_until(lk, deadline, [&] { return static_cast<int>(queue.size()) >= max_batch; });
            } else if (last_batch_n > 1 && last_batch_us > 200.0) {
                // Closed-loop clients fall into two cohorts that take turns (the callers answered by batch A re-enter
                // while batch B runs, and the other way round), so every sweep carries half the clients.  When the store
                // is under concurrent load and sweeps are long, the leader lets the queue SETTLE first: it goes as soon
                // as nobody has joined for one short interval, and never waits longer than an eighth of the last sweep.
                // The callers just released by the previous batch are back by then, the cohorts merge, and a sweep
                // carries everybody (measured with 16 closed-loop clients on 1M x 768: 8.0 -> ~16 queries per sweep).
                // A lone caller, or a corpus small enough to answer in under 200 us, never waits.
                const double cap_us = last_batch_us / 8.0 < 200.0 ? last_batch_us / 8.0 : 200.0;
                const auto deadline = std::chrono::steady_clock::now() + std::chrono::microseconds(static_cast<long>(cap_us));
                size_t seen = queue.size();
                while (static_cast<int>(queue.size()) < max_batch && std::chrono::steady_clock::now() < deadline) {
                    if (!cv.wait_for(lk, std::chrono::microseconds(20), [&] { return queue.size() != seen; })) break;   // settled
                    seen = queue.size();
                }
            }
            const int cls = class_of(queue.front()->top_k);
            std::vector<CoPending*> batch, rest;
            for (CoPending* p : queue) {
                if (static_cast<int>(batch.size()) < max_batch && class_of(p->top_k) == cls)
                    batch.push_back(p);
                else
                    rest.push_back(p);
            }
            queue.swap(rest);
            lk.unlock();
            const auto t_run = std::chrono::steady_clock::now();
            run(batch);
            const double run_us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_run).count();
            lk.lock();
            last_batch_us = run_us;
            last_batch_n = static_cast<int>(batch.size());
            for (CoPending* p : batch) p->done = true;
            leader = false;
            cv.notify_all();
        }
        return me.status;
    }
};

}  // namespace rag
