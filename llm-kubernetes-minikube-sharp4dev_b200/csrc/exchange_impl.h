// exchange_impl.h -- the object behind `ragstore_exchange_t` (csrc/exchange.cu); also used by ragstore_api.cu for the
// fused shard-scan + exchange launch.  Not part of the ABI.
#pragma once
#include "exchange_device.cuh"

struct ragstore_exchange {
    int device = 0, world = 0, rank = 0, max_records = 0, max_queries = 0, sm_count = 0;
    void* base = nullptr;                 // this rank's inbox
    size_t bytes = 0, flag_bytes = 0;
    void* peer[rag::kMaxWorld] = {};           // mapped inboxes (peer[rank] == base)
    bool connected = false;
    unsigned long long seq = 0;
    rag::ExchangeView view{};
    rag::ExchangeView* d_view = nullptr;  // the same view in device memory (the scan kernel's last CTA reads it)
};
