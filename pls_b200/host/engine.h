// pls_b200/host/engine.h — process-wide handle on libplsb.so for the C++ drop-in layer.
//
// The reference has no device, so nothing in its API names one (include/PLS/pls.h).  The engine is a
// lazily created singleton: one plsb context per GPU listed in PLSB_DEVICES (default "0"), an NCCL
// communicator over them when a row-sharded fit is first needed, and a mutex that keeps the
// reference's "const methods are re-entrant" property (threading).
#pragma once
#include <atomic>
#include <cstdint>
#include <functional>
#include <mutex>
#include <vector>

#include "../../include/plsb.h"

namespace PLS {
namespace detail {

// print plsb_last_error() and exit(1): the reference's only failure idiom (pls.cpp:55-58)
[[noreturn]] void die(const char *where, int rc);
inline void ok(int rc, const char *where) {
    if (rc != PLSB_OK) die(where, rc);
}

struct Engine {
    std::vector<int> devs;
    std::vector<plsb_ctx_t> ctx;
    bool comm_ready = false;
    std::atomic<bool> ready{false};   // contexts exist (checked without the lock: worker threads call in)
    int64_t shard_min_rows = 200000;  // fits with at least this many rows are row-sharded when > 1 GPU
    std::recursive_mutex mu;

    static Engine &get();
    void configure(const std::vector<int> &devices);  // drops existing contexts
    void ensure_contexts();
    void ensure_comm();
    size_t ndev() { ensure_contexts(); return ctx.size(); }
    plsb_ctx_t primary() { ensure_contexts(); return ctx[0]; }
    // run fn(device_slot) on one host thread per GPU and join
    void parallel(const std::function<void(size_t)> &fn);
    ~Engine();
};

// Device-resident copies of one (X, Y) pair: a full replica per GPU (cross-validation, single-GPU
// fits) and / or one row shard per GPU (large fits).  Shared by copies of a Model.
struct DeviceMirror {
    int64_t N = 0, K = 0, M = 0;
    std::vector<plsb_dataset_t> full;    // [device slot] or null
    std::vector<plsb_dataset_t> shard;   // [device slot] or null
    std::vector<int64_t> shard_row0;     // ndev + 1 offsets
    plsb_dataset_t full_on(size_t slot, const double *X, int64_t ldx, const double *Y, int64_t ldy);
    void make_shards(const double *X, int64_t ldx, const double *Y, int64_t ldy);
    void drop_shards();
    ~DeviceMirror();
};

}  // namespace detail
}  // namespace PLS
