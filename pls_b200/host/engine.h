// pls_b200/host/engine.h — process-wide handle on libplsb.so for the C++ drop-in layer.
//
// The reference has no device, so nothing in its API names one (include/PLS/pls.h).  The engine is a
// lazily created singleton: one plsb context per GPU listed in PLSB_DEVICES (default "0"), an NCCL
// communicator over them when a row-sharded fit is first needed, and a mutex that keeps the
// reference's "const methods are re-entrant" property (threading).
#pragma once
#include <atomic>
#include <complex>
#include <cstdint>
#include <functional>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/plsb.h"

namespace PLS {
namespace detail {

// print the error and exit(1): the reference's only failure idiom (pls.cpp:55-58).  Called on the thread
// that entered the library, never on a worker (Engine::parallel joins its workers first).
[[noreturn]] void die(const char *where, int rc, const char *msg = nullptr);
inline void ok(int rc, const char *where) {
    if (rc != PLSB_OK) die(where, rc);
}

struct Engine {
    std::vector<int> devs;
    std::vector<plsb_ctx_t> ctx;
    bool comm_ready = false;
    std::atomic<bool> ready{false};   // contexts exist (checked without the lock: worker threads call in)
    // bumped by configure(): device mirrors made under an earlier generation live on contexts the engine has let
    // go of (they stay valid — contexts are reference-counted by their data sets) and are re-homed on next use
    std::atomic<uint64_t> generation{1};
    int64_t shard_min_rows = 200000;  // fits with at least this many rows are row-sharded when > 1 GPU
    std::recursive_mutex mu;

    static Engine &get();
    void configure(const std::vector<int> &devices);  // lets go of the existing contexts
    void ensure_contexts();
    void ensure_comm();
    size_t ndev() { ensure_contexts(); return ctx.size(); }
    plsb_ctx_t primary() { ensure_contexts(); return ctx[0]; }
    // Run fn(device_slot) on one host thread per GPU and join them all.  fn returns a plsb status; a worker never
    // exits the process itself (its siblings may be inside a collective): the first failure, with the message the
    // failing thread saw, is raised on the calling thread after the join.
    void parallel(const std::function<int(size_t)> &fn, const char *where);
    // dst[i] = (src[i], 0) for n elements, on a few host threads (the reference's model matrices are complex-typed,
    // pls.h:251-254; the N x A scores are the one large conversion)
    static void widen(const double *src, std::complex<double> *dst, size_t n);
    ~Engine();
};

// The Model's own copy of its training data (`_X,_Y` in the reference, pls.cpp:344) — kept in HBM only: a full
// replica per GPU (cross-validation, single-GPU fits) and / or one row shard per GPU (large fits).  Replicas on
// further GPUs are made device-to-device from whatever copy exists.  Shared by copies of a Model.
struct DeviceMirror {
    int64_t N = 0, K = 0, M = 0;
    uint64_t gen = 0;                    // engine generation the slots below refer to
    std::vector<plsb_dataset_t> full;    // [device slot] or null
    std::vector<plsb_dataset_t> shard;   // [device slot] or null
    std::vector<int64_t> shard_row0;     // ndev + 1 offsets
    // upload host data (the constructor's fit): one replica on slot 0, or one row shard per GPU
    void load_full(const double *X, int64_t ldx, const double *Y, int64_t ldy);
    void load_shards(const double *X, int64_t ldx, const double *Y, int64_t ldy);
    // the replica on `slot`, made from an existing copy if it is not there yet (call with Engine::mu held).
    // prepare() on the calling thread first: it fixes the copy every missing replica will be made from (the first
    // replica that exists, else the row shards).  Then ensure_full(slot) may run on one worker per slot (returns a
    // status): a worker writes full[slot] only and never looks at its siblings' slots.
    void prepare();
    int ensure_full(size_t slot);
    plsb_dataset_t full_on(size_t slot);
    void drop_shards();
    ~DeviceMirror();

  private:
    std::vector<plsb_dataset_t> source;   // chosen by prepare(); read-only while workers run
    void rehome();   // after Engine::configure: carry the data over to the new contexts
};

}  // namespace detail
}  // namespace PLS
