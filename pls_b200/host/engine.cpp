// pls_b200/host/engine.cpp — see engine.h.
#include "engine.h"

#include <algorithm>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <thread>

#include <PLS/pls.h>

namespace PLS {
namespace detail {

void die(const char *where, int rc, const char *msg) {
    std::cerr << "Error: " << where << " failed (" << rc << "): " << (msg ? msg : plsb_last_error()) << std::endl;
    exit(1);
}

Engine &Engine::get() {
    static Engine e;
    return e;
}

static std::vector<int> parse_devices(const char *s) {
    std::vector<int> out;
    if (!s || !*s) return out;
    std::string tok;
    for (const char *p = s;; ++p) {
        if (*p == ',' || *p == '\0') {
            if (!tok.empty()) out.push_back(atoi(tok.c_str()));
            tok.clear();
            if (*p == '\0') break;
        } else if (*p != ' ') {
            tok.push_back(*p);
        }
    }
    return out;
}

void Engine::ensure_contexts() {
    if (ready.load(std::memory_order_acquire)) return;
    std::lock_guard<std::recursive_mutex> lk(mu);
    if (ready.load(std::memory_order_acquire)) return;
    if (devs.empty()) devs = parse_devices(getenv("PLSB_DEVICES"));
    if (devs.empty()) devs.push_back(getenv("PLSB_DEVICE") ? atoi(getenv("PLSB_DEVICE")) : 0);
    if (const char *m = getenv("PLSB_SHARD_MIN_ROWS")) shard_min_rows = atoll(m);
    for (int d : devs) {
        plsb_ctx_t c = nullptr;
        ok(plsb_ctx_create(d, &c), "plsb_ctx_create");
        ctx.push_back(c);
    }
    ready.store(true, std::memory_order_release);
}

void Engine::configure(const std::vector<int> &devices) {
    std::lock_guard<std::recursive_mutex> lk(mu);
    // Data sets of live Models keep their contexts alive (reference-counted in libplsb.so); the engine only lets go
    // of its own handles.  The generation stamp makes every DeviceMirror carry its data over to the new contexts the
    // next time it is used (DeviceMirror::rehome).
    for (plsb_ctx_t c : ctx) plsb_ctx_destroy(c);
    ctx.clear();
    ready.store(false, std::memory_order_release);
    comm_ready = false;
    devs = devices;
    generation.fetch_add(1, std::memory_order_acq_rel);
}

void Engine::parallel(const std::function<int(size_t)> &fn, const char *where) {
    const size_t n = ndev();
    std::vector<int> rc(n, PLSB_OK);
    std::vector<std::string> msg(n);
    auto run = [&](size_t i) {
        rc[i] = fn(i);
        if (rc[i] != PLSB_OK) msg[i] = plsb_last_error();   // thread-local in libplsb.so: read it on the worker
    };
    if (n == 1) {
        run(0);
    } else {
        std::vector<std::thread> th;
        for (size_t i = 0; i < n; ++i) th.emplace_back(run, i);
        for (auto &t : th) t.join();
    }
    for (size_t i = 0; i < n; ++i)
        if (rc[i] != PLSB_OK) die(where, rc[i], msg[i].c_str());
}

void Engine::widen(const double *src, std::complex<double> *dst, size_t n) {
    size_t nt = n >= (size_t(1) << 20) ? std::min<size_t>(8, std::max(1u, std::thread::hardware_concurrency())) : 1;
    auto run = [&](size_t t) {
        const size_t a = n / nt * t, b = t + 1 == nt ? n : n / nt * (t + 1);
        for (size_t i = a; i < b; ++i) dst[i] = std::complex<double>(src[i], 0.0);
    };
    if (nt == 1) { run(0); return; }
    std::vector<std::thread> th;
    for (size_t t = 1; t < nt; ++t) th.emplace_back(run, t);
    run(0);
    for (auto &t : th) t.join();
}

void Engine::ensure_comm() {
    std::lock_guard<std::recursive_mutex> lk(mu);
    ensure_contexts();
    if (comm_ready || ctx.size() < 2) return;
    char id[128];
    ok(plsb_comm_unique_id(id), "plsb_comm_unique_id");
    const int n = (int)ctx.size();
    // ncclCommInitRank blocks until every rank has joined: one host thread per GPU
    parallel([&](size_t i) { return plsb_comm_init(ctx[i], (int)i, n, id); }, "plsb_comm_init");
    comm_ready = true;
}

Engine::~Engine() {
    // contexts are left to process teardown: static destruction may run after the CUDA runtime's own
}

// ---- DeviceMirror ---------------------------------------------------------------------------------------------------
void DeviceMirror::rehome() {
    Engine &e = Engine::get();
    const size_t n = e.ndev();
    const uint64_t g = e.generation.load(std::memory_order_acquire);
    if (gen == g) {
        if (full.size() < n) full.resize(n, nullptr);
        return;
    }
    // the slots refer to contexts of an earlier device configuration: gather the data (one full replica, else the
    // row shards in order) into a replica on the new primary device, then let the old copies go
    std::vector<plsb_dataset_t> src;
    for (plsb_dataset_t d : full)
        if (d && src.empty()) src.push_back(d);
    if (src.empty())
        for (plsb_dataset_t d : shard)
            if (d) src.push_back(d);
    plsb_dataset_t fresh = nullptr;
    if (!src.empty()) ok(plsb_dataset_concat(e.ctx[0], (int64_t)src.size(), src.data(), &fresh), "plsb_dataset_concat");
    drop_shards();
    for (plsb_dataset_t d : full)
        if (d) plsb_dataset_destroy(d);
    full.assign(n, nullptr);
    full[0] = fresh;
    gen = g;
}

void DeviceMirror::load_full(const double *X, int64_t ldx, const double *Y, int64_t ldy) {
    Engine &e = Engine::get();
    rehome();
    if (full[0]) { plsb_dataset_destroy(full[0]); full[0] = nullptr; }
    ok(plsb_dataset_create(e.ctx[0], X, ldx, N, K, Y, ldy, M, &full[0]), "plsb_dataset_create");
}

void DeviceMirror::load_shards(const double *X, int64_t ldx, const double *Y, int64_t ldy) {
    Engine &e = Engine::get();
    rehome();
    const size_t n = e.ndev();
    drop_shards();
    shard.assign(n, nullptr);
    shard_row0.assign(n + 1, 0);
    for (size_t i = 0; i <= n; ++i) shard_row0[i] = (int64_t)i * N / (int64_t)n;
    e.parallel([&](size_t i) {
        const int64_t r0 = shard_row0[i], rows = shard_row0[i + 1] - r0;
        int rc = plsb_dataset_create(e.ctx[i], X + r0, ldx, rows, K, Y + r0, ldy, M, &shard[i]);
        if (rc == PLSB_OK) rc = plsb_dataset_set_sharded(shard[i], 1);
        return rc;
    }, "plsb_dataset_create(shard)");
}

void DeviceMirror::prepare() {
    rehome();
    source.clear();
    for (plsb_dataset_t d : full)
        if (d && source.empty()) source.push_back(d);
    if (source.empty())
        for (plsb_dataset_t d : shard)
            if (d) source.push_back(d);
}

int DeviceMirror::ensure_full(size_t slot) {
    Engine &e = Engine::get();
    if (full[slot]) return PLSB_OK;
    if (source.empty()) return PLSB_ERR_ARG;   // a Model built without data (the reference's private constructors)
    return plsb_dataset_concat(e.ctx[slot], (int64_t)source.size(), source.data(), &full[slot]);
}

plsb_dataset_t DeviceMirror::full_on(size_t slot) {
    prepare();
    ok(ensure_full(slot), "plsb_dataset_concat");
    return full[slot];
}

void DeviceMirror::drop_shards() {
    for (plsb_dataset_t d : shard)
        if (d) plsb_dataset_destroy(d);
    shard.clear();
}

DeviceMirror::~DeviceMirror() {
    drop_shards();
    for (plsb_dataset_t d : full)
        if (d) plsb_dataset_destroy(d);
}

}  // namespace detail

void set_devices(const std::vector<int> &devices) { detail::Engine::get().configure(devices); }
std::vector<int> devices() {
    detail::Engine &e = detail::Engine::get();
    e.ensure_contexts();
    return e.devs;
}

}  // namespace PLS
