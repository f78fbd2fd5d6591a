// pls_b200/host/engine.cpp — see engine.h.
#include "engine.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <thread>

#include <PLS/pls.h>

namespace PLS {
namespace detail {

void die(const char *where, int rc) {
    std::cerr << "Error: " << where << " failed (" << rc << "): " << plsb_last_error() << std::endl;
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
    for (plsb_ctx_t c : ctx) plsb_ctx_destroy(c);
    ctx.clear();
    ready.store(false, std::memory_order_release);
    comm_ready = false;
    devs = devices;
}

void Engine::parallel(const std::function<void(size_t)> &fn) {
    const size_t n = ndev();
    if (n == 1) { fn(0); return; }
    std::vector<std::thread> th;
    for (size_t i = 0; i < n; ++i) th.emplace_back(fn, i);
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
    parallel([&](size_t i) { ok(plsb_comm_init(ctx[i], (int)i, n, id), "plsb_comm_init"); });
    comm_ready = true;
}

Engine::~Engine() {
    // contexts are left to process teardown: static destruction may run after the CUDA runtime's own
}

plsb_dataset_t DeviceMirror::full_on(size_t slot, const double *X, int64_t ldx, const double *Y, int64_t ldy) {
    Engine &e = Engine::get();
    if (full.size() < e.ndev()) full.resize(e.ndev(), nullptr);
    if (!full[slot])
        ok(plsb_dataset_create(e.ctx[slot], X, ldx, N, K, Y, ldy, M, &full[slot]), "plsb_dataset_create");
    return full[slot];
}

void DeviceMirror::make_shards(const double *X, int64_t ldx, const double *Y, int64_t ldy) {
    Engine &e = Engine::get();
    const size_t n = e.ndev();
    drop_shards();
    shard.assign(n, nullptr);
    shard_row0.assign(n + 1, 0);
    for (size_t i = 0; i <= n; ++i) shard_row0[i] = (int64_t)i * N / (int64_t)n;
    e.parallel([&](size_t i) {
        const int64_t r0 = shard_row0[i], rows = shard_row0[i + 1] - r0;
        ok(plsb_dataset_create(e.ctx[i], X + r0, ldx, rows, K, Y + r0, ldy, M, &shard[i]), "plsb_dataset_create(shard)");
        ok(plsb_dataset_set_sharded(shard[i], 1), "plsb_dataset_set_sharded");
    });
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
