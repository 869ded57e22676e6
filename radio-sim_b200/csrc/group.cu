// group.cu -- rmb_group_*: several GPUs behind ONE object, for a caller that lives in one process and thinks in
// whole node arrays and whole packet lists -- the reference's Simulator (one JVM, one RadioMedium,
// Simulator.java:155-165).  Built purely on the public per-handle C-ABI (rmb_create with rx_base / n_shards,
// rmb_peer_*): one handle per device, receivers in contiguous node-index blocks, the tick's tx records exchanged by
// the handles' own kernels over NVLink peer memory, one host worker thread per device so that the shards' calls --
// and their device-to-host copies -- run side by side.  north_star: "receivers are partitioned across the 8 B200s of
// one box ... the compacted delivery lists are gathered back to the host".
#include "../../include/radiomedium_b200.h"
#include <cuda_runtime.h>
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

struct rmb_group {
    int world = 0;
    int n = 0;                                   // nodes in total
    double inv_block = 0.0;                      // world / n
    std::vector<rmb_handle *> h;
    std::vector<int> lo, hi;                     // shard r owns node indices [lo[r], hi[r])
    std::vector<int> device;
    int max_slots = 0;                           // capacity of the peer buffers (records per shard and tick)
    bool peers_open = false;
    std::string err;
    std::mutex mu;                               // one caller at a time (RadioMedium.transmit comes from any reader thread)

    // worker threads: run `job(r)` on every shard, wait for all
    std::vector<std::thread> workers;
    std::mutex wmu; std::condition_variable wcv, dcv;
    std::function<int(int)> job; std::atomic<unsigned long long> generation{0}; std::atomic<int> remaining{0}; bool quit = false;
    std::vector<int> rc;

    // per tick scratch
    std::vector<std::vector<int32_t>> l_src, l_chan; std::vector<std::vector<double>> l_power;
    std::vector<std::vector<int32_t>> l_packet;  // slot -> original packet index

    // stochastic reception on several devices (tick_with_draws): per-shard device scratch for the draw counts of the
    // shard [positions], of all shards [world][positions] and the packets' positions in arrival order; host copies
    int model = 0;
    std::vector<int32_t *> d_counts, d_all, d_order; std::vector<size_t> scratch_positions;
    std::vector<int32_t> h_counts, h_order;
    std::vector<std::string> shard_err;          // a failure of the group's own CUDA calls on shard r
};

namespace {

thread_local std::string g_group_error;

int owner_of(const rmb_group *g, int node)
{
    // blocks are [r n / G, (r + 1) n / G): a multiplication by the reciprocal lands on the owner or next to it (this
    // runs once per packet on the calling thread: an integer division per packet was 0.1 ms of a 0.6 ms tick)
    int r = std::min(g->world - 1, (int)((double)node * g->inv_block));
    while (r > 0 && node < g->lo[r]) --r;
    while (r + 1 < g->world && node >= g->hi[r]) ++r;
    return r;
}

// Ticks come in quick succession, and a condition-variable wake-up costs tens of microseconds on each side of a
// 0.4 ms tick: a worker that has just finished a job spins on the generation counter for a short while (kSpinUs)
// before it goes to sleep, and the caller spins on the completion count the same way.
constexpr int kSpinUs = 300;

void worker_main(rmb_group *g, int r)
{
    unsigned long long seen = 0;
    for (;;) {
        const auto t0 = std::chrono::steady_clock::now();
        while (g->generation.load(std::memory_order_acquire) == seen &&
               std::chrono::steady_clock::now() - t0 < std::chrono::microseconds(kSpinUs)) { }
        std::function<int(int)> job;
        {
            std::unique_lock<std::mutex> lk(g->wmu);
            g->wcv.wait(lk, [&] { return g->quit || g->generation.load() != seen; });
            if (g->quit) return;
            seen = g->generation.load();
            job = g->job;
        }
        g->rc[r] = job(r);
        if (g->remaining.fetch_sub(1, std::memory_order_acq_rel) == 1) {
            std::lock_guard<std::mutex> lk(g->wmu);
            g->dcv.notify_all();
        }
    }
}

// run job(r) for every shard on its worker thread; first non-zero status (and that handle's text) wins
int run_all(rmb_group *g, std::function<int(int)> job)
{
    if (g->world == 1) { const int rc = job(0); if (rc) g->err = rmb_last_error(g->h[0]); return rc; }
    {
        std::lock_guard<std::mutex> lk(g->wmu);
        g->job = std::move(job); g->remaining.store(g->world);
        std::fill(g->rc.begin(), g->rc.end(), 0);
        g->generation.fetch_add(1, std::memory_order_release);
    }
    g->wcv.notify_all();
    const auto t0 = std::chrono::steady_clock::now();
    while (g->remaining.load(std::memory_order_acquire) != 0 &&
           std::chrono::steady_clock::now() - t0 < std::chrono::milliseconds(20)) { }
    if (g->remaining.load(std::memory_order_acquire) != 0) {
        std::unique_lock<std::mutex> lk(g->wmu);
        g->dcv.wait(lk, [&] { return g->remaining.load() == 0; });
    }
    for (int r = 0; r < g->world; ++r)
        if (g->rc[r]) {
            g->err = std::string("shard ") + std::to_string(r) + ": " + (g->shard_err[r].empty() ? rmb_last_error(g->h[r]) : g->shard_err[r].c_str());
            g->shard_err[r].clear();
            return g->rc[r];
        }
    return RMB_OK;
}

int cuda_failed(rmb_group *g, int r, cudaError_t e, const char *what)
{
    g->shard_err[r] = std::string(what) + ": " + cudaGetErrorString(e);
    return RMB_ERR_CUDA;
}
#define GRP_CK(g, r, call, what) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return cuda_failed(g, r, e_, what); } while (0)

// device scratch of shard r for a tick of `positions` gathered record positions (on the shard's worker thread)
int ensure_draw_scratch(rmb_group *g, int r, size_t positions)
{
    GRP_CK(g, r, cudaSetDevice(g->device[r]), "cudaSetDevice");
    if (positions <= g->scratch_positions[r]) return RMB_OK;
    cudaFree(g->d_counts[r]); cudaFree(g->d_all[r]); cudaFree(g->d_order[r]);
    g->d_counts[r] = g->d_all[r] = g->d_order[r] = nullptr; g->scratch_positions[r] = 0;
    const size_t cap = positions + positions / 2 + 64;
    GRP_CK(g, r, cudaMalloc((void **)&g->d_counts[r], cap * sizeof(int32_t)), "draw counts");
    GRP_CK(g, r, cudaMalloc((void **)&g->d_all[r], cap * (size_t)g->world * sizeof(int32_t)), "draw counts of all shards");
    GRP_CK(g, r, cudaMalloc((void **)&g->d_order[r], cap * sizeof(int32_t)), "arrival order");
    g->scratch_positions[r] = cap;
    return RMB_OK;
}

// Stochastic reception with the receivers on several devices.  The reference draws from ONE java.util.Random: per
// packet, in arrival order, the tx draw and then one draw per receiver that needs it, in node order
// (UDGMRadioMedium.java:85-92,104-111).  Every shard evaluates the tick up to its draw flags and counts its draws per
// packet; the counts of all shards meet on the host; every shard then walks the one uniform stream knowing where its
// receivers' draws sit (rmb_tick_draws_begin / _end in the header).  Two rounds over the worker threads.
int tick_with_draws(rmb_group *g, int n_tx, int slots, const double *power_dbm, const int32_t *channel,
                    const double *uniforms, size_t n_uniforms, rmb_result *results)
{
    const int world = g->world;
    const size_t positions = (size_t)world * (size_t)slots;
    g->h_counts.assign((size_t)world * positions, 0);
    g->h_order.assign((size_t)std::max(n_tx, 1), 0);
    for (int r = 0; r < world; ++r)
        for (size_t i = 0; i < g->l_packet[r].size(); ++i) g->h_order[g->l_packet[r][i]] = (int32_t)((size_t)r * slots + i);
    int rc = run_all(g, [=](int r) -> int {
        int rc = rmb_peer_push(g->h[r], slots, g->l_src[r].data(), power_dbm ? g->l_power[r].data() : nullptr,
                               channel ? g->l_chan[r].data() : nullptr);
        if (rc) return rc;
        if ((rc = ensure_draw_scratch(g, r, positions))) return rc;
        if ((rc = rmb_tick_peers_draws_begin(g->h[r], g->d_counts[r]))) return rc;
        void *st = nullptr;
        if ((rc = rmb_get_stream(g->h[r], &st))) return rc;
        GRP_CK(g, r, cudaMemcpyAsync(g->h_counts.data() + (size_t)r * positions, g->d_counts[r], positions * sizeof(int32_t),
                                     cudaMemcpyDeviceToHost, (cudaStream_t)st), "draw counts to the host");
        GRP_CK(g, r, cudaStreamSynchronize((cudaStream_t)st), "draw counts to the host");
        return RMB_OK;
    });
    if (rc) return rc;
    return run_all(g, [=](int r) -> int {
        void *st = nullptr;
        int rc = rmb_get_stream(g->h[r], &st);
        if (rc) return rc;
        GRP_CK(g, r, cudaSetDevice(g->device[r]), "cudaSetDevice");
        GRP_CK(g, r, cudaMemcpyAsync(g->d_all[r], g->h_counts.data(), (size_t)world * positions * sizeof(int32_t),
                                     cudaMemcpyHostToDevice, (cudaStream_t)st), "draw counts of all shards");
        GRP_CK(g, r, cudaMemcpyAsync(g->d_order[r], g->h_order.data(), (size_t)std::max(n_tx, 1) * sizeof(int32_t),
                                     cudaMemcpyHostToDevice, (cudaStream_t)st), "arrival order");
        return rmb_tick_draws_end(g->h[r], g->d_all[r], world, r, g->d_order[r], n_tx, uniforms, n_uniforms, &results[r]);
    });
}

int open_peers(rmb_group *g, int max_slots)
{
    // handles of one process: the blobs carry plain device pointers, peer access is enabled between the devices
    std::vector<char> blobs((size_t)g->world * RMB_PEER_HANDLE_BYTES);
    for (int r = 0; r < g->world; ++r) {
        const int rc = rmb_peer_export(g->h[r], g->world, r, max_slots, blobs.data() + (size_t)r * RMB_PEER_HANDLE_BYTES);
        if (rc) { g->err = rmb_last_error(g->h[r]); return rc; }
    }
    for (int r = 0; r < g->world; ++r) {
        const int rc = rmb_peer_open(g->h[r], blobs.data());
        if (rc) { g->err = rmb_last_error(g->h[r]); return rc; }
    }
    g->max_slots = max_slots; g->peers_open = true;
    return RMB_OK;
}

}  // namespace

extern "C" {

const char *rmb_group_last_error(const rmb_group *g) { return g ? g->err.c_str() : g_group_error.c_str(); }

int rmb_group_create(const rmb_config *cfg, int32_t n_devices, const int32_t *devices, int32_t max_tx_per_tick, rmb_group **out)
{
    if (!cfg || !out || n_devices < 1 || n_devices > 16) { g_group_error = "rmb_group_create: bad arguments (1..16 devices)"; return RMB_ERR_INVALID; }
    *out = nullptr;
    rmb_group *g = new rmb_group();
    g->world = n_devices;
    g->h.assign(n_devices, nullptr); g->lo.assign(n_devices, 0); g->hi.assign(n_devices, 0); g->device.resize(n_devices);
    g->rc.assign(n_devices, 0); g->shard_err.assign(n_devices, std::string());
    g->model = cfg->model;
    g->d_counts.assign(n_devices, nullptr); g->d_all.assign(n_devices, nullptr); g->d_order.assign(n_devices, nullptr);
    g->scratch_positions.assign(n_devices, 0);
    g->l_src.resize(n_devices); g->l_chan.resize(n_devices); g->l_power.resize(n_devices); g->l_packet.resize(n_devices);
    for (int r = 0; r < n_devices; ++r) {
        rmb_config c = *cfg;
        g->device[r] = devices ? devices[r] : r;
        c.device = g->device[r]; c.rx_base = 0; c.n_shards = n_devices;          // rx_base follows with the node table
        const int rc = rmb_create(&c, &g->h[r]);
        if (rc) {
            g_group_error = std::string("device ") + std::to_string(c.device) + ": " + rmb_last_error(nullptr);
            for (int q = 0; q < r; ++q) rmb_destroy(g->h[q]);
            delete g;
            return rc;
        }
    }
    g->max_slots = std::max(1, max_tx_per_tick);       // worst case: every packet of a tick from one shard
    if (n_devices > 1)
        for (int r = 0; r < n_devices; ++r) g->workers.emplace_back(worker_main, g, r);
    *out = g;
    return RMB_OK;
}

int rmb_group_destroy(rmb_group *g)
{
    if (!g) return RMB_ERR_INVALID;
    {
        std::lock_guard<std::mutex> lk(g->wmu);
        g->quit = true;
    }
    g->wcv.notify_all();
    for (auto &t : g->workers) t.join();
    for (int r = 0; r < g->world; ++r)
        if (g->d_counts[r] || g->d_all[r] || g->d_order[r]) {
            cudaSetDevice(g->device[r]);
            cudaFree(g->d_counts[r]); cudaFree(g->d_all[r]); cudaFree(g->d_order[r]);
        }
    for (auto *h : g->h) if (h) rmb_destroy(h);
    delete g;
    return RMB_OK;
}

int32_t rmb_group_size(const rmb_group *g) { return g ? g->world : 0; }

int rmb_group_shard(const rmb_group *g, int32_t r, rmb_handle **handle, int32_t *first_node, int32_t *n_nodes)
{
    if (!g || r < 0 || r >= g->world) return RMB_ERR_INVALID;
    if (handle) *handle = g->h[r];
    if (first_node) *first_node = g->lo[r];
    if (n_nodes) *n_nodes = g->hi[r] - g->lo[r];
    return RMB_OK;
}

int rmb_group_set_nodes(rmb_group *g, int32_t n, const double *x, const double *y, const double *z,
                        const double *txpower_dbm, const int32_t *channel, const uint8_t *enabled,
                        const double *rxprob, const double *txprob, const int32_t *id_int)
{
    if (!g) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(g->mu);
    if (n < 0 || (n > 0 && (!x || !y))) { g->err = "rmb_group_set_nodes: n < 0 or missing x/y"; return RMB_ERR_INVALID; }
    // the blocks move with n: every handle is told its new first node index before it takes its part of the table
    std::vector<int> lo(g->world), hi(g->world);
    for (int r = 0; r < g->world; ++r) { lo[r] = (int)((long long)r * n / g->world); hi[r] = (int)((long long)(r + 1) * n / g->world); }
    g->n = n; g->lo = lo; g->hi = hi; g->inv_block = n > 0 ? (double)g->world / (double)n : 0.0;
    std::vector<int32_t> ids;
    if (!id_int) { ids.resize((size_t)std::max(n, 1)); for (int i = 0; i < n; ++i) ids[i] = i + 1; id_int = ids.data(); }   // Node.java:52-58
    auto off = [](const auto *p, int o) { return p ? p + o : p; };
    return run_all(g, [=](int r) -> int {
        const int b = g->lo[r], m = g->hi[r] - g->lo[r];
        int rc = rmb_set_shard_base(g->h[r], b);
        if (rc) return rc;
        return rmb_set_nodes(g->h[r], m, off(x, b), off(y, b), off(z, b), off(txpower_dbm, b), off(channel, b), off(enabled, b),
                             off(rxprob, b), off(txprob, b), off(id_int, b));
    });
}

int rmb_group_update_positions(rmb_group *g, int32_t count, const int32_t *idx, const double *x, const double *y, const double *z)
{
    if (!g) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(g->mu);
    if (count < 0 || (count > 0 && (!x || !y))) { g->err = "rmb_group_update_positions: bad arguments"; return RMB_ERR_INVALID; }
    if (!idx) {
        if (count != g->n) { g->err = "dense position update needs count == n"; return RMB_ERR_INVALID; }
        auto off = [](const double *p, int o) { return p ? p + o : p; };
        return run_all(g, [=](int r) { return rmb_update_positions(g->h[r], g->hi[r] - g->lo[r], nullptr, off(x, g->lo[r]), off(y, g->lo[r]), off(z, g->lo[r])); });
    }
    std::vector<std::vector<int32_t>> li(g->world); std::vector<std::vector<double>> lx(g->world), ly(g->world), lz(g->world);
    for (int i = 0; i < count; ++i) {
        if (idx[i] < 0 || idx[i] >= g->n) { g->err = "rmb_group_update_positions: node index out of range"; return RMB_ERR_INVALID; }
        const int r = owner_of(g, idx[i]);
        li[r].push_back(idx[i] - g->lo[r]); lx[r].push_back(x[i]); ly[r].push_back(y[i]); if (z) lz[r].push_back(z[i]);
    }
    return run_all(g, [&](int r) {
        if (li[r].empty()) return (int)RMB_OK;
        return rmb_update_positions(g->h[r], (int)li[r].size(), li[r].data(), lx[r].data(), ly[r].data(), z ? lz[r].data() : nullptr);
    });
}

int rmb_group_update_radio(rmb_group *g, int32_t count, const int32_t *idx, const double *txpower_dbm, const int32_t *channel,
                           const uint8_t *enabled, const double *rxprob, const double *txprob)
{
    if (!g) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(g->mu);
    if (count < 0 || (count > 0 && !idx)) { g->err = "rmb_group_update_radio: sparse updates only (idx required)"; return RMB_ERR_INVALID; }
    struct Part { std::vector<int32_t> i, ch; std::vector<double> pw, rx, tx; std::vector<uint8_t> en; };
    std::vector<Part> parts(g->world);
    for (int k = 0; k < count; ++k) {
        if (idx[k] < 0 || idx[k] >= g->n) { g->err = "rmb_group_update_radio: node index out of range"; return RMB_ERR_INVALID; }
        Part &p = parts[owner_of(g, idx[k])];
        p.i.push_back(idx[k] - g->lo[owner_of(g, idx[k])]);
        if (txpower_dbm) p.pw.push_back(txpower_dbm[k]);
        if (channel) p.ch.push_back(channel[k]);
        if (enabled) p.en.push_back(enabled[k]);
        if (rxprob) p.rx.push_back(rxprob[k]);
        if (txprob) p.tx.push_back(txprob[k]);
    }
    return run_all(g, [&](int r) {
        Part &p = parts[r];
        if (p.i.empty()) return (int)RMB_OK;
        return rmb_update_radio(g->h[r], (int)p.i.size(), p.i.data(), txpower_dbm ? p.pw.data() : nullptr, channel ? p.ch.data() : nullptr,
                                enabled ? p.en.data() : nullptr, rxprob ? p.rx.data() : nullptr, txprob ? p.tx.data() : nullptr);
    });
}

int rmb_group_tick(rmb_group *g, int32_t n_tx, const int32_t *src, const double *power_dbm, const int32_t *channel,
                   const double *uniforms, size_t n_uniforms, rmb_result *results, int32_t *slots_out, int32_t *packet_of)
{
    if (!g) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(g->mu);
    if (n_tx < 0 || (n_tx > 0 && !src) || !results) { g->err = "rmb_group_tick: bad arguments"; return RMB_ERR_INVALID; }
    if (g->world == 1) {
        if (slots_out) *slots_out = n_tx;
        if (packet_of) for (int t = 0; t < n_tx; ++t) packet_of[t] = t;
        const int rc = rmb_tick(g->h[0], n_tx, src, power_dbm, channel, uniforms, n_uniforms, &results[0]);
        if (rc) g->err = rmb_last_error(g->h[0]);
        return rc;
    }
    // every shard contributes the packets whose source it owns, padded to a common slot count; a record's arrival
    // index on the devices is rank * slots + slot (arrival order inside a shard is the tick's own)
    for (int r = 0; r < g->world; ++r) {
        g->l_src[r].clear(); g->l_chan[r].clear(); g->l_power[r].clear(); g->l_packet[r].clear();
        const size_t expect = (size_t)n_tx / g->world * 2 + 64;
        g->l_src[r].reserve(expect); g->l_packet[r].reserve(expect);
        if (power_dbm) g->l_power[r].reserve(expect);
        if (channel) g->l_chan[r].reserve(expect);
    }
    for (int t = 0; t < n_tx; ++t) {
        if (src[t] < 0 || src[t] >= g->n) { g->err = "rmb_group_tick: source node index out of range"; return RMB_ERR_INVALID; }
        const int r = owner_of(g, src[t]);
        g->l_src[r].push_back(src[t]); g->l_packet[r].push_back(t);
        if (power_dbm) g->l_power[r].push_back(power_dbm[t]);
        if (channel) g->l_chan[r].push_back(channel[t]);
    }
    int slots = 1;
    for (int r = 0; r < g->world; ++r) slots = std::max(slots, (int)g->l_src[r].size());
    if (!g->peers_open || slots > g->max_slots) {
        if (g->peers_open) { g->err = "rmb_group_tick: more packets from one shard than max_tx_per_tick"; return RMB_ERR_CAPACITY; }
        const int rc = open_peers(g, std::max(slots, g->max_slots));
        if (rc) return rc;
    }
    for (int r = 0; r < g->world; ++r) {
        g->l_src[r].resize(slots, -1);                                          // -1 = padding
        if (power_dbm) g->l_power[r].resize(slots, 0.0);
        if (channel) g->l_chan[r].resize(slots, 0);
    }
    if (slots_out) *slots_out = slots;
    if (packet_of) {
        std::fill(packet_of, packet_of + (size_t)g->world * slots, -1);
        for (int r = 0; r < g->world; ++r)
            for (size_t i = 0; i < g->l_packet[r].size(); ++i) packet_of[(size_t)r * slots + i] = g->l_packet[r][i];
    }
    // one shard that may draw puts every shard on the draws path (a packet's source may sit on another shard)
    bool draws = false;
    if (g->model == RMB_MODEL_SINR || g->model == RMB_MODEL_UDGM)
        for (int r = 0; r < g->world; ++r) draws = draws || rmb_may_draw(g->h[r]) != 0;
    if (draws) return tick_with_draws(g, n_tx, slots, power_dbm, channel, uniforms, n_uniforms, results);
    // the shards' calls run side by side on the worker threads: the exchange rides in each tick's first kernel
    return run_all(g, [=](int r) {
        return rmb_peer_tick(g->h[r], slots, g->l_src[r].data(), power_dbm ? g->l_power[r].data() : nullptr,
                             channel ? g->l_chan[r].data() : nullptr, &results[r]);
    });
}

int rmb_group_node_info_f64(rmb_group *g, double *rssi_dbm, uint8_t *receiving)
{
    if (!g) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(g->mu);
    return run_all(g, [=](int r) { return rmb_node_info_f64(g->h[r], rssi_dbm ? rssi_dbm + g->lo[r] : nullptr, receiving ? receiving + g->lo[r] : nullptr); });
}

}  // extern "C"
