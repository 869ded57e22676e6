// api.cu -- the C-ABI of include/radiomedium_b200.h: handle, HBM buffers, per-tick pipeline.
//
// No CPU fallback lives here: every evaluation is a CUDA kernel launch; without a device
// rmb_create fails.  One mutex per handle; one stream per handle; the host blocks only where
// it needs a count (list capacity) or the results.
#include "common.cuh"
#include "kernels.h"
#include <algorithm>
#include <cmath>
#include <climits>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <unistd.h>
#include <string>
#include <vector>

using namespace rmb;

namespace {

thread_local std::string g_create_error;

template <class T> struct DevBuf {
    T *p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t need, bool zero = false) {
        if (need <= cap) return cudaSuccess;
        // a quarter of headroom: per-tick sizes (heard pairs, packets) wander by a few per cent from tick to tick, and
        // a re-allocation costs milliseconds
        size_t ncap = std::max(need + need / 4, cap + cap / 2);
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        cudaError_t e = cudaMalloc((void **)&p, std::max<size_t>(ncap, 1) * sizeof(T));
        if (e != cudaSuccess) return e;
        cap = ncap;
        if (zero) return cudaMemset(p, 0, ncap * sizeof(T));
        return cudaSuccess;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

template <class T> struct PinBuf {
    T *p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t need) {
        if (need <= cap) return cudaSuccess;
        size_t ncap = std::max(need + need / 4, cap + cap / 2);      // pinning pages is slow: 17 ms for the lists of a C4 tick
        if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
        cudaError_t e = cudaMallocHost((void **)&p, std::max<size_t>(ncap, 1) * sizeof(T));
        if (e != cudaSuccess) return e;
        cap = ncap;
        return cudaSuccess;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

}  // namespace

struct rmb_handle {
    rmb_config cfg{};
    std::mutex mu;
    std::string err;
    cudaStream_t own_st = nullptr;                 // created by rmb_create; st == own_st unless rmb_set_stream
    cudaStream_t st = nullptr, st2 = nullptr;      // st2: guard-band recheck, overlapped with the pair ordering
    cudaEvent_t ev[6] = {}, ev_fork = nullptr, ev_join = nullptr;
    rmb_timing timing{}; bool timing_pending = false, timing_by_events = true;

    // host mirrors of the radio state (work order + stochastic detection are host-built)
    int n = 0;
    std::vector<int> h_channel; std::vector<uint8_t> h_enabled;
    std::vector<double> h_txpower, h_rxprob, h_txprob;
    bool maybe_draws = false;
    bool planar = true; double planar_z = 0.0;      // all nodes share one z (2-D sweep) and which
    cudaEvent_t ev_h2d = nullptr;                   // "the caller's arrays have been read" (dense position updates)
    // rmb_update_positions_async: the next positions of ALL nodes, uploaded on a copy stream into a second set of
    // arrays while a tick may still be running on the first; swapped in by the next call that settles the handle
    DevBuf<double> x2, y2, z2; cudaStream_t st_copy = nullptr; cudaEvent_t ev_staged = nullptr;
    bool staged = false; int staged_n = 0, staged_planar = 1; double staged_z = 0.0; bool staged_has_z = false;

    // device state
    DevBuf<double> x, y, z, txpower, rxprob, txprob, matrix;
    DevBuf<int> channel, id_int;
    DevBuf<uint8_t> enabled;
    int matrix_n = 0;
    DevBuf<int> perm, gid; DevBuf<float> fx, fy, fz; DevBuf<double> wx, wy, wz, wrxp;
    bool node_info_dirty = false;
    int dense_updates = 0;
    DevBuf<unsigned long long> ro_key, ro_key_alt; DevBuf<int> ro_val, ro_val_alt; DevBuf<char> ro_tmp;
    int n_work = 0;
    Groups groups{};
    DevBuf<DevParams> dp; DevBuf<Geom> geom; DevBuf<unsigned long long> geom_scratch;
    DevBuf<Counters> ctr; PinBuf<Counters> h_ctr;
    DevBuf<float> node_rssi; DevBuf<double> node_rssi64; DevBuf<uint8_t> node_recv;
    DevBuf<int> tx_epoch, hd_cnt; int hd_epoch = 0;      // half duplex (rmb_config.half_duplex): see PrepareAux

    // per tick: tx side
    DevBuf<int> pk_src, pk_channel; DevBuf<double> pk_power; PinBuf<char> pk_stage;
    DevBuf<rmb_tx_record> rec;
    DevBuf<float> txf; DevBuf<float4> stg_f; DevBuf<float> theta, theta_min_g, stg_theta; DevBuf<int> stg_gid; DevBuf<int> arrival, goff, gcount;
    DevBuf<int> blk_cnt, blk_thmin; DevBuf<unsigned> bar; unsigned bar_count = 0, bar2_count = 0; DevBuf<double> recheck_part; DevBuf<int> rx_done;   // fused tx preparation (cooperative launch)
    DevBuf<double> uniforms;

    // slots
    int K = 8;
    DevBuf<unsigned long long> trace;      // debug: per-block time stamps of the sweep (RMB_SWEEP_TRACE=<file>)
    DevBuf<int> slots, cnt, hcnt, hoff, recheck, recheck2, recheck_cnt; DevBuf<float> far_part; SweepSplits splits{1, 1, 0, 0, 512, false}; int forced_splits = 0, taper = 2, fine8 = 4;
    DevBuf<double> sl_prx, sl_lin, sl_sinr, itot; DevBuf<uint8_t> sl_flags;

    // pairs
    DevBuf<unsigned long long> key, key_alt, fb_key, fb_key_alt; DevBuf<unsigned> val, val_alt, fb_val, fb_val_alt;
    DevBuf<double> rec_prx, rec_sinr; DevBuf<uint8_t> rec_flags; DevBuf<int> rec_rank, tx_cnt, first;
    DevBuf<double> pr_p, pr_rssi, pr_sinr; DevBuf<uint8_t> pr_cand;
    DevBuf<int> draw_scan, seg, scan_tmp, blkcnt; DevBuf<long long> tx_base; DevBuf<uint8_t> tx_failed;
    DevBuf<char> cub_tmp;
    DevBuf<int> o_tx, o_rx; DevBuf<float> o_rssi, o_sinr; DevBuf<uint8_t> o_outcome; DevBuf<double> o_rssi64, o_sinr64;
    PinBuf<int> ho_tx, ho_rx, ho_first; DevBuf<int> o_first; PinBuf<float> ho_rssi, ho_sinr; PinBuf<uint8_t> ho_outcome, ho_failed;
    PinBuf<double> ho_rssi64, ho_sinr64;
    DevBuf<char> o_slab; PinBuf<char> ho_slab; bool slab_ready = false;     // finish.cu path: all columns packed, one copy (kernels.h slab_layout)
    // The synchronous calls (rmb_tick, rmb_tick_records, rmb_tick_peers) copy the slab home BEHIND the finish kernel,
    // before the host knows the row count: as many bytes as the previous tick's lists took plus 1/64.  If that
    // covers this tick's lists -- it does whenever consecutive ticks are alike -- the counters and the lists arrive
    // with ONE stream synchronisation instead of two (wait, look at the count, copy, wait again).
    bool want_download = false; size_t slab_pred = 0, slab_copied = 0;

    // multi-GPU exchange over peer memory (rmb_peer_*)
    struct Peer {
        int world = 0, rank = 0, max_slots = 0, slots = 0;
        unsigned epoch = 0;               // ticks pushed so far (flags carry this value)
        unsigned barrier_epoch = 0;       // rmb_peer_barrier_enqueue calls so far (second flag row)
        bool pushed = false;              // a push is waiting for its tick
        bool shares_device = false;       // some peer sits on this handle's device (development on a one-GPU box)
        char *buf = nullptr;              // this handle's gather buffer: flags, then 2 x world x max_slots records
        char *peer[kMaxPeers] = {};       // every rank's buffer as mapped here (peer[rank] == buf)
        bool ipc[kMaxPeers] = {};         // mapped with cudaIpcOpenMemHandle (must be closed)
        unsigned *done = nullptr;         // arrival counter of the push kernel's blocks
        size_t rec_off(unsigned e) const { return kPeerFlagBytes + (size_t)(e & 1u) * (size_t)world * max_slots * sizeof(rmb_tx_record); }
    } peer;
    const unsigned *wait_flags = nullptr; unsigned wait_epoch = 0; int wait_world = 0;   // set around compute() by rmb_tick_peers
    const rmb_tx_record *fused_push = nullptr;   // rmb_peer_exchange_tick_enqueue: the preparation kernel pushes these itself
    // stochastic reception on sharded receivers (rmb_tick_draws_begin / _end): the tick stops after the draw flags,
    // the host side gathers the per-packet draw counts of all shards, the second half walks the global sequence
    int32_t *split_counts = nullptr;      // non-null while compute() runs for rmb_tick_draws_begin
    struct Split { bool active = false; int n_tx = 0; long long m = 0; const rmb_tx_record *rec = nullptr; int launches = 0, sweeps = 0; } split;
    // asynchronous tick (rmb_tick_enqueue / rmb_tick_peers_enqueue): enqueued, counters not looked at yet
    bool async_mode = false, pending = false;
    struct Pending { int n_tx = 0; const rmb_tx_record *rec = nullptr; const unsigned *flags = nullptr; unsigned epoch = 0; int world = 0;
                     int launches = 0, sweeps = 0; bool slab = false; } pend;

    // last tick
    long long last_pairs = 0; int last_ntx = 0;
    rmb_result last{};
    int sweep_variant = 0;
    bool fuse_resolve = true;          // rmb_set_tuning bit 16 turns the fused epilogue off (stand-alone resolve kernel)
    bool fuse_finish = true;           // rmb_set_tuning bit 20: separate recheck / ordering / emit kernels instead of finish.cu
    bool kernel_events = false;        // rmb_set_tuning bit 24: CUDA events BETWEEN the tick's kernels (per-kernel durations by
                                       // event; the launches are then serialised, no programmatic dependent launch)
    bool have_nodes = false;

    int fail(int code, const std::string &m) { err = m; return code; }
    int cuda_fail(cudaError_t e, const char *where) {
        err = std::string(where) + ": " + cudaGetErrorString(e);
        cudaGetLastError();
        return RMB_ERR_CUDA;
    }
};

#define RMB_CK(h, expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return (h)->cuda_fail(e__, #expr); } while (0)
// an enqueued tick (rmb_tick_enqueue) is settled before anything else touches the handle
static int finish_pending(rmb_handle *h);
static int apply_staged(rmb_handle *h);
#define RMB_SETTLE(h) do { int rc__ = finish_pending(h); if (rc__) return rc__; rc__ = apply_staged(h); if (rc__) return rc__; } while (0)

namespace {

NodeArrays node_arrays(rmb_handle *h)
{
    NodeArrays a;
    a.n = h->n; a.x = h->x.p; a.y = h->y.p; a.z = h->z.p; a.txpower = h->txpower.p; a.rxprob = h->rxprob.p;
    a.txprob = h->txprob.p; a.channel = h->channel.p; a.enabled = h->enabled.p; a.id_int = h->id_int.p;
    return a;
}
WorkArrays work_arrays(rmb_handle *h)
{
    WorkArrays w; w.n_work = h->n_work; w.perm = h->perm.p; w.fx = h->fx.p; w.fy = h->fy.p; w.fz = h->fz.p;
    w.x64 = h->wx.p; w.y64 = h->wy.p; w.z64 = h->wz.p; w.rxp = h->wrxp.p;
    w.tx_epoch = h->cfg.half_duplex ? h->tx_epoch.p : nullptr; w.hd_epoch = h->hd_epoch;
    return w;
}
TxTiles tx_tiles(rmb_handle *h)
{
    TxTiles t;
    const size_t q = h->txf.cap / 4;                 // four SoA rows in one allocation, each 256-byte aligned
    t.c = h->txf.p; t.nx = h->txf.p + q; t.ny = h->txf.p + 2 * q; t.nz = h->txf.p + 3 * q;
    t.arrival = h->arrival.p; t.goff = h->goff.p; t.gcount = h->gcount.p;
    return t;
}
SlotArrays slot_arrays(rmb_handle *h)
{
    SlotArrays s; s.K = h->K; s.slots = h->slots.p; s.cnt = h->cnt.p; s.far_part = h->far_part.p; s.splits = h->splits.s1; s.splits2 = h->splits.s2; s.w_taper = h->splits.rb_taper * h->splits.block_rx; s.prx = h->sl_prx.p;
    s.lin = h->sl_lin.p; s.sinr = h->sl_sinr.p; s.flags = h->sl_flags.p; s.hcnt = h->hcnt.p; s.itot = h->itot.p;
    s.recheck = h->recheck.p; s.recheck2 = h->recheck2.p; s.recheck_cnt = h->recheck_cnt.p; s.first = h->first.p;
    return s;
}
PairArrays pair_arrays(rmb_handle *h)
{
    PairArrays p{};
    p.cap = (long long)h->key.cap; p.key = h->key.p; p.key_alt = h->key_alt.p; p.val = h->val.p; p.val_alt = h->val_alt.p;
    p.p = h->pr_p.p; p.rssi = h->pr_rssi.p; p.sinr = h->pr_sinr.p; p.cand = h->pr_cand.p; p.draw_scan = h->draw_scan.p;
    p.seg = h->seg.p; p.tx_base = h->tx_base.p; p.tx_failed = h->tx_failed.p; p.scan_tmp = h->scan_tmp.p;
    p.cub_tmp = h->cub_tmp.p; p.cub_tmp_bytes = h->cub_tmp.cap;
    p.o_tx = h->o_tx.p; p.o_rx = h->o_rx.p; p.o_rssi = h->o_rssi.p; p.o_sinr = h->o_sinr.p; p.o_outcome = h->o_outcome.p;
    p.o_rssi64 = (h->cfg.flags & RMB_FLAG_F64_VALUES) ? h->o_rssi64.p : nullptr;
    p.o_sinr64 = (h->cfg.flags & RMB_FLAG_F64_VALUES) ? h->o_sinr64.p : nullptr;
    int bits = 1;
    while (bits < 31 && (1ll << bits) < (long long)h->cfg.rx_base + h->n + 1) ++bits;
    p.key_shift = bits;
    p.slab = nullptr; p.slab_flags = h->cfg.flags;
    return p;
}

PairRecs pair_recs(rmb_handle *h)
{
    PairRecs r{};
    r.cap = (long long)h->key.cap; r.key = h->key.p; r.prx = h->rec_prx.p; r.sinr = h->rec_sinr.p; r.flags = h->rec_flags.p;
    r.rank = h->rec_rank.p; r.tx_cnt = h->tx_cnt.p;
    int bits = 1;
    while (bits < 31 && (1ll << bits) < (long long)h->cfg.rx_base + h->n + 1) ++bits;
    r.key_shift = bits;
    return r;
}

int bits_for(long long v) { int b = 1; while (b < 62 && (1ll << b) < v + 1) ++b; return b; }

}  // namespace

// ------------------------------------------------------------------------------------------
// work order: enabled receivers grouped by channel (ascending), node order inside a group,
// every group padded to 1024 so that a sweep block never straddles two channels
// ------------------------------------------------------------------------------------------
static int ensure_slots(rmb_handle *h)
{
    const size_t nw = (size_t)h->n_work, nk = nw * (size_t)h->K;
    RMB_CK(h, h->slots.ensure(nk)); RMB_CK(h, h->sl_prx.ensure(nk)); RMB_CK(h, h->sl_lin.ensure(nk));
    RMB_CK(h, h->sl_sinr.ensure(nk)); RMB_CK(h, h->sl_flags.ensure(nk));
    RMB_CK(h, h->cnt.ensure(nw)); RMB_CK(h, h->hcnt.ensure(nw)); RMB_CK(h, h->hoff.ensure(nw + 2));
    RMB_CK(h, h->recheck.ensure(nw)); RMB_CK(h, h->recheck2.ensure(nw)); RMB_CK(h, h->recheck_cnt.ensure(nw)); RMB_CK(h, h->recheck_part.ensure(nw + 1024)); RMB_CK(h, h->rx_done.ensure(2 * (nw / 32 + 1))); RMB_CK(h, h->first.ensure(nw)); RMB_CK(h, h->itot.ensure(nw));
    RMB_CK(h, h->scan_tmp.ensure(nw / 1024 + 8));
    return RMB_OK;
}

static int build_work_order(rmb_handle *h)
{
    const int n = h->n;
    std::vector<int> chans;
    for (int i = 0; i < n; ++i) if (h->h_enabled[i]) chans.push_back(h->h_channel[i]);
    std::sort(chans.begin(), chans.end());
    chans.erase(std::unique(chans.begin(), chans.end()), chans.end());
    if ((int)chans.size() > kMaxGroups)
        return h->fail(RMB_ERR_UNSUPPORTED, "more than 64 distinct receiver channels");
    Groups g{};
    g.ngroups = (int)chans.size();
    std::vector<int> count(g.ngroups, 0);
    std::vector<int> gid(n, -1);
    for (int i = 0; i < n; ++i) if (h->h_enabled[i]) {
        int k = (int)(std::lower_bound(chans.begin(), chans.end(), h->h_channel[i]) - chans.begin());
        gid[i] = k; ++count[k];
    }
    int w = 0;
    for (int k = 0; k < g.ngroups; ++k) {
        g.channel[k] = chans[k]; g.n_rx[k] = count[k]; g.w_begin[k] = w;
        w += (count[k] + 1023) / 1024 * 1024;
    }
    g.w_begin[g.ngroups] = w;
    for (int k = g.ngroups + 1; k <= kMaxGroups; ++k) g.w_begin[k] = w;
    int dense = 0;
    for (int k = 0; k <= kMaxGroups; ++k) { g.dense_begin[k] = dense; if (k < g.ngroups) dense += count[k]; }
    h->groups = g;
    h->n_work = w;
    // perm itself is built on the device (Morton order inside a group): refresh_geometry(h, true)
    const int nn = std::max(n, 1);
    RMB_CK(h, h->perm.ensure(std::max(w, 1))); RMB_CK(h, h->fx.ensure(w)); RMB_CK(h, h->fy.ensure(w)); RMB_CK(h, h->fz.ensure(w));
    RMB_CK(h, h->wx.ensure(std::max(w, 1))); RMB_CK(h, h->wy.ensure(std::max(w, 1))); RMB_CK(h, h->wz.ensure(std::max(w, 1)));
    RMB_CK(h, h->wrxp.ensure(std::max(w, 1)));
    h->node_info_dirty = false;
    RMB_CK(h, h->gid.ensure(nn)); RMB_CK(h, h->ro_key.ensure(nn)); RMB_CK(h, h->ro_key_alt.ensure(nn));
    RMB_CK(h, h->ro_val.ensure(nn)); RMB_CK(h, h->ro_val_alt.ensure(nn)); RMB_CK(h, h->ro_tmp.ensure(reorder_tmp_bytes(nn)));
    if (n > 0) RMB_CK(h, cudaMemcpy(h->gid.p, gid.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice));
    int rc = ensure_slots(h);
    if (rc) return rc;
    // idle nodes report the noise floor (AbstractRadioMedium.java:38,45-52)
    std::vector<float> idle((size_t)std::max(n, 1), (float)h->cfg.noise_floor_dbm);
    std::vector<double> idle64((size_t)std::max(n, 1), h->cfg.noise_floor_dbm);
    RMB_CK(h, h->node_rssi.ensure(std::max(n, 1))); RMB_CK(h, h->node_rssi64.ensure(std::max(n, 1))); RMB_CK(h, h->node_recv.ensure(std::max(n, 1)));
    RMB_CK(h, cudaMemcpy(h->node_rssi.p, idle.data(), idle.size() * sizeof(float), cudaMemcpyHostToDevice));
    RMB_CK(h, cudaMemcpy(h->node_rssi64.p, idle64.data(), idle64.size() * sizeof(double), cudaMemcpyHostToDevice));
    RMB_CK(h, cudaMemset(h->node_recv.p, 0, std::max(n, 1)));
    if (h->cfg.half_duplex) {
        // stamps of earlier ticks never equal a later epoch, so a grown array only has to start from zero
        if ((size_t)nn > h->tx_epoch.cap) { RMB_CK(h, h->tx_epoch.ensure(nn)); RMB_CK(h, cudaMemset(h->tx_epoch.p, 0, h->tx_epoch.cap * sizeof(int))); }
        if (!h->hd_cnt.p) RMB_CK(h, h->hd_cnt.ensure(2 * kMaxGroups, true));
    }
    return RMB_OK;
}

static void update_draw_flag(rmb_handle *h)
{
    bool d = false;
    const int m = h->cfg.model;
    if (m == RMB_MODEL_N2N) d = true;
    if (m == RMB_MODEL_UDGM && h->cfg.success_ratio_rx < 1.0) d = true;
    if (m == RMB_MODEL_UDGM || m == RMB_MODEL_SINR || m == RMB_MODEL_N2N)
        for (int i = 0; i < h->n && !d; ++i) if (h->h_rxprob[i] < 1.0 || h->h_txprob[i] < 1.0) d = true;
    h->maybe_draws = d;
}

// bounding box -> centre / planarity, then the fp32 work-order copy.
// planar_known: 1 / 0 = the caller knows whether all z are equal (it saw them on the host), so nothing is read back
// and nothing waits -- the kernels stay queued behind the upload; -1 = read the device's verdict (one round trip).
static int refresh_geometry(rmb_handle *h, bool reorder, int planar_known = -1)
{
    NodeArrays nd = node_arrays(h);
    RMB_CK(h, launch_geom(nd, h->geom_scratch.p, h->geom.p, h->st));
    Geom g{};
    if (planar_known < 0) RMB_CK(h, cudaMemcpyAsync(&g, h->geom.p, sizeof(Geom), cudaMemcpyDeviceToHost, h->st));
    if (reorder) {
        // the last tick's per-receiver results are in the OLD work order: turn them into per-node values before it goes
        if (h->node_info_dirty && h->n_work > 0 && h->node_rssi.p) {
            RMB_CK(h, launch_node_info(work_arrays(h), slot_arrays(h), h->node_rssi.p, h->node_rssi64.p, h->node_recv.p, h->st));
            h->node_info_dirty = false;
        }
        ReorderScratch sc{h->ro_key.p, h->ro_key_alt.p, h->ro_val.p, h->ro_val_alt.p, h->ro_tmp.p, h->ro_tmp.cap};
        RMB_CK(h, launch_reorder(nd, h->gid.p, h->geom.p, h->groups, h->n_work, sc, h->perm.p, h->st));
    }
    RMB_CK(h, launch_regather(nd, work_arrays(h), h->geom.p, h->st));
    if (planar_known < 0) {
        RMB_CK(h, cudaStreamSynchronize(h->st));
        h->planar = g.planar != 0;
        h->planar_z = g.zmin;
    } else {
        h->planar = planar_known != 0;
    }
    return RMB_OK;
}

template <class T>
static cudaError_t upload(DevBuf<T> &dst, const T *src, T dflt, int n, std::vector<T> *mirror)
{
    cudaError_t e = dst.ensure(std::max(n, 1));
    if (e != cudaSuccess) return e;
    std::vector<T> tmp;
    const T *from = src;
    if (!src) { tmp.assign((size_t)n, dflt); from = tmp.data(); }
    if (mirror) mirror->assign(from, from + n);
    if (n > 0) return cudaMemcpy(dst.p, from, (size_t)n * sizeof(T), cudaMemcpyHostToDevice);
    return cudaSuccess;
}

extern "C" {

int rmb_version(void) { return RMB_ABI_VERSION; }

void rmb_config_default(rmb_config *cfg, int32_t model)
{
    if (!cfg) return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->struct_size = (int32_t)sizeof(rmb_config);
    cfg->model = model;
    cfg->success_ratio_tx = 1.0;      // UDGMRadioMedium.java:17
    cfg->success_ratio_rx = 1.0;      // UDGMRadioMedium.java:18
    cfg->transmission_range = 50.0;   // UDGMRadioMedium.java:23
    cfg->constant_range = 100.0;      // UDGMConstantLossRadioMedium.java:8
    cfg->pl0_db = 40.05; cfg->pl_exponent = 3.0; cfg->d0_m = 1.0;
    cfg->sensitivity_dbm = -95.0; cfg->sinr_threshold_db = 4.0;
    cfg->noise_floor_dbm = -100.0;    // AbstractRadioMedium.java:38
}

const char *rmb_last_error(const rmb_handle *h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int rmb_create(const rmb_config *cfg, rmb_handle **out)
{
    if (!cfg || !out) { g_create_error = "null argument"; return RMB_ERR_INVALID; }
    *out = nullptr;
    if (cfg->struct_size != (int32_t)sizeof(rmb_config)) { g_create_error = "rmb_config.struct_size mismatch"; return RMB_ERR_INVALID; }
    if (cfg->model < RMB_MODEL_NULL || cfg->model > RMB_MODEL_SINR) { g_create_error = "unknown model"; return RMB_ERR_INVALID; }
    if (cfg->model == RMB_MODEL_SINR && !(cfg->pl_exponent > 0.0 && cfg->d0_m > 0.0)) { g_create_error = "pl_exponent and d0_m must be > 0"; return RMB_ERR_INVALID; }
    // row and segment offsets are int32 (seg, tx_first, draw scans; the record index rides in 32 bits of an ordering item)
    if (cfg->max_pairs > (int64_t)INT32_MAX) { g_create_error = "max_pairs must not exceed 2^31 - 1"; return RMB_ERR_INVALID; }
    // N2N looks the SOURCE's integer id up in the node table, and a tx record does not carry it across shards
    if (cfg->model == RMB_MODEL_N2N && cfg->n_shards > 1) { g_create_error = "RMB_MODEL_N2N does not support sharded receivers (n_shards > 1)"; return RMB_ERR_UNSUPPORTED; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        g_create_error = "no CUDA device available (this medium has no CPU fallback)";
        return RMB_ERR_NO_DEVICE;
    }
    if (cfg->device < 0 || cfg->device >= ndev) { g_create_error = "device ordinal out of range"; return RMB_ERR_INVALID; }
    rmb_handle *h = new rmb_handle();
    h->cfg = *cfg;
    if (h->cfg.max_pairs <= 0) h->cfg.max_pairs = 64ll << 20;
    auto bail = [&](cudaError_t ce, const char *w) { g_create_error = std::string(w) + ": " + cudaGetErrorString(ce); delete h; return RMB_ERR_CUDA; };
    if ((e = cudaSetDevice(cfg->device)) != cudaSuccess) return bail(e, "cudaSetDevice");
    if ((e = cudaStreamCreateWithFlags(&h->own_st, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    h->st = h->own_st;
    if ((e = cudaStreamCreateWithFlags(&h->st2, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    for (auto &ev : h->ev) if ((e = cudaEventCreate(&ev)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreateWithFlags(&h->ev_h2d, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreateWithFlags(&h->ev_staged, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaStreamCreateWithFlags(&h->st_copy, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    if ((e = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = h->dp.ensure(1)) != cudaSuccess || (e = h->geom.ensure(1)) != cudaSuccess ||
        (e = h->geom_scratch.ensure(8)) != cudaSuccess || (e = h->ctr.ensure(1)) != cudaSuccess ||
        (e = h->h_ctr.ensure(1)) != cudaSuccess || (e = h->theta_min_g.ensure(kMaxGroups)) != cudaSuccess ||
        (e = h->goff.ensure(kMaxGroups + 1)) != cudaSuccess || (e = h->gcount.ensure(kMaxGroups)) != cudaSuccess ||
        (e = h->blk_cnt.ensure((size_t)kPrepMaxGrid * kMaxGroups)) != cudaSuccess ||
        (e = h->blk_thmin.ensure((size_t)kPrepMaxGrid * kMaxGroups)) != cudaSuccess || (e = h->bar.ensure(4)) != cudaSuccess)
        return bail(e, "cudaMalloc");
    if ((e = cudaMemset(h->bar.p, 0, 4 * sizeof(unsigned))) != cudaSuccess) return bail(e, "cudaMemset");
    DevParams dp{};
    dp.model = cfg->model; dp.success_ratio_rx = cfg->success_ratio_rx; dp.transmission_range = cfg->transmission_range;
    dp.constant_range = cfg->constant_range; dp.pl0_db = cfg->pl0_db; dp.pl_exponent = cfg->pl_exponent; dp.d0_m = cfg->d0_m;
    dp.sensitivity_dbm = cfg->sensitivity_dbm; dp.sinr_threshold_db = cfg->sinr_threshold_db;
    dp.noise_floor_dbm = cfg->noise_floor_dbm; dp.rx_base = cfg->rx_base; dp.half_duplex = cfg->half_duplex ? 1 : 0;
    if ((e = cudaMemcpy(h->dp.p, &dp, sizeof(dp), cudaMemcpyHostToDevice)) != cudaSuccess) return bail(e, "cudaMemcpy");
    if ((e = launch_init_params(h->dp.p, h->st)) != cudaSuccess) return bail(e, "init_params");
    if ((e = cudaStreamSynchronize(h->st)) != cudaSuccess) return bail(e, "init_params sync");
    *out = h;
    return RMB_OK;
}

int rmb_destroy(rmb_handle *h)
{
    if (h) h->pending = false;
    if (!h) return RMB_ERR_INVALID;
    cudaSetDevice(h->cfg.device);
    cudaStreamSynchronize(h->st);
    // DevBuf/PinBuf have no destructors on purpose (explicit lifetime): free through cudaDeviceReset-safe calls
    h->x.release(); h->y.release(); h->z.release(); h->txpower.release(); h->rxprob.release(); h->txprob.release();
    h->matrix.release(); h->channel.release(); h->id_int.release(); h->enabled.release(); h->perm.release(); h->wx.release(); h->wy.release(); h->wz.release(); h->wrxp.release(); h->gid.release(); h->ro_key.release(); h->ro_key_alt.release(); h->ro_val.release(); h->ro_val_alt.release(); h->ro_tmp.release();
    h->fx.release(); h->fy.release(); h->fz.release(); h->dp.release(); h->geom.release(); h->geom_scratch.release();
    h->ctr.release(); h->h_ctr.release(); h->node_rssi.release(); h->node_rssi64.release(); h->tx_epoch.release(); h->hd_cnt.release(); h->node_recv.release(); h->pk_src.release();
    h->pk_channel.release(); h->pk_power.release(); h->pk_stage.release(); h->rec.release(); h->txf.release(); h->stg_f.release(); h->stg_theta.release(); h->stg_gid.release(); h->blk_cnt.release(); h->blk_thmin.release(); h->bar.release(); h->recheck_part.release(); h->rx_done.release();
    h->theta.release(); h->theta_min_g.release(); h->arrival.release(); h->goff.release(); h->gcount.release();
    h->uniforms.release(); h->slots.release(); h->cnt.release(); h->hcnt.release(); h->hoff.release();
    h->recheck.release(); h->recheck2.release(); h->recheck_cnt.release(); h->trace.release(); h->far_part.release(); h->sl_prx.release(); h->sl_lin.release(); h->sl_sinr.release();
    h->itot.release(); h->sl_flags.release(); h->key.release(); h->key_alt.release(); h->val.release();
    h->val_alt.release(); h->pr_p.release(); h->pr_rssi.release(); h->pr_sinr.release(); h->pr_cand.release();
    h->draw_scan.release(); h->seg.release(); h->scan_tmp.release(); h->blkcnt.release(); h->tx_base.release();
    h->tx_failed.release(); h->cub_tmp.release(); h->o_tx.release(); h->o_rx.release(); h->o_rssi.release();
    h->o_sinr.release(); h->o_outcome.release(); h->o_rssi64.release(); h->o_sinr64.release(); h->ho_tx.release();
    h->ho_rx.release(); h->ho_first.release(); h->o_first.release(); h->ho_rssi.release(); h->ho_sinr.release(); h->ho_outcome.release(); h->ho_failed.release();
    h->ho_rssi64.release(); h->ho_sinr64.release(); h->o_slab.release(); h->ho_slab.release();
    h->fb_key.release(); h->fb_key_alt.release(); h->fb_val.release(); h->fb_val_alt.release(); h->rec_prx.release();
    h->rec_sinr.release(); h->rec_flags.release(); h->rec_rank.release(); h->tx_cnt.release(); h->first.release();
    for (auto &ev : h->ev) if (ev) cudaEventDestroy(ev);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_h2d) cudaEventDestroy(h->ev_h2d);
    if (h->ev_staged) cudaEventDestroy(h->ev_staged);
    if (h->st_copy) { cudaStreamSynchronize(h->st_copy); cudaStreamDestroy(h->st_copy); }
    h->x2.release(); h->y2.release(); h->z2.release();
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    for (int r = 0; r < kMaxPeers; ++r) if (h->peer.ipc[r] && h->peer.peer[r]) cudaIpcCloseMemHandle(h->peer.peer[r]);
    if (h->peer.buf) cudaFree(h->peer.buf);
    if (h->peer.done) cudaFree(h->peer.done);
    if (h->st2) { cudaStreamSynchronize(h->st2); cudaStreamDestroy(h->st2); }
    if (h->own_st) cudaStreamDestroy(h->own_st);
    delete h;
    return RMB_OK;
}

int rmb_set_nodes(rmb_handle *h, int32_t n, const double *x, const double *y, const double *z,
                  const double *txpower_dbm, const int32_t *channel, const uint8_t *enabled,
                  const double *rxprob, const double *txprob, const int32_t *id_int)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (n < 0 || (n > 0 && (!x || !y))) return h->fail(RMB_ERR_INVALID, "rmb_set_nodes: n < 0 or missing x/y");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    h->n = n;
    RMB_CK(h, upload<double>(h->x, x, 0.0, n, nullptr));
    RMB_CK(h, upload<double>(h->y, y, 0.0, n, nullptr));
    RMB_CK(h, upload<double>(h->z, z, 0.0, n, nullptr));
    RMB_CK(h, upload<double>(h->txpower, txpower_dbm, 0.0, n, &h->h_txpower));
    RMB_CK(h, upload<int>(h->channel, channel, 26, n, &h->h_channel));          // Transciever.java:12
    RMB_CK(h, upload<uint8_t>(h->enabled, enabled, (uint8_t)1, n, &h->h_enabled));
    RMB_CK(h, upload<double>(h->rxprob, rxprob, 1.0, n, &h->h_rxprob));         // Transciever.java:17-18
    RMB_CK(h, upload<double>(h->txprob, txprob, 1.0, n, &h->h_txprob));
    std::vector<int> ids;
    if (!id_int) { ids.resize(n); for (int i = 0; i < n; ++i) ids[i] = h->cfg.rx_base + i + 1; }
    RMB_CK(h, upload<int>(h->id_int, id_int ? id_int : ids.data(), 0, n, nullptr));
    for (auto &e : h->h_enabled) e = e ? 1 : 0;
    int rc = build_work_order(h);
    if (rc) return rc;
    update_draw_flag(h);
    rc = refresh_geometry(h, true);
    if (rc) return rc;
    h->have_nodes = true;
    return RMB_OK;
}

int rmb_set_shard_base(rmb_handle *h, int32_t rx_base)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (rx_base < 0) return h->fail(RMB_ERR_INVALID, "rmb_set_shard_base: negative index");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    h->cfg.rx_base = rx_base;
    RMB_CK(h, cudaMemcpy(&h->dp.p->rx_base, &rx_base, sizeof(int), cudaMemcpyHostToDevice));
    h->have_nodes = false;                 // the node table that follows is the one that counts
    return RMB_OK;
}

int32_t rmb_may_draw(rmb_handle *h)
{
    if (!h) return 0;
    std::lock_guard<std::mutex> lk(h->mu);
    return h->maybe_draws ? 1 : 0;
}

int rmb_set_matrix(rmb_handle *h, int32_t n, const double *matrix)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (n < 0 || (n > 0 && !matrix)) return h->fail(RMB_ERR_INVALID, "rmb_set_matrix: bad arguments");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    RMB_CK(h, h->matrix.ensure(std::max<size_t>((size_t)n * n, 1)));
    if (n > 0) RMB_CK(h, cudaMemcpy(h->matrix.p, matrix, (size_t)n * n * sizeof(double), cudaMemcpyHostToDevice));
    h->matrix_n = n;
    return RMB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// K4 entry points
// ------------------------------------------------------------------------------------------
extern "C" int rmb_update_positions(rmb_handle *h, int32_t count, const int32_t *idx,
                                    const double *x, const double *y, const double *z)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_update_positions before rmb_set_nodes");
    if (count < 0 || (count > 0 && (!x || !y))) return h->fail(RMB_ERR_INVALID, "rmb_update_positions: bad arguments");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    if (count == 0) return RMB_OK;
    int planar_known = -1;
    if (!idx) {
        if (count != h->n) return h->fail(RMB_ERR_INVALID, "dense position update needs count == n");
        const size_t b = (size_t)count * sizeof(double);
        RMB_CK(h, cudaMemcpyAsync(h->x.p, x, b, cudaMemcpyHostToDevice, h->st));
        RMB_CK(h, cudaMemcpyAsync(h->y.p, y, b, cudaMemcpyHostToDevice, h->st));
        if (z) RMB_CK(h, cudaMemcpyAsync(h->z.p, z, b, cudaMemcpyHostToDevice, h->st));
        else   RMB_CK(h, cudaMemsetAsync(h->z.p, 0, b, h->st));       // Position.set(x, y) zeroes z (Position.java:44-46)
        RMB_CK(h, cudaEventRecord(h->ev_h2d, h->st));
        // planarity is decided here, from the caller's z, so that no device round trip sits on the per-tick path
        planar_known = 1;
        h->planar_z = z ? z[0] : 0.0;
        if (z) for (int i = 1; i < count; ++i) if (z[i] != z[0]) { planar_known = 0; break; }
    } else {
        for (int i = 0; i < count; ++i)
            if (idx[i] < 0 || idx[i] >= h->n) return h->fail(RMB_ERR_INVALID, "rmb_update_positions: node index out of range");
        // staging: idx | x | y | z in the packet staging buffers (reused; ticks are serialised by the mutex)
        RMB_CK(h, h->pk_src.ensure(count)); RMB_CK(h, h->uniforms.ensure((size_t)count * 3));
        RMB_CK(h, cudaMemcpyAsync(h->pk_src.p, idx, (size_t)count * sizeof(int), cudaMemcpyHostToDevice, h->st));
        const size_t b = (size_t)count * sizeof(double);
        double *dx = h->uniforms.p, *dy = dx + count, *dz = dy + count;
        RMB_CK(h, cudaMemcpyAsync(dx, x, b, cudaMemcpyHostToDevice, h->st));
        RMB_CK(h, cudaMemcpyAsync(dy, y, b, cudaMemcpyHostToDevice, h->st));
        if (z) RMB_CK(h, cudaMemcpyAsync(dz, z, b, cudaMemcpyHostToDevice, h->st));
        RMB_CK(h, launch_scatter_positions(node_arrays(h), count, h->pk_src.p, dx, dy, z ? dz : nullptr, h->st));
    }
    // The work order (Morton order inside a channel group) only affects speed -- warp-coherent near re-scans in the
    // sweep -- never results.  A sparse update keeps it; dense updates (per-tick mobility, BASELINE configs[4])
    // re-sort on the first and then every kReorderEvery-th call: nodes that move a little per tick stay sorted
    // well enough, and the re-sort (radix sort of all receivers) costs as much as the upload itself.
    constexpr int kReorderEvery = 8;
    const bool reorder = idx == nullptr && (h->dense_updates++ % kReorderEvery) == 0;
    if (idx && h->planar && (z ? false : h->planar_z == 0.0)) planar_known = 1;     // sparse Position.set(x, y) on a z = 0 plane
    const int rc = refresh_geometry(h, reorder, planar_known);
    if (rc) return rc;
    // the caller may reuse its arrays on return: wait for the upload (not for the kernels queued behind it)
    if (!idx) RMB_CK(h, cudaEventSynchronize(h->ev_h2d));
    else if (planar_known >= 0) RMB_CK(h, cudaStreamSynchronize(h->st));           // staged copies of idx / x / y / z
    return RMB_OK;
}

// positions staged by rmb_update_positions_async become the node table's: the handle's stream waits for the upload,
// the two sets of arrays trade places, the work-order copy is rebuilt -- all enqueued, nothing waits on the host
static int apply_staged(rmb_handle *h)
{
    if (!h->staged) return RMB_OK;
    h->staged = false;
    if (h->staged_n != h->n || !h->have_nodes) return RMB_OK;           // the node table changed meanwhile: the staged set is void
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    RMB_CK(h, cudaStreamWaitEvent(h->st, h->ev_staged, 0));
    std::swap(h->x.p, h->x2.p); std::swap(h->x.cap, h->x2.cap);
    std::swap(h->y.p, h->y2.p); std::swap(h->y.cap, h->y2.cap);
    if (h->staged_has_z) { std::swap(h->z.p, h->z2.p); std::swap(h->z.cap, h->z2.cap); }
    else RMB_CK(h, cudaMemsetAsync(h->z.p, 0, (size_t)h->n * sizeof(double), h->st));      // Position.set(x, y) zeroes z
    h->planar_z = h->staged_z;
    constexpr int kReorderEvery = 8;
    const bool reorder = (h->dense_updates++ % kReorderEvery) == 0;
    return refresh_geometry(h, reorder, h->staged_planar);
}

extern "C" int rmb_update_positions_async(rmb_handle *h, int32_t count, const double *x, const double *y, const double *z)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    // (no settling here: the point is to run beside an enqueued tick)
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_update_positions_async before rmb_set_nodes");
    if (count != h->n || (count > 0 && (!x || !y))) return h->fail(RMB_ERR_INVALID, "rmb_update_positions_async: dense updates only (count == n, x and y given)");
    if (count == 0) return RMB_OK;
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    if (h->staged) RMB_CK(h, cudaStreamSynchronize(h->st_copy));        // a staged set nobody consumed is overwritten
    const size_t b = (size_t)count * sizeof(double);
    RMB_CK(h, h->x2.ensure(count)); RMB_CK(h, h->y2.ensure(count));
    if (z) RMB_CK(h, h->z2.ensure(count));
    // the staging arrays were the node table before the last swap: the tick that read them must be done before they
    // are overwritten -- it is: a swap only happens when the handle is settled, i.e. after that tick's results were taken
    RMB_CK(h, cudaMemcpyAsync(h->x2.p, x, b, cudaMemcpyHostToDevice, h->st_copy));
    RMB_CK(h, cudaMemcpyAsync(h->y2.p, y, b, cudaMemcpyHostToDevice, h->st_copy));
    if (z) RMB_CK(h, cudaMemcpyAsync(h->z2.p, z, b, cudaMemcpyHostToDevice, h->st_copy));
    RMB_CK(h, cudaEventRecord(h->ev_staged, h->st_copy));
    h->staged_planar = 1; h->staged_z = z ? z[0] : 0.0; h->staged_has_z = z != nullptr;
    if (z) for (int i = 1; i < count; ++i) if (z[i] != z[0]) { h->staged_planar = 0; break; }
    h->staged = true; h->staged_n = count;
    return RMB_OK;
}

extern "C" int rmb_update_radio(rmb_handle *h, int32_t count, const int32_t *idx, const double *txpower_dbm,
                                const int32_t *channel, const uint8_t *enabled, const double *rxprob, const double *txprob)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_update_radio before rmb_set_nodes");
    if (count < 0) return h->fail(RMB_ERR_INVALID, "rmb_update_radio: count < 0");
    if (!idx && count != h->n && count != 0) return h->fail(RMB_ERR_INVALID, "dense radio update needs count == n");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    for (int i = 0; i < count; ++i) {
        const int p = idx ? idx[i] : i;
        if (p < 0 || p >= h->n) return h->fail(RMB_ERR_INVALID, "rmb_update_radio: node index out of range");
    }
    bool regroup = false;
    for (int i = 0; i < count; ++i) {
        const int p = idx ? idx[i] : i;
        if (txpower_dbm) h->h_txpower[p] = txpower_dbm[i];
        if (rxprob) h->h_rxprob[p] = rxprob[i];
        if (txprob) h->h_txprob[p] = txprob[i];
        if (channel && h->h_channel[p] != channel[i]) { h->h_channel[p] = channel[i]; regroup = true; }
        if (enabled && h->h_enabled[p] != (enabled[i] ? 1 : 0)) { h->h_enabled[p] = enabled[i] ? 1 : 0; regroup = true; }
    }
    const size_t n = (size_t)h->n;
    if (txpower_dbm) RMB_CK(h, cudaMemcpy(h->txpower.p, h->h_txpower.data(), n * sizeof(double), cudaMemcpyHostToDevice));
    if (rxprob) RMB_CK(h, cudaMemcpy(h->rxprob.p, h->h_rxprob.data(), n * sizeof(double), cudaMemcpyHostToDevice));
    if (txprob) RMB_CK(h, cudaMemcpy(h->txprob.p, h->h_txprob.data(), n * sizeof(double), cudaMemcpyHostToDevice));
    if (channel) RMB_CK(h, cudaMemcpy(h->channel.p, h->h_channel.data(), n * sizeof(int), cudaMemcpyHostToDevice));
    if (enabled) RMB_CK(h, cudaMemcpy(h->enabled.p, h->h_enabled.data(), n, cudaMemcpyHostToDevice));
    update_draw_flag(h);
    if (regroup) {
        int rc = build_work_order(h);
        if (rc) return rc;
        return refresh_geometry(h, true);
    }
    if (rxprob) return refresh_geometry(h, false);        // the work-order copy of rxProbability
    return RMB_OK;
}

extern "C" int rmb_set_tuning(rmb_handle *h, int32_t sweep_variant, int32_t near_slots)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    // low byte: kernel variant; bits 8..15: forced number of tx splits (0 = planned per tick)
    h->sweep_variant = sweep_variant & 0xff;
    h->forced_splits = (sweep_variant >> 8) & 0xff;
    h->fuse_resolve = ((sweep_variant >> 16) & 1) == 0;
    h->fuse_finish = ((sweep_variant >> 20) & 1) == 0;
    h->kernel_events = ((sweep_variant >> 24) & 1) != 0;
    h->fine8 = ((sweep_variant >> 21) & 7) ? ((sweep_variant >> 21) & 7) : 4;       // bits 21..23: eighths of a wave resolved in 32-receiver passes (7: none)
    h->taper = ((sweep_variant >> 17) & 7) ? ((sweep_variant >> 17) & 7) : 2;      // bits 17..19: split multiplier of the final wave
    if (near_slots > 0 && near_slots != h->K) { h->K = near_slots; return ensure_slots(h); }
    return RMB_OK;
}

static int node_info_locked(rmb_handle *h, float *rssi_dbm, double *rssi_dbm64, uint8_t *receiving)
{
    RMB_SETTLE(h);
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_node_info before rmb_set_nodes");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    if (h->node_info_dirty) {
        RMB_CK(h, launch_node_info(work_arrays(h), slot_arrays(h), h->node_rssi.p, h->node_rssi64.p, h->node_recv.p, h->st));
        h->node_info_dirty = false;
    }
    RMB_CK(h, cudaStreamSynchronize(h->st));
    if (rssi_dbm && h->n) RMB_CK(h, cudaMemcpy(rssi_dbm, h->node_rssi.p, (size_t)h->n * sizeof(float), cudaMemcpyDeviceToHost));
    if (rssi_dbm64 && h->n) RMB_CK(h, cudaMemcpy(rssi_dbm64, h->node_rssi64.p, (size_t)h->n * sizeof(double), cudaMemcpyDeviceToHost));
    if (receiving && h->n) RMB_CK(h, cudaMemcpy(receiving, h->node_recv.p, (size_t)h->n, cudaMemcpyDeviceToHost));
    return RMB_OK;
}

extern "C" int rmb_node_info(rmb_handle *h, float *rssi_dbm, uint8_t *receiving)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    return node_info_locked(h, rssi_dbm, nullptr, receiving);
}

extern "C" int rmb_node_info_f64(rmb_handle *h, double *rssi_dbm, uint8_t *receiving)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    return node_info_locked(h, nullptr, rssi_dbm, receiving);
}

extern "C" int rmb_get_stream(rmb_handle *h, void **stream)
{
    if (!h || !stream) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    *stream = (void *)h->st;
    return RMB_OK;
}

extern "C" int rmb_set_stream(rmb_handle *h, void *stream)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    RMB_CK(h, cudaStreamSynchronize(h->st));
    h->st = stream ? (cudaStream_t)stream : h->own_st;
    return RMB_OK;
}

extern "C" int rmb_last_timing(const rmb_handle *h, rmb_timing *out)
{
    if (!h || !out) return RMB_ERR_INVALID;
    rmb_handle *m = const_cast<rmb_handle *>(h);
    std::lock_guard<std::mutex> lk(m->mu);
    finish_pending(m);
    if (m->timing_pending) {
        rmb_timing &t = m->timing;
        cudaEventElapsedTime(&t.ms_total, m->ev[0], m->ev[4]);
        if (m->timing_by_events) {
            cudaEventElapsedTime(&t.ms_prepare, m->ev[0], m->ev[1]);
            cudaEventElapsedTime(&t.ms_sweep, m->ev[1], m->ev[2]);
            cudaEventElapsedTime(&t.ms_resolve, m->ev[2], m->ev[3]);
            cudaEventElapsedTime(&t.ms_emit, m->ev[3], m->ev[4]);
        } else {
            // no event between the kernels (they overlap at their edges): the kernels' own globaltimer stamps
            const Counters &c = *m->h_ctr.p;
            const unsigned long long sb = ~c.t_sweep_begin_inv, fb = ~c.t_finish_begin_inv;
            const bool ok = c.t_prepare_begin && c.t_sweep_end && sb >= c.t_prepare_begin && c.t_sweep_end >= sb;
            t.ms_prepare = ok ? (float)((sb - c.t_prepare_begin) * 1e-6) : 0.f;
            t.ms_sweep = ok ? (float)((c.t_sweep_end - sb) * 1e-6) : 0.f;
            t.ms_resolve = 0.f;
            t.ms_emit = (ok && c.t_finish_begin_inv) ? t.ms_total - (float)((c.t_sweep_end - c.t_prepare_begin) * 1e-6) : 0.f;
            (void)fb;
        }
        cudaGetLastError();
        m->timing_pending = false;
    }
    *out = m->timing;
    return RMB_OK;
}

// ------------------------------------------------------------------------------------------
// the tick
// ------------------------------------------------------------------------------------------
static int ensure_pairs(rmb_handle *h, long long m)
{
    if (m > h->cfg.max_pairs) {
        char b[160]; snprintf(b, sizeof b, "tick produced %lld heard pairs, max_pairs is %lld", m, (long long)h->cfg.max_pairs);
        return h->fail(RMB_ERR_CAPACITY, b);
    }
    const size_t c = (size_t)std::max<long long>(m, 1);
    RMB_CK(h, h->key.ensure(c));
    const size_t cap = h->key.cap;       // grow all pair arrays to the same capacity
    RMB_CK(h, h->pr_p.ensure(cap)); RMB_CK(h, h->pr_rssi.ensure(cap)); RMB_CK(h, h->pr_sinr.ensure(cap));
    RMB_CK(h, h->pr_cand.ensure(cap)); RMB_CK(h, h->draw_scan.ensure(cap + 2));
    RMB_CK(h, h->o_tx.ensure(cap)); RMB_CK(h, h->o_rx.ensure(cap)); RMB_CK(h, h->o_rssi.ensure(cap));
    RMB_CK(h, h->o_sinr.ensure(cap)); RMB_CK(h, h->o_outcome.ensure(cap));
    if (h->cfg.flags & RMB_FLAG_F64_VALUES) { RMB_CK(h, h->o_rssi64.ensure(cap)); RMB_CK(h, h->o_sinr64.ensure(cap)); }
    RMB_CK(h, h->scan_tmp.ensure(cap / 1024 + 8));
    if (h->cfg.model == RMB_MODEL_SINR) {
        RMB_CK(h, h->key_alt.ensure(cap)); RMB_CK(h, h->val.ensure(cap));
        RMB_CK(h, h->rec_prx.ensure(cap)); RMB_CK(h, h->rec_sinr.ensure(cap)); RMB_CK(h, h->rec_flags.ensure(cap));
        RMB_CK(h, h->rec_rank.ensure(cap));
    }
    return RMB_OK;
}

static int read_counters(rmb_handle *h)
{
    RMB_CK(h, cudaMemcpyAsync(h->h_ctr.p, h->ctr.p, sizeof(Counters), cudaMemcpyDeviceToHost, h->st));
    RMB_CK(h, cudaStreamSynchronize(h->st));
    return RMB_OK;
}

// counters (host copy) -> timing bookkeeping and the counts of h->last
static int publish(rmb_handle *h, int n_tx, long long m, bool draws, int launches, int sweeps)
{
    const Counters &c = *h->h_ctr.p;
    rmb_timing &t = h->timing;
    h->timing_pending = true;          // the event durations are read on demand (rmb_last_timing)
    h->timing_by_events = h->kernel_events || h->cfg.model != RMB_MODEL_SINR;
    t.sweep_launches = sweeps; t.kernel_launches = launches; t.recheck_sequential = c.n_recheck2;
    rmb_result &r = h->last;
    memset(&r, 0, sizeof(r));
    r.n_pairs = m; r.n_delivered = c.n_delivered; r.n_interfered = c.n_interfered; r.n_decisions = c.decisions;
    r.n_uniforms_used = draws ? c.uniforms_used : 0; r.n_recheck = c.n_recheck; r.n_tx = n_tx; r.near_slots = h->K;
    h->last_pairs = m; h->last_ntx = n_tx;
    if (draws && c.uniforms_short) { h->err = "uniform stream exhausted; missing draws were read as 0.0"; return RMB_ERR_UNIFORMS; }
    return RMB_OK;
}

static int compute(rmb_handle *h, int n_tx, const rmb_tx_record *d_rec, const double *uniforms, size_t n_uniforms);
static const char *const kNeedsDrawsMsg =
    "a tx record carries a transmission probability strictly between 0 and 1 but this handle evaluates the tick "
    "without draws (its own nodes are all deterministic): sharded receivers need rmb_tick_draws_begin / rmb_tick_draws_end";

// An enqueued tick is settled the first time the host needs anything from the handle: wait for it, look at the
// counters, and in the rare case a capacity was exceeded (or a packet's bucket needs the radix-sort path) redo the
// tick synchronously -- the computation is deterministic, so the redo gives the same lists.
static int finish_pending(rmb_handle *h)
{
    if (!h->pending) return RMB_OK;
    h->pending = false;
    RMB_CK(h, cudaStreamSynchronize(h->st));
    const Counters c = *h->h_ctr.p;
    if (c.peer_timeout) return h->fail(RMB_ERR_PEER, "peer exchange: a rank's tx records did not arrive (no rmb_peer_push on that rank?)");
    if (c.needs_draws) return h->fail(RMB_ERR_UNSUPPORTED, kNeedsDrawsMsg);
    const bool redo = c.max_near > h->K || (long long)c.heard > (long long)h->key.cap || c.max_bucket > kBucketSortMax;
    if (redo) {
        const bool am = h->async_mode;
        h->async_mode = false;
        h->wait_flags = h->pend.flags; h->wait_epoch = h->pend.epoch; h->wait_world = h->pend.world;
        const int rc = compute(h, h->pend.n_tx, h->pend.rec, nullptr, 0);
        h->wait_flags = nullptr; h->wait_world = 0;
        h->async_mode = am;
        return rc;
    }
    h->node_info_dirty = true;
    h->slab_ready = h->pend.slab;
    return publish(h, h->pend.n_tx, (long long)c.heard, false, h->pend.launches, h->pend.sweeps);
}

// d_rec: n_tx records in device memory.  Leaves the result lists on the device.
static int compute(rmb_handle *h, int n_tx, const rmb_tx_record *d_rec, const double *uniforms, size_t n_uniforms)
{
    int rc;
    int launches = 0, sweeps = 0;
    const bool lists_wanted = h->want_download;
    h->want_download = false; h->slab_copied = 0;
    if (h->cfg.half_duplex) ++h->hd_epoch;            // every evaluation stamps its own sources (a redo stamps them again)
    const NodeArrays nd = node_arrays(h);
    const WorkArrays wk = work_arrays(h);
    const int *hd_ep = h->cfg.half_duplex ? h->tx_epoch.p : nullptr;
    // a shard takes the stochastic path whenever the split API is used: a packet's source may sit on another shard
    const bool draws = h->maybe_draws || h->split_counts != nullptr;
    if (draws && h->cfg.n_shards > 1 && !h->split_counts)
        return h->fail(RMB_ERR_UNSUPPORTED, "stochastic reception (rx/tx probability < 1) on sharded receivers needs "
                                            "rmb_tick_draws_begin / rmb_tick_draws_end (the draw order is global)");
    if (h->split_counts && h->cfg.model != RMB_MODEL_SINR && h->cfg.model != RMB_MODEL_UDGM)
        return h->fail(RMB_ERR_UNSUPPORTED, "rmb_tick_draws_begin exists for the SINR and UDGM media (the other media draw nothing "
                                            "or cannot be sharded)");
    h->split.active = false;
    h->slab_ready = false;
    RMB_CK(h, h->seg.ensure((size_t)n_tx + 2)); RMB_CK(h, h->tx_base.ensure((size_t)n_tx + 1));
    RMB_CK(h, h->tx_failed.ensure((size_t)n_tx + 1));
    long long nu = 0;
    if (draws && uniforms && n_uniforms) {
        RMB_CK(h, h->uniforms.ensure(n_uniforms));
        RMB_CK(h, cudaMemcpyAsync(h->uniforms.p, uniforms, n_uniforms * sizeof(double), cudaMemcpyHostToDevice, h->st));
        nu = (long long)n_uniforms;
    } else {
        RMB_CK(h, h->uniforms.ensure(1));
    }
    if (h->cfg.model != RMB_MODEL_SINR) RMB_CK(h, cudaMemsetAsync(h->ctr.p, 0, sizeof(Counters), h->st));   // SINR: zeroed by the preparation kernel
    RMB_CK(h, cudaEventRecord(h->ev[0], h->st));
    long long m = 0;
    bool final_sync = true;           // the SINR path usually synchronises once, inside its attempt loop
    if (h->cfg.model == RMB_MODEL_SINR) {
        const size_t tcap = (size_t)n_tx + (size_t)(h->groups.ngroups + 1) * kTile;
        RMB_CK(h, h->txf.ensure(4 * ((tcap + 63) / 64 * 64))); RMB_CK(h, h->theta.ensure(tcap)); RMB_CK(h, h->arrival.ensure(tcap));
        RMB_CK(h, h->tx_cnt.ensure((size_t)n_tx + 1));
        TxTiles tx = tx_tiles(h);
        const int mode = sweep_mode_for_exponent(h->cfg.pl_exponent);
        // default: 4 receivers per thread, 128 threads (the fastest variant at every shard size once the blocks are short)
        const int n_tx_bucket = n_tx / std::max(1, h->groups.ngroups);
        const int variant = h->sweep_variant ? h->sweep_variant : sweep_auto_variant(mode, h->planar, h->n_work, n_tx_bucket);
        h->splits = sweep_splits(mode, h->planar, variant, h->n_work, n_tx_bucket,
                                 sweep_plan(mode, h->planar, variant, h->n_work, n_tx_bucket, h->forced_splits), h->taper, h->fine8);
        RMB_CK(h, h->far_part.ensure((size_t)h->splits.s2 * (size_t)std::max(h->n_work, 1)));
        RMB_CK(h, h->stg_f.ensure(std::max(n_tx, 1))); RMB_CK(h, h->stg_theta.ensure(std::max(n_tx, 1))); RMB_CK(h, h->stg_gid.ensure(std::max(n_tx, 1)));
        if (h->key.cap == 0 && (rc = ensure_pairs(h, std::min<long long>(h->cfg.max_pairs, std::max(2ll * h->n_work, 1024ll))))) return rc;
        TxStage stg{h->stg_f.p, h->stg_theta.p, h->stg_gid.p};
        const int pgrid = prepare_grid(n_tx);
        PrepareAux aux{};
        aux.stochastic = draws ? 1 : 0;
        if (h->cfg.half_duplex) { aux.tx_epoch = h->tx_epoch.p; aux.hd_epoch = h->hd_epoch; aux.hd_cnt = h->hd_cnt.p; aux.node_gid = h->gid.p; }
        aux.cnt = h->cnt.p; aux.n_work = h->n_work; aux.tx_cnt = h->tx_cnt.p; aux.tx_failed = h->tx_failed.p; aux.tx_base = h->tx_base.p;
        aux.blk_cnt = h->blk_cnt.p; aux.blk_thmin = h->blk_thmin.p; aux.bar = h->bar.p; aux.bar_target = h->bar_count + (unsigned)pgrid;
        aux.rx_done = h->rx_done.p; aux.n_rx_done = 2 * (h->n_work / 32 + 1);
        aux.peer_flags = h->wait_flags; aux.world = h->wait_world; aux.epoch = h->wait_epoch;
        if (h->fused_push) {
            aux.push_local = h->fused_push; aux.push_slots = h->peer.slots; aux.push_rank = h->peer.rank;
            aux.push_off = h->peer.rec_off(h->peer.epoch); aux.push_done = h->peer.done;
            for (int r = 0; r < h->peer.world; ++r) aux.push_peer[r] = h->peer.peer[r];
        }
        RMB_CK(h, launch_prepare(nd, h->dp.p, h->geom.p, h->groups, n_tx, d_rec, stg, tx, h->theta.p, h->theta_min_g.p, h->ctr.p, aux, pgrid, h->st));
        h->bar_count += (unsigned)pgrid;      // only a launch that went out has arrivals: host and device targets stay in step
        launches += 1;
        const bool kev = h->kernel_events;
        if (kev) RMB_CK(h, cudaEventRecord(h->ev[1], h->st));
        const bool use_finish = !draws && h->fuse_finish && n_tx > 0;
        // The whole tick is enqueued without a host round trip: the kernels downstream of the resolve read
        // the pair / recheck counts on the device.  The host looks at the counters once, at the end, and
        // redoes the tick only when a capacity was exceeded (near slots, pair records) -- both grow
        // geometrically, so this happens a handful of times in the life of a handle.
        bool fallback = false;
        for (int attempt = 0;; ++attempt) {
            SweepArgs a;
            a.fx = h->fx.p; a.fy = h->fy.p; a.fz = h->fz.p; a.groups = h->groups; a.tx = tx; a.theta = h->theta.p;
            a.theta_min_g = h->theta_min_g.p; a.neg_h = (float)(-0.5 * h->cfg.pl_exponent);
            a.far_part = h->far_part.p; a.cnt = h->cnt.p; a.slots = h->slots.p; a.K = h->K;
            if (attempt) {               // attempt 0: the preparation kernel zeroed them
                RMB_CK(h, cudaMemsetAsync(h->cnt.p, 0, (size_t)h->n_work * sizeof(int), h->st));
                RMB_CK(h, cudaMemsetAsync(h->tx_cnt.p, 0, ((size_t)n_tx + 1) * sizeof(int), h->st));
                RMB_CK(h, cudaMemsetAsync(h->rx_done.p, 0, 2 * ((size_t)h->n_work / 32 + 1) * sizeof(int), h->st));
                RMB_CK(h, cudaMemsetAsync(h->bar.p + 2, 0, sizeof(unsigned), h->st));
                if (kev) RMB_CK(h, cudaEventRecord(h->ev[1], h->st));
            }
            // exact resolve (decisions + dense pair records, then the per-packet segment offsets): fused into the
            // sweep as the epilogue of every rx block's last tx split; a kernel of its own for the other variants
            SlotArrays sl = slot_arrays(h);
            PairRecs pr = pair_recs(h);
            PairArrays pa = pair_arrays(h);
            SweepFuse fuse{};
            fuse.rb_spin = h->splits.rb_taper; fuse.rb_fine = h->splits.rb_fine;
            fuse.enabled = (h->fuse_resolve && h->splits.can_fuse) ? 1 : 0;
            fuse.wk = wk; fuse.dp = h->dp.p; fuse.rec = d_rec; fuse.sl = sl; fuse.pr = pr; fuse.ctr = h->ctr.p;
            fuse.n_tx = n_tx; fuse.seg = (use_finish && finish_scans(n_tx)) ? nullptr : h->seg.p; fuse.rx_done = h->rx_done.p; fuse.rx_pass = h->rx_done.p + (h->n_work / 32 + 1);
            static const char *trace_path = getenv("RMB_SWEEP_TRACE");
            const size_t trace_blocks = (size_t)h->splits.rb_taper * h->splits.s1 +
                                        (size_t)(h->n_work / h->splits.block_rx - h->splits.rb_taper) * h->splits.s2;
            if (trace_path) { RMB_CK(h, h->trace.ensure(12 * trace_blocks + 12)); fuse.trace = h->trace.p; }
            // programmatic dependent launch behind the preparation kernel: worth ~3 us on a launch that is several waves
            // deep, but a single-wave launch (16 channel groups on a 1/8 shard) ran 30 us LONGER with it (measured,
            // tools/c5_shard_probe.py with RMB_PDL=0/1), so only deep launches take it
            const bool deep = h->splits.s2 > h->splits.s1;
            RMB_CK(h, launch_sweep(mode, h->planar, variant, h->n_work, h->splits, a, fuse, attempt == 0 && !kev && deep, h->st));
            if (trace_path) {
                std::vector<unsigned long long> tr(12 * trace_blocks);
                RMB_CK(h, cudaMemcpyAsync(tr.data(), h->trace.p, tr.size() * 8, cudaMemcpyDeviceToHost, h->st));
                RMB_CK(h, cudaStreamSynchronize(h->st));
                if (FILE *f = fopen(trace_path, "wb")) {
                    const int hdr[4] = {h->splits.s1, h->splits.s2, h->splits.rb_taper, h->n_work / h->splits.block_rx};
                    fwrite(hdr, sizeof(int), 4, f); fwrite(tr.data(), 8, tr.size(), f); fclose(f);
                }
            }
            ++launches; ++sweeps;
            if (kev) RMB_CK(h, cudaEventRecord(h->ev[2], h->st));
            if (!fuse.enabled) {
                RMB_CK(h, launch_resolve(nd, wk, h->dp.p, tx, d_rec, sl, pr, h->ctr.p, n_tx, h->seg.p, 0, h->st));
                launches += 1;
            }
            const bool dev_publish = use_finish;
            if (kev) RMB_CK(h, cudaEventRecord(h->ev[3], h->st));
            if (!draws && h->fuse_finish && n_tx <= 0) {
                // nothing on the air: no launch (and no barrier arrivals to account for)
            } else if (!draws && h->fuse_finish) {
                // deterministic reception: guard-band recheck, ordering and output rows in one cooperative kernel,
                // whose last block stores the counters into the pinned host copy
                const int fgrid = finish_grid(n_tx);
                RMB_CK(h, h->o_slab.ensure((size_t)slab_layout((long long)h->key.cap, n_tx, h->cfg.flags).total));
                pa.slab = h->o_slab.p;
                RMB_CK(h, launch_finish(wk, h->dp.p, h->groups, tx, d_rec, sl, pr, pa, n_tx, h->seg.p, h->key_alt.p, h->recheck_part.p,
                                        h->ctr.p, h->h_ctr.p, fgrid, h->bar.p + 2, fuse.enabled && !kev && !trace_path, h->st));
                launches += 1;
            } else {
                // guard-band receivers: exact sums on a side stream, overlapped with the ordering
                RMB_CK(h, cudaEventRecord(h->ev_fork, h->st));
                RMB_CK(h, cudaStreamWaitEvent(h->st2, h->ev_fork, 0));
                const int rgrid = recheck_grid();
                h->bar2_count += (unsigned)rgrid;
                RMB_CK(h, launch_recheck(nd, wk, h->dp.p, h->groups, tx, d_rec, sl, pr, h->ctr.p, h->recheck_part.p, rgrid,
                                         h->bar.p + 1, h->bar2_count, h->st2)); launches += 2;
                RMB_CK(h, cudaEventRecord(h->ev_join, h->st2));
                RMB_CK(h, launch_order_buckets(pr, h->ctr.p, n_tx, h->seg.p, h->key_alt.p, h->val.p, h->st)); launches += 2;
                RMB_CK(h, cudaStreamWaitEvent(h->st, h->ev_join, 0));
                if (!draws) { RMB_CK(h, launch_emit_sinr(nd, h->dp.p, d_rec, n_tx, pr, h->val.p, pa, -1, false, h->ctr.p, nullptr, h->st)); launches += 1; }
            }
            RMB_CK(h, cudaEventRecord(h->ev[4], h->st));
            h->slab_copied = 0;
            static const bool guess_off = getenv("RMB_SLAB_GUESS") && atoi(getenv("RMB_SLAB_GUESS")) == 0;    // A/B knob
            if (dev_publish && lists_wanted && !guess_off && !h->async_mode && !draws && h->slab_pred > 0 && n_tx > 0) {
                const size_t guess = std::min(h->slab_pred, h->o_slab.cap);
                RMB_CK(h, h->ho_slab.ensure(guess));
                RMB_CK(h, cudaMemcpyAsync(h->ho_slab.p, h->o_slab.p, guess, cudaMemcpyDeviceToHost, h->st));
                h->slab_copied = guess;
            }
            if (h->async_mode && attempt == 0 && !draws) {
                // enqueue only: the counters travel to the host behind the tick, finish_pending looks at them
                if (!dev_publish) RMB_CK(h, cudaMemcpyAsync(h->h_ctr.p, h->ctr.p, sizeof(Counters), cudaMemcpyDeviceToHost, h->st));
                h->pending = true;
                h->pend.n_tx = n_tx; h->pend.rec = d_rec; h->pend.flags = h->wait_flags; h->pend.epoch = h->wait_epoch;
                h->pend.world = h->wait_world; h->pend.launches = launches; h->pend.sweeps = sweeps; h->pend.slab = use_finish;
                memset(&h->last, 0, sizeof(h->last));
                h->last.n_tx = n_tx; h->last_pairs = 0; h->last_ntx = n_tx;
                return RMB_OK;
            }
            if (dev_publish) RMB_CK(h, cudaStreamSynchronize(h->st));
            else if ((rc = read_counters(h))) return rc;
            const Counters c = *h->h_ctr.p;
            if (c.peer_timeout) return h->fail(RMB_ERR_PEER, "peer exchange: a rank's tx records did not arrive (no rmb_peer_push on that rank?)");
            if (c.needs_draws) return h->fail(RMB_ERR_UNSUPPORTED, kNeedsDrawsMsg);
            const bool slots_short = c.max_near > h->K, recs_short = (long long)c.heard > (long long)h->key.cap;
            if (!slots_short && !recs_short) { fallback = c.max_bucket > kBucketSortMax; break; }
            if (attempt > 5) return h->fail(RMB_ERR_CAPACITY, "near-slot / pair-record capacity could not be established");
            if (slots_short) {
                // a receiver had more near candidates than slots: grow and redo the tick (deterministic)
                int k = h->K;
                while (k < c.max_near) k *= 2;
                if ((double)k * h->n_work > 3.0e9)
                    return h->fail(RMB_ERR_CAPACITY, "too many near candidates per receiver (dense all-hear-all tick)");
                h->K = k;
                if ((rc = ensure_slots(h))) return rc;
            }
            // with too few slots the heard count is a lower bound only: the next attempt re-checks it
            if (recs_short && (rc = ensure_pairs(h, (long long)c.heard))) return rc;
            Counters keep{};
            keep.decisions = c.decisions;
            RMB_CK(h, cudaMemcpyAsync(h->ctr.p, &keep, sizeof(keep), cudaMemcpyHostToDevice, h->st));
            RMB_CK(h, cudaStreamSynchronize(h->st));
        }
        m = (long long)h->h_ctr.p->heard;
        h->node_info_dirty = true;
        h->slab_ready = use_finish && !(fallback && m > 0);       // the finish kernel packed the lists (one copy takes them home)
        final_sync = (fallback && m > 0) || draws;
        PairRecs pr = pair_recs(h);
        PairArrays pa = pair_arrays(h);
        if (fallback && m > 0) {
            // some packet was heard by more receivers than a block sorts: radix sort of the whole key instead
            const size_t cap = h->key.cap;
            RMB_CK(h, h->fb_key.ensure(cap)); RMB_CK(h, h->fb_key_alt.ensure(cap)); RMB_CK(h, h->fb_val.ensure(cap));
            RMB_CK(h, h->fb_val_alt.ensure(cap)); RMB_CK(h, h->cub_tmp.ensure(sort_pairs_tmp_bytes((long long)cap)));
            OrderScratch sc{h->key_alt.p, h->fb_key.p, h->fb_key_alt.p, h->fb_val.p, h->fb_val_alt.p, h->cub_tmp.p, h->cub_tmp.cap};
            RMB_CK(h, launch_order_fallback(pr, m, pa.key_shift + bits_for(n_tx), sc, h->val.p, h->st));
            launches += 7;
            if (!draws) { RMB_CK(h, launch_emit_sinr(nd, h->dp.p, d_rec, n_tx, pr, h->val.p, pa, m, false, h->ctr.p, nullptr, h->st)); launches += 1; }
        }
        if (draws) {
            pa.key = h->key_alt.p;         // sorted keys for the draw-order stage (the bucket items are dead by now)
            RMB_CK(h, launch_emit_sinr(nd, h->dp.p, d_rec, n_tx, pr, h->val.p, pa, m, true, h->ctr.p, nullptr, h->st));
            if (h->split_counts) {
                RMB_CK(h, launch_emit_draw_prepare(n_tx, pa, m, h->split_counts, h->st));
                h->split.active = true; h->split.n_tx = n_tx; h->split.m = m; h->split.rec = d_rec;
                h->split.launches = launches + 6; h->split.sweeps = sweeps;
                if (h->st == h->own_st) RMB_CK(h, cudaStreamSynchronize(h->st));    // the caller gathers the counts on another stream
                return RMB_OK;
            }
            RMB_CK(h, launch_emit(h->dp.p, d_rec, n_tx, pa, m, true, h->uniforms.p, nu, h->ctr.p, h->st));
            launches += 7;
        }
    } else {
        const int nblk = refmedia_blocks_per_tx(h->n);
        const size_t nb = (size_t)nblk * (size_t)n_tx;
        if (nb > 0x7fffffffull) return h->fail(RMB_ERR_CAPACITY, "n_tx * n too large for the reference-media path");
        RMB_CK(h, h->blkcnt.ensure(nb + 2)); RMB_CK(h, h->scan_tmp.ensure(nb / 1024 + 8));
        RMB_CK(h, cudaMemsetAsync(h->node_recv.p, 0, std::max(h->n, 1), h->st));
        // sharded receivers: the peers' records must have landed before anything reads them (the SINR medium's
        // preparation kernel does this wait itself)
        if (h->wait_flags) RMB_CK(h, launch_peer_wait(h->wait_flags, h->wait_world, h->wait_epoch, h->ctr.p, h->st));
        RMB_CK(h, cudaEventRecord(h->ev[1], h->st));
        RMB_CK(h, launch_refmedia_mark(nd, h->dp.p, n_tx, d_rec, h->cfg.half_duplex ? h->tx_epoch.p : nullptr, h->hd_epoch, h->st));
        RMB_CK(h, launch_refmedia_count(nd, h->dp.p, h->matrix.p, h->matrix_n, n_tx, d_rec, h->blkcnt.p, h->ctr.p, draws, hd_ep, h->hd_epoch, h->st));
        ++launches; ++sweeps;
        RMB_CK(h, cudaEventRecord(h->ev[2], h->st));
        if ((rc = read_counters(h))) return rc;
        if (h->h_ctr.p->peer_timeout) return h->fail(RMB_ERR_PEER, "peer exchange: a rank's tx records did not arrive (no rmb_peer_push on that rank?)");
        if (h->h_ctr.p->needs_draws) return h->fail(RMB_ERR_UNSUPPORTED, kNeedsDrawsMsg);
        m = (long long)h->h_ctr.p->heard;
        if ((rc = ensure_pairs(h, m))) return rc;
        RMB_CK(h, cudaEventRecord(h->ev[3], h->st));
        if (m > 0) {
            RMB_CK(h, launch_exclusive_scan(h->blkcnt.p, h->blkcnt.p, (int)nb, h->scan_tmp.p, nullptr, h->st)); launches += 3;
            RMB_CK(h, launch_refmedia_scatter(nd, h->dp.p, h->matrix.p, h->matrix_n, n_tx, d_rec, h->blkcnt.p, pair_arrays(h),
                                              h->node_recv.p, hd_ep, h->hd_epoch, h->st));
            ++launches;
        }
    }
    if (h->cfg.model != RMB_MODEL_SINR && h->split_counts) {
        // sharded receivers with draws: the first half of the draw stage only (flags, scan, per-packet counts)
        RMB_CK(h, launch_emit_draw_prepare(n_tx, pair_arrays(h), m, h->split_counts, h->st));
        h->split.active = true; h->split.n_tx = n_tx; h->split.m = m; h->split.rec = d_rec;
        h->split.launches = launches + 6; h->split.sweeps = sweeps;
        if (h->st == h->own_st) RMB_CK(h, cudaStreamSynchronize(h->st));
        return RMB_OK;
    }
    if (h->cfg.model != RMB_MODEL_SINR) {
        RMB_CK(h, launch_emit(h->dp.p, d_rec, n_tx, pair_arrays(h), m, draws, h->uniforms.p, nu, h->ctr.p, h->st));
        launches += draws ? 7 : 2;
    }
    if (final_sync) {
        RMB_CK(h, cudaEventRecord(h->ev[4], h->st));
        if ((rc = read_counters(h))) return rc;
    }
    return publish(h, n_tx, m, draws, launches, sweeps);
}

static int download(rmb_handle *h, rmb_result *out)
{
    const long long m = h->last_pairs;
    const int nt = h->last_ntx;
    const size_t c = (size_t)std::max<long long>(m, 1);
    const bool f64 = (h->cfg.flags & RMB_FLAG_F64_VALUES) != 0, compact = (h->cfg.flags & RMB_FLAG_COMPACT) != 0;
    if (h->slab_ready) {
        // the finish kernel packed every column for exactly m rows: ONE copy, then pointers into the pinned slab
        const SlabLayout L = slab_layout(m, nt, h->cfg.flags);
        if (h->slab_copied < (size_t)L.total) {          // no guess, or this tick's lists are longer than the guess: the rest
            size_t have = h->slab_copied;
            if (h->ho_slab.cap < (size_t)L.total) have = 0;                 // the pinned buffer moves: everything again
            RMB_CK(h, h->ho_slab.ensure((size_t)L.total));
            RMB_CK(h, cudaMemcpyAsync(h->ho_slab.p + have, h->o_slab.p + have, (size_t)L.total - have, cudaMemcpyDeviceToHost, h->st));
            RMB_CK(h, cudaStreamSynchronize(h->st));
        }
        h->slab_copied = 0;
        h->slab_pred = (size_t)L.total + (size_t)L.total / 64 + 4096;       // every byte of the guess is PCIe time
        char *b = h->ho_slab.p;
        rmb_result &r = h->last;
        r.rx = (const int32_t *)(b + L.rx); r.rssi_dbm = (const float *)(b + L.rssi); r.outcome = (const uint8_t *)(b + L.outcome);
        r.tx = compact ? nullptr : (const int32_t *)(b + L.tx); r.sinr_db = compact ? nullptr : (const float *)(b + L.sinr);
        r.tx_first = compact ? (const int32_t *)(b + L.first) : nullptr;
        r.rssi_dbm_f64 = f64 ? (const double *)(b + L.rssi64) : nullptr; r.sinr_db_f64 = f64 ? (const double *)(b + L.sinr64) : nullptr;
        r.tx_failed = (const uint8_t *)(b + L.failed);
        if (out) *out = r;
        return RMB_OK;
    }
    RMB_CK(h, h->ho_rx.ensure(c)); RMB_CK(h, h->ho_rssi.ensure(c));
    RMB_CK(h, h->ho_outcome.ensure(c)); RMB_CK(h, h->ho_failed.ensure((size_t)nt + 1));
    if (!compact) { RMB_CK(h, h->ho_tx.ensure(c)); RMB_CK(h, h->ho_sinr.ensure(c)); }
    else { RMB_CK(h, h->ho_first.ensure((size_t)nt + 2)); RMB_CK(h, h->o_first.ensure((size_t)nt + 2)); }
    if (f64) { RMB_CK(h, h->ho_rssi64.ensure(c)); RMB_CK(h, h->ho_sinr64.ensure(c)); }
    if (m > 0) {
        RMB_CK(h, cudaMemcpyAsync(h->ho_rx.p, h->o_rx.p, m * sizeof(int), cudaMemcpyDeviceToHost, h->st));
        RMB_CK(h, cudaMemcpyAsync(h->ho_rssi.p, h->o_rssi.p, m * sizeof(float), cudaMemcpyDeviceToHost, h->st));
        RMB_CK(h, cudaMemcpyAsync(h->ho_outcome.p, h->o_outcome.p, m, cudaMemcpyDeviceToHost, h->st));
        if (!compact) {
            RMB_CK(h, cudaMemcpyAsync(h->ho_tx.p, h->o_tx.p, m * sizeof(int), cudaMemcpyDeviceToHost, h->st));
            RMB_CK(h, cudaMemcpyAsync(h->ho_sinr.p, h->o_sinr.p, m * sizeof(float), cudaMemcpyDeviceToHost, h->st));
        }
        if (f64) {
            RMB_CK(h, cudaMemcpyAsync(h->ho_rssi64.p, h->o_rssi64.p, m * sizeof(double), cudaMemcpyDeviceToHost, h->st));
            RMB_CK(h, cudaMemcpyAsync(h->ho_sinr64.p, h->o_sinr64.p, m * sizeof(double), cudaMemcpyDeviceToHost, h->st));
        }
    }
    if (compact) {
        RMB_CK(h, launch_tx_first(h->o_tx.p, m, nt, h->o_first.p, h->st));
        RMB_CK(h, cudaMemcpyAsync(h->ho_first.p, h->o_first.p, ((size_t)nt + 1) * sizeof(int), cudaMemcpyDeviceToHost, h->st));
    }
    if (nt > 0) RMB_CK(h, cudaMemcpyAsync(h->ho_failed.p, h->tx_failed.p, nt, cudaMemcpyDeviceToHost, h->st));
    RMB_CK(h, cudaStreamSynchronize(h->st));
    rmb_result &r = h->last;
    r.rx = h->ho_rx.p; r.rssi_dbm = h->ho_rssi.p; r.outcome = h->ho_outcome.p;
    r.tx = compact ? nullptr : h->ho_tx.p; r.sinr_db = compact ? nullptr : h->ho_sinr.p;
    r.tx_first = compact ? h->ho_first.p : nullptr;
    r.rssi_dbm_f64 = f64 ? h->ho_rssi64.p : nullptr; r.sinr_db_f64 = f64 ? h->ho_sinr64.p : nullptr;
    r.tx_failed = h->ho_failed.p;
    if (out) *out = r;
    return RMB_OK;
}

// packets (host) -> tx records (device)
static int stage_records(rmb_handle *h, int n_tx, const int32_t *src, const double *power, const int32_t *channel,
                         rmb_tx_record *d_out, bool require_local)
{
    if (n_tx == 0) return RMB_OK;
    if (require_local)
        for (int t = 0; t < n_tx; ++t)
            if (src[t] < h->cfg.rx_base || src[t] >= h->cfg.rx_base + h->n)
                return h->fail(RMB_ERR_INVALID, "rmb_tick: source node index out of range");
    RMB_CK(h, h->pk_src.ensure(n_tx)); RMB_CK(h, h->pk_power.ensure(n_tx)); RMB_CK(h, h->pk_channel.ensure(n_tx));
    RMB_CK(h, cudaMemcpyAsync(h->pk_src.p, src, (size_t)n_tx * sizeof(int), cudaMemcpyHostToDevice, h->st));
    if (power) RMB_CK(h, cudaMemcpyAsync(h->pk_power.p, power, (size_t)n_tx * sizeof(double), cudaMemcpyHostToDevice, h->st));
    if (channel) RMB_CK(h, cudaMemcpyAsync(h->pk_channel.p, channel, (size_t)n_tx * sizeof(int), cudaMemcpyHostToDevice, h->st));
    RMB_CK(h, launch_gather_records(node_arrays(h), h->cfg.rx_base, n_tx, h->pk_src.p, power ? h->pk_power.p : nullptr,
                                    channel ? h->pk_channel.p : nullptr, d_out, h->st));
    return RMB_OK;
}

extern "C" {

int rmb_tick(rmb_handle *h, int32_t n_tx, const int32_t *src, const double *power_dbm, const int32_t *channel,
             const double *uniforms, size_t n_uniforms, rmb_result *out)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_tick before rmb_set_nodes");
    if (n_tx < 0 || (n_tx > 0 && !src) || !out) return h->fail(RMB_ERR_INVALID, "rmb_tick: bad arguments");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    RMB_CK(h, h->rec.ensure(std::max(n_tx, 1)));
    int rc = stage_records(h, n_tx, src, power_dbm, channel, h->rec.p, true);
    if (rc) return rc;
    h->want_download = true;
    rc = compute(h, n_tx, h->rec.p, uniforms, n_uniforms);
    if (rc && rc != RMB_ERR_UNIFORMS) return rc;
    int rc2 = download(h, out);
    return rc2 ? rc2 : rc;
}

int rmb_build_tx_records(rmb_handle *h, int32_t n_tx, const int32_t *src, const double *power_dbm,
                         const int32_t *channel, void *d_records)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_build_tx_records before rmb_set_nodes");
    if (n_tx < 0 || (n_tx > 0 && (!src || !d_records))) return h->fail(RMB_ERR_INVALID, "rmb_build_tx_records: bad arguments");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    int rc = stage_records(h, n_tx, src, power_dbm, channel, (rmb_tx_record *)d_records, false);
    if (rc) return rc;
    if (h->st == h->own_st) RMB_CK(h, cudaStreamSynchronize(h->st));      // the caller hands d_records to another stream (NCCL)
    return RMB_OK;
}

int rmb_tick_records(rmb_handle *h, int32_t n_tx, const void *d_records, const double *uniforms, size_t n_uniforms,
                     rmb_result *out)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_tick_records before rmb_set_nodes");
    if (n_tx < 0 || (n_tx > 0 && !d_records) || !out) return h->fail(RMB_ERR_INVALID, "rmb_tick_records: bad arguments");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    h->want_download = true;
    int rc = compute(h, n_tx, (const rmb_tx_record *)d_records, uniforms, n_uniforms);
    if (rc && rc != RMB_ERR_UNIFORMS) return rc;
    int rc2 = download(h, out);
    return rc2 ? rc2 : rc;
}

int rmb_tick_resident(rmb_handle *h, int32_t n_tx, const void *d_records, rmb_result *out)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_tick_resident before rmb_set_nodes");
    if (n_tx < 0 || (n_tx > 0 && !d_records)) return h->fail(RMB_ERR_INVALID, "rmb_tick_resident: bad arguments");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    int rc = compute(h, n_tx, (const rmb_tx_record *)d_records, nullptr, 0);
    if (out) *out = h->last;
    return rc;
}

int rmb_fetch(rmb_handle *h, rmb_result *out)
{
    if (!h || !out) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    return download(h, out);
}

// ---- stochastic reception on sharded receivers ------------------------------------------------------
int rmb_tick_draws_begin(rmb_handle *h, int32_t n_tx, const void *d_records, int32_t *d_counts_out)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_tick_draws_begin before rmb_set_nodes");
    if (n_tx < 0 || (n_tx > 0 && (!d_records || !d_counts_out))) return h->fail(RMB_ERR_INVALID, "rmb_tick_draws_begin: bad arguments");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    int32_t dummy = 0;
    h->split_counts = d_counts_out ? d_counts_out : &dummy;
    const int rc = compute(h, n_tx, (const rmb_tx_record *)d_records, nullptr, 0);
    h->split_counts = nullptr;
    return rc;
}

int rmb_tick_draws_end(rmb_handle *h, const int32_t *d_all_counts, int32_t world, int32_t rank,
                       const int32_t *d_order, int32_t n_order,
                       const double *uniforms, size_t n_uniforms, rmb_result *out)
{
    if (!h || !out) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    if (!h->split.active) return h->fail(RMB_ERR_STATE, "rmb_tick_draws_end without rmb_tick_draws_begin");
    if (world < 1 || rank < 0 || rank >= world || (h->split.n_tx > 0 && !d_all_counts) || n_order < 0 || n_order > h->split.n_tx)
        return h->fail(RMB_ERR_INVALID, "rmb_tick_draws_end: bad arguments");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    h->split.active = false;
    long long nu = 0;
    if (uniforms && n_uniforms) {
        RMB_CK(h, h->uniforms.ensure(n_uniforms));
        RMB_CK(h, cudaMemcpyAsync(h->uniforms.p, uniforms, n_uniforms * sizeof(double), cudaMemcpyHostToDevice, h->st));
        nu = (long long)n_uniforms;
    }
    PairArrays pa = pair_arrays(h);
    if (h->cfg.model == RMB_MODEL_SINR) pa.key = h->key_alt.p;      // the ordering stage left the sorted keys there
    RMB_CK(h, launch_emit_draw_finish(h->dp.p, h->split.rec, h->split.n_tx, pa, h->split.m, h->uniforms.p, nu, h->ctr.p,
                                      d_all_counts, world, rank, d_order, n_order, h->st));
    RMB_CK(h, cudaEventRecord(h->ev[4], h->st));
    int rc = read_counters(h);
    if (rc) return rc;
    rc = publish(h, h->split.n_tx, h->split.m, true, h->split.launches + 2, h->split.sweeps);
    if (rc && rc != RMB_ERR_UNIFORMS) return rc;
    const int rc2 = download(h, out);
    return rc2 ? rc2 : rc;
}

// ---- asynchronous tick ---------------------------------------------------------------------------
int rmb_tick_enqueue(rmb_handle *h, int32_t n_tx, const void *d_records)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_tick_enqueue before rmb_set_nodes");
    if (n_tx < 0 || (n_tx > 0 && !d_records)) return h->fail(RMB_ERR_INVALID, "rmb_tick_enqueue: bad arguments");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    h->async_mode = true;
    const int rc = compute(h, n_tx, (const rmb_tx_record *)d_records, nullptr, 0);
    h->async_mode = false;
    return rc;
}

int rmb_tick_finish(rmb_handle *h, rmb_result *out)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    const int rc = finish_pending(h);
    if (out) *out = h->last;
    return rc;
}

// ---- multi-GPU exchange over peer memory ------------------------------------------------------
namespace {
struct PeerBlob {                     // what rmb_peer_export writes (RMB_PEER_HANDLE_BYTES)
    uint32_t magic; int32_t world, rank, max_slots, device; int32_t pad;
    int64_t pid; uint64_t raw;        // same-process peers use the raw device pointer
    cudaIpcMemHandle_t ipc;           // other processes map this
};
static_assert(sizeof(PeerBlob) <= RMB_PEER_HANDLE_BYTES, "peer blob does not fit");
constexpr uint32_t kPeerMagic = 0x524d4250u;   // "RMBP"
}

int rmb_peer_export(rmb_handle *h, int32_t world, int32_t rank, int32_t max_slots, void *blob_out)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    if (!blob_out || world < 1 || world > kMaxPeers || rank < 0 || rank >= world || max_slots < 1)
        return h->fail(RMB_ERR_INVALID, "rmb_peer_export: bad arguments (world <= 16)");
    if (h->peer.buf) return h->fail(RMB_ERR_STATE, "rmb_peer_export: already exported");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    auto &P = h->peer;
    P.world = world; P.rank = rank; P.max_slots = max_slots;
    const size_t bytes = kPeerFlagBytes + 2 * (size_t)world * max_slots * sizeof(rmb_tx_record);
    RMB_CK(h, cudaMalloc((void **)&P.buf, bytes));
    RMB_CK(h, cudaMemset(P.buf, 0, bytes));
    RMB_CK(h, cudaMalloc((void **)&P.done, sizeof(unsigned)));
    RMB_CK(h, cudaMemset(P.done, 0, sizeof(unsigned)));
    PeerBlob b{};
    b.magic = kPeerMagic; b.world = world; b.rank = rank; b.max_slots = max_slots; b.device = h->cfg.device;
    b.pid = (int64_t)getpid(); b.raw = (uint64_t)(uintptr_t)P.buf;
    RMB_CK(h, cudaIpcGetMemHandle(&b.ipc, P.buf));
    memset(blob_out, 0, RMB_PEER_HANDLE_BYTES);
    memcpy(blob_out, &b, sizeof(b));
    return RMB_OK;
}

int rmb_peer_open(rmb_handle *h, const void *blobs)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    auto &P = h->peer;
    if (!P.buf) return h->fail(RMB_ERR_STATE, "rmb_peer_open before rmb_peer_export");
    if (!blobs) return h->fail(RMB_ERR_INVALID, "rmb_peer_open: null blobs");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    P.shares_device = false;
    for (int r = 0; r < P.world; ++r) {
        PeerBlob b;
        memcpy(&b, (const char *)blobs + (size_t)r * RMB_PEER_HANDLE_BYTES, sizeof(b));
        if (b.magic != kPeerMagic || b.world != P.world || b.rank != r || b.max_slots != P.max_slots)
            return h->fail(RMB_ERR_INVALID, "rmb_peer_open: blob does not match (world / rank order / max_slots)");
        if (r == P.rank) { P.peer[r] = P.buf; continue; }
        if (b.device == h->cfg.device) P.shares_device = true;
        if (b.pid == (int64_t)getpid()) {
            // a peer handle in this very process: use its pointer (peer access between the two devices)
            if (b.device != h->cfg.device) {
                cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return h->cuda_fail(e, "cudaDeviceEnablePeerAccess");
                cudaGetLastError();
            }
            P.peer[r] = (char *)(uintptr_t)b.raw;
        } else {
            void *ptr = nullptr;
            RMB_CK(h, cudaIpcOpenMemHandle(&ptr, b.ipc, cudaIpcMemLazyEnablePeerAccess));
            P.peer[r] = (char *)ptr; P.ipc[r] = true;
        }
    }
    return RMB_OK;
}

static int peer_push_locked(rmb_handle *h, int slots, const rmb_tx_record *d_local)
{
    auto &P = h->peer;
    for (int r = 0; r < P.world; ++r) if (!P.peer[r]) return h->fail(RMB_ERR_STATE, "rmb_peer_push before rmb_peer_open");
    if (P.pushed) return h->fail(RMB_ERR_STATE, "rmb_peer_push twice without a tick in between");
    P.slots = slots; ++P.epoch; P.pushed = true;
    PeerPush pp{};
    for (int r = 0; r < P.world; ++r) pp.peer[r] = P.peer[r];
    pp.world = P.world; pp.rank = P.rank; pp.rec_off = P.rec_off(P.epoch); pp.epoch = P.epoch; pp.done = P.done;
    RMB_CK(h, launch_peer_push(d_local, slots, pp, h->st));
    return RMB_OK;
}

int rmb_peer_push(rmb_handle *h, int32_t slots, const int32_t *src, const double *power_dbm, const int32_t *channel)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_peer_push before rmb_set_nodes");
    if (slots < 0 || slots > h->peer.max_slots || (slots > 0 && !src)) return h->fail(RMB_ERR_INVALID, "rmb_peer_push: bad arguments (slots <= max_slots)");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    RMB_CK(h, h->rec.ensure(std::max(slots, 1)));
    int rc = stage_records(h, slots, src, power_dbm, channel, h->rec.p, false);
    if (rc) return rc;
    return peer_push_locked(h, slots, h->rec.p);
}

int rmb_peer_push_records(rmb_handle *h, int32_t slots, const void *d_records)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (slots < 0 || slots > h->peer.max_slots || (slots > 0 && !d_records) || ((uintptr_t)d_records & 15))
        return h->fail(RMB_ERR_INVALID, "rmb_peer_push_records: bad arguments (slots <= max_slots, 16-byte aligned records)");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    return peer_push_locked(h, slots, (const rmb_tx_record *)d_records);
}

static int tick_peers_locked(rmb_handle *h, const double *uniforms, size_t n_uniforms)
{
    auto &P = h->peer;
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_tick_peers before rmb_set_nodes");
    if (!P.pushed) return h->fail(RMB_ERR_STATE, "rmb_tick_peers without a preceding rmb_peer_push");
    P.pushed = false;
    h->wait_flags = (const unsigned *)P.buf; h->wait_world = P.world; h->wait_epoch = P.epoch;
    const rmb_tx_record *gathered = (const rmb_tx_record *)(P.buf + P.rec_off(P.epoch));
    int rc = compute(h, P.world * P.slots, gathered, uniforms, n_uniforms);
    h->wait_flags = nullptr; h->wait_world = 0;
    return rc;
}

int rmb_tick_peers(rmb_handle *h, const double *uniforms, size_t n_uniforms, rmb_result *out)
{
    if (!h || !out) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    h->want_download = true;
    int rc = tick_peers_locked(h, uniforms, n_uniforms);
    if (rc && rc != RMB_ERR_UNIFORMS) return rc;
    int rc2 = download(h, out);
    return rc2 ? rc2 : rc;
}

int rmb_peer_tick(rmb_handle *h, int32_t slots, const int32_t *src, const double *power_dbm, const int32_t *channel, rmb_result *out)
{
    if (!h || !out) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    auto &P = h->peer;
    if (!h->have_nodes) return h->fail(RMB_ERR_STATE, "rmb_peer_tick before rmb_set_nodes");
    if (slots < 0 || slots > P.max_slots || (slots > 0 && !src)) return h->fail(RMB_ERR_INVALID, "rmb_peer_tick: bad arguments (slots <= max_slots)");
    if (P.world < 1) return h->fail(RMB_ERR_STATE, "rmb_peer_tick before rmb_peer_export");
    for (int r = 0; r < P.world; ++r) if (!P.peer[r]) return h->fail(RMB_ERR_STATE, "rmb_peer_tick before rmb_peer_open");
    if (P.pushed) return h->fail(RMB_ERR_STATE, "a pushed tick is still waiting for rmb_tick_peers");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    RMB_CK(h, h->rec.ensure(std::max(slots, 1)));
    int rc = stage_records(h, slots, src, power_dbm, channel, h->rec.p, false);
    if (rc) return rc;
    h->want_download = true;
    if (h->cfg.model == RMB_MODEL_SINR && !h->maybe_draws && !P.shares_device) {
        // the preparation kernel stores this rank's records into every peer's buffer and raises the flags itself
        // (two handles on ONE device take the separate push: their preparation kernels are cooperative launches
        // that wait for each other's records, and nothing says both are resident at once)
        P.slots = slots; ++P.epoch; P.pushed = true;
        h->fused_push = h->rec.p;
        rc = tick_peers_locked(h, nullptr, 0);
        h->fused_push = nullptr;
    } else {
        if ((rc = peer_push_locked(h, slots, h->rec.p))) return rc;
        rc = tick_peers_locked(h, nullptr, 0);
    }
    if (rc) return rc;
    return download(h, out);
}

int rmb_tick_peers_enqueue(rmb_handle *h)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    h->async_mode = true;
    const int rc = tick_peers_locked(h, nullptr, 0);
    h->async_mode = false;
    return rc;
}

int rmb_tick_peers_draws_begin(rmb_handle *h, int32_t *d_counts_out)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    if (!d_counts_out) return h->fail(RMB_ERR_INVALID, "rmb_tick_peers_draws_begin: null counts");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    h->split_counts = d_counts_out;
    const int rc = tick_peers_locked(h, nullptr, 0);
    h->split_counts = nullptr;
    return rc;
}

int rmb_peer_exchange_tick_enqueue(rmb_handle *h, int32_t slots, const void *d_records)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    auto &P = h->peer;
    if (slots < 0 || slots > P.max_slots || (slots > 0 && !d_records) || ((uintptr_t)d_records & 15))
        return h->fail(RMB_ERR_INVALID, "rmb_peer_exchange_tick_enqueue: bad arguments (slots <= max_slots, 16-byte aligned records)");
    for (int r = 0; r < P.world; ++r) if (!P.peer[r]) return h->fail(RMB_ERR_STATE, "rmb_peer_exchange_tick_enqueue before rmb_peer_open");
    if (P.world < 1) return h->fail(RMB_ERR_STATE, "rmb_peer_exchange_tick_enqueue before rmb_peer_export");
    if (P.pushed) return h->fail(RMB_ERR_STATE, "a pushed tick is still waiting for rmb_tick_peers");
    if (h->cfg.model != RMB_MODEL_SINR) return h->fail(RMB_ERR_UNSUPPORTED, "the fused exchange exists for the SINR medium only");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    P.slots = slots; ++P.epoch; P.pushed = true;
    h->async_mode = true;
    h->fused_push = (const rmb_tx_record *)d_records;
    const int rc = tick_peers_locked(h, nullptr, 0);
    h->fused_push = nullptr;
    h->async_mode = false;
    return rc;
}

int rmb_peer_barrier_enqueue(rmb_handle *h)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    auto &P = h->peer;
    if (P.world < 1) return h->fail(RMB_ERR_STATE, "rmb_peer_barrier_enqueue before rmb_peer_export");
    for (int r = 0; r < P.world; ++r) if (!P.peer[r]) return h->fail(RMB_ERR_STATE, "rmb_peer_barrier_enqueue before rmb_peer_open");
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    PeerPush pp{};
    for (int r = 0; r < P.world; ++r) pp.peer[r] = P.peer[r];
    pp.world = P.world; pp.rank = P.rank; pp.epoch = ++P.barrier_epoch;
    RMB_CK(h, launch_peer_barrier(pp, h->ctr.p, h->st));
    return RMB_OK;
}

int rmb_tick_peers_resident(rmb_handle *h, rmb_result *out)
{
    if (!h) return RMB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(h->mu);
    RMB_SETTLE(h);
    RMB_CK(h, cudaSetDevice(h->cfg.device));
    int rc = tick_peers_locked(h, nullptr, 0);
    if (out) *out = h->last;
    return rc;
}

}  // extern "C"
