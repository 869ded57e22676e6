// resolve_body.cuh -- the exact (double, strict IEEE) evaluation of a receiver's near candidates, as device
// code shared by the sweep kernel's epilogue (sweep.cu: the last tx-split block of an rx block resolves
// its receivers while the other blocks keep the XU pipe busy) and the stand-alone resolve kernel
// (resolve.cu).  Specification: see resolve.cu.
#pragma once
#include "common.cuh"
#include "kernels.h"
#include "exact_math.cuh"
#include <climits>

namespace rmb {

__device__ __forceinline__ double rx_power_dbm(const DevParams *dp, double power_dbm, double d)
{
    double dd = d < dp->d0_m ? dp->d0_m : d;
    double pl = xm::add(dp->pl0_db, xm::mul(xm::mul(10.0, dp->pl_exponent), xm::log10_strict(xm::dvd(dd, dp->d0_m))));
    return xm::sub(power_dbm, pl);
}

__device__ __forceinline__ double sinr_db(double prx, double itot, double lin)
{
    double ioth = xm::sub(itot, lin);
    return ioth > 0.0 ? xm::sub(prx, xm::mw_to_dbm(ioth)) : __longlong_as_double(0x7ff0000000000000ll);
}

// One warp per 32 receivers (work order).  The exact evaluation of the near candidates is the
// expensive part (fdlibm log10 + exp in strict double) and the candidate counts differ between
// receivers, so the warp pools its candidates: candidate i of the warp (prefix sum of the
// per-lane counts) is evaluated by lane i % 32, whoever owns it.  The per-receiver parts (far
// sum, the ordered addition of the near terms, flags) stay on the owning lane.
constexpr uint8_t kGuard = 4;         // slot flag, resolve-internal: SINR within the guard band

__device__ __forceinline__ int owner_of(int i, int P)
{
    // largest lane o with P_o <= i (P = exclusive prefix of the per-lane candidate counts)
    int lo = 0, hi = 31;
#pragma unroll
    for (int step = 0; step < 5; ++step) {
        const int mid = (lo + hi + 1) >> 1;
        const int pm = __shfl_sync(0xffffffffu, P, mid);
        if (pm <= i) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// seg[t] = first pair of packet t (exclusive scan of the per-packet heard counts), largest bucket ->
// ctr->max_bucket.  Run by the LAST block of the resolve kernel to finish (128 threads, every thread a
// contiguous range), which saves a launch on the tick's critical path.
// One block of 128 threads.  Warp v owns a quarter of the packets; pass 1 sums the quarters (loads independent
// of each other), pass 2 writes the offsets: 128 counts per warp iteration, the only serial dependence between
// iterations is one shuffle + add on the carry.
__device__ __forceinline__ void scan_tx_counts(const int *tx_cnt, int n_tx, int *seg, Counters *ctr)
{
    __shared__ int s_w[4], s_m[4];
    const int tid = threadIdx.x, lane = tid & 31, wv = tid >> 5;
    const int q = ((n_tx + 3) / 4 + 127) / 128 * 128;
    const int b = min(n_tx, wv * q), e = min(n_tx, b + q);
    int sum = 0, mx = 0;
#pragma unroll 8
    for (int i = b + lane; i < e; i += 32) { const int v = __ldcg(tx_cnt + i); sum += v; mx = max(mx, v); }
    for (int o = 16; o > 0; o >>= 1) { sum += __shfl_xor_sync(0xffffffffu, sum, o); mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o)); }
    if (lane == 0) { s_w[wv] = sum; s_m[wv] = mx; }
    __syncthreads();
    int carry = 0;
    for (int k = 0; k < wv; ++k) carry += s_w[k];
#pragma unroll 2
    for (int c0 = b; c0 < e; c0 += 128) {
        const int i0 = c0 + 4 * lane;
        int v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = i0 + k < e ? __ldcg(tx_cnt + i0 + k) : 0;
        const int t = v[0] + v[1] + v[2] + v[3];
        int incl = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
        int ex = carry + incl - t;
#pragma unroll
        for (int k = 0; k < 4; ++k) { if (i0 + k < e) seg[i0 + k] = ex; ex += v[k]; }
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (tid == 0) { seg[n_tx] = s_w[0] + s_w[1] + s_w[2] + s_w[3]; ctr->max_bucket = max(max(s_m[0], s_m[1]), max(s_m[2], s_m[3])); }
}

// One warp = up to 32 consecutive receivers in work order: lane l stands for receiver w = w_lane0 + l and owns it
// iff `active` (the active lanes are the first ones).  All 32 lanes must call and all of them evaluate candidates,
// so a warp that owns only 8 receivers gets through its pooled candidates four times sooner -- what the sweep's
// epilogue wants at the end of the launch, where latency and not throughput is what counts.
// fresh: the candidate slots, counters and far sums were written by OTHER blocks of the running kernel
// (fused epilogue) and are read past the L1.
template <bool FRESH>
__device__ __forceinline__ void resolve_receiver(int w, bool active, const WorkArrays &wk, const DevParams *dp, const TxTiles &tx,
                                                 const rmb_tx_record *rec, const SlotArrays &sl, const PairRecs &pr, Counters *ctr,
                                                 unsigned long long *dbg = nullptr)
{
    const int lane = threadIdx.x & 31;
    const int K = sl.K;
    if (dbg && threadIdx.x == 0) dbg[0] = global_timer_ns();
    const int p = active ? wk.perm[w] : -1;
    const size_t base = (size_t)w * K;
    int c = 0;
    double itot = dp->n0_mw;
    if (p >= 0) {
        // (the first eight slots are fetched before anyone knows how many are valid: independent of the count and
        // of the far sums, so the three round trips overlap)
        int sv[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) sv[k] = k < K ? __ldcg(&sl.slots[base + k]) : INT_MAX;
        const int found = FRESH ? __ldcg(&sl.cnt[w]) : sl.cnt[w];
        if (found > K) atomicMax(&ctr->max_near, found);      // slot overflow: the host grows K and redoes the sweep
        c = min(found, K);
        // far field: the sweep's per-split partial sums, added in split order (deterministic)
        // (twenty loads in flight at a time; a missing split adds +0.0, which changes nothing)
        double far = 0.0;
        const int nsplit = w < sl.w_taper ? sl.splits : sl.splits2;
        for (int s0 = 0; s0 < nsplit; s0 += 20) {      // (19 rows, or 38 in the launch's final wave: one or two round trips)
            float v[20];
#pragma unroll
            for (int k = 0; k < 20; ++k) v[k] = s0 + k < nsplit ? __ldcg(&sl.far_part[(size_t)(s0 + k) * wk.n_work + w]) : 0.f;
#pragma unroll
            for (int k = 0; k < 20; ++k) far = xm::add(far, (double)v[k]);
        }
        itot = xm::add(itot, far);
        // the sweep appends near candidates in arbitrary order: sort by bucket position (= arrival order)
        if (c > 1 && c <= 8) {
            // the common case in registers: independent loads, odd-even transposition network, one write back
            int v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = k < c ? sv[k] : INT_MAX;
#pragma unroll
            for (int round = 0; round < 8; ++round)
#pragma unroll
                for (int k = round & 1; k + 1 < 8; k += 2) { const int lo = min(v[k], v[k + 1]), hi = max(v[k], v[k + 1]); v[k] = lo; v[k + 1] = hi; }
#pragma unroll
            for (int k = 0; k < 8; ++k) if (k < c) sl.slots[base + k] = v[k];
        } else
        for (int a = 1; a < c; ++a) {
            const int v = __ldcg(&sl.slots[base + a]);          // slots are read past the L1 everywhere (other blocks wrote them)
            int b = a - 1;
            while (b >= 0 && __ldcg(&sl.slots[base + b]) > v) { sl.slots[base + b + 1] = __ldcg(&sl.slots[base + b]); --b; }
            sl.slots[base + b + 1] = v;
        }
    }
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    const int P = incl - c;
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    const size_t wbase = (size_t)(w - lane) * K;
    __syncwarp();
    if (dbg && threadIdx.x == 0) dbg[1] = global_timer_ns();

    // pass 1 (pooled): link budget of every candidate
    for (int i0 = 0; i0 < total; i0 += 32) {
        const int i = i0 + lane;
        const int o = owner_of(i, P);
        const int po = __shfl_sync(0xffffffffu, p, o);
        const int so = i - __shfl_sync(0xffffffffu, P, o);
        if (i < total) {
            const size_t q = wbase + (size_t)o * K + so;
            const rmb_tx_record r = rec[tx.arrival[__ldcg(&sl.slots[q])]];
            uint8_t f = 0;
            double prx = 0.0, lin = 0.0;
            if (r.src != dp->rx_base + po) {
                const int wo = w - lane + o;
                const double d = xm::distance(r.x, r.y, r.z, wk.x64[wo], wk.y64[wo], wk.z64[wo]);
                prx = rx_power_dbm(dp, r.power_dbm, d);
                lin = xm::dbm_to_mw(prx);
                if (prx >= dp->sensitivity_dbm && wk.rxp[wo] > 0.0) f = kHeard;
                if (wk.tx_epoch && wk.tx_epoch[po] == wk.hd_epoch) f = 0;     // half duplex: the receiver is on the air itself
            }
            sl.prx[q] = prx; sl.lin[q] = lin; sl.flags[q] = f;
        }
    }
    __syncwarp();

    if (dbg && threadIdx.x == 0) dbg[2] = global_timer_ns();
    // per receiver: total in-band power, near terms added in arrival order
    int nh = 0;
    for (int s = 0; s < c; ++s) {
        itot = xm::add(itot, sl.lin[base + s]);       // the source's own packet contributes lin = 0
        nh += sl.flags[base + s] & kHeard;
    }

    // the warp's heard pairs take one contiguous range of the record list (one atomic per warp);
    // a receiver's records are contiguous in slot order, starting at first[w]
    int hincl = nh;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, hincl, o); if (lane >= o) hincl += y; }
    const int wtotal = __shfl_sync(0xffffffffu, hincl, 31);
    unsigned long long rbase = 0;
    if (lane == 0 && wtotal) rbase = atomicAdd(&ctr->heard, (unsigned long long)wtotal);
    rbase = __shfl_sync(0xffffffffu, rbase, 0);
    if (active) sl.first[w] = (int)rbase + hincl - nh;

    if (dbg && threadIdx.x == 0) { dbg[3] = global_timer_ns(); dbg[5] = (unsigned long long)total; }
    // pass 2 (pooled): SINR of every heard candidate, pair record
    int running = 0;
    for (int i0 = 0; i0 < total; i0 += 32) {
        const int i = i0 + lane;
        const int o = owner_of(i, P);
        const int po = __shfl_sync(0xffffffffu, p, o);
        const int so = i - __shfl_sync(0xffffffffu, P, o);
        const double it = __shfl_sync(0xffffffffu, itot, o);
        size_t q = 0;
        uint8_t f = 0;
        if (i < total) { q = wbase + (size_t)o * K + so; f = sl.flags[q]; }
        const unsigned hb = __ballot_sync(0xffffffffu, (f & kHeard) != 0);
        if (f & kHeard) {
            const double prx = sl.prx[q];
            const double sinr = sinr_db(prx, it, sl.lin[q]);
            if (sinr >= dp->sinr_threshold_db) f |= kDecodable;
            if (fabs(sinr - dp->sinr_threshold_db) < dp->sinr_guard_db) f |= kGuard;
            sl.sinr[q] = sinr; sl.flags[q] = f;
            const long long ri = (long long)rbase + running + __popc(hb & ((1u << lane) - 1));
            const int t = tx.arrival[__ldcg(&sl.slots[q])];
            const int rank = atomicAdd(&pr.tx_cnt[t], 1);
            if (ri < pr.cap) {
                pr.key[ri] = ((unsigned long long)t << pr.key_shift) | (unsigned)(dp->rx_base + po);
                pr.prx[ri] = prx; pr.sinr[ri] = sinr; pr.flags[ri] = f & (kHeard | kDecodable); pr.rank[ri] = rank;
            }
        }
        running += __popc(hb);
    }
    __syncwarp();

    if (p >= 0) {
        bool guard = false;
        for (int s = 0; s < c; ++s) guard |= (sl.flags[base + s] & kGuard) != 0;
        if (wk.tx_epoch && wk.tx_epoch[p] == wk.hd_epoch) itot = dp->n0_mw;     // half duplex: a node on the air reports the noise floor
        sl.itot[w] = itot;
        if (guard && sl.cnt[w] <= K) {
            const int e = atomicAdd(&ctr->n_recheck, 1);
            sl.recheck[e] = w; sl.recheck_cnt[e] = 0;
        }
    }
    if (active) sl.hcnt[w] = nh;
}

// ---- guard-band recheck helpers (resolve.cu, finish.cu) ----------------------------------------
__device__ __forceinline__ double exact_lin(const DevParams *dp, const rmb_tx_record &r, int self, double X, double Y, double Z)
{
    // itot + 0.0 == itot: skipping the source is the same as adding zero
    if (r.src == self) return 0.0;
    return xm::dbm_to_mw(rx_power_dbm(dp, r.power_dbm, xm::distance(r.x, r.y, r.z, X, Y, Z)));
}

// writes the receiver's total and the SINR / flags of its heard slots and pair records; returns
// true if some decision is closer to the threshold than tol_rel (relative uncertainty of itot) allows
__device__ __forceinline__ bool apply_itot(const DevParams *dp, SlotArrays sl, PairRecs pr, int w, double itot, double tol_rel)
{
    bool close = false;
    const int c = min(sl.cnt[w], sl.K);
    const size_t base = (size_t)w * sl.K;
    sl.itot[w] = itot;
    int h = 0;                                   // the receiver's records follow first[w] in slot order (heard slots only)
    for (int s = 0; s < c; ++s) {
        if (!(sl.flags[base + s] & kHeard)) continue;
        const double lin = sl.lin[base + s];
        const double sinr = sinr_db(sl.prx[base + s], itot, lin);
        const double ioth = itot - lin;
        const double tol_db = ioth > 0.0 ? 4.342944819032518 * tol_rel * (itot / ioth) + 1e-12 : 0.0;
        if (fabs(sinr - dp->sinr_threshold_db) <= tol_db) close = true;
        const uint8_t f = kHeard | (sinr >= dp->sinr_threshold_db ? kDecodable : 0);
        sl.sinr[base + s] = sinr; sl.flags[base + s] = f;
        const long long ri = (long long)sl.first[w] + h;
        if (ri < pr.cap) { pr.sinr[ri] = sinr; pr.flags[ri] = f; }
        ++h;
    }
    return close;
}


}  // namespace rmb
