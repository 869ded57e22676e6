// radio-sim_b200/csrc/finish.cu -- everything behind the sweep in ONE cooperative kernel.
//
// The deterministic-reception tick used to end in five launches on two streams (guard-band tree sum and
// sequential tier on a side stream; bucket placement, per-packet sort and emission on the main one).  On
// a 1/8 shard those are ~45 us of mostly launch latency and stream hand-overs behind a 370 us sweep, so
// they are one kernel here, its phases separated by grid barriers:
//   1  every pair record goes to its packet's bucket (items = rx << 32 | record), and -- alongside -- the
//      guard-band receivers' exact interference sums are formed as tree partials, one (receiver, part) per block;
//      the block that completes a receiver adds its partials in part order and rewrites the decisions of its pair
//      records; a receiver still within 2 n u of the threshold is queued for tier 2;
//   2  tier 2 (practically only constructed cases; skipped, with its barrier, when the queue is empty): the sum
//      redone strictly in arrival order;
//   3  every packet's bucket is ranked by rx (a warp for buckets <= 64 pairs, a block above) and its output
//      rows are written straight from the ranked items: (packet arrival, node index) order, the order in which
//      UDGMRadioMedium.java:78-110 walks `nodes` for each packet of the tick.
// The barrier counter is zeroed by the tx preparation kernel of the same tick, so the targets are k * gridDim.x.
// The last block to finish stores the tick's counters into pinned host memory.
// The stochastic rule and the radix-sort fallback (a packet heard by > kBucketSortMax receivers) keep the
// separate kernels of resolve.cu / emit.cu.
#include "kernels.h"
#include "resolve_body.cuh"

namespace rmb {

constexpr int kFinThreads = 256;
constexpr int kFinWarps = kFinThreads / 32;
constexpr int kWarpBucket = 64;                 // buckets up to this size are ranked by one warp
constexpr int kCountRankMax = 512;              // ... up to this size by counting in a block, above by a bitonic sort
constexpr int kFinTreeMaxParts = 64;
constexpr int kFinSeqChunk = (kFinThreads - 32) * 2;

struct EmitCount { int delivered, interfered; };

__device__ __forceinline__ void emit_row(const PairRecs &pr, const PairArrays &pa, int t, bool failed,
                                         unsigned long long item, long long pos, EmitCount &n)
{
    const unsigned ri = (unsigned)item;
    const double rssi = __ldcg(&pr.prx[ri]), sinr = __ldcg(&pr.sinr[ri]);
    const bool cand = (__ldcg(&pr.flags[ri]) & kDecodable) != 0;
    const int outcome = (failed || !cand) ? RMB_INTERFERED : RMB_DELIVERED;
    if (pa.o_tx) { pa.o_tx[pos] = t; pa.o_sinr[pos] = (float)sinr; }       // (the compact slab has neither column)
    pa.o_rx[pos] = (int)(item >> 32);
    pa.o_rssi[pos] = (float)rssi;
    if (pa.o_rssi64) { pa.o_rssi64[pos] = rssi; pa.o_sinr64[pos] = sinr; }
    pa.o_outcome[pos] = (uint8_t)outcome;
    n.delivered += outcome == RMB_DELIVERED; n.interfered += outcome != RMB_DELIVERED;
}

// apply_itot (resolve_body.cuh) by a whole warp: a lane per candidate slot, the heard ordinal from a ballot prefix.
// Same arithmetic per slot; returns (to every lane) whether some decision is within tol_rel of the threshold.
__device__ __forceinline__ bool apply_itot_warp(const DevParams *dp, SlotArrays sl, PairRecs pr, int w, double itot, double tol_rel, int lane)
{
    const int c = min(sl.cnt[w], sl.K);
    const size_t base = (size_t)w * sl.K;
    if (lane == 0) sl.itot[w] = itot;
    const long long first = (long long)sl.first[w];
    bool close = false;
    int running = 0;
    for (int s0 = 0; s0 < c; s0 += 32) {
        const int s = s0 + lane;
        const bool heard = s < c && (sl.flags[base + s] & kHeard) != 0;
        const unsigned hb = __ballot_sync(0xffffffffu, heard);
        if (heard) {
            const double lin = sl.lin[base + s];
            const double sinr = sinr_db(sl.prx[base + s], itot, lin);
            const double ioth = itot - lin;
            const double tol_db = ioth > 0.0 ? 4.342944819032518 * tol_rel * (itot / ioth) + 1e-12 : 0.0;
            if (fabs(sinr - dp->sinr_threshold_db) <= tol_db) close = true;
            const uint8_t f = kHeard | (sinr >= dp->sinr_threshold_db ? kDecodable : 0);
            sl.sinr[base + s] = sinr; sl.flags[base + s] = f;
            const long long ri = first + running + __popc(hb & ((1u << lane) - 1u));
            if (ri < pr.cap) { pr.sinr[ri] = sinr; pr.flags[ri] = f; }
        }
        running += __popc(hb);
    }
    return __any_sync(0xffffffffu, close);
}

// one packet's bucket (nb <= 32 * PER pairs) ranked by rx by one warp -- counting: the rx of one packet's pairs are
// distinct -- and its rows written
template <int PER>
__device__ __forceinline__ void warp_bucket(const PairRecs &pr, const PairArrays &pa, const unsigned long long *items,
                                            unsigned long long *sw, int t, int b0, int nb, int lane, EmitCount &n)
{
    unsigned long long v[PER];
    int r[PER];
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        const int i = lane + 32 * k;
        v[k] = i < nb ? __ldcg(&items[b0 + i]) : ~0ull;
        sw[i] = v[k];
        r[k] = 0;
    }
    __syncwarp();
    // (the rx index -- the items' high word -- alone orders a bucket: 32-bit compares)
    const unsigned *sk = reinterpret_cast<const unsigned *>(sw) + 1;
    unsigned vk[PER];
#pragma unroll
    for (int k = 0; k < PER; ++k) vk[k] = (unsigned)(v[k] >> 32);
#pragma unroll 4
    for (int j = 0; j < nb; ++j) {
        const unsigned x = sk[2 * j];
#pragma unroll
        for (int k = 0; k < PER; ++k) r[k] += x < vk[k];
    }
    const bool failed = pa.tx_failed[t] != 0;
#pragma unroll
    for (int k = 0; k < PER; ++k)
        if (lane + 32 * k < nb) emit_row(pr, pa, t, failed, v[k], (long long)b0 + r[k], n);
    __syncwarp();
}

// seg[t] = first pair of packet t (exclusive scan of the per-packet heard counts) for n <= kFinScanMax packets, by ONE
// block into shared memory; returns the largest count.  Every block of the finish kernel does this for itself -- the
// sweep used to leave the scan to the one block that resolved last, 10 us at the very end of its launch with the rest
// of the GPU idle (profiles/r02_trace8.txt).  Warp w owns the contiguous segment [w * seg_len, (w + 1) * seg_len):
// 32 counts per step straight from global memory (coalesced, the next steps' loads already in flight), a shuffle scan,
// the running carry in a register -- no barrier inside the loop, no bank conflict.  One barrier, then the warps'
// offsets are added in place (coalesced again).
__device__ __forceinline__ int block_scan_counts(const int *cnt, int n, int *s_seg, int *s_warp /* [2 * kFinWarps] */)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int seg_len = ((n + kFinWarps - 1) / kFinWarps + 127) / 128 * 128;
    const int b = min(n, warp * seg_len), e = min(n, b + seg_len);
    int carry = 0, mx = 0;
    // four counts per lane and step (one 16-byte load), five steps' loads in flight: the loop is two L2 round trips
    // long instead of ten
#pragma unroll 5
    for (int i0 = b; i0 < e; i0 += 128) {
        const int i = i0 + 4 * lane;
        int4 v = make_int4(0, 0, 0, 0);
        if (i + 3 < e) v = __ldcg(reinterpret_cast<const int4 *>(cnt + i));
        else {
            if (i < e) v.x = __ldcg(cnt + i);
            if (i + 1 < e) v.y = __ldcg(cnt + i + 1);
            if (i + 2 < e) v.z = __ldcg(cnt + i + 2);
        }
        mx = max(mx, max(max(v.x, v.y), max(v.z, v.w)));
        const int t = v.x + v.y + v.z + v.w;
        int incl = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
        const int ex = carry + incl - t;
        if (i + 3 < e) *reinterpret_cast<int4 *>(s_seg + i) = make_int4(ex, ex + v.x, ex + v.x + v.y, ex + v.x + v.y + v.z);
        else {
            if (i < e) s_seg[i] = ex;
            if (i + 1 < e) s_seg[i + 1] = ex + v.x;
            if (i + 2 < e) s_seg[i + 2] = ex + v.x + v.y;
        }
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0) { s_warp[warp] = carry; s_warp[kFinWarps + warp] = mx; }
    __syncthreads();
    int off = 0, total = 0, gmx = 0;
#pragma unroll
    for (int k = 0; k < kFinWarps; ++k) { const int v = s_warp[k]; if (k < warp) off += v; total += v; gmx = max(gmx, s_warp[kFinWarps + k]); }
    if (off)
        for (int i = b + lane; i < e; i += 32) s_seg[i] += off;
    if (tid == 0) s_seg[n] = total;
    __syncthreads();
    return gmx;
}

__global__ void __launch_bounds__(kFinThreads, 4)
finish_kernel(WorkArrays wk, const DevParams *dp, Groups groups, TxTiles tx, const rmb_tx_record *rec, SlotArrays sl,
              PairRecs pr, PairArrays pa_in, int n_tx, int *seg_global, unsigned long long *items, double *part,
              Counters *ctr, Counters *host_ctr, unsigned *bar, int scan_here)
{
    PairArrays pa = pa_in;
    // one dynamic buffer, two lives: the segment offsets of all packets while the records go to their buckets
    // (scan_here: [n_tx + 1] ints), the ranking scratch of phase 3 behind the grid barrier ([kBucketSortMax] items;
    // the offsets are read from global memory there -- block 0 stored them in front of the barrier)
    extern __shared__ unsigned long long s_items[];
    int *const s_seg_dyn = reinterpret_cast<int *>(s_items);
    __shared__ double s_part[kFinWarps];
    __shared__ int s_last, s_scanw[2 * kFinWarps];
    double *s_lin = reinterpret_cast<double *>(s_items);            // tier 2: [2][kFinSeqChunk]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // launched with programmatic stream serialization: everything above ran while the sweep's last blocks were
    // still at work; its results (and the counters) may be read from here on
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0 && blockIdx.x == 0) atomicMax(&ctr->t_finish_begin_inv, ~global_timer_ns());
    const long long m_all = (long long)ctr->heard;
    const bool stand_down = m_all > pr.cap;                         // the host grows the record lists and redoes the tick
    const long long m = stand_down ? 0 : m_all;
    // the output columns: packed for exactly m rows in the slab (one copy takes them to the host)
    SlabLayout L{};
    if (pa.slab) {
        L = slab_layout(m, n_tx, pa.slab_flags);
        pa.o_rx = reinterpret_cast<int *>(pa.slab + L.rx); pa.o_rssi = reinterpret_cast<float *>(pa.slab + L.rssi);
        pa.o_tx = L.tx >= 0 ? reinterpret_cast<int *>(pa.slab + L.tx) : nullptr;
        pa.o_sinr = L.sinr >= 0 ? reinterpret_cast<float *>(pa.slab + L.sinr) : nullptr;
        pa.o_rssi64 = L.rssi64 >= 0 ? reinterpret_cast<double *>(pa.slab + L.rssi64) : nullptr;
        pa.o_sinr64 = L.sinr64 >= 0 ? reinterpret_cast<double *>(pa.slab + L.sinr64) : nullptr;
        pa.o_outcome = reinterpret_cast<uint8_t *>(pa.slab + L.outcome);
    }
    // ---- 0: segment offsets.  scan_here: every block scans the per-packet counts for itself (block 0 also stores
    //         them for the host side / later kernels); otherwise the sweep's last resolving block did it
    const int *seg = seg_global;
    int max_bucket;
    if (scan_here) {
        max_bucket = block_scan_counts(pr.tx_cnt, n_tx, s_seg_dyn, s_scanw);
        seg = s_seg_dyn;
        if (blockIdx.x == 0) {
            for (int i = threadIdx.x; i <= n_tx; i += kFinThreads) seg_global[i] = s_seg_dyn[i];
            if (threadIdx.x == 0) ctr->max_bucket = max_bucket;
        }
    } else {
        max_bucket = __ldcg(&ctr->max_bucket);
    }

    // ---- 1a: records -> buckets, at the rank their atomic drew -----------------------------------
    for (long long i = (long long)blockIdx.x * kFinThreads + threadIdx.x; i < m; i += (long long)gridDim.x * kFinThreads) {
        const unsigned long long k = pr.key[i];
        const int t = (int)(k >> pr.key_shift);
        const unsigned long long rx = k & ((1ull << pr.key_shift) - 1);
        items[seg[t] + pr.rank[i]] = (rx << 32) | (unsigned long long)i;
    }
    // ---- 1b: guard-band receivers: exact tree partials, one (receiver, part) per block; the block that stores a
    //          receiver's last partial adds them up in part order and rewrites the receiver's decisions --------
    const int n1 = stand_down ? 0 : ctr->n_recheck;
    const int Q = n1 > 0 ? max(1, min(kFinTreeMaxParts, (int)gridDim.x / n1)) : 1;
    for (long long u = blockIdx.x; u < (long long)n1 * Q; u += gridDim.x) {
        const int b = (int)(u / Q), q = (int)(u % Q);
        const int w = sl.recheck[b];
        const int p = wk.perm[w];
        int g = 0;
        while (g + 1 < groups.ngroups && w >= groups.w_begin[g + 1]) ++g;
        const int j0 = tx.goff[g], nj = tx.gcount[g];
        const int len = (nj + Q - 1) / Q, i0 = min(nj, q * len), i1 = min(nj, i0 + len);
        const double X = wk.x64[w], Y = wk.y64[w], Z = wk.z64[w];
        const int self = dp->rx_base + p;
        double acc = 0.0;
        for (int i = i0 + threadIdx.x; i < i1; i += kFinThreads) acc += exact_lin(dp, rec[tx.arrival[j0 + i]], self, X, Y, Z);
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) s_part[warp] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double v = 0.0;
            for (int i = 0; i < kFinWarps; ++i) v += s_part[i];
            part[u] = v;
            __threadfence();
            s_last = atomicAdd(&sl.recheck_cnt[b], 1) == Q - 1;
        }
        __syncthreads();
        if (s_last && warp == 0) {
            __threadfence();
            double itot = dp->n0_mw;
            for (int k = 0; k < Q; ++k) itot += __ldcg(&part[(long long)b * Q + k]);
            const double tol_rel = 2.0 * (double)(nj + 64) * 1.1102230246251565e-16;
            const bool close = apply_itot_warp(dp, sl, pr, w, itot, tol_rel, lane);
            if (close && lane == 0) sl.recheck2[atomicAdd(&ctr->n_recheck2, 1)] = w;
        }
        __syncthreads();
    }
    grid_barrier(bar, gridDim.x);
    seg = seg_global;                                   // the shared copy is about to be overwritten by the ranking scratch

    // ---- 2: tier 2 (rare), the addition in arrival order: one adder thread, the other warps evaluate ahead ----
    const int n2 = __ldcg(&ctr->n_recheck2);
    if (n2 > 0) {
        for (int b = blockIdx.x; b < n2; b += gridDim.x) {
            const int w = __ldcg(&sl.recheck2[b]);
            const int p = wk.perm[w];
            int g = 0;
            while (g + 1 < groups.ngroups && w >= groups.w_begin[g + 1]) ++g;
            const int j0 = tx.goff[g], nj = tx.gcount[g];
            const double X = wk.x64[w], Y = wk.y64[w], Z = wk.z64[w];
            const int self = dp->rx_base + p;
            const int nchunks = (nj + kFinSeqChunk - 1) / kFinSeqChunk;
            double itot = dp->n0_mw;
            for (int k = 0; k <= nchunks; ++k) {
                if (threadIdx.x >= 32) {
                    if (k < nchunks) {
                        const int c0 = k * kFinSeqChunk, nc = min(kFinSeqChunk, nj - c0);
                        for (int i = threadIdx.x - 32; i < nc; i += kFinThreads - 32)
                            s_lin[(k & 1) * kFinSeqChunk + i] = exact_lin(dp, rec[tx.arrival[j0 + c0 + i]], self, X, Y, Z);
                    }
                } else if (threadIdx.x == 0 && k > 0) {
                    const int c0 = (k - 1) * kFinSeqChunk, nc = min(kFinSeqChunk, nj - c0);
                    const double *v = s_lin + ((k - 1) & 1) * kFinSeqChunk;
#pragma unroll 8
                    for (int i = 0; i < nc; ++i) itot = xm::add(itot, v[i]);
                }
                __syncthreads();
            }
            if (threadIdx.x == 0) apply_itot(dp, sl, pr, w, itot, 0.0);
            __syncthreads();
        }
        grid_barrier(bar, 2u * gridDim.x);      // n2 was read behind the first barrier: every block takes this branch or none
    }

    // ---- 3: rank every packet's bucket by rx, write its rows -----------------------------------------
    if (pa.slab) {
        // first row of every packet (== the segment offsets: rows are written at seg[t] + rank) and the tx-failure flags
        int *first = reinterpret_cast<int *>(pa.slab + L.first);
        uint8_t *failed = reinterpret_cast<uint8_t *>(pa.slab + L.failed);
        for (int i = blockIdx.x * kFinThreads + threadIdx.x; i <= n_tx; i += gridDim.x * kFinThreads) {
            first[i] = m > 0 ? seg[i] : 0;
            if (i < n_tx) failed[i] = pa.tx_failed[i];
        }
    }
    EmitCount n{0, 0};
    // (m == 0: nothing was heard -- with no enabled receiver at all the sweep did not even run and `seg` is stale)
    if (m > 0 && max_bucket <= kBucketSortMax) {
        // buckets of up to kWarpBucket pairs: a warp each
        unsigned long long *sw = s_items + warp * kWarpBucket;
        for (int t = blockIdx.x * kFinWarps + warp; t < n_tx; t += gridDim.x * kFinWarps) {
            const int b0 = seg[t], nb = seg[t + 1] - b0;
            if (nb <= 0 || nb > kWarpBucket) continue;
            if (nb <= 32) warp_bucket<1>(pr, pa, items, sw, t, b0, nb, lane, n);
            else warp_bucket<2>(pr, pa, items, sw, t, b0, nb, lane, n);
        }
        __syncthreads();
        // larger buckets: a block each -- ranks by counting up to kCountRankMax pairs, a bitonic sort above
        for (int t = blockIdx.x; t < n_tx; t += gridDim.x) {
            const int b0 = seg[t], nb = seg[t + 1] - b0;
            if (nb <= kWarpBucket) continue;
            const bool failed = pa.tx_failed[t] != 0;
            if (nb <= kCountRankMax) {
                for (int i = threadIdx.x; i < nb; i += kFinThreads) s_items[i] = __ldcg(&items[b0 + i]);
                __syncthreads();
                const unsigned *sk = reinterpret_cast<const unsigned *>(s_items) + 1;     // high words: the rx indices
                for (int i = threadIdx.x; i < nb; i += kFinThreads) {
                    const unsigned long long mine = s_items[i];
                    const unsigned mk = (unsigned)(mine >> 32);
                    int rank = 0;
#pragma unroll 8
                    for (int j = 0; j < nb; ++j) rank += sk[2 * j] < mk;
                    emit_row(pr, pa, t, failed, mine, (long long)b0 + rank, n);
                }
            } else {
                int np2 = 1024;
                while (np2 < nb) np2 <<= 1;
                for (int i = threadIdx.x; i < np2; i += kFinThreads) s_items[i] = i < nb ? __ldcg(&items[b0 + i]) : ~0ull;
                __syncthreads();
                for (int k = 2; k <= np2; k <<= 1)
                    for (int j = k >> 1; j > 0; j >>= 1) {
                        for (int i = threadIdx.x; i < np2; i += kFinThreads) {
                            const int l = i ^ j;
                            if (l > i) {
                                const unsigned long long a = s_items[i], b = s_items[l];
                                if ((a > b) == ((i & k) == 0)) { s_items[i] = b; s_items[l] = a; }
                            }
                        }
                        __syncthreads();
                    }
                for (int i = threadIdx.x; i < nb; i += kFinThreads) emit_row(pr, pa, t, failed, s_items[i], (long long)b0 + i, n);
            }
            __syncthreads();
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        n.delivered += __shfl_xor_sync(0xffffffffu, n.delivered, o);
        n.interfered += __shfl_xor_sync(0xffffffffu, n.interfered, o);
    }
    if (lane == 0) {
        if (n.delivered) atomicAdd(&ctr->n_delivered, n.delivered);
        if (n.interfered) atomicAdd(&ctr->n_interfered, n.interfered);
    }
    if (host_ctr) {
        __syncthreads();
        if (threadIdx.x == 0) { __threadfence(); s_last = atomicAdd(&ctr->emit_done, 1u) == gridDim.x - 1; }
        __syncthreads();
        if (s_last && threadIdx.x < sizeof(Counters) / sizeof(unsigned)) {
            __threadfence();
            reinterpret_cast<volatile unsigned *>(host_ctr)[threadIdx.x] = __ldcg(reinterpret_cast<const unsigned *>(ctr) + threadIdx.x);
            __threadfence_system();
        }
    }
}

// co-resident blocks of the kernel on the current device for a given dynamic shared memory size
constexpr int kFinScanMax = 16 * 1024 - 1;         // packets whose offsets a block scans for itself (64 KB of shared memory)
static size_t finish_smem(int n_tx, bool scan_here)
{
    const size_t items = (size_t)kBucketSortMax * sizeof(unsigned long long);
    return scan_here ? std::max(items, ((size_t)n_tx + 1) * sizeof(int)) : items;
}
bool finish_scans(int n_tx) { return n_tx <= kFinScanMax; }

int finish_grid(int n_tx)
{
    static int cache[64][2] = {};
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    const int slot = (dev >= 0 && dev < 64) ? dev : 0;
    if (!attr_set[slot]) {                              // per device: more than 48 KB of dynamic shared memory is opt-in
        cudaFuncSetAttribute(finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (kFinScanMax + 1) * (int)sizeof(int));
        attr_set[slot] = true;
    }
    // occupancy depends on the dynamic size only through a few steps: cache the two common cases (none / <= 48 KB),
    // query the others
    const bool scan = finish_scans(n_tx);
    const size_t smem = finish_smem(n_tx, scan);
    const int kind = smem <= 16 * 1024 ? 0 : (smem <= 40 * 1024 + 4 ? 1 : -1);
    if (kind >= 0 && cache[slot][kind]) return cache[slot][kind];
    int per_sm = 1;
    const size_t q = kind == 0 ? 16 * 1024 : kind == 1 ? 40 * 1024 + 4 : smem;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, finish_kernel, kFinThreads, q) != cudaSuccess || per_sm < 1) {
        per_sm = 1; cudaGetLastError();
    }
    const int g = device_sms() * std::min(per_sm, 4);
    if (kind >= 0) cache[slot][kind] = g;
    return g;
}

cudaError_t launch_finish(const WorkArrays &wk, const DevParams *dp, Groups groups, TxTiles tx, const rmb_tx_record *rec,
                          SlotArrays sl, PairRecs pr, PairArrays pa, int n_tx, int *seg, unsigned long long *items,
                          double *part, Counters *ctr, Counters *host_ctr, int grid, unsigned *bar, bool overlap, cudaStream_t st)
{
    if (n_tx <= 0) return cudaSuccess;
    WorkArrays a_wk = wk; Groups a_groups = groups;
    int scan_here = finish_scans(n_tx) ? 1 : 0;
    void *args[] = {(void *)&a_wk, (void *)&dp, (void *)&a_groups, (void *)&tx, (void *)&rec, (void *)&sl, (void *)&pr,
                    (void *)&pa, (void *)&n_tx, (void *)&seg, (void *)&items, (void *)&part, (void *)&ctr, (void *)&host_ctr,
                    (void *)&bar, (void *)&scan_here};
    // (finish_grid(n_tx) was computed for the same dynamic size; it also opted the kernel in to > 48 KB)
    return launch_grid_synced_ex((const void *)&finish_kernel, dim3((unsigned)grid), dim3(kFinThreads), args,
                                 finish_smem(n_tx, scan_here != 0), overlap, st);
}

}  // namespace rmb
