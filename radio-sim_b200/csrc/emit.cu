// emit.cu -- K3: order-preserving compaction of the heard pairs, (tx arrival, rx index)
// ordering, the reference's stochastic reception rule with host-supplied uniforms, outputs.
//
// The result list reproduces the Java call order of Simulator.generateReceptionEvents
// (UDGMRadioMedium.java:99-114, Simulator.java:321-335): packets in arrival order, receivers
// in node-array order.  Compaction is warp ballot / shuffle prefix sums (no atomic appends).
// The uniforms stand in for simulator.getRandom().nextDouble() and are consumed in the
// reference's draw order (UDGMRadioMedium.java:88-92,104-111): per packet first the tx draw
// (only if 0 < p_tx < 1), then one draw per receiver that needs it, in node order.
#include "common.cuh"
#include "kernels.h"
#include <cub/device/device_radix_sort.cuh>

namespace rmb {

// ---- device-wide exclusive scan of int32 (1024 elements per block) ------------------------
__device__ __forceinline__ int block_exclusive_scan_1024(int v, int *s_warp, int &block_total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int x = s_warp[lane], xi = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, xi, o); if (lane >= o) xi += y; }
        s_warp[lane] = xi - x;
        if (lane == 31) s_warp[32] = xi;
    }
    __syncthreads();
    block_total = s_warp[32];
    return incl - v + s_warp[wid];
}

__global__ void __launch_bounds__(1024) scan_sums_kernel(const int *in, int n, int *sums)
{
    __shared__ int s_warp[33];
    int i = blockIdx.x * 1024 + threadIdx.x;
    int total;
    block_exclusive_scan_1024(i < n ? in[i] : 0, s_warp, total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

// single block: exclusive scan of the block sums in place, grand total to sums[nb]
__global__ void __launch_bounds__(1024) scan_top_kernel(int *sums, int nb, int *total_out)
{
    __shared__ int s_warp[33];
    __shared__ int s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (int c0 = 0; c0 < nb; c0 += 1024) {
        int i = c0 + threadIdx.x, total;
        int v = i < nb ? sums[i] : 0;
        int ex = block_exclusive_scan_1024(v, s_warp, total);
        int carry = s_carry;
        if (i < nb) sums[i] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) s_carry = carry + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) { sums[nb] = s_carry; if (total_out) *total_out = s_carry; }
}

__global__ void __launch_bounds__(1024) scan_final_kernel(const int *in, int *out, int n, const int *sums, int nb)
{
    __shared__ int s_warp[33];
    int i = blockIdx.x * 1024 + threadIdx.x, total;
    int ex = block_exclusive_scan_1024(i < n ? in[i] : 0, s_warp, total);
    if (i < n) out[i] = sums[blockIdx.x] + ex;
    if (i == n) out[n] = sums[nb];           // out has n+1 entries: out[n] = grand total
}

cudaError_t launch_exclusive_scan(const int *in, int *out, int n, int *tmp, int *total_out, cudaStream_t st)
{
    int nb = n / 1024 + 1;                    // covers index n as well (out[n] = total)
    scan_sums_kernel<<<nb, 1024, 0, st>>>(in, n, tmp);
    scan_top_kernel<<<1, 1024, 0, st>>>(tmp, nb, total_out);
    scan_final_kernel<<<nb, 1024, 0, st>>>(in, out, n, tmp, nb);
    return cudaGetLastError();
}

// ---- SINR medium: pair records -> (packet arrival, node index) order --------------------------
// (seg[t] = first pair of packet t comes from the resolve: scan_tx_counts in resolve_body.cuh)
// record i -> its packet's bucket, at the rank its atomic drew: items = (global rx << 32) | record index
// (grid-stride over the device-side record count: the host does not know it yet)
__global__ void place_items_kernel(PairRecs pr, const Counters *ctr, const int *seg, unsigned long long *items)
{
    const long long m = (long long)ctr->heard;
    if (m > pr.cap) return;                      // record overflow: the host grows the lists and redoes the tick
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) {
        const unsigned long long k = pr.key[i];
        const int t = (int)(k >> pr.key_shift);
        const unsigned long long rx = k & ((1ull << pr.key_shift) - 1);
        items[seg[t] + pr.rank[i]] = (rx << 32) | (unsigned long long)i;
    }
}

// one block per packet: bitonic sort of the bucket by rx in shared memory
__global__ void __launch_bounds__(128) bucket_sort_kernel(PairRecs pr, const Counters *ctr, const int *seg,
                                                          const unsigned long long *items, unsigned *order)
{
    __shared__ unsigned long long s[kBucketSortMax];
    if (ctr->heard == 0 || (long long)ctr->heard > pr.cap || ctr->max_bucket > kBucketSortMax) return;   // nothing heard (seg may be stale: no receiver, no resolve) / host: redo / radix-sort fallback
    const int t = blockIdx.x;
    const int b0 = seg[t], n = seg[t + 1] - b0;
    if (n <= 0) return;
    if (n == 1) { if (threadIdx.x == 0) order[b0] = (unsigned)items[b0]; return; }
    if (n <= 256) {
        // small bucket (the common case): rank by counting, no barriers beyond the load.  The rx keys of one
        // packet's bucket are distinct, so the ranks are a permutation.
        for (int i = threadIdx.x; i < n; i += blockDim.x) s[i] = items[b0 + i];
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const unsigned long long mine = s[i];
            int rank = 0;
#pragma unroll 8
            for (int j = 0; j < n; ++j) rank += s[j] < mine;
            order[b0 + rank] = (unsigned)mine;
        }
        return;
    }
    int np2 = 2;
    while (np2 < n) np2 <<= 1;
    for (int i = threadIdx.x; i < np2; i += blockDim.x) s[i] = i < n ? items[b0 + i] : ~0ull;
    __syncthreads();
    for (int k = 2; k <= np2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < np2; i += blockDim.x) {
                const int l = i ^ j;
                if (l > i) {
                    const unsigned long long a = s[i], b = s[l];
                    if ((a > b) == ((i & k) == 0)) { s[i] = b; s[l] = a; }
                }
            }
            __syncthreads();
        }
    for (int i = threadIdx.x; i < n; i += blockDim.x) order[b0 + i] = (unsigned)s[i];
}

// fallback for buckets that do not fit a block: radix sort of the whole (t, rx) key
__global__ void fallback_fill_kernel(PairRecs pr, long long m, unsigned long long *key, unsigned *val)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) { key[i] = pr.key[i]; val[i] = (unsigned)i; }
}

size_t sort_pairs_tmp_bytes(long long n)
{
    size_t bytes = 0;
    cub::DoubleBuffer<unsigned long long> k(nullptr, nullptr);
    cub::DoubleBuffer<unsigned> v(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, k, v, (int)n, 0, 64, 0);
    return bytes;
}

cudaError_t launch_order_buckets(PairRecs pr, const Counters *ctr, int n_tx, const int *seg, unsigned long long *items,
                                 unsigned *order, cudaStream_t st)
{
    if (n_tx <= 0) return cudaSuccess;
    place_items_kernel<<<148 * 8, 256, 0, st>>>(pr, ctr, seg, items);
    bucket_sort_kernel<<<n_tx, 128, 0, st>>>(pr, ctr, seg, items, order);
    return cudaGetLastError();
}

cudaError_t launch_order_fallback(PairRecs pr, long long m, int key_bits, OrderScratch sc, unsigned *order, cudaStream_t st)
{
    if (m <= 0) return cudaSuccess;
    const unsigned blocks = (unsigned)((m + 255) / 256);
    fallback_fill_kernel<<<blocks, 256, 0, st>>>(pr, m, sc.fb_key, sc.fb_val);
    cub::DoubleBuffer<unsigned long long> k(sc.fb_key, sc.fb_key_alt);
    cub::DoubleBuffer<unsigned> v(sc.fb_val, sc.fb_val_alt);
    size_t bytes = sc.cub_tmp_bytes;
    RMB_CUDA_OK(cub::DeviceRadixSort::SortPairs(sc.cub_tmp, bytes, k, v, (int)m, 0, key_bits, st));
    return cudaMemcpyAsync(order, v.Current(), (size_t)m * sizeof(unsigned), cudaMemcpyDeviceToDevice, st);
}

// ---- stochastic rule + outputs ---------------------------------------------------------------
// first sorted pair of every packet (lower bound of t << key_shift)
__global__ void seg_kernel(const unsigned long long *key, long long n, int n_tx, int shift, int *seg)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > n_tx) return;
    const unsigned long long target = (unsigned long long)t << shift;
    long long lo = 0, hi = n;
    while (lo < hi) { long long mid = (lo + hi) >> 1; if (key[mid] < target) lo = mid + 1; else hi = mid; }
    seg[t] = (int)lo;
}

// first result row of every packet (lower bound of t in the sorted tx column): the compact result form
__global__ void tx_first_kernel(const int *o_tx, long long n, int n_tx, int *first)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > n_tx) return;
    long long lo = 0, hi = n;
    while (lo < hi) { long long mid = (lo + hi) >> 1; if (o_tx[mid] < t) lo = mid + 1; else hi = mid; }
    first[t] = (int)lo;
}

cudaError_t launch_tx_first(const int *o_tx, long long n, int n_tx, int *first, cudaStream_t st)
{
    tx_first_kernel<<<(n_tx + 1 + 255) / 256, 256, 0, st>>>(o_tx, n, n_tx, first);
    return cudaGetLastError();
}

// 1 where the pair would consume a uniform if its packet did not fail
__global__ void drawflag_kernel(const uint8_t *cand, const double *p, long long n, int *flag)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = (cand[i] && p[i] < 1.0) ? 1 : 0;
}

__device__ __forceinline__ double tx_success(const DevParams *dp, const rmb_tx_record &r)
{
    if (dp->model == RMB_MODEL_UDGM) return dp->success_ratio_rx * r.txprob;   // UDGMRadioMedium.java:63-65 (sic: Rx)
    if (dp->model == RMB_MODEL_N2N || dp->model == RMB_MODEL_SINR) return r.txprob;   // N2NRadioMedium.java:24-26
    return 1.0;                                                                 // Null / ConstantLoss: no loss model
}

// no draws possible: the tx rule degenerates to p_tx <= 0
__global__ void txfail_det_kernel(const DevParams *dp, const rmb_tx_record *rec, int n_tx, uint8_t *failed, long long *base)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tx) return;
    const rmb_tx_record r = rec[t];
    failed[t] = (r.src >= 0 && tx_success(dp, r) <= 0.0) ? 1 : 0;
    base[t] = 0;
}

// The draw order is inherently sequential over packets (whether packet t consumes its
// receivers' uniforms depends on its own tx draw).  One block: all threads stage a chunk of
// per-packet (p_tx, draws) into shared memory, thread 0 walks it.
// Sharded receivers: all_counts[g][t] = draws the receivers of shard g would make for packet t (every shard's
// per-packet counts, gathered by the host side); a packet's draws are laid out tx draw, shard 0, shard 1, ...
// as in the unsharded node order, and every shard walks the same global sequence.  all_counts == nullptr:
// unsharded, the counts come from this handle's own scan.
__global__ void __launch_bounds__(1024)
txloop_kernel(const DevParams *dp, const rmb_tx_record *rec, int n_tx, const int *seg, const int *draw_scan,
              const double *uniforms, long long n_uniforms, uint8_t *failed, long long *base, Counters *ctr,
              const int *all_counts, int world, int rank, const int *order, int n_order)
{
    // order (may be null): the records in the tick's ORIGINAL packet arrival order -- the gathered record list of
    // a sharded tick is grouped by owner rank, but the reference draws packet by packet as they arrived
    __shared__ int s_before[1024], s_t[1024];
    const int n_rec = n_tx;
    if (order) n_tx = n_order;
    __shared__ double s_ptx[1024];
    __shared__ int s_draws[1024];
    __shared__ long long s_u;
    __shared__ int s_short;
    if (threadIdx.x == 0) { s_u = 0; s_short = 0; }
    for (int c0 = 0; c0 < n_tx; c0 += 1024) {
        const int k = c0 + threadIdx.x;
        if (k < n_tx) {
            const int t = order ? order[k] : k;
            s_t[threadIdx.x] = t;
            const rmb_tx_record r = rec[t];
            s_ptx[threadIdx.x] = r.src >= 0 ? tx_success(dp, r) : 1.0;
            if (all_counts) {
                int tot = 0, before = 0;
                for (int g = 0; g < world; ++g) { const int c = all_counts[(size_t)g * n_rec + t]; if (g < rank) before += c; tot += c; }
                s_draws[threadIdx.x] = tot; s_before[threadIdx.x] = before;
            } else {
                s_draws[threadIdx.x] = draw_scan[seg[t + 1]] - draw_scan[seg[t]];
                s_before[threadIdx.x] = 0;
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            long long u = s_u;
            int sh = s_short;
            const int nc = min(1024, n_tx - c0);
            for (int i = 0; i < nc; ++i) {
                const double ptx = s_ptx[i];
                int f = 0;
                if (ptx <= 0.0) f = 1;
                else if (ptx < 1.0) {
                    double uu = 0.0;
                    if (u < n_uniforms) uu = uniforms[u]; else sh = 1;
                    ++u;
                    if (uu > ptx) f = 1;
                }
                failed[s_t[i]] = (uint8_t)f;
                base[s_t[i]] = u + s_before[i];
                if (!f) u += s_draws[i];
            }
            s_u = u; s_short = sh;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        ctr->uniforms_used = s_u;
        if (s_u > n_uniforms) s_short = 1;
        ctr->uniforms_short = s_short;
    }
}

__global__ void finalize_kernel(PairArrays pa, long long n, bool stochastic, const double *uniforms, long long n_uniforms,
                                Counters *ctr)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int delivered = 0, interfered = 0;
    if (i < n) {
        const unsigned long long k = pa.key[i];
        const int t = (int)(k >> pa.key_shift);
        const double p = pa.p[i];
        int outcome;
        if (pa.tx_failed[t] || !pa.cand[i]) outcome = RMB_INTERFERED;
        else if (p < 1.0 && stochastic) {
            const long long idx = pa.tx_base[t] + (long long)(pa.draw_scan[i] - pa.draw_scan[pa.seg[t]]);
            const double uu = idx < n_uniforms ? uniforms[idx] : 0.0;
            outcome = uu > p ? RMB_INTERFERED : RMB_DELIVERED;
        } else outcome = RMB_DELIVERED;
        pa.o_tx[i] = t;
        pa.o_rx[i] = (int)(k & ((1ull << pa.key_shift) - 1));
        const double rssi = pa.rssi[i], sinr = pa.sinr[i];
        pa.o_rssi[i] = (float)rssi;
        pa.o_sinr[i] = (float)sinr;
        if (pa.o_rssi64) { pa.o_rssi64[i] = rssi; pa.o_sinr64[i] = sinr; }
        pa.o_outcome[i] = (uint8_t)outcome;
        delivered = outcome == RMB_DELIVERED; interfered = !delivered;
    }
    for (int o = 16; o > 0; o >>= 1) {
        delivered += __shfl_xor_sync(0xffffffffu, delivered, o);
        interfered += __shfl_xor_sync(0xffffffffu, interfered, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (delivered) atomicAdd(&ctr->n_delivered, delivered);
        if (interfered) atomicAdd(&ctr->n_interfered, interfered);
    }
}

// draws the local receivers of packet t would make if the packet does not fail
__global__ void draw_counts_kernel(const int *seg, const int *draw_scan, int n_tx, int *counts)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n_tx) counts[t] = draw_scan[seg[t + 1]] - draw_scan[seg[t]];
}

// stochastic rule, first half: segments, draw flags and their scan; optionally the per-packet draw counts
cudaError_t launch_emit_draw_prepare(int n_tx, PairArrays pa, long long n, int *counts_out, cudaStream_t st)
{
    if (n_tx <= 0) return cudaSuccess;
    seg_kernel<<<(n_tx + 1 + 255) / 256, 256, 0, st>>>(pa.key, n, n_tx, pa.key_shift, pa.seg);
    if (n > 0) drawflag_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(pa.cand, pa.p, n, pa.draw_scan);
    // in-place exclusive scan (n+1 entries)
    cudaError_t e = launch_exclusive_scan(pa.draw_scan, pa.draw_scan, (int)n, pa.scan_tmp, nullptr, st);
    if (e != cudaSuccess) return e;
    if (counts_out) draw_counts_kernel<<<(n_tx + 255) / 256, 256, 0, st>>>(pa.seg, pa.draw_scan, n_tx, counts_out);
    return cudaGetLastError();
}

// second half: the sequential walk over the packets (tx draws, bases) and the per-pair outcomes
cudaError_t launch_emit_draw_finish(const DevParams *dp, const rmb_tx_record *rec, int n_tx, PairArrays pa, long long n,
                                    const double *uniforms, long long n_uniforms, Counters *ctr,
                                    const int *all_counts, int world, int rank, const int *order, int n_order,
                                    cudaStream_t st)
{
    if (n_tx <= 0) return cudaSuccess;
    txloop_kernel<<<1, 1024, 0, st>>>(dp, rec, n_tx, pa.seg, pa.draw_scan, uniforms, n_uniforms, pa.tx_failed,
                                      pa.tx_base, ctr, all_counts, world, rank, order, n_order);
    if (n > 0)
        finalize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(pa, n, true, uniforms, n_uniforms, ctr);
    return cudaGetLastError();
}

cudaError_t launch_emit(const DevParams *dp, const rmb_tx_record *rec, int n_tx, PairArrays pa, long long n,
                        bool stochastic, const double *uniforms, long long n_uniforms, Counters *ctr, cudaStream_t st)
{
    if (n_tx <= 0) return cudaSuccess;
    if (stochastic) {
        RMB_CUDA_OK(launch_emit_draw_prepare(n_tx, pa, n, nullptr, st));
        return launch_emit_draw_finish(dp, rec, n_tx, pa, n, uniforms, n_uniforms, ctr, nullptr, 1, 0, nullptr, 0, st);
    }
    txfail_det_kernel<<<(n_tx + 255) / 256, 256, 0, st>>>(dp, rec, n_tx, pa.tx_failed, pa.tx_base);
    if (n > 0)
        finalize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(pa, n, false, uniforms, n_uniforms, ctr);
    return cudaGetLastError();
}

// ---- SINR medium: records in (packet, node) order -> outputs ------------------------------------
// Deterministic reception (no probability < 1 anywhere): the three-way outcome is final here.
// Stochastic reception: only the sorted pair arrays are produced; launch_emit applies the draws.
// m_host < 0: the pair count is read on the device (the host has not synchronised yet) and the kernel
// stands down when the host will have to redo the tick or take the radix-sort fallback.
// host_ctr (pinned host memory, may be null): the last block to finish stores the tick's counters there, so
// that the host needs no copy operation behind the tick -- a stream synchronisation is enough.
template <bool STOCHASTIC>
__global__ void emit_sinr_kernel(NodeArrays nd, const DevParams *dp, PairRecs pr, const unsigned *order, PairArrays pa,
                                 long long m_host, Counters *ctr, Counters *host_ctr)
{
    long long m = m_host;
    if (m_host < 0) {
        m = (long long)ctr->heard;
        if (m > pr.cap || ctr->max_bucket > kBucketSortMax) m = 0;       // stand down: the host redoes / re-orders
    }
    int delivered = 0, interfered = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) {
        const unsigned ri = order[i];
        const unsigned long long k = pr.key[ri];
        const double rssi = pr.prx[ri], sinr = pr.sinr[ri];
        const bool cand = (pr.flags[ri] & kDecodable) != 0;
        if (STOCHASTIC) {
            const int p = (int)(k & ((1ull << pa.key_shift) - 1)) - dp->rx_base;
            pa.key[i] = k; pa.p[i] = nd.rxprob[p]; pa.rssi[i] = rssi; pa.sinr[i] = sinr; pa.cand[i] = cand ? 1 : 0;
        } else {
            const int t = (int)(k >> pa.key_shift);
            const int outcome = (pa.tx_failed[t] || !cand) ? RMB_INTERFERED : RMB_DELIVERED;
            pa.o_tx[i] = t;
            pa.o_rx[i] = (int)(k & ((1ull << pa.key_shift) - 1));
            pa.o_rssi[i] = (float)rssi;
            pa.o_sinr[i] = (float)sinr;
            if (pa.o_rssi64) { pa.o_rssi64[i] = rssi; pa.o_sinr64[i] = sinr; }
            pa.o_outcome[i] = (uint8_t)outcome;
            delivered += outcome == RMB_DELIVERED; interfered += outcome != RMB_DELIVERED;
        }
    }
    if (!STOCHASTIC) {
        for (int o = 16; o > 0; o >>= 1) {
            delivered += __shfl_xor_sync(0xffffffffu, delivered, o);
            interfered += __shfl_xor_sync(0xffffffffu, interfered, o);
        }
        if ((threadIdx.x & 31) == 0) {
            if (delivered) atomicAdd(&ctr->n_delivered, delivered);
            if (interfered) atomicAdd(&ctr->n_interfered, interfered);
        }
    }
    if (host_ctr) {
        __shared__ int s_last;
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

cudaError_t launch_emit_sinr(const NodeArrays &nd, const DevParams *dp, const rmb_tx_record *rec, int n_tx, PairRecs pr,
                             const unsigned *order, PairArrays pa, long long m, bool stochastic, Counters *ctr,
                             Counters *host_ctr, cudaStream_t st)
{
    if (n_tx <= 0) return cudaSuccess;
    const unsigned blocks = m < 0 ? 148u * 8u : (unsigned)((m + 255) / 256);
    if (stochastic) {
        if (m > 0) emit_sinr_kernel<true><<<blocks, 256, 0, st>>>(nd, dp, pr, order, pa, m, ctr, nullptr);
        return cudaGetLastError();
    }
    // tx_failed / tx_base of the deterministic rule were written by the tx preparation kernel
    if (m != 0) emit_sinr_kernel<false><<<blocks, 256, 0, st>>>(nd, dp, pr, order, pa, m, ctr, host_ctr);
    return cudaGetLastError();
}

}  // namespace rmb
