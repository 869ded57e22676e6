// resolve.cu -- exact (double, strict IEEE) evaluation of the near candidates and the
// guard-band recheck.  This is where every heard / decodable DECISION of the SINR medium is
// taken; the fp32 sweep only supplies the weak far-field interference sum.
//
// Specification implemented here (the CPU oracle states the same):
//   d    = Position.getDistance(src, rx)              (Position.java:56-64, unfused)
//   Prx  = Ptx - (PL0 + (10 n) log10(max(d,d0)/d0))   [dBm]
//   lin  = exp(Prx * ln10/10)                         [mW]
//   Itot = N0 + sum of lin over co-channel packets (arrival order, source != rx)
//   heard iff Prx >= sensitivity && rxProbability > 0
//   SINR = Prx - 10 log10(Itot - lin);  decodable iff SINR >= threshold
#include "resolve_body.cuh"

namespace rmb {

__global__ void __launch_bounds__(128)
resolve_kernel(NodeArrays nd, WorkArrays wk, const DevParams *dp, TxTiles tx, const rmb_tx_record *rec,
               SlotArrays sl, PairRecs pr, Counters *ctr, int n_tx, int *seg, int w_first)
{
    __shared__ int s_last;
    const int w = w_first + blockIdx.x * blockDim.x + threadIdx.x;       // n_work is a multiple of 1024: no partial warps
    resolve_receiver<false>(w, true, wk, dp, tx, rec, sl, pr, ctr);

    // the last block to finish turns the per-packet counts into segment offsets
    // (barrier, then ONE fence + arrival by thread 0: the release pattern of a grid-wide semaphore)
    __syncthreads();
    // (resolve_done counts receivers: the sweep's fused epilogue resolves the others)
    if (threadIdx.x == 0) { __threadfence(); s_last = atomicAdd(&ctr->resolve_done, 128u) + 128u == (unsigned)wk.n_work; }
    __syncthreads();
    if (s_last) { __threadfence(); scan_tx_counts(pr.tx_cnt, n_tx, seg, ctr); }
}


cudaError_t launch_resolve(const NodeArrays &nd, const WorkArrays &wk, const DevParams *dp, TxTiles tx,
                           const rmb_tx_record *rec, SlotArrays sl, PairRecs pr, Counters *ctr, int n_tx, int *seg,
                           int w_first, cudaStream_t st)
{
    if (wk.n_work - w_first <= 0) return cudaSuccess;
    resolve_kernel<<<(wk.n_work - w_first) / 128, 128, 0, st>>>(nd, wk, dp, tx, rec, sl, pr, ctr, n_tx, seg, w_first);
    return cudaGetLastError();
}

// Transciever.getRSSI / getReceivingState for the per-tick node-info (Transciever.java:52-78,
// JSONClientConnection.java:331-341), produced on demand from the work-order results.
__global__ void node_info_kernel(WorkArrays wk, SlotArrays sl, float *node_rssi, double *node_rssi64, uint8_t *node_recv)
{
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= wk.n_work) return;
    const int p = wk.perm[w];
    if (p < 0) return;
    const double v = xm::mw_to_dbm(sl.itot[w]);
    node_rssi[p] = (float)v;
    node_rssi64[p] = v;
    node_recv[p] = sl.hcnt[w] > 0;
}

cudaError_t launch_node_info(const WorkArrays &wk, SlotArrays sl, float *node_rssi, double *node_rssi64, uint8_t *node_recv, cudaStream_t st)
{
    if (wk.n_work <= 0) return cudaSuccess;
    node_info_kernel<<<(wk.n_work + 255) / 256, 256, 0, st>>>(wk, sl, node_rssi, node_rssi64, node_recv);
    return cudaGetLastError();
}

// ---- guard-band recheck ----------------------------------------------------------------------
// Receivers whose SINR lies within the guard band of the threshold cannot be decided from the
// fp32 far-field sum.  Two tiers, both one block per receiver:
//   tier 1: ALL co-channel terms in exact double, summed as a fixed tree (parallel).  A tree
//           sum and the oracle's sequential sum of n positive terms differ by at most
//           2 n u relative (u = 2^-53), i.e. ~1e-11 dB of SINR -- so the decision is certain
//           unless |SINR - threshold| is below that bound;
//   tier 2: for those (practically only constructed test cases) the ADDITION is redone strictly
//           sequentially in arrival order, bit-identical to the Java/C loop, pipelined over
//           double-buffered chunks (one adder thread, 31 warps evaluating the next chunk).
constexpr int kRecheckThreads = 1024;          // tier 2 (sequential sum)
constexpr int kTreeThreads = 256;              // tier 1: work unit = (flagged receiver, part of its bucket)
constexpr int kTreeMaxParts = 64;

// Tier 1.  Few receivers are flagged (C4: ~1e-4 of them), each needs ALL its co-channel terms in
// double: a receiver's bucket is cut into Q parts so that the flagged receivers fill the GPU; a part
// is summed by one block as a fixed tree, the parts in order by one thread after a grid barrier.
__global__ void __launch_bounds__(kTreeThreads, 2)
recheck_tree_kernel(WorkArrays wk, const DevParams *dp, Groups groups, TxTiles tx, const rmb_tx_record *rec,
                    SlotArrays sl, PairRecs pr, Counters *ctr, double *part, unsigned *bar, unsigned bar_target)
{
    __shared__ double s_part[kTreeThreads / 32];
    const bool stand_down = (long long)ctr->heard > pr.cap;      // the host grows the record lists and redoes the tick
    const int n1 = stand_down ? 0 : ctr->n_recheck;               // the host launches a fixed grid before it knows the count
    const int Q = n1 > 0 ? max(1, min(kTreeMaxParts, (int)gridDim.x / n1)) : 1;
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
        for (int i = i0 + threadIdx.x; i < i1; i += kTreeThreads) acc += exact_lin(dp, rec[tx.arrival[j0 + i]], self, X, Y, Z);
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double v = 0.0;
            for (int i = 0; i < kTreeThreads / 32; ++i) v += s_part[i];
            part[u] = v;
        }
        __syncthreads();
    }
    grid_barrier(bar, bar_target);
    for (int b = blockIdx.x * kTreeThreads + threadIdx.x; b < n1; b += gridDim.x * kTreeThreads) {
        const int w = sl.recheck[b];
        int g = 0;
        while (g + 1 < groups.ngroups && w >= groups.w_begin[g + 1]) ++g;
        double itot = dp->n0_mw;
        for (int q = 0; q < Q; ++q) itot += __ldcg(&part[(long long)b * Q + q]);
        const double tol_rel = 2.0 * (double)(tx.gcount[g] + 64) * 1.1102230246251565e-16;
        if (apply_itot(dp, sl, pr, w, itot, tol_rel)) sl.recheck2[atomicAdd(&ctr->n_recheck2, 1)] = w;
    }
}

constexpr int kRecheckChunk = 1984;           // 31 warps x 64 terms
constexpr int kRecheckSeqBlocks = 32;

__global__ void __launch_bounds__(kRecheckThreads)
recheck_seq_kernel(WorkArrays wk, const DevParams *dp, Groups groups, TxTiles tx, const rmb_tx_record *rec,
                   SlotArrays sl, PairRecs pr, const Counters *ctr)
{
    __shared__ double s_lin[2][kRecheckChunk];
    const int n2 = ctr->n_recheck2;
    for (int b = blockIdx.x; b < n2; b += gridDim.x) {
        const int w = sl.recheck2[b];
        const int p = wk.perm[w];
        int g = 0;
        while (g + 1 < groups.ngroups && w >= groups.w_begin[g + 1]) ++g;
        const int j0 = tx.goff[g], nj = tx.gcount[g];
        const double X = wk.x64[w], Y = wk.y64[w], Z = wk.z64[w];
        const int self = dp->rx_base + p;
        const int nchunks = (nj + kRecheckChunk - 1) / kRecheckChunk;
        double itot = dp->n0_mw;
        for (int k = 0; k <= nchunks; ++k) {
            if (threadIdx.x >= 32) {
                if (k < nchunks) {                                   // evaluate chunk k
                    const int c0 = k * kRecheckChunk, nc = min(kRecheckChunk, nj - c0);
                    for (int i = threadIdx.x - 32; i < nc; i += kRecheckThreads - 32)
                        s_lin[k & 1][i] = exact_lin(dp, rec[tx.arrival[j0 + c0 + i]], self, X, Y, Z);
                }
            } else if (threadIdx.x == 0 && k > 0) {                  // add chunk k-1, in order
                const int c0 = (k - 1) * kRecheckChunk, nc = min(kRecheckChunk, nj - c0);
                const double *v = s_lin[(k - 1) & 1];
#pragma unroll 8
                for (int i = 0; i < nc; ++i) itot = xm::add(itot, v[i]);
            }
            __syncthreads();
        }
        if (threadIdx.x == 0) apply_itot(dp, sl, pr, w, itot, 0.0);
        __syncthreads();
    }
}

int recheck_grid()
{
    return 2 * device_sms();                            // __launch_bounds__(kTreeThreads, 2): always co-resident
}

cudaError_t launch_recheck(const NodeArrays &nd, const WorkArrays &wk, const DevParams *dp, Groups groups, TxTiles tx,
                           const rmb_tx_record *rec, SlotArrays sl, PairRecs pr, Counters *ctr, double *part, int grid,
                           unsigned *bar, unsigned bar_target, cudaStream_t st)
{
    (void)nd;
    if (wk.n_work <= 0) return cudaSuccess;
    WorkArrays a_wk = wk; Groups a_groups = groups;
    void *args[] = {(void *)&a_wk, (void *)&dp, (void *)&a_groups, (void *)&tx, (void *)&rec, (void *)&sl, (void *)&pr,
                    (void *)&ctr, (void *)&part, (void *)&bar, (void *)&bar_target};
    RMB_CUDA_OK(launch_grid_synced((const void *)&recheck_tree_kernel, dim3((unsigned)grid), dim3(kTreeThreads), args, st));
    recheck_seq_kernel<<<kRecheckSeqBlocks, kRecheckThreads, 0, st>>>(wk, dp, groups, tx, rec, sl, pr, ctr);
    return cudaGetLastError();
}

}  // namespace rmb
