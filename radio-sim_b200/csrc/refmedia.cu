// refmedia.cu -- K2: the reference's own four media, evaluated exactly (double, unfused) so
// that the delivered / interfered sets are bit-identical to the Java loops:
//   NullRadioMedium.transmit              (NullRadioMedium.java:48-77)      filter only
//   UDGMRadioMedium.transmit              (UDGMRadioMedium.java:63-116)     ratio = d*d / (R*R) <= 1
//   UDGMConstantLossRadioMedium.transmit  (UDGMConstantLossRadioMedium.java:16-36)  d < range (strict)
//   N2NRadioMedium.transmit               (N2NRadioMedium.java:24-72)       matrix[srcId-1][dstId-1]
// Receiver filter (UDGMRadioMedium.java:100-102): node != source, enabled, same channel as the
// PACKET.  One thread per (packet, receiver); two passes (count, scan, scatter) give the heard
// pairs directly in (packet arrival, node index) order with ballot/prefix compaction.
#include "common.cuh"
#include "kernels.h"
#include "exact_math.cuh"

namespace rmb {

constexpr int kRefThreads = 256;

int refmedia_blocks_per_tx(int n) { return (n + kRefThreads - 1) / kRefThreads; }

// returns the reception probability p (> 0 => heard), and whether the filter passed
__device__ __forceinline__ double ref_probability(const NodeArrays &nd, const DevParams *dp, const double *matrix,
                                                  int matrix_n, const rmb_tx_record &r, int node, bool &filter,
                                                  const int *tx_epoch, int hd_epoch)
{
    filter = (dp->rx_base + node) != r.src && nd.enabled[node] && nd.channel[node] == r.channel;
    if (tx_epoch && tx_epoch[node] == hd_epoch) filter = false;       // half duplex: the node is on the air itself
    if (!filter) return 0.0;
    switch (dp->model) {
    case RMB_MODEL_NULL:
        return 1.0;
    case RMB_MODEL_UDGM_CONSTANT: {
        const double d = xm::distance(r.x, r.y, r.z, nd.x[node], nd.y[node], nd.z[node]);
        return d < dp->constant_range ? 1.0 : 0.0;
    }
    case RMB_MODEL_UDGM: {
        // getRxSuccessProbability (UDGMRadioMedium.java:67-81); Math.pow(v, 2.0) == v*v
        const double d = xm::distance(r.x, r.y, r.z, nd.x[node], nd.y[node], nd.z[node]);
        const double d2 = xm::mul(d, d);
        const double R = dp->transmission_range;
        if (R == 0.0) return 0.0;
        double ratio = xm::dvd(d2, xm::mul(R, R));
        if (ratio > 1.0) return 0.0;
        ratio = xm::sub(1.0, xm::mul(ratio, xm::sub(1.0, dp->success_ratio_rx)));
        return xm::mul(ratio, nd.rxprob[node]);
    }
    case RMB_MODEL_N2N: {
        // ids are 1-based integers (N2NRadioMedium.java:28-37, Node.java:52-58)
        const int s = r.src - dp->rx_base;
        const int sid = (s >= 0 && s < nd.n) ? nd.id_int[s] : -1, did = nd.id_int[node];
        if (!matrix || sid <= 0 || did <= 0 || sid > matrix_n || did > matrix_n) return 0.0;
        return xm::mul(matrix[(size_t)(sid - 1) * matrix_n + (did - 1)], nd.rxprob[node]);
    }
    }
    return 0.0;
}

__global__ void __launch_bounds__(kRefThreads)
refmedia_count_kernel(NodeArrays nd, const DevParams *dp, const double *matrix, int matrix_n, int nblk,
                      const rmb_tx_record *rec, int *blkcnt, Counters *ctr, int stochastic, const int *tx_epoch, int hd_epoch)
{
    __shared__ int s_h[kRefThreads / 32], s_f[kRefThreads / 32];
    const int t = blockIdx.x / nblk, b = blockIdx.x % nblk;
    const int node = b * kRefThreads + threadIdx.x;
    const rmb_tx_record r = rec[t];
    bool filter = false, heard = false;
    if (node < nd.n && r.src >= 0) heard = ref_probability(nd, dp, matrix, matrix_n, r, node, filter, tx_epoch, hd_epoch) > 0.0;
    if (b == 0 && threadIdx.x == 0 && r.src >= 0 && !stochastic) {
        // a record whose tx rule needs a draw must not be evaluated on the deterministic path (see prepare_all_kernel)
        const double pt = dp->model == RMB_MODEL_UDGM ? dp->success_ratio_rx * r.txprob : dp->model == RMB_MODEL_N2N ? r.txprob : 1.0;
        if (pt > 0.0 && pt < 1.0) ctr->needs_draws = 1;
    }
    const unsigned bh = __ballot_sync(0xffffffffu, heard), bf = __ballot_sync(0xffffffffu, filter);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) { s_h[wid] = __popc(bh); s_f[wid] = __popc(bf); }
    __syncthreads();
    if (threadIdx.x == 0) {
        int h = 0, f = 0;
        for (int i = 0; i < kRefThreads / 32; ++i) { h += s_h[i]; f += s_f[i]; }
        blkcnt[blockIdx.x] = h;
        if (f) atomicAdd(&ctr->decisions, (unsigned long long)f);
        if (h) atomicAdd(&ctr->heard, (unsigned long long)h);
    }
}

__global__ void __launch_bounds__(kRefThreads)
refmedia_scatter_kernel(NodeArrays nd, const DevParams *dp, const double *matrix, int matrix_n, int nblk,
                        const rmb_tx_record *rec, const int *blkoff, PairArrays pa, uint8_t *node_recv,
                        const int *tx_epoch, int hd_epoch)
{
    __shared__ int s_w[kRefThreads / 32];
    const int t = blockIdx.x / nblk, b = blockIdx.x % nblk;
    const int node = b * kRefThreads + threadIdx.x;
    const rmb_tx_record r = rec[t];
    bool filter = false;
    double p = 0.0;
    if (node < nd.n && r.src >= 0) p = ref_probability(nd, dp, matrix, matrix_n, r, node, filter, tx_epoch, hd_epoch);
    const bool heard = p > 0.0;
    const unsigned bh = __ballot_sync(0xffffffffu, heard);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) s_w[wid] = __popc(bh);
    __syncthreads();
    int woff = 0;
    for (int i = 0; i < wid; ++i) woff += s_w[i];
    if (heard) {
        const long long o = (long long)blkoff[blockIdx.x] + woff + __popc(bh & ((1u << lane) - 1));
        pa.key[o] = ((unsigned long long)(unsigned)t << pa.key_shift) | (unsigned)(dp->rx_base + node);
        pa.p[o] = p;
        pa.rssi[o] = r.power_dbm;          // RSSI handed on = packet tx power (UDGMRadioMedium.java:95,108-110)
        pa.sinr[o] = __longlong_as_double(0x7ff8000000000000ll);
        pa.cand[o] = 1;
        node_recv[node] = 1;
    }
}

// half duplex: stamp the tick's local sources (the count / scatter kernels then drop them as receivers)
__global__ void refmedia_mark_kernel(NodeArrays nd, const DevParams *dp, int n_tx, const rmb_tx_record *rec, int *tx_epoch, int hd_epoch)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tx) return;
    const int s = rec[t].src - dp->rx_base;
    if (rec[t].src >= 0 && s >= 0 && s < nd.n) tx_epoch[s] = hd_epoch;
}

cudaError_t launch_refmedia_mark(const NodeArrays &nd, const DevParams *dp, int n_tx, const rmb_tx_record *rec,
                                 int *tx_epoch, int hd_epoch, cudaStream_t st)
{
    if (n_tx <= 0 || !tx_epoch) return cudaSuccess;
    refmedia_mark_kernel<<<(n_tx + 255) / 256, 256, 0, st>>>(nd, dp, n_tx, rec, tx_epoch, hd_epoch);
    return cudaGetLastError();
}

cudaError_t launch_refmedia_count(const NodeArrays &nd, const DevParams *dp, const double *matrix, int matrix_n,
                                  int n_tx, const rmb_tx_record *rec, int *blkcnt, Counters *ctr, bool stochastic,
                                  const int *tx_epoch, int hd_epoch, cudaStream_t st)
{
    const int nblk = refmedia_blocks_per_tx(nd.n);
    if (n_tx <= 0 || nblk <= 0) return cudaSuccess;
    refmedia_count_kernel<<<(unsigned)((long long)nblk * n_tx), kRefThreads, 0, st>>>(nd, dp, matrix, matrix_n, nblk, rec, blkcnt, ctr,
                                                                                     stochastic ? 1 : 0, tx_epoch, hd_epoch);
    return cudaGetLastError();
}

cudaError_t launch_refmedia_scatter(const NodeArrays &nd, const DevParams *dp, const double *matrix, int matrix_n,
                                    int n_tx, const rmb_tx_record *rec, const int *blkoff, PairArrays pa,
                                    uint8_t *node_recv, const int *tx_epoch, int hd_epoch, cudaStream_t st)
{
    const int nblk = refmedia_blocks_per_tx(nd.n);
    if (n_tx <= 0 || nblk <= 0) return cudaSuccess;
    refmedia_scatter_kernel<<<(unsigned)((long long)nblk * n_tx), kRefThreads, 0, st>>>(nd, dp, matrix, matrix_n, nblk, rec, blkoff, pa, node_recv,
                                                                                       tx_epoch, hd_epoch);
    return cudaGetLastError();
}

}  // namespace rmb
