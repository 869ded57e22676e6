// static_tick.cu -- rmb_udgm_tick_static: the deterministic UDGM tick (refmedia.cu's count -> scan -> scatter) for
// callers that cannot allocate: no handle, no cudaMalloc / cudaFree / allocating library call anywhere below.  The
// per-(packet, node block) counts that refmedia.cu keeps in a growing device buffer live here in ONE file-scope
// __device__ array of RMB_STATIC_MAX_CELLS ints (2 MiB, held by every GPU that loads the library until exit).
// No counterpart in the reference.
//
// The array is shared by every call on a device and keeps the last call's contents.  Nothing in it has to start at
// zero: every cell a call reads (n_tx * ceil(n / 256) of them) is first written by that call's count kernel, and the
// scan runs in place.  Two calls that overlap on one GPU (different streams, several threads) race on it.
//
// Host code may not hand the array's NAME to a kernel -- in host code that is a host-side stand-in, not a device
// address -- so the device address is looked up once per device with cudaGetSymbolAddress (device current) and cached.
#include "../../include/radiomedium_b200.h"
#include "exact_math.cuh"
#include <cuda_runtime.h>
#include <cstdio>
#include <mutex>

namespace {

constexpr int kThreads = 256;                 // nodes per cell (one block)
constexpr int kScanThreads = 1024, kScanPer = 4;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) int g_static_cells[RMB_STATIC_MAX_CELLS];

std::mutex g_mu;
int *g_cells_on[kMaxDevices] = {};            // device address of g_static_cells, per device (set-up on first use)

// Receiver filter + UDGM range rule with every probability at 1 (UDGMRadioMedium.java:67-81,100-102):
// heard <=> node != source, enabled, on the packet's channel, d*d / (R*R) <= 1  (Position.java:56-64, unfused double)
__device__ __forceinline__ bool heard_by(int node, int src, double sx, double sy, double sz, int pkt_channel,
                                         const double *x, const double *y, const double *z, const int32_t *channel,
                                         const uint8_t *enabled, double R)
{
    if (node == src || (enabled && !enabled[node]) || channel[node] != pkt_channel) return false;
    if (R == 0.0) return false;
    const double d = rmb::xm::distance(sx, sy, sz, x[node], y[node], z ? z[node] : 0.0);
    const double ratio = rmb::xm::dvd(rmb::xm::mul(d, d), rmb::xm::mul(R, R));
    return !(ratio > 1.0);
}

struct Packet { int src, channel; double x, y, z; bool valid; };

__device__ __forceinline__ Packet packet_of(int t, int n, const int32_t *src, const int32_t *pkt_channel,
                                            const double *x, const double *y, const double *z, const int32_t *channel)
{
    Packet p;
    p.src = src[t];
    p.valid = p.src >= 0 && p.src < n;           // anything else (e.g. -1 padding) is heard by nobody
    p.x = p.y = p.z = 0.0; p.channel = 0;
    if (p.valid) {
        p.x = x[p.src]; p.y = y[p.src]; p.z = z ? z[p.src] : 0.0;
        p.channel = pkt_channel ? pkt_channel[t] : channel[p.src];
    }
    return p;
}

// cell = (packet t, node block b): how many of the block's nodes hear the packet
__global__ void __launch_bounds__(kThreads)
static_count_kernel(int n, int nblk, const double *x, const double *y, const double *z, const int32_t *channel,
                    const uint8_t *enabled, const int32_t *src, const int32_t *pkt_channel, double R, int *cells)
{
    const int t = blockIdx.x / nblk, b = blockIdx.x % nblk;
    const int node = b * kThreads + threadIdx.x;
    const Packet p = packet_of(t, n, src, pkt_channel, x, y, z, channel);
    const bool heard = p.valid && node < n && heard_by(node, p.src, p.x, p.y, p.z, p.channel, x, y, z, channel, enabled, R);
    const int c = __syncthreads_count(heard ? 1 : 0);
    if (threadIdx.x == 0) cells[blockIdx.x] = c;
}

// in-place exclusive scan of the cells by ONE block (tiles of 4096 with a running carry); total -> *n_pairs and
// tx_first[n_tx]
__global__ void __launch_bounds__(kScanThreads)
static_scan_kernel(int *cells, int n_cells, int n_tx, int32_t *tx_first, int64_t *n_pairs)
{
    __shared__ int s_warp[kScanThreads / 32];
    __shared__ int s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int base = 0; base < n_cells; base += kScanThreads * kScanPer) {
        const int i0 = base + threadIdx.x * kScanPer;
        int v[kScanPer], sum = 0;
#pragma unroll
        for (int k = 0; k < kScanPer; ++k) { v[k] = i0 + k < n_cells ? cells[i0 + k] : 0; sum += v[k]; }
        int inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += u; }
        if (lane == 31) s_warp[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            int w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += u; }
            s_warp[lane] = w;                                   // inclusive over the warps
        }
        __syncthreads();
        const int carry = s_carry;
        int ex = carry + (wid ? s_warp[wid - 1] : 0) + inc - sum;
#pragma unroll
        for (int k = 0; k < kScanPer; ++k) { if (i0 + k < n_cells) cells[i0 + k] = ex; ex += v[k]; }
        __syncthreads();
        if (threadIdx.x == kScanThreads - 1) s_carry = carry + s_warp[kScanThreads / 32 - 1];
        __syncthreads();
    }
    if (threadIdx.x == 0) { *n_pairs = (int64_t)s_carry; tx_first[n_tx] = s_carry; }
}

// second pass: the same decisions, written at cell offset + rank inside the block => rows in (packet arrival, node
// index) order, the reference's generateReceptionEvents call order (UDGMRadioMedium.java:97-114)
__global__ void __launch_bounds__(kThreads)
static_scatter_kernel(int n, int nblk, const double *x, const double *y, const double *z, const int32_t *channel,
                      const uint8_t *enabled, const int32_t *src, const double *power_dbm, const int32_t *pkt_channel,
                      double R, const int *cells, long long capacity, int32_t *o_tx, int32_t *o_rx, float *o_rssi,
                      uint8_t *o_outcome, int32_t *tx_first)
{
    __shared__ int s_w[kThreads / 32];
    const int t = blockIdx.x / nblk, b = blockIdx.x % nblk;
    const int node = b * kThreads + threadIdx.x;
    const Packet p = packet_of(t, n, src, pkt_channel, x, y, z, channel);
    const bool heard = p.valid && node < n && heard_by(node, p.src, p.x, p.y, p.z, p.channel, x, y, z, channel, enabled, R);
    const unsigned bh = __ballot_sync(0xffffffffu, heard);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) s_w[wid] = __popc(bh);
    __syncthreads();
    const int off = cells[blockIdx.x];
    if (b == 0 && threadIdx.x == 0) tx_first[t] = off;
    if (heard) {
        int woff = 0;
        for (int i = 0; i < wid; ++i) woff += s_w[i];
        const long long o = (long long)off + woff + __popc(bh & ((1u << lane) - 1));
        if (o < capacity) {
            if (o_tx) o_tx[o] = t;
            o_rx[o] = node;
            o_rssi[o] = (float)power_dbm[t];     // RSSI handed on = the packet's tx power (UDGMRadioMedium.java:95,108-110)
            o_outcome[o] = RMB_DELIVERED;        // no loss model active: heard == delivered
        }
    }
}

int fail(char *text, size_t bytes, int code, const char *msg)
{
    if (text && bytes) snprintf(text, bytes, "%s", msg);
    return code;
}

}  // namespace

extern "C" int rmb_udgm_tick_static(int32_t device, void *stream, int32_t n, const double *d_x, const double *d_y, const double *d_z,
                                    const int32_t *d_channel, const uint8_t *d_enabled, int32_t n_tx, const int32_t *d_src,
                                    const double *d_power_dbm, const int32_t *d_pkt_channel, double transmission_range,
                                    int64_t capacity, int32_t *d_tx, int32_t *d_rx, float *d_rssi_dbm, uint8_t *d_outcome,
                                    int32_t *d_tx_first, int64_t *d_n_pairs, char *error_text, size_t error_text_bytes)
{
    char msg[256];
    if (error_text && error_text_bytes) error_text[0] = 0;
    if (device < 0 || device >= kMaxDevices) return fail(error_text, error_text_bytes, RMB_ERR_INVALID, "rmb_udgm_tick_static: device index out of range (0..63)");
    if (n < 0 || n_tx < 0 || capacity < 0 || !(transmission_range >= 0.0))
        return fail(error_text, error_text_bytes, RMB_ERR_INVALID, "rmb_udgm_tick_static: negative n / n_tx / capacity / range");
    if (!d_tx_first || !d_n_pairs || (n > 0 && (!d_x || !d_y || !d_channel)) || (n_tx > 0 && (!d_src || !d_power_dbm)) ||
        (capacity > 0 && (!d_rx || !d_rssi_dbm || !d_outcome)))
        return fail(error_text, error_text_bytes, RMB_ERR_INVALID, "rmb_udgm_tick_static: a required device pointer is NULL");
    const long long nblk = ((long long)n + kThreads - 1) / kThreads;
    const long long n_cells = nblk * (long long)n_tx;
    if (n_cells > (long long)RMB_STATIC_MAX_CELLS) {
        snprintf(msg, sizeof msg, "rmb_udgm_tick_static: n_tx * ceil(n / 256) = %lld exceeds RMB_STATIC_MAX_CELLS = %d "
                                  "(the scratch is a fixed __device__ array; use rmb_tick for larger ticks)", n_cells, (int)RMB_STATIC_MAX_CELLS);
        return fail(error_text, error_text_bytes, RMB_ERR_CAPACITY, msg);
    }
    cudaError_t e = cudaSetDevice(device);
    int *cells = nullptr;
    if (e == cudaSuccess) {
        std::lock_guard<std::mutex> lk(g_mu);
        if (!g_cells_on[device]) {
            // set-up, once per device, that device current: the DEVICE address of the array
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_static_cells);
            if (e == cudaSuccess) g_cells_on[device] = (int *)p;
        }
        cells = g_cells_on[device];
    }
    if (e != cudaSuccess || !cells) {
        snprintf(msg, sizeof msg, "rmb_udgm_tick_static: set-up on device %d failed: %s", (int)device, cudaGetErrorString(e));
        return fail(error_text, error_text_bytes, e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver ? RMB_ERR_NO_DEVICE : RMB_ERR_CUDA, msg);
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (n_cells > 0)
        static_count_kernel<<<(unsigned)n_cells, kThreads, 0, st>>>(n, (int)nblk, d_x, d_y, d_z, d_channel, d_enabled, d_src, d_pkt_channel,
                                                                     transmission_range, cells);
    static_scan_kernel<<<1, kScanThreads, 0, st>>>(cells, (int)n_cells, n_tx, d_tx_first, d_n_pairs);
    if (n_cells > 0)
        static_scatter_kernel<<<(unsigned)n_cells, kThreads, 0, st>>>(n, (int)nblk, d_x, d_y, d_z, d_channel, d_enabled, d_src, d_power_dbm,
                                                                       d_pkt_channel, transmission_range, cells, (long long)capacity,
                                                                       d_tx, d_rx, d_rssi_dbm, d_outcome, d_tx_first);
    else if (n_tx > 0) {
        // no nodes: every packet's first row is 0 (the scan kernel wrote tx_first[n_tx] and the count)
        e = cudaMemsetAsync(d_tx_first, 0, (size_t)n_tx * sizeof(int32_t), st);
    }
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(msg, sizeof msg, "rmb_udgm_tick_static: launch failed: %s", cudaGetErrorString(e));
        return fail(error_text, error_text_bytes, RMB_ERR_CUDA, msg);
    }
    return RMB_OK;
}
