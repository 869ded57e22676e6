// common.cuh -- shared types between the C-ABI layer (api.cu) and the kernels.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../../include/radiomedium_b200.h"

namespace rmb {

constexpr int kMaxGroups = 64;        // distinct receiver channels handled per tick
constexpr int kTile = 16;             // tx per unrolled tile in the sweep (near test granularity)

// model parameters as the kernels see them
struct DevParams {
    int    model;
    double success_ratio_rx, transmission_range, constant_range;
    double pl0_db, pl_exponent, d0_m, sensitivity_dbm, sinr_threshold_db, noise_floor_dbm;
    double n0_mw;                     // dbm_to_mw(noise_floor) computed on device once
    double sinr_guard_db;             // |sinr - threshold| below this => exact recheck
    int    rx_base;
    int    half_duplex;               // rmb_config.half_duplex
};

// bounding box / fast-path geometry, written by geom_kernel after every position update
struct Geom {
    double cx, cy, cz;                // centre subtracted before the fp32 conversion
    double maxabs;                    // max |coordinate - centre|
    double zmin, zmax;
    int    planar;                    // all z equal => 2-D sweep
    int    pad;
};

// per-tick tx side in sweep order (bucketed by channel group, arrival order inside a bucket)
struct TxTiles {
    // SoA, exactly as a sweep block wants it in shared memory (so that a stage is a handful of 1-D bulk copies):
    // c = P_t^(-1/n), nx / ny / nz = MINUS the scaled centred coordinates (dx' = c * x_rx + nx)
    float *c, *nx, *ny, *nz;
    int    *arrival;                  // bucket position -> arrival index t
    int    *goff;                     // [ngroups+1] bucket offsets (multiples of kTile)
    int    *gcount;                   // [ngroups] real tx per bucket
};

// receiver channel groups in work order (host-built, passed by value)
struct Groups {
    int ngroups;
    int channel[kMaxGroups];
    int w_begin[kMaxGroups + 1];      // work-order range of group g, padded to the block size
    int n_rx[kMaxGroups];             // enabled receivers in the group
    int dense_begin[kMaxGroups + 1];  // first enabled receiver of group g in the unpadded (sorted) list
};

// counters the host reads back once per tick
struct Counters {
    unsigned long long decisions;
    unsigned long long heard;         // total heard pairs (sum of per-receiver heard counts)
    int max_near;                     // max near candidates of any receiver (slot overflow check)
    int n_recheck;
    int n_delivered;
    int n_interfered;
    long long uniforms_used;
    int uniforms_short;
    int max_bucket;                   // most heard pairs of any one packet (chooses the ordering path)
    int n_recheck2;                   // guard-band receivers that needed the sequential sum
    int peer_timeout;                 // multi-GPU exchange: a peer's records did not arrive (see peer_wait)
    unsigned resolve_done;            // blocks of the resolve kernel that have finished (the last one scans tx_cnt)
    unsigned recheck_done;            // work units of the guard-band recheck that have finished
    unsigned emit_done;               // blocks of the emit kernel that have finished (the last one publishes the counters)
    int needs_draws;                  // a record carries 0 < p_tx < 1 but the tick runs on the deterministic path
    // device-side time stamps of the tick (globaltimer, ns): no CUDA event sits between the tick's kernels (an event
    // record between two launches would serialise them and undo the programmatic dependent launch), so the kernels
    // stamp themselves.  *_inv holds ~t so that one atomicMax gives the EARLIEST start.
    unsigned long long t_prepare_begin, t_sweep_begin_inv, t_sweep_end, t_finish_begin_inv;
};

// slot flags
constexpr uint8_t kHeard = 1, kDecodable = 2;

// ---- grid-wide barrier for cooperative launches ------------------------------------------------
// A monotonic arrival counter in global memory; the host passes the value the counter reaches when
// every block of THIS launch has arrived (it adds the grid size per launch), so the counter is
// never reset.  All blocks are co-resident (cudaLaunchCooperativeKernel), so spinning is safe.
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
constexpr unsigned long long kBarrierWaitNs = 5ull * 1000 * 1000 * 1000;
__device__ __forceinline__ void grid_barrier(unsigned *bar, unsigned target)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1u);
        // a barrier that never completes (host and device disagreeing about the arrival count) must not hang
        // the GPU: after kBarrierWaitNs the kernel traps and the host sees a launch failure
        unsigned spins = 0;
        unsigned long long t0 = 0;
        while ((int)(ld_acquire_gpu(bar) - target) < 0) {
            if ((++spins & 0xfffu) == 0) {
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (t0 == 0) t0 = now;
                else if (now - t0 > kBarrierWaitNs) __trap();
            }
        }
        __threadfence();
    }
    __syncthreads();
}

// ---- multi-GPU exchange of the tick's tx records over NVLink peer memory ------------------------
// Every rank stores its records straight into every peer's gather buffer and then raises its flag
// there to the tick's epoch (release, system scope).  The consumer spins on its LOCAL flags.  A peer
// that never arrives (crashed rank) must not hang the GPU: after kPeerWaitNs the wait gives up and
// the tick is reported as failed.
constexpr int kMaxPeers = 16;
constexpr unsigned long long kPeerWaitNs = 20ull * 1000 * 1000 * 1000;
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned *p, unsigned v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
// a flag store BEHIND a __threadfence_system(): the fence has ordered the data, the store itself need not wait for
// anything -- seven release stores in a row wait for one another's acknowledgement over NVLink, seven of these pipeline
__device__ __forceinline__ void st_relaxed_sys(unsigned *p, unsigned v)
{
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ bool peer_wait(const unsigned *flag, unsigned epoch)
{
    if ((int)(ld_acquire_sys(flag) - epoch) >= 0) return true;
    const unsigned long long t0 = global_timer_ns();
    while ((int)(ld_acquire_sys(flag) - epoch) < 0)
        if (global_timer_ns() - t0 > kPeerWaitNs) return false;
    return true;
}

// everything the fused tx-side preparation kernel needs beyond the tx arrays
constexpr int kPrepThreads = 256;
constexpr int kPrepMaxGrid = 256;
struct PrepareAux {
    int *cnt; int n_work;             // near-candidate counters of the receivers (zeroed here)
    int *tx_cnt;                      // [n_tx+1] heard pairs per packet (zeroed here)
    int *rx_done; int n_rx_done;      // per-rx-block split counters of the sweep's fused epilogue (zeroed here)
    uint8_t *tx_failed; long long *tx_base;   // [n_tx] deterministic tx-failure rule (p_tx <= 0), draw base 0
    int *blk_cnt, *blk_thmin;         // [grid][kMaxGroups] per-block bucket counts / min thresholds
    unsigned *bar; unsigned bar_target;
    const unsigned *peer_flags; int world; unsigned epoch;   // peer_flags == nullptr: no exchange to wait for
    // exchange fused into this kernel (push_local != nullptr): store this rank's records into every peer's gather
    // buffer, raise the flags, then wait -- one kernel does the "all-gather" and the preparation that consumes it
    const rmb_tx_record *push_local; int push_slots, push_rank; size_t push_off; unsigned *push_done;
    char *push_peer[kMaxPeers];
    int stochastic;                   // the tick takes the draw path (records with 0 < p_tx < 1 are fine)
    // half duplex (rmb_config.half_duplex): tx_epoch[node] = hd_epoch for every local source of this tick,
    // hd_cnt[parity][group] = distinct local sources per receiver group; tx_epoch == nullptr: off
    int *tx_epoch; int hd_epoch; int *hd_cnt; const int *node_gid;
};

// SM count of the current device (queried once per device: this sits on the tick's launch path)
inline int device_sms()
{
    static int cache[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return 148;
    if (cache[dev] == 0) {
        int sms = 148;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) { sms = 148; cudaGetLastError(); }
        cache[dev] = sms;
    }
    return cache[dev];
}

// Kernels with a grid barrier are launched cooperatively (co-residency guaranteed).  Measured on B200 the
// cooperative launch costs nothing extra on the tick's critical path (471 vs 472 us on a 1/8 shard), so there is no
// reason to weaken the guarantee; RMB_PLAIN_LAUNCH=1 switches to plain launches (an experiment knob: it lets the
// barrier kernels of several handles run side by side on ONE device).
inline bool use_cooperative_launch()
{
    static int v = -1;
    if (v < 0) { const char *e = getenv("RMB_PLAIN_LAUNCH"); v = (e && e[0] == '1') ? 0 : 1; }
    return v == 1;
}
inline cudaError_t launch_grid_synced(const void *fn, dim3 grid, dim3 block, void **args, cudaStream_t st)
{
    return use_cooperative_launch() ? cudaLaunchCooperativeKernel(fn, grid, block, args, 0, st)
                                    : cudaLaunchKernel(fn, grid, block, args, 0, st);
}

// The same with dynamic shared memory and, optionally, programmatic dependent launch: the kernel's blocks may become
// resident as soon as every block of the kernel in front of it on the stream has executed griddepcontrol.launch_dependents
// (or exited); the kernel itself must execute griddepcontrol.wait before it touches what its predecessor wrote.
// RMB_NO_PDL=1 turns the attribute off (an experiment knob).
inline int pdl_mode()          // RMB_PDL: 0 off, 1 sweep behind the preparation kernel only (default), 2 also finish behind the sweep
{
    static int v = -1;
    if (v < 0) { const char *e = getenv("RMB_PDL"); v = (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 1; }
    return v;
}
inline bool use_pdl() { return pdl_mode() >= 1; }
inline cudaError_t launch_grid_synced_ex(const void *fn, dim3 grid, dim3 block, void **args, size_t smem, bool overlap, cudaStream_t st)
{
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (use_cooperative_launch()) { attr[na].id = cudaLaunchAttributeCooperative; attr[na].val.cooperative = 1; ++na; }
    if (overlap && pdl_mode() >= 2) { attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization; attr[na].val.programmaticStreamSerializationAllowed = 1; ++na; }
    cfg.attrs = attr; cfg.numAttrs = (unsigned)na;
    return cudaLaunchKernelExC(&cfg, fn, args);
}

#define RMB_CUDA_OK(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return e__; } while (0)

}  // namespace rmb
