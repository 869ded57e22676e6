// common.cuh -- shared types between the C-ABI layer (api.cu) and the kernels.
#pragma once
#include <cstdint>
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
    float4 *f;                        // {x', y', z', theta}: scaled centred coords, near threshold
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
    int pad;
};

// slot flags
constexpr uint8_t kHeard = 1, kDecodable = 2;

#define RMB_CUDA_OK(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return e__; } while (0)

}  // namespace rmb
