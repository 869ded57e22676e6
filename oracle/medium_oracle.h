/*
 * medium_oracle.h -- TEST INFRASTRUCTURE ONLY (see oracle/README.md).
 *
 * CPU restatement of the reference radio media (radio-medium/java/se/sics/emul8/radiomedium/)
 * plus the double-precision specification of the SINR medium the reference does not have.
 * Nothing under radio-sim_b200/ may include, link or call this.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
 *
 * Parity status: the reference has NO golden vectors or assertions for this path
 *  and no JVM exists here, so the reference-pinned media are pinned by
 * known-answer cases derived by hand from the Java source (tests/golden/kat_*.json);
 * the SINR medium is "parity unpinned" -- this file IS its specification.
 */
#ifndef MEDIUM_ORACLE_H
#define MEDIUM_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- java.util.Random (JDK spec: 48-bit LCG) -- Simulator.java:64,83-94 ---- */
typedef struct { uint64_t seed; } orc_jrandom;
void     orc_jrandom_init(orc_jrandom *r, int64_t seed);        /* new Random(seed)   */
int32_t  orc_jrandom_next(orc_jrandom *r, int bits);            /* protected next()   */
double   orc_jrandom_next_double(orc_jrandom *r);               /* nextDouble()       */
int32_t  orc_jrandom_next_int_bound(orc_jrandom *r, int32_t n); /* nextInt(n)         */
void     orc_jrandom_fill_doubles(int64_t seed, double *out, size_t n);

/* ---- node state as the media read it (Node/Position/Transciever.java) ---- */
typedef struct {
    int32_t        n;
    const double  *x, *y, *z;      /* Position.java:37-39                       */
    const int32_t *channel;        /* Transciever.wirelessChannel, default 26   */
    const uint8_t *enabled;        /* Transciever.isEnabled                     */
    const double  *rxprob;         /* Transciever.rxProbability, default 1.0    */
    const double  *txprob;         /* Transciever.txProbability, default 1.0    */
    const int32_t *id_int;         /* Node.getIdAsInteger (N2N only; may be NULL => index+1) */
} orc_nodes;

/* one RadioPacket (RadioPacket.java:46-52): source index into the node array,
 * snapshotted/overridden power and channel */
typedef struct {
    int32_t src;
    int32_t channel;
    double  power_dbm;
} orc_packet;

enum { ORC_NOT_HEARD = 0, ORC_INTERFERED = 1, ORC_DELIVERED = 2 };

/* one heard (tx,rx) pair == one generateReceptionEvents / deliverRadioPacket call */
typedef struct {
    int32_t tx;        /* index of the packet in this tick's arrival-ordered list */
    int32_t rx;        /* node-array index ("neighbour index")                    */
    double  rssi_dbm;  /* value handed to generateReceptionEvents                 */
    double  sinr_db;   /* SINR medium only; NaN for the reference media           */
    int32_t outcome;   /* ORC_INTERFERED / ORC_DELIVERED                          */
} orc_pair;

enum {
    ORC_MODEL_NULL = 0,            /* NullRadioMedium.java:48-77               */
    ORC_MODEL_UDGM = 1,            /* UDGMRadioMedium.java:63-116              */
    ORC_MODEL_UDGM_CONSTANT = 2,   /* UDGMConstantLossRadioMedium.java:16-36   */
    ORC_MODEL_N2N = 3,             /* N2NRadioMedium.java:24-72                */
    ORC_MODEL_SINR = 4             /* new: log-distance + interference + SINR  */
};

typedef struct {
    int32_t model;
    /* UDGM (UDGMRadioMedium.java:17-24) */
    double  success_ratio_tx;      /* 1.0; never read by transmit() (quirk)    */
    double  success_ratio_rx;      /* 1.0                                      */
    double  transmission_range;    /* 50.0                                     */
    /* ConstantLoss (UDGMConstantLossRadioMedium.java:8) */
    double  constant_range;        /* 100.0                                    */
    /* N2N */
    const double *matrix;          /* flat row-major n x n, may be NULL        */
    int32_t matrix_n;
    /* SINR */
    double  pl0_db;                /* 40.05                                    */
    double  pl_exponent;           /* 3.0                                      */
    double  d0_m;                  /* 1.0                                      */
    double  sensitivity_dbm;       /* -95.0                                    */
    double  sinr_threshold_db;     /* 4.0                                      */
    double  noise_floor_dbm;       /* -100.0 == AbstractRadioMedium.baseRSSI   */
    /* spec flag, all media (SURVEY.md 8(a) "open" row).  0 = the reference's behaviour: only node == source is
     * excluded (UDGMRadioMedium.java:100), a node that is itself transmitting still receives.  1: a node that is
     * the source of ANY packet of the tick passes no receiver filter during that tick (one transceiver). */
    int32_t half_duplex;
} orc_params;

void orc_params_default(orc_params *p, int32_t model);

typedef struct {
    size_t   n_pairs;              /* heard pairs written to out[]             */
    size_t   n_needed;             /* heard pairs that exist (>= n_pairs if cap hit) */
    size_t   n_uniforms_used;
    uint64_t n_decisions;          /* filter-passing (tx,rx) pairs evaluated   */
    int32_t  status;               /* 0 ok, -1 capacity, -2 uniforms exhausted */
} orc_stats;

/*
 * Evaluate one tick: packets[0..T) in arrival order, each processed as the Java
 * transmit() would (reference media: in isolation; SINR: jointly).  Heard pairs are
 * appended to out[] in (tx arrival, rx index) order == the Java call order of
 * generateReceptionEvents.  tx_failed[t] (may be NULL) receives the per-packet
 * "interference" flag (UDGMRadioMedium.java:88-92).  uniforms[] stands in for
 * simulator.getRandom().nextDouble(), consumed strictly in Java draw order.
 * total_dbm (may be NULL, SINR only): per node 10*log10(N0 + sum of co-channel power).
 * threads: 1 = sequential; >1 parallelises the SINR receiver loop (same arithmetic).
 */
void orc_tick(const orc_params *p, const orc_nodes *nodes,
              const orc_packet *packets, int32_t T,
              const double *uniforms, size_t n_uniforms,
              orc_pair *out, size_t cap, uint8_t *tx_failed,
              double *total_dbm, int threads, orc_stats *stats);

/* Position.getDistance (Position.java:56-64), exposed for known-answer tests */
double orc_distance(double x1, double y1, double z1, double x2, double y2, double z2);
/* UDGMRadioMedium.getRxSuccessProbability (UDGMRadioMedium.java:67-81) */
double orc_udgm_rx_probability(const orc_params *p, const orc_nodes *nodes, int32_t src, int32_t dst);
/* SINR spec pieces, exposed for unit tests */
double orc_sinr_rx_power_dbm(const orc_params *p, double power_dbm, double distance);

#ifdef __cplusplus
}
#endif
#endif
