/*
 * medium_oracle.c -- TEST INFRASTRUCTURE ONLY (see medium_oracle.h / oracle/README.md).
 * Build: gcc -O2 -ffp-contract=off -pthread (oracle/Makefile).  No fast-math, no FMA.
 */
#include "medium_oracle.h"
#include "oracle_math.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

/* ------------------------------------------------------------------ java.util.Random */
#define JR_MULT 0x5DEECE66DULL
#define JR_ADD  0xBULL
#define JR_MASK ((1ULL << 48) - 1)

void orc_jrandom_init(orc_jrandom *r, int64_t seed) { r->seed = ((uint64_t)seed ^ JR_MULT) & JR_MASK; }

int32_t orc_jrandom_next(orc_jrandom *r, int bits)
{
    r->seed = (r->seed * JR_MULT + JR_ADD) & JR_MASK;
    return (int32_t)(int64_t)(r->seed >> (48 - bits));
}

double orc_jrandom_next_double(orc_jrandom *r)
{
    int64_t a = (int64_t)orc_jrandom_next(r, 26);
    int64_t b = (int64_t)orc_jrandom_next(r, 27);
    return (double)((a << 27) + b) * 0x1.0p-53;
}

int32_t orc_jrandom_next_int_bound(orc_jrandom *r, int32_t n)
{
    if ((n & -n) == n) return (int32_t)(((int64_t)n * (int64_t)orc_jrandom_next(r, 31)) >> 31);
    int32_t bits, val;
    do {
        bits = orc_jrandom_next(r, 31);
        val = bits % n;
    } while ((int32_t)((uint32_t)bits - (uint32_t)val + (uint32_t)(n - 1)) < 0);
    return val;
}

void orc_jrandom_fill_doubles(int64_t seed, double *out, size_t n)
{
    orc_jrandom r; orc_jrandom_init(&r, seed);
    for (size_t i = 0; i < n; i++) out[i] = orc_jrandom_next_double(&r);
}

/* ------------------------------------------------------------------ parameters */
void orc_params_default(orc_params *p, int32_t model)
{
    memset(p, 0, sizeof(*p));
    p->model = model;
    p->success_ratio_tx = 1.0;       /* UDGMRadioMedium.java:17 */
    p->success_ratio_rx = 1.0;       /* UDGMRadioMedium.java:18 */
    p->transmission_range = 50.0;    /* UDGMRadioMedium.java:23 */
    p->constant_range = 100.0;       /* UDGMConstantLossRadioMedium.java:8 */
    p->pl0_db = 40.05;
    p->pl_exponent = 3.0;
    p->d0_m = 1.0;
    p->sensitivity_dbm = -95.0;
    p->sinr_threshold_db = 4.0;
    p->noise_floor_dbm = -100.0;     /* AbstractRadioMedium.java:38 */
}

/* ------------------------------------------------------------------ geometry */
/* Position.java:56-64 -- dx = this - other, squares, unfused adds, sqrt */
double orc_distance(double x1, double y1, double z1, double x2, double y2, double z2)
{
    double dx = x1 - x2;
    double dy = y1 - y2;
    double dz = z1 - z2;
    dx = dx * dx;
    dy = dy * dy;
    dz = dz * dz;
    return sqrt(dx + dy + dz);
}

static double node_distance(const orc_nodes *nd, int32_t a, int32_t b)
{
    return orc_distance(nd->x[a], nd->y[a], nd->z[a], nd->x[b], nd->y[b], nd->z[b]);
}

/* UDGMRadioMedium.java:67-81.  Math.pow(v, 2.0) == v*v (fdlibm special case). */
double orc_udgm_rx_probability(const orc_params *p, const orc_nodes *nd, int32_t src, int32_t dst)
{
    double distance = node_distance(nd, src, dst);
    double distanceSquared = distance * distance;
    double distanceMax = p->transmission_range;
    if (distanceMax == 0.0) return 0.0;
    double distanceMaxSquared = distanceMax * distanceMax;
    double ratio = distanceSquared / distanceMaxSquared;
    if (ratio > 1.0) return 0.0;
    ratio = 1.0 - ratio * (1.0 - p->success_ratio_rx);
    return ratio * nd->rxprob[dst];
}

/* N2NRadioMedium.java:28-37 */
static double n2n_rx_probability(const orc_params *p, const orc_nodes *nd, int32_t src, int32_t dst)
{
    int32_t sid = nd->id_int ? nd->id_int[src] : src + 1;
    int32_t did = nd->id_int ? nd->id_int[dst] : dst + 1;
    if (p->matrix == NULL || sid <= 0 || did <= 0 || sid > p->matrix_n || did > p->matrix_n) return 0.0;
    return p->matrix[(size_t)(sid - 1) * (size_t)p->matrix_n + (size_t)(did - 1)] * nd->rxprob[dst];
}

/* ------------------------------------------------------------------ output helpers */
typedef struct {
    const double *u; size_t n, cur; int exhausted;
} ustream;

static double draw(ustream *s)
{
    if (s->cur >= s->n) { s->exhausted = 1; s->cur++; return 0.0; }
    return s->u[s->cur++];
}

typedef struct { orc_pair *out; size_t cap, n, needed; } sink;

static void emit(sink *s, int32_t tx, int32_t rx, double rssi, double sinr, int32_t outcome)
{
    if (s->n < s->cap) {
        orc_pair *q = &s->out[s->n++];
        q->tx = tx; q->rx = rx; q->rssi_dbm = rssi; q->sinr_db = sinr; q->outcome = outcome;
    }
    s->needed++;
}

/* receiver filter shared by all four Java media (UDGMRadioMedium.java:100-102 etc.).
 * txing (NULL unless orc_params.half_duplex): nodes that are the source of some packet of this tick.  The
 * reference excludes only node == source, so a node that is itself on the air still "receives"; with the
 * half_duplex spec flag (SURVEY.md 8(a) "open" row) such a node passes no filter during the tick. */
static int passes_filter(const orc_nodes *nd, int32_t node, const orc_packet *pk, const uint8_t *txing)
{
    if (txing && txing[node]) return 0;
    return node != pk->src && nd->enabled[node] && nd->channel[node] == pk->channel;
}

/* ------------------------------------------------------------------ reference media */
/* UDGMRadioMedium.transmit (java:83-116) and N2NRadioMedium.transmit (java:40-72):
 * identical control flow, different probability functions. */
static void stochastic_transmit(const orc_params *p, const orc_nodes *nd, const orc_packet *pk, int32_t t,
                                ustream *us, sink *sk, uint8_t *tx_failed, uint64_t *decisions, const uint8_t *txing)
{
    double txSuccess = (p->model == ORC_MODEL_UDGM)
        ? p->success_ratio_rx * nd->txprob[pk->src]      /* java:63-65 -- uses the Rx ratio (sic) */
        : nd->txprob[pk->src];                            /* N2N java:24-26 */
    int interference = 0;
    if (txSuccess <= 0.0 || (txSuccess < 1.0 && draw(us) > txSuccess)) interference = 1;
    if (tx_failed) tx_failed[t] = (uint8_t)interference;
    double rssi = pk->power_dbm;
    for (int32_t node = 0; node < nd->n; node++) {
        if (!passes_filter(nd, node, pk, txing)) continue;
        (*decisions)++;
        double rxSuccess = (p->model == ORC_MODEL_UDGM)
            ? orc_udgm_rx_probability(p, nd, pk->src, node)
            : n2n_rx_probability(p, nd, pk->src, node);
        if (rxSuccess <= 0.0) {
            /* not heard */
        } else if (interference || (rxSuccess < 1.0 && draw(us) > rxSuccess)) {
            emit(sk, t, node, rssi, NAN, ORC_INTERFERED);
        } else {
            emit(sk, t, node, rssi, NAN, ORC_DELIVERED);
        }
    }
}

/* UDGMConstantLossRadioMedium.transmit (java:16-36): strict '<', immediate delivery */
static void constant_transmit(const orc_params *p, const orc_nodes *nd, const orc_packet *pk, int32_t t,
                              sink *sk, uint64_t *decisions, const uint8_t *txing)
{
    double rssi = pk->power_dbm;
    for (int32_t node = 0; node < nd->n; node++) {
        if (!passes_filter(nd, node, pk, txing)) continue;
        (*decisions)++;
        double distance = node_distance(nd, pk->src, node);
        if (distance < p->constant_range) emit(sk, t, node, rssi, NAN, ORC_DELIVERED);
    }
}

/* NullRadioMedium.transmit (java:48-77) */
static void null_transmit(const orc_nodes *nd, const orc_packet *pk, int32_t t, sink *sk, uint64_t *decisions,
                          const uint8_t *txing)
{
    double rssi = pk->power_dbm;
    for (int32_t node = 0; node < nd->n; node++) {
        if (!passes_filter(nd, node, pk, txing)) continue;
        (*decisions)++;
        emit(sk, t, node, rssi, NAN, ORC_DELIVERED);
    }
}

/* ------------------------------------------------------------------ SINR medium (the spec) */
/*
 * SPEC (DESIGN.md section "SINR medium").  All double, no FMA.
 *   d      = Position.getDistance(src, rx)                       (Position.java:56-64)
 *   dd     = d < d0 ? d0 : d
 *   PL     = PL0 + (10*n) * log10(dd / d0)
 *   Prx    = Ptx - PL                                            [dBm]
 *   lin    = exp(Prx * ln10/10)                                  [mW]
 *   Itot_r = N0 + sum over co-channel packets t (arrival order, src(t) != r) of lin(t,r)
 *   heard(t,r)     iff filter(t,r) && Prx >= sensitivity && rxProbability[r] > 0
 *   Ioth           = Itot_r - lin ;  SINR = Ioth > 0 ? Prx - 10*log10(Ioth) : +inf
 *   decodable(t,r) iff SINR >= threshold
 * Stochastic rule = the reference's (UDGMRadioMedium.java:85-92,104-111) with
 *   p_tx = txProbability[src], p_rx = rxProbability[r]:
 *   tx fails   iff p_tx <= 0 || (p_tx < 1 && U > p_tx)
 *   interfered iff failed || !decodable || (p_rx < 1 && U > p_rx)   (short-circuit, so the
 *   draw happens only for decodable pairs of non-failed packets with 0 < p_rx < 1)
 * Draw order: packets in arrival order; per packet first the tx draw, then receivers in
 * node-array order.  The RSSI handed on is Prx (the reference hands on Ptx unattenuated).
 */
double orc_sinr_rx_power_dbm(const orc_params *p, double power_dbm, double distance)
{
    double dd = distance < p->d0_m ? p->d0_m : distance;
    double pl = p->pl0_db + (10.0 * p->pl_exponent) * orc_log10(dd / p->d0_m);
    return power_dbm - pl;
}

typedef struct { int32_t tx, rx; double prx, sinr; int32_t decodable; } cand;
typedef struct { cand *v; size_t n, cap; } candvec;

static void cv_push(candvec *c, cand x)
{
    if (c->n == c->cap) {
        c->cap = c->cap ? c->cap * 2 : 256;
        c->v = (cand *)realloc(c->v, c->cap * sizeof(cand));
    }
    c->v[c->n++] = x;
}

static void sinr_receiver(const orc_params *p, const orc_nodes *nd, const orc_packet *pk, int32_t T,
                          int32_t r, double n0, candvec *cv, double *total_dbm, uint64_t *decisions, const uint8_t *txing)
{
    if (!nd->enabled[r]) { if (total_dbm) total_dbm[r] = p->noise_floor_dbm; return; }
    size_t first = cv->n;
    double itot = n0;
    uint64_t dec = 0;
    for (int32_t t = 0; t < T; t++) {
        if (!passes_filter(nd, r, &pk[t], txing)) continue;
        dec++;
        double d = node_distance(nd, pk[t].src, r);
        double prx = orc_sinr_rx_power_dbm(p, pk[t].power_dbm, d);
        double lin = orc_dbm_to_mw(prx);
        itot = itot + lin;
        if (prx >= p->sensitivity_dbm && nd->rxprob[r] > 0.0) {
            cand c; c.tx = t; c.rx = r; c.prx = prx; c.sinr = lin; /* lin parked in .sinr */ c.decodable = 0;
            cv_push(cv, c);
        }
    }
    for (size_t i = first; i < cv->n; i++) {
        double lin = cv->v[i].sinr;
        double ioth = itot - lin;
        double sinr = ioth > 0.0 ? cv->v[i].prx - orc_mw_to_dbm(ioth) : INFINITY;
        cv->v[i].sinr = sinr;
        cv->v[i].decodable = sinr >= p->sinr_threshold_db;
    }
    if (total_dbm) total_dbm[r] = orc_mw_to_dbm(itot);
    *decisions += dec;
}

typedef struct {
    const orc_params *p; const orc_nodes *nd; const orc_packet *pk; int32_t T; double n0;
    candvec *cvs; uint64_t *decs; double *total_dbm; int nchunk; int next; const uint8_t *txing;
} sinr_job;

/* receivers are independent given the tick's packets: chunks of the node array are handed to
 * worker threads; each chunk's arithmetic is exactly the sequential one */
static void *sinr_worker(void *arg)
{
    sinr_job *j = (sinr_job *)arg;
    for (;;) {
        int c = __atomic_fetch_add(&j->next, 1, __ATOMIC_RELAXED);
        if (c >= j->nchunk) break;
        int32_t lo = (int32_t)((int64_t)j->nd->n * c / j->nchunk), hi = (int32_t)((int64_t)j->nd->n * (c + 1) / j->nchunk);
        for (int32_t r = lo; r < hi; r++)
            sinr_receiver(j->p, j->nd, j->pk, j->T, r, j->n0, &j->cvs[c], j->total_dbm, &j->decs[c], j->txing);
    }
    return NULL;
}

static void sinr_tick(const orc_params *p, const orc_nodes *nd, const orc_packet *pk, int32_t T,
                      ustream *us, sink *sk, uint8_t *tx_failed, double *total_dbm, int threads,
                      uint64_t *decisions, const uint8_t *txing)
{
    double n0 = orc_dbm_to_mw(p->noise_floor_dbm);
    int nchunk = threads > 1 ? threads * 8 : 1;
    if (nchunk > nd->n) nchunk = nd->n > 0 ? nd->n : 1;
    candvec *cvs = (candvec *)calloc((size_t)nchunk, sizeof(candvec));
    uint64_t *decs = (uint64_t *)calloc((size_t)nchunk, sizeof(uint64_t));
    sinr_job job = { p, nd, pk, T, n0, cvs, decs, total_dbm, nchunk, 0, txing };
    int nthr = threads > 1 ? threads : 1;
    if (nthr == 1) {
        sinr_worker(&job);
    } else {
        pthread_t *th = (pthread_t *)malloc((size_t)nthr * sizeof(pthread_t));
        for (int i = 0; i < nthr; i++) pthread_create(&th[i], NULL, sinr_worker, &job);
        for (int i = 0; i < nthr; i++) pthread_join(th[i], NULL);
        free(th);
    }
    /* rx-major candidate list -> stable counting sort by tx == (tx arrival, rx index) order */
    size_t total = 0;
    for (int c = 0; c < nchunk; c++) { total += cvs[c].n; *decisions += decs[c]; }
    size_t *off = (size_t *)calloc((size_t)T + 1, sizeof(size_t));
    for (int c = 0; c < nchunk; c++) for (size_t i = 0; i < cvs[c].n; i++) off[cvs[c].v[i].tx + 1]++;
    for (int32_t t = 0; t < T; t++) off[t + 1] += off[t];
    cand *sorted = (cand *)malloc((total ? total : 1) * sizeof(cand));
    size_t *cur = (size_t *)malloc(((size_t)T + 1) * sizeof(size_t));
    memcpy(cur, off, ((size_t)T + 1) * sizeof(size_t));
    for (int c = 0; c < nchunk; c++) for (size_t i = 0; i < cvs[c].n; i++) sorted[cur[cvs[c].v[i].tx]++] = cvs[c].v[i];
    /* stochastic decisions in Java draw order */
    for (int32_t t = 0; t < T; t++) {
        double ptx = nd->txprob[pk[t].src];
        int failed = 0;
        if (ptx <= 0.0 || (ptx < 1.0 && draw(us) > ptx)) failed = 1;
        if (tx_failed) tx_failed[t] = (uint8_t)failed;
        for (size_t i = off[t]; i < off[t + 1]; i++) {
            const cand *c = &sorted[i];
            double prx_p = nd->rxprob[c->rx];
            if (failed || !c->decodable || (prx_p < 1.0 && draw(us) > prx_p))
                emit(sk, t, c->rx, c->prx, c->sinr, ORC_INTERFERED);
            else
                emit(sk, t, c->rx, c->prx, c->sinr, ORC_DELIVERED);
        }
    }
    for (int c = 0; c < nchunk; c++) free(cvs[c].v);
    free(cvs); free(decs); free(off); free(sorted); free(cur);
}

/* ------------------------------------------------------------------ entry point */
void orc_tick(const orc_params *p, const orc_nodes *nodes, const orc_packet *packets, int32_t T,
              const double *uniforms, size_t n_uniforms, orc_pair *out, size_t cap, uint8_t *tx_failed,
              double *total_dbm, int threads, orc_stats *stats)
{
    ustream us = { uniforms, uniforms ? n_uniforms : 0, 0, 0 };
    sink sk = { out, out ? cap : 0, 0, 0 };
    uint64_t decisions = 0;
    uint8_t *txing = NULL;
    if (p->half_duplex) {
        txing = (uint8_t *)calloc((size_t)(nodes->n > 0 ? nodes->n : 1), 1);
        for (int32_t t = 0; t < T; t++) if (packets[t].src >= 0 && packets[t].src < nodes->n) txing[packets[t].src] = 1;
    }
    if (p->model == ORC_MODEL_SINR) {
        sinr_tick(p, nodes, packets, T, &us, &sk, tx_failed, total_dbm, threads, &decisions, txing);
    } else {
        for (int32_t t = 0; t < T; t++) {
            switch (p->model) {
            case ORC_MODEL_NULL:
                if (tx_failed) tx_failed[t] = 0;
                null_transmit(nodes, &packets[t], t, &sk, &decisions, txing); break;
            case ORC_MODEL_UDGM_CONSTANT:
                if (tx_failed) tx_failed[t] = 0;
                constant_transmit(p, nodes, &packets[t], t, &sk, &decisions, txing); break;
            default:
                stochastic_transmit(p, nodes, &packets[t], t, &us, &sk, tx_failed, &decisions, txing); break;
            }
        }
    }
    free(txing);
    if (stats) {
        stats->n_pairs = sk.n;
        stats->n_needed = sk.needed;
        stats->n_uniforms_used = us.cur;
        stats->n_decisions = decisions;
        stats->status = us.exhausted ? -2 : (sk.needed > sk.n && out ? -1 : 0);
    }
}
