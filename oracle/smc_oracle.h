/* TEST INFRASTRUCTURE ONLY -- CPU restatement (plain C) of the SMCSMC particle-filter hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library.  The product path
 * (smcsmc_b200/csrc, libsmcgpu.so) never includes, links or calls anything in oracle/.
 *
 * Parity status: PINNED.  The deterministic pieces (fastexp, pruning likelihood, branch length with data, systematic
 * resampling, normalisation/ESS, lagged tallies) are checked against the unmodified reference compiled from
 * /root/reference (oracle/_ref/libsmcref.so, see oracle/ref_harness.cpp) and against the committed fixtures in
 * tests/golden/ generated from it.  The stochastic pieces (prior tree, SMC' move, next recombination position) use a
 * Philox4x32-10 stream instead of the reference's mt19937_64, so they are pinned to the reference DISTRIBUTIONALLY
 * (tests/test_oracle_vs_reference.py) and the CUDA kernels are compared with this file draw for draw.
 *
 * Each function cites the reference file:line (under /root/reference/src) whose behaviour it restates.
 */
#ifndef SMC_ORACLE_H
#define SMC_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAXN 64          /* max haplotypes supported by the oracle */
#define ORC_MAXNODES (2 * ORC_MAXN - 1)
#define ORC_MAXE 64          /* max epochs */

/* Flat demographic model (single population in this round; see DESIGN.md "out of scope"). Times in generations. */
typedef struct {
  int n;                    /* haplotypes */
  int E;                    /* epochs */
  double mu;                /* mutation rate per bp per generation   (Model::mutation_rate) */
  double rho;               /* recombination rate per bp per generation (Model::recombination_rate) */
  double seqlen;            /* locus length in bp (Model::loci_length) */
  double t_start[ORC_MAXE]; /* epoch start times, t_start[0] = 0     (Model::change_times_) */
  double inv2N[ORC_MAXE];   /* 1/(2 N_e) per epoch                   (Model::inv_double_pop_size, model.h:175) */
  double vb_logfac[ORC_MAXE];/* log(exp_digamma(c)/c) per epoch, 0 when -vb is off (particle.cpp:65-74,266-272) */
  uint8_t rec_mask[ORC_MAXE];/* bit0: record recombination, bit1: record coal/migr (pfparam.cpp:82-99) */
} orc_model;

/* One particle: flat local tree + scalars.  Leaves are slots 0..n-1 (height 0), internal nodes n..2n-2 stored in
 * increasing height order; parent[root] = -1. */
typedef struct {
  int parent[ORC_MAXNODES];
  double height[ORC_MAXNODES];
  double logw;        /* log posterior weight (reference keeps linear weights, particle.hpp:237) */
  double updated_to;  /* site_where_weight_was_updated_ (particle.hpp:236) */
  double cur_base;    /* Forest::current_base(): position of the last recombination */
  double next_base;   /* Forest::next_base(): position of the next recombination */
} orc_particle;

/* Tally records emitted while a particle is extended (the content of the reference's EvolutionaryEvent records,
 * coalevent.hpp:71-369, without the linked tree):
 *   ORC_EV_COAL   coalescence opportunity opp = k*dt of the floating lineage in `epoch`, stamped at position x0 (= x1);
 *                 cnt = 1 if the interval ended in a coalescence          (particle.cpp:251-300)
 *   ORC_EV_RECOPP recombination opportunity per bp (= local branch length inside `epoch`) over [x0, x1); opp is the
 *                 RATE per bp, the opportunity is opp*(x1-x0)            (particle.cpp:305-357)
 *   ORC_EV_REC    recombination event at position x0 with height in `epoch`, cnt = 1   (particle.cpp:360-377) */
enum { ORC_EV_COAL = 0, ORC_EV_RECOPP = 1, ORC_EV_REC = 2 };
typedef struct {
  int type, epoch, cnt;
  double x0, x1, opp;
} orc_event;
typedef struct {
  orc_event* ev;
  int count, cap;      /* events beyond cap are dropped and `overflow` set */
  int overflow;
} orc_events;

/* A tally record in channel form (smc_tally_oracle.c: orc_channel_tallies): POINT (rect = 0) adds val at position x0,
 * RECT (rect = 1) adds val per bp over [x0, x1). */
typedef struct { int row, slot, epoch, chan, rect; double x0, x1, val; } orc_chan_event;
void orc_channel_tallies(int E, int Q, const double* lags, int R, int N, const double* row_end, const double* weights,
                         const int* anc, const double* w_final, double seq_end, const orc_chan_event* evs, long nev,
                         double* out);

/* ---- deterministic pieces ---- */
double orc_fastexp(double x);                                               /* particle.cpp:30-40 */
double orc_exp_digamma(double x);                                           /* particle.cpp:65-74 */
double orc_prune_likelihood(int n, const int* parent, const double* height, const int* hap, double mu,
                            int ancestral_aware);                           /* particle.cpp:625-680 */
double orc_site_likelihood(int n, const int* parent, const double* height, const int* data, double mu,
                           int ancestral_aware, int dephase);               /* particleContainer.cpp:138-224 */
double orc_tree_length(int n, const int* parent, const double* height);     /* forest.h:188 */
double orc_branch_length_with_data(int n, const int* parent, const double* height, const int* data); /* particle.cpp:699-730 */
void orc_epoch_tree_length(const orc_model* m, const int* parent, const double* height, double* len_e);

void orc_systematic_resampling(const double* partial_sum, int R, int N, double u, int* counts); /* particleContainer.cpp:474-504 */
/* normalise linear weights in place, returns log(total)  (particleContainer.cpp:420-438) */
double orc_normalize(double* w, int N);
/* ESS and partial sums over pilot weights (particleContainer.cpp:256-270); partial_sum has N+1 entries */
double orc_ess_partial_sums(const double* w, int N, double* partial_sum);
/* log-domain versions used by the CUDA path: lse normalisation of log-weights; returns log total */
double orc_normalize_log(double* logw, int N);

/* ---- Philox4x32-10 counter-based stream (replaces mersenne_twister.cc / random_generator.cc) ---- */
typedef struct {
  uint32_t key[2];
  uint32_t ctr[4];   /* ctr[0] = block counter, ctr[1] = slot, ctr[2] = step, ctr[3] = chunk | purpose<<24 */
  double buf[2];
  int have;
} orc_rng;
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void orc_rng_init(orc_rng* r, uint64_t seed, uint32_t slot, uint32_t step, uint32_t chunk, uint32_t purpose);
double orc_rng_uniform(orc_rng* r);   /* in (0,1), 53 bits */

/* ---- stochastic pieces ---- */
void orc_init_particle(const orc_model* m, orc_particle* p, orc_rng* r, double start_pos); /* forest.cc:299-347 */
/* One SMC' move at position x (forest.cc:423-598 under window 0); records coal tallies + recombination event. */
void orc_smc_move(const orc_model* m, orc_particle* p, orc_rng* r, int max_rec_epoch, orc_events* t,
                  double* rec_height_out);
/* Extend one particle over a row ending at x1 with alleles data[n] (particle.cpp:743-918) and, if is_site, apply the
 * site likelihood (particleContainer.cpp:187-224).  Tally records of the row are appended to t (may be NULL). */
void orc_extend_row(const orc_model* m, orc_particle* p, orc_rng* r, double x1, const int* data, int is_site,
                    int ancestral_aware, int max_rec_epoch, orc_events* t);

#ifdef __cplusplus
}
#endif
#endif
