/* smcsmc_b200 -- C ABI of the B200-native SMCSMC particle-filter engine (libsmcgpu.so).
 *
 * The reference (luntergroup/smcsmc) has no in-process plugin API: its hot path is reached through the C++ classes
 * ParticleContainer / ForestState / CountModel driven by pfARG_core (src/smcsmc.cpp:278-400).  This header is the
 * boundary a maintainer binds instead of those classes; every entry point cites the reference call it replaces
 * (paths relative to /root/reference/src).  INTEGRATION.md shows the replacement of pfARG_core's loop body.
 *
 * Conventions: plain C, no C++ types, no exceptions across the boundary.  All pointers are HOST memory; the library
 * owns every device allocation and does the host<->device copies itself.  Functions return 0 on success and a
 * non-zero smc_status otherwise; smc_last_error() returns the message of the last failure on the calling thread.
 * One engine per chunk; an engine is not thread-safe (the reference is single-threaded as well).  Times are in
 * generations, positions in bp, rates per bp per generation -- the units of scrm's Model (scrm/src/model.h:175-188).
 * There is NO CPU fallback: every compute entry point fails with SMC_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef SMCSMC_B200_H
#define SMCSMC_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define SMC_MAX_HAPLOTYPES 32   /* n: local trees are flat arrays of 2n-1 nodes */
#define SMC_MAX_EPOCHS 64
#define SMC_MAX_POPULATIONS 4

typedef enum {
  SMC_OK = 0,
  SMC_ERR_ARG = 1,       /* invalid argument (message says which) */
  SMC_ERR_CUDA = 2,      /* CUDA runtime / no usable device */
  SMC_ERR_STATE = 3,     /* call out of order (e.g. sweep before rows were uploaded) */
  SMC_ERR_NUMERIC = 4    /* zero or non-finite total weight (ParticleContainer::normalize_probability throws,
                            particleContainer.cpp:428-429) */
} smc_status;

typedef struct smc_engine smc_engine;

/* Flattened Model + PfParam of one sweep (one population; smc_set_populations adds structure).
 * Replaces: Model (scrm/src/model.h:55), PfParam (pfparam.hpp), CountModel::lags (count.hpp). */
typedef struct {
  int32_t n;                /* -nsam: haplotypes, 2..SMC_MAX_HAPLOTYPES */
  int32_t N;                /* -Np: particles held by THIS engine (a shard when particles are split over GPUs) */
  int32_t E;                /* number of epochs (Model::change_times_.size()) */
  int32_t device;           /* CUDA device ordinal */
  uint64_t seed;            /* -seed; Philox key */
  uint32_t chunk;           /* RNG sub-stream id of this chunk (24 bits) */
  uint32_t slot_offset;     /* global index of this engine's first particle (RNG slot = slot_offset + local index) */
  double mu;                /* Model::mutation_rate() */
  double rho;               /* Model::recombination_rate() */
  double seqlen;            /* Model::loci_length() */
  double start_pos;         /* initial_position of ParticleContainer (particleContainer.cpp:37) */
  double ess_fraction;      /* -ESS (pfparam.cpp:323: ESSthreshold = N * ESS_fraction) */
  int32_t ancestral_aware;  /* -ancestral_aware */
  int32_t parity_scan;      /* 1: order-faithful sequential partial sums (particleContainer.cpp:256-262) so that
                               offspring counts are bit-exact given the weights; 0: blocked parallel scan */
  const double* t_start;    /* [E] epoch start times, t_start[0] = 0 */
  const double* inv2N;      /* [E] 1/(2 N_e)  (Model::inv_double_pop_size) */
  const double* vb_logfac;  /* [E] log(exp_digamma(c)/c) per epoch or NULL  (particle.cpp:65-74,266-272) */
  const uint8_t* rec_mask;  /* [E] bit0 record recombination, bit1 record coalescence (pfparam.cpp:82-99) or NULL=3 */
  const double* lags;       /* [E] CountModel::lags (count.cpp:230-264) */
} smc_config;

/* ---- engine life cycle -------------------------------------------------------------------------------------- */
int smc_create(const smc_config* cfg, smc_engine** out);   /* ParticleContainer ctor minus the trees */
void smc_destroy(smc_engine* e);
const char* smc_last_error(void);
const char* smc_version(void);

/* Structured populations (scrm -I / -ej / -em / -eM / -ema / -en; Model::population_number() > 1).  Call after
 * smc_create and before smc_upload_rows; calling it again between sweeps (after smc_reset) replaces the rates, which is
 * what the M-step does.  With it the engine runs the structured transition: the floating lineage coalesces with the
 * contemporaries of its own population, migrates at rate sum_j mig[e][i][j], is moved by the fixed-time events at epoch
 * starts, and above the old root both lineages float (forest.cc:496-756, 1100-1175).
 *   npop                 2..SMC_MAX_POPULATIONS
 *   inv2N[E][npop]       Model::inv_double_pop_size(pop, t) per epoch   (cfg.inv2N is ignored from here on)
 *   mig[E][npop][npop]   Model::migration_rate(i, j): per-generation rate of a lineage in i moving to j, backwards in
 *                        time; diagonal ignored; NULL = no migration
 *   single_mig[E][npop][npop]  Model::single_mig_pop(i, j): probability that a lineage in i is moved to j at the START
 *                        of the epoch (-ej t i j sets [i][j] = 1); NULL = none
 *   leaf_pop[n]          Model::sample_population
 *   max_marks            capacity of the per-particle pool of migration marks (<= 120), 0 = 2n + 8; a particle that
 *                        needs more makes smc_get_lnZ fail
 * Tallies then have 2Q rows, Q = 2 + 3*npop + npop^2 (smc_tally_rows): rows [0,Q) are sums of w * cell, rows [Q,2Q)
 * sums of w^2 * cell (the reference's *_weight arrays, count.cpp:519-526), cells in the order
 *   0 recombination count, 1 recombination opportunity, 2+3p coalescence count in p, 3+3p coalescence opportunity in p,
 *   4+3p migration opportunity in p, 2+3*npop + p*npop + q migration count p -> q
 * (CountModel::total_coal_count / total_coal_opportunity / total_mig_count / total_mig_opportunity, count.hpp).
 * Combinable with smc_shard_attach in either order (both before smc_upload_rows; blocks of 64 or 128 particles).
 * cfg.vb_logfac is not used by the structured move: see smc_set_population_vb. */
int smc_set_populations(smc_engine* e, int32_t npop, const double* inv2N, const double* mig, const double* single_mig,
                        const int32_t* leaf_pop, int32_t max_marks);
/* -vb with populations (particle.cpp:262-272): the log of exp_digamma(c) / c, c = Model::coalescent_event_count(pop) resp.
 * Model::migration_event_count(from, to) of the epoch, is added to a particle's log weight at every coalescence
 * (pairwise ones included) resp. migration of its moves.  coal_logfac[E][npop], mig_logfac[E][npop][npop]; NULL = zeros.
 * Call after smc_set_populations, between sweeps. */
int smc_set_population_vb(smc_engine* e, const double* coal_logfac, const double* mig_logfac);
/* rows of the array smc_get_tallies fills: 6 for a single population, 2Q after smc_set_populations */
int smc_tally_rows(smc_engine* e);
/* populations and migration marks of the current particles (tests): pop[N][2n-1] population at the bottom of the branch
 * above each node, nmarks[N], and per particle max_marks entries (branch, time, population above the mark), unused
 * entries -1 / 0 / -1.  Any pointer may be NULL. */
int smc_get_population_state(smc_engine* e, int32_t* pop, int32_t* nmarks, int32_t* mark_branch, double* mark_time,
                             int32_t* mark_pop);

/* Upload the rows of the .seg chunk (Segment, segdata.cpp:55-223) and derive the lagged-tally schedule
 * (CountModel::extract_and_update_count's per-epoch update positions, count.cpp:363-385).
 *   x_end[R]            segment_end() of each row, already clipped to seqlen (particleContainer.cpp:452)
 *   alleles[R*n]        allelic_state_at_Segment_end: 0, 1, -1 missing, 2 unphased het (pairs)
 *   is_site[R]          1 when segment_state == SEGMENT_INVARIANT (the row ends in a typed site)
 *   max_rec_epoch[R]    max_epoch_to_update (smcsmc.cpp:266-275) or NULL (= E-1 everywhere) */
int smc_upload_rows(smc_engine* e, const double* x_end, const int8_t* alleles, const uint8_t* is_site,
                    const int32_t* max_rec_epoch, int64_t R);

/* N prior trees at start_pos + first recombination positions + weights 1/N
 * (ParticleContainer ctor particleContainer.cpp:33-65 -> Forest::buildInitialTree forest.cc:299-347). */
int smc_init_particles(smc_engine* e);

/* The hot loop of pfARG_core (smcsmc.cpp:328-361) for rows [row_begin, row_end): per row
 *   extend_ARGs + update_weight_at_site + normalize_probability   (particleContainer.cpp:441-466)
 *   CountModel::extract_and_update_count                          (count.cpp:355-415)
 *   ParticleContainer::resample                                   (particleContainer.cpp:247-311)
 * Returns when the rows are done (the launches themselves queue back to back without host round trips). */
int smc_sweep(smc_engine* e, int64_t row_begin, int64_t row_end);

/* Sweep ALL remaining rows of `count` engines (one per chunk; the reference runs chunks as independent processes,
 * smcsmc/model.py:563-662,1094-1100).  Launches are issued round-robin, rows_per_turn rows of each engine at a time, on
 * the engines' own streams, so the chunks' row kernels run concurrently on the device.  Returns when all are done; the
 * device time of the whole group is booked on engines[0] (smc_get_timing). */
int smc_sweep_many(smc_engine** engines, int32_t count, int64_t rows_per_turn);

/* Particle sharding over GPUs (no counterpart in the reference, whose only parallelism is chunks): the
 * population of world * N particles is split evenly, one engine (one process, one GPU) per rank, each created with its
 * local N.  Rank 0 obtains an id, every rank attaches with it after smc_create and before smc_upload_rows (all ranks
 * upload the same rows).  From then on smc_upload_rows, smc_init_particles, smc_sweep, smc_finish, smc_get_particles and
 * smc_reset are collective: every rank calls them with the same arguments.  Per row the ranks exchange four doubles
 * INSIDE the row kernel (each rank stores them and a flag into every peer's mailbox over NVLink and waits for its own
 * to fill: one launch per row, no host round trip, no NCCL call on the data path); offspring whose parent lives on
 * another rank read its state straight from that GPU's memory (CUDA IPC mappings over NVLink).  The NCCL communicator
 * -- used to exchange the IPC handles and for the one all-reduce of smc_finish -- and the peers' mappings are cached per
 * process and id: engines attached with the same id share them, so an engine per EM iteration costs an allocation.  Philox streams are keyed by the global particle slot, so the result does not
 * depend on the number of ranks beyond the order of floating-point sums in the normaliser.  After smc_finish every
 * rank holds the global tallies, ln normaliser, resampling log and counters; smc_get_particles returns the local slice,
 * smc_get_last_ancestry global parent indices.  Needs libnccl.so.2 at run time (bound with dlopen). */
int smc_shard_unique_id(uint8_t id[128]);
int smc_shard_attach(smc_engine* e, int32_t rank, int32_t world, const uint8_t id[128]);

/* Start another EM iteration on the same rows (the reference re-enters pfARG_core, smcsmc.cpp:83): clears particles,
 * tallies, ln normaliser and counters; optionally replaces the rates the M-step changed (CountModel::reset_Ne /
 * reset_recomb_rate, count.cpp:267-325).  inv2N / vb_logfac may be NULL (unchanged); rho <= 0 means unchanged.
 * The lags (and with them the tally schedule) stay as uploaded. */
int smc_reset(smc_engine* e, const double* inv2N, double rho, const double* vb_logfac, uint64_t seed);

/* End of data: final normalisation and flush of all pending tallies with lag 0 (smcsmc.cpp:371-373). */
int smc_finish(smc_engine* e);

/* ---- results ------------------------------------------------------------------------------------------------ */
/* out[6*E]: coal_count, coal_opportunity, coal_weight, recomb_count, recomb_opportunity, recomb_weight per epoch,
 * WITHOUT CountModel's initial pseudo-counts (count.cpp:161-193).  *_weight is sum(W^2 * opportunity) with W the weight
 * pooled over the particles that share the record (count.cpp:519-526): pooled per ancestor slot of every closed lag
 * bucket, per particle for the bucket still open at the end of the data and for particle-sharded populations (an upper
 * bound of the resulting ESS).  The *_weight rows are complete after smc_finish (the second moments of a
 * row's buckets are booked one launch later and added up at the end).  Environment SMC_NO_POOL=1: per-particle squares. */
int smc_get_tallies(smc_engine* e, double* out);
/* out[R]: per .seg row the posterior-mean recombination opportunity sum_i w_i * integral over the row of the local tree
 * length of particle i (generations x bp), weighted with the particles' weights right after the row -- the
 * opportunity column of the local recombination map (CountModel::record_local_recomb_events, count.cpp:559-592; the
 * reference weights every stretch with the pooled weight of its descendants a lag later: same expectation per locus up to
 * the difference between the filtering and the fixed-lag smoothing posterior).  The driver spreads it over 100-bp bins
 * into <prefix>.recomb.gz (CountModel::dump_local_recomb_logs, count.cpp:616-654).  Rows not swept yet read 0. */
int smc_get_row_opportunity(smc_engine* e, double* out);
int smc_get_lnZ(smc_engine* e, double* lnZ);                       /* ParticleContainer::ln_normalization_factor */
/* counters[0] rows swept, [1] resampling events, [2] recombinations sampled, [3] numeric-error flag */
int smc_get_counters(smc_engine* e, int64_t counters[4]);
/* (position, ESS) of resampling events as PfParam::append_resample_file (pfparam.cpp:530-538); returns count via *n */
int smc_get_resample_log(smc_engine* e, double* position, double* ess, int64_t cap, int64_t* n);
/* particle state, AoS on the host side: parent[N*(2n-1)], height[N*(2n-1)], logw[N] (normalised), next_base[N] */
int smc_get_particles(smc_engine* e, int32_t* parent, double* height, double* logw, double* next_base);
int smc_set_particles(smc_engine* e, const int32_t* parent, const double* height, const double* logw,
                      const double* next_base);
/* ancestry of the last resampling event: anc[j] = slot (before) that slot j (after) was copied from; -1 if none yet */
int smc_get_last_ancestry(smc_engine* e, int32_t* anc);
/* device time spent in the engine's kernels since creation (CUDA events on the engine's stream), and launches */
int smc_get_timing(smc_engine* e, double* kernel_ms, int64_t* launches);

/* ---- stand-alone entry points (function-level parity; host buffers in, host buffers out) ---------------------- */
/* ForestState::calculate_likelihood over phasings as ParticleContainer::update_weight_at_site
 * (particle.cpp:625-680, particleContainer.cpp:138-224): lik[count], linear domain like the reference. */
int smc_site_likelihood(int device, int32_t n, int64_t count, const int32_t* parent, const double* height,
                        const int8_t* alleles, double mu, int32_t ancestral_aware, double* lik);
/* ForestState::trackLocalTreeBranchLength and Forest::getLocalTreeLength (particle.cpp:699-730, forest.h:188) */
int smc_branch_lengths(int device, int32_t n, int64_t count, const int32_t* parent, const double* height,
                       const int8_t* alleles, double* with_data, double* total);
/* ParticleContainer::systematic_resampling (particleContainer.cpp:474-504): counts[N] from partial_sum[N+1] and the
 * uniform u; bit-exact with the reference. */
int smc_systematic_offspring(int device, const double* partial_sum, int64_t N, double u, int32_t* counts);
/* resample (particleContainer.cpp:247-311) from weights: partial sums (sequential when parity_scan) + ESS + counts */
int smc_resample_counts(int device, const double* w, int64_t N, double u, int32_t parity_scan, double* ess,
                        int32_t* counts);
/* normalize_probability in log space (particleContainer.cpp:420-438): logw -= log(sum exp(logw)); returns lnZ, ESS */
int smc_normalize_log(int device, double* logw, int64_t N, double* lnZ, double* ess);
/* Lag calibration: median survival distance (bp) of local-tree nodes per epoch under the prior, from n_args
 * unconditioned ARGs simulated on the device (calculate_median_survival_distances, smcsmc.cpp:169-263, including its
 * fall-backs for sparsely sampled epochs).  CountModel::reset_lag then sets lag = lag_fraction * median. */
int smc_calibrate_survival(int device, int32_t n, int32_t E, const double* t_start, const double* inv2N, double rho,
                           double seqlen, uint64_t seed, int64_t n_args, double* median_survival, int64_t* n_events);
/* The same calibration under a structured model (tables as in smc_set_populations): the reference forward-simulates with
 * whatever Model the command line describes (smcsmc.cpp:169-189), populations and migration included. */
int smc_calibrate_survival_struct(int device, int32_t n, int32_t E, const double* t_start, int32_t npop,
                                  const double* inv2N, const double* mig, const double* single_mig,
                                  const int32_t* leaf_pop, double rho, double seqlen, uint64_t seed, int64_t n_args,
                                  double* median_survival, int64_t* n_events);
/* Terminal branch length quantiles of prior trees (calculate_terminal_branch_length_quantiles, smcsmc.cpp:128-167; the
 * reference computes them at the start of every pfARG_core for the auxiliary particle filter's look-ahead): n_trees
 * prior trees on the device (the reference draws 10^6), per sample the height of its parent coalescence ignoring
 * migrations, lengths[n][nq] = the sorted values at index (size_t)(quantiles[k] * n_trees), and the mean tree length.
 * npop < 2: single population (inv2N[E]); else the tables of smc_set_populations. */
int smc_terminal_branch_quantiles(int device, int32_t n, int32_t E, const double* t_start, int32_t npop,
                                  const double* inv2N, const double* mig, const double* single_mig,
                                  const int32_t* leaf_pop, uint64_t seed, int64_t n_trees, int32_t nq,
                                  const double* quantiles, double* lengths, double* mean_total_length);
/* Measurement helper (no reference counterpart): the device's sustained fp64 FMA rate in TFLOP/s from a register-only
 * probe kernel; bench.py quotes the row kernel's fp64 work against it beside the HBM roofline. */
int smc_measure_fp64_fma(int device, double* tflops);
/* fastexp (particle.cpp:30-40) evaluated on the device, for the 1e-12 no-mutation-factor parity test */
int smc_fastexp(int device, const double* x, int64_t count, double* out);

/* ---- allocation-free entry point ------------------------------------------------------------------------------- */
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same computation as
 * smc_resample_counts (partial sums -- sequential when parity_scan -- then ESS and the bit-exact systematic offspring
 * counts), but
 *   - d_w[N], d_ess[1] and d_counts[N] are DEVICE pointers on `device`, supplied by the caller;
 *   - the scratch (u-chain control block + N+1 partial sums) is a file-scope static __device__ array of
 *     (SMC_NOALLOC_MAX_N + 1) * 8 bytes plus a ~6 KB control block (about 1.06 MB) that every GPU loading the library
 *     holds until exit; N > SMC_NOALLOC_MAX_N is rejected with SMC_ERR_ARG;
 *   - the work is queued on `stream` (a cudaStream_t passed as void*, NULL = the default stream) and the call returns
 *     without synchronising; it calls no cudaMalloc / cudaFree / allocating library routine.
 * Every call on a device shares the one array (the control block is cleared on the stream at the start of each call):
 * calls that overlap on the GPU -- different streams, or graph replays in flight together -- race; serialise them.
 * The first call on a device resolves the array's address (cudaGetSymbolAddress); make it outside graph capture. */
#define SMC_NOALLOC_MAX_N 131072
int smc_resample_counts_noalloc(int device, const double* d_w, int64_t N, double u, int32_t parity_scan,
                                double* d_ess, int32_t* d_counts, void* stream);

#ifdef __cplusplus
}
#endif
#endif
