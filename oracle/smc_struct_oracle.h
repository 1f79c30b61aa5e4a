/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the STRUCTURED-population part of the SMCSMC hot path (several
 * populations, migration, population joins): the prior tree and the SMC' move of scrm's Forest with more than one
 * population (forest.cc:299-347,423-756,995-1175; time_interval.cc:60-141; contemporaries_container.h:249-379) and the
 * records ForestState::record_all_event writes for them (particle.cpp:193-302).
 *
 * Status: the checker of the structured CUDA path (smc_struct_device.cuh, tests/test_gpu_struct.py compares draw for
 * draw); the file itself is pinned against the unmodified reference DISTRIBUTIONALLY
 * (tests/test_struct_oracle.py: moments of ARG walks under structured models, reference side from
 * oracle/_ref/libsmcref.so, committed as tests/golden/struct_moments.npz).
 *
 * Representation (the one DESIGN.md section 10 plans for the device): the binary local tree exactly as in smc_oracle.h
 * (leaves 0..n-1 at height 0, internal slots height-sorted), plus pop[v] = population at the bottom of the branch
 * above v, plus a pool of migration marks (branch, time, population above the mark).  scrm stores a mark as a unary
 * node; nothing else in the path looks at it.
 */
#ifndef SMC_STRUCT_ORACLE_H
#define SMC_STRUCT_ORACLE_H
#include "smc_oracle.h"
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAXP 4
#define ORC_MAXMARKS 256

typedef struct {
  orc_model base;                                   /* n, E, mu, rho, seqlen, t_start, rec_mask (inv2N unused) */
  int P;                                            /* populations */
  double inv2N[ORC_MAXE][ORC_MAXP];                 /* Model::inv_double_pop_size(pop, t)        model.h:175 */
  double mig[ORC_MAXE][ORC_MAXP][ORC_MAXP];         /* Model::migration_rate(i, j), per generation, backwards i -> j */
  double single_mig[ORC_MAXE][ORC_MAXP][ORC_MAXP];  /* Model::single_mig_pop(i, j) at the START of the epoch (-ej/-es) */
  int leaf_pop[ORC_MAXN];                           /* Model::sample_population */
  /* -vb (particle.cpp:262-272): log(exp_digamma(c) / c) added to the particle's log weight at every coalescence in
   * (epoch, pop) resp. migration (epoch, from, to) of a move; all zero = no correction */
  double vb_coal[ORC_MAXE][ORC_MAXP];
  double vb_mig[ORC_MAXE][ORC_MAXP][ORC_MAXP];
} orc_smodel;

typedef struct {
  orc_particle tree;
  int pop[ORC_MAXNODES];
  int nmarks;
  int mark_branch[ORC_MAXMARKS];
  double mark_time[ORC_MAXMARKS];
  int mark_pop[ORC_MAXMARKS];
  int overflow;                                     /* more marks than ORC_MAXMARKS were needed */
} orc_sparticle;

/* records of one move, the content of the EvolutionaryEvents of particle.cpp:251-300:
 *   opportunity of lineage `who` (0 = the floating lineage, 1 = the old root's lineage once the sweep is above it)
 *   in (epoch, pop) over dt with coalescence weight `weight` (contemporaries, +1 for the pairwise term), ended by
 *   event 0 = none, 1 = coalescence, 2 = migration to `to_pop` */
typedef struct { int who, epoch, pop, weight, event, to_pop; double dt; } orc_srecord;
typedef struct { orc_srecord* rec; int count, cap; } orc_srecords;

int orc_spop_at(const orc_sparticle* p, int v, double t);
/* prior tree by sequential insertion of the leaves (Forest::buildInitialTree, forest.cc:299-347) + first next_base */
void orc_sinit_particle(const orc_smodel* m, orc_sparticle* p, orc_rng* r, double start_pos);
/* one SMC' move at tree.next_base (Forest::sampleNextGenealogy under window 0); rec_height_out may be NULL */
void orc_ssmc_move(const orc_smodel* m, orc_sparticle* p, orc_rng* r, orc_srecords* recs, double* rec_height_out);
/* consistency check used by the tests: 0 if the structure is a valid tree with consistent populations */
int orc_scheck(const orc_smodel* m, const orc_sparticle* p);

/* walks for the distribution-level tests; return 0 / the number of moves done, or an orc_scheck failure code */
int orc_sprior_stats(const orc_smodel* m, uint64_t seed, int count, double* tmrca, double* tree_len, double* marks);
int orc_sarg_walk(const orc_smodel* m, uint64_t seed, int count, double* pos, double* rec_h, double* tmrca, double* tlen,
                  double* marks);

/* the whole filter with populations (mirror of orc_particle_filter); tallies [2][Q][E], Q = orc_schannels(P) */
int orc_schannels(int P);
int orc_sparticle_filter(const orc_smodel* m, int N, uint64_t seed, uint32_t chunk, uint32_t slot_offset,
                         double start_pos, double ess_fraction, int ancestral_aware, const double* lags, int R,
                         const double* row_end, const int* data, const int* is_site, const int* max_rec,
                         orc_sparticle* parts, double* lnZ_out, double* tallies, int* n_resample_out, int* last_anc,
                         long* n_recomb_out);

#ifdef __cplusplus
}
#endif
#endif
