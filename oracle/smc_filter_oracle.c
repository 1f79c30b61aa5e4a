/* TEST INFRASTRUCTURE ONLY -- see smc_oracle.h.
 *
 * The whole per-chunk particle filter on the CPU, built from the restated pieces: the loop body of pfARG_core
 * (smcsmc.cpp:328-361) = update_state_to_data (particleContainer.cpp:441-466) + extract_and_update_count
 * (count.cpp:355-415) + resample (particleContainer.cpp:247-392), then the end-of-data flush (smcsmc.cpp:371-373).
 * It consumes the SAME Philox streams as the CUDA engine (stream = (seed, particle slot, row, chunk, purpose)), so the
 * engine can be compared with it draw for draw: trees, weights, ln normaliser, resampling decisions and ancestry,
 * lagged tallies.  Physical copies replace the reference's multiplicity counter (SURVEY.md 7 "Multiplicity vs
 * physical copies"): a parent with k > 1 offspring gives k copies that each re-draw their next recombination position.
 */
#include "smc_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct { int row, slot; orc_event ev; } orc_rec_event;
void orc_lagged_tallies(int E, const double* lags, int R, int N, const double* row_end, const double* weights,
                        const int* anc, const double* w_final, double seq_end, const orc_rec_event* evs, long nev,
                        double* out);
void orc_redraw_next_base(const orc_model* m, orc_particle* p, orc_rng* r);

enum { PURPOSE_INIT = 1, PURPOSE_EXTEND = 2, PURPOSE_RESAMPLE = 3, PURPOSE_REDRAW = 4 };

typedef struct {
  int N;
  uint64_t seed;
  uint32_t chunk, slot_offset;
  double start_pos, ess_fraction;
  int ancestral_aware;
  const double* lags;     /* [E] */
} orc_pf_config;

/* Optional side output: out[R], per row sum_i w_i * (local tree length x bp covered in the row) with the weights right
 * after the row -- the opportunity the reference spreads over its local recombination map (count.cpp:559-587), here under
 * the filtering weights.  NULL switches it off.  Only meaningful with nothing masked (rec_mask 3, max_rec NULL). */
static double* g_row_opp_out = 0;
void orc_set_row_opportunity_buffer(double* out) { g_row_opp_out = out; }

/* returns 0, or 1 on overflow of the event buffer / zero total weight.
 * data[R][n] alleles; rs_pos/rs_ess sized R (may be NULL); last_anc[N] ancestry of the last APPLIED resampling. */
int orc_particle_filter(const orc_model* m, const orc_pf_config* c, int R, const double* row_end, const int* data,
                        const int* is_site, const int* max_rec, orc_particle* parts, double* lnZ_out, double* tallies,
                        int* n_resample_out, double* rs_pos, double* rs_ess, int* last_anc, long* n_recomb_out) {
  const int N = c->N, n = m->n, E = m->E;
  orc_rng r;
  orc_particle* next = (orc_particle*)malloc(sizeof(orc_particle) * (size_t)N);
  double* weights = (double*)malloc(sizeof(double) * (size_t)R * N);
  int* anc = (int*)malloc(sizeof(int) * (size_t)R * N);
  double* lw = (double*)malloc(sizeof(double) * (size_t)N);
  double* ps = (double*)malloc(sizeof(double) * (size_t)(N + 1));
  int* counts = (int*)malloc(sizeof(int) * (size_t)N);
  long cap = (long)R * N * (E + 6) + 1024, nev = 0;
  orc_rec_event* evs = (orc_rec_event*)malloc(sizeof(orc_rec_event) * (size_t)cap);
  static orc_event rowbuf[16384];
  int rc = 0, n_resample = 0;
  long n_recomb = 0;
  double lnZ = 0.0;
  for (int i = 0; i < N; i++) {
    orc_rng_init(&r, c->seed, c->slot_offset + (uint32_t)i, 0, c->chunk, PURPOSE_INIT);
    orc_init_particle(m, &parts[i], &r, c->start_pos);
    parts[i].logw = -log((double)N);
    if (last_anc) last_anc[i] = -1;
  }
  double* area = (double*)calloc((size_t)N, sizeof(double));
  for (int row = 0; row < R; row++) {
    for (int i = 0; i < N; i++) {
      orc_events t = { rowbuf, 0, 16384, 0 };
      area[i] = 0.0;
      orc_rng_init(&r, c->seed, c->slot_offset + (uint32_t)i, (uint32_t)row, c->chunk, PURPOSE_EXTEND);
      orc_extend_row(m, &parts[i], &r, row_end[row], data + (size_t)row * n, is_site[row], c->ancestral_aware,
                     max_rec ? max_rec[row] : E - 1, &t);
      if (t.overflow) { rc = 1; goto done; }
      if (nev + t.count > cap) {
        cap = cap * 2 + t.count;
        evs = (orc_rec_event*)realloc(evs, sizeof(orc_rec_event) * (size_t)cap);
      }
      for (int k = 0; k < t.count; k++) {
        evs[nev].row = row; evs[nev].slot = i; evs[nev].ev = t.ev[k]; nev++;
        if (t.ev[k].type == ORC_EV_REC) n_recomb++;
        if (t.ev[k].type == ORC_EV_RECOPP) area[i] += t.ev[k].opp * (t.ev[k].x1 - t.ev[k].x0);
      }
    }
    for (int i = 0; i < N; i++) lw[i] = parts[i].logw;
    double lse = orc_normalize_log(lw, N);
    if (!isfinite(lse)) { rc = 1; goto done; }
    lnZ += lse;
    for (int i = 0; i < N; i++) { parts[i].logw = lw[i]; weights[(size_t)row * N + i] = exp(lw[i]); }
    if (g_row_opp_out) {
      double acc = 0.0;
      for (int i = 0; i < N; i++) acc += weights[(size_t)row * N + i] * area[i];
      g_row_opp_out[row] = acc;
    }
    double ess = orc_ess_partial_sums(weights + (size_t)row * N, N, ps);
    anc[(size_t)row * N] = -1;
    if (ess < N * c->ess_fraction - 1e-6) {
      if (rs_pos) rs_pos[n_resample] = row_end[row];
      if (rs_ess) rs_ess[n_resample] = ess;
      n_resample++;
      if (row + 1 == R) break;     /* decided, but the sweep ends here: the flush uses the un-resampled set */
      orc_rng_init(&r, c->seed, 0, (uint32_t)row, c->chunk, PURPOSE_RESAMPLE);
      double u = orc_rng_uniform(&r);
      orc_systematic_resampling(ps, N, N, u, counts);
      int j = 0;
      for (int i = 0; i < N; i++)
        for (int k = 0; k < counts[i]; k++) {
          next[j] = parts[i];
          next[j].logw = log(ps[N] / (double)N);          /* particleContainer.cpp:350-351 */
          anc[(size_t)row * N + j] = i;
          if (counts[i] > 1) {                            /* particleContainer.cpp:357-367 */
            orc_rng_init(&r, c->seed, c->slot_offset + (uint32_t)j, (uint32_t)row, c->chunk, PURPOSE_REDRAW);
            orc_redraw_next_base(m, &next[j], &r);
          }
          j++;
        }
      memcpy(parts, next, sizeof(orc_particle) * (size_t)N);
      if (last_anc) memcpy(last_anc, anc + (size_t)row * N, sizeof(int) * (size_t)N);
    }
  }
  if (tallies)
    orc_lagged_tallies(E, c->lags, R, N, row_end, weights, anc, weights + (size_t)(R - 1) * N, m->seqlen, evs, nev,
                       tallies);
done:
  *lnZ_out = lnZ;
  if (n_resample_out) *n_resample_out = n_resample;
  if (n_recomb_out) *n_recomb_out = n_recomb;
  free(next); free(weights); free(anc); free(lw); free(ps); free(counts); free(evs); free(area);
  return rc;
}
