/* TEST INFRASTRUCTURE ONLY -- see smc_oracle.h.
 *
 * Restatement of the reference's fixed-lag, particle-weighted event tallies:
 *   CountModel::extract_and_update_count            count.cpp:355-415   (update schedule per epoch, lag, 0.1*lag rule)
 *   CountModel::update_all_counts                   count.cpp:448-491   (a record fires with the summed posterior
 *                                                                         weight of all particle records that share it)
 *   CountModel::update_all_counts_single_evolevent  count.cpp:495-555   (what is added, clipped to [counted_to,update_to))
 *   EvolutionaryEvent::*_between                    coalevent.hpp:209-235
 * The reference stores records in a ref-counted tree shared between a particle and its resampled copies; a record's
 * weight at update time is therefore the summed weight of the particles alive NOW that descend from the particle
 * that wrote it.  Here that is computed directly from the resampling ancestry (anc), with no linked structure.
 */
#include "smc_oracle.h"
#include <stdlib.h>
#include <string.h>

/* count.cpp:363-385: per-epoch update position for a sweep that has reached current_base.  Returns the first epoch
 * that updates (E if none).  counted_to is NOT modified. */
int orc_update_schedule(int E, const double* lags, const double* counted_to, double current_base, int end_data,
                        double* update_to) {
  int first = E;
  for (int e = 0; e < E; e++) {
    double lagging = end_data ? 0 : lags[e];
    double x_end = current_base - lagging;
    if ((x_end - counted_to[e]) < lagging * 0.1 && first > e) {
      update_to[e] = counted_to[e];
    } else {
      update_to[e] = x_end;
      if (e < first) first = e;
    }
  }
  return first;
}

/* smcsmc.cpp:266-275 max_epoch_to_update */
int orc_max_epoch_to_record(int E, const double* lags, double distance_to_mutation) {
  int epoch = 0;
  while (epoch < E && distance_to_mutation < 0.5 * lags[epoch]) epoch++;
  return epoch - 1;
}

typedef struct { int row, slot; orc_event ev; } orc_rec_event;

/* out: [6][E] = coal_count, coal_opp, coal_weight(w^2), rec_count, rec_opp, rec_weight(w^2) -- WITHOUT the initial
 * pseudo-counts of count.cpp:161-193.
 * rows 0..R-1: row_end[r], weights[r*N+i] = normalised posterior weight of slot i when row r's update runs
 * (after normalize_probability, before resampling); anc[r*N+i] = slot (in row r's indexing) that slot i of row r+1 was
 * copied from, or anc[r*N] = -1 when no resampling followed row r.  w_final = weights for the end-of-data flush
 * (smcsmc.cpp:371-373), in the indexing after the last row's resampling. */
void orc_lagged_tallies(int E, const double* lags, int R, int N, const double* row_end, const double* weights,
                        const int* anc, const double* w_final, double seq_end, const orc_rec_event* evs, long nev,
                        double* out) {
  double counted_to[ORC_MAXE], update_to[ORC_MAXE];
  for (int e = 0; e < E; e++) counted_to[e] = 0.0;
  memset(out, 0, sizeof(double) * 6 * E);
  double* W = (double*)malloc(sizeof(double) * (size_t)(R + 1) * N);   /* W[r][s] descendant weight */
  for (int step = 0; step <= R; step++) {
    int end_data = step == R;
    double cur = end_data ? seq_end : row_end[step];
    int first = orc_update_schedule(E, lags, counted_to, cur, end_data, update_to);
    if (first < E) {
      /* descendant-summed weights of every (row, slot) as seen from this step */
      const double* wnow = end_data ? w_final : weights + (size_t)step * N;
      int top = end_data ? R : step;          /* indexing of `top`: row `step`, or the virtual row after the last */
      for (int i = 0; i < N; i++) W[(size_t)top * N + i] = wnow[i];
      for (int r = top - 1; r >= 0; r--) {
        double* dst = W + (size_t)r * N;
        const double* src = W + (size_t)(r + 1) * N;
        if (anc[(size_t)r * N] < 0) { memcpy(dst, src, sizeof(double) * N); continue; }
        for (int i = 0; i < N; i++) dst[i] = 0.0;
        for (int i = 0; i < N; i++) dst[anc[(size_t)r * N + i]] += src[i];
      }
      for (long k = 0; k < nev; k++) {
        const orc_rec_event* re = &evs[k];
        int e = re->ev.epoch;
        if (e < first) continue;
        if (re->row > top) continue;
        if (!end_data && re->row > step) continue;
        double w = W[(size_t)re->row * N + re->slot];
        double xs = counted_to[e], xe = update_to[e];
        if (re->ev.type == ORC_EV_COAL) {
          double x = re->ev.x0;
          if (xs <= x && x < xe) {            /* count.cpp:464,512: counted once, when update_to passes it */
            out[0 * E + e] += w * re->ev.cnt;
            out[1 * E + e] += w * re->ev.opp;
            out[2 * E + e] += w * w * re->ev.opp;
          }
        } else if (re->ev.type == ORC_EV_RECOPP) {
          double a = re->ev.x0 > xs ? re->ev.x0 : xs, b = re->ev.x1 < xe ? re->ev.x1 : xe;
          if (b > a) {                        /* coalevent.hpp:224-226 clipped rectangle */
            double opp = re->ev.opp * (b - a);
            out[4 * E + e] += w * opp;
            out[5 * E + e] += w * w * opp;
          }
        } else {
          double x = re->ev.x0;
          if (xs <= x && x < xe) out[3 * E + e] += w * re->ev.cnt;   /* coalevent.hpp:231-235 */
        }
      }
    }
    for (int e = 0; e < E; e++) counted_to[e] = update_to[e];
  }
  free(W);
}
