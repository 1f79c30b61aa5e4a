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

/* The same rule for any number of tallied quantities ("channels", Q per epoch): a POINT record adds val to its channel
 * when the update window [counted_to, update_to) passes its position x0; a RECT record adds val * |[x0,x1) n window|.
 * out: [2][Q][E] = sum of w * amount, then sum of w^2 * amount (the latter feeds the ESS column for the opportunity
 * channels).  orc_lagged_tallies below is this function with the single-population channel map; the structured
 * filter (smc_struct_oracle.c) uses it with per-population coalescence / migration channels. */
/* Optional side output [Q][E]: the second moments as the CUDA engine defines them -- the weight is pooled
 * over the particles that share the record's LAG BUCKET, i.e. that have the same ancestor in the row in which the
 * update window the record falls into was passed by the sweep (the engine freezes a bucket's cells there), not the same
 * ancestor in the row that wrote the record as in the reference.  A lower bound of the reference's value (block 1 of
 * `out`), equal to it when no resampling separates writing and closing.  The window counted at the end of the data
 * never closes: per-particle squares.  NULL switches it off. */
static double* g_pooled_at_close = 0;
void orc_set_pooled_at_close_buffer(double* out) { g_pooled_at_close = out; }

void orc_channel_tallies(int E, int Q, const double* lags, int R, int N, const double* row_end, const double* weights,
                         const int* anc, const double* w_final, double seq_end, const orc_chan_event* evs, long nev,
                         double* out) {
  double counted_to[ORC_MAXE], update_to[ORC_MAXE];
  for (int e = 0; e < E; e++) counted_to[e] = 0.0;
  memset(out, 0, sizeof(double) * 2 * (size_t)Q * E);
  double* pooled = g_pooled_at_close;
  double* V = pooled ? (double*)malloc(sizeof(double) * (size_t)(R + 1) * N) : 0;   /* V[r][s], see below */
  if (pooled) memset(pooled, 0, sizeof(double) * (size_t)Q * E);
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
      /* engine-defined second moments: per epoch the row in which this step's window closed, and for every (row, slot)
       * at or before it the sum of squared descendant weights over its descendants in that row */
      for (int ep = first; pooled && ep < E; ep++) {
        int rc = top;                                   /* end of data: the window never closes -> per particle */
        if (!end_data) { rc = 0; while (rc < step && row_end[rc] < update_to[ep]) rc++; }
        for (int i = 0; i < N; i++) { double w = W[(size_t)rc * N + i]; V[(size_t)rc * N + i] = w * w; }
        for (int r = rc - 1; r >= 0; r--) {
          double* dst = V + (size_t)r * N;
          const double* src = V + (size_t)(r + 1) * N;
          if (anc[(size_t)r * N] < 0) { memcpy(dst, src, sizeof(double) * N); continue; }
          for (int i = 0; i < N; i++) dst[i] = 0.0;
          for (int i = 0; i < N; i++) dst[anc[(size_t)r * N + i]] += src[i];
        }
        for (long k = 0; k < nev; k++) {
          const orc_chan_event* re = &evs[k];
          if (re->epoch != ep || re->row > top || (!end_data && re->row > step)) continue;
          double xs = counted_to[ep], xe = update_to[ep], amount;
          if (re->rect) {
            double a = re->x0 > xs ? re->x0 : xs, b = re->x1 < xe ? re->x1 : xe;
            if (!(b > a)) continue;
            amount = re->val * (b - a);
          } else {
            if (!(xs <= re->x0 && re->x0 < xe)) continue;
            amount = re->val;
          }
          int rw = re->row < rc ? re->row : rc;          /* a record written in the closing row itself: its own slot */
          pooled[(size_t)re->chan * E + ep] += V[(size_t)rw * N + re->slot] * amount;
        }
      }
      for (long k = 0; k < nev; k++) {
        const orc_chan_event* re = &evs[k];
        int e = re->epoch;
        if (e < first) continue;
        if (re->row > top) continue;
        if (!end_data && re->row > step) continue;
        double w = W[(size_t)re->row * N + re->slot];
        double xs = counted_to[e], xe = update_to[e];
        double amount;
        if (re->rect) {
          double a = re->x0 > xs ? re->x0 : xs, b = re->x1 < xe ? re->x1 : xe;
          if (!(b > a)) continue;             /* coalevent.hpp:224-226 clipped rectangle */
          amount = re->val * (b - a);
        } else {
          if (!(xs <= re->x0 && re->x0 < xe)) continue;   /* count.cpp:464,512: counted once, when update_to passes it */
          amount = re->val;
        }
        out[((size_t)0 * Q + re->chan) * E + e] += w * amount;
        out[((size_t)1 * Q + re->chan) * E + e] += w * w * amount;
      }
    }
    for (int e = 0; e < E; e++) counted_to[e] = update_to[e];
  }
  free(W);
  if (V) free(V);
}

/* out: [6][E] = coal_count, coal_opp, coal_weight(w^2), rec_count, rec_opp, rec_weight(w^2) -- WITHOUT the initial
 * pseudo-counts of count.cpp:161-193.
 * rows 0..R-1: row_end[r], weights[r*N+i] = normalised posterior weight of slot i when row r's update runs
 * (after normalize_probability, before resampling); anc[r*N+i] = slot (in row r's indexing) that slot i of row r+1 was
 * copied from, or anc[r*N] = -1 when no resampling followed row r.  w_final = weights for the end-of-data flush
 * (smcsmc.cpp:371-373), in the indexing after the last row's resampling.
 * Channels: 0 coalescence count, 1 coalescence opportunity, 2 recombination count, 3 recombination opportunity. */
void orc_lagged_tallies(int E, const double* lags, int R, int N, const double* row_end, const double* weights,
                        const int* anc, const double* w_final, double seq_end, const orc_rec_event* evs, long nev,
                        double* out) {
  orc_chan_event* g = (orc_chan_event*)malloc(sizeof(orc_chan_event) * (size_t)(2 * nev + 1));
  long ng = 0;
  for (long k = 0; k < nev; k++) {
    const orc_rec_event* re = &evs[k];
    orc_chan_event c = { re->row, re->slot, re->ev.epoch, 0, 0, re->ev.x0, re->ev.x1, 0.0 };
    if (re->ev.type == ORC_EV_COAL) {
      c.chan = 0; c.val = re->ev.cnt; g[ng++] = c;
      c.chan = 1; c.val = re->ev.opp; g[ng++] = c;
    } else if (re->ev.type == ORC_EV_RECOPP) {
      c.chan = 3; c.rect = 1; c.val = re->ev.opp; g[ng++] = c;
    } else {
      c.chan = 2; c.val = re->ev.cnt; g[ng++] = c;
    }
  }
  double* o = (double*)malloc(sizeof(double) * 2 * 4 * (size_t)E);
  double* user = g_pooled_at_close;                       /* [2][E]: coalescence and recombination opportunity */
  double* p4 = user ? (double*)malloc(sizeof(double) * 4 * (size_t)E) : 0;
  g_pooled_at_close = p4;
  orc_channel_tallies(E, 4, lags, R, N, row_end, weights, anc, w_final, seq_end, g, ng, o);
  g_pooled_at_close = user;
  if (user) {
    for (int e = 0; e < E; e++) { user[e] = p4[1 * E + e]; user[E + e] = p4[3 * E + e]; }
    free(p4);
  }
  for (int e = 0; e < E; e++) {
    out[0 * E + e] = o[(0 * 4 + 0) * E + e];
    out[1 * E + e] = o[(0 * 4 + 1) * E + e];
    out[2 * E + e] = o[(1 * 4 + 1) * E + e];
    out[3 * E + e] = o[(0 * 4 + 2) * E + e];
    out[4 * E + e] = o[(0 * 4 + 3) * E + e];
    out[5 * E + e] = o[(1 * 4 + 3) * E + e];
  }
  free(g); free(o);
}
