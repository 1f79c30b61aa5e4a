/* TEST INFRASTRUCTURE ONLY -- see smc_struct_oracle.h. */
#include "smc_struct_oracle.h"
#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---- marks ------------------------------------------------------------------------------------------------- */
int orc_spop_at(const orc_sparticle* p, int v, double t) {
  int pop = p->pop[v];
  double best = -1.0;
  for (int k = 0; k < p->nmarks; k++)
    if (p->mark_branch[k] == v && p->mark_time[k] <= t && p->mark_time[k] > best) { best = p->mark_time[k]; pop = p->mark_pop[k]; }
  return pop;
}
static void add_mark(orc_sparticle* p, int branch, double time, int pop) {
  if (p->nmarks >= ORC_MAXMARKS) { p->overflow = 1; return; }
  p->mark_branch[p->nmarks] = branch; p->mark_time[p->nmarks] = time; p->mark_pop[p->nmarks] = pop;
  p->nmarks++;
}
/* drop the marks of `branch` with lo <= time < hi */
static void drop_marks(orc_sparticle* p, int branch, double lo, double hi) {
  int w = 0;
  for (int k = 0; k < p->nmarks; k++) {
    int gone = p->mark_branch[k] == branch && p->mark_time[k] >= lo && p->mark_time[k] < hi;
    if (!gone) { p->mark_branch[w] = p->mark_branch[k]; p->mark_time[w] = p->mark_time[k]; p->mark_pop[w] = p->mark_pop[k]; w++; }
  }
  p->nmarks = w;
}
/* hand the marks of `from` with lo <= time < hi to branch `to` */
static void move_marks(orc_sparticle* p, int from, int to, double lo, double hi) {
  for (int k = 0; k < p->nmarks; k++)
    if (p->mark_branch[k] == from && p->mark_time[k] >= lo && p->mark_time[k] < hi) p->mark_branch[k] = to;
}

static int sepoch_of(const orc_smodel* m, double t) {
  int e = 0;
  while (e + 1 < m->base.E && m->base.t_start[e + 1] <= t) e++;
  return e;
}
static double mig_total(const orc_smodel* m, int e, int i) {
  double s = 0.0;
  for (int j = 0; j < m->P; j++) if (j != i) s += m->mig[e][i][j];
  return s;
}

/* a lineage's path of population changes while it floats */
typedef struct { int n; double t[256]; int pop[256]; } spath;
static void path_add(spath* a, double t, int pop) { if (a->n < 256) { a->t[a->n] = t; a->pop[a->n] = pop; a->n++; } }

static void srecord(orc_srecords* recs, int who, int epoch, int pop, int weight, double dt, int event, int to_pop) {
  if (!recs || recs->count >= recs->cap) return;
  orc_srecord* q = &recs->rec[recs->count++];
  q->who = who; q->epoch = epoch; q->pop = pop; q->weight = weight; q->dt = dt; q->event = event; q->to_pop = to_pop;
}

/* Fixed-time events at the start of epoch e for a floating lineage in population *pop (forest.cc:1139-1175): one
 * uniform per round, chained until no move happens. */
static void fixed_time_events(const orc_smodel* m, int e, double t, int* pop, spath* path, orc_rng* r) {
  int any = 0;
  for (int i = 0; i < m->P && !any; i++) for (int j = 0; j < m->P; j++) if (m->single_mig[e][i][j] > 0.0) any = 1;
  if (!any) return;                                       /* Model::hasFixedTimeEvent */
  for (int chain = 0; chain < 10000; chain++) {
    double sample = orc_rng_uniform(r);
    int moved = 0;
    for (int j = 0; j < m->P; j++) {
      sample -= m->single_mig[e][*pop][j];
      if (sample < 0) { *pop = j; path_add(path, t, j); moved = 1; break; }
    }
    if (!moved) break;
  }
}

/* The sweep of a floating lineage that starts at (t0, pop0) over the tree made of the nodes with present[v] != 0
 * (forest.cc:496-598 with calcRates :601-664, sampleEvent :677-696, sampleEventType :705-756).  Returns the
 * coalescence time; *target is the branch it coalesces into, or -1 for the pairwise coalescence with the old root's
 * lineage above the root.  fa / ra collect the population changes of the floating and of the root lineage. */
static double sfloat(const orc_smodel* m, const orc_sparticle* p, const char* present, int root, orc_rng* r, double t0,
                     int pop0, spath* fa, spath* ra, int* target, int* coal_pop, orc_srecords* recs, double* vb_logw) {
  const int n = m->base.n, nn = 2 * n - 1;
  const int* parent = p->tree.parent; const double* height = p->tree.height;
  const double hroot = height[root];
  double tt = t0;
  int pf = pop0, pr = p->pop[root], e = sepoch_of(m, tt);
  fa->n = 0; ra->n = 0;
  for (;;) {
    const int above = !(tt < hroot);
    /* contemporaries of the floating lineage's population, and the end of the interval */
    int k = 0;
    double tend = e + 1 < m->base.E ? m->base.t_start[e + 1] : DBL_MAX;
    const double tep = tend;
    if (!above) {
      for (int v = 0; v < nn; v++) {
        if (!present[v]) continue;
        if (v >= n && height[v] > tt && height[v] < tend) tend = height[v];
        if (parent[v] >= 0 && height[v] <= tt && tt < height[parent[v]] && orc_spop_at(p, v, tt) == pf) k++;
      }
      for (int q = 0; q < p->nmarks; q++) if (p->mark_time[q] > tt && p->mark_time[q] < tend) tend = p->mark_time[q];
    }
    const double coal_rate = above ? 0.0 : k * m->inv2N[e][pf];
    const double mig_f = mig_total(m, e, pf);
    const double mig_r = above ? mig_total(m, e, pr) : 0.0;
    const double pw = (above && pf == pr) ? m->inv2N[e][pf] : 0.0;
    const double rate = coal_rate + mig_f + mig_r + pw;
    double w = rate > 0.0 ? -log(orc_rng_uniform(r)) / rate : DBL_MAX;
    const int hit = tt + w < tend;
    if (!hit && tend == DBL_MAX) { *target = -3; *coal_pop = pf; return tt; }   /* "Lines did not coalescence" (forest.cc:856-871) */
    const double tev = hit ? tt + w : tend;
    int ev_f = 0, to_f = -1, ev_r = 0, to_r = -1;
    if (hit) {
      double u = orc_rng_uniform(r) * rate;
      if (u < coal_rate) ev_f = 1;
      else if ((u -= coal_rate) < mig_f) {
        ev_f = 2;
        for (int j = 0; j < m->P; j++) { if (j == pf) continue; to_f = j; u -= m->mig[e][pf][j]; if (u < 0) break; }
      } else if ((u -= mig_f) < mig_r) {
        ev_r = 2;
        for (int j = 0; j < m->P; j++) { if (j == pr) continue; to_r = j; u -= m->mig[e][pr][j]; if (u < 0) break; }
      } else ev_f = 1;                                   /* pairwise coalescence, booked on the floating lineage */
    }
    srecord(recs, 0, e, pf, k + (pw > 0.0 ? 1 : 0), tev - tt, ev_f, to_f);
    if (above) srecord(recs, 1, e, pr, 0, tev - tt, ev_r, to_r);
    if (vb_logw) {                                        /* particle.cpp:262-272, before the record masks apply */
      if (ev_f == 1) *vb_logw += m->vb_coal[e][pf];
      if (ev_f == 2) *vb_logw += m->vb_mig[e][pf][to_f];
      if (ev_r == 2) *vb_logw += m->vb_mig[e][pr][to_r];
    }
    const double ts = tt;                                 /* start of the interval: where the contemporaries were counted */
    tt = tev;
    if (hit) {
      if (ev_f == 1) {
        *coal_pop = pf;
        if (above) { *target = -1; return tt; }
        int idx = (int)(orc_rng_uniform(r) * k);
        if (idx >= k) idx = k - 1;
        /* the set of contemporaries is constant inside an interval: pick at its start */
        for (int v = 0; v < nn; v++) {
          if (!present[v] || parent[v] < 0) continue;
          if (height[v] <= ts && ts < height[parent[v]] && orc_spop_at(p, v, ts) == pf) { if (idx == 0) { *target = v; return tt; } idx--; }
        }
        *target = -2;                                     /* cannot happen */
        return tt;
      }
      if (ev_f == 2) { pf = to_f; path_add(fa, tt, pf); }
      if (ev_r == 2) { pr = to_r; path_add(ra, tt, pr); }
      continue;
    }
    if (tend == tep) {
      e++;
      fixed_time_events(m, e, tt, &pf, fa, r);
      if (!(tt < hroot)) fixed_time_events(m, e, tt, &pr, ra, r);
    }
  }
}

/* move internal slot a to its place in the height order, renaming parents, marks and populations with it */
static void srotate(orc_sparticle* p, int n, int nn_used, int a) {
  int* parent = p->tree.parent; double* height = p->tree.height;
  double ha = height[a];
  int d = a;
  while (d + 1 < nn_used && height[d + 1] < ha) d++;
  if (d == a) while (d - 1 >= n && height[d - 1] > ha) d--;
  if (d == a) return;
  int pa = parent[a], popa = p->pop[a];
  if (d > a) for (int j = a; j < d; j++) { height[j] = height[j + 1]; parent[j] = parent[j + 1]; p->pop[j] = p->pop[j + 1]; }
  else for (int j = a; j > d; j--) { height[j] = height[j - 1]; parent[j] = parent[j - 1]; p->pop[j] = p->pop[j - 1]; }
  height[d] = ha; parent[d] = pa; p->pop[d] = popa;
  const int lo = d > a ? a : d, hi = d > a ? d : a, shift = d > a ? -1 : 1;
  for (int v = 0; v < nn_used; v++) {
    int q = parent[v];
    if (q >= lo && q <= hi) parent[v] = q == a ? d : q + shift;
  }
  for (int k = 0; k < p->nmarks; k++) {
    int q = p->mark_branch[k];
    if (q >= lo && q <= hi) p->mark_branch[k] = q == a ? d : q + shift;
  }
}

static double stree_length(int nn, const int* parent, const double* height) {
  double len = 0.0;
  for (int v = 0; v < nn; v++) if (parent[v] >= 0) len += height[parent[v]] - height[v];
  return len;
}
static void sdraw_next_base(const orc_smodel* m, orc_sparticle* p, orc_rng* r, double from) {
  double rate = m->base.rho * stree_length(2 * m->base.n - 1, p->tree.parent, p->tree.height);
  double next = from + -log(orc_rng_uniform(r)) / rate;
  if (next == from) next = nextafter(from, from * 2);
  if (!(next < m->base.seqlen)) next = m->base.seqlen;
  p->tree.next_base = next;
}

/* Forest::buildInitialTree (forest.cc:299-347): leaf 0 is the tree; every further leaf floats up from its own
 * population and coalesces into the tree built so far (or pairwise with its root's lineage). */
void orc_sinit_particle(const orc_smodel* m, orc_sparticle* p, orc_rng* r, double start_pos) {
  const int n = m->base.n, nn = 2 * n - 1;
  char present[ORC_MAXNODES];
  memset(present, 0, sizeof present);
  memset(p, 0, sizeof *p);
  for (int v = 0; v < nn; v++) { p->tree.parent[v] = -1; p->tree.height[v] = 0.0; }
  for (int i = 0; i < n; i++) p->pop[i] = m->leaf_pop[i];
  present[0] = 1;
  int root = 0, used = n;                       /* internal slots n .. used-1 are in use */
  for (int i = 1; i < n; i++) {
    spath fa, ra;
    int target = -1, cpop = 0;
    double tc = sfloat(m, p, present, root, r, 0.0, m->leaf_pop[i], &fa, &ra, &target, &cpop, 0, 0);
    const int nw = used++;
    present[i] = 1; present[nw] = 1;
    p->tree.height[nw] = tc; p->pop[nw] = cpop;
    for (int q = 0; q < fa.n; q++) add_mark(p, i, fa.t[q], fa.pop[q]);
    p->tree.parent[i] = nw;
    if (target < 0) {                           /* new root above the old one */
      for (int q = 0; q < ra.n; q++) add_mark(p, root, ra.t[q], ra.pop[q]);
      p->tree.parent[root] = nw;
      p->tree.parent[nw] = -1;
      root = nw;
    } else {
      p->tree.parent[nw] = p->tree.parent[target];
      p->tree.parent[target] = nw;
      move_marks(p, target, nw, tc, DBL_MAX);
    }
    /* sort the new slot into place among the used internal slots; the root is the one present node without parent */
    srotate(p, n, used, nw);
    for (int v = 0; v < nn; v++) if (present[v] && p->tree.parent[v] < 0) root = v;
  }
  p->tree.logw = 0.0;
  p->tree.updated_to = start_pos;
  p->tree.cur_base = start_pos;
  sdraw_next_base(m, p, r, start_pos);
}

/* Forest::sampleNextGenealogy under window 0 with populations (forest.cc:423-598): the P = 1 case is orc_smc_move. */
void orc_ssmc_move(const orc_smodel* m, orc_sparticle* p, orc_rng* r, orc_srecords* recs, double* rec_height_out) {
  const int n = m->base.n, nn = 2 * n - 1;
  int* parent = p->tree.parent; double* height = p->tree.height;
  char present[ORC_MAXNODES];
  memset(present, 1, sizeof present);
  /* 1. recombination point, uniform on the tree */
  double tl = stree_length(nn, parent, height);
  double goal = orc_rng_uniform(r) * tl;
  int b = -1; double hh = 0.0, cum = 0.0;
  for (int v = 0; v < nn; v++) {
    if (parent[v] < 0) continue;
    double bl = height[parent[v]] - height[v];
    b = v; hh = height[v] + (goal - cum);
    if (goal < cum + bl) break;
    cum += bl;
  }
  if (hh > height[parent[b]]) hh = height[parent[b]];
  if (rec_height_out) *rec_height_out = hh;
  const int Pn = parent[b];
  int root = nn - 1;
  int S = -1;
  for (int v = 0; v < nn; v++) if (parent[v] == Pn && v != b) S = v;
  /* 2. float */
  spath fa, ra;
  int target = -1, cpop = 0;
  const double tc = sfloat(m, p, present, root, r, hh, orc_spop_at(p, b, hh), &fa, &ra, &target, &cpop, recs,
                           recs ? &p->tree.logw : 0);   /* the -vb factors belong to the filter, not to the prior walks */
  /* 3. surgery.  The stub of b above the cut disappears unless the lineage coalesces back into it. */
  if (target == b) {
    drop_marks(p, b, hh, tc);
    for (int q = 0; q < fa.n; q++) add_mark(p, b, fa.t[q], fa.pop[q]);
    return;
  }
  drop_marks(p, b, hh, DBL_MAX);
  for (int q = 0; q < fa.n; q++) add_mark(p, b, fa.t[q], fa.pop[q]);
  if (target < 0) {                                  /* pairwise above the old root */
    if (Pn != root) {
      move_marks(p, Pn, S, 0.0, DBL_MAX);            /* S's lineage now runs through where Pn was */
      parent[S] = parent[Pn];
      for (int q = 0; q < ra.n; q++) add_mark(p, root, ra.t[q], ra.pop[q]);
      parent[root] = Pn;
      parent[Pn] = -1;
    } else {
      for (int q = 0; q < ra.n; q++) add_mark(p, S, ra.t[q], ra.pop[q]);   /* the root lineage is S's */
    }
    height[Pn] = tc; p->pop[Pn] = cpop;
  } else if (target == Pn) {                         /* into the branch above Pn: Pn slides up, S inherits the part below */
    move_marks(p, Pn, S, 0.0, tc);
    height[Pn] = tc; p->pop[Pn] = cpop;
  } else if (target == S) {                          /* into S's branch: Pn slides down, taking S's marks above tc */
    move_marks(p, S, Pn, tc, DBL_MAX);
    height[Pn] = tc; p->pop[Pn] = cpop;
  } else {
    if (parent[Pn] >= 0) move_marks(p, Pn, S, 0.0, DBL_MAX);
    else drop_marks(p, S, 0.0, DBL_MAX);             /* Pn was the root: S becomes the root and has no branch above */
    parent[S] = parent[Pn];
    parent[Pn] = parent[target];
    parent[target] = Pn;
    move_marks(p, target, Pn, tc, DBL_MAX);
    height[Pn] = tc; p->pop[Pn] = cpop;
  }
  srotate(p, n, nn, Pn);
  drop_marks(p, nn - 1, 0.0, DBL_MAX);               /* the root has no branch above it (e.g. Pn was the root and slid down) */
}

int orc_scheck(const orc_smodel* m, const orc_sparticle* p) {
  const int n = m->base.n, nn = 2 * n - 1;
  int kids[ORC_MAXNODES];
  memset(kids, 0, sizeof kids);
  int roots = 0;
  for (int v = 0; v < nn; v++) {
    int q = p->tree.parent[v];
    if (q < 0) { roots++; continue; }
    if (q < n || q >= nn) return 1;
    if (p->tree.height[q] < p->tree.height[v]) return 2;
    kids[q]++;
    /* population at the top of the branch must be the population the parent's coalescence happened in */
    if (orc_spop_at(p, v, p->tree.height[q]) != p->pop[q]) return 3;
  }
  if (roots != 1 || p->tree.parent[nn - 1] >= 0) return 4;
  for (int v = n; v < nn; v++) { if (kids[v] != 2) return 5; if (v > n && p->tree.height[v] < p->tree.height[v - 1]) return 6; }
  for (int i = 0; i < n; i++) if (p->pop[i] != m->leaf_pop[i] || p->tree.height[i] != 0.0) return 7;
  for (int k = 0; k < p->nmarks; k++) {
    int v = p->mark_branch[k];
    if (v < 0 || v >= nn || p->tree.parent[v] < 0) return 8;
    if (!(p->mark_time[k] >= p->tree.height[v] && p->mark_time[k] <= p->tree.height[p->tree.parent[v]])) return 9;
  }
  return p->overflow ? 10 : 0;
}

/* ---- walks used by the distribution-level tests (the reference side is ref_prior_stats / ref_arg_walk) ---------- */
int orc_sprior_stats(const orc_smodel* m, uint64_t seed, int count, double* tmrca, double* tree_len, double* marks) {
  const int nn = 2 * m->base.n - 1;
  for (int t = 0; t < count; t++) {
    orc_sparticle p;
    orc_rng r;
    orc_rng_init(&r, seed, (uint32_t)t, 0, 0, 1);
    orc_sinit_particle(m, &p, &r, 0.0);
    int bad = orc_scheck(m, &p);
    if (bad) return 100 + bad;
    tmrca[t] = p.tree.height[nn - 1];
    tree_len[t] = stree_length(nn, p.tree.parent, p.tree.height);
    if (marks) marks[t] = p.nmarks;
  }
  return 0;
}

int orc_sarg_walk(const orc_smodel* m, uint64_t seed, int count, double* pos, double* rec_h, double* tmrca, double* tlen,
                  double* marks) {
  const int nn = 2 * m->base.n - 1;
  orc_sparticle p;
  orc_rng r;
  orc_rng_init(&r, seed, 0, 0, 0, 2);
  orc_sinit_particle(m, &p, &r, 0.0);
  for (int i = 0; i < count; i++) {
    if (p.tree.next_base >= m->base.seqlen) return i;
    p.tree.cur_base = p.tree.next_base;
    orc_ssmc_move(m, &p, &r, 0, &rec_h[i]);
    int bad = orc_scheck(m, &p);
    if (bad) return -(100 + bad);
    pos[i] = p.tree.cur_base;
    sdraw_next_base(m, &p, &r, p.tree.cur_base);
    tmrca[i] = p.tree.height[nn - 1];
    tlen[i] = stree_length(nn, p.tree.parent, p.tree.height);
    if (marks) marks[i] = p.nmarks;
  }
  return count;
}

/* ---- the structured particle filter (test-side mirror of the engine with populations) --------------------------- */
/* channel layout of the per-epoch tallies with P populations, Q = 2 + 3P + P*P:
 *   0 recombination count, 1 recombination opportunity, 2+3p coalescence count in p, 3+3p coalescence opportunity in p,
 *   4+3p migration opportunity in p, 2+3P + p*P + q migration count p -> q (backwards in time) */
int orc_schannels(int P) { return 2 + 3 * P + P * P; }

typedef struct { orc_chan_event* ev; long count, cap; } chan_buf;
static void chan_emit(chan_buf* b, int row, int slot, int epoch, int chan, int rect, double x0, double x1, double val) {
  if (b->count >= b->cap) {
    b->cap = b->cap * 2 + 1024;
    b->ev = (orc_chan_event*)realloc(b->ev, sizeof(orc_chan_event) * (size_t)b->cap);
  }
  orc_chan_event* e = &b->ev[b->count++];
  e->row = row; e->slot = slot; e->epoch = epoch; e->chan = chan; e->rect = rect; e->x0 = x0; e->x1 = x1; e->val = val;
}

/* ForestState::extend_ARG for one row with populations: orc_extend_row with orc_ssmc_move as the transition and the
 * records of particle.cpp:251-300 (coalescence AND migration opportunity of the floating lineages) as channel events.
 * Returns the number of recombinations. */
static int sextend_row(const orc_smodel* m, orc_sparticle* p, orc_rng* r, double x1, const int* data, int is_site,
                       int ancestral_aware, int max_rec_epoch, chan_buf* out, int row, int slot) {
  const orc_model* b = &m->base;
  const int n = b->n, P = m->P;
  orc_particle* t = &p->tree;
  int nmiss = 0, nrec = 0;
  for (int i = 0; i < n; i++) nmiss += data[i] == -1;
  int leaf_status = nmiss == 0 ? 1 : (nmiss == n ? -1 : 0);
  double L_obs = leaf_status == -1 ? 0.0
               : leaf_status == 1 ? orc_tree_length(n, t->parent, t->height)
                                  : orc_branch_length_with_data(n, t->parent, t->height, data);
  double len_e[ORC_MAXE];
  orc_srecord rbuf[512];
  orc_epoch_tree_length(b, t->parent, t->height, len_e);
  while (t->updated_to < x1) {
    double new_to = x1 < t->next_base ? x1 : t->next_base;
    t->logw += log(orc_fastexp(-b->mu * L_obs * (new_to - t->updated_to)));
    for (int e = 0; e < b->E; e++)
      if (len_e[e] > 0.0 && (b->rec_mask[e] & 1) && e <= max_rec_epoch)
        chan_emit(out, row, slot, e, 1, 1, t->updated_to, new_to, len_e[e]);
    t->updated_to = new_to;
    if (t->updated_to < x1) {
      t->cur_base = t->updated_to;
      orc_srecords recs = { rbuf, 0, 512 };
      double hh = 0.0;
      orc_ssmc_move(m, p, r, &recs, &hh);
      nrec++;
      int e = sepoch_of(m, hh);
      if ((b->rec_mask[e] & 1) && e <= max_rec_epoch) chan_emit(out, row, slot, e, 0, 0, t->cur_base, t->cur_base, 1.0);
      for (int k = 0; k < recs.count; k++) {
        const orc_srecord* q = &recs.rec[k];
        if (!((b->rec_mask[q->epoch] & 2) && q->epoch <= max_rec_epoch)) continue;
        const double x = t->cur_base;
        if (q->event == 1) chan_emit(out, row, slot, q->epoch, 2 + 3 * q->pop, 0, x, x, 1.0);
        if (q->weight > 0) chan_emit(out, row, slot, q->epoch, 3 + 3 * q->pop, 0, x, x, q->weight * q->dt);
        chan_emit(out, row, slot, q->epoch, 4 + 3 * q->pop, 0, x, x, q->dt);
        if (q->event == 2) chan_emit(out, row, slot, q->epoch, 2 + 3 * P + q->pop * P + q->to_pop, 0, x, x, 1.0);
      }
      if (leaf_status == 0) L_obs = orc_branch_length_with_data(n, t->parent, t->height, data);
      if (leaf_status == 1) L_obs = orc_tree_length(n, t->parent, t->height);
      orc_epoch_tree_length(b, t->parent, t->height, len_e);
      sdraw_next_base(m, p, r, t->cur_base);
    }
    t->cur_base = t->updated_to;
  }
  if (is_site) t->logw += log(orc_site_likelihood(n, t->parent, t->height, data, b->mu, ancestral_aware, 0));
  return nrec;
}

/* orc_particle_filter (smc_filter_oracle.c) with populations: same streams, same resampling, channel tallies
 * out [2][Q][E].  Returns 0, 1 on zero total weight, 2 when a particle failed orc_scheck. */
int orc_sparticle_filter(const orc_smodel* m, int N, uint64_t seed, uint32_t chunk, uint32_t slot_offset,
                         double start_pos, double ess_fraction, int ancestral_aware, const double* lags, int R,
                         const double* row_end, const int* data, const int* is_site, const int* max_rec,
                         orc_sparticle* parts, double* lnZ_out, double* tallies, int* n_resample_out, int* last_anc,
                         long* n_recomb_out) {
  const int n = m->base.n, E = m->base.E, Q = orc_schannels(m->P);
  orc_rng r;
  orc_sparticle* next = (orc_sparticle*)malloc(sizeof(orc_sparticle) * (size_t)N);
  double* weights = (double*)malloc(sizeof(double) * (size_t)R * N);
  int* anc = (int*)malloc(sizeof(int) * (size_t)R * N);
  double* lw = (double*)malloc(sizeof(double) * (size_t)N);
  double* ps = (double*)malloc(sizeof(double) * (size_t)(N + 1));
  int* counts = (int*)malloc(sizeof(int) * (size_t)N);
  chan_buf evs = { 0, 0, 0 };
  int rc = 0, n_resample = 0;
  long n_recomb = 0;
  double lnZ = 0.0;
  for (int i = 0; i < N; i++) {
    orc_rng_init(&r, seed, slot_offset + (uint32_t)i, 0, chunk, 1);
    orc_sinit_particle(m, &parts[i], &r, start_pos);
    parts[i].tree.logw = -log((double)N);
    if (last_anc) last_anc[i] = -1;
  }
  for (int row = 0; row < R; row++) {
    for (int i = 0; i < N; i++) {
      orc_rng_init(&r, seed, slot_offset + (uint32_t)i, (uint32_t)row, chunk, 2);
      n_recomb += sextend_row(m, &parts[i], &r, row_end[row], data + (size_t)row * n, is_site[row], ancestral_aware,
                              max_rec ? max_rec[row] : E - 1, &evs, row, i);
      if (orc_scheck(m, &parts[i])) { rc = 2; goto done; }
    }
    for (int i = 0; i < N; i++) lw[i] = parts[i].tree.logw;
    double lse = orc_normalize_log(lw, N);
    if (!isfinite(lse)) { rc = 1; goto done; }
    lnZ += lse;
    for (int i = 0; i < N; i++) { parts[i].tree.logw = lw[i]; weights[(size_t)row * N + i] = exp(lw[i]); }
    double ess = orc_ess_partial_sums(weights + (size_t)row * N, N, ps);
    anc[(size_t)row * N] = -1;
    if (ess < N * ess_fraction - 1e-6) {
      n_resample++;
      if (row + 1 == R) break;
      orc_rng_init(&r, seed, 0, (uint32_t)row, chunk, 3);
      double u = orc_rng_uniform(&r);
      orc_systematic_resampling(ps, N, N, u, counts);
      int j = 0;
      for (int i = 0; i < N; i++)
        for (int k = 0; k < counts[i]; k++) {
          next[j] = parts[i];
          next[j].tree.logw = log(ps[N] / (double)N);
          anc[(size_t)row * N + j] = i;
          if (counts[i] > 1) {
            orc_rng_init(&r, seed, slot_offset + (uint32_t)j, (uint32_t)row, chunk, 4);
            sdraw_next_base(m, &next[j], &r, next[j].tree.cur_base);
          }
          j++;
        }
      memcpy(parts, next, sizeof(orc_sparticle) * (size_t)N);
      if (last_anc) memcpy(last_anc, anc + (size_t)row * N, sizeof(int) * (size_t)N);
    }
  }
  if (tallies)
    orc_channel_tallies(E, Q, lags, R, N, row_end, weights, anc, weights + (size_t)(R - 1) * N, m->base.seqlen, evs.ev,
                        evs.count, tallies);
done:
  *lnZ_out = lnZ;
  if (n_resample_out) *n_resample_out = n_resample;
  if (n_recomb_out) *n_recomb_out = n_recomb;
  free(next); free(weights); free(anc); free(lw); free(ps); free(counts); free(evs.ev);
  return rc;
}
