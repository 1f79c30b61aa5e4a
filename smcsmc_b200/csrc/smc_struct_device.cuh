// Structured populations (several demes, migration, joins) for the SMCSMC particle filter on sm_100a.
//
// The particle's local tree stays the binary tree of smc_device.cuh (leaves 0..n-1 at height 0, internal slots sorted by
// height).  Added per particle: pop[v] = population at the bottom of the branch above v, and a small pool of migration
// marks (branch, time, population above the mark).  scrm stores a mark as a unary node inside the tree; nothing else on
// the hot path looks at those nodes, so a side pool keeps the likelihood, branch-length and recombination-opportunity
// code of the single-population case untouched.
//
// Reference behaviour restated here (paths under /root/reference/src/scrm/src unless noted):
//   Forest::buildInitialTree                 forest.cc:299-347     sequential insertion of the leaves
//   Forest::sampleNextGenealogy              forest.cc:423-467
//   Forest::sampleCoalescences               forest.cc:496-598     two active lineages: the cut one and the old root's
//   calcRates / sampleEvent / sampleEventType forest.cc:601-756    coalescence + migration + pairwise rates
//   implementMigration / FixedTimeEvent      forest.cc:1100-1175   -em/-eM rates, -ej/-es moves at epoch starts
//   ContemporariesContainer                  contemporaries_container.h:249-379   per-population lineage sets
//   ForestState::record_all_event            ../../particle.cpp:193-302    opportunity + event per interval and lineage
// The test-side CPU restatement (oracle/smc_struct_oracle.c, pinned to the unmodified reference in law) consumes the
// same Philox draws in the same order: the two are compared draw for draw.
#pragma once
#include "smc_device.cuh"

namespace smc {

enum { SFLAG_MARKS = 1, SFLAG_NOCOAL = 4 };
// While a lineage floats, its population changes are written to the mark pool under a provisional branch id; the
// surgery that follows renames them to the branch the lineage ends up as (no per-move path buffer, no second cap).
enum { MARK_FLOATING = -2, MARK_ROOT_LINEAGE = -3 };

// Tables of the structured model; they live in HBM (read-only, cached) -- [E][P][P] does not fit the parameter bank.
struct SModelDev {
  int P;                        // populations (1 = structure off)
  int K;                        // mark capacity per particle
  int Q;                        // tally cells per epoch: 2 + 3P + P*P
  const double* inv2N;          // [E][P]    Model::inv_double_pop_size(pop, t)
  const double* mig;            // [E][P][P] Model::migration_rate(i, j), backwards in time i -> j
  const double* migtot;         // [E][P]    sum over j != i, added in index order
  const double* single_mig;     // [E][P][P] Model::single_mig_pop(i, j) at the START of the epoch
  const unsigned char* fixed;   // [E]       the epoch start has a fixed-time event
  const signed char* leaf_pop;  // [n]       Model::sample_population
  const double* vb_coal;        // [E][P]    -vb: log(exp_digamma(c)/c) per coalescence in (epoch, pop), or NULL
  const double* vb_mig;         // [E][P][P] -vb: the same per migration (epoch, from, to), or NULL
};

// cell indices inside an epoch's tally record
__host__ __device__ inline int cell_rec_cnt() { return 0; }
__host__ __device__ inline int cell_rec_opp() { return 1; }
__host__ __device__ inline int cell_coal_cnt(int p) { return 2 + 3 * p; }
__host__ __device__ inline int cell_coal_opp(int p) { return 3 + 3 * p; }
__host__ __device__ inline int cell_mig_opp(int p) { return 4 + 3 * p; }
__host__ __device__ inline int cell_mig_cnt(int P, int p, int q) { return 2 + 3 * P + p * P + q; }

template <int BS_>
struct STreeViewT : TreeViewT<BS_> {
  using TreeViewT<BS_>::t;
  static constexpr int BS = BS_;
  signed char* POP;     // [nn][BS]
  double* MT;           // [K][BS] mark times
  signed char* MB;      // [K][BS] mark branches
  signed char* MP;      // [K][BS] population above the mark
  signed char* NM;      // [BS]    marks in use
  unsigned char* FLAG;  // [BS]    SFLAG_* raised during this launch
  const SModelDev* sm;
  __device__ __forceinline__ signed char& pop(int v) const { return POP[v * BS + t]; }
  __device__ __forceinline__ double& mt(int k) const { return MT[k * BS + t]; }
  __device__ __forceinline__ signed char& mb(int k) const { return MB[k * BS + t]; }
  __device__ __forceinline__ signed char& mp(int k) const { return MP[k * BS + t]; }
  __device__ __forceinline__ signed char& nm() const { return NM[t]; }
  __device__ __forceinline__ void raise(int f) const { FLAG[t] |= (unsigned char)f; }
};

__host__ __device__ inline size_t stree_bytes_per_thread(int nn, int E, int K) {
  return tree_bytes_per_thread(nn, E) + (size_t)nn + (size_t)K * 10 + 2;
}
template <int BS>
__device__ __forceinline__ STreeViewT<BS> make_stree_view(unsigned char* smem, int n, int nn, int E, int t,
                                                          const SModelDev* sm) {
  STreeViewT<BS> T;
  T.t = t; T.n = n; T.sm = sm;
  const int K = sm->K;
  double* base = reinterpret_cast<double*>(smem);
  const size_t in = (size_t)(nn - n) * BS;
  T.H = base - (size_t)n * BS;
  T.L0 = base + in - (size_t)n * BS;
  T.L1 = base + 2 * in - (size_t)n * BS;
  T.ACC2 = base + in;
  T.ACC = base + in + tree_scratch_rows(nn, E) * BS;
  T.MT = T.ACC + (size_t)E * BS;
  T.PAR = reinterpret_cast<signed char*>(T.MT + (size_t)K * BS);
  T.POP = T.PAR + (size_t)nn * BS;
  T.MB = T.POP + (size_t)nn * BS;
  T.MP = T.MB + (size_t)K * BS;
  T.NM = T.MP + (size_t)K * BS;
  T.FLAG = reinterpret_cast<unsigned char*>(T.NM + BS);
  return T;
}

// ---- marks ---------------------------------------------------------------------------------------------------------
template <class TV>
__device__ inline int s_pop_at(const TV& T, int v, double t) {
  int pop = T.pop(v);
  double best = -1.0;
  const int nm = T.nm();
  for (int k = 0; k < nm; k++) {
    double mt = T.mt(k);
    if (T.mb(k) == v && mt <= t && mt > best) { best = mt; pop = T.mp(k); }
  }
  return pop;
}
template <class TV>
__device__ inline void s_add_mark(const TV& T, int branch, double time, int pop) {
  int nm = T.nm();
  if (nm >= T.sm->K) { T.raise(SFLAG_MARKS); return; }
  T.mb(nm) = (signed char)branch; T.mt(nm) = time; T.mp(nm) = (signed char)pop;
  T.nm() = (signed char)(nm + 1);
}
// drop the marks of `branch` with lo <= time < hi
template <class TV>
__device__ inline void s_drop_marks(const TV& T, int branch, double lo, double hi) {
  const int nm = T.nm();
  int w = 0;
  for (int k = 0; k < nm; k++) {
    double mt = T.mt(k);
    bool gone = T.mb(k) == branch && mt >= lo && mt < hi;
    if (!gone) {
      if (w != k) { T.mb(w) = T.mb(k); T.mt(w) = mt; T.mp(w) = T.mp(k); }
      w++;
    }
  }
  T.nm() = (signed char)w;
}
// hand the marks of `from` with lo <= time < hi to branch `to`
template <class TV>
__device__ inline void s_move_marks(const TV& T, int from, int to, double lo, double hi) {
  const int nm = T.nm();
  for (int k = 0; k < nm; k++) {
    double mt = T.mt(k);
    if (T.mb(k) == from && mt >= lo && mt < hi) T.mb(k) = (signed char)to;
  }
}

// forest.cc:1139-1175: fixed-time moves at the start of epoch e for a floating lineage in population pop; one uniform
// per round, chained until no move happens
template <class TV>
__device__ inline void s_fixed_time_events(const ModelDev& m, const TV& T, int e, double t, int& pop, int who, Rng& r) {
  const SModelDev& s = *T.sm;
  if (!s.fixed[e]) return;
  const int P = s.P;
  for (int chain = 0; chain < 10000; chain++) {
    double sample = rng_uniform(r);
    bool moved = false;
    for (int j = 0; j < P; j++) {
      sample -= s.single_mig[(e * P + pop) * P + j];
      if (sample < 0) { pop = j; s_add_mark(T, who, t, j); moved = true; break; }
    }
    if (!moved) break;
  }
}

// Null sink for the structured records (prior trees are not recorded)
struct SNullSink {
  __device__ void scoal(int, int, double, int, double) const {}
  __device__ void smig(int, int, double, int, double) const {}
  __device__ void srec(int, int, double, int, double, int, double) const {}
};

// The sweep of a floating lineage that starts at (t0, pop0) over the tree made of the nodes in `present`
// (forest.cc:496-598).  Returns the coalescence time; target = branch it coalesces into, -1 = pairwise coalescence with
// the old root's lineage above the root, -3 = the lineages cannot coalesce under this model.  The population changes of
// the floating and of the root lineage are left in the mark pool under MARK_FLOATING / MARK_ROOT_LINEAGE.
template <class TV, class Sink>
__device__ inline double s_float(const ModelDev& m, const TV& T, unsigned long long present, int root, Rng& r, double t0,
                                 int pop0, int& target, int& coal_pop, Sink& sink, double x, int max_rec_epoch,
                                 double* vb_logw) {
  const SModelDev& s = *T.sm;
  const int n = m.n, nn = m.nn, P = s.P;
  const double hroot = T.h(root);
  double tt = t0;
  int pf = pop0, pr = T.pop(root), e = epoch_of(m, tt);
  // branches that carry marks: only those need the mark pool to tell their population at a given time (the marks this
  // sweep adds belong to the two floating lineages, which are not branches of the tree yet)
  unsigned long long marked = 0ull;
  {
    const int nm0 = T.nm();
    for (int q = 0; q < nm0; q++) { const int mb = T.mb(q); if (mb >= 0) marked |= 1ull << mb; }
  }
  // The internal slots in use are height-sorted and contiguous from n (s_rotate keeps them so while the prior tree
  // grows): c = first slot above tt.  A branch v -> p is alive at tt exactly when v is a present leaf or a slot below c
  // and p >= c, and the next node above tt is slot c -- no height is compared inside the scans.
  double a_opp = 0.0, a_dt = 0.0, r_dt = 0.0;          // records not yet written: floating lineage (e, pf), root lineage (e, pr)
  int used = n;
  while (used < nn && ((present >> used) & 1ull)) used++;
  const unsigned long long leaves = present & ((1ull << n) - 1ull);
  int c = n;
  for (;;) {
    const bool above = !(tt < hroot);
    int k = 0;
    double tend = e + 1 < m.E ? m.t_start[e + 1] : DBL_MAX;
    const double tep = tend;
    if (!above) {
      while (c < used && T.h(c) <= tt) c++;
      if (c < used) { const double hc = T.h(c); if (hc < tend) tend = hc; }
      for (int v = 0; v < c; v++) {
        if (v < n && !((leaves >> v) & 1ull)) continue;
        if (T.par(v) >= c && (((marked >> v) & 1ull) ? s_pop_at(T, v, tt) : (int)T.pop(v)) == pf) k++;
      }
      const int nm = T.nm();
      for (int q = 0; q < nm; q++) { double mt = T.mt(q); if (mt > tt && mt < tend) tend = mt; }
    }
#ifdef SMC_CLOCKS
    if (threadIdx.x == 0 && vb_logw) atomicAdd(&g_clk[15], 1ull);      // intervals of thread 0's floats
#endif
    const double inv = s.inv2N[e * P + pf];
    const double coal_rate = above ? 0.0 : k * inv;
    const double mig_f = s.migtot[e * P + pf];
    const double mig_r = above ? s.migtot[e * P + pr] : 0.0;
    const double pw = (above && pf == pr) ? inv : 0.0;
    const double rate = coal_rate + mig_f + mig_r + pw;
    double w = rate > 0.0 ? -log(rng_uniform(r)) / rate : DBL_MAX;
    const bool hit = tt + w < tend;
    if (!hit && tend == DBL_MAX) { target = -3; coal_pop = pf; T.raise(SFLAG_NOCOAL); return tt; }
    const double tev = hit ? tt + w : tend;
    int ev_f = 0, to_f = -1, ev_r = 0, to_r = -1;
    if (hit) {
      double u = rng_uniform(r) * rate;
      if (u < coal_rate) ev_f = 1;
      else if ((u -= coal_rate) < mig_f) {
        ev_f = 2;
        for (int j = 0; j < P; j++) { if (j == pf) continue; to_f = j; u -= s.mig[(e * P + pf) * P + j]; if (u < 0) break; }
      } else if ((u -= mig_f) < mig_r) {
        ev_r = 2;
        for (int j = 0; j < P; j++) { if (j == pr) continue; to_r = j; u -= s.mig[(e * P + pr) * P + j]; if (u < 0) break; }
      } else ev_f = 1;                                   // pairwise coalescence, booked on the floating lineage
    }
    if (vb_logw && s.vb_coal) {                          // particle.cpp:262-272, before the record masks apply
      if (ev_f == 1) *vb_logw += s.vb_coal[e * P + pf];
      if (ev_f == 2) *vb_logw += s.vb_mig[(e * P + pf) * P + to_f];
      if (ev_r == 2) *vb_logw += s.vb_mig[(e * P + pr) * P + to_r];
    }
    if ((m.rec_mask[e] & 2) && e <= max_rec_epoch) {     // particle.cpp:251-300
      // Consecutive intervals of one lineage in one (epoch, population) add to the same cells: they are summed in
      // registers and written when an event has to be recorded with them or the epoch changes -- a float that crosses
      // ten nodes makes one round trip to its cells, not ten.
      const double dt = tev - tt;
      const int weight = k + (pw > 0.0 ? 1 : 0);
      a_opp += weight * dt;
      a_dt += dt;
      if (above) r_dt += dt;
      if (hit || tend == tep) {
        sink.srec(e, pf, x, ev_f == 1, a_opp, ev_f == 2 ? to_f : -1, a_dt);
        if (r_dt > 0.0 || ev_r == 2) sink.smig(e, pr, x, ev_r == 2 ? to_r : -1, r_dt);
        a_opp = 0.0; a_dt = 0.0; r_dt = 0.0;
      }
    }
    const double ts = tt;                                // start of the interval: where the contemporaries were counted
    tt = tev;
    if (hit) {
      if (ev_f == 1) {
        coal_pop = pf;
        if (above) { target = -1; return tt; }
        int idx = (int)(rng_uniform(r) * k);
        if (idx >= k) idx = k - 1;
        for (int v = 0; v < c; v++) {                     // c still is the first slot above ts
          if (v < n && !((leaves >> v) & 1ull)) continue;
          if (T.par(v) >= c && (((marked >> v) & 1ull) ? s_pop_at(T, v, ts) : (int)T.pop(v)) == pf) { if (idx == 0) { target = v; return tt; } idx--; }
        }
        target = -2;                                     // cannot happen
        return tt;
      }
      if (ev_f == 2) { pf = to_f; s_add_mark(T, MARK_FLOATING, tt, pf); }
      if (ev_r == 2) { pr = to_r; s_add_mark(T, MARK_ROOT_LINEAGE, tt, pr); }
      continue;
    }
    if (tend == tep) {
      e++;
      s_fixed_time_events(m, T, e, tt, pf, MARK_FLOATING, r);
      if (!(tt < hroot)) s_fixed_time_events(m, T, e, tt, pr, MARK_ROOT_LINEAGE, r);
    }
  }
}

// move internal slot a to its place in the height order of slots n .. used-1, renaming parents, populations and marks
template <class TV>
__device__ inline void s_rotate(const ModelDev& m, const TV& T, int used, int a) {
  const int n = m.n;
  const double ha = T.h(a);
  int d = a;
  while (d + 1 < used && T.h(d + 1) < ha) d++;
  if (d == a) while (d - 1 >= n && T.h(d - 1) > ha) d--;
  if (d == a) return;
  const signed char pa = T.par(a), popa = T.pop(a);
  if (d > a) for (int j = a; j < d; j++) { T.hw(j) = T.h(j + 1); T.par(j) = T.par(j + 1); T.pop(j) = T.pop(j + 1); }
  else for (int j = a; j > d; j--) { T.hw(j) = T.h(j - 1); T.par(j) = T.par(j - 1); T.pop(j) = T.pop(j - 1); }
  T.hw(d) = ha; T.par(d) = pa; T.pop(d) = popa;
  const int lo = d > a ? a : d, hi = d > a ? d : a, shift = d > a ? -1 : 1;
  for (int v = 0; v < used; v++) {
    int q = T.par(v);
    if (q >= lo && q <= hi) T.par(v) = (signed char)(q == a ? d : q + shift);
  }
  const int nm = T.nm();
  for (int k = 0; k < nm; k++) {
    int q = T.mb(k);
    if (q >= lo && q <= hi) T.mb(k) = (signed char)(q == a ? d : q + shift);
  }
}

// Forest::buildInitialTree (forest.cc:299-347): leaf 0 is the tree; every further leaf floats up from its own
// population and coalesces into the tree built so far (or pairwise with its root's lineage)
template <int BS>
__device__ inline void init_tree(const ModelDev& m, const STreeViewT<BS>& T, Rng& r) {
  const int n = m.n, nn = m.nn;
  const SModelDev& s = *T.sm;
  for (int v = 0; v < nn; v++) { T.par(v) = -1; T.pop(v) = 0; if (v >= n) T.hw(v) = 0.0; }
  for (int i = 0; i < n; i++) T.pop(i) = s.leaf_pop[i];
  T.nm() = 0;
  unsigned long long present = 1ull;
  int root = 0, used = n;
  SNullSink none;
  for (int i = 1; i < n; i++) {
    int target = -1, cpop = 0;
    double tc = s_float(m, T, present, root, r, 0.0, s.leaf_pop[i], target, cpop, none, 0.0, -1, nullptr);
    if (target == -3) return;
    const int nw = used++;
    present |= (1ull << i) | (1ull << nw);
    T.hw(nw) = tc; T.pop(nw) = (signed char)cpop;
    s_move_marks(T, MARK_FLOATING, i, 0.0, DBL_MAX);
    T.par(i) = (signed char)nw;
    if (target < 0) {                           // new root above the old one
      s_move_marks(T, MARK_ROOT_LINEAGE, root, 0.0, DBL_MAX);
      T.par(root) = (signed char)nw;
      T.par(nw) = -1;
      root = nw;
    } else {
      T.par(nw) = T.par(target);
      T.par(target) = (signed char)nw;
      s_move_marks(T, target, nw, tc, DBL_MAX);
    }
    s_rotate(m, T, used, nw);
    for (int v = 0; v < nn; v++) if (((present >> v) & 1ull) && T.par(v) < 0) root = v;
    // s_rotate renames slots: `present` is a set of the first `used` internal slots and the first i+1 leaves either way
  }
}

// One SMC' transition at position x with populations (Forest::sampleNextGenealogy under window 0)
template <int BS, class Sink>
__device__ inline void smc_move(const ModelDev& m, const STreeViewT<BS>& T, Rng& r, double x, int max_rec_epoch,
                                Sink& sink, double& logw, double tl) {
  const int n = m.n, nn = m.nn;
  // 1. recombination point, uniform on the local tree
  double goal = rng_uniform(r) * tl;
  int b = -1;
  double hh = 0.0, cum = 0.0;
  for (int v = 0; v < nn; v++) {
    int p = T.par(v);
    if (p < 0) continue;
    double hv = T.h(v);
    double bl = T.h(p) - hv;
    b = v; hh = hv + (goal - cum);
    if (goal < cum + bl) break;
    cum += bl;
  }
  const int Pn = T.par(b);
  if (hh > T.h(Pn)) hh = T.h(Pn);
  {
    int e = epoch_of(m, hh);
    if ((m.rec_mask[e] & 1) && e <= max_rec_epoch) sink.rec_event(e, x);
  }
  const int root = nn - 1;
  int S = -1;
  for (int v = 0; v < nn; v++) if (T.par(v) == Pn && v != b) S = v;
  // 2. float
  int target = -1, cpop = 0;
  const unsigned long long all = nn >= 64 ? ~0ull : ((1ull << nn) - 1ull);
#ifdef SMC_CLOCKS
  long long clk3 = clock64();
#endif
  const double tc = s_float(m, T, all, root, r, hh, s_pop_at(T, b, hh), target, cpop, sink, x, max_rec_epoch, &logw);
  SMC_CLK(7, clk3);                                    // the float alone
  if (target <= -2) return;
  // 3. surgery.  The stub of b above the cut disappears unless the lineage coalesces back into it.
  if (target == b) {
    s_drop_marks(T, b, hh, tc);
    s_move_marks(T, MARK_FLOATING, b, 0.0, DBL_MAX);
    return;
  }
  s_drop_marks(T, b, hh, DBL_MAX);
  s_move_marks(T, MARK_FLOATING, b, 0.0, DBL_MAX);
  if (target < 0) {                                    // pairwise above the old root
    if (Pn != root) {
      s_move_marks(T, Pn, S, 0.0, DBL_MAX);            // S's lineage now runs through where Pn was
      T.par(S) = T.par(Pn);
      s_move_marks(T, MARK_ROOT_LINEAGE, root, 0.0, DBL_MAX);
      T.par(root) = (signed char)Pn;
      T.par(Pn) = -1;
    } else {
      s_move_marks(T, MARK_ROOT_LINEAGE, S, 0.0, DBL_MAX);   // the root lineage is S's
    }
  } else if (target == Pn) {                           // into the branch above Pn: Pn slides up, S inherits the part below
    s_move_marks(T, Pn, S, 0.0, tc);
  } else if (target == S) {                            // into S's branch: Pn slides down, taking S's marks above tc
    s_move_marks(T, S, Pn, tc, DBL_MAX);
  } else {
    if (T.par(Pn) >= 0) s_move_marks(T, Pn, S, 0.0, DBL_MAX);
    else s_drop_marks(T, S, 0.0, DBL_MAX);             // Pn was the root: S becomes the root and has no branch above
    T.par(S) = T.par(Pn);
    T.par(Pn) = T.par(target);
    T.par(target) = (signed char)Pn;
    s_move_marks(T, target, Pn, tc, DBL_MAX);
  }
  T.hw(Pn) = tc; T.pop(Pn) = (signed char)cpop;
  s_rotate(m, T, nn, Pn);
  s_drop_marks(T, nn - 1, 0.0, DBL_MAX);               // the root has no branch above it
}

}  // namespace smc
