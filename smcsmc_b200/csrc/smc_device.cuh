// Device functions of the SMCSMC particle filter for sm_100a: one thread per particle, the particle's local tree
// staged in shared memory ([node][thread] so that a warp touches consecutive banks).  Compiled with -fmad=false: the
// arithmetic is operation-for-operation the reference's (an x86-64 build has no FMA contraction), which is what lets
// the pruning likelihood and the no-mutation factor meet 1e-12 against the reference (they are in fact bit-identical).
//
// Reference behaviour restated here (paths under /root/reference/src):
//   fastexp                         particle.cpp:30-40
//   calculate_likelihood            particle.cpp:625-680   + phasing enumeration particleContainer.cpp:138-224
//   trackLocalTreeBranchLength      particle.cpp:699-730
//   extend_ARG                      particle.cpp:743-918
//   sampleNextGenealogy (SMC')      scrm/src/forest.cc:423-598, time_interval.cc:84-134, forest.cc:601-756
//   sampleNextBase                  particle.cpp:1195-1254
//   record_all_event / record_recomb_extension / record_recomb_event   particle.cpp:193-390
#pragma once
#include <cstdint>
#include <cfloat>
#include <cmath>

namespace smc {

// ------------------------------------------------------------------------------------------------------------------
// Philox4x32-10, counter = (block, slot, step, chunk | purpose << 24), key = seed.  Replaces the reference's single
// global mt19937_64 (scrm/src/random/mersenne_twister.cc): streams differ by design, every (particle, row) has its own.
// ------------------------------------------------------------------------------------------------------------------
enum { RNG_INIT = 1, RNG_EXTEND = 2, RNG_RESAMPLE = 3, RNG_REDRAW = 4 };

struct Rng {
  uint32_t k0, k1, c0, c1, c2, c3;
  double spare;
  int have;
};

// One out-of-line copy of the Philox rounds and of the fp64 logarithm: the row kernel calls them from half a dozen
// places; out of line the kernel needs 96 instead of 122 registers (5 instead of 4 blocks of 128 per SM) and is ~5 %
// faster end to end (profiles/r1_row_kernel_summary.md).  -DSMC_INLINE_MATH restores the inlined build.
#if defined(__CUDA_ARCH__) && !defined(SMC_INLINE_MATH)
#define SMC_RNG_ATTR __noinline__
#else
#define SMC_RNG_ATTR inline
#endif
__host__ __device__ SMC_RNG_ATTR void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll
  for (int round = 0; round < 10; round++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
    uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__host__ __device__ inline double u53(uint32_t a, uint32_t b) {
  // 53 random bits + half an ulp: strictly inside (0,1)
  return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6) + 0.5) * (1.0 / 9007199254740992.0);
}

__host__ __device__ inline void rng_init(Rng& r, uint64_t seed, uint32_t slot, uint32_t step, uint32_t chunk,
                                         uint32_t purpose) {
  r.k0 = (uint32_t)seed; r.k1 = (uint32_t)(seed >> 32);
  r.c0 = 0; r.c1 = slot; r.c2 = step; r.c3 = (chunk & 0xFFFFFFu) | (purpose << 24);
  r.have = 0; r.spare = 0.0;
}

__host__ __device__ inline double rng_uniform(Rng& r) {
  if (r.have) { r.have = 0; return r.spare; }
  uint32_t o[4];
  philox4x32_10(r.c0, r.c1, r.c2, r.c3, r.k0, r.k1, o);
  r.c0++;
  r.spare = u53(o[2], o[3]);
  r.have = 1;
  return u53(o[0], o[1]);
}

// one out-of-line copy of the fp64 logarithm for the hot loop (the inlined expansion is ~100 instructions per call site)
#ifndef SMC_INLINE_MATH
__device__ __noinline__ double dlog(double x) { return log(x); }
#else
__device__ __forceinline__ double dlog(double x) { return log(x); }
#endif
#ifdef SMC_NOINLINE_EXP
__device__ __noinline__ double dexp(double x) { return exp(x); }
#else
__device__ __forceinline__ double dexp(double x) { return exp(x); }
#endif

// particle.cpp:30-40 -- continued-fraction exp; same rational expression, no contraction
__host__ __device__ inline double fastexp(double x) {
  double xx = x * x;
  if (xx < 0.516167859) return 1 + 2 * x / (2 - x + xx / (6 + xx * 0.1));
  return exp(x);
}
// The same function with the two divisions folded into one: 1 + 2x(6 + 0.1xx) / ((2 - x)(6 + 0.1xx) + xx).  Differs from
// fastexp in the last bits only (a few 1e-16 relative); used for the 2n-2 branch factors of the pruning pass, whose
// parity bar against the reference is 1e-12 relative (tests/test_gpu_parity.py).  The no-mutation factor and the
// stand-alone entry point keep the reference's exact expression.
__host__ __device__ inline double fastexp1(double x) {
  double xx = x * x;
  if (xx < 0.516167859) { double d = 6 + xx * 0.1; return 1 + 2 * x * d / ((2 - x) * d + xx); }
  return exp(x);
}

// ------------------------------------------------------------------------------------------------------------------
// Model constants shared by all threads (kernel parameter block; lives in the constant bank)
// ------------------------------------------------------------------------------------------------------------------
struct ModelDev {
  int n, nn, E;
  int ancestral_aware;
  double mu, rho, seqlen;
  double t_start[64];
  double inv2N[64];
  double vb_logfac[64];
  unsigned char rec_mask[64];
};

// Per-thread view of the particle's tree in shared memory.  Only what varies is stored: the heights of the n-1 internal
// nodes (leaves sit at 0), the parent slots, likelihood scratch for the internal nodes, and one accumulator per epoch.
template <int BS_>
struct TreeViewT {
  static constexpr int BS = BS_;   // compile-time block size: column addressing v * BS + t folds into shifts
  double* H;          // [n-1][BS] internal-node heights, addressed by node id (base shifted by -n rows)
  signed char* PAR;   // [nn][BS]
  double* L0;         // [n-1][BS] scratch for the internal nodes, addressed by node id (base shifted by -n rows)
  double* L1;         // [n-1][BS]
  double* ACC;        // [E][BS]  recombination opportunity summed over the current row
  double* ACC2;       // [E][BS]  the same beyond the first bucket bound inside the row; shares the storage of L0 / L1
  int t, n;
  __device__ __forceinline__ double h(int v) const { return v < n ? 0.0 : H[v * BS + t]; }
  __device__ __forceinline__ double& hw(int v) const { return H[v * BS + t]; }        // internal nodes only
  __device__ __forceinline__ signed char& par(int v) const { return PAR[v * BS + t]; }
  __device__ __forceinline__ double& l0(int v) const { return L0[v * BS + t]; }
  __device__ __forceinline__ double& l1(int v) const { return L1[v * BS + t]; }
};

// Carve a block's dynamic shared memory into the per-thread columns of a TreeView.
// scratch rows shared by L0 / L1 (2 (n-1) rows, used by the pruning passes) and ACC2 (E rows, used between them)
__host__ __device__ inline size_t tree_scratch_rows(int nn, int E) { return (size_t)(nn - 1) > (size_t)E ? (size_t)(nn - 1) : (size_t)E; }
__host__ __device__ inline size_t tree_bytes_per_thread(int nn, int E) {
  return (size_t)(nn - 1) / 2 * 8 + (size_t)nn + tree_scratch_rows(nn, E) * 8 + (size_t)E * 8;
}
template <int BS>
__device__ __forceinline__ TreeViewT<BS> make_tree_view(unsigned char* smem, int n, int nn, int E, int t) {
  TreeViewT<BS> T;
  T.t = t; T.n = n;
  double* base = reinterpret_cast<double*>(smem);
  const size_t in = (size_t)(nn - n) * BS;      // one row per internal node
  T.H = base - (size_t)n * BS;
  T.L0 = base + in - (size_t)n * BS;
  T.L1 = base + 2 * in - (size_t)n * BS;
  T.ACC2 = base + in;
  T.ACC = base + in + tree_scratch_rows(nn, E) * BS;
  T.PAR = reinterpret_cast<signed char*>(T.ACC + (size_t)E * BS);
  return T;
}

// forest.h:188 getLocalTreeLength: sum of branch lengths in slot order (the order the parity tests assume)
template <class TV>
__device__ inline double tree_length(const ModelDev& m, const TV& T) {
  // leaves sit at height 0: their term is h(parent) - 0.0 == h(parent) exactly, so splitting
  // the loop changes no bit and spares the leaf test on every height access
  double len = 0.0;
  const int n = m.n;
  for (int v = 0; v < n; v++) { int p = T.par(v); if (p >= 0) len += T.hw(p); }   // p < 0: a tree that failed to join
  for (int v = n; v < m.nn; v++) {
    int p = T.par(v);
    if (p >= 0) len += T.hw(p) - T.hw(v);
  }
  return len;
}

// particle.cpp:699-730 trackLocalTreeBranchLength: length of the subtree spanning the leaves that carry data.
// Iterative form of the reference's recursion: nodes are visited children-first (leaves, then internal slots in
// increasing height), each node hands its value to its parent; the lower-numbered child arrives first, which is the
// reference's left/right order.
template <class TV>
__device__ inline double branch_length_with_data(const ModelDev& m, const TV& T, const signed char* data) {
  const int n = m.n, nn = m.nn;
  for (int v = n; v < nn; v++) { T.l0(v) = -2.0; T.l1(v) = -2.0; }   // -2 = child value not yet delivered
  double total = 0.0;
  for (int v = 0; v < nn; v++) {
    double s;
    if (v < n) {
      s = data[v] >= 0 ? 0.0 : -1.0;
    } else {
      double sl = T.l0(v), sr = T.l1(v);
      if (sl >= 0.0 && sr >= 0.0) { total = sl + sr; s = total; }
      else if (sl >= 0.0) s = sl;
      else s = sr;
    }
    int p = T.par(v);
    if (p >= 0) {
      double up = s;
      if (up >= 0.0) up += T.hw(p) - T.h(v);
      if (T.l0(p) == -2.0) T.l0(p) = up; else T.l1(p) = up;
    }
  }
  return total;
}

// particle.cpp:625-680: two-state Felsenstein pruning for one fully specified haplotype (0 / 1 / -1 missing).
template <class TV>
__device__ inline double prune_likelihood(const ModelDev& m, const TV& T, const signed char* hap) {
  const int n = m.n, nn = m.nn;
  for (int v = n; v < nn; v++) { T.l0(v) = 1.0; T.l1(v) = 1.0; }
  double r0 = 1.0, r1 = 1.0;
  for (int v = 0; v < nn; v++) {
    double a0, a1;
    if (v < n) {
      int s = hap[v];
      a0 = s == 1 ? 0.0 : 1.0;
      a1 = s == 0 ? 0.0 : 1.0;
    } else {
      a0 = T.l0(v); a1 = T.l1(v);
    }
    int p = T.par(v);
    if (p < 0) { r0 = a0; r1 = a1; continue; }
    double t = T.hw(p) - T.h(v);
    double pr = fastexp1(-t * m.mu);
    double q = 1 - pr;
    // 1.0 * x == x exactly, so the first child's factor is stored as is and the second multiplies it: (left)*(right)
    T.l0(p) = T.l0(p) * (a0 * pr + a1 * q);
    T.l1(p) = T.l1(p) * (a1 * pr + a0 * q);
  }
  double prior0 = m.ancestral_aware ? 1.0 : 0.5, prior1 = m.ancestral_aware ? 0.0 : 0.5;
  return r0 * prior0 + r1 * prior1;
}

// particleContainer.cpp:138-181,204-221: mean likelihood over the 2^k phasings of the unphased het pairs (code 2).
template <class TV>
__device__ inline double site_likelihood(const ModelDev& m, const TV& T, const signed char* data) {
  signed char hap[32];
  const int n = m.n;
  int ncfg = 1;
  for (int i = 0; i < n; i++) hap[i] = data[i];
  for (int i = 0; i + 1 < n; i += 2)
    if (data[i] == 2) { ncfg *= 2; hap[i] = 0; hap[i + 1] = 1; }
  double norm = 1.0 / ncfg, acc = 0.0;
  for (;;) {
    acc += prune_likelihood(m, T, hap);
    if (ncfg <= 1) break;
    bool advanced = false;
    for (int i = 0; i + 1 < n; i += 2) {
      if (data[i] != 2) continue;
      if (hap[i] == 0) { hap[i] = 1; hap[i + 1] = 0; advanced = true; break; }
      hap[i] = 0; hap[i + 1] = 1;
    }
    if (!advanced) break;
  }
  return acc * norm;
}

// Recombination opportunity of the stretch [xa, xb): the local branch length inside each epoch times the stretch
// (record_recomb_extension, particle.cpp:305-357: sum over time slices of #lineages * slice length).  Internal slots are
// sorted by height and all leaves sit at 0, so the slice between consecutive internal nodes carries n, n-1, ...
// lineages; one ascending sweep completes the epochs in order and hands each one to the sink as soon as it is known.
// Recombination opportunity per bp of the current tree inside each epoch (= the tree's branch length in the epoch),
// handed to the sink for the stretch [xa, xb).  One merged walk over the height-sorted internal slots and the epoch
// bounds: every iteration consumes whichever comes first, so the lanes of a warp stay in step (a loop over the nodes
// with an inner loop over the bounds diverged on every tree: 25 % of a block's life at 11 epochs).
template <class TV, class Sink>
__device__ inline void record_recomb_opportunity(const ModelDev& m, const TV& T, Sink& sink, double xa, double xb,
                                                 int max_rec_epoch) {
  const int n = m.n, nn = m.nn, E = m.E;
  int e = 0, k = n, j = n;
  double lo = 0.0, len = 0.0;
  double hi = T.hw(j);
  double b = e + 1 < E ? m.t_start[e + 1] : DBL_MAX;
  for (;;) {
    if (b < hi) {                        // the epoch ends below the next node
      if (b > lo) { len += k * (b - lo); lo = b; }
      if (len > 0.0 && (m.rec_mask[e] & 1) && e <= max_rec_epoch) sink.rec_opp(e, xa, xb, len);
      len = 0.0;
      e++;
      b = e + 1 < E ? m.t_start[e + 1] : DBL_MAX;
    } else {                             // the next node
      if (hi > lo) len += k * (hi - lo);
      lo = hi;
      k--;
      if (++j == nn) break;
      hi = T.hw(j);
    }
  }
  if (len > 0.0 && (m.rec_mask[e] & 1) && e <= max_rec_epoch) sink.rec_opp(e, xa, xb, len);
}

__device__ inline int epoch_of(const ModelDev& m, double t) {
  int e = 0;
  while (e + 1 < m.E && m.t_start[e + 1] <= t) e++;
  return e;
}

// Slot s holds a changed height: move it to its place in the ascending order of the internal slots and rename the parent
// pointers accordingly.  One rotation; the result is that of the chain of adjacent slot swaps it replaces (5 % of the
// sweep: every swap was a pass over all parent pointers).
template <class TV>
__device__ inline void resort_slot(const ModelDev& m, const TV& T, int s) {
  const int n = m.n, nn = m.nn;
  const double hs = T.hw(s);
  int d = s;
  while (d + 1 < nn && T.hw(d + 1) < hs) d++;
  if (d == s) while (d - 1 >= n && T.hw(d - 1) > hs) d--;
  if (d == s) return;
  const signed char ps = T.par(s);
  if (d > s) { for (int k = s; k < d; k++) { T.hw(k) = T.hw(k + 1); T.par(k) = T.par(k + 1); } }
  else { for (int k = s; k > d; k--) { T.hw(k) = T.hw(k - 1); T.par(k) = T.par(k - 1); } }
  T.hw(d) = hs; T.par(d) = ps;
  const int lo = d > s ? s : d, hi = d > s ? d : s, shift = d > s ? -1 : 1;
  for (int v = 0; v < nn; v++) {
    const int p = T.par(v);
    if (p == s) T.par(v) = (signed char)d;
    else if (p >= lo && p <= hi) T.par(v) = (signed char)(p + shift);
  }
}

// Sink for the tally records a move / a stretch produces (the content of the reference's EvolutionaryEvent records).
// Implemented by the engine (lag-bucket ring in HBM) and by a null sink for the stand-alone kernels.
struct NullSink {
  __device__ void coal(int, double, int, double) const {}
  __device__ void rec_event(int, double) const {}
  __device__ void rec_opp(int, double, double, double) const {}
  __device__ void row_begin(double) const {}
  __device__ void row_end() const {}
};

// One SMC' transition at position x (Forest::sampleNextGenealogy with exact window 0).  The test-side CPU restatement
// consumes the same Philox draws in the same order, so the two can be compared draw for draw.
template <class TV, class Sink>
__device__ inline void smc_move(const ModelDev& m, const TV& T, Rng& r, double x, int max_rec_epoch, Sink& sink,
                                double& logw, double tl) {
  const int n = m.n, nn = m.nn;
  // 1. recombination point, uniform on the local tree (forest.cc:370-412); tl = tree_length of the current tree
  double target = rng_uniform(r) * tl;
  int b = -1;
  double hh = 0.0, cum = 0.0;
  for (int v = 0; v < nn; v++) {
    int p = T.par(v);
    if (p < 0) continue;
    double hv = T.h(v);
    double bl = T.hw(p) - hv;
    b = v; hh = hv + (target - cum);
    if (target < cum + bl) break;
    cum += bl;
  }
  int Pn = T.par(b);
  if (hh > T.hw(Pn)) hh = T.hw(Pn);
  int e = epoch_of(m, hh);
  if ((m.rec_mask[e] & 1) && e <= max_rec_epoch) sink.rec_event(e, x);
  // 2. the cut lineage floats upwards (time_interval.cc:84-134, forest.cc:601-696)
  const int root = nn - 1;                 // internal slots are height-sorted
  const double hroot = T.hw(root);
  int c = n;                               // next internal slot above tt
  while (c < nn && T.hw(c) <= hh) c++;
  double tt = hh, tc = 0.0;
  int k = 0;
  bool above_root = false;
  // coalescence opportunity of consecutive intervals inside one epoch is summed in a register and recorded once
  double opp = 0.0;
  bool rec_on = (m.rec_mask[e] & 2) && e <= max_rec_epoch;
  // ONE unit exponential for the whole float: an interval of length d at rate r uses up r * d of it, the lineage
  // coalesces in the interval in which it runs out (memoryless, the same law as a fresh draw per interval; the reference
  // re-uses its unit exponential the same way, random_generator.cc:32-78)
  double ex = -dlog(rng_uniform(r));
  for (;;) {
    double tnode = DBL_MAX;
    above_root = !(tt < hroot);
    if (!above_root) { tnode = T.hw(c); k = n - (c - n); }
    else k = 1;
    double tep = e + 1 < m.E ? m.t_start[e + 1] : DBL_MAX;
    double tend = tnode < tep ? tnode : tep;
    double rate = k * m.inv2N[e];
    double need = rate * (tend - tt);
    bool coal = ex < need;
    if (coal) { tc = tt + ex / rate; if (!(tc < tend)) tc = nextafter(tend, 0.0); }
    double dt = (coal ? tc : tend) - tt;
    opp += k * dt;
    if (coal) break;
    ex -= need;
    tt = tend;
    if (tend == tep) {
      if (rec_on) sink.coal(e, x, 0, opp);
      opp = 0.0;
      e++;
      rec_on = (m.rec_mask[e] & 2) && e <= max_rec_epoch;
    }
    if (!above_root && tend == tnode) c++;
  }
  if (rec_on) sink.coal(e, x, 1, opp);
  logw += m.vb_logfac[e];
  // 3. target among the k contemporaries (contemporaries_container.h:343-365) and 4. surgery; the unary node left at
  // the cut is pruned (forest.cc:1245-1257): the parent slot of the cut branch becomes the new coalescence node.
  int S = -1;
  for (int v = 0; v < nn; v++) if (T.par(v) == Pn && v != b) S = v;
  if (above_root) {
    if (Pn != root) {
      T.par(S) = T.par(Pn);
      T.par(root) = (signed char)Pn;
      T.par(Pn) = -1;
    }
    T.hw(Pn) = tc;
  } else {
    int idx = (int)(rng_uniform(r) * k);
    if (idx >= k) idx = k - 1;
    int Tg = -1;
    // slots are height-sorted and c is the first one above tt: the branches alive at tt are the v < c with parent >= c
    for (int v = 0; v < c; v++) {
      if (T.par(v) >= c) { if (idx == 0) { Tg = v; break; } idx--; }
    }
    if (Tg == b) {
      // back-coalescence into the stub: tree unchanged (SMC')
    } else if (Tg == S || Tg == Pn) {
      T.hw(Pn) = tc;
    } else {
      T.par(S) = T.par(Pn);
      T.par(Pn) = T.par(Tg);
      T.par(Tg) = (signed char)Pn;
      T.hw(Pn) = tc;
    }
  }
  resort_slot(m, T, Pn);
}

// particle.cpp:1195-1254 with multiplicity 1 and one rho segment; forest.h:140-151 for the re-draw after resampling
__device__ inline double draw_next_base(const ModelDev& m, double tree_len, Rng& r, double from) {
  double rate = m.rho * tree_len;
  double len = -dlog(rng_uniform(r)) / rate;
  double next = from + len;
  if (next == from) next = nextafter(from, from * 2);
  if (!(next < m.seqlen)) next = m.seqlen;
  return next;
}

// ForestState::extend_ARG for one row (x0, x1] with alleles data[n], followed by update_weight_at_site when is_site.
// Returns the number of recombinations sampled.
template <class TV, class Sink>
__device__ inline int extend_row(const ModelDev& m, const TV& T, Rng& r, double x0, double x1,
                                 const signed char* data, int is_site, int max_rec_epoch, Sink& sink, double& logw,
                                 double& next_base) {
  const int n = m.n;
  int nmiss = 0;
  for (int i = 0; i < n; i++) nmiss += data[i] == -1;
  const int leaf_status = nmiss == 0 ? 1 : (nmiss == n ? -1 : 0);   // particleContainer.cpp:103-111
  double tl = 0.0, L_obs = 0.0;
  sink.row_begin(x1);
  double updated_to = x0;
  int nrec = 0;
  bool moved = false;
  while (updated_to < x1) {
    // (re)derive what depends on the tree: once at the row start and once after every move -- a single copy of this
    // code keeps the kernel's instruction footprint down
    tl = tree_length(m, T);
    L_obs = leaf_status == -1 ? 0.0 : leaf_status == 1 ? tl : branch_length_with_data(m, T, data);
    if (moved) next_base = draw_next_base(m, tl, r, updated_to);
    double new_to = x1 < next_base ? x1 : next_base;
    logw += dlog(fastexp(-m.mu * L_obs * (new_to - updated_to)));     // particle.cpp:770
    record_recomb_opportunity(m, T, sink, updated_to, new_to, max_rec_epoch);
    updated_to = new_to;
    moved = false;
    // a stretch without a move ends the row: the loop condition fails and nothing tree-dependent is recomputed
    if (updated_to < x1) {
      smc_move(m, T, r, updated_to, max_rec_epoch, sink, logw, tl);
      nrec++;
      moved = true;
    }
  }
  sink.row_end();
  if (is_site) logw += dlog(site_likelihood(m, T, data));
  return nrec;
}

// ------------------------------------------------------------------------------------------------------------------
// extend_row with the SMC' moves compacted over the block.
//
// Recombinations per particle and row are Poisson (at the bench shape 35 % none, 37 % one, 28 % more): with one thread
// walking its own particle through `while (updated_to < x1)` a warp runs as long as its unluckiest lane, and the move
// code -- the bulk of the row -- executes with 8 of 32 lanes on average.  The trees sit in shared memory as
// [node][column], so any thread of the block can run any resident particle's move.  Here every thread does its own first
// stretch; then, round by round, the particles that still owe a move are listed (ballot + prefix over the warps) and
// threads 0..count-1 take one each: full warps run the move and the stretch after it, the other warps wait at the
// barrier without issuing.  Per particle the draws, the operations and their order are those of extend_row (the oracle
// stays the checker draw for draw); what a particle carries from round to round lives in shared memory (MoveSmem).
// ------------------------------------------------------------------------------------------------------------------
#ifdef SMC_CLOCKS
// build variant for profiles/: thread 0 of every block adds the cycles of its phases to g_clk (smc_debug_clocks)
__device__ unsigned long long g_clk[16];
#define SMC_CLK(slot, t0) do { if (threadIdx.x == 0) { long long now_ = clock64(); atomicAdd(&g_clk[slot], (unsigned long long)(now_ - (t0))); (t0) = now_; } } while (0)
#else
#define SMC_CLK(slot, t0) do { } while (0)
#endif

struct MoveSmem {
  double* X;              // [BS] position of the particle's next recombination (>= x1: no move owed)
  double* LOGW;           // [BS]
  double* TL;             // [BS] tree length of the current tree
  double* SPARE;          // [BS] second uniform of the last Philox call
  double* AREA;           // [BS] tree length x bp summed over the row's stretches (the row's recombination opportunity)
  uint32_t* CTR;          // [BS] Philox call counter of the (particle, row) stream
  unsigned char* HAVE;    // [BS] spare is valid
  unsigned char* LIST;    // [2][BS] columns that owe a move, by round parity
};
__host__ __device__ inline size_t move_bytes_per_thread() { return 5 * 8 + 4 + 1 + 2 + 1; }   // 48
template <int BS>
__device__ __forceinline__ MoveSmem make_move_smem(unsigned char* base) {
  MoveSmem ms;
  double* d = reinterpret_cast<double*>(base);
  ms.X = d; ms.LOGW = d + BS; ms.TL = d + 2 * BS; ms.SPARE = d + 3 * BS; ms.AREA = d + 4 * BS;
  ms.CTR = reinterpret_cast<uint32_t*>(d + 5 * BS);
  ms.HAVE = reinterpret_cast<unsigned char*>(ms.CTR + BS);
  ms.LIST = ms.HAVE + BS;
  return ms;
}

template <int BS, class TV, class Sink>
__device__ inline int extend_row_block(const ModelDev& m, const TV& T, Rng& r, double x0, double x1,
                                       const signed char* data, int is_site, int max_rec_epoch, Sink& sink, double& logw,
                                       double& next_base, bool live, const MoveSmem& ms, int* wcnt /* shared [2][BS/32] */,
                                       double& row_area) {
  const int n = m.n, t = threadIdx.x, lane = t & 31, warp = t >> 5;
  constexpr int nwarp = BS >> 5;
  int nmiss = 0;
  for (int i = 0; i < n; i++) nmiss += data[i] == -1;
  const int leaf_status = nmiss == 0 ? 1 : (nmiss == n ? -1 : 0);   // particleContainer.cpp:103-111
  int nrec = 0;
#ifdef SMC_CLOCKS
  long long clk = clock64();
#endif
  if (live) {
    sink.row_begin(x1);
    ms.X[t] = next_base; ms.LOGW[t] = logw; ms.TL[t] = 0.0; ms.AREA[t] = 0.0;
    ms.CTR[t] = r.c0; ms.SPARE[t] = r.spare; ms.HAVE[t] = (unsigned char)r.have;
  }
  // Round 0 is the stretch up to the particle's first recombination (every thread its own particle, no move); the
  // rounds after it run the compacted list of particles that owe a move.  One body serves both: the kernel is bound by
  // instruction fetch, a second copy of the stretch code costs more than the `first` tests.
  bool first = true;
  int total = BS, cur = 0;
  while (total > 0) {                  // uniform over the block
    __syncthreads();                   // the list of this round is complete (round 0: the columns above are written)
#ifdef SMC_CLOCKS
    if (threadIdx.x == 0 && !first) { atomicAdd(&g_clk[8], 1ull); atomicAdd(&g_clk[9], (unsigned long long)total); }
#endif
    bool again = false;
    int tp = t;
    if (first ? live : t < total) {
      if (!first) tp = ms.LIST[cur * BS + t];
      TV Tp = T; Tp.t = tp;
      Sink sp = sink; sp.t = tp; sp.i = sink.i - t + tp;
      Rng rp = r;
      rp.c1 = r.c1 - (uint32_t)t + (uint32_t)tp; rp.c0 = ms.CTR[tp]; rp.spare = ms.SPARE[tp]; rp.have = ms.HAVE[tp];
      double lw = ms.LOGW[tp];
      const double x = first ? x0 : ms.X[tp];          // where the stretch starts
#ifdef SMC_CLOCKS
      long long clk2 = clock64();
#endif
      if (!first) {
        smc_move(m, Tp, rp, x, max_rec_epoch, sp, lw, ms.TL[tp]);
        nrec++;
      }
      SMC_CLK(11, clk2);               // the move
      const double tl = tree_length(m, Tp);
      const double L_obs = leaf_status == -1 ? 0.0 : leaf_status == 1 ? tl : branch_length_with_data(m, Tp, data);
      SMC_CLK(12, clk2);               // tree length, observed length
      const double nb = first ? ms.X[tp] : draw_next_base(m, tl, rp, x);
      const double new_to = x1 < nb ? x1 : nb;
      if (x < x1) {                    // a row of zero length (two typed sites at one position) has no stretch
        lw += dlog(fastexp(-m.mu * L_obs * (new_to - x)));     // particle.cpp:770
        SMC_CLK(13, clk2);             // next base, no-mutation factor
        record_recomb_opportunity(m, Tp, sp, x, new_to, max_rec_epoch);
        SMC_CLK(14, clk2);             // recombination opportunity of the stretch
        ms.AREA[tp] += tl * (new_to - x);
      }
      again = new_to < x1;
      ms.X[tp] = nb; ms.LOGW[tp] = lw; ms.TL[tp] = tl;
      ms.CTR[tp] = rp.c0; ms.SPARE[tp] = rp.spare; ms.HAVE[tp] = (unsigned char)rp.have;
    }
    const unsigned bal = __ballot_sync(0xffffffffu, again);
    if (lane == 0) wcnt[(cur ^ 1) * nwarp + warp] = __popc(bal);
    __syncthreads();                   // counts of the next round are visible, every read of this round's list is done
    total = 0;
    int off = 0;
    for (int w = 0; w < nwarp; w++) { const int c = wcnt[(cur ^ 1) * nwarp + w]; if (w < warp) off += c; total += c; }
    if (again) ms.LIST[(cur ^ 1) * BS + off + __popc(bal & ((1u << lane) - 1u))] = (unsigned char)tp;
    cur ^= 1;
    if (first) SMC_CLK(3, clk); else SMC_CLK(4, clk);   // first stretch / move rounds
    first = false;
  }
  __syncthreads();                     // the last round's results are visible to their owners
  if (live) {
    logw = ms.LOGW[t]; next_base = ms.X[t]; row_area = ms.AREA[t];
    sink.row_end();
    if (is_site) logw += dlog(site_likelihood(m, T, data));
  }
  SMC_CLK(5, clk);                     // row-end flush + site likelihood (thread 0's own)
  return nrec;
}

// Prior tree at the left end (forest.cc:299-347): Kingman coalescent under piecewise-constant epochs; mirrors
// the test-side CPU restatement draw for draw.
template <class TV>
__device__ inline void init_tree(const ModelDev& m, const TV& T, Rng& r) {
  const int n = m.n, nn = m.nn;
  signed char active[32];
  for (int i = 0; i < n; i++) active[i] = (signed char)i;
  for (int v = 0; v < nn; v++) T.par(v) = -1;
  int k = n, e = 0, next_slot = n;
  double t = 0.0;
  while (k > 1) {
    double pairs = 0.5 * k * (k - 1);
    double rate = pairs * m.inv2N[e];
    double w = -log(rng_uniform(r)) / rate;
    double tend = e + 1 < m.E ? m.t_start[e + 1] : DBL_MAX;
    if (t + w < tend) {
      t += w;
      int idx = (int)(rng_uniform(r) * pairs);
      if (idx >= (int)pairs) idx = (int)pairs - 1;
      int i = 0;
      while (idx >= k - 1 - i) { idx -= k - 1 - i; i++; }
      int j = i + 1 + idx;
      T.hw(next_slot) = t;
      T.par(active[i]) = (signed char)next_slot;
      T.par(active[j]) = (signed char)next_slot;
      active[i] = (signed char)next_slot;
      active[j] = active[k - 1];
      next_slot++; k--;
    } else {
      t = tend; e++;
    }
  }
}

}  // namespace smc
