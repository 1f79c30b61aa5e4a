// libsmcgpu.so -- B200-native SMCSMC particle-filter engine: kernels + the C ABI of include/smcsmc_b200.h.
//
// One kernel launch per .seg row does everything the reference's pfARG_core loop body does (smcsmc.cpp:328-361):
//   phase 0  lagged tallies due after the previous row (CountModel::extract_and_update_count, count.cpp:355-415)
//   phase 1  if the previous row triggered resampling: systematic offspring assignment + gather of the parent's state
//            (ParticleContainer::resample / systematic_resampling / implement_resampling, particleContainer.cpp:247-392)
//   phase 2  extend_ARG over the row, no-mutation factor, SMC' moves, tally records, pruning likelihood at the site
//            (particle.cpp:743-918, forest.cc:423-598, particle.cpp:625-680)
//   phase 3  log-sum-exp partials per block; the LAST block to arrive combines them: ln normaliser, ESS, resampling
//            decision, the exact u_j chain of the systematic sampler (normalize_probability :420-438, resample :256-270)
// Nothing returns to the host inside the sweep: the resampling decision, the buffer flip and the row counter live in a
// device-side control block, so the R launches of a chunk are identical and queue back to back on one stream.
//
// HBM layout (structure of arrays, particle index fastest => every warp access is coalesced):
//   height  [n-1][N] f64   internal-node heights (leaves sit at 0), slots height-sorted
//   parent  [2n-1][N] i8
//   logw[N], next_base[N] f64
//   open    [E][N][4] f64      tallies of the lag bucket each epoch is writing now: coal count, coal opp, recomb count,
//                              recomb opp -- travels with the particle
// all of it twice (ping-pong: resampling gathers from one copy into the other), plus, for the CLOSED lag buckets,
//   frozen  [E][B][N][4] f64   written once, at the slot of the particle that owned the record when the bucket closed
//   map     [E][B][N] i32      (twice) which slot holds the record of this particle's ancestor: a resampling copies
//                              4 bytes per closed bucket instead of the 32-byte record
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>   // types only: the library is bound with dlopen when an engine is sharded
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <type_traits>
#include <vector>
#include "../../include/smcsmc_b200.h"
#include "smc_device.cuh"
#include "smc_struct_device.cuh"

using namespace smc;

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CUDA_TRY(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return fail(SMC_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(_e)); } while (0)

constexpr int MAXSEG = 160;

constexpr int SMC_MAX_RANKS = 8;
constexpr int AUX_BS = 64;   // block size of the one-off kernels (prior trees, stand-alone tree functions, calibration)

struct USeg { long long j0; double u0; double step; };   // u_j = u0 + (j - j0) * step for j in [j0, next j0)

struct Ctrl {
  long long row;            // next row to process
  int cur;                  // which copy of the state is current
  int resampled;            // the row just finished asked for resampling: the next launch pulls offspring
  double lse;               // ln normaliser of the last row; stored log-weights are un-normalised by this
  double lnZ;
  double norm;              // last partial sum (== total normalised weight)
  double child_logw;        // log(norm / N): weight of every offspring (particleContainer.cpp:350-351)
  double resample_pos;      // row end at which the resampling was decided (offspring re-draw from here)
  long long resample_row;
  long long cons_done;      // row whose leading tally consumption has already been done (by a materialise pass)
  int pool_c0, pool_c1;     // range of P.cons whose descendants' weights sit pooled in P.pool[pool_par] and still owe their
  int pool_par;             // second-moment tallies (the launch after the one that pooled them adds those)
  unsigned int ticket;
  int err;
  int sflags;               // SFLAG_* raised by any particle (structured populations)
  long long n_resample, n_recomb;
  long long gen;            // sweep generation of a sharded engine (smc_reset): tags the ranks' per-row mailbox flags
  int nseg;
  USeg seg[MAXSEG];
};

struct StateBuf {
  double* height;       // [n-1][N]
  signed char* parent;  // [nn][N]
  double* logw;
  double* next_base;
  double* open;         // [E][N][Q]   Q = 4 cells per epoch (2 + 3P + P*P with P > 1 populations)
  int* map;             // [E][B][N] global slot of the ancestor's frozen record
  // structured populations only (smc_struct_device.cuh)
  signed char* pop;     // [nn][N]
  double* mtime;        // [K][N]
  signed char* mbranch; // [K][N]
  signed char* mpop;    // [K][N]
  signed char* nmarks;  // [N]
};

struct Params {
  ModelDev m;
  SModelDev sm;            // sm.P > 1: structured populations (the <.., true> kernels)
  int N, B, parity_scan, nblocks, flat_search;
  uint64_t seed;
  uint32_t chunk, slot_offset;
  double ess_fraction, start_pos;
  StateBuf buf[2];
  Ctrl* ctrl;
  // rows
  long long R;
  const double* x_end;
  const signed char* alleles;
  const unsigned char* is_site;
  const int* max_rec;
  const int* kstart;       // [R+1][E] bucket index at the row's left end (row R: after the last row)
  const int* kslot;        // [R+1][E] kstart modulo the ring depth B
  const int* kfirst;       // [R+1][E] oldest bucket still pending once the row's own hand-over is done
  const unsigned long long* closes;   // [R] bit e: the bucket open at the row's left end closes by the row end
  double* frozen;          // [E][B][N][Q]
  const int* ub_off;       // [E+1]
  const double* ub;        // bucket upper bounds per epoch (last = +inf: the end-of-data bucket)
  const int* cons_off;     // [R+1]
  const int2* cons;        // (epoch, bucket) pairs due after each row
  // per-row weight partials, double-buffered by row parity
  double* local_cum[2];    // [N] inclusive prefix of exp(logw - bmax) inside the block
  double* bmax[2];
  double* bsum[2];
  double* bsq[2];
  double* barea[2];        // [nblocks] block sums of exp(logw - block max) * (tree length x bp of the row)
  double* row_opp;         // [R] posterior-mean recombination opportunity of each row (sum_i w_i int_row L_i dx)
  double* boff[2];         // [nblocks+1] exclusive prefix of normalised block totals
  double* bscale[2];
  double* psum;            // [N+1] parity-mode sequential partial sums
  double* tallies;         // [6][E]; structured: [2][Q][E] (sum of w * cell, sum of w^2 * cell)
  double* pool;            // [2][E][N] weight of the current particles pooled per ancestor slot (see phase 0), by parity
  int pool_on;             // 0: per-particle w^2 (tuning knob SMC_NO_POOL; sharded populations)
  double* pool2;           // [E][QW][N] per slot: sum of pool^2 * cell over the buckets whose record the slot owned (QW = 2
                           // opportunity cells, or all Q cells with populations); smc_finish adds it up into the tallies
  int* anc;                // [N] ancestry of the last resampling (global particle index when sharded)
  double* rs_pos; double* rs_ess; long long rs_cap;
  // particle sharding over ranks: this engine owns particles [base, base + N) of Ntot.  What a
  // pulling offspring reads of its parent -- tree, next_base, tallies, and the weight partials -- lives in two slabs
  // whose peer mappings differ from the local addresses by a constant byte offset per rank.
  int rank, world;
  long long Ntot, base;
  long long peer_state[SMC_MAX_RANKS];   // byte offset: peer r's state slab - mine
  long long peer_tally[SMC_MAX_RANKS];   // byte offset: peer r's tally slab (open, map, frozen) - mine
  double* mbox;             // [2][SMC_MAX_RANKS][8] per-row mailbox in the state slab, by row parity: peers store their
                            // (max, sum, sum of squares, prefix total) and a flag here over NVLink
  double* rank_off[2];      // [world+1] normalised exclusive prefix of the rank totals, by row parity
  double* rank_fac[2];      // [world] exp(rank max - ln normaliser), by row parity
};

template <class Tp>
__device__ __forceinline__ Tp* at_peer(Tp* p, long long delta) {
  return reinterpret_cast<Tp*>(reinterpret_cast<char*>(const_cast<typename std::remove_const<Tp>::type*>(p)) + delta);
}

// the frozen record (4 doubles) of closed bucket (e, slot s) that belongs to particle i's ancestor
template <bool SH>
__device__ __forceinline__ int load_frozen(const Params& P, const int* map, int e, int s, long long i, double out[4]) {
  const size_t es = (size_t)(e * P.B + s);
  const int m = map[es * P.N + i];
  const double* f = P.frozen;
  long long l = m;
  if constexpr (SH) { const int r = m / P.N; l = m - r * P.N; f = at_peer(f, P.peer_tally[r]); }
  const double2* c2 = reinterpret_cast<const double2*>(f + (es * P.N + l) * 4);
  double2 a = c2[0], b = c2[1];
  out[0] = a.x; out[1] = a.y; out[2] = b.x; out[3] = b.y;
  return m;
}

// Tally sink.  Records of the bucket that stays open beyond this row go to the particle's `open` cells, records of
// buckets that close by the row end go straight to their `frozen` cells (phase 1 has moved the carried values there).
// Recombination opportunity (the dense record: every epoch, every stretch) is summed in shared memory over the row and
// written once per epoch at the row end, unless a bucket bound of that epoch falls inside the row (then each stretch
// is split at the bound and written directly).
template <int BS, bool ST = false>
struct RingSink {
  const Params* P;
  double* open;    // destination copy of the open cells
  int i;           // particle index
  const int* ks;   // bucket index of this row's left end [E]
  const int* ksl;  // the same modulo the ring depth [E]
  const int* kn;   // bucket index of the NEXT row's left end [E]: the bucket that stays open
  double* ACC;     // [E][BS] shared
  int t;
  const double* ubf;          // [E] shared: upper bound of the bucket open at the row's left end
  unsigned long long split;   // bit e: a bound of epoch e lies inside this row
  const double* ubf2;         // [E] shared: the bound after ubf (the bucket born at ubf ends there)
  double* ACC2;               // [E][BS] shared: recombination opportunity of the row in [ubf, ubf2), or NULL (row with
                              // partly missing data: the scratch it aliases is in use, the cells are written through)
  // cells per epoch and where the recombination cells sit (single population: coal cnt, coal opp, rec cnt, rec opp)
  __device__ __forceinline__ int Q() const { return ST ? P->sm.Q : 4; }
  static constexpr int REC_CNT = ST ? 0 : 2, REC_OPP = ST ? 1 : 3;
  __device__ __forceinline__ int bucket(int e, double x) const {
    int k = ks[e];
    if (x < ubf[e]) return k;
    if (x < ubf2[e]) return k + 1;
    k += 2;
    const double* ub = P->ub + P->ub_off[e];
    while (ub[k] <= x) k++;          // last bound is +inf
    return k;
  }
  __device__ __forceinline__ double* cell(int e, int k, int q) const {
    if (k == kn[e]) return open + ((size_t)e * P->N + i) * Q() + q;
    int s = ksl[e] + (k - ks[e]);
    if (s >= P->B) s -= P->B;
    return P->frozen + ((size_t)(e * P->B + s) * P->N + i) * Q() + q;
  }
  __device__ void row_begin(double x1) {
    split = 0ull;
    for (int e = 0; e < P->m.E; e++) {
      ACC[e * BS + t] = 0.0;
      if (ubf[e] < x1) { split |= 1ull << e; if (ACC2) ACC2[e * BS + t] = 0.0; }
    }
  }
  __device__ void row_end() const {
    // four independent read-modify-writes in flight at a time (the cells of different epochs never alias)
    const int E = P->m.E;
    for (int e0 = 0; e0 < E; e0 += 4) {
      double a[4], v[4];
      double* cp[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        int e = e0 + j;
        a[j] = e < E ? ACC[e * BS + t] : 0.0;
        cp[j] = cell(e < E ? e : 0, ks[e < E ? e : 0], REC_OPP);
        v[j] = a[j] != 0.0 ? *cp[j] : 0.0;
      }
#pragma unroll
      for (int j = 0; j < 4; j++) if (a[j] != 0.0) *cp[j] = v[j] + a[j];
    }
    if (ACC2 && split) {               // the buckets born at the first bound inside the row
      for (int e0 = 0; e0 < E; e0 += 4) {
        double a[4], v[4];
        double* cp[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int e = e0 + j;
          const bool on = e < E && ((split >> e) & 1ull);
          a[j] = on ? ACC2[e * BS + t] : 0.0;
          cp[j] = cell(on ? e : 0, ks[on ? e : 0] + (on ? 1 : 0), REC_OPP);
          v[j] = a[j] != 0.0 ? *cp[j] : 0.0;
        }
#pragma unroll
        for (int j = 0; j < 4; j++) if (a[j] != 0.0) *cp[j] = v[j] + a[j];
      }
    }
  }
  __device__ void coal(int e, double x, int cnt, double opp) const {
    int k = (split >> e) & 1ull ? bucket(e, x) : ks[e];
    // count and opportunity sit side by side (16-byte aligned): one read-modify-write instead of two dependent ones.
    // (Tried and slower, profiles/r2_kernel_notes.md: reductions -- atomicAdd without a return value -- for these records
    // (9 %) and for the split path of rec_opp (1.9 x: the loop around an atomic no longer reconverges); queueing a move's
    // records and committing them together with all loads in flight (20 %: the queue lives in local memory).)
    double2* c2 = reinterpret_cast<double2*>(cell(e, k, 0));
    double2 v = *c2;
    v.x += cnt ? 1.0 : 0.0;
    v.y += opp;
    *c2 = v;
  }
  // structured records (particle.cpp:251-300): coalescence opportunity / event of a floating lineage in population p,
  // and its migration opportunity dt / event p -> to
  __device__ void scoal(int e, int p, double x, int cnt, double opp) const {
    int k = (split >> e) & 1ull ? bucket(e, x) : ks[e];
    if (cnt) *cell(e, k, cell_coal_cnt(p)) += 1.0;
    if (opp != 0.0) *cell(e, k, cell_coal_opp(p)) += opp;
  }
  __device__ void smig(int e, int p, double x, int to, double dt) const {
    int k = (split >> e) & 1ull ? bucket(e, x) : ks[e];
    double* base = cell(e, k, 0);
    double* po = base + cell_mig_opp(p);
    double* pm = base + cell_mig_cnt(P->sm.P, p, to >= 0 ? to : 0);
    const double vo = *po, vm = to >= 0 ? *pm : 0.0;      // both loads in flight before the first store
    *po = vo + dt;
    if (to >= 0) *pm = vm + 1.0;
  }
  // One interval of the floating lineage in population p: coalescence opportunity (and event), migration opportunity
  // (and event p -> to).  The three cells of a population are neighbours; everything is loaded before anything is stored,
  // so the interval costs one round trip to L2 instead of three or four dependent ones (the same additions to the same
  // cells as scoal + smig).
  __device__ void srec(int e, int p, double x, int cnt, double opp, int to, double dt) const {
    int k = (split >> e) & 1ull ? bucket(e, x) : ks[e];
    double* base = cell(e, k, 0);
    double* pc = base + cell_coal_cnt(p);                  // count, opportunity, migration opportunity
    double* pm = base + cell_mig_cnt(P->sm.P, p, to >= 0 ? to : 0);
    const double v0 = pc[0], v1 = pc[1], v2 = pc[2], vm = to >= 0 ? *pm : 0.0;
    if (cnt) pc[0] = v0 + 1.0;
    if (opp != 0.0) pc[1] = v1 + opp;
    pc[2] = v2 + dt;
    if (to >= 0) *pm = vm + 1.0;
  }
  __device__ void rec_event(int e, double x) const {
    int k = (split >> e) & 1ull ? bucket(e, x) : ks[e];
    *cell(e, k, REC_CNT) += 1.0;
  }
  __device__ void rec_opp(int e, double xa, double xb, double rate) const {
    // a stretch that ends before the first bound inside the row belongs to the bucket of the row's left end, whose
    // cell row_end() adds the accumulator to -- only what lies beyond that bound is written through
    if (!((split >> e) & 1ull) || xb <= ubf[e]) { ACC[e * BS + t] += rate * (xb - xa); return; }
    if (xa < ubf[e]) { ACC[e * BS + t] += rate * (ubf[e] - xa); xa = ubf[e]; }
    if (ACC2 && xa < ubf2[e]) {        // the common case: one bound inside the row
      const double end = xb < ubf2[e] ? xb : ubf2[e];
      ACC2[e * BS + t] += rate * (end - xa);
      xa = end;
      if (!(xa < xb)) return;
    }
    int k = bucket(e, xa);
    const double* ub = P->ub + P->ub_off[e];
    while (xa < xb) {
      double bend = ub[k] < xb ? ub[k] : xb;
      *cell(e, k, REC_OPP) += rate * (bend - xa);
      xa = bend;
      k++;
    }
  }
};

__device__ __forceinline__ double warp_sum(double v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// pool[key] += w, once per run of equal keys in the warp (segmented inclusive scan; the run's last lane adds).  Keys are
// ancestor slots: systematic resampling keeps the particles' order, so the descendants of one slot are neighbours and a
// slot with thousands of descendants costs a few atomics per warp instead of thousands on one address.  key < 0: no part.
__device__ __noinline__ void pool_add_runs(double* pool, int key, double w) {
  const int lane = threadIdx.x & 31;
  for (int d = 1; d < 32; d <<= 1) {
    const double ow = __shfl_up_sync(0xffffffffu, w, d);
    const int ok = __shfl_up_sync(0xffffffffu, key, d);
    if (lane >= d && ok == key) w += ow;
  }
  const int nk = __shfl_down_sync(0xffffffffu, key, 1);
  if (key >= 0 && (lane == 31 || nk != key) && w != 0.0) atomicAdd(pool + key, w);
}
__device__ __forceinline__ void pool_add(double* pool, int key, double w) {
  const int key0 = __shfl_sync(0xffffffffu, key, 0);
  if (__all_sync(0xffffffffu, key == key0)) {          // the whole warp descends from one slot (the usual case at depth)
    w = warp_sum(w);
    if ((threadIdx.x & 31) == 0 && key >= 0 && w != 0.0) atomicAdd(pool + key, w);
    return;
  }
  const int pk = __shfl_up_sync(0xffffffffu, key, 1);
  if (__all_sync(0xffffffffu, (threadIdx.x & 31) == 0 || pk != key)) {   // no two neighbours share a slot (shallow buckets)
    if (key >= 0 && w != 0.0) atomicAdd(pool + key, w);
    return;
  }
  pool_add_runs(pool, key, w);                         // out of line: the row kernel is bound by instruction fetch
}
// normalised cumulative weight c_i = P_{i+1} (un-divided) of particle i after row parity p
__device__ __forceinline__ double cum_at(const Params& P, int p, long long i, int bs) {
  if (P.parity_scan) return P.psum[i + 1];
  int b = (int)(i / bs);
  double v = P.boff[p][b] + __ldcg(P.local_cum[p] + i) * P.bscale[p][b];
  double cap = P.boff[p][b + 1];
  return v < cap ? v : cap;
}

// the same for global particle g when the population is sharded: the owner's arrays hold partials relative to the
// owner's own maximum (all written before the per-row exchange), rank_fac rescales them and rank_off places them; the
// last particle of a rank sits exactly on the rank's upper bound, so the owner of a parent follows from rank_off alone
__device__ __forceinline__ double cum_at_sharded(const Params& P, int p, int r, long long l, int bs) {
  const long long d = P.peer_state[r];
  const double hi = P.rank_off[p][r + 1];
  if (l == P.N - 1) return hi;
  int b = (int)(l / bs);
  const double* boff = at_peer(P.boff[p], d);
  double v = boff[b] + __ldcg(at_peer(P.local_cum[p], d) + l) * at_peer(P.bscale[p], d)[b];
  double cap = boff[b + 1];
  v = v < cap ? v : cap;
  v = P.rank_off[p][r] + v * P.rank_fac[p][r];
  return v < hi ? v : hi;
}

// global index of the parent of offspring with threshold uj (sharded population)
__device__ __forceinline__ long long parent_of_sharded(const Params& P, int p, double uj, double norm, int bs) {
  int r = 0;
  while (r + 1 < P.world && !(P.rank_off[p][r + 1] / norm > uj)) r++;
  long long lo = 0, hi = P.N;
  while (lo < hi) {
    long long mid = (lo + hi) >> 1;
    if (cum_at_sharded(P, p, r, mid, bs) / norm > uj) hi = mid; else lo = mid + 1;
  }
  if (lo >= P.N) lo = P.N - 1;
  return (long long)r * P.N + lo;
}
__device__ __forceinline__ double cum_at_global(const Params& P, int p, long long g, int bs) {
  return cum_at_sharded(P, p, (int)(g / P.N), g % P.N, bs);
}

__device__ __forceinline__ double u_at(const Ctrl* c, long long j) {
  int lo = 0, hi = c->nseg - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (c->seg[mid].j0 <= j) lo = mid; else hi = mid - 1;
  }
  return c->seg[lo].u0 + (double)(j - c->seg[lo].j0) * c->seg[lo].step;
}

// parent of offspring j: the first interval i with c_i / norm > u_j (particleContainer.cpp:474-504)
__device__ __forceinline__ long long parent_of(const Params& P, int p, double uj, double norm, int bs) {
  long long lo = 0, hi = P.N;   // count of i with c_i/norm <= uj
  if (!P.parity_scan && !P.flat_search) {
    // two levels: c is non-decreasing and capped by boff[b+1] inside block b, so the block that holds the answer is
    // the first b with boff[b+1]/norm > uj -- found in the small (cache-resident) offset array -- and only the search
    // inside that block touches the per-particle prefixes.  Same answer as the flat search.
    int bl = 0, bh = P.nblocks;
    while (bl < bh) {
      int mid = (bl + bh) >> 1;
      if (P.boff[p][mid + 1] / norm > uj) bh = mid; else bl = mid + 1;
    }
    if (bl >= P.nblocks) return P.N - 1;
    // everything before block bl fails the test and the first element after it passes (its c is >= boff[bl+1]), so the
    // flat search's answer lies in [start of bl, end of bl] -- the end itself only if rounding left the block's last
    // prefix a hair under boff[bl+1]
    lo = (long long)bl * bs;
    hi = lo + bs < P.N ? lo + bs : P.N;
    while (lo < hi) {
      long long mid = (lo + hi) >> 1;
      if (cum_at(P, p, mid, bs) / norm > uj) hi = mid; else lo = mid + 1;
    }
    return lo < P.N ? lo : P.N - 1;
  }
  while (lo < hi) {
    long long mid = (lo + hi) >> 1;
    if (cum_at(P, p, mid, bs) / norm > uj) hi = mid; else lo = mid + 1;
  }
  return lo < P.N ? lo : P.N - 1;
}

// The exact chain u_0 = U/N, u_{j+1} = fl(u_j + 1/N) (particleContainer.cpp:479,494) in closed form: inside a binade
// repeated addition of d advances by a constant multiple of the ulp (after at most one tie-resolving step), so the
// chain is piecewise an exact arithmetic progression.  One thread builds the <= 2*log2(N)+O(1) pieces.
__device__ void build_u_chain(Ctrl* c, double U, long long N) {
  const double d = 1.0 / (double)N;
  double u = U / (double)N;
  long long j = 0;
  int ns = 0;
  while (j < N && ns < MAXSEG - 2) {
    c->seg[ns].j0 = j; c->seg[ns].u0 = u; c->seg[ns].step = 0.0; ns++;    // element j on its own
    double u1 = u + d;
    if (j + 1 >= N) { j++; break; }
    int e0, e1, e2;
    frexp(u, &e0); frexp(u1, &e1);
    double u2 = u1 + d;
    frexp(u2, &e2);
    if (e0 != e1 || e1 != e2 || u <= 0.0) { j++; u = u1; continue; }
    double s = u2 - u1;                 // exact: same binade
    double ulp = ldexp(1.0, e1 - 53);   // u1 in [2^(e1-1), 2^e1)
    double top = ldexp(1.0, e1);
    // number of elements u1 + i*s (i >= 0) below the top of the binade; all quantities are integer multiples of ulp
    unsigned long long room = (unsigned long long)((top - u1) / ulp);
    unsigned long long si = (unsigned long long)(s / ulp);
    if (si == 0) { j++; u = u1; continue; }
    long long cnt = (long long)((room + si - 1) / si);
    if (cnt > N - (j + 1)) cnt = N - (j + 1);
    if (cnt < 1) { j++; u = u1; continue; }
    c->seg[ns].j0 = j + 1; c->seg[ns].u0 = u1; c->seg[ns].step = s; ns++;
    double ulast = u1 + (double)(cnt - 1) * s;
    j = j + 1 + cnt;
    u = ulast + d;
  }
  // (cannot happen for N < 2^52) fall back to explicit steps being cut short: extend the last piece
  c->nseg = ns;
}

// End of a row (one thread): book the ln normaliser, decide on resampling (particleContainer.cpp:281) and, if so,
// draw the systematic-resampling uniform and build the u-chain over all Ntot particles.
__device__ void commit_row(const Params& P, Ctrl* c, long long row, double new_lse, double ess, bool ok, double norm,
                           long long Ntot) {
  if (!ok || !isfinite(new_lse)) c->err = 1;
  c->lnZ += new_lse;
  c->lse = new_lse;
  const bool rs = ess < (double)Ntot * P.ess_fraction - 1e-6;
  if (c->resampled) c->cur ^= 1;
  c->resampled = rs ? 1 : 0;
  if (rs) {
    Rng rr;
    rng_init(rr, P.seed, 0, (uint32_t)row, P.chunk, RNG_RESAMPLE);
    double U = rng_uniform(rr);
    build_u_chain(c, U, Ntot);
    c->norm = norm;
    c->child_logw = log(norm / (double)Ntot);
    c->resample_pos = P.x_end[row];
    c->resample_row = row;
    if (c->n_resample < P.rs_cap) { P.rs_pos[c->n_resample] = P.x_end[row]; P.rs_ess[c->n_resample] = ess; }
    c->n_resample++;
  }
  c->ticket = 0;
  c->row = row + 1;
}

// second half of a sharded row, once the ranks' summaries are in: every rank derives the same ln normaliser, ESS, rank
// offsets and resampling decision from the same numbers in the same order
__device__ void shard_commit(const Params& P, Ctrl* c, long long row, const double* all) {
  const int pnow = (int)(row & 1), G = P.world;
  double gm = -DBL_MAX;
  for (int r = 0; r < G; r++) gm = fmax(gm, all[r * 4]);
  double S = 0.0, Q = 0.0;
  for (int r = 0; r < G; r++) {
    double f = exp(all[r * 4] - gm);
    S += all[r * 4 + 1] * f;
    Q += all[r * 4 + 2] * f * f;
  }
  const double new_lse = gm + log(S);
  const double ess = S * S / Q;
  double acc = 0.0;
  for (int r = 0; r < G; r++) {
    double fac = exp(all[r * 4] - new_lse);
    P.rank_fac[pnow][r] = fac;
    P.rank_off[pnow][r] = acc;
    acc += all[r * 4 + 3] * fac;
  }
  P.rank_off[pnow][G] = acc;
  commit_row(P, c, row, new_lse, ess, S > 0.0, acc, P.Ntot);
}

// ---------------------------------------------------------------------------------------------------------------
// the row kernel
// ---------------------------------------------------------------------------------------------------------------
// mode 0: one row; mode 1: materialise only (apply pending normalisation / resampling, no extension, no finalise)
template <int BS>
__device__ __forceinline__ TreeViewT<BS> tree_view_of(const Params& P, unsigned char* smem, int t, std::false_type) {
  return make_tree_view<BS>(smem, P.m.n, P.m.nn, P.m.E, t);
}
template <int BS>
__device__ __forceinline__ STreeViewT<BS> tree_view_of(const Params& P, unsigned char* smem, int t, std::true_type) {
  return make_stree_view<BS>(smem, P.m.n, P.m.nn, P.m.E, t, &P.sm);
}

// pointer to the frozen record (Q doubles) of closed bucket (e, slot s) that belongs to particle i's ancestor
template <bool SH>
__device__ __forceinline__ const double* frozen_record(const Params& P, const int* map, int e, int s, long long i, int Q) {
  const size_t es = (size_t)(e * P.B + s);
  const int m = map[es * P.N + i];
  const double* f = P.frozen;
  long long l = m;
  if constexpr (SH) { const int r = m / P.N; l = m - r * P.N; f = at_peer(f, P.peer_tally[r]); }
  return f + (es * P.N + l) * Q;
}

// SH: the population is sharded over ranks (smc_shard_attach); the single-GPU instantiation carries none of that code
#ifdef SMC_MINBLOCKS
#define SMC_ROW_BOUNDS __launch_bounds__(BS, SMC_MINBLOCKS)     // occupancy experiments
#else
#define SMC_ROW_BOUNDS __launch_bounds__(BS)
#endif
template <int BS, bool ST, bool SH>
__global__ void SMC_ROW_BOUNDS
smc_row_kernel(const __grid_constant__ Params P, int mode) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int t = threadIdx.x;
  // The model tables are read with a per-lane epoch index all through the row.  In the parameter bank such a load is
  // replayed once per distinct address in the warp (constant-cache broadcast); a copy in shared memory is not.
  __shared__ ModelDev s_model;
  {
    static_assert(sizeof(ModelDev) % 8 == 0, "ModelDev is copied as 8-byte words");
    const unsigned long long* srcw = reinterpret_cast<const unsigned long long*>(&P.m);
    unsigned long long* dstw = reinterpret_cast<unsigned long long*>(&s_model);
    for (int w = t; w < (int)(sizeof(ModelDev) / 8); w += BS) dstw[w] = srcw[w];
  }
  __syncthreads();
  const ModelDev& m = s_model;
  const int n = m.n, nn = m.nn, E = m.E;
  auto T = tree_view_of<BS>(P, smem_raw, t, std::integral_constant<bool, ST>());
  const int QC = ST ? P.sm.Q : 4;      // tally cells per epoch
  __shared__ double red[4][4];
  __shared__ bool is_last;

  Ctrl* c = P.ctrl;
  const long long row = c->row;
  if (mode == 0 && row >= P.R) return;
#ifdef SMC_CLOCKS
  long long clk = clock64();
  if (threadIdx.x == 0) atomicAdd(&g_clk[10], 1ull);
#endif
  const int cur = c->cur;
  const int resampled = c->resampled;
  const double lse = c->lse;
  const int pprev = (int)((row + 1) & 1);     // parity of the previous row's partials
  const int pnow = (int)(row & 1);
  const long long i = (long long)blockIdx.x * BS + t;
  const bool live = i < P.N;
  const StateBuf& src = P.buf[cur];
  const StateBuf& dst = P.buf[resampled ? cur ^ 1 : cur];
  constexpr int nwarp = BS >> 5;
  const int warp = t >> 5, lane = t & 31;

  // ---- phase 0: tallies due after the previous row, weighted with the previous row's normalised weights ----------
  // A materialising getter AFTER the last row (mode 1, row == R) leaves what the last row owes to smc_finish, whose flush
  // starts from the buckets consumed after rows 0..R-2; consuming them here as well would count them twice.  The pull
  // below then has to carry the ancestor slots of those owed buckets too: it lists pending buckets as of row R-1.
  const bool tail = mode == 1 && row >= P.R;
  int c0 = 0, c1 = 0;
  if (row > 0 && c->cons_done != row && !tail) { c0 = P.cons_off[row - 1]; c1 = P.cons_off[row]; }
  // Second moments (the *_weight tallies behind the ESS column of .out, count.cpp:519-526): the reference squares the
  // weight POOLED over the particles that share a record.  Particles share a closed bucket's record exactly when their
  // map entries are equal, so the launch that consumes a bucket pools the weights per ancestor slot (pool_add) and the
  // NEXT launch -- when every block's contribution has landed -- lets the owner of each slot add pool^2 * opportunity
  // to its own running sums (pool2: plain read-modify-writes, no shuffles or atomics; a warp reduction and an atomic per
  // bucket cost 5 % of the sweep) and clear the pool; smc_finish adds the running sums into the tallies.
  // Two pools by parity, the ring keeps a consumed bucket one launch longer (smc_upload_rows).
  // Not pooled: what siblings born while the bucket was still open share of it (they carry it by value), the bucket
  // open at the end of the data, and sharded populations (per-particle w^2, a lower bound).
  const int pc0 = c->pool_c0, pc1 = c->pool_c1, ppar = c->pool_par;
  const bool pooling = !SH && P.pool_on;
  if (pooling && pc1 > pc0) {
    const int QW = ST ? QC : 2;
    for (int q0 = pc0; q0 < pc1; q0 += 4) {
      // the pool values of four buckets are loaded together (each wait would otherwise be a round trip to L2)
      double pw4[4];
#pragma unroll
      for (int j = 0; j < 4; j++)
        pw4[j] = (live && q0 + j < pc1) ? P.pool[((size_t)ppar * E + P.cons[q0 + j].x) * P.N + i] : 0.0;
      if constexpr (ST) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const double pw = pw4[j];
          if (pw == 0.0) continue;         // few slots have descendants left by the time a bucket falls due
          const int2 ek = P.cons[q0 + j];
          P.pool[((size_t)ppar * E + ek.x) * P.N + i] = 0.0;
          const double* fr = P.frozen + ((size_t)(ek.x * P.B + ek.y % P.B) * P.N + i) * QC;
          double* s2 = P.pool2 + (size_t)ek.x * QW * P.N + i;
          for (int a = 0; a < QW; a++) s2[(size_t)a * P.N] += pw * pw * fr[a];
        }
      } else {
        // one population: cells 1 and 3 (the opportunities).  All loads of the four buckets first, then the stores: done
        // one bucket after the other, each read-modify-write waited out its own round trip (4 % of the sweep)
        double f1[4], f3[4], a1[4], a3[4];
        double* s2[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          f1[j] = f3[j] = a1[j] = a3[j] = 0.0;
          s2[j] = nullptr;
          if (pw4[j] != 0.0) {
            const int2 ek = P.cons[q0 + j];
            const double2* fr = reinterpret_cast<const double2*>(P.frozen + ((size_t)(ek.x * P.B + ek.y % P.B) * P.N + i) * 4);
            s2[j] = P.pool2 + (size_t)ek.x * 2 * P.N + i;
            f1[j] = fr[0].y; f3[j] = fr[1].y;
            a1[j] = s2[j][0]; a3[j] = s2[j][P.N];
            P.pool[((size_t)ppar * E + ek.x) * P.N + i] = 0.0;
          }
        }
#pragma unroll
        for (int j = 0; j < 4; j++)
          if (pw4[j] != 0.0) {
            s2[j][0] = a1[j] + pw4[j] * pw4[j] * f1[j];
            s2[j][P.N] = a3[j] + pw4[j] * pw4[j] * f3[j];
          }
      }
    }
  }
  if (c1 > c0) {
    // at most one bucket per epoch falls due per row; the per-warp sums of all of them are staged in the (still
    // unused) tree area and handed to the global tallies with one pass
    double* scr = reinterpret_cast<double*>(smem_raw);           // [c1 - c0][6][nwarp]  (structured: [..][2Q][nwarp])
    const double w = live ? dexp(src.logw[i] - lse) : 0.0;
    const double w2 = pooling ? 0.0 : w * w;
    double* pl = P.pool + (size_t)(ppar ^ 1) * E * P.N;
    if (pooling && ST) {
      for (int q = c0; q < c1; q++) {
        const int2 ek = P.cons[q];
        const int key = live ? src.map[(size_t)(ek.x * P.B + ek.y % P.B) * P.N + i] : -1;
        pool_add(pl + (size_t)ek.x * P.N, key, w);
      }
    }
    if constexpr (ST) {
      // every cell is summed with w and with w^2: rows [0, QC) and [QC, 2Q) of the tallies
      for (int q = c0; q < c1; q++) {
        int2 ek = P.cons[q];
        const double* fr = live ? frozen_record<SH>(P, src.map, ek.x, ek.y % P.B, i, QC) : nullptr;
        for (int a = 0; a < QC; a++) {
          const double f = live ? fr[a] : 0.0;
          const double v0 = warp_sum(w * f), v1 = warp_sum(w2 * f);
          if (lane == 0) {
            scr[((q - c0) * 2 * QC + a) * nwarp + warp] = v0;
            scr[((q - c0) * 2 * QC + QC + a) * nwarp + warp] = v1;
          }
        }
      }
      __syncthreads();
      for (int idx = t; idx < (c1 - c0) * 2 * QC; idx += BS) {
        double s = 0;
        for (int w_ = 0; w_ < nwarp; w_++) s += scr[idx * nwarp + w_];
        if (s != 0.0) atomicAdd(P.tallies + (idx % (2 * QC)) * E + P.cons[c0 + idx / (2 * QC)].x, s);
      }
    } else {
      for (int q = c0; q < c1; q++) {
        int2 ek = P.cons[q];
        double v[6] = {0, 0, 0, 0, 0, 0};
        int key = -1;
        if (live) {
          double f[4];
          key = load_frozen<SH>(P, src.map, ek.x, ek.y % P.B, i, f);
          v[0] = w * f[0]; v[1] = w * f[1]; v[2] = w2 * f[1]; v[3] = w * f[2]; v[4] = w * f[3]; v[5] = w2 * f[3];
        }
        if (pooling) pool_add(pl + (size_t)ek.x * P.N, key, w);
#pragma unroll
        for (int a = 0; a < 6; a++) {
          v[a] = warp_sum(v[a]);
          if (lane == 0) scr[((q - c0) * 6 + a) * nwarp + warp] = v[a];
        }
      }
      __syncthreads();
      for (int idx = t; idx < (c1 - c0) * 6; idx += BS) {
        double s = 0;
        for (int w_ = 0; w_ < nwarp; w_++) s += scr[idx * nwarp + w_];
        if (s != 0.0) atomicAdd(P.tallies + (idx % 6) * E + P.cons[c0 + idx / 6].x, s);
      }
    }
    __syncthreads();
  }

  SMC_CLK(0, clk);
  // ---- phase 1: load own state, or pull the parent's after a resampling ---------------------------------------
  // After a resampling every particle copies its parent's ancestor slot of every pending closed bucket (~150 of them
  // at the bench shape).  The (epoch, ring slot) pairs are the same for all particles of the launch: they are listed
  // once per block in shared memory (the per-row accumulators of phase 2 are not in use yet), so that the copy below is
  // a flat loop with eight independent gathers in flight instead of one dependent load -> store per bucket.
  int npend = -1;                                             // -1: no list (not resampled, or it does not fit)
  int* plist = reinterpret_cast<int*>(T.ACC);
  if (resampled && (size_t)E * P.B * sizeof(int) <= (size_t)E * BS * sizeof(double)) {
    __shared__ int pl_off[SMC_MAX_EPOCHS + 1];
    const long long rowc = row < P.R ? row : P.R;
    const int* ks = P.kstart + rowc * E;
    const int* ksl = P.kslot + rowc * E;
    const int* kf = P.kfirst + (tail && rowc > 0 ? rowc - 1 : rowc) * E;
    if (t == 0) {
      int acc = 0;
      for (int e = 0; e < E; e++) { pl_off[e] = acc; int cnt = ks[e] - kf[e]; acc += cnt > 0 ? cnt : 0; }
      pl_off[E] = acc;
    }
    __syncthreads();
    for (int e = t; e < E; e += BS) {
      int s = ksl[e], o = pl_off[e];
      for (int k = ks[e] - 1; k >= kf[e]; k--) {
        if (--s < 0) s = P.B - 1;
        plist[o++] = e * P.B + s;
      }
    }
    __syncthreads();
    npend = pl_off[E];
  }
  double logw = 0.0, next_base = 0.0;
  long long a = i;
  long long dS = 0, dR = 0;      // byte offsets to the parent's owner (0: this rank)
  bool redraw = false;
  if (live) {
    if (resampled) {
      const double norm = c->norm;
      // did the parent get more than one offspring?  (then every copy re-draws its next recombination position,
      // particleContainer.cpp:357-367 / forest.h:140-151)
      if constexpr (!SH) {
        double uj = u_at(c, i);
        a = parent_of(P, pprev, uj, norm, BS);
        if (i > 0) {
          double up = u_at(c, i - 1);
          if (a == 0 || !(cum_at(P, pprev, a - 1, BS) / norm > up)) redraw = true;
        }
        if (!redraw && i + 1 < P.N) {
          double un = u_at(c, i + 1);
          if (a == P.N - 1 || cum_at(P, pprev, a, BS) / norm > un) redraw = true;
        }
        P.anc[i] = (int)a;
      } else {
        const long long gi = P.base + i;
        double uj = u_at(c, gi);
        const long long ga = parent_of_sharded(P, pprev, uj, norm, BS);
        if (gi > 0) {
          double up = u_at(c, gi - 1);
          if (ga == 0 || !(cum_at_global(P, pprev, ga - 1, BS) / norm > up)) redraw = true;
        }
        if (!redraw && gi + 1 < P.Ntot) {
          double un = u_at(c, gi + 1);
          if (ga == P.Ntot - 1 || cum_at_global(P, pprev, ga, BS) / norm > un) redraw = true;
        }
        P.anc[i] = (int)ga;
        const int ra = (int)(ga / P.N);
        a = ga % P.N;
        dS = P.peer_state[ra];
        dR = P.peer_tally[ra];
      }
      logw = c->child_logw;
    } else {
      logw = src.logw[i] - lse;
    }
    next_base = at_peer(src.next_base, dS)[a];
    {
      const double* __restrict__ sh_ = at_peer(src.height, dS);
      const signed char* __restrict__ sp_ = at_peer(src.parent, dS);
      for (int v = n; v < nn; v++) T.hw(v) = sh_[(size_t)(v - n) * P.N + a];
      for (int v = 0; v < nn; v++) T.par(v) = sp_[(size_t)v * P.N + a];
      if constexpr (ST) {
        // populations and marks sit in the tally slab (smc_upload_rows): a peer's copy is reached through peer_tally
        const signed char* __restrict__ spop = at_peer(src.pop, dR);
        for (int v = 0; v < nn; v++) T.pop(v) = spop[(size_t)v * P.N + a];
        const int nm = at_peer(src.nmarks, dR)[a];
        T.nm() = (signed char)nm;
        T.FLAG[t] = 0;
        const double* __restrict__ smt = at_peer(src.mtime, dR);
        const signed char* __restrict__ smb = at_peer(src.mbranch, dR);
        const signed char* __restrict__ smp = at_peer(src.mpop, dR);
        for (int k = 0; k < nm; k++) {
          T.mt(k) = smt[(size_t)k * P.N + a];
          T.mb(k) = smb[(size_t)k * P.N + a];
          T.mp(k) = smp[(size_t)k * P.N + a];
        }
      }
    }
    // tallies: the open cells travel by value, closed buckets by reference (4-byte slot ids); a bucket that closes
    // during this row is frozen now -- its carried value moves to this particle's own frozen cell and the buckets born
    // inside the row start from zero there -- so that phase 2 only ever adds to cells that exist
    {
      const long long rowc = row < P.R ? row : P.R;
      const int* ks = P.kstart + rowc * E;
      const int* ksl = P.kslot + rowc * E;
      const int* kf = P.kfirst + (tail && rowc > 0 ? rowc - 1 : rowc) * E;
      const unsigned long long closes = (mode == 0) ? P.closes[row] : 0ull;
      const double* __restrict__ osrc = at_peer(src.open, dR);
      const int* __restrict__ msrc = at_peer(src.map, dR);
      const int self = (int)(P.base + i);
      const size_t NN = (size_t)P.N;
      if constexpr (ST) {
        for (int e = 0; e < E; e++) {
          const bool closing = (closes >> e) & 1ull;
          if (resampled || closing) {
            const double2* o2 = reinterpret_cast<const double2*>(osrc + ((size_t)e * NN + a) * QC);
            double2* d2 = reinterpret_cast<double2*>(dst.open + ((size_t)e * NN + i) * QC);
            const int QC2 = QC >> 1;                       // QC = 2 + P(P+3) is even
            const double2 zero = make_double2(0.0, 0.0);
            if (closing) {
              int s = ksl[e];
              double2* f2 = reinterpret_cast<double2*>(P.frozen + ((size_t)(e * P.B + s) * NN + i) * QC);
              for (int c2 = 0; c2 < QC2; c2++) { f2[c2] = o2[c2]; }
              dst.map[(size_t)(e * P.B + s) * NN + i] = self;
              const int kn = P.kstart[(row + 1) * E + e];
              for (int k = ks[e] + 1; k < kn; k++) {
                if (++s >= P.B) s = 0;
                double2* z2 = reinterpret_cast<double2*>(P.frozen + ((size_t)(e * P.B + s) * NN + i) * QC);
                for (int c2 = 0; c2 < QC2; c2++) z2[c2] = zero;
                dst.map[(size_t)(e * P.B + s) * NN + i] = self;
              }
              for (int c2 = 0; c2 < QC2; c2++) d2[c2] = zero;
            } else {
              for (int c2 = 0; c2 < QC2; c2++) d2[c2] = o2[c2];
            }
          }
        }
      } else {
        // four epochs at a time: the loads of a batch are all in flight before the first store
        for (int e0 = 0; e0 < E; e0 += 4) {
          double2 o0[4], o1[4];
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const int e = e0 + j;
            if (e < E && (resampled || ((closes >> e) & 1ull))) {
              const double2* o2 = reinterpret_cast<const double2*>(osrc + ((size_t)e * NN + a) * 4);
              o0[j] = o2[0]; o1[j] = o2[1];
            }
          }
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const int e = e0 + j;
            if (e >= E) continue;
            const bool closing = (closes >> e) & 1ull;
            if (!(resampled || closing)) continue;
            double2* d2 = reinterpret_cast<double2*>(dst.open + ((size_t)e * NN + i) * 4);
            if (closing) {
              int s = ksl[e];
              double2* f2 = reinterpret_cast<double2*>(P.frozen + ((size_t)(e * P.B + s) * NN + i) * 4);
              f2[0] = o0[j]; f2[1] = o1[j];
              dst.map[(size_t)(e * P.B + s) * NN + i] = self;
              const int kn = P.kstart[(row + 1) * E + e];
              for (int k = ks[e] + 1; k < kn; k++) {
                if (++s >= P.B) s = 0;
                double2* z2 = reinterpret_cast<double2*>(P.frozen + ((size_t)(e * P.B + s) * NN + i) * 4);
                z2[0] = make_double2(0.0, 0.0); z2[1] = make_double2(0.0, 0.0);
                dst.map[(size_t)(e * P.B + s) * NN + i] = self;
              }
              o0[j] = make_double2(0.0, 0.0); o1[j] = o0[j];
            }
            d2[0] = o0[j]; d2[1] = o1[j];
          }
        }
      }
      if (resampled && npend < 0) {
        for (int e = 0; e < E; e++) {
          int s = ksl[e];
          for (int k = ks[e] - 1; k >= kf[e]; k--) {
            if (--s < 0) s = P.B - 1;
            const size_t at = (size_t)(e * P.B + s) * NN;
            dst.map[at + i] = msrc[at + a];
          }
        }
      }
      if (npend > 0) {
        const int* __restrict__ ms = msrc + a;
        int* __restrict__ md = dst.map + i;
        for (int q0 = 0; q0 < npend; q0 += 8) {
          int val[8];
#pragma unroll
          for (int j = 0; j < 8; j++) val[j] = q0 + j < npend ? ms[(size_t)plist[q0 + j] * NN] : 0;
#pragma unroll
          for (int j = 0; j < 8; j++) if (q0 + j < npend) md[(size_t)plist[q0 + j] * NN] = val[j];
        }
      }
    }
    if (redraw) {
      Rng rr;
      rng_init(rr, P.seed, P.slot_offset + (uint32_t)i, (uint32_t)c->resample_row, P.chunk, RNG_REDRAW);
      next_base = draw_next_base(m, tree_length(m, T), rr, c->resample_pos);
    }
  }

  // the row's tally schedule is the same for every particle: one copy per block in shared memory
  __shared__ int s_ks[SMC_MAX_EPOCHS], s_ksl[SMC_MAX_EPOCHS], s_kn[SMC_MAX_EPOCHS];
  __shared__ double s_ubf[SMC_MAX_EPOCHS], s_ubf2[SMC_MAX_EPOCHS];
  if (mode == 0) {
    for (int e = t; e < E; e += BS) {
      const int k0 = P.kstart[row * E + e];
      s_ks[e] = k0; s_ksl[e] = P.kslot[row * E + e]; s_kn[e] = P.kstart[(row + 1) * E + e];
      s_ubf[e] = P.ub[P.ub_off[e] + k0];
      s_ubf2[e] = s_ubf[e] < DBL_MAX ? P.ub[P.ub_off[e] + k0 + 1] : s_ubf[e];      // the last bound of an epoch is +inf
    }
  }
  __syncthreads();   // also: the slot list shares its storage with the accumulators phase 2 is about to zero

  SMC_CLK(1, clk);
  // ---- phase 2: the row (every thread takes part: the moves are compacted over the block) ------------------------
  int nrec = 0;
  double row_area = 0.0;
  if (mode == 0) {
    __shared__ int s_wcnt[2 * (BS >> 5)];
    const double x0 = row > 0 ? P.x_end[row - 1] : P.start_pos;
    const double x1 = P.x_end[row];
    signed char data[32];
    for (int k = 0; k < n; k++) data[k] = P.alleles[row * n + k];
    Rng r;
    rng_init(r, P.seed, P.slot_offset + (uint32_t)i, (uint32_t)row, P.chunk, RNG_EXTEND);
    int nmiss_row = 0;
    for (int k = 0; k < n; k++) nmiss_row += data[k] == -1;
    // ACC2 shares the pruning scratch: free during the row unless branch_length_with_data runs (partly missing data)
    RingSink<BS, ST> sink{&P, dst.open, (int)i, s_ks, s_ksl, s_kn, T.ACC, t, s_ubf, 0ull, s_ubf2,
                          (nmiss_row == 0 || nmiss_row == n) ? T.ACC2 : nullptr};
    const size_t tree_bytes = ST ? stree_bytes_per_thread(nn, E, P.sm.K) : tree_bytes_per_thread(nn, E);
    const MoveSmem ms = make_move_smem<BS>(smem_raw + (((size_t)BS * tree_bytes + 15) & ~(size_t)15));
    nrec = extend_row_block<BS>(m, T, r, x0, x1, data, P.is_site[row], P.max_rec[row], sink, logw, next_base, live, ms,
                                s_wcnt, row_area);
  }

  SMC_CLK(2, clk);      // phase 2 as a whole (slots 3-5 split it)
  // ---- phase 3: store, weight partials, last block finalises ----------------------------------------------------
  if (live) {
    dst.logw[i] = logw;
    dst.next_base[i] = next_base;
    if (resampled || mode == 0) {
      for (int v = n; v < nn; v++) dst.height[(size_t)(v - n) * P.N + i] = T.h(v);
      for (int v = 0; v < nn; v++) dst.parent[(size_t)v * P.N + i] = T.par(v);
      if constexpr (ST) {
        for (int v = 0; v < nn; v++) dst.pop[(size_t)v * P.N + i] = T.pop(v);
        const int nm = T.nm();
        dst.nmarks[i] = (signed char)nm;
        for (int k = 0; k < nm; k++) {
          dst.mtime[(size_t)k * P.N + i] = T.mt(k);
          dst.mbranch[(size_t)k * P.N + i] = T.mb(k);
          dst.mpop[(size_t)k * P.N + i] = T.mp(k);
        }
        if (T.FLAG[t]) atomicOr(&c->sflags, (int)T.FLAG[t]);
      }
    }
  }
  if (mode != 0) return;

  // block max, then inclusive scan of exp(logw - max) and the sum of squares
  double mx = warp_max(live ? logw : -DBL_MAX);
  if (lane == 0) red[0][warp] = mx;
  __syncthreads();
  mx = red[0][0];
  for (int w = 1; w < nwarp; w++) mx = fmax(mx, red[0][w]);
  double ex = live ? dexp(logw - mx) : 0.0;
  double incl = ex;
  for (int o = 1; o < 32; o <<= 1) {
    double y = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += y;
  }
  double sq = warp_sum(ex * ex);
  double ar = warp_sum(live ? ex * row_area : 0.0);
  __syncthreads();
  if (lane == 31) red[1][warp] = incl;
  if (lane == 0) { red[2][warp] = sq; red[3][warp] = ar; }
  __syncthreads();
  double woff = 0.0, btot = 0.0, bsq = 0.0, barea = 0.0;
  for (int w = 0; w < nwarp; w++) {
    if (w < warp) woff += red[1][w];
    btot += red[1][w];
    bsq += red[2][w];
    barea += red[3][w];
  }
  if (live) P.local_cum[pnow][i] = woff + incl;
  int nr = warp_sum((double)nrec) + 0.5;
  if (lane == 0 && nr > 0) atomicAdd((unsigned long long*)&c->n_recomb, (unsigned long long)nr);
  if (t == 0) {
    P.bmax[pnow][blockIdx.x] = mx;
    P.bsum[pnow][blockIdx.x] = btot;
    P.bsq[pnow][blockIdx.x] = bsq;
    P.barea[pnow][blockIdx.x] = barea;
  }
  __threadfence();
  __syncthreads();
  SMC_CLK(6, clk);      // store + partials
  if (t == 0) {
    unsigned int tk = atomicAdd(&c->ticket, 1u);
    is_last = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();

  // ---- finalise (one block): ln normaliser, ESS, block offsets, resampling decision -----------------------------
  const int nb = gridDim.x;
  double* sh = reinterpret_cast<double*>(smem_raw);   // tree staging is dead now: reuse it
  double gm = -DBL_MAX;
  for (int b = t; b < nb; b += BS) gm = fmax(gm, __ldcg(P.bmax[pnow] + b));
  gm = warp_max(gm);
  if (lane == 0) red[0][warp] = gm;
  __syncthreads();
  gm = red[0][0];
  for (int w = 1; w < nwarp; w++) gm = fmax(gm, red[0][w]);
  double S = 0.0, Q = 0.0, A = 0.0;
  for (int b = t; b < nb; b += BS) {
    double f = dexp(__ldcg(P.bmax[pnow] + b) - gm);
    S += __ldcg(P.bsum[pnow] + b) * f;
    Q += __ldcg(P.bsq[pnow] + b) * f * f;
    A += __ldcg(P.barea[pnow] + b) * f;
  }
  S = warp_sum(S); Q = warp_sum(Q); A = warp_sum(A);
  __syncthreads();
  if (lane == 0) { red[1][warp] = S; red[2][warp] = Q; red[3][warp] = A; }
  __syncthreads();
  S = 0.0; Q = 0.0; A = 0.0;
  for (int w = 0; w < nwarp; w++) { S += red[1][w]; Q += red[2][w]; A += red[3][w]; }
  constexpr bool sharded = SH;
  const double new_lse = gm + log(S);
  const double ess = S * S / Q;
  // a sharded rank scales its partials by its own maximum only: peers read them before the global normaliser is known
  const double ref = sharded ? gm : new_lse;
  // exclusive prefix over blocks of the normalised block totals: each thread owns a contiguous run of blocks
  const int per = (nb + BS - 1) / BS;
  const int b0 = t * per, b1 = min(nb, b0 + per);
  double run = 0.0;
  for (int b = b0; b < b1; b++) {
    double sc = dexp(__ldcg(P.bmax[pnow] + b) - ref);
    P.bscale[pnow][b] = sc;
    run += __ldcg(P.bsum[pnow] + b) * sc;
  }
  sh[t] = run;
  __syncthreads();
  if (t == 0) {
    double accu = 0.0;
    for (int k = 0; k < BS; k++) { double v = sh[k]; sh[k] = accu; accu += v; }
    sh[BS] = accu;
  }
  __syncthreads();
  run = sh[t];
  for (int b = b0; b < b1; b++) {
    P.boff[pnow][b] = run;
    run += __ldcg(P.bsum[pnow] + b) * P.bscale[pnow][b];
  }
  if (t == 0) P.boff[pnow][nb] = sh[BS];
  __syncthreads();
  if (t != 0) return;
  if constexpr (sharded) {
    // The ranks exchange (max, sum, sum of squares, prefix total) inside this launch: every rank stores its four numbers
    // and then a flag into slot [row parity][own rank] of EVERY rank's mailbox (peer stores over NVLink; the mailbox
    // sits in the IPC-mapped state slab), waits until all slots of its own mailbox carry this row's flag, and finishes
    // the row from the same numbers in the same order as everybody else.  One launch per row, graph-capturable like
    // the unsharded sweep; it replaces launch -> ncclAllGather(4 doubles) -> 1-thread commit launch.  The fence before
    // the flag is also what orders this rank's state, written by all its blocks before their tickets, against the
    // pulls of its peers' next launch; two slots by row parity suffice because no rank can reach row
    // k + 2 before every rank has consumed row k.
    const int G = P.world, par = (int)(row & 1);
    const long long flag = (c->gen << 40) | (row + 1);
    const size_t mine = (size_t)(par * SMC_MAX_RANKS + P.rank) * 8;
    for (int r = 0; r < G; r++) {
      double* mb = at_peer(P.mbox, P.peer_state[r]) + mine;
      mb[0] = gm; mb[1] = S; mb[2] = Q; mb[3] = sh[BS]; mb[5] = A;
    }
    __threadfence_system();
    for (int r = 0; r < G; r++)
      *reinterpret_cast<volatile long long*>(at_peer(P.mbox, P.peer_state[r]) + mine + 4) = flag;
    double all[SMC_MAX_RANKS * 4];
    for (int r = 0; r < G; r++) {
      const volatile long long* f = reinterpret_cast<const volatile long long*>(P.mbox + (size_t)(par * SMC_MAX_RANKS + r) * 8 + 4);
      while (*f != flag) __nanosleep(64);
    }
    __threadfence_system();
    for (int r = 0; r < G; r++) {
      const volatile double* v = P.mbox + (size_t)(par * SMC_MAX_RANKS + r) * 8;
      for (int k = 0; k < 4; k++) all[r * 4 + k] = v[k];
    }
    {   // the row's posterior-mean recombination opportunity over all ranks (weights relative to the global maximum)
      double gmax = -DBL_MAX, sa = 0.0, ss = 0.0;
      for (int r = 0; r < G; r++) gmax = fmax(gmax, all[r * 4]);
      for (int r = 0; r < G; r++) {
        const double f = exp(all[r * 4] - gmax);
        ss += all[r * 4 + 1] * f;
        sa += reinterpret_cast<const volatile double*>(P.mbox + (size_t)(par * SMC_MAX_RANKS + r) * 8)[5] * f;
      }
      P.row_opp[row] = sa / ss;
    }
    shard_commit(P, c, row, all);
    return;
  }
  P.row_opp[row] = A / S;
  double norm = sh[BS];
  if (P.parity_scan) {
    // order-faithful partial sums (particleContainer.cpp:256-262): left to right, one thread
    double acc = 0.0;
    for (long long k = 0; k < P.N; k++) { P.psum[k] = acc; acc += exp(__ldcg(dst.logw + k) - new_lse); }
    P.psum[P.N] = acc;
    norm = acc;
  }
  if (pooling) {        // what this launch pooled waits for the next one; what it found pooled is booked
    c->pool_c0 = c0; c->pool_c1 = c1;
    if (c1 > c0) c->pool_par = ppar ^ 1;
  }
  commit_row(P, c, row, new_lse, ess, S > 0.0, norm, (long long)P.N);
}

// commit of a materialise-only pass (its phase 0 did what a row launch's does: see the pool hand-over there)
__global__ void smc_commit_kernel(const __grid_constant__ Params P, Ctrl* c) {
  if (P.world == 1 && P.pool_on) {
    const long long row = c->row;
    int c0 = 0, c1 = 0;
    if (row > 0 && row < P.R && c->cons_done != row) { c0 = P.cons_off[row - 1]; c1 = P.cons_off[row]; }
    c->pool_c0 = c0; c->pool_c1 = c1;
    if (c1 > c0) c->pool_par ^= 1;
  }
  c->cons_done = c->row;
  if (c->resampled) c->cur ^= 1;
  c->resampled = 0;
  c->lse = 0.0;
}

// end of data: every pending bucket is weighted with the final normalised weights (smcsmc.cpp:371-373, lag 0)
__global__ void __launch_bounds__(128) smc_flush_kernel(const __grid_constant__ Params P, const int* first_pending) {
  __shared__ double tal[6][4];
  Ctrl* c = P.ctrl;
  const int BS = blockDim.x, t = threadIdx.x, warp = t >> 5, lane = t & 31, nwarp = BS >> 5;
  const long long i = (long long)blockIdx.x * BS + t;
  const bool live = i < P.N;
  const StateBuf& src = P.buf[c->cur];
  const int E = P.m.E;
  double w = live ? exp(src.logw[i] - c->lse) : 0.0;
  for (int e = 0; e < E; e++) {
    int nbk = P.ub_off[e + 1] - P.ub_off[e];
    double v[6] = {0, 0, 0, 0, 0, 0};
    if (live) {
      // closed buckets the schedule never reached, through the ancestor maps; then the bucket still open (+inf bound)
      for (int k = first_pending[e]; k < nbk; k++) {
        double f[4];
        if (k < nbk - 1) {
          if (P.world > 1) load_frozen<true>(P, src.map, e, k % P.B, i, f); else load_frozen<false>(P, src.map, e, k % P.B, i, f);
        } else {
          const double2* o2 = reinterpret_cast<const double2*>(src.open + ((size_t)e * P.N + i) * 4);
          double2 a = o2[0], b = o2[1];
          f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y;
        }
        // second moments of closed buckets are pooled per ancestor slot by smc_flush_pool_kernel (one population of one GPU)
        const double w2 = (P.world > 1 || !P.pool_on || k == nbk - 1) ? w * w : 0.0;
        v[0] += w * f[0]; v[1] += w * f[1]; v[2] += w2 * f[1]; v[3] += w * f[2]; v[4] += w * f[3]; v[5] += w2 * f[3];
      }
    }
    for (int a = 0; a < 6; a++) { v[a] = warp_sum(v[a]); if (lane == 0) tal[a][warp] = v[a]; }
    __syncthreads();
    if (t < 6) {
      double s = 0;
      for (int ww = 0; ww < nwarp; ww++) s += tal[t][ww];
      if (s != 0.0) atomicAdd(P.tallies + t * E + e, s);
    }
    __syncthreads();
  }
}

// structured populations: the end-of-data flush over all Q cells, with w and with w^2
__global__ void __launch_bounds__(128) smc_flush_struct_kernel(const __grid_constant__ Params P, const int* first_pending) {
  __shared__ double tal[2][4];
  Ctrl* c = P.ctrl;
  const int BS = blockDim.x, t = threadIdx.x, warp = t >> 5, lane = t & 31, nwarp = BS >> 5;
  const long long i = (long long)blockIdx.x * BS + t;
  const bool live = i < P.N;
  const StateBuf& src = P.buf[c->cur];
  const int E = P.m.E, Q = P.sm.Q;
  const double w = live ? exp(src.logw[i] - c->lse) : 0.0;
  for (int e = 0; e < E; e++) {
    const int nbk = P.ub_off[e + 1] - P.ub_off[e];
    for (int a = 0; a < Q; a++) {
      double f = 0.0, fo = 0.0;
      if (live) {
        for (int k = first_pending[e]; k < nbk - 1; k++)
          f += P.world > 1 ? frozen_record<true>(P, src.map, e, k % P.B, i, Q)[a]
                           : frozen_record<false>(P, src.map, e, k % P.B, i, Q)[a];
        fo = src.open[((size_t)e * P.N + i) * Q + a];
      }
      // second moments of closed buckets are pooled per ancestor slot by smc_flush_pool_kernel (one GPU)
      const double v0 = warp_sum(w * (f + fo)), v1 = warp_sum(w * w * ((P.world > 1 || !P.pool_on) ? f + fo : fo));
      if (lane == 0) { tal[0][warp] = v0; tal[1][warp] = v1; }
      __syncthreads();
      if (t < 2) {
        double s = 0;
        for (int ww = 0; ww < nwarp; ww++) s += tal[t][ww];
        if (s != 0.0) atomicAdd(P.tallies + ((size_t)t * Q + a) * E + e, s);
      }
      __syncthreads();
    }
  }
}

// End of data, second moments of the closed buckets (see phase 0 of the row kernel): pass 0 books what the last row launch
// left pooled; pass 1 pools the final weights per ancestor slot for bucket first_pending[e] + j of every epoch, pass 2
// books it (owner of the slot: pool^2 * cell) and clears the pool.  One launch per pass: the launch boundary is the
// grid-wide barrier between pooling and booking.
__global__ void __launch_bounds__(128) smc_flush_pool_kernel(const __grid_constant__ Params P, const int* first_pending, int j,
                                                             int pass) {
  Ctrl* c = P.ctrl;
  const int BS = blockDim.x, t = threadIdx.x;
  const long long i = (long long)blockIdx.x * BS + t;
  const bool live = i < P.N;
  const StateBuf& src = P.buf[c->cur];
  const int E = P.m.E;
  const bool st = P.sm.P > 1;
  const int Q = st ? P.sm.Q : 4, QW = st ? Q : 2;
  const int items = pass == 0 ? c->pool_c1 - c->pool_c0 : E;
  const double* pool = P.pool + (pass == 0 ? (size_t)c->pool_par * E * P.N : 0);
  const double w = live && pass == 1 ? exp(src.logw[i] - c->lse) : 0.0;
  for (int it = 0; it < items; it++) {
    int e, k;
    if (pass == 0) { const int2 ek = P.cons[c->pool_c0 + it]; e = ek.x; k = ek.y; }
    else { e = it; k = first_pending[e] + j; if (k >= P.ub_off[e + 1] - P.ub_off[e] - 1) continue; }
    const size_t es = (size_t)(e * P.B + k % P.B);
    double* pp = const_cast<double*>(pool) + (size_t)e * P.N;
    if (pass == 1) {
      pool_add(pp, live ? src.map[es * P.N + i] : -1, w);
      continue;
    }
    const double pw = live ? pp[i] : 0.0;
    if (pw == 0.0) continue;
    pp[i] = 0.0;
    const double* fr = P.frozen + (es * P.N + i) * Q;
    double* s2 = P.pool2 + (size_t)e * QW * P.N + i;
    for (int a = 0; a < QW; a++) s2[(size_t)a * P.N] += pw * pw * fr[st ? a : 2 * a + 1];
  }
}

// the slots' running second-moment sums into the tallies (end of data)
__global__ void __launch_bounds__(128) smc_pool_reduce_kernel(const __grid_constant__ Params P) {
  const int BS = blockDim.x, t = threadIdx.x, lane = t & 31;
  const long long i = (long long)blockIdx.x * BS + t;
  const bool live = i < P.N;
  const int E = P.m.E;
  const bool st = P.sm.P > 1;
  const int Q = st ? P.sm.Q : 4, QW = st ? Q : 2;
  for (int e = 0; e < E; e++)
    for (int a = 0; a < QW; a++) {
      double* s2 = P.pool2 + ((size_t)e * QW + a) * P.N;
      double v = live ? s2[i] : 0.0;
      if (live && v != 0.0) s2[i] = 0.0;
      v = warp_sum(v);
      if (lane == 0 && v != 0.0) atomicAdd(P.tallies + (size_t)(st ? Q + a : (a == 0 ? 2 : 5)) * E + e, v);
    }
}

// prior trees + first recombination position + weight 1/N (particleContainer.cpp:33-65)
template <bool ST>
__global__ void __launch_bounds__(128) smc_init_kernel(const __grid_constant__ Params P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int BS = AUX_BS;
  const int t = threadIdx.x;
  const ModelDev& m = P.m;
  const int n = m.n, nn = m.nn;
  auto T = tree_view_of<BS>(P, smem_raw, t, std::integral_constant<bool, ST>());
  const long long i = (long long)blockIdx.x * BS + t;
  if (i >= P.N) return;
  Rng r;
  rng_init(r, P.seed, P.slot_offset + (uint32_t)i, 0, P.chunk, RNG_INIT);
  if constexpr (ST) T.FLAG[t] = 0;
  init_tree(m, T, r);
  double nb = draw_next_base(m, tree_length(m, T), r, P.start_pos);
  const StateBuf& d = P.buf[0];
  for (int v = n; v < nn; v++) d.height[(size_t)(v - n) * P.N + i] = T.h(v);
  for (int v = 0; v < nn; v++) d.parent[(size_t)v * P.N + i] = T.par(v);
  if constexpr (ST) {
    for (int v = 0; v < nn; v++) d.pop[(size_t)v * P.N + i] = T.pop(v);
    const int nm = T.nm();
    d.nmarks[i] = (signed char)nm;
    for (int k = 0; k < nm; k++) {
      d.mtime[(size_t)k * P.N + i] = T.mt(k);
      d.mbranch[(size_t)k * P.N + i] = T.mb(k);
      d.mpop[(size_t)k * P.N + i] = T.mp(k);
    }
    if (T.FLAG[t]) atomicOr(&P.ctrl->sflags, (int)T.FLAG[t]);
  }
  d.logw[i] = -log((double)P.Ntot);
  d.next_base[i] = nb;
}

// ---------------------------------------------------------------------------------------------------------------
// stand-alone kernels (function-level parity entry points)
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_tree_functions(ModelDev m, long long count, const int* parent,
                                                        const double* height, const signed char* alleles, int what,
                                                        double* out0, double* out1) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int BS = AUX_BS;
  const int t = threadIdx.x;
  const int n = m.n, nn = m.nn;
  TreeViewT<BS> T = make_tree_view<BS>(smem_raw, m.n, nn, m.E, t);
  const long long i = (long long)blockIdx.x * BS + t;
  if (i >= count) return;
  for (int v = 0; v < nn; v++) {      // leaves sit at height 0 (checked on the host)
    if (v >= n) T.hw(v) = height[i * nn + v];
    T.par(v) = (signed char)parent[i * nn + v];
  }
  signed char data[32];
  for (int k = 0; k < n; k++) data[k] = alleles[i * n + k];
  if (what == 0) {
    out0[i] = site_likelihood(m, T, data);
  } else {
    out0[i] = branch_length_with_data(m, T, data);
    out1[i] = tree_length(m, T);
  }
}

// Lag calibration (calculate_median_survival_distances, smcsmc.cpp:169-263): one thread simulates one unconditioned
// ARG from the prior -- prior tree, then SMC' moves along the sequence -- and reports, for every internal node of the
// initial tree, the position at which a node of exactly that height is no longer in the local tree (or -1 if it
// survived to 0.6 * seqlen).  The host takes per-epoch medians.
struct SurvivalSink : NullSink, SNullSink {};
template <bool ST>
__global__ void __launch_bounds__(128) k_survival(const __grid_constant__ ModelDev m, const __grid_constant__ SModelDev sm,
                                                  long long count, uint64_t seed, double* dist, int* epoch, int* flags) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int BS = AUX_BS;
  const int t = threadIdx.x;
  const int n = m.n, nn = m.nn;
  auto T = [&]() {
    if constexpr (ST) return make_stree_view<BS>(smem_raw, m.n, nn, m.E, t, &sm);
    else return make_tree_view<BS>(smem_raw, m.n, nn, m.E, t);
  }();
  const long long i = (long long)blockIdx.x * BS + t;
  if (i >= count) return;
  Rng r;
  rng_init(r, seed, (uint32_t)i, 0, 0xCA11B, RNG_INIT);
  if constexpr (ST) T.FLAG[t] = 0;
  init_tree(m, T, r);
  // original heights live in the L0 scratch column; L1 marks them alive (1) / gone (0)
  int alive = n - 1;
  for (int j = 0; j < n - 1; j++) {
    double h = T.h(n + j);
    T.l0(n + j) = h; T.l1(n + j) = 1.0;
    dist[i * (n - 1) + j] = -1.0;
    epoch[i * (n - 1) + j] = epoch_of(m, h);
  }
  SurvivalSink sink;
  double logw = 0.0;
  double tl = tree_length(m, T);
  double next = draw_next_base(m, tl, r, 0.0);
  const double stop = m.seqlen * 0.6;
  bool ok = true;
  if constexpr (ST) ok = T.FLAG[t] == 0;
  while (ok && alive > 0 && next < stop) {
    smc_move(m, T, r, next, m.E, sink, logw, tl);
    if constexpr (ST) ok = T.FLAG[t] == 0;
    tl = tree_length(m, T);
    double pos = next;
    next = draw_next_base(m, tl, r, pos);
    // the reference records at most one vanished height per move (the loop breaks after the first, smcsmc.cpp:218-234)
    for (int j = 0; j < n - 1; j++) {
      if (T.l1(n + j) == 0.0) continue;
      double h = T.l0(n + j);
      bool present = false;
      for (int v = n; v < nn; v++) present |= (T.h(v) == h);
      if (!present) { dist[i * (n - 1) + j] = pos; T.l1(n + j) = 0.0; alive--; break; }
    }
  }
  if constexpr (ST) { if (T.FLAG[t]) atomicOr(flags, (int)T.FLAG[t]); }
}

// Terminal branch lengths of prior trees (calculate_terminal_branch_length_quantiles, smcsmc.cpp:128-167): one thread draws
// one prior tree and reports, per sample, the height of its parent coalescence (migrations are marks here, never nodes,
// so parent_height_ignoring_migrations is the parent slot's height) and the tree length.  dist[(leaf * count) + i].
template <bool ST>
__global__ void __launch_bounds__(128) k_terminal(const __grid_constant__ ModelDev m, const __grid_constant__ SModelDev sm,
                                                  long long count, uint64_t seed, double* dist, double* tlen, int* flags) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int BS = AUX_BS;
  const int t = threadIdx.x;
  const int n = m.n, nn = m.nn;
  auto T = [&]() {
    if constexpr (ST) return make_stree_view<BS>(smem_raw, m.n, nn, m.E, t, &sm);
    else return make_tree_view<BS>(smem_raw, m.n, nn, m.E, t);
  }();
  const long long i = (long long)blockIdx.x * BS + t;
  if (i >= count) return;
  Rng r;
  rng_init(r, seed, (uint32_t)i, 0, 0x7B1, RNG_INIT);
  if constexpr (ST) T.FLAG[t] = 0;
  init_tree(m, T, r);
  for (int v = 0; v < n; v++) { const int p = T.par(v); dist[(size_t)v * count + i] = p >= 0 ? T.hw(p) : 0.0; }
  tlen[i] = tree_length(m, T);
  if constexpr (ST) { if (T.FLAG[t]) atomicOr(flags, (int)T.FLAG[t]); }
}

__global__ void k_fastexp(const double* x, long long count, double* out) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) out[i] = fastexp(x[i]);
}

__global__ void k_build_chain(Ctrl* c, double U, long long N) { build_u_chain(c, U, N); }

// sequential partial sums + ESS on one thread (parity mode), or taken as given
__global__ void k_seq_partial_sums(const double* w, long long N, double* psum, double* ess) {
  double acc = 0.0;
  for (long long k = 0; k < N; k++) { psum[k] = acc; acc += w[k]; }
  psum[N] = acc;
  double scaling = 1.0 / acc, sq = 0.0;
  for (long long k = 0; k < N; k++) sq += (w[k] * scaling) * (w[k] * scaling);
  *ess = 1.0 / sq;
}

// blocked scan variant of the above (three tiny kernels folded into one single-block kernel; test sizes only)
__global__ void __launch_bounds__(1024) k_par_partial_sums(const double* w, long long N, double* psum, double* ess) {
  __shared__ double part[1024];
  __shared__ double sqp[1024];
  const int t = threadIdx.x, BS = blockDim.x;
  const long long per = (N + BS - 1) / BS, b0 = t * per, b1 = min(N, b0 + per);
  double run = 0.0, sq = 0.0;
  for (long long k = b0; k < b1; k++) { run += w[k]; sq += w[k] * w[k]; }
  part[t] = run; sqp[t] = sq;
  __syncthreads();
  if (t == 0) {
    double acc = 0.0, q = 0.0;
    for (int k = 0; k < BS; k++) { double v = part[k]; part[k] = acc; acc += v; q += sqp[k]; }
    psum[N] = acc;
    *ess = acc * acc / q;
  }
  __syncthreads();
  run = part[t];
  for (long long k = b0; k < b1; k++) { psum[k] = run; run += w[k]; }
}

// offspring counts from partial sums: count_i = #{ j : parent(j) == i }, by two binary searches over the exact u chain
__global__ void k_offspring_counts(const Ctrl* c, const double* psum, long long N, int* counts) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  const double norm = psum[N];
  // first j with u_j >= c_{i-1}/norm ... offspring of i are the j with c_{i-1}/norm <= u_j < c_i/norm, where
  // "c_i/norm > u_j" is the reference's comparison
  auto first_not_below = [&](double bound, bool last) -> long long {   // #{ j : !(bound > u_j) } complement count
    // returns number of j with bound > u_j  (u_j increasing)
    if (last) return N;   // the reference's loop assigns every remaining sample to the last interval reached
    long long lo = 0, hi = N;
    while (lo < hi) {
      long long mid = (lo + hi) >> 1;
      if (bound > u_at(c, mid)) lo = mid + 1; else hi = mid;
    }
    return lo;
  };
  long long hi_cnt = first_not_below(psum[i + 1] / norm, false);
  long long lo_cnt = i == 0 ? 0 : first_not_below(psum[i] / norm, false);
  counts[i] = (int)(hi_cnt - lo_cnt);
}

__global__ void __launch_bounds__(1024) k_normalize_log(double* logw, long long N, double* out) {
  __shared__ double red[32];
  const int t = threadIdx.x, BS = blockDim.x, lane = t & 31, warp = t >> 5, nwarp = BS >> 5;
  double mx = -DBL_MAX;
  for (long long k = t; k < N; k += BS) mx = fmax(mx, logw[k]);
  mx = warp_max(mx);
  if (lane == 0) red[warp] = mx;
  __syncthreads();
  mx = red[0];
  for (int w = 1; w < nwarp; w++) mx = fmax(mx, red[w]);
  __syncthreads();
  double s = 0.0, q = 0.0;
  for (long long k = t; k < N; k += BS) { double e = exp(logw[k] - mx); s += e; q += e * e; }
  s = warp_sum(s); q = warp_sum(q);
  if (lane == 0) red[warp] = s;
  __syncthreads();
  double S = 0.0;
  for (int w = 0; w < nwarp; w++) S += red[w];
  __syncthreads();
  if (lane == 0) red[warp] = q;
  __syncthreads();
  double Q = 0.0;
  for (int w = 0; w < nwarp; w++) Q += red[w];
  double lz = mx + log(S);
  for (long long k = t; k < N; k += BS) logw[k] -= lz;
  if (t == 0) { out[0] = lz; out[1] = S * S / Q; }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------
struct smc_engine {
  smc_config cfg;
  Params P;
  int BS = 128;
  size_t smem = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  double kernel_ms = 0.0;
  long long launches = 0;
  std::vector<void*> allocs;
  std::vector<size_t> alloc_bytes;
  bool rows_ready = false, inited = false, finished = false;
  long long gen = 0;                    // sweeps started on this engine (smc_reset); tags a sharded row's mailbox flags
  long long swept = 0;
  // GRAPH_ROWS consecutive row launches as one CUDA graph: the launches of a sweep are identical (the row index lives in
  // the device control block), so the graph is captured once and replayed; rebuilt when a parameter changed
  cudaGraphExec_t row_graph = nullptr;
  bool graph_dirty = true;
  std::vector<int> first_pending_h;     // per epoch: number of buckets the regular schedule consumes
  size_t pool2_count = 0;               // doubles in P.pool2
  int* first_pending_d = nullptr;
  // particle sharding
  void* state_slab = nullptr;
  void* ring_slab = nullptr;            // tally slab: open[2], map[2], frozen
  int Q = 4;                            // tally cells per epoch
  int tally_rows = 6;                   // rows of the tallies array: 6, or 2Q with structured populations
  size_t open_bytes = 0;
  ncclComm_t comm = nullptr;
  std::vector<void*> peer_maps;         // cudaIpcOpenMemHandle results, closed in smc_destroy
};

namespace {

// NCCL, bound at run time (torch's bundled libnccl.so.2 is already in the process when the caller is Python)
struct NcclApi {
  void* h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;

int nccl_load() {
  if (g_nccl.h) return 0;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return fail(SMC_ERR_CUDA, std::string("cannot load libnccl.so.2: ") + dlerror());
  g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(h, "ncclGetUniqueId");
  g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(h, "ncclCommInitRank");
  g_nccl.AllGather = (decltype(g_nccl.AllGather))dlsym(h, "ncclAllGather");
  g_nccl.AllReduce = (decltype(g_nccl.AllReduce))dlsym(h, "ncclAllReduce");
  g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(h, "ncclCommDestroy");
  g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(h, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllGather || !g_nccl.AllReduce || !g_nccl.CommDestroy ||
      !g_nccl.GetErrorString)
    return fail(SMC_ERR_CUDA, "libnccl.so.2 lacks a required symbol");
  g_nccl.h = h;
  return 0;
}
#define NCCL_TRY(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) return fail(SMC_ERR_CUDA, std::string("NCCL: ") + g_nccl.GetErrorString(r_)); } while (0)

}  // namespace

namespace {

// Device blocks of destroyed engines are kept for the next engine of the same shape (every EM iteration builds new
// engines with identical sizes; cudaMalloc / cudaFree of the multi-MB slabs cost up to hundreds of ms and synchronise the
// device).  Bounded: beyond POOL_CAP bytes per device blocks are really freed.  Slabs peers have mapped are not pooled.
struct BlockPool {
  std::mutex mu;
  std::multimap<std::pair<int, size_t>, void*> free_blocks;   // (device, bytes) -> block
  std::map<int, size_t> held;
  static constexpr size_t POOL_CAP = (size_t)16 << 30;
  void* take(int device, size_t bytes) {
    std::lock_guard<std::mutex> g(mu);
    auto it = free_blocks.find({device, bytes});
    if (it == free_blocks.end()) return nullptr;
    void* p = it->second;
    free_blocks.erase(it);
    held[device] -= bytes;
    return p;
  }
  bool give(int device, size_t bytes, void* p) {
    std::lock_guard<std::mutex> g(mu);
    if (held[device] + bytes > POOL_CAP) return false;
    free_blocks.insert({{device, bytes}, p});
    held[device] += bytes;
    return true;
  }
};
BlockPool g_pool;

// every cudaMalloc'ed block gets a serial number: a peer's cached mapping of (pointer, serial) stays valid exactly as long
// as that block lives, also when the allocator hands the same address out again later
std::mutex g_serial_mu;
std::map<void*, unsigned long long> g_block_serial;
unsigned long long g_next_serial = 1;

int raw_alloc(smc_engine* e, void** q, size_t bytes) {
  *q = g_pool.take(e->cfg.device, bytes);
  if (!*q) {
    CUDA_TRY(cudaMalloc(q, bytes));
    std::lock_guard<std::mutex> g(g_serial_mu);
    g_block_serial[*q] = g_next_serial++;
  }
  e->allocs.push_back(*q);
  e->alloc_bytes.push_back(bytes);
  return 0;
}

template <class Tp>
int dev_alloc(smc_engine* e, Tp** p, size_t count, bool zero = true) {
  void* q = nullptr;
  size_t bytes = std::max<size_t>(count, 1) * sizeof(Tp);
  int rc_ = raw_alloc(e, &q, bytes);
  if (rc_) return rc_;
  if (zero) CUDA_TRY(cudaMemsetAsync(q, 0, bytes, e->stream));
  *p = (Tp*)q;
  return 0;
}

int use_device(int device) {
  int count = 0;
  cudaError_t err = cudaGetDeviceCount(&count);
  if (err != cudaSuccess || count == 0)
    return fail(SMC_ERR_CUDA, std::string("no CUDA device available: ") + cudaGetErrorString(err) +
                                  " (libsmcgpu has no CPU fallback)");
  if (device < 0 || device >= count) return fail(SMC_ERR_ARG, "device ordinal out of range");
  CUDA_TRY(cudaSetDevice(device));
  return 0;
}

size_t tree_smem(int nn, int E, int BS) { return (size_t)BS * smc::tree_bytes_per_thread(nn, E) + 16; }
size_t stree_smem(int nn, int E, int K, int BS) { return (size_t)BS * smc::stree_bytes_per_thread(nn, E, K) + 16; }
// the row kernel: the tree columns plus what a particle carries between the rounds of a row (MoveSmem)
// dynamic shared memory a row kernel may ask for: the 227 KB a block can have minus its static arrays (4 KB bound)
#define SMC_MAX_DYN_SMEM 228352
size_t row_smem(int nn, int E, int BS) { return tree_smem(nn, E, BS) + (size_t)BS * smc::move_bytes_per_thread() + 16; }
size_t srow_smem(int nn, int E, int K, int BS) { return stree_smem(nn, E, K, BS) + (size_t)BS * smc::move_bytes_per_thread() + 16; }

int fill_model(ModelDev& m, int n, int E, double mu, double rho, double seqlen, int aa, const double* t_start,
               const double* inv2N, const double* vb, const uint8_t* mask) {
  if (n < 2 || n > SMC_MAX_HAPLOTYPES) return fail(SMC_ERR_ARG, "n must be in 2..32");
  if (E < 1 || E > SMC_MAX_EPOCHS) return fail(SMC_ERR_ARG, "E must be in 1..64");
  memset(&m, 0, sizeof(m));
  m.n = n; m.nn = 2 * n - 1; m.E = E; m.ancestral_aware = aa; m.mu = mu; m.rho = rho; m.seqlen = seqlen;
  for (int e = 0; e < E; e++) {
    m.t_start[e] = t_start ? t_start[e] : 0.0;
    m.inv2N[e] = inv2N ? inv2N[e] : 1.0;
    m.vb_logfac[e] = vb ? vb[e] : 0.0;
    m.rec_mask[e] = mask ? mask[e] : 3;
    if (e > 0 && !(m.t_start[e] > m.t_start[e - 1])) return fail(SMC_ERR_ARG, "t_start must be strictly increasing");
    if (!(m.inv2N[e] > 0)) return fail(SMC_ERR_ARG, "inv2N must be positive");
  }
  if (m.t_start[0] != 0.0) return fail(SMC_ERR_ARG, "t_start[0] must be 0");
  return 0;
}

// All ranks publish one slab (address, allocation serial, IPC handle) and map everybody else's; delta[r] = peer r's
// address - mine.  Mappings are cached per communicator and (peer, address, serial): the engines of one process come and
// go (one per EM iteration in the drop-in driver), their slabs return to the block pool and come back with the same
// address, and cudaIpcOpenMemHandle -- tens of milliseconds per peer -- is paid once per slab instead of once per engine.
struct SlabNote { unsigned long long ptr, serial; cudaIpcMemHandle_t handle; };
struct PeerMapKey {
  ncclComm_t comm; int peer; unsigned long long ptr, serial;
  bool operator<(const PeerMapKey& o) const {
    return std::tie(comm, peer, ptr, serial) < std::tie(o.comm, o.peer, o.ptr, o.serial);
  }
};
std::map<PeerMapKey, void*> g_peer_maps;
std::map<void*, cudaIpcMemHandle_t> g_own_handles;

int exchange_slab(smc_engine* e, void* slab, long long* delta) {
  const int G = e->P.world, me = e->P.rank;
  SlabNote mine;
  memset(&mine, 0, sizeof(mine));
  mine.ptr = (unsigned long long)slab;
  {
    std::lock_guard<std::mutex> g(g_serial_mu);
    mine.serial = g_block_serial[slab];
    auto it = g_own_handles.find(slab);
    if (it == g_own_handles.end() || mine.serial == 0) {
      cudaIpcMemHandle_t h;
      CUDA_TRY(cudaIpcGetMemHandle(&h, slab));
      g_own_handles[slab] = h;
      it = g_own_handles.find(slab);
    }
    mine.handle = it->second;
  }
  char* d_in = nullptr; char* d_out = nullptr;
  int rc;
  if ((rc = dev_alloc(e, &d_in, sizeof(mine), false)) || (rc = dev_alloc(e, &d_out, sizeof(mine) * (size_t)G, false))) return rc;
  CUDA_TRY(cudaMemcpyAsync(d_in, &mine, sizeof(mine), cudaMemcpyHostToDevice, e->stream));
  NCCL_TRY(g_nccl.AllGather(d_in, d_out, sizeof(mine), ncclChar, e->comm, e->stream));
  std::vector<SlabNote> all(G);
  CUDA_TRY(cudaMemcpyAsync(all.data(), d_out, sizeof(mine) * (size_t)G, cudaMemcpyDeviceToHost, e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  std::lock_guard<std::mutex> g(g_serial_mu);
  for (int r = 0; r < G; r++) {
    delta[r] = 0;
    if (r == me) continue;
    const PeerMapKey key{e->comm, r, all[r].ptr, all[r].serial};
    auto it = g_peer_maps.find(key);
    void* ptr = nullptr;
    if (it != g_peer_maps.end()) ptr = it->second;
    else {
      CUDA_TRY(cudaIpcOpenMemHandle(&ptr, all[r].handle, cudaIpcMemLazyEnablePeerAccess));
      g_peer_maps[key] = ptr;
    }
    delta[r] = (long long)((char*)ptr - (char*)slab);
  }
  return 0;
}

int launch_row(smc_engine* e, int mode) {
  int nb = e->P.nblocks;
  if (e->P.sm.P > 1 && e->P.world > 1) {
    switch (e->BS) {
      case 64: smc_row_kernel<64, true, true><<<nb, 64, e->smem, e->stream>>>(e->P, mode); break;
      default: smc_row_kernel<128, true, true><<<nb, 128, e->smem, e->stream>>>(e->P, mode); break;
    }
  } else if (e->P.sm.P > 1) {
    switch (e->BS) {
      case 32: smc_row_kernel<32, true, false><<<nb, 32, e->smem, e->stream>>>(e->P, mode); break;
      case 64: smc_row_kernel<64, true, false><<<nb, 64, e->smem, e->stream>>>(e->P, mode); break;
      default: smc_row_kernel<128, true, false><<<nb, 128, e->smem, e->stream>>>(e->P, mode); break;
    }
  } else if (e->P.world > 1) {
    switch (e->BS) {
      case 32: smc_row_kernel<32, false, true><<<nb, 32, e->smem, e->stream>>>(e->P, mode); break;
      case 64: smc_row_kernel<64, false, true><<<nb, 64, e->smem, e->stream>>>(e->P, mode); break;
      default: smc_row_kernel<128, false, true><<<nb, 128, e->smem, e->stream>>>(e->P, mode); break;
    }
  } else {
    switch (e->BS) {
      case 32: smc_row_kernel<32, false, false><<<nb, 32, e->smem, e->stream>>>(e->P, mode); break;
      case 64: smc_row_kernel<64, false, false><<<nb, 64, e->smem, e->stream>>>(e->P, mode); break;
      default: smc_row_kernel<128, false, false><<<nb, 128, e->smem, e->stream>>>(e->P, mode); break;
    }
  }
  e->launches++;
  // a launch that cannot be queued (bad configuration, captured stream invalidated) is reported where it happens
  const cudaError_t le = cudaPeekAtLastError();
  if (le != cudaSuccess) return fail(SMC_ERR_CUDA, std::string("row kernel launch: ") + cudaGetErrorString(le));
  return 0;
}

// `count` consecutive rows of an unsharded engine: batches of GRAPH_ROWS through the captured graph (one host call per
// batch instead of one ~11 us launch per row -- the host launch rate bounds small populations), the rest directly
int GRAPH_ROWS = 32;
int launch_rows(smc_engine* e, int64_t count) {
  static const bool no_graph = getenv("SMC_NO_GRAPH") != nullptr;
  static const bool tuned = [] { if (const char* g = getenv("SMC_GRAPH_ROWS")) { int v = atoi(g); if (v >= 2 && v <= 4096) GRAPH_ROWS = v; } return true; }();
  (void)tuned;
  if (!no_graph && count >= GRAPH_ROWS) {
    if (!e->row_graph || e->graph_dirty) {
      if (e->row_graph) { CUDA_TRY(cudaGraphExecDestroy(e->row_graph)); e->row_graph = nullptr; }
      cudaGraph_t g = nullptr;
      const long long before = e->launches;
      CUDA_TRY(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
      int lrc = 0;
      for (int k = 0; k < GRAPH_ROWS && !lrc; k++) lrc = launch_row(e, 0);
      const cudaError_t ce_end = cudaStreamEndCapture(e->stream, &g);      // always end the capture, also after a failure
      e->launches = before;
      if (lrc) { if (g) cudaGraphDestroy(g); return lrc; }
      if (ce_end != cudaSuccess) return fail(SMC_ERR_CUDA, std::string("cudaStreamEndCapture: ") + cudaGetErrorString(ce_end));
      cudaError_t ce = cudaGraphInstantiate(&e->row_graph, g, 0);
      cudaGraphDestroy(g);
      if (ce != cudaSuccess) { e->row_graph = nullptr; return fail(SMC_ERR_CUDA, cudaGetErrorString(ce)); }
      e->graph_dirty = false;
    }
    for (; count >= GRAPH_ROWS; count -= GRAPH_ROWS) {
      CUDA_TRY(cudaGraphLaunch(e->row_graph, e->stream));
      e->launches += GRAPH_ROWS;
    }
  }
  for (; count > 0; count--) {
    int rc = launch_row(e, 0);
    if (rc) return rc;
  }
  return 0;
}

int materialise(smc_engine* e) {
  int rc = launch_row(e, 1);
  if (rc) return rc;
  smc_commit_kernel<<<1, 1, 0, e->stream>>>(e->P, e->P.ctrl);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

}  // namespace

extern "C" {

static int read_ctrl(smc_engine* e, Ctrl* c);

const char* smc_last_error(void) { return g_err.c_str(); }
const char* smc_version(void) { return "smcsmc_b200 0.1 (sm_100a)"; }

// SMC_DEBUG_ARGS=1: one line per call with the scalar arguments and an FNV-1a checksum of every array, to compare what two
// host programs hand to the engine
static unsigned long long dbg_hash(const void* p, size_t bytes) {
  unsigned long long h = 1469598103934665603ull;
  const unsigned char* c = (const unsigned char*)p;
  for (size_t k = 0; k < bytes; k++) { h ^= c[k]; h *= 1099511628211ull; }
  return p ? h : 0ull;
}
static bool dbg_args() { static const bool on = getenv("SMC_DEBUG_ARGS") != nullptr; return on; }

int smc_create(const smc_config* cfg, smc_engine** out) {
  if (cfg && dbg_args())
    fprintf(stderr, "[smc args] create n=%d N=%d E=%d seed=%llu chunk=%u slot=%u mu=%.17g rho=%.17g seqlen=%.17g start=%.17g ess=%.17g aa=%d "
            "t_start=%llx inv2N=%llx vb=%llx mask=%llx lags=%llx\n", cfg->n, cfg->N, cfg->E, (unsigned long long)cfg->seed, cfg->chunk,
            cfg->slot_offset, cfg->mu, cfg->rho, cfg->seqlen, cfg->start_pos, cfg->ess_fraction, cfg->ancestral_aware,
            dbg_hash(cfg->t_start, cfg->E * 8), dbg_hash(cfg->inv2N, cfg->E * 8), dbg_hash(cfg->vb_logfac, cfg->vb_logfac ? cfg->E * 8 : 0),
            dbg_hash(cfg->rec_mask, cfg->rec_mask ? cfg->E : 0), dbg_hash(cfg->lags, cfg->E * 8));
  if (!cfg || !out) return fail(SMC_ERR_ARG, "null argument");
  *out = nullptr;
  if (cfg->N < 1) return fail(SMC_ERR_ARG, "N must be >= 1");
  if (!cfg->lags) return fail(SMC_ERR_ARG, "lags required");
  if (!(cfg->mu >= 0) || !(cfg->rho > 0) || !(cfg->seqlen > 0)) return fail(SMC_ERR_ARG, "mu/rho/seqlen invalid");
  int rc = use_device(cfg->device);
  if (rc) return rc;
  smc_engine* e = new smc_engine();
  e->cfg = *cfg;
  Params& P = e->P;
  memset(&P, 0, sizeof(P));
  rc = fill_model(P.m, cfg->n, cfg->E, cfg->mu, cfg->rho, cfg->seqlen, cfg->ancestral_aware, cfg->t_start, cfg->inv2N,
                  cfg->vb_logfac, cfg->rec_mask);
  if (rc) { delete e; return rc; }
  P.N = cfg->N; P.parity_scan = cfg->parity_scan; P.seed = cfg->seed; P.chunk = cfg->chunk;
  P.slot_offset = cfg->slot_offset; P.ess_fraction = cfg->ess_fraction; P.start_pos = cfg->start_pos;
  // block size: keep every SM busy when there are few particles
  e->BS = (cfg->N < 8192) ? 64 : 128;     // measured: 128 is 3-6 % faster from 10^4 particles on (profiles/r1_occupancy_probe.txt)
  if (const char* s = getenv("SMC_BLOCK")) { int b = atoi(s); if (b == 32 || b == 64 || b == 128) e->BS = b; }   // tuning knob
  P.nblocks = (cfg->N + e->BS - 1) / e->BS;
  e->smem = row_smem(P.m.nn, P.m.E, e->BS);
  P.pool_on = getenv("SMC_NO_POOL") ? 0 : 1;   // knob: per-particle w^2 in the *_weight tallies (ESS column too high), 4 % faster
  if (getenv("SMC_FLAT_SEARCH")) P.flat_search = 1;   // tests: the one-level parent search the two-level one must reproduce
  if (const char* s = getenv("SMC_EXTRA_SMEM")) e->smem += (size_t)atoi(s);                                       // occupancy experiments
  cudaError_t ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
  if (ce != cudaSuccess) { delete e; return fail(SMC_ERR_CUDA, cudaGetErrorString(ce)); }
  cudaEventCreate(&e->ev0); cudaEventCreate(&e->ev1);
  // opt in to the full 227 KB once (the attribute is per function, not per engine)
  if (e->smem > SMC_MAX_DYN_SMEM) { smc_destroy(e); return fail(SMC_ERR_ARG, "n and E need more shared memory than one SM has"); }
  ce = cudaFuncSetAttribute(smc_row_kernel<32, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<64, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<128, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<32, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<64, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<128, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (const char* cv = getenv("SMC_CARVEOUT")) {   // tuning knob: shared-memory share of the 256 KB L1/shared array, in percent
    const int pct = atoi(cv);
    if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<128, false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
    if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<64, false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
  }
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_init_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<32, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<64, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<128, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<64, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_row_kernel<128, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce == cudaSuccess) ce = cudaFuncSetAttribute(smc_init_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM);
  if (ce != cudaSuccess) { std::string msg = cudaGetErrorString(ce); smc_destroy(e); return fail(SMC_ERR_CUDA, "shared memory opt-in: " + msg); }
  const size_t N = cfg->N;
  const int n = cfg->n, nn = 2 * n - 1;
  // everything a pulling offspring may read of a parent on another rank sits in one slab (one IPC handle per rank)
  {
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t s_h = up((size_t)(n - 1) * N * 8), s_p = up((size_t)nn * N), s_d = up(N * 8),
                 s_b1 = up(((size_t)P.nblocks + 1) * 8), s_b = up((size_t)P.nblocks * 8);
    const size_t per = s_h + s_p + s_d + s_d + s_b1 + s_b;
    const size_t s_mbox = up(2 * (size_t)SMC_MAX_RANKS * 8 * sizeof(double));   // per-row mailbox of a sharded engine
    void* slab = nullptr;
    if ((rc = raw_alloc(e, &slab, 2 * per + s_mbox))) { smc_destroy(e); return rc; }
    cudaMemsetAsync(slab, 0, 2 * per + s_mbox, e->stream);
    e->state_slab = slab;
    P.mbox = (double*)((char*)slab + 2 * per);
    char* q = (char*)slab;
    for (int b = 0; b < 2; b++) {
      P.buf[b].height = (double*)q; q += s_h;
      P.buf[b].parent = (signed char*)q; q += s_p;
      P.buf[b].next_base = (double*)q; q += s_d;
      P.local_cum[b] = (double*)q; q += s_d;
      P.boff[b] = (double*)q; q += s_b1;
      P.bscale[b] = (double*)q; q += s_b;
    }
  }
  for (int b = 0; b < 2; b++) {
    if ((rc = dev_alloc(e, &P.buf[b].logw, N))) { smc_destroy(e); return rc; }
    if ((rc = dev_alloc(e, &P.bmax[b], P.nblocks))) { smc_destroy(e); return rc; }
    if ((rc = dev_alloc(e, &P.bsum[b], P.nblocks))) { smc_destroy(e); return rc; }
    if ((rc = dev_alloc(e, &P.bsq[b], P.nblocks))) { smc_destroy(e); return rc; }
    if ((rc = dev_alloc(e, &P.barea[b], P.nblocks))) { smc_destroy(e); return rc; }
  }
  P.rank = 0; P.world = 1; P.Ntot = (long long)N; P.base = 0;
  if ((rc = dev_alloc(e, &P.psum, cfg->parity_scan ? N + 1 : 1))) { smc_destroy(e); return rc; }
  if ((rc = dev_alloc(e, &P.tallies, 6 * (size_t)cfg->E))) { smc_destroy(e); return rc; }
  if ((rc = dev_alloc(e, &P.pool, 2 * (size_t)cfg->E * N))) { smc_destroy(e); return rc; }
  if ((rc = dev_alloc(e, &P.pool2, 2 * (size_t)cfg->E * N))) { smc_destroy(e); return rc; }
  e->pool2_count = 2 * (size_t)cfg->E * N;
  if ((rc = dev_alloc(e, &P.anc, N))) { smc_destroy(e); return rc; }
  if ((rc = dev_alloc(e, &P.ctrl, 1))) { smc_destroy(e); return rc; }
  cudaMemsetAsync(P.anc, 0xff, N * sizeof(int), e->stream);
  {
    Ctrl init;
    memset(&init, 0, sizeof(init));
    init.cons_done = -1;
    cudaMemcpyAsync(P.ctrl, &init, sizeof(init), cudaMemcpyHostToDevice, e->stream);
    ce = cudaStreamSynchronize(e->stream);
    if (ce != cudaSuccess) { std::string msg = cudaGetErrorString(ce); smc_destroy(e); return fail(SMC_ERR_CUDA, msg); }
  }
  *out = e;
  return 0;
}

void smc_destroy(smc_engine* e) {
  if (!e) return;
  cudaSetDevice(e->cfg.device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  if (e->row_graph) cudaGraphExecDestroy(e->row_graph);
  // The communicator and the peers' mappings of this engine's slabs belong to the process-wide caches (g_comms,
  // g_peer_maps) and outlive the engine; its blocks -- slabs included -- go back to the pool.  A block the pool turns
  // down is really freed: its serial number dies with it, so no peer will ever match its cached mapping again.
  for (size_t k = 0; k < e->allocs.size(); k++)
    if (!g_pool.give(e->cfg.device, e->alloc_bytes[k], e->allocs[k])) {
      {
        std::lock_guard<std::mutex> g(g_serial_mu);
        g_block_serial.erase(e->allocs[k]);
        g_own_handles.erase(e->allocs[k]);
      }
      cudaFree(e->allocs[k]);
    }
  if (e->ev0) cudaEventDestroy(e->ev0);
  if (e->ev1) cudaEventDestroy(e->ev1);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

int smc_shard_unique_id(uint8_t id[128]) {
  if (!id) return fail(SMC_ERR_ARG, "null argument");
  int rc = nccl_load();
  if (rc) return rc;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId u;
  NCCL_TRY(g_nccl.GetUniqueId(&u));
  memcpy(id, &u, 128);
  return 0;
}

int smc_shard_attach(smc_engine* e, int32_t rank, int32_t world, const uint8_t id[128]) {
  if (e) e->graph_dirty = true;
  if (!e || !id) return fail(SMC_ERR_ARG, "null argument");
  if (world < 2 || world > SMC_MAX_RANKS || rank < 0 || rank >= world) return fail(SMC_ERR_ARG, "rank/world out of range (2..8 ranks)");
  if (e->rows_ready || e->inited) return fail(SMC_ERR_STATE, "attach the shard before uploading rows");
  if (e->P.world > 1) return fail(SMC_ERR_STATE, "engine is already sharded");
  if (e->P.sm.P > 1 && e->BS == 32) return fail(SMC_ERR_ARG, "a sharded structured engine needs blocks of 64 or 128 particles");
  if (e->cfg.parity_scan) return fail(SMC_ERR_ARG, "parity_scan is a single-rank mode");
  if ((long long)e->cfg.N * world > 0x7fffffffLL) return fail(SMC_ERR_ARG, "more than 2^31 particles in total");
  int rc = nccl_load();
  if (rc) return rc;
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  {
    // one communicator per (id, rank, world, device) and process: engines attached with the same id share it
    static std::mutex mu;
    static std::map<std::string, ncclComm_t> comms;
    std::lock_guard<std::mutex> g(mu);
    std::string key((const char*)id, 128);
    key += "/" + std::to_string(rank) + "/" + std::to_string(world) + "/" + std::to_string(e->cfg.device);
    auto it = comms.find(key);
    if (it == comms.end()) {
      ncclUniqueId u;
      memcpy(&u, id, 128);
      ncclComm_t comm = nullptr;
      NCCL_TRY(g_nccl.CommInitRank(&comm, world, u, rank));
      it = comms.emplace(key, comm).first;
    }
    e->comm = it->second;
  }
  Params& P = e->P;
  P.rank = rank; P.world = world;
  P.Ntot = (long long)e->cfg.N * world;
  P.base = (long long)e->cfg.N * rank;
  P.slot_offset = e->cfg.slot_offset + (uint32_t)P.base;      // Philox streams are keyed by the global particle slot
  // a slab from the pool may carry an earlier engine's flags: clear the mailbox before any peer can reach it (the slab
  // exchange below is a collective, so no peer maps this slab before the memset is queued)
  CUDA_TRY(cudaMemsetAsync(P.mbox, 0, 2 * (size_t)SMC_MAX_RANKS * 8 * sizeof(double), e->stream));
  for (int b = 0; b < 2; b++)
    if ((rc = dev_alloc(e, &P.rank_off[b], world + 1)) || (rc = dev_alloc(e, &P.rank_fac[b], world))) return rc;
  return exchange_slab(e, e->state_slab, P.peer_state);
}

int smc_set_populations(smc_engine* e, int32_t npop, const double* inv2N, const double* mig, const double* single_mig,
                        const int32_t* leaf_pop, int32_t max_marks) {
  if (e) e->graph_dirty = true;
  if (!e || !inv2N || !leaf_pop) return fail(SMC_ERR_ARG, "null argument");
  if (npop < 2 || npop > SMC_MAX_POPULATIONS) return fail(SMC_ERR_ARG, "populations must be in 2..4");
  Params& P = e->P;
  if (P.world > 1 && e->BS == 32) return fail(SMC_ERR_ARG, "a sharded structured engine needs blocks of 64 or 128 particles");
  const bool update = P.sm.P > 1;
  if (update && P.sm.P != npop) return fail(SMC_ERR_STATE, "the number of populations cannot change");
  if (!update && e->rows_ready) return fail(SMC_ERR_STATE, "set the populations before uploading rows");
  if (e->inited && e->swept > 0) return fail(SMC_ERR_STATE, "populations cannot change during a sweep");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  const int E = e->cfg.E, n = e->cfg.n, nn = 2 * n - 1, Pp = npop;
  std::vector<double> h_inv((size_t)E * Pp), h_mig((size_t)E * Pp * Pp, 0.0), h_tot((size_t)E * Pp, 0.0), h_single((size_t)E * Pp * Pp, 0.0);
  std::vector<unsigned char> h_fixed(E, 0);
  std::vector<signed char> h_leaf(n);
  for (int ep = 0; ep < E; ep++)
    for (int i = 0; i < Pp; i++) {
      h_inv[(size_t)ep * Pp + i] = inv2N[(size_t)ep * Pp + i];
      if (!(h_inv[(size_t)ep * Pp + i] > 0)) return fail(SMC_ERR_ARG, "inv2N must be positive");
      double tot = 0.0, stot = 0.0;
      for (int j = 0; j < Pp; j++) {
        const size_t at = ((size_t)ep * Pp + i) * Pp + j;
        double mr = (mig && j != i) ? mig[at] : 0.0;
        double sm = single_mig ? single_mig[at] : 0.0;
        if (!(mr >= 0)) return fail(SMC_ERR_ARG, "migration rates must be non-negative");
        if (!(sm >= 0 && sm <= 1)) return fail(SMC_ERR_ARG, "single_mig entries are probabilities");
        h_mig[at] = mr; h_single[at] = sm;
        if (j != i) tot += mr;
        stot += sm;
        if (sm > 0) h_fixed[ep] = 1;
      }
      if (stot > 1.0 + 1e-12) return fail(SMC_ERR_ARG, "single_mig rows must sum to at most 1");
      h_tot[(size_t)ep * Pp + i] = tot;
    }
  for (int i = 0; i < n; i++) {
    if (leaf_pop[i] < 0 || leaf_pop[i] >= Pp) return fail(SMC_ERR_ARG, "leaf population out of range");
    h_leaf[i] = (signed char)leaf_pop[i];
  }
  int rc;
  if (!update) {
    const int K = max_marks > 0 ? max_marks : 2 * n + 8;
    if (K > 120) return fail(SMC_ERR_ARG, "max_marks must be at most 120");
    const int Q = 2 + 3 * Pp + Pp * Pp;
    size_t smem = srow_smem(nn, E, K, e->BS);
    smem = std::max(smem, (size_t)E * 2 * Q * (e->BS / 32) * sizeof(double) + 16);   // hand-over staging of phase 0
    if (smem > SMC_MAX_DYN_SMEM || stree_smem(nn, E, K, AUX_BS) > SMC_MAX_DYN_SMEM)
      return fail(SMC_ERR_ARG, "n, E and max_marks need more shared memory than one SM has (fewer particles per block: SMC_BLOCK=32)");
    double *d_inv, *d_mig, *d_tot, *d_single, *d_vbc, *d_vbm; unsigned char* d_fixed; signed char* d_leaf;
    if ((rc = dev_alloc(e, &d_vbc, h_inv.size(), false)) || (rc = dev_alloc(e, &d_vbm, h_mig.size(), false))) return rc;
    CUDA_TRY(cudaMemsetAsync(d_vbc, 0, h_inv.size() * sizeof(double), e->stream));
    CUDA_TRY(cudaMemsetAsync(d_vbm, 0, h_mig.size() * sizeof(double), e->stream));
    if ((rc = dev_alloc(e, &d_inv, h_inv.size(), false)) || (rc = dev_alloc(e, &d_mig, h_mig.size(), false)) ||
        (rc = dev_alloc(e, &d_tot, h_tot.size(), false)) || (rc = dev_alloc(e, &d_single, h_single.size(), false)) ||
        (rc = dev_alloc(e, &d_fixed, (size_t)E, false)) || (rc = dev_alloc(e, &d_leaf, (size_t)n, false)))
      return rc;
    // the per-particle populations and migration marks are laid out with the tallies (smc_upload_rows): one slab that
    // the peers of a sharded population map
    if ((rc = dev_alloc(e, &P.tallies, 2 * (size_t)Q * E))) return rc;
    if ((rc = dev_alloc(e, &P.pool2, (size_t)Q * E * P.N))) return rc;
    e->pool2_count = (size_t)Q * E * P.N;
    P.sm.K = K; P.sm.Q = Q;
    P.sm.inv2N = d_inv; P.sm.mig = d_mig; P.sm.migtot = d_tot; P.sm.single_mig = d_single; P.sm.fixed = d_fixed;
    P.sm.leaf_pop = d_leaf;
    P.sm.vb_coal = d_vbc; P.sm.vb_mig = d_vbm;
    e->Q = Q; e->tally_rows = 2 * Q; e->smem = smem;
  }
  CUDA_TRY(cudaMemcpyAsync((void*)P.sm.inv2N, h_inv.data(), h_inv.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync((void*)P.sm.mig, h_mig.data(), h_mig.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync((void*)P.sm.migtot, h_tot.data(), h_tot.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync((void*)P.sm.single_mig, h_single.data(), h_single.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync((void*)P.sm.fixed, h_fixed.data(), h_fixed.size(), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync((void*)P.sm.leaf_pop, h_leaf.data(), h_leaf.size(), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  P.sm.P = Pp;
  return 0;
}

int smc_set_population_vb(smc_engine* e, const double* coal_logfac, const double* mig_logfac) {
  if (!e) return fail(SMC_ERR_ARG, "null engine");
  Params& P = e->P;
  if (P.sm.P < 2) return fail(SMC_ERR_STATE, "call smc_set_populations first");
  if (e->inited && e->swept > 0 && !e->finished) return fail(SMC_ERR_STATE, "the -vb factors cannot change during a sweep");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  const size_t E = (size_t)e->cfg.E, Pp = (size_t)P.sm.P;
  std::vector<double> c(E * Pp, 0.0), mg(E * Pp * Pp, 0.0);
  for (size_t k = 0; k < c.size(); k++) if (coal_logfac) { c[k] = coal_logfac[k]; if (!std::isfinite(c[k])) return fail(SMC_ERR_ARG, "non-finite -vb factor"); }
  for (size_t k = 0; k < mg.size(); k++) if (mig_logfac) { mg[k] = mig_logfac[k]; if (!std::isfinite(mg[k])) return fail(SMC_ERR_ARG, "non-finite -vb factor"); }
  CUDA_TRY(cudaMemcpyAsync((void*)P.sm.vb_coal, c.data(), c.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync((void*)P.sm.vb_mig, mg.data(), mg.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  return 0;
}

int smc_tally_rows(smc_engine* e) { return e ? e->tally_rows : 0; }

int smc_get_population_state(smc_engine* e, int32_t* pop, int32_t* nmarks, int32_t* mark_branch, double* mark_time,
                             int32_t* mark_pop) {
  if (!e) return fail(SMC_ERR_ARG, "null engine");
  if (e->P.sm.P < 2) return fail(SMC_ERR_STATE, "the engine has a single population");
  if (!e->inited) return fail(SMC_ERR_STATE, "particles not initialised");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  if (!e->finished) { int rc = materialise(e); if (rc) return rc; }
  Ctrl c;
  int rc = read_ctrl(e, &c);
  if (rc) return rc;
  const size_t N = e->cfg.N;
  const int nn = 2 * e->cfg.n - 1, K = e->P.sm.K;
  const StateBuf& s = e->P.buf[c.cur];
  std::vector<signed char> hp((size_t)nn * N), hn(N), hb((size_t)K * N), hq((size_t)K * N);
  std::vector<double> ht((size_t)K * N);
  CUDA_TRY(cudaMemcpy(hp.data(), s.pop, hp.size(), cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(hn.data(), s.nmarks, hn.size(), cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(hb.data(), s.mbranch, hb.size(), cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(hq.data(), s.mpop, hq.size(), cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(ht.data(), s.mtime, ht.size() * sizeof(double), cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < N; i++) {
    if (pop) for (int v = 0; v < nn; v++) pop[i * nn + v] = hp[(size_t)v * N + i];
    if (nmarks) nmarks[i] = hn[i];
    for (int k = 0; k < K; k++) {
      const bool used = k < hn[i];
      if (mark_branch) mark_branch[i * K + k] = used ? hb[(size_t)k * N + i] : -1;
      if (mark_time) mark_time[i * K + k] = used ? ht[(size_t)k * N + i] : 0.0;
      if (mark_pop) mark_pop[i * K + k] = used ? hq[(size_t)k * N + i] : -1;
    }
  }
  return 0;
}

int smc_upload_rows(smc_engine* e, const double* x_end, const int8_t* alleles, const uint8_t* is_site,
                    const int32_t* max_rec_epoch, int64_t R) {
  if (e && dbg_args())
    fprintf(stderr, "[smc args] rows R=%lld x_end=%llx alleles=%llx is_site=%llx max_rec=%llx\n", (long long)R,
            dbg_hash(x_end, (size_t)R * 8), dbg_hash(alleles, (size_t)R * e->cfg.n), dbg_hash(is_site, (size_t)R),
            dbg_hash(max_rec_epoch, max_rec_epoch ? (size_t)R * 4 : 0));
  if (e) e->graph_dirty = true;
  if (!e || !x_end || !alleles || !is_site) return fail(SMC_ERR_ARG, "null argument");
  if (R < 1) return fail(SMC_ERR_ARG, "R must be >= 1");
  if (e->rows_ready) return fail(SMC_ERR_STATE, "rows already uploaded");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  const int E = e->cfg.E, n = e->cfg.n;
  for (int64_t r = 0; r < R; r++) {
    double prev = r ? x_end[r - 1] : e->cfg.start_pos;
    // a zero-length row is legal: two typed sites at one position (the reference just skips the extension loop)
    if (!(x_end[r] >= prev)) return fail(SMC_ERR_ARG, "row ends must be non-decreasing and not before start_pos");
    if (x_end[r] > e->cfg.seqlen) return fail(SMC_ERR_ARG, "row end beyond seqlen");
    for (int k = 0; k < n; k++) {
      int a = alleles[r * n + k];
      if (a < -1 || a > 2) return fail(SMC_ERR_ARG, "allele code must be -1, 0, 1 or 2");
      if (a == 2 && !((k % 2 == 0 && k + 1 < n && alleles[r * n + k + 1] == 2) || (k % 2 == 1 && alleles[r * n + k - 1] == 2)))
        return fail(SMC_ERR_ARG, "unphased het code 2 must come in (even, odd) pairs");
    }
  }
  // ---- lagged-tally schedule: CountModel::extract_and_update_count's update positions (count.cpp:363-385) ----
  // bucket k of epoch e holds the records whose position lies in [ub[e][k-1], ub[e][k]); it is consumed after the row
  // whose update made update_to = ub[e][k].  The last bucket (+inf) is consumed by smc_finish.
  std::vector<std::vector<double>> ub(E);
  std::vector<double> counted_to(E, 0.0);
  std::vector<int> cons_off(R + 1, 0);
  std::vector<int2> cons;
  for (int64_t r = 0; r < R; r++) {
    int first = E;
    for (int ep = 0; ep < E; ep++) {
      double lagging = e->cfg.lags[ep];
      double x_e = x_end[r] - lagging;
      if ((x_e - counted_to[ep]) < lagging * 0.1 && first > ep) continue;   // no update
      if (ep < first) first = ep;
      counted_to[ep] = x_e;
      // a bound at or below the previous one closes an empty window: nothing can be in it, skip the bucket
      if (!ub[ep].empty() && !(x_e > ub[ep].back())) continue;
      if (ub[ep].empty() && !(x_e > e->cfg.start_pos)) continue;
      cons.push_back(make_int2(ep, (int)ub[ep].size()));
      ub[ep].push_back(x_e);
    }
    cons_off[r + 1] = (int)cons.size();
  }
  e->first_pending_h.assign(E, 0);
  std::vector<int> ub_off(E + 1, 0);
  std::vector<double> ub_flat;
  for (int ep = 0; ep < E; ep++) {
    e->first_pending_h[ep] = (int)ub[ep].size();
    ub[ep].push_back(HUGE_VAL);
    ub_off[ep + 1] = ub_off[ep] + (int)ub[ep].size();
    ub_flat.insert(ub_flat.end(), ub[ep].begin(), ub[ep].end());
  }
  // bucket at each row's left end (row R: after the last row), which buckets close inside each row, the oldest pending
  // bucket per launch, and the ring depth: a slot is re-used only by a bucket born after the launch whose hand-over
  // consumed the old occupant (other blocks of that launch may still be reading it)
  std::vector<int> kstart((size_t)(R + 1) * E), kslot((size_t)(R + 1) * E), kfirst((size_t)(R + 1) * E);
  std::vector<unsigned long long> closes(R, 0ull);
  std::vector<int> kcur(E, 0);
  for (int64_t r = 0; r <= R; r++) {
    double x0 = r ? x_end[r - 1] : e->cfg.start_pos;
    for (int ep = 0; ep < E; ep++) {
      while (ub[ep][kcur[ep]] <= x0) kcur[ep]++;
      kstart[(size_t)r * E + ep] = kcur[ep];
    }
  }
  {
    std::vector<int> done(E, 0);
    for (int64_t r = 0; r <= R; r++) {       // launch r hands over cons[cons_off[r-1] .. cons_off[r])
      if (r > 0) for (int q = cons_off[r - 1]; q < cons_off[r]; q++) done[cons[q].x] = cons[q].y + 1;
      for (int ep = 0; ep < E; ep++) kfirst[(size_t)r * E + ep] = done[ep];
    }
  }
  int B = 2;
  for (int64_t r = 0; r < R; r++)
    for (int ep = 0; ep < E; ep++) {
      const int ks = kstart[(size_t)r * E + ep], kn = kstart[(size_t)(r + 1) * E + ep];
      if (kn > ks) closes[r] |= 1ull << ep;
      const int oldest = r > 0 ? kfirst[(size_t)(r - 1) * E + ep] : 0;
      B = std::max(B, kn - oldest + 1);
    }
  B += 1;      // a consumed bucket's records are read once more by the next launch (pooled second moments, phase 0)
  Params& P = e->P;
  P.B = B;
  P.R = R;
  std::vector<int> mr(R);
  for (int64_t r = 0; r < R; r++) mr[r] = max_rec_epoch ? max_rec_epoch[r] : E - 1;
  int rc;
  double* d_x; signed char* d_al; unsigned char* d_site; int* d_mr; int* d_ks; int* d_ksl; int* d_uboff; double* d_ub; int* d_co; int2* d_cons; int* d_kf; unsigned long long* d_cl;
  if ((rc = dev_alloc(e, &d_x, R, false)) || (rc = dev_alloc(e, &d_al, (size_t)R * n, false)) ||
      (rc = dev_alloc(e, &d_site, R, false)) || (rc = dev_alloc(e, &d_mr, R, false)) ||
      (rc = dev_alloc(e, &d_ks, (size_t)(R + 1) * E, false)) || (rc = dev_alloc(e, &d_ksl, (size_t)(R + 1) * E, false)) ||
      (rc = dev_alloc(e, &d_kf, (size_t)(R + 1) * E, false)) || (rc = dev_alloc(e, &d_cl, R, false)) || (rc = dev_alloc(e, &d_uboff, E + 1, false)) ||
      (rc = dev_alloc(e, &d_ub, ub_flat.size(), false)) || (rc = dev_alloc(e, &d_co, R + 1, false)) ||
      (rc = dev_alloc(e, &d_cons, cons.size(), false)) || (rc = dev_alloc(e, &e->first_pending_d, E, false)))
    return rc;
  CUDA_TRY(cudaMemcpyAsync(d_x, x_end, R * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync(d_al, alleles, (size_t)R * n, cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync(d_site, is_site, R, cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync(d_mr, mr.data(), R * sizeof(int), cudaMemcpyHostToDevice, e->stream));
  for (size_t q = 0; q < kstart.size(); q++) kslot[q] = kstart[q] % B;
  CUDA_TRY(cudaMemcpyAsync(d_ks, kstart.data(), kstart.size() * sizeof(int), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync(d_ksl, kslot.data(), kslot.size() * sizeof(int), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync(d_kf, kfirst.data(), kfirst.size() * sizeof(int), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync(d_cl, closes.data(), closes.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync(d_uboff, ub_off.data(), ub_off.size() * sizeof(int), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync(d_ub, ub_flat.data(), ub_flat.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync(d_co, cons_off.data(), cons_off.size() * sizeof(int), cudaMemcpyHostToDevice, e->stream));
  if (!cons.empty())
    CUDA_TRY(cudaMemcpyAsync(d_cons, cons.data(), cons.size() * sizeof(int2), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaMemcpyAsync(e->first_pending_d, e->first_pending_h.data(), E * sizeof(int), cudaMemcpyHostToDevice, e->stream));
  P.x_end = d_x; P.alleles = d_al; P.is_site = d_site; P.max_rec = d_mr; P.kstart = d_ks; P.kslot = d_ksl; P.ub_off = d_uboff; P.ub = d_ub;
  P.cons_off = d_co; P.cons = d_cons; P.kfirst = d_kf; P.closes = d_cl;
  // tally storage (one slab, one IPC handle): open cells and ancestor maps twice, frozen records once
  {
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    const size_t NN = (size_t)e->cfg.N;
    const size_t QQ = (size_t)e->Q;
    const size_t s_open = up((size_t)E * NN * QQ * sizeof(double)), s_map = up((size_t)E * B * NN * sizeof(int)),
                 s_frozen = up((size_t)E * B * NN * QQ * sizeof(double));
    // structured populations: pop[nn][N], mark times / branches / populations [K][N], marks in use [N], twice
    const bool st = P.sm.P > 1;
    const size_t KK = st ? (size_t)P.sm.K : 0, nnn = (size_t)P.m.nn;
    const size_t s_pop = st ? up(nnn * NN) : 0, s_mt = st ? up(KK * NN * sizeof(double)) : 0, s_mb = st ? up(KK * NN) : 0,
                 s_nm = st ? up(NN) : 0;
    char* slab = nullptr;
    if ((rc = dev_alloc(e, &slab, 2 * s_open + 2 * s_map + s_frozen + 2 * (s_pop + s_mt + 2 * s_mb + s_nm)))) return rc;
    e->ring_slab = slab;
    e->open_bytes = s_open;
    char* q = slab;
    for (int b = 0; b < 2; b++) { P.buf[b].open = (double*)q; q += s_open; }
    for (int b = 0; b < 2; b++) { P.buf[b].map = (int*)q; q += s_map; }
    P.frozen = (double*)q; q += s_frozen;
    if (st)
      for (int b = 0; b < 2; b++) {
        P.buf[b].mtime = (double*)q; q += s_mt;
        P.buf[b].pop = (signed char*)q; q += s_pop;
        P.buf[b].mbranch = (signed char*)q; q += s_mb;
        P.buf[b].mpop = (signed char*)q; q += s_mb;
        P.buf[b].nmarks = (signed char*)q; q += s_nm;
      }
    if (P.world > 1 && (rc = exchange_slab(e, slab, P.peer_tally))) return rc;
  }
  P.rs_cap = R;
  if ((rc = dev_alloc(e, &P.rs_pos, R)) || (rc = dev_alloc(e, &P.rs_ess, R)) || (rc = dev_alloc(e, &P.row_opp, R))) return rc;
  CUDA_TRY(cudaStreamSynchronize(e->stream));   // host vectors go out of scope
  e->rows_ready = true;
  return 0;
}

int smc_init_particles(smc_engine* e) {
  if (e) e->graph_dirty = true;
  if (!e) return fail(SMC_ERR_ARG, "null engine");
  if (!e->rows_ready) return fail(SMC_ERR_STATE, "upload rows before initialising particles");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  const unsigned init_blocks = (unsigned)((e->cfg.N + AUX_BS - 1) / AUX_BS);
  if (e->P.sm.P > 1)
    smc_init_kernel<true><<<init_blocks, AUX_BS, stree_smem(e->P.m.nn, e->P.m.E, e->P.sm.K, AUX_BS), e->stream>>>(e->P);
  else
    smc_init_kernel<false><<<init_blocks, AUX_BS, tree_smem(e->P.m.nn, e->P.m.E, AUX_BS), e->stream>>>(e->P);
  e->launches++;
  CUDA_TRY(cudaGetLastError());
  e->inited = true;
  return 0;
}

int smc_sweep(smc_engine* e, int64_t row_begin, int64_t row_end) {
  if (!e) return fail(SMC_ERR_ARG, "null engine");
  if (!e->inited) return fail(SMC_ERR_STATE, "particles not initialised");
  if (e->finished) return fail(SMC_ERR_STATE, "sweep after finish");
  if (row_begin != e->swept) return fail(SMC_ERR_STATE, "rows must be swept in order, starting where the last sweep ended");
  if (row_end < row_begin || row_end > e->P.R) return fail(SMC_ERR_ARG, "row range out of bounds");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  CUDA_TRY(cudaEventRecord(e->ev0, e->stream));
  {
    // sharded engines take the same path: the ranks' per-row exchange happens inside the row kernel (peer mailbox)
    int rc = launch_rows(e, row_end - row_begin);
    if (rc) return rc;
  }
  CUDA_TRY(cudaEventRecord(e->ev1, e->stream));
  CUDA_TRY(cudaGetLastError());
  // timing is collected lazily: the events are read at the next synchronising call
  e->swept = row_end;
  cudaEventSynchronize(e->ev1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e->ev0, e->ev1);
  e->kernel_ms += ms;
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int smc_sweep_many(smc_engine** engines, int32_t count, int64_t rows_per_turn) {
  if (!engines || count < 1) return fail(SMC_ERR_ARG, "no engines");
  if (rows_per_turn < 1) rows_per_turn = 1;
  for (int k = 0; k < count; k++) {
    smc_engine* e = engines[k];
    if (!e) return fail(SMC_ERR_ARG, "null engine");
    if (!e->inited) return fail(SMC_ERR_STATE, "particles not initialised");
    if (e->finished) return fail(SMC_ERR_STATE, "sweep after finish");
    if (e->P.world > 1) return fail(SMC_ERR_STATE, "smc_sweep_many takes unsharded engines; sweep a sharded one with smc_sweep on every rank");
  }
  // every engine has its own stream: launches of different chunks interleave and run concurrently on the device
  CUDA_TRY(cudaSetDevice(engines[0]->cfg.device));
  CUDA_TRY(cudaEventRecord(engines[0]->ev0, engines[0]->stream));
  for (int k = 1; k < count; k++) {
    CUDA_TRY(cudaSetDevice(engines[k]->cfg.device));
    if (engines[k]->cfg.device == engines[0]->cfg.device) cudaStreamWaitEvent(engines[k]->stream, engines[0]->ev0, 0);
  }
  bool more = true;
  const auto host_t0 = std::chrono::steady_clock::now();
  while (more) {
    more = false;
    for (int k = 0; k < count; k++) {
      smc_engine* e = engines[k];
      if (e->swept >= e->P.R) continue;
      CUDA_TRY(cudaSetDevice(e->cfg.device));
      // a turn is at least one graph batch; the streams keep the chunks concurrent on the device either way
      int64_t stop = std::min<int64_t>(e->P.R, e->swept + std::max<int64_t>(rows_per_turn, (int64_t)GRAPH_ROWS));
      int rc = launch_rows(e, stop - e->swept);
      if (rc) return rc;
      e->swept = stop;
      if (e->swept < e->P.R) more = true;
    }
  }
  double span = 0.0;
  const double host_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - host_t0).count();
  for (int k = 0; k < count; k++) {
    smc_engine* e = engines[k];
    CUDA_TRY(cudaSetDevice(e->cfg.device));
    CUDA_TRY(cudaEventRecord(e->ev1, e->stream));
  }
  for (int k = 0; k < count; k++) {
    smc_engine* e = engines[k];
    CUDA_TRY(cudaSetDevice(e->cfg.device));
    CUDA_TRY(cudaEventSynchronize(e->ev1));
    if (e->cfg.device == engines[0]->cfg.device) {
      float ms = 0;
      cudaEventElapsedTime(&ms, engines[0]->ev0, e->ev1);
      span = std::max(span, (double)ms);
    }
  }
  engines[0]->kernel_ms += span;     // device span of the whole group, booked on the first engine
  if (getenv("SMC_DEBUG_TIMING")) fprintf(stderr, "[smc] sweep_many: %d engines, host launch loop %.1f ms, device span %.1f ms\n", count, host_ms, span);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

int smc_reset(smc_engine* e, const double* inv2N, double rho, const double* vb_logfac, uint64_t seed) {
  if (e) e->graph_dirty = true;
  if (!e) return fail(SMC_ERR_ARG, "null engine");
  if (!e->rows_ready) return fail(SMC_ERR_STATE, "upload rows first");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  Params& P = e->P;
  const int E = e->cfg.E;
  for (int ep = 0; ep < E; ep++) {
    if (inv2N) { if (!(inv2N[ep] > 0)) return fail(SMC_ERR_ARG, "inv2N must be positive"); P.m.inv2N[ep] = inv2N[ep]; }
    if (vb_logfac) P.m.vb_logfac[ep] = vb_logfac[ep];
  }
  if (rho > 0) P.m.rho = rho;
  P.seed = seed;
  // frozen records and maps are (re)initialised when their bucket is born; only the open cells start from zero
  for (int b = 0; b < 2; b++) CUDA_TRY(cudaMemsetAsync(P.buf[b].open, 0, e->open_bytes, e->stream));
  CUDA_TRY(cudaMemsetAsync(P.tallies, 0, (size_t)e->tally_rows * E * sizeof(double), e->stream));
  CUDA_TRY(cudaMemsetAsync(P.pool, 0, 2 * (size_t)E * P.N * sizeof(double), e->stream));
  CUDA_TRY(cudaMemsetAsync(P.pool2, 0, e->pool2_count * sizeof(double), e->stream));
  CUDA_TRY(cudaMemsetAsync(P.anc, 0xff, (size_t)P.N * sizeof(int), e->stream));
  Ctrl init;
  memset(&init, 0, sizeof(init));
  init.cons_done = -1;
  init.gen = ++e->gen;          // every rank of a sharded population resets the same number of times
  CUDA_TRY(cudaMemcpyAsync(P.ctrl, &init, sizeof(init), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  e->swept = 0; e->inited = false; e->finished = false;
  return 0;
}

#ifdef SMC_CLOCKS
static void clocks_print();
#endif
int smc_finish(smc_engine* e) {
  if (!e) return fail(SMC_ERR_ARG, "null engine");
  if (!e->inited) return fail(SMC_ERR_STATE, "particles not initialised");
  if (e->finished) return 0;
#ifdef SMC_CLOCKS
  cudaDeviceSynchronize();
  clocks_print();
#endif
  if (e->swept != e->P.R) return fail(SMC_ERR_STATE, "finish before the last row was swept");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  // buckets the regular schedule still owes for the last row are folded into the flush: both use the final weights
  std::vector<int> fp(e->cfg.E);
  for (int ep = 0; ep < e->cfg.E; ep++) fp[ep] = 0;
  // number of buckets consumed so far per epoch = those scheduled after rows 0..R-2
  {
    std::vector<int> co(e->P.R + 1);
    CUDA_TRY(cudaMemcpy(co.data(), e->P.cons_off, (e->P.R + 1) * sizeof(int), cudaMemcpyDeviceToHost));
    int upto = co[e->P.R - 1];
    std::vector<int2> cs(std::max(upto, 1));
    if (upto) CUDA_TRY(cudaMemcpy(cs.data(), e->P.cons, upto * sizeof(int2), cudaMemcpyDeviceToHost));
    for (int q = 0; q < upto; q++) fp[cs[q].x] = cs[q].y + 1;
  }
  CUDA_TRY(cudaMemcpyAsync(e->first_pending_d, fp.data(), fp.size() * sizeof(int), cudaMemcpyHostToDevice, e->stream));
  CUDA_TRY(cudaEventRecord(e->ev0, e->stream));
  if (e->P.world == 1 && e->P.pool_on) {       // pooled second moments: first what the last row launch left, then bucket by bucket
    smc_flush_pool_kernel<<<e->P.nblocks, e->BS, 0, e->stream>>>(e->P, e->first_pending_d, 0, 0);
    e->launches++;
    int deepest = 0;
    for (int ep = 0; ep < e->cfg.E; ep++) deepest = std::max(deepest, e->first_pending_h[ep] - fp[ep]);
    for (int j = 0; j < deepest; j++) {
      smc_flush_pool_kernel<<<e->P.nblocks, e->BS, 0, e->stream>>>(e->P, e->first_pending_d, j, 1);
      smc_flush_pool_kernel<<<e->P.nblocks, e->BS, 0, e->stream>>>(e->P, e->first_pending_d, j, 2);
      e->launches += 2;
    }
    smc_pool_reduce_kernel<<<e->P.nblocks, e->BS, 0, e->stream>>>(e->P);
    e->launches++;
  }
  if (e->P.sm.P > 1) smc_flush_struct_kernel<<<e->P.nblocks, e->BS, 0, e->stream>>>(e->P, e->first_pending_d);
  else smc_flush_kernel<<<e->P.nblocks, e->BS, 0, e->stream>>>(e->P, e->first_pending_d);
  e->launches++;
  if (e->P.world > 1) {
    // the ranks' tallies and recombination counts add up (the ln normaliser and the resampling log are already global)
    NCCL_TRY(g_nccl.AllReduce(e->P.tallies, e->P.tallies, (size_t)e->tally_rows * e->cfg.E, ncclDouble, ncclSum, e->comm, e->stream));
    NCCL_TRY(g_nccl.AllReduce(&e->P.ctrl->n_recomb, &e->P.ctrl->n_recomb, 1, ncclUint64, ncclSum, e->comm, e->stream));
  }
  CUDA_TRY(cudaEventRecord(e->ev1, e->stream));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  float ms = 0;
  cudaEventElapsedTime(&ms, e->ev0, e->ev1);
  e->kernel_ms += ms;
  CUDA_TRY(cudaGetLastError());
  e->finished = true;
  return 0;
}

int smc_get_tallies(smc_engine* e, double* out) {
  if (!e || !out) return fail(SMC_ERR_ARG, "null argument");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  CUDA_TRY(cudaMemcpy(out, e->P.tallies, (size_t)e->tally_rows * e->cfg.E * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

static int read_ctrl(smc_engine* e, Ctrl* c) {
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  CUDA_TRY(cudaMemcpy(c, e->P.ctrl, sizeof(Ctrl), cudaMemcpyDeviceToHost));
  return 0;
}

int smc_get_row_opportunity(smc_engine* e, double* out) {
  if (!e || !out) return fail(SMC_ERR_ARG, "null argument");
  if (!e->rows_ready) return fail(SMC_ERR_STATE, "upload rows first");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  CUDA_TRY(cudaMemcpy(out, e->P.row_opp, (size_t)e->P.R * sizeof(double), cudaMemcpyDeviceToHost));
  for (long long r = e->swept; r < e->P.R; r++) out[r] = 0.0;       // rows not swept yet
  return 0;
}

int smc_get_lnZ(smc_engine* e, double* lnZ) {
  if (!e || !lnZ) return fail(SMC_ERR_ARG, "null argument");
  Ctrl c;
  int rc = read_ctrl(e, &c);
  if (rc) return rc;
  *lnZ = c.lnZ;
  // an overflowing mark pool loses population changes, after which a lineage may well be stranded: report the cause
  if (c.sflags & SFLAG_MARKS) return fail(SMC_ERR_NUMERIC, "a particle needed more migration marks than max_marks (smc_set_populations)");
  if (c.sflags & SFLAG_NOCOAL) return fail(SMC_ERR_NUMERIC, "Lines did not coalescence: the populations of the model never join and do not exchange migrants");
  if (c.err) return fail(SMC_ERR_NUMERIC, "Zero or negative probabilities");
  return 0;
}

int smc_get_counters(smc_engine* e, int64_t counters[4]) {
  if (!e || !counters) return fail(SMC_ERR_ARG, "null argument");
  Ctrl c;
  int rc = read_ctrl(e, &c);
  if (rc) return rc;
  counters[0] = c.row; counters[1] = c.n_resample; counters[2] = c.n_recomb; counters[3] = c.err;
  return 0;
}

int smc_get_resample_log(smc_engine* e, double* position, double* ess, int64_t cap, int64_t* n) {
  if (!e || !n) return fail(SMC_ERR_ARG, "null argument");
  Ctrl c;
  int rc = read_ctrl(e, &c);
  if (rc) return rc;
  int64_t have = std::min<int64_t>(c.n_resample, e->P.rs_cap);
  *n = have;
  int64_t take = std::min(have, cap);
  if (take > 0 && position) CUDA_TRY(cudaMemcpy(position, e->P.rs_pos, take * sizeof(double), cudaMemcpyDeviceToHost));
  if (take > 0 && ess) CUDA_TRY(cudaMemcpy(ess, e->P.rs_ess, take * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

int smc_get_particles(smc_engine* e, int32_t* parent, double* height, double* logw, double* next_base) {
  if (!e) return fail(SMC_ERR_ARG, "null engine");
  if (!e->inited) return fail(SMC_ERR_STATE, "particles not initialised");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  if (!e->finished) { int rc = materialise(e); if (rc) return rc; }
  Ctrl c;
  int rc = read_ctrl(e, &c);
  if (rc) return rc;
  const size_t N = e->cfg.N;
  const int n = e->cfg.n, nn = 2 * n - 1;
  const StateBuf& s = e->P.buf[c.cur];
  std::vector<double> h((size_t)(n - 1) * N);
  std::vector<signed char> p((size_t)nn * N);
  CUDA_TRY(cudaMemcpy(h.data(), s.height, h.size() * sizeof(double), cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(p.data(), s.parent, p.size(), cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < N; i++)
    for (int v = 0; v < nn; v++) {
      if (height) height[i * nn + v] = v < n ? 0.0 : h[(size_t)(v - n) * N + i];
      if (parent) parent[i * nn + v] = p[(size_t)v * N + i];
    }
  if (logw) {
    CUDA_TRY(cudaMemcpy(logw, s.logw, N * sizeof(double), cudaMemcpyDeviceToHost));
    if (e->finished) for (size_t i = 0; i < N; i++) logw[i] -= c.lse;
  }
  if (next_base) CUDA_TRY(cudaMemcpy(next_base, s.next_base, N * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

int smc_set_particles(smc_engine* e, const int32_t* parent, const double* height, const double* logw,
                      const double* next_base) {
  if (e) e->graph_dirty = true;
  if (!e || !parent || !height || !logw || !next_base) return fail(SMC_ERR_ARG, "null argument");
  if (!e->rows_ready) return fail(SMC_ERR_STATE, "upload rows first");
  if (e->swept != 0) return fail(SMC_ERR_STATE, "set_particles only before the first row");
  if (e->P.sm.P > 1) return fail(SMC_ERR_STATE, "set_particles is not available with structured populations");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  const size_t N = e->cfg.N;
  const int n = e->cfg.n, nn = 2 * n - 1;
  std::vector<double> h((size_t)(n - 1) * N);
  std::vector<signed char> p((size_t)nn * N);
  for (size_t i = 0; i < N; i++) {
    for (int v = 0; v < nn; v++) {
      int par = parent[i * nn + v];
      if (par < -1 || par >= nn) return fail(SMC_ERR_ARG, "parent index out of range");
      p[(size_t)v * N + i] = (signed char)par;
      if (v >= n) {
        h[(size_t)(v - n) * N + i] = height[i * nn + v];
        if (v > n && height[i * nn + v] < height[i * nn + v - 1]) return fail(SMC_ERR_ARG, "internal slots must be height-sorted");
      } else if (height[i * nn + v] != 0.0) return fail(SMC_ERR_ARG, "leaves must sit at height 0");
    }
    if (parent[i * nn + nn - 1] != -1) return fail(SMC_ERR_ARG, "the last slot must be the root");
  }
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  const StateBuf& s = e->P.buf[0];
  CUDA_TRY(cudaMemcpy(s.height, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(s.parent, p.data(), p.size(), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(s.logw, logw, N * sizeof(double), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(s.next_base, next_base, N * sizeof(double), cudaMemcpyHostToDevice));
  e->inited = true;
  return 0;
}

int smc_get_last_ancestry(smc_engine* e, int32_t* anc) {
  if (!e || !anc) return fail(SMC_ERR_ARG, "null argument");
  CUDA_TRY(cudaSetDevice(e->cfg.device));
  if (e->inited && !e->finished) { int rc = materialise(e); if (rc) return rc; }
  CUDA_TRY(cudaStreamSynchronize(e->stream));
  CUDA_TRY(cudaMemcpy(anc, e->P.anc, (size_t)e->cfg.N * sizeof(int), cudaMemcpyDeviceToHost));
  return 0;
}

int smc_get_timing(smc_engine* e, double* kernel_ms, int64_t* launches) {
  if (!e) return fail(SMC_ERR_ARG, "null engine");
  if (kernel_ms) *kernel_ms = e->kernel_ms;
  if (launches) *launches = e->launches;
  return 0;
}

// ---- stand-alone entry points -----------------------------------------------------------------------------------
namespace {
struct DevBuf {
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  int alloc(size_t bytes) { CUDA_TRY(cudaMalloc(&p, std::max<size_t>(bytes, 8))); return 0; }
};

int tree_fn(int device, int n, int64_t count, const int32_t* parent, const double* height, const int8_t* alleles,
            double mu, int aa, int what, double* o0, double* o1) {
  if (!parent || !height || !alleles || !o0) return fail(SMC_ERR_ARG, "null argument");
  if (count < 0) return fail(SMC_ERR_ARG, "negative count");
  if (count == 0) return 0;
  int rc = use_device(device);
  if (rc) return rc;
  ModelDev m;
  double t0 = 0.0, inv = 1.0;
  if ((rc = fill_model(m, n, 1, mu, 1.0, 1.0, aa, &t0, &inv, nullptr, nullptr))) return rc;
  const int nn = 2 * n - 1, BS = 64;
  for (int64_t i = 0; i < count; i++)
    for (int v = 0; v < n; v++)
      if (height[i * nn + v] != 0.0) return fail(SMC_ERR_ARG, "leaves must sit at height 0");
  DevBuf dp, dh, da, d0, d1;
  if ((rc = dp.alloc(count * nn * sizeof(int))) || (rc = dh.alloc(count * nn * sizeof(double))) ||
      (rc = da.alloc(count * n)) || (rc = d0.alloc(count * sizeof(double))) || (rc = d1.alloc(count * sizeof(double))))
    return rc;
  CUDA_TRY(cudaMemcpy(dp.p, parent, count * nn * sizeof(int), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(dh.p, height, count * nn * sizeof(double), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(da.p, alleles, count * n, cudaMemcpyHostToDevice));
  size_t smem = tree_smem(nn, 1, BS);
  CUDA_TRY(cudaFuncSetAttribute(k_tree_functions, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM));
  k_tree_functions<<<(unsigned)((count + BS - 1) / BS), BS, smem>>>(m, count, (const int*)dp.p, (const double*)dh.p,
                                                                   (const signed char*)da.p, what, (double*)d0.p, (double*)d1.p);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpy(o0, d0.p, count * sizeof(double), cudaMemcpyDeviceToHost));
  if (o1) CUDA_TRY(cudaMemcpy(o1, d1.p, count * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}
}  // namespace

int smc_site_likelihood(int device, int32_t n, int64_t count, const int32_t* parent, const double* height,
                        const int8_t* alleles, double mu, int32_t ancestral_aware, double* lik) {
  return tree_fn(device, n, count, parent, height, alleles, mu, ancestral_aware, 0, lik, nullptr);
}

int smc_branch_lengths(int device, int32_t n, int64_t count, const int32_t* parent, const double* height,
                       const int8_t* alleles, double* with_data, double* total) {
  if (!total) return fail(SMC_ERR_ARG, "null argument");
  return tree_fn(device, n, count, parent, height, alleles, 0.0, 0, 1, with_data, total);
}

static int offspring_from_psum(const double* d_psum, int64_t N, double u, int32_t* counts) {
  DevBuf dc, dn;
  int rc;
  if ((rc = dc.alloc(sizeof(Ctrl))) || (rc = dn.alloc(N * sizeof(int)))) return rc;
  CUDA_TRY(cudaMemset(dc.p, 0, sizeof(Ctrl)));
  k_build_chain<<<1, 1>>>((Ctrl*)dc.p, u, N);
  k_offspring_counts<<<(unsigned)((N + 127) / 128), 128>>>((const Ctrl*)dc.p, d_psum, N, (int*)dn.p);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpy(counts, dn.p, N * sizeof(int), cudaMemcpyDeviceToHost));
  return 0;
}

int smc_systematic_offspring(int device, const double* partial_sum, int64_t N, double u, int32_t* counts) {
  if (!partial_sum || !counts) return fail(SMC_ERR_ARG, "null argument");
  if (N < 1) return fail(SMC_ERR_ARG, "N must be >= 1");
  if (!(u >= 0.0 && u < 1.0)) return fail(SMC_ERR_ARG, "u must be in [0,1)");
  int rc = use_device(device);
  if (rc) return rc;
  DevBuf dp;
  if ((rc = dp.alloc((N + 1) * sizeof(double)))) return rc;
  CUDA_TRY(cudaMemcpy(dp.p, partial_sum, (N + 1) * sizeof(double), cudaMemcpyHostToDevice));
  return offspring_from_psum((const double*)dp.p, N, u, counts);
}

int smc_resample_counts(int device, const double* w, int64_t N, double u, int32_t parity_scan, double* ess,
                        int32_t* counts) {
  if (!w || !counts || !ess) return fail(SMC_ERR_ARG, "null argument");
  if (N < 1) return fail(SMC_ERR_ARG, "N must be >= 1");
  if (!(u >= 0.0 && u < 1.0)) return fail(SMC_ERR_ARG, "u must be in [0,1)");
  int rc = use_device(device);
  if (rc) return rc;
  DevBuf dw, dp, de;
  if ((rc = dw.alloc(N * sizeof(double))) || (rc = dp.alloc((N + 1) * sizeof(double))) || (rc = de.alloc(sizeof(double))))
    return rc;
  CUDA_TRY(cudaMemcpy(dw.p, w, N * sizeof(double), cudaMemcpyHostToDevice));
  if (parity_scan) k_seq_partial_sums<<<1, 1>>>((const double*)dw.p, N, (double*)dp.p, (double*)de.p);
  else k_par_partial_sums<<<1, 1024>>>((const double*)dw.p, N, (double*)dp.p, (double*)de.p);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpy(ess, de.p, sizeof(double), cudaMemcpyDeviceToHost));
  return offspring_from_psum((const double*)dp.p, N, u, counts);
}

int smc_normalize_log(int device, double* logw, int64_t N, double* lnZ, double* ess) {
  if (!logw || !lnZ || !ess) return fail(SMC_ERR_ARG, "null argument");
  if (N < 1) return fail(SMC_ERR_ARG, "N must be >= 1");
  int rc = use_device(device);
  if (rc) return rc;
  DevBuf dw, dout;
  if ((rc = dw.alloc(N * sizeof(double))) || (rc = dout.alloc(2 * sizeof(double)))) return rc;
  CUDA_TRY(cudaMemcpy(dw.p, logw, N * sizeof(double), cudaMemcpyHostToDevice));
  k_normalize_log<<<1, 1024>>>((double*)dw.p, N, (double*)dout.p);
  CUDA_TRY(cudaGetLastError());
  double o[2];
  CUDA_TRY(cudaMemcpy(o, dout.p, sizeof(o), cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(logw, dw.p, N * sizeof(double), cudaMemcpyDeviceToHost));
  *lnZ = o[0]; *ess = o[1];
  if (!std::isfinite(o[0])) return fail(SMC_ERR_NUMERIC, "Zero or negative probabilities");
  return 0;
}

// shared by the two calibration entry points; npop <= 1: single population (inv2N[E]), else structured tables
static int calibrate_impl(int device, int32_t n, int32_t E, const double* t_start, int32_t npop, const double* inv2N,
                          const double* mig, const double* single_mig, const int32_t* leaf_pop, double rho, double seqlen,
                          uint64_t seed, int64_t n_args, double* median_survival, int64_t* n_events,
                          int32_t nq = 0, const double* quantiles = nullptr, double* tbl = nullptr, double* mean_tl = nullptr) {
  const bool terminal = quantiles != nullptr;      // terminal branch length quantiles instead of survival distances
  if (!t_start || !inv2N || (!terminal && !median_survival) || (terminal && (!tbl || nq < 1))) return fail(SMC_ERR_ARG, "null argument");
  if (n_args < 1) return fail(SMC_ERR_ARG, "n_args must be >= 1");
  int rc = use_device(device);
  if (rc) return rc;
  const bool st = npop > 1;
  if (st && (npop > SMC_MAX_POPULATIONS || !leaf_pop)) return fail(SMC_ERR_ARG, "populations must be in 2..4 with leaf_pop given");
  ModelDev m;
  std::vector<double> inv0(E);
  for (int e = 0; e < E; e++) inv0[e] = st ? inv2N[(size_t)e * npop] : inv2N[e];
  if ((rc = fill_model(m, n, E, 0.0, rho, seqlen, 0, t_start, inv0.data(), nullptr, nullptr))) return rc;
  const int BS = AUX_BS, per = terminal ? n : n - 1;
  DevBuf dd, de, df, t_inv, t_mig, t_tot, t_single, t_fixed, t_leaf;
  if ((rc = dd.alloc(n_args * per * sizeof(double))) || (rc = de.alloc(n_args * (terminal ? 2 : per) * sizeof(int))) || (rc = df.alloc(sizeof(int))))
    return rc;
  CUDA_TRY(cudaMemset(df.p, 0, sizeof(int)));
  SModelDev sm;
  memset(&sm, 0, sizeof(sm));
  sm.P = 1;
  size_t smem = tree_smem(m.nn, E, BS);
  if (st) {
    const int Pp = npop;
    std::vector<double> h_inv((size_t)E * Pp), h_mig((size_t)E * Pp * Pp, 0.0), h_tot((size_t)E * Pp, 0.0), h_single((size_t)E * Pp * Pp, 0.0);
    std::vector<unsigned char> h_fixed(E, 0);
    std::vector<signed char> h_leaf(n);
    for (int ep = 0; ep < E; ep++)
      for (int i = 0; i < Pp; i++) {
        h_inv[(size_t)ep * Pp + i] = inv2N[(size_t)ep * Pp + i];
        if (!(h_inv[(size_t)ep * Pp + i] > 0)) return fail(SMC_ERR_ARG, "inv2N must be positive");
        double tot = 0.0;
        for (int j = 0; j < Pp; j++) {
          const size_t at = ((size_t)ep * Pp + i) * Pp + j;
          double mr = (mig && j != i) ? mig[at] : 0.0;
          double s1 = single_mig ? single_mig[at] : 0.0;
          if (!(mr >= 0)) return fail(SMC_ERR_ARG, "migration rates must be non-negative");
          if (!(s1 >= 0 && s1 <= 1)) return fail(SMC_ERR_ARG, "single_mig entries are probabilities");
          h_mig[at] = mr; h_single[at] = s1;
          if (j != i) tot += mr;
          if (s1 > 0) h_fixed[ep] = 1;
        }
        h_tot[(size_t)ep * Pp + i] = tot;
      }
    for (int i = 0; i < n; i++) {
      if (leaf_pop[i] < 0 || leaf_pop[i] >= Pp) return fail(SMC_ERR_ARG, "leaf population out of range");
      h_leaf[i] = (signed char)leaf_pop[i];
    }
    auto up = [&](DevBuf& d, const void* src, size_t bytes) -> int {
      int r2 = d.alloc(bytes);
      if (r2) return r2;
      CUDA_TRY(cudaMemcpy(d.p, src, bytes, cudaMemcpyHostToDevice));
      return 0;
    };
    if ((rc = up(t_inv, h_inv.data(), h_inv.size() * 8)) || (rc = up(t_mig, h_mig.data(), h_mig.size() * 8)) ||
        (rc = up(t_tot, h_tot.data(), h_tot.size() * 8)) || (rc = up(t_single, h_single.data(), h_single.size() * 8)) ||
        (rc = up(t_fixed, h_fixed.data(), h_fixed.size())) || (rc = up(t_leaf, h_leaf.data(), h_leaf.size())))
      return rc;
    sm.P = Pp; sm.K = 120; sm.Q = 2 + 3 * Pp + Pp * Pp;
    sm.inv2N = (const double*)t_inv.p; sm.mig = (const double*)t_mig.p; sm.migtot = (const double*)t_tot.p;
    sm.single_mig = (const double*)t_single.p; sm.fixed = (const unsigned char*)t_fixed.p;
    sm.leaf_pop = (const signed char*)t_leaf.p;
    sm.vb_coal = nullptr; sm.vb_mig = nullptr;
    smem = stree_smem(m.nn, E, sm.K, BS);
    if (smem > SMC_MAX_DYN_SMEM) { sm.K = 2 * n + 8; smem = stree_smem(m.nn, E, sm.K, BS); }
    if (smem > SMC_MAX_DYN_SMEM) return fail(SMC_ERR_ARG, "n and E need more shared memory than one SM has");
    CUDA_TRY(cudaFuncSetAttribute(k_survival<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM));
    CUDA_TRY(cudaFuncSetAttribute(k_terminal<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM));
    if (terminal) k_terminal<true><<<(unsigned)((n_args + BS - 1) / BS), BS, smem>>>(m, sm, n_args, seed, (double*)dd.p, (double*)de.p, (int*)df.p);
    else k_survival<true><<<(unsigned)((n_args + BS - 1) / BS), BS, smem>>>(m, sm, n_args, seed, (double*)dd.p, (int*)de.p, (int*)df.p);
  } else {
    CUDA_TRY(cudaFuncSetAttribute(k_survival<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM));
    CUDA_TRY(cudaFuncSetAttribute(k_terminal<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMC_MAX_DYN_SMEM));
    if (terminal) k_terminal<false><<<(unsigned)((n_args + BS - 1) / BS), BS, smem>>>(m, sm, n_args, seed, (double*)dd.p, (double*)de.p, (int*)df.p);
    else k_survival<false><<<(unsigned)((n_args + BS - 1) / BS), BS, smem>>>(m, sm, n_args, seed, (double*)dd.p, (int*)de.p, (int*)df.p);
  }
  CUDA_TRY(cudaGetLastError());
  if (terminal) {
    std::vector<double> d(n_args * per), tl(n_args);
    int fl = 0;
    CUDA_TRY(cudaMemcpy(d.data(), dd.p, d.size() * sizeof(double), cudaMemcpyDeviceToHost));
    CUDA_TRY(cudaMemcpy(tl.data(), de.p, tl.size() * sizeof(double), cudaMemcpyDeviceToHost));
    CUDA_TRY(cudaMemcpy(&fl, df.p, sizeof(int), cudaMemcpyDeviceToHost));
    if (fl & SFLAG_NOCOAL) return fail(SMC_ERR_NUMERIC, "Lines did not coalescence: the populations of the model never join and do not exchange migrants");
    for (int v = 0; v < n; v++) {
      std::sort(d.begin() + (size_t)v * n_args, d.begin() + (size_t)(v + 1) * n_args);
      for (int k = 0; k < nq; k++) {
        if (!(quantiles[k] >= 0.0 && quantiles[k] < 1.0)) return fail(SMC_ERR_ARG, "quantiles must be in [0,1)");
        tbl[(size_t)v * nq + k] = d[(size_t)v * n_args + (size_t)(quantiles[k] * (double)n_args)];   // smcsmc.cpp:158
      }
    }
    double sum = 0.0;
    for (double x : tl) sum += x;
    if (mean_tl) *mean_tl = sum / (double)n_args;
    return 0;
  }
  std::vector<double> dist(n_args * per);
  std::vector<int> ep(n_args * per);
  int flags = 0;
  CUDA_TRY(cudaMemcpy(dist.data(), dd.p, dist.size() * sizeof(double), cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(ep.data(), de.p, ep.size() * sizeof(int), cudaMemcpyDeviceToHost));
  CUDA_TRY(cudaMemcpy(&flags, df.p, sizeof(int), cudaMemcpyDeviceToHost));
  if (flags & SFLAG_NOCOAL) return fail(SMC_ERR_NUMERIC, "Lines did not coalescence: the populations of the model never join and do not exchange migrants");
  std::vector<std::vector<double>> by(E);
  for (size_t k = 0; k < dist.size(); k++) if (dist[k] >= 0.0) by[ep[k]].push_back(dist[k]);
  // medians with the reference's fall-backs for sparsely sampled epochs (smcsmc.cpp:238-262)
  const int max_safe_sample_size = 10;
  double earliest = -1;
  for (int e = 0; e < E; e++) {
    std::sort(by[e].begin(), by[e].end());
    if (n_events) n_events[e] = (int64_t)by[e].size();
    int median_idx = by[e].empty() ? -1 : (int)((by[e].size() - 1) / 2);
    if (median_idx < max_safe_sample_size) median_survival[e] = -1;
    else { median_survival[e] = by[e][median_idx]; if (earliest < 0) earliest = median_survival[e]; }
  }
  for (int e = 0; e < E; e++)
    if (median_survival[e] < 0) median_survival[e] = e > 0 ? median_survival[e - 1] : earliest;
  if (median_survival[0] < 0) return fail(SMC_ERR_NUMERIC, "no node survival distances could be sampled; check the model");
  return 0;
}

int smc_calibrate_survival(int device, int32_t n, int32_t E, const double* t_start, const double* inv2N, double rho,
                           double seqlen, uint64_t seed, int64_t n_args, double* median_survival, int64_t* n_events) {
  return calibrate_impl(device, n, E, t_start, 1, inv2N, nullptr, nullptr, nullptr, rho, seqlen, seed, n_args,
                        median_survival, n_events);
}

int smc_calibrate_survival_struct(int device, int32_t n, int32_t E, const double* t_start, int32_t npop,
                                  const double* inv2N, const double* mig, const double* single_mig,
                                  const int32_t* leaf_pop, double rho, double seqlen, uint64_t seed, int64_t n_args,
                                  double* median_survival, int64_t* n_events) {
  if (npop < 2) return fail(SMC_ERR_ARG, "populations must be in 2..4");
  return calibrate_impl(device, n, E, t_start, npop, inv2N, mig, single_mig, leaf_pop, rho, seqlen, seed, n_args,
                        median_survival, n_events);
}

int smc_fastexp(int device, const double* x, int64_t count, double* out) {
  if (!x || !out) return fail(SMC_ERR_ARG, "null argument");
  if (count <= 0) return count == 0 ? 0 : fail(SMC_ERR_ARG, "negative count");
  int rc = use_device(device);
  if (rc) return rc;
  DevBuf dx, dout;
  if ((rc = dx.alloc(count * sizeof(double))) || (rc = dout.alloc(count * sizeof(double)))) return rc;
  CUDA_TRY(cudaMemcpy(dx.p, x, count * sizeof(double), cudaMemcpyHostToDevice));
  k_fastexp<<<(unsigned)((count + 255) / 256), 256>>>((const double*)dx.p, count, (double*)dout.p);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpy(out, dout.p, count * sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}

}  // extern "C"

// ---- allocation-free variant of smc_resample_counts (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving loops) ------------
// The scratch smc_resample_counts takes from cudaMalloc (the u-chain control block and the N+1 partial sums) lives in a
// file-scope device array instead.  Every GPU that loads libsmcgpu.so holds these SMC_NOALLOC_SCRATCH_BYTES until exit.
// In host code the array's name is a host-side stand-in, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, asked once per device with that device current (noalloc_scratch below).
constexpr size_t NOALLOC_CTRL_BYTES = (sizeof(Ctrl) + 255) / 256 * 256;
constexpr size_t NOALLOC_SCRATCH_BYTES = NOALLOC_CTRL_BYTES + ((size_t)SMC_NOALLOC_MAX_N + 1) * sizeof(double);
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mutex;
void* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

// device address of g_noalloc_scratch on `device`, which must be the current device
int noalloc_scratch(int device, unsigned char** out) {
  if (device >= NOALLOC_MAX_DEVICES) return fail(SMC_ERR_ARG, "device ordinal beyond the no-alloc scratch table (64)");
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_ptr[device]) {
    void* ptr = nullptr;
    CUDA_TRY(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
    if (!ptr) return fail(SMC_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the no-alloc scratch");
    g_noalloc_ptr[device] = ptr;
  }
  *out = (unsigned char*)g_noalloc_ptr[device];
  return 0;
}
}  // namespace

extern "C" int smc_resample_counts_noalloc(int device, const double* d_w, int64_t N, double u, int32_t parity_scan,
                                           double* d_ess, int32_t* d_counts, void* stream) {
  if (!d_w || !d_counts || !d_ess) return fail(SMC_ERR_ARG, "null argument");
  if (N < 1) return fail(SMC_ERR_ARG, "N must be >= 1");
  if (N > SMC_NOALLOC_MAX_N)
    return fail(SMC_ERR_ARG, "smc_resample_counts_noalloc: N = " + std::to_string((long long)N) +
                                 " exceeds the static scratch capacity SMC_NOALLOC_MAX_N = " +
                                 std::to_string((long long)SMC_NOALLOC_MAX_N) + " (use smc_resample_counts)");
  if (!(u >= 0.0 && u < 1.0)) return fail(SMC_ERR_ARG, "u must be in [0,1)");
  int rc = use_device(device);
  if (rc) return rc;
  unsigned char* scratch = nullptr;
  if ((rc = noalloc_scratch(device, &scratch))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  Ctrl* ctrl = (Ctrl*)scratch;
  double* psum = (double*)(scratch + NOALLOC_CTRL_BYTES);
  // the array keeps the previous call's contents: the control block must start at zero, the partial sums are
  // overwritten in full ([0, N]) before they are read
  CUDA_TRY(cudaMemsetAsync(ctrl, 0, sizeof(Ctrl), st));
  if (parity_scan) k_seq_partial_sums<<<1, 1, 0, st>>>(d_w, N, psum, d_ess);
  else k_par_partial_sums<<<1, 1024, 0, st>>>(d_w, N, psum, d_ess);
  k_build_chain<<<1, 1, 0, st>>>(ctrl, u, N);
  k_offspring_counts<<<(unsigned)((N + 127) / 128), 128, 0, st>>>(ctrl, psum, N, d_counts);
  CUDA_TRY(cudaGetLastError());
  return 0;
}

extern "C" int smc_terminal_branch_quantiles(int device, int32_t n, int32_t E, const double* t_start, int32_t npop,
                                             const double* inv2N, const double* mig, const double* single_mig,
                                             const int32_t* leaf_pop, uint64_t seed, int64_t n_trees, int32_t nq,
                                             const double* quantiles, double* lengths, double* mean_total_length) {
  if (!quantiles || !lengths) return fail(SMC_ERR_ARG, "null argument");
  return calibrate_impl(device, n, E, t_start, npop < 2 ? 1 : npop, inv2N, mig, single_mig, leaf_pop, 1e-8, 1.0, seed, n_trees,
                        nullptr, nullptr, nq, quantiles, lengths, mean_total_length);
}

// ---- measurement helper: the fp64 FMA rate of the device (bench.py reports the row kernel against it) -------------------
namespace {
__global__ void __launch_bounds__(256) k_dfma_probe(double* out, int iters, double a, double b) {
  // eight independent chains per thread: enough to cover the DFMA latency at 8 resident warps per scheduler
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
    x0 = __fma_rn(x0, a, b); x1 = __fma_rn(x1, a, b); x2 = __fma_rn(x2, a, b); x3 = __fma_rn(x3, a, b);
    x4 = __fma_rn(x4, a, b); x5 = __fma_rn(x5, a, b); x6 = __fma_rn(x6, a, b); x7 = __fma_rn(x7, a, b);
  }
  const double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  if (s == 12345.678) out[0] = s;      // never true: keeps the chains alive
}
}  // namespace

extern "C" int smc_measure_fp64_fma(int device, double* tflops) {
  if (!tflops) return fail(SMC_ERR_ARG, "null argument");
  int rc = use_device(device);
  if (rc) return rc;
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  double* d = nullptr;
  CUDA_TRY(cudaMalloc(&d, sizeof(double)));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int blocks = prop.multiProcessorCount * 8, iters = 1 << 15;
  k_dfma_probe<<<blocks, 256>>>(d, 1 << 10, 0.999999, 1e-9);           // warm-up
  cudaEventRecord(e0);
  k_dfma_probe<<<blocks, 256>>>(d, iters, 0.999999, 1e-9);
  cudaEventRecord(e1);
  cudaError_t ce = cudaEventSynchronize(e1);
  float ms = 0.f;
  if (ce == cudaSuccess) ce = cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
  if (ce != cudaSuccess) return fail(SMC_ERR_CUDA, cudaGetErrorString(ce));
  *tflops = 2.0 * 8.0 * (double)iters * 256.0 * (double)blocks / ((double)ms * 1e-3) * 1e-12;
  return 0;
}

#ifdef SMC_CLOCKS
// SMC_CLOCKS_PRINT=1: every smc_finish prints the accumulated per-phase cycles (tools/struct_time.sh with the clk build)
static void clocks_print() {
  if (!getenv("SMC_CLOCKS_PRINT")) return;
  unsigned long long v[16];
  if (cudaMemcpyFromSymbol(v, smc::g_clk, sizeof(v)) != cudaSuccess) return;
  const double blocks = (double)v[10], tot = (double)(v[0] + v[1] + v[2] + v[6]);
  fprintf(stderr, "[clk] blocks %.0f cycles/block %.0f rounds/block %.2f tasks/round %.1f cycles/round %.0f\n", blocks, tot / blocks,
          v[8] / blocks, (double)v[9] / (double)(v[8] ? v[8] : 1), (double)v[4] / (double)(v[8] ? v[8] : 1));
  const char* nm[16] = {"phase0 due tallies", "phase1 load/pull", "phase2 total", " first stretch", " move rounds", " flush+site lik",
                        "phase3 store+partials", "", "", "", "", "  move", "  tree length", "  next base+factor", "  rec opp", ""};
  fprintf(stderr, "[clk] structured: float %.0f cycles/block (%.1f %%), %.2f intervals per round-0-thread float, %.0f cycles per interval\n",
          v[7] / blocks, 100.0 * v[7] / tot, (double)v[15] / (double)(v[8] ? v[8] : 1), (double)v[7] / (double)(v[15] ? v[15] : 1));
  for (int k : {0, 1, 2, 3, 4, 11, 12, 13, 14, 5, 6}) fprintf(stderr, "[clk] %-24s %10.0f cycles/block %5.1f %%\n", nm[k], v[k] / blocks, 100.0 * v[k] / tot);
}
extern "C" int smc_debug_clocks(unsigned long long out[16], int reset) {
  CUDA_TRY(cudaMemcpyFromSymbol(out, smc::g_clk, sizeof(unsigned long long) * 16));
  if (reset) { unsigned long long z[16] = {0}; CUDA_TRY(cudaMemcpyToSymbol(smc::g_clk, z, sizeof(z))); }
  return 0;
}
#endif
