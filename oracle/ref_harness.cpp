// TEST INFRASTRUCTURE ONLY -- never linked or loaded by the product path.
//
// Function-level oracle built FROM THE UNMODIFIED REFERENCE SOURCES (luntergroup/smcsmc, /root/reference/src).
// This file contains no restated algorithm: it only opens up the reference's classes (private -> public) and calls
// the reference's own functions on caller-supplied flat trees / weights, so that oracle/smc_oracle.c (the C
// restatement) and the CUDA kernels can be pinned against the real code.  Built by oracle/build_ref.sh into
// oracle/_ref/libsmcref.so; tests generate the committed fixtures in tests/golden/ from it (tests/golden/make_golden.py).
//
// Reference entry points used (file:line in /root/reference/src):
//   ForestState::calculate_likelihood            particle.cpp:667
//   ForestState::trackLocalTreeBranchLength      particle.cpp:699
//   Forest::getLocalTreeLength                   scrm/src/forest.h:188
//   ParticleContainer::systematic_resampling     particleContainer.cpp:474
//   ParticleContainer::resample                  particleContainer.cpp:247
//   ParticleContainer::normalize_probability     particleContainer.cpp:420
//   ParticleContainer::calculate_initial_haplotype_configuration / next_haplotype   particleContainer.cpp:138,163
//   Forest::buildInitialTree / sampleNextGenealogy / sampleRecSeqPosition   scrm/src/forest.cc:299,423,474
//   CountModel::extract_and_update_count         count.cpp:355
//   ForestState copy-ctor / resample_recombination_position / record_recomb_extension   particle.cpp:113,393,305
#include <vector>
#include <string>
#include <queue>
#include <deque>
#include <map>
#include <set>
#include <valarray>
#include <sstream>
#include <iostream>
#include <fstream>
#include <memory>
#include <random>
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <algorithm>
#include <unordered_map>
#include <unordered_set>
#include <iomanip>
#include <climits>
#include <cfloat>
#include <cassert>
#include <functional>
#include <list>
#include <ctime>
#include <zlib.h>

#define private public
#define protected public
#include "particleContainer.hpp"
#include "count.hpp"
#include "pfparam.hpp"
#undef private
#undef protected

class Arena* Arena::globalArena;   // defined in the reference's main file (smcsmc.cpp:38), which is not linked here

// A MersenneTwister whose uniforms can be scripted (sample() is virtual, random_generator.h:50).
class ScriptedRNG : public MersenneTwister {
 public:
  ScriptedRNG(size_t seed) : MersenneTwister(true, seed) {}
  std::deque<double> script;
  double sample() override {
    if (!script.empty()) { double u = script.front(); script.pop_front(); return u; }
    return MersenneTwister::sample();
  }
};

struct RefHandle {
  PfParam* pf;
  ScriptedRNG* rng;
  Forest* walker;       // persistent forest for prior-ARG walks
  std::string err;
};

static thread_local std::string g_err;

extern "C" const char* ref_last_error() { return g_err.c_str(); }

extern "C" void* ref_open(int argc, const char** argv) {
  try {
    RefHandle* h = new RefHandle();
    h->pf = new PfParam();
    std::vector<char*> av;
    for (int i = 0; i < argc; i++) av.push_back(const_cast<char*>(argv[i]));
    h->pf->parse(argc, av.data());
    h->rng = new ScriptedRNG(h->pf->rg->seed());
    h->walker = NULL;
    if (Arena::globalArena == NULL) new Arena(h->pf->model.getNumEpochs(), 1000000);   // smcsmc.cpp:73
    return h;
  } catch (std::exception& e) {
    g_err = e.what();
    return NULL;
  }
}

extern "C" void ref_close(void* hv) {
  RefHandle* h = (RefHandle*)hv;
  if (!h) return;
  delete h->walker;
  delete h;   // PfParam owns model/rg/Segfile; leak the rest (test tool)
}

extern "C" int ref_num_epochs(void* hv) { return (int)((RefHandle*)hv)->pf->model.change_times_.size(); }
extern "C" double ref_mutation_rate(void* hv) { return ((RefHandle*)hv)->pf->model.mutation_rate(); }
extern "C" double ref_recombination_rate(void* hv) { return ((RefHandle*)hv)->pf->model.recombination_rate(0); }
extern "C" double ref_max_seg_len(void* hv) { RefHandle* h = (RefHandle*)hv; return (int)(h->pf->max_segment_length_factor_ / (h->pf->model.recombination_rate() * 4 * h->pf->model.default_pop_size())); }
extern "C" void ref_epoch_table(void* hv, double* t_start, double* inv2N /*[E][P]*/) {
  Model& m = ((RefHandle*)hv)->pf->model;
  size_t E = m.change_times_.size(), P = m.population_number();
  m.resetTime();
  for (size_t e = 0; e < E; e++) {
    if (e > 0) m.increaseTime();
    t_start[e] = m.change_times_[e];
    for (size_t p = 0; p < P; p++) inv2N[e * P + p] = 1.0 / (2.0 * m.population_size(p));
  }
  m.resetTime();
}

// Structured models: populations, sample populations, migration rates and fixed-time moves per epoch.
extern "C" int ref_num_pops(void* hv) { return (int)((RefHandle*)hv)->pf->model.population_number(); }
extern "C" void ref_sample_pops(void* hv, int* pops) {
  Model& m = ((RefHandle*)hv)->pf->model;
  for (size_t i = 0; i < m.sample_size(); i++) pops[i] = (int)m.sample_population(i);
}
extern "C" void ref_mig_table(void* hv, double* mig /*[E][P][P]*/, double* single /*[E][P][P]*/) {
  Model& m = ((RefHandle*)hv)->pf->model;
  size_t E = m.change_times_.size(), P = m.population_number();
  m.resetTime();
  for (size_t e = 0; e < E; e++) {
    if (e > 0) m.increaseTime();
    for (size_t i = 0; i < P; i++)
      for (size_t j = 0; j < P; j++) {
        mig[(e * P + i) * P + j] = i == j ? 0.0 : m.migration_rate(i, j);
        single[(e * P + i) * P + j] = i == j ? 0.0 : m.single_mig_pop(i, j);
      }
  }
  m.resetTime();
}
// Prior trees of any model (migration nodes allowed): height of the local root and local tree length only.
extern "C" int ref_prior_stats(void* hv, int count, double* tmrca, double* tree_len) {
  RefHandle* h = (RefHandle*)hv;
  try {
    for (int t = 0; t < count; t++) {
      Forest f(&h->pf->model, h->rng);
      f.buildInitialTree(false);
      tmrca[t] = f.local_root()->height();
      tree_len[t] = f.getLocalTreeLength();
    }
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}

// Build the flat tree (n leaves = slots 0..n-1 with label slot+1, internal slots n..nn-1, parent[root] = -1) inside a
// Forest using the reference's own node API, the way Forest::createExampleTree does (forest-debug.cc:57-121).
static void build_tree(Forest* f, int n, int nn, const int* parent, const double* height) {
  f->clear();
  f->rec_bases_.push_back(0.0);
  f->current_rec_ = 1;
  std::vector<Node*> nd(nn, (Node*)NULL);
  for (int i = n - 1; i >= 0; i--) {
    nd[i] = f->nodes()->createNode(height[i], i + 1);
    nd[i]->set_label(i + 1);
    f->nodes()->add(nd[i]);
  }
  // internal nodes in increasing height order
  std::vector<int> order;
  for (int i = n; i < nn; i++) order.push_back(i);
  std::sort(order.begin(), order.end(), [&](int a, int b) { return height[a] < height[b]; });
  for (int v : order) {
    Node* c1 = NULL; Node* c2 = NULL;
    for (int c = 0; c < nn; c++) if (parent[c] == v) { if (!c1) c1 = nd[c]; else c2 = nd[c]; }
    nd[v] = f->nodes()->createNode(height[v]);
    f->addNodeToTree(nd[v], NULL, c1, c2);
  }
  Node* root = NULL;
  for (int i = 0; i < nn; i++) if (parent[i] < 0) root = nd[i];
  f->set_local_root(root);
  f->set_primary_root(root);
  f->set_sample_size(n);
  for (int i = 0; i < n; i++) f->updateAbove(nd[i]);
  f->contemporaries_ = ContemporariesContainer(f->model().population_number(), f->model().sample_size(),
                                               f->random_generator());
  f->tmp_event_time_ = -1;
  f->coalescence_finished_ = true;
}

// Export the local tree of a Forest as flat arrays.  Leaves -> slot label-1; internal (two local children) nodes ->
// slots n.. in increasing height order.  Unary / migration nodes are skipped (Node::getLocalChild1/2, node.cc:117-149).
// pop[] is the population label stored on the node.  Returns number of nodes written (2n-1) or -1.
static int export_tree(const Forest* f, int n, int* parent, double* height, int* pop) {
  std::vector<Node*> internal;
  std::vector<Node*> stack;
  stack.push_back(f->local_root());
  std::map<Node*, Node*> up;   // local-tree parent
  up[f->local_root()] = NULL;
  std::vector<Node*> leaves(n, (Node*)NULL);
  while (!stack.empty()) {
    Node* v = stack.back(); stack.pop_back();
    if (v->in_sample()) { leaves[v->label() - 1] = v; continue; }
    Node* a = v->getLocalChild1();
    Node* b = v->getLocalChild2();
    if (!a || !b) return -1;
    internal.push_back(v);
    up[a] = v; up[b] = v;
    stack.push_back(a); stack.push_back(b);
  }
  if ((int)internal.size() != n - 1) return -1;
  std::stable_sort(internal.begin(), internal.end(), [](Node* x, Node* y) { return x->height() < y->height(); });
  std::map<Node*, int> slot;
  for (int i = 0; i < n; i++) { if (!leaves[i]) return -1; slot[leaves[i]] = i; }
  for (int i = 0; i < n - 1; i++) slot[internal[i]] = n + i;
  for (auto& kv : slot) {
    int s = kv.second;
    height[s] = kv.first->height();
    if (pop) pop[s] = (int)kv.first->population();
    Node* p = up[kv.first];
    parent[s] = p ? slot[p] : -1;
  }
  return 2 * n - 1;
}

static ForestState* new_state(RefHandle* h) {
  ForestState* fs = new ForestState(&h->pf->model, h->rng, h->pf, false);
  fs->init_EventContainers(&h->pf->model);
  return fs;
}

// Pruning likelihood of one fully specified haplotype (alleles in {0,1,-1}) on a caller-supplied tree.
extern "C" int ref_prune(void* hv, int n, int nn, const int* parent, const double* height, const int* alleles,
                         int ancestral_aware, double* lik) {
  RefHandle* h = (RefHandle*)hv;
  try {
    ForestState* fs = new_state(h);
    build_tree(fs, n, nn, parent, height);
    std::vector<int> hap(alleles, alleles + n);
    *lik = fs->calculate_likelihood(ancestral_aware != 0, hap);
    delete fs;
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}

// Site likelihood with unphased hets (2) marginalised, as ParticleContainer::update_weight_at_site (particleContainer.cpp:204-221).
extern "C" int ref_site_likelihood(void* hv, int n, int nn, const int* parent, const double* height,
                                   const int* data, int ancestral_aware, int dephase, double* lik) {
  RefHandle* h = (RefHandle*)hv;
  try {
    ForestState* fs = new_state(h);
    build_tree(fs, n, nn, parent, height);
    ParticleContainer pc(&h->pf->model, h->rng, h->pf, 0, 0.0, false, std::vector<int>());
    std::vector<int> data_at_tips(data, data + n), hap;
    int ncfg = pc.calculate_initial_haplotype_configuration(data_at_tips, hap, dephase != 0);
    double norm = 1.0 / ncfg, acc = 0;
    do {
      acc += fs->calculate_likelihood(ancestral_aware != 0, hap);
    } while ((ncfg > 1) && pc.next_haplotype(hap, data_at_tips, dephase != 0));
    *lik = acc * norm;
    delete fs;
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}

// Branch length of the subtree spanning leaves with data (trackLocalTreeBranchLength) and total local tree length.
extern "C" int ref_branch_lengths(void* hv, int n, int nn, const int* parent, const double* height,
                                  const int* data, double* with_data, double* total) {
  RefHandle* h = (RefHandle*)hv;
  try {
    ForestState* fs = new_state(h);
    build_tree(fs, n, nn, parent, height);
    std::vector<int> d(data, data + n);
    *with_data = fs->trackLocalTreeBranchLength(d);
    *total = fs->getLocalTreeLength();
    delete fs;
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}

// The no-mutation factor of extend_ARG over one recombination-free stretch: fastexp(-mu * L * dx) (particle.cpp:770)
// fastexp is file-static in particle.cpp, so evaluate it through extend_ARG on a tree whose next recombination is
// beyond the stretch.
extern "C" int ref_extend_nomut(void* hv, int n, int nn, const int* parent, const double* height,
                                const int* data, double x0, double x1, double* factor) {
  RefHandle* h = (RefHandle*)hv;
  try {
    ForestState* fs = new_state(h);
    build_tree(fs, n, nn, parent, height);
    fs->rec_bases_.back() = x0;
    fs->set_next_base(x1 + 1.0);
    fs->setSiteWhereWeightWasUpdated(x0);
    fs->setPosteriorWeight(1.0);
    fs->setPilotWeight(1.0);
    int nmiss = 0;
    for (int i = 0; i < n; i++) nmiss += data[i] == -1;
    int leaf_status = nmiss == 0 ? 1 : (nmiss == n ? -1 : 0);
    std::vector<int> d(data, data + n);
    fs->extend_ARG(x1, leaf_status, d, NULL);
    *factor = fs->posteriorWeight();
    delete fs;
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}

// ParticleContainer::systematic_resampling (particleContainer.cpp:474) with the single uniform scripted to `u`.
extern "C" int ref_systematic(void* hv, const double* partial_sum /*[R+1]*/, int R, int N, double u, int* counts) {
  RefHandle* h = (RefHandle*)hv;
  try {
    ParticleContainer pc(&h->pf->model, h->rng, h->pf, 0, 0.0, false, std::vector<int>());
    std::valarray<double> ps(partial_sum, R + 1);
    std::valarray<int> sc(R);
    h->rng->script.clear();
    h->rng->script.push_back(u);
    pc.systematic_resampling(ps, sc, N);
    for (int i = 0; i < R; i++) counts[i] = sc[i];
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}

// Full weight path on N real particles (prior trees; only their weights matter):
//   normalize_probability (particleContainer.cpp:420) on the given posterior weights, then resample
//   (particleContainer.cpp:247) with ESS threshold from the handle's -ESS flag and the uniform scripted to `u`.
// Outputs: ln normaliser, ESS decision (1 = resampled), per-input-particle multiplicity, post-resampling posterior
// weight of each surviving record (0 for dead ones).
extern "C" int ref_weights_path(void* hv, int N, const double* w_post, const double* w_pilot, double u,
                                double* lnZ, int* resampled, int* multiplicity, double* w_out) {
  RefHandle* h = (RefHandle*)hv;
  try {
    h->pf->N = N;
    h->pf->ESSthreshold = N * h->pf->ESS_fraction;
    ParticleContainer pc(&h->pf->model, h->rng, h->pf, N, 0.0, false, std::vector<int>());
    std::map<ForestState*, int> idx;
    for (int i = 0; i < N; i++) {
      pc.particles[i]->setPosteriorWeight(w_post[i]);
      pc.particles[i]->setPilotWeight(w_pilot[i]);
      idx[pc.particles[i]] = i;
    }
    pc.normalize_probability();
    *lnZ = pc.ln_normalization_factor();
    h->rng->script.clear();
    h->rng->script.push_back(u);
    bool rec = h->pf->record_resample_file;
    h->pf->record_resample_file = false;
    *resampled = pc.resample(0, *h->pf);
    h->pf->record_resample_file = rec;
    h->rng->script.clear();
    for (int i = 0; i < N; i++) { multiplicity[i] = 0; w_out[i] = 0; }
    for (size_t k = 0; k < pc.particles.size(); k++) {
      int i = idx[pc.particles[k]];
      multiplicity[i] = *resampled ? pc.particles[k]->multiplicity() : 1;
      w_out[i] = pc.particles[k]->posteriorWeight();
    }
    pc.clear();
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}

// Prior trees: Forest::buildInitialTree(false) (forest.cc:299) -> flat tree.  count trees, arrays [count][2n-1].
extern "C" int ref_prior_trees(void* hv, int count, int* parent, double* height, double* tree_len) {
  RefHandle* h = (RefHandle*)hv;
  try {
    int n = (int)h->pf->model.sample_size(), nn = 2 * n - 1;
    for (int t = 0; t < count; t++) {
      Forest f(&h->pf->model, h->rng);
      f.buildInitialTree(false);
      if (export_tree(&f, n, parent + (size_t)t * nn, height + (size_t)t * nn, NULL) < 0) {
        g_err = "export_tree failed"; return 1;
      }
      tree_len[t] = f.getLocalTreeLength();
    }
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}

// One SMC' transition from a caller-supplied tree using the reference's sampleNextGenealogy (forest.cc:423), repeated
// `reps` times from the same start tree; returns the flat result trees.  Used for distribution-level parity of the
// SMC' move (recombination point + re-coalescence) against the C restatement / CUDA kernel.
extern "C" int ref_smc_moves(void* hv, int n, const int* parent0, const double* height0, int reps,
                             int* parent, double* height, double* rec_height) {
  RefHandle* h = (RefHandle*)hv;
  try {
    int nn = 2 * n - 1;
    for (int r = 0; r < reps; r++) {
      Forest f(&h->pf->model, h->rng);
      build_tree(&f, n, nn, parent0, height0);
      f.set_next_base(100.0);             // pretend the next recombination is at x=100
      f.rec_bases_.push_back(100.0);      // sampleNextGenealogy does ++current_rec_
      f.sampleNextGenealogy(false);
      rec_height[r] = f.rec_point.height();
      // force pruning of the old stub so that the local tree can be exported
      for (TimeIntervalIterator ti(&f, f.nodes()->at(0)); ti.good(); ++ti) {}
      if (export_tree(&f, n, parent + (size_t)r * nn, height + (size_t)r * nn, NULL) < 0) {
        g_err = "export_tree failed after move"; return 1;
      }
    }
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}

// Walk the prior ARG along the sequence with the reference (buildInitialTree + sampleNextGenealogy +
// sampleRecSeqPosition, as smcsmc.cpp:193-215) and report for each of `count` recombinations its position, the
// recombination height, the tree height and tree length after the move.
extern "C" int ref_arg_walk(void* hv, int count, double* pos, double* rec_h, double* tmrca, double* tlen) {
  RefHandle* h = (RefHandle*)hv;
  try {
    Forest f(&h->pf->model, h->rng);
    f.buildInitialTree(false);
    for (int i = 0; i < count; i++) {
      if (f.next_base() >= h->pf->model.loci_length()) { count = i; break; }
      f.sampleNextGenealogy(false);
      pos[i] = f.current_base();
      rec_h[i] = f.rec_point.height();
      f.sampleRecSeqPosition(false);
      tmrca[i] = f.local_root()->height();
      tlen[i] = f.getLocalTreeLength();
    }
    return count;
  } catch (std::exception& e) { g_err = e.what(); return -1; }
}

// Replay a scripted particle history through the reference's own event records and CountModel.
// Script entries (doubles, 8 per op), op codes:
//   0 COAL    : slot, epoch, t0, t1, x, k, has_event      -> record as particle.cpp:279-290
//   1 RECOPP  : slot, epoch, t0, t1, x0, x1, k            -> record as particle.cpp:348-353
//   2 RECEV   : slot, epoch, t0, t1, x0, x1, k  (+ event at height (t0+t1)/2 on that opportunity, particle.cpp:373-376)
//   3 RESAMPLE: followed by N ancestor indices packed in ceil(N/7) following ops' payload  (physical copies,
//               ForestState copy-ctor particle.cpp:113, shares the event chains via copyEventContainers :139)
//   4 UPDATE  : current_base, end_data, then N weights in following ops' payload   (count.cpp:355)
// lags[E] overrides CountModel::lags.  out[6][E] as orc_lagged_tallies (initial pseudo-counts subtracted).
extern "C" int ref_tally_script(void* hv, int N, const double* lags, const double* script, long nops, double* out) {
  RefHandle* h = (RefHandle*)hv;
  try {
    Model* model = &h->pf->model;
    int E = (int)model->change_times_.size();
    h->pf->N = N;
    ParticleContainer pc(model, h->rng, h->pf, N, 0.0, false, std::vector<int>());
    // drop the records made by the constructor (record_recomb_extension of the prior tree)
    for (int i = 0; i < N; i++) {
      pc.particles[i]->clear_eventContainer();
      for (int e = 0; e <= E; e++) if (e < (int)pc.particles[i]->eventTrees.size()) pc.particles[i]->eventTrees[e] = NULL;
    }
    CountModel cm(*model, h->pf->lag);
    cm.init();
    for (int e = 0; e < E; e++) cm.lags[e] = lags[e];
    std::vector<double> base(6 * E);
    for (int e = 0; e < E; e++) {
      base[0 * E + e] = cm.total_coal_count[e][0];   base[1 * E + e] = cm.total_coal_opportunity[e][0];
      base[2 * E + e] = cm.total_coal_weight[e][0];  base[3 * E + e] = cm.total_recomb_count[e][0];
      base[4 * E + e] = cm.total_recomb_opportunity[e][0]; base[5 * E + e] = cm.total_recomb_weight[e][0];
    }
    long k = 0;
    auto payload = [&](int count, std::vector<double>& dst) {
      dst.clear();
      while ((int)dst.size() < count) {
        const double* op = script + 8 * k; k++;
        for (int j = 1; j < 8 && (int)dst.size() < count; j++) dst.push_back(op[j]);
      }
    };
    while (k < nops) {
      const double* op = script + 8 * k; k++;
      int code = (int)op[0];
      if (code == 0) {
        ForestState* fs = pc.particles[(int)op[1]];
        int e = (int)op[2];
        void* mem = Arena::allocate(e);
        EvolutionaryEvent* ev = new (mem) EvolutionaryEvent(op[3], (size_t)e, op[4], (size_t)e, op[5], (size_t)0, (int)op[6]);
        if (op[7] != 0) ev->set_coal_event();
        ev->add_leaf_to_tree(&fs->eventTrees[e]);
      } else if (code == 1 || code == 2) {
        ForestState* fs = pc.particles[(int)op[1]];
        int e = (int)op[2];
        void* mem = Arena::allocate(e);
        EvolutionaryEvent* ev = new (mem) EvolutionaryEvent(op[3], (size_t)e, op[4], (size_t)e, op[5], op[6], (int)op[7]);
        if (code == 2) ev->set_recomb_event_time(0.5 * (op[3] + op[4]));
        ev->add_leaf_to_tree(&fs->eventTrees[e]);
      } else if (code == 3) {
        std::vector<double> a;
        payload(N, a);
        std::vector<ForestState*> np(N);
        std::vector<int> used(N, 0);
        for (int i = 0; i < N; i++) {
          int src = (int)a[i];
          if (!used[src]) { np[i] = pc.particles[src]; used[src] = 1; }
          else np[i] = new ForestState(*pc.particles[src]);
        }
        for (int i = 0; i < N; i++) if (!used[i]) delete pc.particles[i];
        pc.particles = np;
      } else if (code == 4) {
        double cur = op[1]; bool end_data = op[2] != 0;
        std::vector<double> w;
        payload(N, w);
        for (int i = 0; i < N; i++) { pc.particles[i]->setPosteriorWeight(w[i]); pc.particles[i]->setMultiplicity(1); }
        cm.extract_and_update_count(pc, cur, end_data);
      }
    }
    for (int e = 0; e < E; e++) {
      out[0 * E + e] = cm.total_coal_count[e][0] - base[0 * E + e];
      out[1 * E + e] = cm.total_coal_opportunity[e][0] - base[1 * E + e];
      out[2 * E + e] = cm.total_coal_weight[e][0] - base[2 * E + e];
      out[3 * E + e] = cm.total_recomb_count[e][0] - base[3 * E + e];
      out[4 * E + e] = cm.total_recomb_opportunity[e][0] - base[4 * E + e];
      out[5 * E + e] = cm.total_recomb_weight[e][0] - base[5 * E + e];
    }
    pc.clear();
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}


// The reference's REAL recording path around a spawn, with the heap block of the copy pre-filled with a byte pattern.
// extend_ARG, when a particle of multiplicity > 1 recombines (particle.cpp:836-854), runs
//     spawn()  ->  new ForestState(*this)            (particle.hpp:137, copy constructor particle.cpp:113-136)
//     copy->set_current_base(updated_to); copy->setMultiplicity(mult - 1);
//     copy->resample_recombination_position();       (particle.cpp:393-436)
// and only later does ParticleContainer::extend_ARGs assign copy->max_epoch_to_record_ (particleContainer.cpp:123).  The
// copy constructor never writes that field, so the guard `epoch <= max_epoch_to_record_` (particle.cpp:402) reads whatever
// the allocator handed out.  The probe makes "whatever" explicit: the copy is placement-constructed into a block filled
// with fill_byte.  Outputs: the field as the copy sees it, the number of epochs that hold a recombination-opportunity
// record of the parent, the number of those whose record the copy extended to its new next_base, and the opportunity
// (lineages x generations x bp) recorded for the copy's stretch [spawn point, new next_base) next to the tree length
// times that distance.
extern "C" int ref_spawn_probe(void* hv, int fill_byte, int* field_seen, int* epochs_recorded, int* epochs_extended,
                               double* recorded, double* expected) {
  RefHandle* h = (RefHandle*)hv;
  try {
    Model* model = &h->pf->model;
    const int E = (int)model->change_times_.size();
    ForestState* p = new ForestState(model, h->rng, h->pf, false);
    p->init_EventContainers(model);
    p->buildInitialTree(true);
    p->set_max_epoch_to_record(E - 1);
    p->record_recomb_extension();            // as the ParticleContainer constructor does (particleContainer.cpp:54)
    p->save_recomb_state();
    p->setMultiplicity(3);
    const double updated_to = p->next_base();   // the parent's recombination point
    void* mem = operator new(sizeof(ForestState));
    memset(mem, fill_byte, sizeof(ForestState));
    ForestState* c = new (mem) ForestState(*p);  // == spawn()
    *field_seen = c->max_epoch_to_record_;
    c->set_current_base(updated_to);
    c->setMultiplicity(2);
    p->save_recomb_state();
    c->restore_recomb_state();
    c->resample_recombination_position();
    c->save_recomb_state();
    p->restore_recomb_state();
    const double x1 = c->next_base();
    *epochs_recorded = 0; *epochs_extended = 0; *recorded = 0.0;
    for (int e = 0; e < E; e++) {
      if (p->eventTrees[e] && p->eventTrees[e]->is_recomb()) (*epochs_recorded)++;
      bool extended = false;
      for (EvolutionaryEvent* ev = c->eventTrees[e]; ev; ev = ev->parent()) {
        if (!ev->is_recomb()) continue;
        if (ev->end_base() == x1) extended = true;
        *recorded += ev->recomb_opportunity_between(0.0, DBL_MAX, updated_to, x1);
      }
      if (extended) (*epochs_extended)++;
    }
    *expected = c->getLocalTreeLength() * (x1 - updated_to);
    c->~ForestState();
    operator delete(mem);
    delete p;
    return 0;
  } catch (std::exception& e) { g_err = e.what(); return 1; }
}
