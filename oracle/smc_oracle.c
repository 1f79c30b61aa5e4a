/* TEST INFRASTRUCTURE ONLY -- see smc_oracle.h.  Plain C restatement of the reference's hot path; compiled with
 * -ffp-contract=off so that no FMA contraction changes the arithmetic relative to the reference's x86-64 build. */
#include "smc_oracle.h"
#include <math.h>
#include <string.h>
#include <float.h>

/* ------------------------------------------------------------------------------------------------------------
 * deterministic pieces
 * ---------------------------------------------------------------------------------------------------------- */

/* particle.cpp:30-40 -- continued-fraction exp, exact rational form */
double orc_fastexp(double x) {
  double xx = x * x;
  if (xx < 0.516167859) return 1 + 2 * x / (2 - x + xx / (6 + xx * 0.1));
  return exp(x);
}

/* particle.cpp:65-74 */
double orc_exp_digamma(double x) {
  if (x > 10) return x - 0.5 + (x + 0.5) / (24 * x * x);
  double f = 0.0;
  while (x < 6) { f = f + 1.0 / x; x = x + 1.0; }
  double psi = log(x) - 1 / (2 * x) - 1 / (12 * x * x);
  return exp(psi - f);
}

static void children_of(int nn, const int* parent, int v, int* c1, int* c2) {
  *c1 = -1; *c2 = -1;
  for (int c = 0; c < nn; c++)
    if (parent[c] == v) { if (*c1 < 0) *c1 = c; else *c2 = c; }
}

/* particle.cpp:625-658 cal_partial_likelihood_infinite (recursive, two states, symmetric mutation model) */
static void partial_lik(int n, int nn, const int* parent, const double* height, const int* hap, double mu, int v,
                        double out[2]) {
  if (v < n) {
    int s = hap[v];
    out[0] = s == 1 ? 0.0 : 1.0;   /* missing (-1) encodes as (1,1) */
    out[1] = s == 0 ? 0.0 : 1.0;
    return;
  }
  int l, r;
  children_of(nn, parent, v, &l, &r);
  double t_left = height[v] - height[l];
  double t_right = height[v] - height[r];
  double p_left = orc_fastexp(-t_left * mu);
  double p_right = orc_fastexp(-t_right * mu);
  double ml[2], mr[2];
  partial_lik(n, nn, parent, height, hap, mu, l, ml);
  partial_lik(n, nn, parent, height, hap, mu, r, mr);
  out[0] = (ml[0] * p_left + ml[1] * (1 - p_left)) * (mr[0] * p_right + mr[1] * (1 - p_right));
  out[1] = (ml[1] * p_left + ml[0] * (1 - p_left)) * (mr[1] * p_right + mr[0] * (1 - p_right));
}

static int root_of(int nn, const int* parent) {
  for (int v = 0; v < nn; v++) if (parent[v] < 0) return v;
  return -1;
}

/* particle.cpp:667-680 calculate_likelihood */
double orc_prune_likelihood(int n, const int* parent, const double* height, const int* hap, double mu,
                            int ancestral_aware) {
  int nn = 2 * n - 1;
  double m[2];
  partial_lik(n, nn, parent, height, hap, mu, root_of(nn, parent), m);
  double prior0 = ancestral_aware ? 1.0 : 0.5, prior1 = ancestral_aware ? 0.0 : 0.5;
  return m[0] * prior0 + m[1] * prior1;
}

/* particleContainer.cpp:138-181 (phasing enumeration) + :204-221 (mean over configurations) */
double orc_site_likelihood(int n, const int* parent, const double* height, const int* data, double mu,
                           int ancestral_aware, int dephase) {
  int hap[ORC_MAXN];
  int ncfg = 1;
  for (int i = 0; i < n; i++) hap[i] = data[i];
  for (int i = 0; i + 1 < n; i += 2) {
    int het = (data[i] == 2) || (dephase && data[i] + data[i + 1] == 1);
    if (het) { ncfg *= 2; hap[i] = 0; hap[i + 1] = 1; }
  }
  double norm = 1.0 / ncfg, acc = 0.0;
  for (;;) {
    acc += orc_prune_likelihood(n, parent, height, hap, mu, ancestral_aware);
    if (ncfg <= 1) break;
    /* next_haplotype: binary counter over the unphased pairs */
    int advanced = 0;
    for (int i = 0; i + 1 < n; i += 2) {
      int het = (data[i] == 2) || (dephase && data[i] + data[i + 1] == 1);
      if (!het) continue;
      if (hap[i] == 0) { hap[i] = 1; hap[i + 1] = 0; advanced = 1; break; }
      hap[i] = 0; hap[i + 1] = 1;
    }
    if (!advanced) break;
  }
  return acc * norm;
}

/* forest.h:188 getLocalTreeLength == sum of branch lengths of the local tree */
double orc_tree_length(int n, const int* parent, const double* height) {
  int nn = 2 * n - 1;
  double len = 0.0;
  for (int v = 0; v < nn; v++) if (parent[v] >= 0) len += height[parent[v]] - height[v];
  return len;
}

/* particle.cpp:706-730 trackSubtreeBranchLength */
static double subtree_bl(int n, int nn, const int* parent, const double* height, const int* data, int v,
                         double* total) {
  if (v < n) return data[v] >= 0 ? 0.0 : -1.0;
  int l, r;
  children_of(nn, parent, v, &l, &r);
  double sl = subtree_bl(n, nn, parent, height, data, l, total);
  double sr = subtree_bl(n, nn, parent, height, data, r, total);
  if (sl >= 0.0) sl += height[v] - height[l];
  if (sr >= 0.0) sr += height[v] - height[r];
  if (sl >= 0.0 && sr >= 0.0) { *total = sl + sr; return *total; }
  if (sl >= 0.0) return sl;
  return sr;
}

/* particle.cpp:699-703 trackLocalTreeBranchLength */
double orc_branch_length_with_data(int n, const int* parent, const double* height, const int* data) {
  int nn = 2 * n - 1;
  double total = 0.0;
  subtree_bl(n, nn, parent, height, data, root_of(nn, parent), &total);
  return total;
}

/* Local branch length inside each epoch = recombination opportunity per bp, the sum over the time slices of
 * record_recomb_extension (particle.cpp:305-357) of  weight * (end_height - start_height). */
void orc_epoch_tree_length(const orc_model* m, const int* parent, const double* height, double* len_e) {
  int nn = 2 * m->n - 1;
  for (int e = 0; e < m->E; e++) len_e[e] = 0.0;
  for (int v = 0; v < nn; v++) {
    if (parent[v] < 0) continue;
    double lo = height[v], hi = height[parent[v]];
    for (int e = 0; e < m->E; e++) {
      double es = m->t_start[e], ee = e + 1 < m->E ? m->t_start[e + 1] : DBL_MAX;
      double a = lo > es ? lo : es, b = hi < ee ? hi : ee;
      if (b > a) len_e[e] += b - a;
    }
  }
}

/* particleContainer.cpp:474-504 systematic_resampling, with the uniform passed in (u = rg.sample()) */
void orc_systematic_resampling(const double* partial_sum, int R, int N, double u, int* counts) {
  int sample_i = 0;
  double u_j = u / N;
  int interval_j = 0;
  double norm = partial_sum[R];
  counts[0] = 0;
  while (sample_i < N) {
    if (partial_sum[interval_j + 1] / norm > u_j) {
      counts[interval_j] += 1;
      sample_i++;
      u_j += 1.0 / (double)N;
    } else {
      interval_j++;
      counts[interval_j] = 0;
    }
  }
  interval_j++;
  for (; interval_j < R; interval_j++) counts[interval_j] = 0;
}

/* particleContainer.cpp:420-438 */
double orc_normalize(double* w, int N) {
  double total = 0;
  for (int i = 0; i < N; i++) total += w[i];
  double inv = 1.0 / total;
  for (int i = 0; i < N; i++) w[i] *= inv;
  return log(total);
}

/* particleContainer.cpp:256-270 */
double orc_ess_partial_sums(const double* w, int N, double* partial_sum) {
  double wi_sum = 0, wi_sq_sum = 0;
  for (int i = 0; i < N; i++) { partial_sum[i] = wi_sum; wi_sum += w[i]; }
  double scaling = 1.0 / wi_sum;
  for (int i = 0; i < N; i++) wi_sq_sum += (w[i] * scaling) * (w[i] * scaling);
  partial_sum[N] = wi_sum;
  return 1.0 / wi_sq_sum;
}

/* log-domain restatement of normalize_probability: logw_i -= log(sum exp(logw)); returns log(sum) */
double orc_normalize_log(double* logw, int N) {
  double mx = -DBL_MAX;
  for (int i = 0; i < N; i++) if (logw[i] > mx) mx = logw[i];
  double s = 0;
  for (int i = 0; i < N; i++) s += exp(logw[i] - mx);
  double lz = mx + log(s);
  for (int i = 0; i < N; i++) logw[i] -= lz;
  return lz;
}

/* ------------------------------------------------------------------------------------------------------------
 * Philox4x32-10 (Salmon et al. 2011), counter-based.  Replaces the reference's single global mt19937_64 stream
 * (scrm/src/random/mersenne_twister.cc) -- streams differ BY DESIGN, see DESIGN.md "RNG contract".
 * ---------------------------------------------------------------------------------------------------------- */
void orc_philox4x32_10(const uint32_t ctr_in[4], const uint32_t key_in[2], uint32_t out[4]) {
  uint32_t c0 = ctr_in[0], c1 = ctr_in[1], c2 = ctr_in[2], c3 = ctr_in[3];
  uint32_t k0 = key_in[0], k1 = key_in[1];
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

void orc_rng_init(orc_rng* r, uint64_t seed, uint32_t slot, uint32_t step, uint32_t chunk, uint32_t purpose) {
  r->key[0] = (uint32_t)seed; r->key[1] = (uint32_t)(seed >> 32);
  r->ctr[0] = 0; r->ctr[1] = slot; r->ctr[2] = step; r->ctr[3] = (chunk & 0xFFFFFFu) | (purpose << 24);
  r->have = 0;
}

static double u53(uint32_t a, uint32_t b) {
  /* 53 random bits + half an ulp: strictly inside (0,1), so -log(u) is finite and positive */
  return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6) + 0.5) * (1.0 / 9007199254740992.0);
}

double orc_rng_uniform(orc_rng* r) {
  if (r->have == 0) {
    uint32_t o[4];
    orc_philox4x32_10(r->ctr, r->key, o);
    r->ctr[0]++;
    r->buf[0] = u53(o[0], o[1]);
    r->buf[1] = u53(o[2], o[3]);
    r->have = 2;
  }
  double u = r->buf[2 - r->have];
  r->have--;
  return u;
}

/* ------------------------------------------------------------------------------------------------------------
 * stochastic pieces
 * ---------------------------------------------------------------------------------------------------------- */
static void emit(orc_events* t, int type, int epoch, int cnt, double x0, double x1, double opp) {
  if (!t) return;
  if (t->count >= t->cap) { t->overflow = 1; return; }
  orc_event* e = &t->ev[t->count++];
  e->type = type; e->epoch = epoch; e->cnt = cnt; e->x0 = x0; e->x1 = x1; e->opp = opp;
}

static int epoch_of(const orc_model* m, double t) {
  int e = 0;
  while (e + 1 < m->E && m->t_start[e + 1] <= t) e++;
  return e;
}

/* Next recombination position: Exp(rho * tree_len) from `from`, limited to the locus end
 * (particle.cpp:1195-1254 with multiplicity 1, one rho segment; forest.h:140-151 for the re-draw). */
static void draw_next_base(const orc_model* m, orc_particle* p, orc_rng* r, double from) {
  double rate = m->rho * orc_tree_length(m->n, p->parent, p->height);
  double len = -log(orc_rng_uniform(r)) / rate;
  double next = from + len;
  if (next == from) next = nextafter(from, from * 2);   /* particle.cpp:1238-1244 underflow guard */
  if (!(next < m->seqlen)) next = m->seqlen;
  p->next_base = next;
}
void orc_redraw_next_base(const orc_model* m, orc_particle* p, orc_rng* r) { draw_next_base(m, p, r, p->cur_base); }

/* Prior tree at the left end: Kingman coalescent under piecewise-constant epochs.  The reference adds the leaves one
 * at a time and coalesces each into the growing tree (forest.cc:299-347); the law of the resulting tree is the
 * standard coalescent, which is what is sampled here (n lineages, rate C(k,2)/(2N(e)), uniform pair). */
void orc_init_particle(const orc_model* m, orc_particle* p, orc_rng* r, double start_pos) {
  int n = m->n, nn = 2 * n - 1;
  int active[ORC_MAXN];
  for (int i = 0; i < n; i++) { active[i] = i; p->height[i] = 0.0; }
  for (int v = 0; v < nn; v++) p->parent[v] = -1;
  int k = n, e = 0, next_slot = n;
  double t = 0.0;
  while (k > 1) {
    double pairs = 0.5 * k * (k - 1);
    double rate = pairs * m->inv2N[e];
    double w = -log(orc_rng_uniform(r)) / rate;
    double tend = e + 1 < m->E ? m->t_start[e + 1] : DBL_MAX;
    if (t + w < tend) {
      t += w;
      int idx = (int)(orc_rng_uniform(r) * pairs);
      if (idx >= (int)pairs) idx = (int)pairs - 1;
      int i = 0;
      while (idx >= k - 1 - i) { idx -= k - 1 - i; i++; }
      int j = i + 1 + idx;
      p->height[next_slot] = t;
      p->parent[active[i]] = next_slot;
      p->parent[active[j]] = next_slot;
      active[i] = next_slot;
      active[j] = active[k - 1];
      next_slot++; k--;
    } else {
      t = tend; e++;
    }
  }
  p->logw = 0.0;
  p->updated_to = start_pos;
  p->cur_base = start_pos;
  draw_next_base(m, p, r, start_pos);
}

static void swap_slots(int nn, int* parent, double* height, int a, int b) {
  double th = height[a]; height[a] = height[b]; height[b] = th;
  int tp = parent[a]; parent[a] = parent[b]; parent[b] = tp;
  for (int v = 0; v < nn; v++) {
    if (parent[v] == a) parent[v] = b;
    else if (parent[v] == b) parent[v] = a;
  }
}

/* SMC' transition at position p->cur_base (Forest::sampleNextGenealogy, forest.cc:423-467, with exact window 0):
 *  1. recombination point uniform on the local tree            (forest.cc:370-412 / particle.cpp:1060 unbiased case)
 *  2. the cut lineage floats upwards through the intervals bounded by node heights and epoch changes
 *     (time_interval.cc:84-134); rate = k * 1/(2N(e)) with k = branches of the ORIGINAL tree crossing the interval,
 *     the stub above the cut included (forest.cc:601-664,805-808); above the old root the two remaining lineages
 *     coalesce pairwise at rate 1/(2N(e)) (forest.cc:626-636,811-814)
 *  3. target uniform among the k contemporaries (contemporaries_container.h:343-365)
 *  4. the unary node left at the cut is pruned (forest.cc:1245-1257): the parent slot of the cut branch is re-used
 *     as the new coalescence node.
 * Tally records as ForestState::record_all_event (particle.cpp:193-302) and record_recomb_event (:360-377). */
void orc_smc_move(const orc_model* m, orc_particle* p, orc_rng* r, int max_rec_epoch, orc_events* t,
                  double* rec_height_out) {
  int n = m->n, nn = 2 * n - 1;
  int* parent = p->parent; double* height = p->height;
  double x = p->cur_base;
  /* 1. recombination point */
  double tl = orc_tree_length(n, parent, height);
  double target = orc_rng_uniform(r) * tl;
  int b = -1; double hh = 0.0, cum = 0.0;
  for (int v = 0; v < nn; v++) {
    if (parent[v] < 0) continue;
    double bl = height[parent[v]] - height[v];
    b = v; hh = height[v] + (target - cum);
    if (target < cum + bl) break;
    cum += bl;
  }
  if (hh > height[parent[b]]) hh = height[parent[b]];
  if (rec_height_out) *rec_height_out = hh;
  int e = epoch_of(m, hh);
  if ((m->rec_mask[e] & 1) && e <= max_rec_epoch) emit(t, ORC_EV_REC, e, 1, x, x, 0.0);
  /* 2. float */
  int Pn = parent[b];
  int root = root_of(nn, parent);
  double hroot = height[root];
  double tt = hh, tc = 0.0;
  int k = 0, above_root = 0;
  /* one unit exponential for the whole float, used up at rate * length per interval (memoryless; the reference's
     sampleExpoExpoLimit re-uses its unit exponential across intervals the same way, random_generator.cc:32-78) */
  double ex = -log(orc_rng_uniform(r));
  for (;;) {
    double tnode = DBL_MAX;
    above_root = !(tt < hroot);
    if (!above_root) {
      for (int v = n; v < nn; v++) if (height[v] > tt && height[v] < tnode) tnode = height[v];
      k = 0;
      for (int v = 0; v < nn; v++) if (parent[v] >= 0 && height[v] <= tt && tt < height[parent[v]]) k++;
    } else {
      k = 1;
    }
    double tep = e + 1 < m->E ? m->t_start[e + 1] : DBL_MAX;
    double tend = tnode < tep ? tnode : tep;
    double rate = k * m->inv2N[e];
    double need = rate * (tend - tt);
    int coal = ex < need;
    if (coal) { tc = tt + ex / rate; if (!(tc < tend)) tc = nextafter(tend, 0.0); }
    double dt = (coal ? tc : tend) - tt;
    if ((m->rec_mask[e] & 2) && e <= max_rec_epoch) emit(t, ORC_EV_COAL, e, coal, x, x, k * dt);
    if (coal) break;
    ex -= need;
    tt = tend;
    if (tend == tep) e++;
  }
  p->logw += m->vb_logfac[e];
  /* 3. target + 4. surgery */
  int S = -1;
  for (int v = 0; v < nn; v++) if (parent[v] == Pn && v != b) S = v;
  if (above_root) {
    if (Pn != root) {
      parent[S] = parent[Pn];
      parent[root] = Pn;
      parent[Pn] = -1;
    }
    height[Pn] = tc;
  } else {
    int idx = (int)(orc_rng_uniform(r) * k);
    if (idx >= k) idx = k - 1;
    int T = -1;
    for (int v = 0; v < nn; v++)
      if (parent[v] >= 0 && height[v] <= tt && tt < height[parent[v]]) { if (idx == 0) { T = v; break; } idx--; }
    if (T == b) {
      /* back-coalescence into the stub: tree unchanged */
    } else if (T == S || T == Pn) {
      height[Pn] = tc;
    } else {
      parent[S] = parent[Pn];
      parent[Pn] = parent[T];
      parent[T] = Pn;
      height[Pn] = tc;
    }
  }
  /* keep internal slots sorted by height */
  int s = Pn;
  while (s + 1 < nn && height[s + 1] < height[s]) { swap_slots(nn, parent, height, s, s + 1); s++; }
  while (s - 1 >= n && height[s - 1] > height[s]) { swap_slots(nn, parent, height, s, s - 1); s--; }
}

/* ForestState::extend_ARG (particle.cpp:743-918) for one particle over one .seg row, followed by the site likelihood
 * of ParticleContainer::update_weight_at_site (particleContainer.cpp:187-224) when is_site != 0. */
void orc_extend_row(const orc_model* m, orc_particle* p, orc_rng* r, double x1, const int* data, int is_site,
                    int ancestral_aware, int max_rec_epoch, orc_events* t) {
  int n = m->n;
  int nmiss = 0;
  for (int i = 0; i < n; i++) nmiss += data[i] == -1;
  int leaf_status = nmiss == 0 ? 1 : (nmiss == n ? -1 : 0);   /* particleContainer.cpp:103-111 */
  double L_obs = leaf_status == -1 ? 0.0
               : leaf_status == 1 ? orc_tree_length(n, p->parent, p->height)
                                  : orc_branch_length_with_data(n, p->parent, p->height, data);
  double len_e[ORC_MAXE];
  orc_epoch_tree_length(m, p->parent, p->height, len_e);
  while (p->updated_to < x1) {
    double new_to = x1 < p->next_base ? x1 : p->next_base;
    p->logw += log(orc_fastexp(-m->mu * L_obs * (new_to - p->updated_to)));   /* particle.cpp:770 */
    for (int e = 0; e < m->E; e++)
      if (len_e[e] > 0.0 && (m->rec_mask[e] & 1) && e <= max_rec_epoch)
        emit(t, ORC_EV_RECOPP, e, 0, p->updated_to, new_to, len_e[e]);
    p->updated_to = new_to;
    if (p->updated_to < x1) {
      p->cur_base = p->updated_to;
      orc_smc_move(m, p, r, max_rec_epoch, t, 0);
      if (leaf_status == 0) L_obs = orc_branch_length_with_data(n, p->parent, p->height, data);
      if (leaf_status == 1) L_obs = orc_tree_length(n, p->parent, p->height);
      orc_epoch_tree_length(m, p->parent, p->height, len_e);
      draw_next_base(m, p, r, p->cur_base);
    }
    p->cur_base = p->updated_to;   /* particle.cpp:907 */
  }
  if (is_site) p->logw += log(orc_site_likelihood(n, p->parent, p->height, data, m->mu, ancestral_aware, 0));
}
