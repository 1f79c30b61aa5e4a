// Command-line parser: PfParam::parse (pfparam.cpp:51-169) + the subset of scrm's Param::parse (param.cc:37-520) that
// the piecewise-constant engine supports: sizes per epoch and population, migration matrices, joins (-ej / -eps).
#include "params.hpp"
#include <iostream>
#include <algorithm>
#include <cstdio>

namespace host {

namespace {

struct Args {
  std::vector<std::string> v;
  size_t i = 0;
  bool more() const { return i < v.size(); }
  const std::string& cur() const { return v[i]; }
  std::string next(const std::string& flag) {
    if (i + 1 >= v.size()) throw std::invalid_argument("Not enough parameters when parsing options: " + flag);
    return v[++i];
  }
  double next_double(const std::string& flag) { return to_double(next(flag), flag); }
};

void unsupported(const std::string& flag, const char* what) {
  throw std::invalid_argument("option " + flag + " (" + what + ") is not supported by the B200 engine yet");
}

// -xr / -xc take "a" or "a-b" (0-based closed range), pfparam.cpp readRange
int read_range(const std::string& s, int& last) {
  size_t dash = s.find('-');
  int first;
  if (dash == std::string::npos) { first = last = (int)to_double(s, "-xr/-xc"); }
  else { first = (int)to_double(s.substr(0, dash), "-xr/-xc"); last = (int)to_double(s.substr(dash + 1), "-xr/-xc"); }
  if (first < 0 || last < first) throw std::invalid_argument("Out of range: -xr/-xc " + s);
  return first;
}

const double kNaN = std::nan("");

// Model::addChangeTime (model.cc:96-130): position of the epoch starting at `t` generations, inserted if new
Epoch& epoch_at(Params& P, double t) {
  size_t pos = 0;
  for (; pos < P.epochs.size(); pos++) {
    if (P.epochs[pos].t_start == t) return P.epochs[pos];
    if (P.epochs[pos].t_start > t) break;
  }
  Epoch e;
  e.t_start = t;
  P.epochs.insert(P.epochs.begin() + pos, e);
  return P.epochs[pos];
}

// Populations exist in command-line order: -I declares the first ones, every -es adds one (Model::addPopulation,
// model.cc:672-686).  Tables are sized for the final number from the start; an option that addresses "all populations"
// touches the ones declared so far and leaves the later ones as "not given" (NaN), which is what scrm's
// addPopToVectorList / addPopToMatrixList produce.
thread_local int g_cur_pops = 1;   // populations in existence while the scrm flags are read (-I, then -es adds one)

void check_pop(const Params& P, int pop, const std::string& flag) {
  if (pop < 0 || pop >= g_cur_pops) throw std::invalid_argument("Invalid population in option " + flag);
}

// Model::addPopulationSizes / addPopulationSize (model.cc:202-281); pop < 0: all populations
void add_size_change(Params& P, double t_scaled, int pop, double rel_size, double count) {
  if (!(rel_size * P.N0 > 0.0)) throw std::invalid_argument("population size <= 0");
  Epoch& e = epoch_at(P, t_scaled * 4 * P.N0);
  if (pop < 0) {
    e.pop_size.assign(P.npop, kNaN);
    e.vb_count.assign(P.npop, kNaN);
    for (int p = 0; p < g_cur_pops; p++) { e.pop_size[p] = rel_size * P.N0; e.vb_count[p] = count; }
  } else {
    if (e.pop_size.empty()) { e.pop_size.assign(P.npop, kNaN); e.vb_count.assign(P.npop, kNaN); }
    e.pop_size[pop] = rel_size * P.N0;
    e.vb_count[pop] = count;
  }
}

// Model::addMigrationRates / addMigrationRate / addSymmetricMigration (model.cc:370-459); M = 4 N0 m
void add_mig_all(Params& P, double t_scaled, double M, double count = 1e10) {
  Epoch& e = epoch_at(P, t_scaled * 4 * P.N0);
  e.mig.assign((size_t)P.npop * P.npop, kNaN);
  e.mig_vb.assign((size_t)P.npop * P.npop, kNaN);
  if (!std::isnan(M))
    for (int i = 0; i < g_cur_pops; i++)
      for (int j = 0; j < g_cur_pops; j++) {
        e.mig[(size_t)i * P.npop + j] = M / (4 * P.N0);
        e.mig_vb[(size_t)i * P.npop + j] = count;
      }
}
void add_mig_one(Params& P, double t_scaled, int i, int j, double M, double count = 1e10) {
  Epoch& e = epoch_at(P, t_scaled * 4 * P.N0);
  if (e.mig.empty()) { e.mig.assign((size_t)P.npop * P.npop, kNaN); e.mig_vb.assign((size_t)P.npop * P.npop, kNaN); }
  e.mig[(size_t)i * P.npop + j] = M / (4 * P.N0);
  e.mig_vb[(size_t)i * P.npop + j] = count;
}
// Model::addSingleMigrationEvent (model.cc:461-480)
void add_single_mig(Params& P, double t_scaled, int i, int j, double fraction) {
  if (t_scaled < 0.0) throw std::invalid_argument("Single migration event: Negative time");
  if (fraction < 0.0 || fraction > 1.0) throw std::invalid_argument("Single migration event: Fraction out of range");
  Epoch& e = epoch_at(P, t_scaled * 4 * P.N0);
  if (e.single_mig.empty()) e.single_mig.assign((size_t)P.npop * P.npop, 0.0);
  e.single_mig[(size_t)i * P.npop + j] = fraction;
  P.has_migration = true;
}

}  // namespace

void parse(int argc, char** argv, Params& P) {
  Args a;
  for (int k = 1; k < argc; k++) a.v.push_back(argv[k]);
  P.argv = a.v;
  if (a.v.empty()) { P.help = true; return; }
  std::vector<std::string> scrm;
  for (; a.more(); a.i++) {
    const std::string f = a.cur();
    if (f == "-Np") P.N = (size_t)a.next_double(f);
    else if (f == "-nsam") P.nsam = (size_t)a.next_double(f);
    else if (f == "-ESS") {
      P.ESS_fraction = a.next_double(f); P.ESS_default = false;
      if (P.ESS_fraction > 1.0 || P.ESS_fraction < 0.0) throw std::invalid_argument("Out of range: -ESS " + a.cur());
    }
    else if (f == "-arg") unsupported(f, "posterior ARG output");
    else if (f == "-EM") P.EM_steps = (int)a.next_double(f);
    else if (f == "-chunks") {
      P.chunks = (int)a.next_double(f);
      if (P.chunks < 1 || P.chunks > 4096) throw std::invalid_argument("Out of range: -chunks " + a.cur());
    }
    else if (f == "-xr" || f == "-xc") {
      int last, first = read_range(a.next(f), last);
      last++;
      for (int i = 0; i < last; i++) {
        if ((int)P.record_event_in_epoch.size() <= i) P.record_event_in_epoch.push_back(3);
        if (i >= first) P.record_event_in_epoch[i] &= (f == "-xc") ? ~2 : ~1;
      }
    }
    else if (f == "-cap") { P.Ne_cap = a.next_double(f); P.useCap = true; }
    else if (f == "-tmax") P.top_t = a.next_double(f);
    else if (f == "-p") P.pattern = a.next(f);
    else if (f == "-seg") P.seg_file = a.next(f);
    else if (f == "-guide") unsupported(f, "recombination guide");
    else if (f == "-startpos") {
      P.start_position = a.next_double(f);
      if (P.start_position < 1) throw std::invalid_argument("Out of range: -startpos " + a.cur());
    }
    else if (f == "-lag") { P.lag = a.next_double(f); P.calibrate_lag = false; }
    else if (f == "-lags") {
      // (B200 engine) one lag per epoch, comma separated, bp: what the reference prints as "Lags set to:"
      std::string list = a.next(f), item;
      P.lags.clear();
      for (size_t q = 0; q <= list.size(); q++) {
        if (q == list.size() || list[q] == ',') {
          if (!item.empty()) {
            char* end = nullptr;
            const double v = strtod(item.c_str(), &end);
            if (*end != 0 || !(v > 0)) throw std::invalid_argument("Out of range: -lags " + list);
            P.lags.push_back(v);
          }
          item.clear();
        } else item += list[q];
      }
      P.calibrate_lag = false;
    }
    else if (f == "-calibrate_lag") {
      P.calibrate_lag = true; P.lag_fraction = a.next_double(f);
      if (P.lag_fraction < 0.0) throw std::invalid_argument("Out of range: -calibrate_lag " + a.cur());
    }
    else if (f == "-delay") {
      // the delay of an importance weight only exists with -bias_* / -guide, which the engine rejects: accepted, no effect
      (void)a.next_double(f);
      std::cerr << "Note: -delay has no effect (importance-weight delays belong to -bias_heights / -guide, which the B200 "
                   "engine does not support yet)" << std::endl;
    }
    else if (f == "-delay_coal" || f == "-delay_migr")
      std::cerr << "Note: " << f << " has no effect (importance-weight delays belong to -bias_heights / -guide, which the "
                   "B200 engine does not support yet)" << std::endl;
    else if (f == "-ancestral_aware") P.ancestral_aware = true;
    else if (f == "-dephase") P.dephase = true;
    else if (f == "-apf") { if ((int)a.next_double(f) != 0) unsupported(f, "auxiliary particle filter"); }
    else if (f == "-o") P.out_prefix = a.next(f);
    else if (f == "-log") {}
    else if (f == "-record_ess") P.record_resample_file = true;
    else if (f == "-dump_model") P.dump_model = true;
    else if (f == "-h" || f == "-help") P.help = true;
    else if (f == "-v" || f == "-version") P.version = true;
    else scrm.push_back(f);
  }
  if (P.help || P.version) return;

  // ---- the scrm part (param.cc), options in the order given: like scrm, -N0 scales what FOLLOWS it (theta, times,
  // sizes, migration rates), so a command line that names it late is read the way the reference reads it ----
  // -I fixes the number of populations; every table row is sized by it, so it is read ahead of the other options
  // (scrm itself needs -I to come first for the same reason, param.cc:148-160)
  g_cur_pops = 1;
  for (size_t k = 0; k + 1 < scrm.size(); k++)
    if (scrm[k] == "-I") {
      P.npop = (int)to_double(scrm[k + 1], "-I");
      if (P.npop < 1) throw std::invalid_argument("Failed to parse option: -I");
      g_cur_pops = P.npop;
    }
  for (const std::string& f : scrm) if (f == "-es") P.npop++;      // every admixture event adds a population
  if (P.npop > 4) throw std::invalid_argument("-I / -es: the B200 engine takes 1 to 4 populations");
  {
    Epoch e0;
    e0.t_start = 0.0;
    P.epochs.push_back(e0);
  }
  if (!P.pattern.empty())
    for (double t : expand_pattern(P.pattern, P.top_t)) {
      char buf[64];
      snprintf(buf, sizeof buf, "%f", t);                       // the reference round-trips through to_string()
      scrm.push_back("-eN"); scrm.push_back(buf); scrm.push_back("1");
    }
  Args s;
  s.v = scrm;
  bool have_t = false, have_r = false;
  double min_time = 0.0;
  size_t L_cur = 1;                            // the default Model's locus length (model.cc:31)
  bool n0_used = false;                        // Model::default_pop_size() was consulted (scaling_factor_ > 0)
  double mu_bits = 0.0, rho_bits = 0.0;
  auto time_of = [&](const std::string& f, bool timed) {
    double t = timed ? s.next_double(f) : 0.0;
    if (t < min_time)
      throw std::invalid_argument("If you use '" + f + "' in a model with population merges ('-es'), then you need to sort both arguments by time.");
    return t;
  };
  auto next_pop = [&](const std::string& f) {
    int pop = (int)s.next_double(f) - 1;
    check_pop(P, pop, f);
    return pop;
  };
  for (; s.more(); s.i++) {
    const std::string f = s.cur();
    if (f != "-N0" && f != "-vb" && f != "-VB" && f != "-seed" && f != "--seed" && f != "-seeds" && f != "--seeds" &&
        f != "-l" && f != "-I")
      n0_used = true;                            // every other option scales a rate, a size or a time with N0
    if (f == "-t") {
      // Model::setMutationRate(theta, per_locus, scaled) at the locus length known so far (model.cc:749-759)
      P.theta = s.next_double(f); have_t = true;
      double rate = P.theta;
      rate /= 4.0 * P.N0;
      mu_bits = rate / (double)L_cur;
    }
    else if (f == "-r") {
      // Model::setLocusLength rescales the per-base mutation rate, then setRecombinationRate (param.cc:102-106,
      // model.h:455-462): reproduced operation for operation so that the rates carry the reference's last bits
      P.rho_per_locus = s.next_double(f); P.loci_length = s.next_double(f); have_r = true;
      const size_t L_new = (size_t)P.loci_length;
      if (L_new > 1) {
        mu_bits *= (double)L_cur / L_new;
        L_cur = L_new;
        double rate = P.rho_per_locus;
        rate /= 4.0 * P.N0;
        rho_bits = rate / (double)(L_cur - 1);
      }
    }
    else if (f == "-N0") {
      // Model::set_default_pop_size(int) (model.h:85-88): an integer, and only before anything was scaled with it
      const double v = s.next_double(f);
      if (n0_used) throw std::invalid_argument("Default population size (-N0) must be set before other population-related parameters");
      P.N0 = (double)(int)v;
    }
    else if (f == "-vb" || f == "-VB") {
      if (P.epochs.size() > 1) throw std::invalid_argument("-vb must occur before -eN / -en / -eM / -ema");
      P.vb = true;
    }
    else if (f == "-eN" || f == "-N") {
      double t = time_of(f, f == "-eN");
      double size = s.next_double(f);
      double count = P.vb ? s.next_double(f) : 1e10;
      add_size_change(P, t, -1, size, count);
    }
    else if (f == "-en" || f == "-n") {
      double t = f == "-en" ? s.next_double(f) : 0.0;
      int pop = next_pop(f);
      double size = s.next_double(f);
      double count = P.vb ? s.next_double(f) : 1e10;
      add_size_change(P, t, pop, size, count);
    }
    else if (f == "-seed" || f == "--seed" || f == "-seeds" || f == "--seeds") {
      P.seed = (uint64_t)s.next_double(f); P.seed_set = true;
      for (int k = 0; k < 2 && s.i + 1 < s.v.size(); k++) {      // up to three seeds (ms compatibility)
        char* end;
        std::strtod(s.v[s.i + 1].c_str(), &end);
        if (*end == '\0' && s.v[s.i + 1][0] != '-') s.i++; else break;
      }
    }
    else if (f == "-l") { (void)s.next(f); }                    // exact window: the driver always uses 0 (pfparam.cpp:315)
    else if (f == "-I") {
      (void)s.next(f);
      P.leaf_pop.clear();
      for (int p = 0; p < g_cur_pops; p++) {
        int cnt = (int)s.next_double(f);
        if (cnt < 0) throw std::invalid_argument("Failed to parse option: -I");
        for (int k = 0; k < cnt; k++) P.leaf_pop.push_back(p);
      }
      // an optional symmetric migration rate may follow (param.cc:156-160)
      if (s.i + 1 < s.v.size()) {
        char* end = nullptr;
        double M = std::strtod(s.v[s.i + 1].c_str(), &end);
        if (end != s.v[s.i + 1].c_str() && *end == '\0') {
          s.i++;
          n0_used = true;
          if (g_cur_pops > 1) { add_mig_all(P, 0.0, M / (g_cur_pops - 1)); if (M > 0) P.has_migration = true; }
        }
      }
    }
    else if (f == "-ma" || f == "-ema") {
      double t = time_of(f, f == "-ema");
      std::vector<double> M((size_t)P.npop * P.npop, 0.0), C((size_t)P.npop * P.npop, 1e10);
      for (int i = 0; i < g_cur_pops; i++)
        for (int j = 0; j < g_cur_pops; j++) {
          if (i == j) { (void)s.next(f); if (P.vb) (void)s.next(f); continue; }
          M[(size_t)i * P.npop + j] = s.next_double(f);
          if (P.vb) C[(size_t)i * P.npop + j] = s.next_double(f);
        }
      add_mig_all(P, t, 0.0);
      for (int i = 0; i < g_cur_pops; i++)
        for (int j = 0; j < g_cur_pops; j++)
          if (i != j) {
            add_mig_one(P, t, i, j, M[(size_t)i * P.npop + j], C[(size_t)i * P.npop + j]);
            if (M[(size_t)i * P.npop + j] > 0) P.has_migration = true;
          }
    }
    else if (f == "-m" || f == "-em") {
      double t = time_of(f, f == "-em");
      int i = next_pop(f), j = next_pop(f);
      double M = s.next_double(f);
      add_mig_one(P, t, i, j, M);
      if (M > 0) P.has_migration = true;
    }
    else if (f == "-M" || f == "-eM") {
      double t = time_of(f, f == "-eM");
      if (g_cur_pops < 2) throw std::invalid_argument(f + " needs more than one population (-I)");
      double M = s.next_double(f) / (g_cur_pops - 1);
      double count = P.vb ? s.next_double(f) / (g_cur_pops - 1) : 1e10;     // param.cc:288-290
      add_mig_all(P, t, M, count);
      if (M > 0) P.has_migration = true;
    }
    else if (f == "-ej") {
      double t = time_of(f, true);
      min_time = t;
      int i = next_pop(f), j = next_pop(f);
      add_single_mig(P, t, i, j, 1.0);
      for (int k = 0; k < g_cur_pops; k++)
        if (k != i) add_mig_one(P, t, k, i, 0.0);               // nobody migrates into the dissolved population
    }
    else if (f == "-eps") {
      double t = time_of(f, true);
      min_time = t;
      int i = next_pop(f), j = next_pop(f);
      double fraction = s.next_double(f);
      if (fraction < 0.0) throw std::invalid_argument("Probability in `-eps` argument is negative.");
      if (fraction > 1.0) throw std::invalid_argument("Probability in `-eps` argument greater than one.");
      add_single_mig(P, t, i, j, 1 - fraction);
    }
    else if (f == "-es") {                                      // param.cc:299-317: admixture into a NEW population
      double t = time_of(f, true);
      min_time = t;
      int i = next_pop(f);
      double fraction = s.next_double(f);
      if (fraction < 0.0) throw std::invalid_argument("Probability in `-es` argument is negative.");
      if (fraction > 1.0) throw std::invalid_argument("Probability in `-es` argument greater than one.");
      const int fresh = g_cur_pops++;
      add_single_mig(P, t, i, fresh, 1 - fraction);
    }
    else if (f == "-eI") unsupported(f, "ancient samples");
    else if (f == "-G" || f == "-eG" || f == "-g" || f == "-eg") unsupported(f, "exponential growth");
    else if (f == "-bias_heights" || f == "-bias_strengths") unsupported(f, "biased recombination sampling");
    else if (f == "-sr" || f == "-st") unsupported(f, "position-dependent rates");
    else throw std::invalid_argument("unknown/unexpected argument: " + f);
  }
  if (!have_r) throw std::invalid_argument("The option -r is required");
  if (!have_t) throw std::invalid_argument("The option -t is required");
  if (P.loci_length <= 1) throw std::invalid_argument("Locus length must be at least two");
  if (P.epochs.back().t_start >= P.top_t * 40000)
    throw std::invalid_argument("Problem: -tmax must be larger than bottom of final epoch");
  while (P.record_event_in_epoch.size() < P.epochs.size()) P.record_event_in_epoch.push_back(3);
  if (P.record_event_in_epoch.size() > P.epochs.size())
    throw std::invalid_argument("Problem: epochs specified in -xr/-xc options out of range");
  if (P.nsam < 2 || P.nsam > 32) throw std::invalid_argument("-nsam must be between 2 and 32 for the B200 engine");
  if (!P.seed_set) { P.seed = (uint64_t)time(nullptr); }
  if (P.leaf_pop.empty()) P.leaf_pop.assign(P.nsam, 0);
  if (P.leaf_pop.size() != P.nsam)
    throw std::invalid_argument("-I: the sample sizes of the populations do not add up to -nsam");
  finalize_model(P);
  if (P.npop > 1 && !P.has_migration)      // Model::check (model.cc:630-633)
    throw std::invalid_argument("Model has multiple populations but no migration. Coalescence impossible");
  finalize_rates(P);
  P.mutation_rate = mu_bits;               // the same numbers as finalize_rates, with the reference's rounding
  P.recombination_rate = P.original_recombination_rate = rho_bits;
}

// Model::finalize (model.cc:553-670): sizes not given at a change time continue from the epoch before (default N0),
// migration rates likewise (default 0); fixed-time moves hold for their own epoch start only
void finalize_model(Params& P) {
  const size_t np = (size_t)P.npop;
  const std::vector<double>* last_mig = nullptr;
  const std::vector<double>* last_vb = nullptr;
  for (size_t e = 0; e < P.epochs.size(); e++) {
    Epoch& ep = P.epochs[e];
    if (ep.pop_size.empty()) { ep.pop_size.assign(np, kNaN); ep.vb_count.assign(np, kNaN); }
    for (size_t p = 0; p < np; p++) {
      if (std::isnan(ep.pop_size[p])) ep.pop_size[p] = e == 0 ? P.N0 : P.epochs[e - 1].pop_size[p];
      if (std::isnan(ep.vb_count[p])) ep.vb_count[p] = 1e10;
    }
    if (ep.mig.empty()) {
      if (last_mig) { ep.mig = *last_mig; ep.mig_vb = *last_vb; }
      else { ep.mig.assign(np * np, 0.0); ep.mig_vb.assign(np * np, 1e10); }
    } else {
      for (size_t k = 0; k < np * np; k++) {
        if (std::isnan(ep.mig[k])) ep.mig[k] = last_mig ? (*last_mig)[k] : 0.0;
        // Model::finalize does not fill the event counts of entries that were not given (they stay NaN and are never
        // read for a rate of zero); here they fall back to "no correction"
        if (std::isnan(ep.mig_vb[k])) ep.mig_vb[k] = last_vb ? (*last_vb)[k] : 1e10;
      }
    }
    for (size_t p = 0; p < np; p++) ep.mig[p * np + p] = 0.0;
    last_mig = &ep.mig;
    last_vb = &ep.mig_vb;
    if (ep.single_mig.empty()) ep.single_mig.assign(np * np, 0.0);
  }
}

void finalize_rates(Params& P) {
  // Model::setMutationRate / setRecombinationRate with per_locus = scaled = true (model.cc:718-759)
  P.mutation_rate = P.theta / (4.0 * P.N0) / P.loci_length;
  P.recombination_rate = P.rho_per_locus / (4.0 * P.N0) / (P.loci_length - 1);
  P.original_recombination_rate = P.recombination_rate;
}

}  // namespace host
