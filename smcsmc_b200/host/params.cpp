// Command-line parser: PfParam::parse (pfparam.cpp:51-169) + the subset of scrm's Param::parse (param.cc:37-520) that
// the single-population, piecewise-constant engine supports.
#include "params.hpp"
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

void add_size_change(Params& P, double t_scaled, double rel_size, double count) {
  double t = t_scaled * 4 * P.N0;
  for (auto& e : P.epochs)
    if (e.t_start == t) { e.pop_size = rel_size * P.N0; e.vb_count = count; return; }
  // a new change time inherits nothing: scrm's addPopulationSizes sets the size from this time onwards until the
  // next explicit change; sizes of later, already-declared epochs stay as declared
  P.epochs.push_back({t, rel_size * P.N0, count});
  std::sort(P.epochs.begin(), P.epochs.end(), [](const Epoch& a, const Epoch& b) { return a.t_start < b.t_start; });
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
    else if (f == "-calibrate_lag") {
      P.calibrate_lag = true; P.lag_fraction = a.next_double(f);
      if (P.lag_fraction < 0.0) throw std::invalid_argument("Out of range: -calibrate_lag " + a.cur());
    }
    else if (f == "-delay") { (void)a.next_double(f); }        // only used with -bias_* / -guide (not supported): no effect
    else if (f == "-delay_coal" || f == "-delay_migr") {}
    else if (f == "-ancestral_aware") P.ancestral_aware = true;
    else if (f == "-dephase") unsupported(f, "dephasing");
    else if (f == "-apf") { if ((int)a.next_double(f) != 0) unsupported(f, "auxiliary particle filter"); }
    else if (f == "-o") P.out_prefix = a.next(f);
    else if (f == "-log") {}
    else if (f == "-record_ess") P.record_resample_file = true;
    else if (f == "-h" || f == "-help") P.help = true;
    else if (f == "-v" || f == "-version") P.version = true;
    else scrm.push_back(f);
  }
  if (P.help || P.version) return;

  // ---- the scrm part (param.cc): two passes so that -N0 applies wherever it stands ----
  for (size_t k = 0; k + 1 < scrm.size(); k++)
    if (scrm[k] == "-N0") P.N0 = to_double(scrm[k + 1], "-N0");
  P.epochs.push_back({0.0, P.N0, 1e10});
  if (!P.pattern.empty())
    for (double t : expand_pattern(P.pattern, P.top_t)) {
      char buf[64];
      snprintf(buf, sizeof buf, "%f", t);                       // the reference round-trips through to_string()
      scrm.push_back("-eN"); scrm.push_back(buf); scrm.push_back("1");
    }
  Args s;
  s.v = scrm;
  bool have_t = false, have_r = false;
  for (; s.more(); s.i++) {
    const std::string f = s.cur();
    if (f == "-t") { P.theta = s.next_double(f); have_t = true; }
    else if (f == "-r") { P.rho_per_locus = s.next_double(f); P.loci_length = s.next_double(f); have_r = true; }
    else if (f == "-N0") { (void)s.next(f); }
    else if (f == "-vb" || f == "-VB") {
      if (P.epochs.size() > 1) throw std::invalid_argument("-vb must occur before -eN / -en / -eM / -ema");
      P.vb = true;
    }
    else if (f == "-eN" || f == "-N") {
      double t = f == "-eN" ? s.next_double(f) : 0.0;
      double size = s.next_double(f);
      double count = P.vb ? s.next_double(f) : 1e10;
      add_size_change(P, t, size, count);
    }
    else if (f == "-en" || f == "-n") {
      double t = f == "-en" ? s.next_double(f) : 0.0;
      int pop = (int)s.next_double(f);
      if (pop != 1) unsupported(f, "more than one population");
      double size = s.next_double(f);
      double count = P.vb ? s.next_double(f) : 1e10;
      add_size_change(P, t, size, count);
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
      int npop = (int)s.next_double(f);
      if (npop != 1) unsupported(f, "structured populations");
      (void)s.next(f);
    }
    else if (f == "-eI" || f == "-M" || f == "-eM" || f == "-m" || f == "-em" || f == "-ma" || f == "-ema" ||
             f == "-es" || f == "-ej" || f == "-eps") unsupported(f, "migration / population splits");
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
  finalize_rates(P);
}

void finalize_rates(Params& P) {
  // Model::setMutationRate / setRecombinationRate with per_locus = scaled = true (model.cc:718-759)
  P.mutation_rate = P.theta / (4.0 * P.N0) / P.loci_length;
  P.recombination_rate = P.rho_per_locus / (4.0 * P.N0) / (P.loci_length - 1);
  P.original_recombination_rate = P.recombination_rate;
}

}  // namespace host
