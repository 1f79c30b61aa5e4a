// Host side of the drop-in boundary: the reference's command line (PfParam, src/pfparam.cpp:51-340, forwarding the
// rest to scrm's Param, src/scrm/src/param.cc:37-520) parsed into the flat model the engine takes.
// Same flags, same defaults, same error behaviour (exceptions -> "Error: ..." on stderr, exit 1).  Flags whose
// features the B200 engine does not implement yet are recognised and rejected with a clear message instead of being
// silently ignored.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <ctime>
#include <stdexcept>
#include <string>
#include <vector>

namespace host {

// One row of scrm's Model tables (model.h:55: change_times_, pop_sizes_list_, mig_rates_list_, single_mig_probs_list_).
// While the command line is parsed NaN means "as in the previous epoch" and an empty vector "not mentioned at this
// time"; finalize_model() resolves both the way Model::finalize does (model.cc:553-670).
struct Epoch {
  double t_start;                  // generations
  std::vector<double> pop_size;    // [P] absolute N_e
  std::vector<double> vb_count;    // [P] -vb event counts (1e10 = no correction)
  std::vector<double> mig;         // [P*P] per-generation rate of a lineage in i moving to j (backwards in time)
  std::vector<double> mig_vb;      // [P*P] -vb event counts of the migration rates (1e10 = no correction)
  std::vector<double> single_mig;  // [P*P] probability of being moved i -> j at the START of the epoch (-ej / -eps)
};

struct Params {
  // smcsmc flags (defaults: pfparam.cpp:193-255)
  size_t N = 100;
  size_t nsam = 2;
  double ESS_fraction = 0.5;
  bool ESS_default = true;
  int EM_steps = 0;
  int chunks = 1;                 // extension: sweep this many equal windows of the locus concurrently on the device
  double lag = 0.0;
  bool calibrate_lag = true;
  double lag_fraction = 2.0;
  std::vector<double> lags;                 // -lags: explicit lag per epoch (B200 engine extension)
  bool ancestral_aware = false;
  bool dephase = false;           // -dephase: heterozygous genotypes are summed over both phasings
  double Ne_cap = 200000;
  bool useCap = false;
  double top_t = 2.0;
  std::string pattern;
  std::string seg_file;
  std::string out_prefix = "smcsmc";
  double start_position = 1;
  bool record_resample_file = false;
  bool help = false, version = false;
  std::vector<int> record_event_in_epoch;   // bit0 recomb, bit1 coal/migr (pfparam.hpp RECORD_*)
  // scrm flags
  double N0 = 10000;
  double theta = -1, rho_per_locus = -1;
  double loci_length = 2e7;
  uint64_t seed = 0;
  bool seed_set = false;
  bool vb = false;
  std::vector<Epoch> epochs;                // sorted by time, epochs[0].t_start == 0
  int npop = 1;                             // -I
  std::vector<int> leaf_pop;                // [nsam] Model::sample_population (0-based)
  bool has_migration = false;               // Model::has_migration()
  bool dump_model = false;                  // -dump_model (extension): print the flattened tables and exit
  int marks_scale = 1;                      // doubled by the driver when a sweep ran out of migration-mark slots
  std::vector<std::string> argv;
  // derived
  double mutation_rate = 0, recombination_rate = 0, original_recombination_rate = 0;

  int num_epochs() const { return (int)epochs.size(); }
};

inline double to_double(const std::string& s, const std::string& flag) {
  char* end = nullptr;
  double v = std::strtod(s.c_str(), &end);
  if (end == s.c_str() || *end != '\0') throw std::invalid_argument("Failed to parse option: " + flag + " " + s);
  return v;
}

// "-p" pattern -> epoch boundaries (src/pattern.cpp:45-162): PSMC-style "a*b+c*d+e", log-spaced in [0, tmax]
inline std::vector<double> expand_pattern(const std::string& pattern, double top_t) {
  std::vector<size_t> l1, l2;
  size_t num_seg = 0;
  const char* p = pattern.c_str();
  auto number = [&]() -> size_t {
    if (!isdigit(*p)) throw std::invalid_argument(std::string("Pattern: digits expected at ") + p);
    char* end;
    size_t v = strtol(p, &end, 10);
    p = end;
    return v;
  };
  auto factor = [&]() -> size_t {
    size_t a = number();
    if (*p == '+' || *p == '\0') { l1.push_back(1); l2.push_back(a); return a; }
    if (*p != '*') throw std::invalid_argument(std::string("Pattern: * expected at ") + p);
    p++;
    size_t b = number();
    l1.push_back(a); l2.push_back(b);
    if (*p != '+' && *p != '\0') throw std::invalid_argument(std::string("Pattern: + expected at ") + p);
    return a * b;
  };
  num_seg = factor();
  while (*p) { p++; num_seg += factor(); }
  std::vector<double> out;
  if (num_seg < 2) return out;
  std::vector<double> t(num_seg);
  for (size_t i = 0; i < num_seg; i++)
    t[i] = 0.1 * std::exp((double)i / (double)(num_seg - 1) * std::log(1 + 10 * top_t)) - 0.1;
  size_t index = 0;
  for (size_t s = 0; s < l1.size(); s++)
    for (size_t i = 0; i < l1[s]; i++) { out.push_back(t[index] / 2); index += l2[s]; }
  return out;   // times in units of 4*N0, each becomes "-eN t 1"
}

void parse(int argc, char** argv, Params& P);      // params.cpp
void finalize_rates(Params& P);
void finalize_model(Params& P);                    // Model::finalize (model.cc:553-670)

}  // namespace host
