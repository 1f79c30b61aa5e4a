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

struct Epoch {
  double t_start;     // generations
  double pop_size;    // absolute N_e
  double vb_count;    // -vb event count (1e10 = no correction)
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
  bool ancestral_aware = false;
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

}  // namespace host
