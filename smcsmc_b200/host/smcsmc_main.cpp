// smcsmc -- drop-in host driver for the B200 particle-filter engine.
//
// Same surface as the reference executable (src/smcsmc.cpp:46-103 main, :278-400 pfARG_core): scrm-style command line,
// .seg input, <prefix>.out / .log / .recomb.gz / .resample outputs in the reference's formats, in-binary EM
// (CountModel::reset_model_parameters, count.cpp:44-63,267-325), exceptions -> "Error: ..." + exit 1.  Everything
// between "read the rows" and "write the tallies" is the CUDA engine behind include/smcsmc_b200.h.
#include <algorithm>
#include <zlib.h>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <thread>
#include "../../include/smcsmc_b200.h"
#include "params.hpp"
#include "segment.hpp"

using namespace std;
using host::Params;

#ifndef SMCSMC_B200_VERSION
#define SMCSMC_B200_VERSION "b200-0.1"
#endif

namespace {

void check(int rc) {
  if (rc != 0) throw runtime_error(string("engine: ") + smc_last_error());
}

// a particle needed more migration marks than the engine was given room for: the sweep is repeated with twice the room
struct MarksOverflow : runtime_error { MarksOverflow() : runtime_error("migration marks") {} };

// particle.cpp:65-74
double exp_digamma(double x) {
  if (x > 10) return x - 0.5 + (x + 0.5) / (24 * x * x);
  double f = 0.0;
  while (x < 6) { f = f + 1.0 / x; x = x + 1.0; }
  double psi = log(x) - 1 / (2 * x) - 1 / (12 * x * x);
  return exp(psi - f);
}

// FormatDouble (pfparam.cpp:482-497)
string fmt(double d, double scientific_bound = 0.1, int precision = 2) {
  const int field_length = 14;
  const double maxdouble = exp((field_length - precision - 1) * log(10.0));
  ostringstream o;
  if (d < maxdouble && (d > scientific_bound || d == 0.0)) o << setw(field_length) << fixed << setprecision(precision) << d;
  else o << setw(field_length) << scientific << setprecision(field_length - 7) << d;
  return o.str();
}

struct Out {
  string out_name, log_name, recomb_name, resample_name;
  void header() const {   // PfParam::outFileHeader (pfparam.cpp:459-479)
    ofstream f(out_name.c_str(), ios::binary);
    f << setw(6) << "Iter" << " " << setw(6) << "Epoch" << " " << setw(14) << "Start" << " " << setw(14) << "End" << " "
      << setw(6) << "Type" << " " << setw(6) << "From" << " " << setw(6) << "To" << " " << setw(14) << "Opp" << " "
      << setw(14) << "Count" << " " << setw(14) << "Rate" << " " << setw(14) << "Ne" << " " << setw(14) << "ESS" << endl;
  }
  void row(size_t em, int epoch, double b, double e, const string& type, int from, int to, double opp, double count,
           double weight) const {   // PfParam::appendToOutFile (pfparam.cpp:500-527)
    ofstream f(out_name.c_str(), ios::out | ios::app | ios::binary);
    f << setw(6) << em << " " << setw(6) << epoch << " " << fmt(b) << " " << fmt(e) << " " << setw(6) << type << " "
      << setw(6) << from << " " << setw(6) << to << " " << fmt(opp) << " " << fmt(count) << " "
      << fmt(count / (opp + 1e-10)) << " " << fmt(type == "Coal" ? (opp + 1e-10) / (2.0 * count) : 0.0) << " "
      << fmt(1.0 / (weight / opp + 1e-10), 1.0, 3) << endl;
  }
};

void write_log(ostream& w, const Params& P, const Out& o) {   // PfParam::writeLog (pfparam.cpp:403-456)
  w << "###########################\n#        smcsmc log       #\n###########################\n";
  w << "Program was compiled on: " << __DATE__ << endl;
  w << "smcsmc version: " << SMCSMC_B200_VERSION << endl;
  w << "scrm version:   " << "none (B200 engine: " << smc_version() << ")" << endl;
  w << "smcsmc parameters: \n";
  w << "Segment Data file: " << (P.seg_file.empty() ? "empty" : P.seg_file.c_str()) << "\n";
  w << "Recombination bias file: None\n";
  w << setw(15) << " EM steps =" << setw(10) << P.EM_steps << "\n";
  if (P.lag > 0) w << setw(15) << "Constant lag =" << setw(10) << P.lag << "\n";
  w << setw(15) << "N =" << setw(10) << P.N << "\n";
  w << setw(15) << "ESS =" << setw(10) << P.ESS_fraction;
  if (P.ESS_default) w << " (by default)";
  w << "\n";
  w << "scrm model parameters: \n";
  w << setw(17) << "Extract window =" << setw(10) << 0 << "\n";
  w << setw(17) << "Extract window R =" << setw(10) << 500 << "\n";
  w << setw(17) << "Sample size =" << setw(10) << P.nsam << "\n";
  w << setw(17) << "Seq length =" << setw(10) << (size_t)P.loci_length << "\n";
  w << setw(17) << "mutation rate =" << setw(10) << P.mutation_rate << "\n";
  w << setw(17) << "recomb rate =" << setw(10) << P.original_recombination_rate << "\n";
  w << setw(17) << "inferred recomb rate = " << setw(10) << P.recombination_rate << "\n";
  w << setw(17) << "Pop size (at Generation):\n";
  for (const auto& e : P.epochs) {
    w << setw(3) << "(" << setw(8) << e.t_start << " )";
    for (double sz : e.pop_size) w << " | " << setw(10) << sz;
    w << "\n";
  }
  if (P.npop > 1) {
    w << setw(17) << "Migration rate (per generation, backwards in time; at Generation):\n";
    for (const auto& e : P.epochs) {
      w << setw(3) << "(" << setw(8) << e.t_start << " )";
      for (int i = 0; i < P.npop; i++) {
        w << " |";
        for (int j = 0; j < P.npop; j++) w << " " << setw(10) << e.mig[(size_t)i * P.npop + j];
      }
      w << "\n";
    }
  }
  w << "Out file is saved in file: " << o.out_name << "\n";
}

void print_help(const Params& P) {
  cout << "smcsmc (B200 engine) -- command line of luntergroup/smcsmc -- Version " << SMCSMC_B200_VERSION << endl;
  auto opt = [](const char* f, const char* t, const string& d) {
    cout << setw(15) << f << setw(8) << t << "  --  " << d << endl;
  };
  cout << "Options:" << endl;
  opt("-Np", "INT", "Number of particles [ " + to_string(P.N) + " ]");
  opt("-seg", "STR", "Data file in seg format [ Chrom1.seg ]");
  opt("-o", "STR", "Prefix for output files");
  opt("-EM", "INT", "EM iterations [ 20 ]");
  opt("-chunks", "INT", "(B200 engine) sweep INT equal windows of the locus concurrently, tallies added [ 1 ]");
  opt("-startpos", "INT", "First nucleotide position to analyze [ 1 ]");
  opt("-nsam", "INT", "Number of haplotypes (2..32)");
  opt("-log", " ", "Generate *.log file");
  opt("-v", " ", "Display version");
  cout << endl << "Inference tuning:" << endl;
  opt("-calibrate_lag", "FLT", "Lag before extracting events (multiple of survival time) [ 2 ]");
  opt("-lag", "FLT", "Constant lag (bp) instead of calibration");
  opt("-lags", "F,F,..", "(B200 engine) one lag (bp) per epoch instead of calibration");
  opt("-tmax", "FLT", "Maximum tree height, in unit of 4N0 [ 2 ]");
  opt("-p", "STR", "Pattern of time segments, e.g. 1*3+15*4+1");
  opt("-ESS", "FLT", "Fractional ESS threshold for resampling [ 0.5 ]");
  opt("-xr", "INT", "Epoch or epoch range to exclude from recombination EM (0-based, closed)");
  opt("-xc", "INT", "Epoch or epoch range (e.g. 0-10) to exclude from coalescent EM");
  opt("-cap", "FLT", "Cap on inferred population sizes");
  opt("-ancestral_aware", " ", "Alleles are polarised (0 = ancestral)");
  opt("-dephase", " ", "Dephase heterozygous sites");
  opt("-record_ess", " ", "Generate *.resample file");
  cout << endl << "Model (scrm syntax): -t THETA  -r RHO LENGTH  -N0 N  -eN T SIZE  -en T POP SIZE  -vb  -seed S" << endl;
  cout << "  -I NPOP n1..nP [M]  -eM T M  -em T i j M  -ema T M11..  -ej T i j  -eps T i j f  -es T i f   (up to 4 populations)" << endl;
  cout << "Not supported by this engine yet (rejected with an error): -eI, -eG (growth), -guide," << endl
       << "  -bias_heights/-bias_strengths, -apf > 0, -arg" << endl;
}

// the flattened Model the engine takes: [E] times, [E][P] 1/(2N), [E][P][P] migration rates and fixed-time moves
struct FlatModel {
  int E, np;
  vector<double> t_start, inv2N, mig, single_mig;
  vector<int32_t> leaf_pop;
  explicit FlatModel(const Params& P) : E(P.num_epochs()), np(P.npop) {
    for (const auto& e : P.epochs) {
      t_start.push_back(e.t_start);
      for (int p = 0; p < np; p++) {
        // the same roundings as the reference's own round trip: Model stores 1/(2N), population_size() returns
        // 0.5 / stored, the engine takes 1 / (2 * population_size()) (model.cc:268-281, model.h:167-169)
        const double stored = 1.0 / (2.0 * e.pop_size[p]);
        const double size = 0.5 / stored;
        inv2N.push_back(1.0 / (2.0 * size));
      }
      mig.insert(mig.end(), e.mig.begin(), e.mig.end());
      single_mig.insert(single_mig.end(), e.single_mig.begin(), e.single_mig.end());
    }
    for (int v : P.leaf_pop) leaf_pop.push_back(v);
  }
};

// -dump_model (extension, used by the CPU tests): the tables exactly as they are handed to the engine
void dump_model(const Params& P) {
  FlatModel M(P);
  cout << setprecision(17);
  cout << "npop " << M.np << "\nepochs " << M.E << "\nmutation_rate " << P.mutation_rate << "\nrecombination_rate "
       << P.recombination_rate << "\nloci_length " << P.loci_length << "\nleaf_pop";
  for (int v : M.leaf_pop) cout << " " << v;
  cout << "\n";
  for (int e = 0; e < M.E; e++) {
    cout << "epoch " << e << " t_start " << M.t_start[e] << " inv2N";
    for (int p = 0; p < M.np; p++) cout << " " << M.inv2N[(size_t)e * M.np + p];
    cout << " mig";
    for (int k = 0; k < M.np * M.np; k++) cout << " " << M.mig[(size_t)e * M.np * M.np + k];
    cout << " single_mig";
    for (int k = 0; k < M.np * M.np; k++) cout << " " << M.single_mig[(size_t)e * M.np * M.np + k];
    cout << "\n";
  }
}

// one EM iteration: pfARG_core (smcsmc.cpp:278-400)
// ---- <prefix>.recomb.gz rows (CountModel::dump_local_recomb_logs, count.cpp:630-653) -----------------------------------
// "%.5e" of a positive double, digits from integer arithmetic; anything near a rounding tie or out of the usual range goes
// through snprintf, so the text is what the reference's iostream prints
static int format_sci5(char* o, double v) {
  if (v == 0.0) { memcpy(o, "0.00000e+00", 11); return 11; }
  if (!(v > 1e-290 && v < 1e290)) return snprintf(o, 32, "%.5e", v);
  int ex = (int)floor(log10(v));
  double m = v / pow(10.0, ex);
  if (m >= 10.0) { m /= 10.0; ex++; }
  if (m < 1.0) { m *= 10.0; ex--; }
  const double scaled = m * 1e5;
  const double r = floor(scaled + 0.5);
  if (fabs(scaled + 0.5 - floor(scaled + 0.5)) < 1e-6 || fabs(scaled + 0.5 - floor(scaled + 0.5)) > 1.0 - 1e-6 || r >= 1e6 || ex < -99 || ex > 99)
    return snprintf(o, 32, "%.5e", v);
  long long d = (long long)r;
  o[0] = (char)('0' + d / 100000); o[1] = '.';
  for (int k = 5; k >= 1; k--) { o[1 + k] = (char)('0' + d % 10); d /= 10; }
  o[7] = 'e'; o[8] = ex < 0 ? '-' : '+';
  const int ax = ex < 0 ? -ex : ex;
  o[9] = (char)('0' + ax / 10); o[10] = (char)('0' + ax % 10);
  return 11;
}

static thread g_recomb_writer;
static string g_recomb_error;
static void recomb_writer_join() {
  if (g_recomb_writer.joinable()) g_recomb_writer.join();
  if (!g_recomb_error.empty()) { string m = g_recomb_error; g_recomb_error.clear(); throw runtime_error(m); }
}

static void write_recomb_table(string name, int em, int n, double start_position, double bin, size_t nbins, vector<double> diff) {
  gzFile g = gzopen(name.c_str(), "ab1");
  if (!g) { g_recomb_error = "cannot open " + name; return; }
  if (em == 0) {
    string h = "iter\tlocus\tsize\topp_per_nt";
    for (int s = 0; s < n; s++) h += "\t" + to_string(s + 1);
    h += "\ttime\tlog_time\n";
    gzwrite(g, h.data(), (unsigned)h.size());
  }
  string zeros;
  for (int s = 0; s < n + 2; s++) zeros += "\t0.00000e+00";
  zeros += "\n";
  char head[48], binsz[32];
  const int nh = snprintf(head, sizeof head, "%d\t", em), nb = snprintf(binsz, sizeof binsz, "\t%.0f\t", bin);
  string buf;
  buf.reserve((1u << 20) + 256);
  char num[40];
  double cur = 0.0;
  for (size_t i = 0; i < nbins; i++) {
    cur += diff[i];
    buf.append(head, nh);
    const double locus = i * bin + start_position;
    if (locus >= 0 && locus < 9e15 && locus == floor(locus)) {
      char* e_ = num + sizeof num; char* q = e_;
      unsigned long long u = (unsigned long long)locus;
      do { *--q = (char)('0' + u % 10); u /= 10; } while (u);
      buf.append(q, e_ - q);
    } else buf.append(num, snprintf(num, sizeof num, "%.0f", locus));
    buf.append(binsz, nb);
    buf.append(num, format_sci5(num, cur / bin));
    buf += zeros;
    if (buf.size() > (1u << 20)) { gzwrite(g, buf.data(), (unsigned)buf.size()); buf.clear(); }
  }
  if (!buf.empty()) gzwrite(g, buf.data(), (unsigned)buf.size());
  gzclose(g);
}

void pfARG_core(Params& P, const vector<host::SegRows>& chunks, const vector<double>& chunk_start, double chunk_len,
                const Out& out, int em, const vector<int>& devices) {
  const int device = devices[0];
  const int E = P.num_epochs(), n = (int)P.nsam, np = P.npop;
  const FlatModel M(P);
  const vector<double>& t_start = M.t_start;
  vector<double> inv2N0(E), vb(E), lags(E);
  vector<uint8_t> mask(E);
  for (int e = 0; e < E; e++) {
    inv2N0[e] = M.inv2N[(size_t)e * np];
    vb[e] = P.vb ? log(exp_digamma(P.epochs[e].vb_count[0]) / P.epochs[e].vb_count[0]) : 0.0;
    mask[e] = (uint8_t)(P.record_event_in_epoch[e] & 3);
  }
  // CountModel::init_lags (count.cpp:230-249), then reset_lag from the calibration (smcsmc.cpp:314-316)
  for (int e = 0; e < E; e++) {
    if (E == 1) lags[e] = 20000;
    else {
      double top_t = e == E - 1 ? t_start[E - 1] : t_start[e + 1];
      lags[e] = P.lag > 0 ? P.lag : 4.0 / (P.recombination_rate * top_t);
    }
  }
  if (!P.lags.empty()) {
    if ((int)P.lags.size() != E) throw invalid_argument("-lags needs one value per epoch (" + to_string(E) + ")");
    lags = P.lags;
  }
  if (P.calibrate_lag) {
    vector<double> surv(E);
    vector<int64_t> cnt(E);
    // The reference keeps simulating until every epoch holds 200 survival events (up to 10^6 trees, smcsmc.cpp:169-236);
    // here the batch of ARGs is quadrupled until that holds (rarely visited epochs: fine -p patterns, two samples)
    for (int64_t n_args = 8192;; n_args *= 4) {
      if (np > 1)
        check(smc_calibrate_survival_struct(device, n, E, t_start.data(), np, M.inv2N.data(), M.mig.data(),
                                            M.single_mig.data(), M.leaf_pop.data(), P.recombination_rate, chunk_len, 1,
                                            n_args, surv.data(), cnt.data()));
      else
        check(smc_calibrate_survival(device, n, E, t_start.data(), inv2N0.data(), P.recombination_rate, chunk_len, 1,
                                     n_args, surv.data(), cnt.data()));
      const int64_t fewest = *min_element(cnt.begin(), cnt.end());
      if (fewest >= 200 || n_args >= 524288) {
        if (fewest < 22)
          clog << " Warning: an epoch saw " << fewest << " survival events in " << n_args
               << " simulated ARGs; its lag is taken from a neighbouring epoch" << endl;
        break;
      }
    }
    for (int e = 0; e < E; e++) {
      clog << " Epoch " << e << " (" << cnt[e] << " events): survival distance " << surv[e] << endl;
      lags[e] = surv[e] * P.lag_fraction;
    }
  }
  clog << "    Lags set to: ";
  for (double l : lags) clog << l << " ";
  clog << endl;
  clog << " Starting position: " << fixed << setprecision(0) << P.start_position << setprecision(6) << scientific << endl;
  clog.unsetf(ios::floatfield);

  // One engine per chunk.  The reference runs one process per chunk (smc2 -chunks, model.py:563-662) and adds the
  // .out files; here the chunks of one process are swept concurrently on the device (smc_sweep_many) and added up.
  const size_t C = chunks.size();
  int marks_in_use = 0;                 // capacity of the migration-mark pool this sweep runs with
  vector<smc_engine*> eng(C, nullptr);
  struct Guard { vector<smc_engine*>& v; ~Guard() { for (smc_engine* e : v) smc_destroy(e); } } guard{eng};
  size_t Rtot = 0;
  for (size_t k = 0; k < C; k++) {
    const host::SegRows& rows = chunks[k];
    const size_t R = rows.rows();
    Rtot += R;
    vector<int32_t> max_rec(R);
    for (size_t r = 0; r < R; r++) {   // max_epoch_to_update (smcsmc.cpp:266-275)
      int e = 0;
      while (e < E && rows.dist_to_mutation[r] < 0.5 * lags[e]) e++;
      max_rec[r] = e - 1;
    }
    smc_config c;
    memset(&c, 0, sizeof c);
    c.n = n; c.N = (int32_t)P.N; c.E = E; c.device = devices[k % devices.size()];   // chunks round-robin over the GPUs
    // RNG sub-stream of the chunk: the low 24 bits of its start go into the Philox counter, the bits above into the key,
    // so that no two windows of a locus share a stream (starts that differ by a multiple of 2^24 bp used to)
    c.seed = P.seed + (uint64_t)em * 0x9E3779B97F4A7C15ull + (uint64_t)(((long long)chunk_start[k]) >> 24) * 0xD1B54A32D192ED03ull;
    c.chunk = (uint32_t)(((long long)chunk_start[k]) & 0xFFFFFF);
    c.mu = P.mutation_rate; c.rho = P.recombination_rate; c.seqlen = chunk_len; c.start_pos = 0.0;
    c.ess_fraction = P.ESS_fraction; c.ancestral_aware = P.ancestral_aware; c.parity_scan = 0;
    c.t_start = t_start.data(); c.inv2N = inv2N0.data(); c.vb_logfac = vb.data(); c.rec_mask = mask.data(); c.lags = lags.data();
    check(smc_create(&c, &eng[k]));
    if (np > 1) {
      // capacity of the per-particle pool of migration marks.  Marks live in shared memory (10 bytes each), so room that
      // is never used costs resident blocks: start from a pool that fits the typical tree (a branch per sample that
      // changes population once, plus what the summed migration rate adds) and, should any particle overflow it -- the
      // engine raises a flag, nothing is silently dropped -- repeat the sweep with twice the room (engine maximum 120).
      double msum = 0;
      for (int e = 0; e < E; e++) {
        double s = 0;
        for (int q = 0; q < np * np; q++) s += M.mig[(size_t)e * np * np + q];
        msum = max(msum, s * 4 * P.N0);
      }
      int marks = min(120, (n + 4 * (int)ceil(msum)) * P.marks_scale);
      if (const char* mb = getenv("SMCSMC_MARKS_BASE")) marks = min(120, atoi(mb) * P.marks_scale);   // tests: force the retry
      marks_in_use = marks;
      if (const char* mm = getenv("SMCSMC_MAX_MARKS")) marks = atoi(mm);
      check(smc_set_populations(eng[k], np, M.inv2N.data(), M.mig.data(), M.single_mig.data(), M.leaf_pop.data(), marks));
      if (P.vb) {   // particle.cpp:262-272: exp_digamma(c) / c per coalescence resp. migration
        vector<double> vc((size_t)E * np), vm((size_t)E * np * np, 0.0);
        for (int e = 0; e < E; e++)
          for (int p = 0; p < np; p++) {
            const double c = P.epochs[e].vb_count[p];
            vc[(size_t)e * np + p] = log(exp_digamma(c) / c);
            for (int q = 0; q < np; q++) {
              const double cm = P.epochs[e].mig_vb[(size_t)p * np + q];
              if (p != q) vm[((size_t)e * np + p) * np + q] = log(exp_digamma(cm) / cm);
            }
          }
        check(smc_set_population_vb(eng[k], vc.data(), vm.data()));
      }
    }
    check(smc_upload_rows(eng[k], rows.x_end.data(), rows.alleles.data(), rows.is_site.data(), max_rec.data(), (int64_t)R));
    check(smc_init_particles(eng[k]));
  }
  if (C == 1) {
    const host::SegRows& rows = chunks[0];
    const int64_t R = (int64_t)rows.rows();
    const int64_t step = max<int64_t>(1, R / 100);
    for (int64_t r0 = 0; r0 < R; r0 += step) {
      int64_t r1 = min<int64_t>(R, r0 + step);
      check(smc_sweep(eng[0], r0, r1));
      cout << "\r Particle filtering " << setw(4) << int(rows.x_end[r1 - 1] * 100 / chunk_len) << "% completed." << flush;
    }
  } else {
    cout << " Particle filtering " << C << " chunks (" << Rtot << " segments) concurrently." << endl;
    check(smc_sweep_many(eng.data(), (int32_t)C, 1));
  }
  cout << "\r Particle filtering step 100% completed." << endl;
  const int rows_t = smc_tally_rows(eng[0]);           // 6 for one population, 2Q with structure
  vector<double> t((size_t)rows_t * E, 0.0);
  double lnZ = 0;
  int64_t counters[4] = {0, 0, 0, 0};
  for (size_t k = 0; k < C; k++) {
    check(smc_finish(eng[k]));
    vector<double> tk((size_t)rows_t * E);
    double lz = 0;
    int64_t ck[4];
    check(smc_get_tallies(eng[k], tk.data()));
    int rc = smc_get_lnZ(eng[k], &lz);
    if (rc == SMC_ERR_NUMERIC && string(smc_last_error()).find("migration marks") != string::npos &&
        !getenv("SMCSMC_MAX_MARKS") && marks_in_use < 120)      // a pool already at the engine's maximum cannot grow
      throw MarksOverflow();
    if (rc == SMC_ERR_NUMERIC) throw runtime_error(string(smc_last_error()).find("probabilit") != string::npos
                                                       ? string("Zero or negative probabilities") : string(smc_last_error()));
    check(rc);
    check(smc_get_counters(eng[k], ck));
    for (size_t q = 0; q < tk.size(); q++) t[q] += tk[q];
    lnZ += lz;
    for (int q = 0; q < 4; q++) counters[q] += ck[q];
  }
  {
    double dev_ms = 0;
    int64_t launches = 0;
    for (size_t k = 0; k < C; k++) {
      double ms = 0;
      int64_t ln = 0;
      check(smc_get_timing(eng[k], &ms, &ln));
      dev_ms += ms; launches += ln;
    }
    clog << " Sweep: device time " << fixed << setprecision(1) << dev_ms << " ms, " << launches << " kernel launches"
         << setprecision(6) << scientific << endl;
    clog.unsetf(ios::floatfield);
  }
  if (P.record_resample_file)
    for (size_t k = 0; k < C; k++) {
      int64_t ck[4];
      check(smc_get_counters(eng[k], ck));
      vector<double> pos(ck[1] + 1), ess(ck[1] + 1);
      int64_t got = 0;
      check(smc_get_resample_log(eng[k], pos.data(), ess.data(), ck[1], &got));
      ofstream f(out.resample_name.c_str(), ios::out | ios::app | ios::binary);
      for (int64_t q = 0; q < got; q++) f << (long long)(pos[q] + (C > 1 ? chunk_start[k] - 1 : 0)) << "\t" << ess[q] << endl;
    }
  {
    double dev_ms = 0;
    int64_t launches = 0;
    for (size_t k = 0; k < C; k++) {
      double ms = 0;
      int64_t l = 0;
      check(smc_get_timing(eng[k], &ms, &l));
      dev_ms += ms; launches += l;
    }
    clog << " Device time of the sweep: " << dev_ms << " ms in " << launches << " kernel launches ("
         << (double)P.N * chunk_len * C / (dev_ms * 1e-3) << " particle*sites/s)" << endl;
  }
  clog << "Got to end of sequence; resampled " << counters[1] << " times" << endl;
  clog << " Inference step completed." << endl;

  // CountModel totals incl. the initial pseudo-counts (count.cpp:161-226): count = the rate, opportunity = weight = 1
  vector<double> coal_count((size_t)E * np), coal_opp((size_t)E * np), coal_weight((size_t)E * np);
  vector<double> rec_count(E), rec_opp(E), rec_weight(E);
  vector<double> mig_count((size_t)E * np * np, 0.0), mig_opp((size_t)E * np, 1.0), mig_weight((size_t)E * np, 1.0);
  if (np == 1) {
    for (int e = 0; e < E; e++) {
      coal_count[e] = t[0 * E + e] + inv2N0[e];
      coal_opp[e] = t[1 * E + e] + 1.0;
      coal_weight[e] = t[2 * E + e] + 1.0;
      rec_count[e] = t[3 * E + e] + P.recombination_rate;
      rec_opp[e] = t[4 * E + e] + 1.0;
      rec_weight[e] = t[5 * E + e] + 1.0;
    }
  } else {
    // cells of smc_set_populations: rows [0,Q) sums of w * cell, rows [Q,2Q) sums of w^2 * cell
    const int Q = rows_t / 2;
    auto W = [&](int cell, int e) { return t[(size_t)cell * E + e]; };
    auto W2 = [&](int cell, int e) { return t[(size_t)(Q + cell) * E + e]; };
    for (int e = 0; e < E; e++) {
      rec_count[e] = W(0, e) + P.recombination_rate;
      rec_opp[e] = W(1, e) + 1.0;
      rec_weight[e] = W2(1, e) + 1.0;
      for (int p = 0; p < np; p++) {
        coal_count[(size_t)e * np + p] = W(2 + 3 * p, e) + M.inv2N[(size_t)e * np + p];
        coal_opp[(size_t)e * np + p] = W(3 + 3 * p, e) + 1.0;
        coal_weight[(size_t)e * np + p] = W2(3 + 3 * p, e) + 1.0;
        mig_opp[(size_t)e * np + p] = W(4 + 3 * p, e) + 1.0;
        mig_weight[(size_t)e * np + p] = W2(4 + 3 * p, e) + 1.0;
        for (int q = 0; q < np; q++)
          mig_count[((size_t)e * np + p) * np + q] = W(2 + 3 * np + p * np + q, e) + M.mig[((size_t)e * np + p) * np + q];
      }
    }
  }
  const double resample_count = (double)counters[1], delayed_opp = P.loci_length;

  {   // <prefix>.recomb.gz (CountModel::dump_local_recomb_logs, count.cpp:616-654): per 100-bp bin the posterior-mean
      // recombination opportunity per nucleotide.  Each .seg row's opportunity (smc_get_row_opportunity) is spread
      // evenly over the row with the reference's differential book-keeping (record_local_recomb_events,
      // count.cpp:559-587).  The per-sample / time / log_time columns (recombination events attributed to their
      // descendants) are row f2 of the scope table and read 0.
    const double bin = 100.0;
    bool any_rec = false, all_rec = true;
    for (int e = 0; e < E; e++) {
      const bool on = e < (int)P.record_event_in_epoch.size() ? (P.record_event_in_epoch[e] & 1) : true;
      any_rec = any_rec || on; all_rec = all_rec && on;
    }
    if (any_rec && !all_rec && em == 0)
      clog << " Note: " << out.recomb_name << " reports the opportunity of all epochs, including those excluded with -xr" << endl;
    const size_t nbins = any_rec ? (size_t)(P.loci_length / bin) : 0;     // no recorded recombinations: header only
    vector<double> diff(nbins + 3, 0.0);
    for (size_t k = 0; k < C && nbins; k++) {
      const host::SegRows& rows = chunks[k];
      vector<double> ro(rows.rows());
      check(smc_get_row_opportunity(eng[k], ro.data()));
      const double off = chunk_start[k] - P.start_position;
      for (size_t r = 0; r < rows.rows(); r++) {
        const double xs = off + (r ? rows.x_end[r - 1] : 0.0), xe = off + rows.x_end[r];
        if (!(xs < xe)) continue;
        const size_t first = (size_t)(xs / bin), last = (size_t)(1 + xe / bin);
        if (last + 1 > diff.size()) diff.resize(last + 1, 0.0);
        const double fi = min((first + 1) * bin, xe) - xs, li = xe - max((last - 1) * bin, xs), dens = ro[r] / (xe - xs);
        if (first == last - 1) { diff[first] += fi * dens; diff[first + 1] -= fi * dens; }
        else {
          diff[first] += fi * dens; diff[first + 1] += (bin - fi) * dens;
          diff[last - 1] += (li - bin) * dens; diff[last] -= li * dens;
        }
      }
    }
    // formatting and compressing 10 bins per kb takes longer than the sweep of a 100-Mb locus: a writer thread does it
    // while the next EM iteration runs (joined before the next table and at exit)
    recomb_writer_join();
    g_recomb_writer = thread(write_recomb_table, out.recomb_name, em, n, (double)P.start_position, bin, nbins, std::move(diff));
  }

  // reset_model_parameters(force_update = true): the in-binary M-step (count.cpp:44-63,267-352)
  clog << " MODEL IS RESET at base " << P.loci_length << endl;
  double ro = 0, rc_ = 0, rw = 0;
  for (int e = 0; e < E; e++) { ro += rec_opp[e]; rc_ += rec_count[e]; rw += rec_weight[e]; }
  P.recombination_rate = rc_ / ro;
  clog << " Setting recombination rate to " << P.recombination_rate << " ( " << rc_ << " / " << ro << "; post-lag ESS "
       << 1.0 / (rw / ro) << " )" << endl;
  for (int e = 0; e < E; e++)
    for (int p = 0; p < np; p++) {
      const size_t at = (size_t)e * np + p;
      double rate = coal_count[at] / coal_opp[at];
      double pop = 1.0 / (2.0 * rate);
      P.epochs[e].pop_size[p] = (P.useCap && pop >= P.Ne_cap) ? P.Ne_cap : pop;
      P.epochs[e].vb_count[p] = 1e10;   // Model::addPopulationSize's default event count: the in-binary M-step drops -vb counts
      clog << " Setting size of population " << p << " @ " << setw(8) << t_start[e] << " to " << setw(8) << pop
           << " ( 0.5 * " << coal_opp[at] << " / " << coal_count[at] << "; post-lag ESS "
           << 1.0 / (coal_weight[at] / coal_opp[at]) << " )" << endl;
    }
  if (P.has_migration && np > 1)   // CountModel::reset_mig_rate (count.cpp:327-352)
    for (int e = 0; e < E; e++)
      for (int p = 0; p < np; p++)
        for (int q = 0; q < np; q++)
          if (p != q) {
            P.epochs[e].mig[(size_t)p * np + q] = mig_count[((size_t)e * np + p) * np + q] / mig_opp[(size_t)e * np + p];
            P.epochs[e].mig_vb[(size_t)p * np + q] = 1e10;                   // Model::addMigrationRate's default
          }

  // CountModel::log_counts (count.cpp:66-158) and the LogL row (smcsmc.cpp:391)
  auto t_end = [&](int e) { return e == E - 1 ? 1e+99 : t_start[e + 1]; };
  for (int e = 0; e < E; e++)
    for (int p = 0; p < np; p++)
      out.row(em, e, t_start[e], t_end(e), "Coal", p, -1, coal_opp[(size_t)e * np + p], coal_count[(size_t)e * np + p],
              coal_weight[(size_t)e * np + p]);
  out.row(em, -1, 0.0, 1e+99, "Recomb", -1, -1, ro, rc_, rw);
  for (int e = 0; e < E; e++)
    for (int p = 0; p < np; p++)
      for (int q = 0; q < np; q++)
        if (p != q)
          out.row(em, e, t_start[e], t_end(e), "Migr", p, q, mig_opp[(size_t)e * np + p],
                  mig_count[((size_t)e * np + p) * np + q], mig_weight[(size_t)e * np + p]);
  out.row(em, -1, 0.0, 1e+99, "Delay", -1, -1, delayed_opp, 0.0 / (double)P.N, delayed_opp);
  out.row(em, -1, 0.0, 1e+99, "Resamp", -1, -1, delayed_opp, resample_count, delayed_opp);
  out.row(em, -1, 0, 1e+99, "LogL", -1, -1, 1.0, lnZ, 1.0);
  clog << " Estimated log likelihood: " << lnZ << endl;
}

}  // namespace

int main(int argc, char* argv[]) {
  try {
    auto start = chrono::system_clock::now();
    Params P;
    host::parse(argc, argv, P);
    if (P.version) {
      cout << "Program was compiled on: " << __DATE__ << endl;
      cout << "smcsmc version: " << SMCSMC_B200_VERSION << endl;
      cout << "scrm version:   none (B200 engine: " << smc_version() << ")" << endl;
      return EXIT_SUCCESS;
    }
    if (P.help) { print_help(P); return EXIT_SUCCESS; }
    if (P.dump_model) { dump_model(P); return EXIT_SUCCESS; }
    Out out{P.out_prefix + ".out", P.out_prefix + ".log", P.out_prefix + ".recomb.gz", P.out_prefix + ".resample"};
    remove(out.out_name.c_str());
    remove(out.log_name.c_str());
    remove(out.recomb_name.c_str());
    if (P.record_resample_file) remove(out.resample_name.c_str());
    clog << "Command line:-" << endl << argv[0];
    for (int i = 1; i < argc; i++) clog << " " << argv[i];
    clog << endl;
    if (P.seg_file.empty()) throw invalid_argument("a .seg file is required (-seg); the empty-file mode is not supported");
    int max_seg_len = (int)(2.0 / (P.recombination_rate * 4 * P.N0));       // pfparam.cpp:364
    // -chunks C (extension): C equal windows of the locus, as smc2 would hand them to C processes via -startpos
    vector<host::SegRows> chunks;
    vector<double> chunk_start;
    const double chunk_len = P.chunks > 1 ? floor(P.loci_length / P.chunks) : P.loci_length;
    if (chunk_len < 2) throw invalid_argument("Out of range: -chunks (windows shorter than two bases)");
    for (int c = 0; c < P.chunks; c++) {
      double s0 = P.start_position + c * chunk_len;
      chunk_start.push_back(s0);
      chunks.push_back(host::read_seg(P.seg_file, P.nsam, chunk_len, (long long)s0, max_seg_len));
      if (P.dephase) {
        // -dephase (particleContainer.cpp:138-181): a genotype whose two haplotypes differ is treated as an unphased
        // heterozygote in the site likelihood, i.e. exactly like the pair "//" (code 2); nothing else reads the phase
        vector<int8_t>& al = chunks.back().alleles;
        for (size_t r = 0; r < chunks.back().rows(); r++)
          for (size_t i = 0; i + 1 < P.nsam; i += 2) {
            int8_t& a = al[r * P.nsam + i];
            int8_t& b = al[r * P.nsam + i + 1];
            if ((a == 0 && b == 1) || (a == 1 && b == 0)) a = b = 2;
          }
      }
    }
    // SMCSMC_DEVICE=3 selects one GPU; SMCSMC_DEVICES=0,1,2,3 spreads the chunks of -chunks over several
    vector<int> devices;
    if (const char* d = getenv("SMCSMC_DEVICES")) {
      string list(d), item;
      for (size_t q = 0; q <= list.size(); q++) {
        if (q == list.size() || list[q] == ',') { if (!item.empty()) devices.push_back(atoi(item.c_str())); item.clear(); }
        else item += list[q];
      }
    }
    if (devices.empty()) devices.push_back(getenv("SMCSMC_DEVICE") ? atoi(getenv("SMCSMC_DEVICE")) : 0);
    out.header();
    for (int i = 0; i <= P.EM_steps; i++) {
      clog << "EM step " << i << endl;
      for (;;) {
        try { pfARG_core(P, chunks, chunk_start, chunk_len, out, i, devices); break; }
        catch (const MarksOverflow&) {
          P.marks_scale *= 2;
          clog << " A particle ran out of migration-mark slots; repeating the sweep with room for "
               << "twice the room (the engine holds at most 120 marks per particle)" << endl;
        }
      }
      clog << "End of EM step " << i << endl;
      auto dur = chrono::duration_cast<chrono::seconds>(chrono::system_clock::now() - start);
      clog << "Elapsed time: " << setprecision(3) << (double)dur.count() << endl;
    }
    {
      ofstream lf(out.log_name.c_str(), ios::out | ios::app | ios::binary);
      write_log(lf, P, out);
    }
    write_log(cout, P, out);
    recomb_writer_join();
    return EXIT_SUCCESS;
  } catch (const exception& e) {
    if (g_recomb_writer.joinable()) g_recomb_writer.join();
    cerr << "Error: " << e.what() << endl;
    return EXIT_FAILURE;
  }
}
