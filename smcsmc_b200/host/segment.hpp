// The .seg reader of the reference (Segment, src/segdata.cpp:55-223,413-451) producing the flat row arrays the engine
// uploads once per chunk.  Same accepted formats (3 or 6 tab-separated columns, alleles 0 1 . /), same validation
// messages, same splitting of long rows into INVARIANT_PARTIAL pieces, same -startpos windowing.
#pragma once
#include <cstdint>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace host {

struct SegRows {
  std::vector<double> x_end;            // min(segment_end, seqlen), offsets from data_start
  std::vector<int8_t> alleles;          // [R][nsam]
  std::vector<uint8_t> is_site;         // 1 = SEGMENT_INVARIANT, 0 = SEGMENT_INVARIANT_PARTIAL
  std::vector<double> dist_to_mutation; // Segment::distance_to_mutation (segdata.cpp:234-262)
  size_t rows() const { return x_end.size(); }
};

inline std::vector<int8_t> parse_alleles(const std::string& field, size_t nsam) {
  if (nsam > field.size()) throw std::invalid_argument("Number of variant site is wrong: " + field);
  std::vector<int8_t> out(nsam);
  for (size_t i = 0; i < nsam; i++) {
    switch (field[i]) {
      case '.': out[i] = -1; break;
      case '/': out[i] = 2; break;
      case '0': out[i] = 0; break;
      case '1': out[i] = 1; break;
      default: throw std::invalid_argument("Unknown character found in .seg file; expect one of '.', '/', '0' or '1'.");
    }
    if (out[i] == -1 && (i % 2) == 1 && out[i - 1] != -1)
      throw std::invalid_argument("Found inconsistent unphased heterozygous marks");
  }
  return out;
}

inline SegRows read_seg(const std::string& path, size_t nsam, double seqlen, long long data_start,
                        double max_segment_length) {
  std::ifstream in(path.c_str());
  if (!in.good()) throw std::invalid_argument("Invalid input file: " + path);
  struct Datum { long long start, len; int state; std::vector<int8_t> al; };
  std::vector<Datum> buf;
  std::string line;
  long long next_start = -1;
  while (std::getline(in, line) && !line.empty()) {
    if (line[0] == '#') continue;
    std::vector<size_t> cols;
    size_t pos = 0;
    while (pos < line.size()) {
      cols.push_back(pos);
      while (pos < line.size() && line[pos] != '\t') pos++;
      if (pos < line.size()) pos++;
    }
    if (cols.size() < 3) throw std::invalid_argument("Require 3 or 6 columns");
    char* end;
    long long seg_start = strtoll(line.c_str(), &end, 10);
    if (*end != '\t') throw std::invalid_argument("Segment start position at:" + line);
    long long seg_len = strtol(line.c_str() + cols[1], &end, 10);   // the last length may carry a decimal point
    std::vector<int8_t> al;
    auto field = [&](size_t c) {
      size_t e = c + 1 < cols.size() ? cols[c + 1] - 1 : line.size();
      return line.substr(cols[c], e - cols[c]);
    };
    char c2 = line[cols[2]];
    if ((c2 == 'T' || c2 == 'F') && cols[2] + 1 < line.size() && line[cols[2] + 1] == '\t') {
      if (cols.size() != 6) throw std::invalid_argument("Require 6 (or 3) columns");
      char c3 = line[cols[3]];
      if ((c3 != 'T' && c3 != 'F') || line[cols[3] + 1] != '\t')
        throw std::invalid_argument("Expected T or F in .seg file column 3 and 4");
      strtol(line.c_str() + cols[4], &end, 10);
      if (*end != '\t') throw std::invalid_argument("Bad chromosome (not an integer) in column 5");
      al = parse_alleles(field(5), nsam);
    } else {
      if (cols.size() != 3) throw std::invalid_argument("Require 3 (or 6) columns");
      al = parse_alleles(field(2), nsam);
    }
    if (next_start > -1 && next_start != seg_start) throw std::invalid_argument("Segments are not consecutive");
    next_start = seg_start + seg_len;
    if ((double)seg_start >= (double)data_start + seqlen) break;
    if (seg_start + seg_len > data_start) {
      do {
        int state;
        if ((double)seg_len > max_segment_length) { seg_len = (long long)max_segment_length; state = 0; }
        else state = 1;
        if (seg_start + seg_len > data_start) buf.push_back({seg_start, seg_len, state, al});
        seg_start += seg_len;
        seg_len = next_start - seg_start;
      } while (seg_start < next_start);
    }
  }
  if (buf.empty())
    throw std::invalid_argument("No data found in file " + path + " between positions " + std::to_string(data_start) +
                                " and " + std::to_string((long long)(data_start + seqlen)));
  SegRows out;
  const size_t R = buf.size();
  std::vector<char> allmiss(R);
  for (size_t r = 0; r < R; r++) {
    bool m = true;
    for (int8_t a : buf[r].al) m &= (a == -1);
    allmiss[r] = m;
  }
  for (size_t r = 0; r < R; r++) {
    double e = (double)(buf[r].start - data_start + buf[r].len);
    out.x_end.push_back(e < seqlen ? e : seqlen);
    out.alleles.insert(out.alleles.end(), buf[r].al.begin(), buf[r].al.end());
    out.is_site.push_back((uint8_t)buf[r].state);
    double d = 0.0;
    if (allmiss[r]) {
      for (long long i = (long long)r; i >= 0 && allmiss[i]; i--) d = (double)(buf[r].start - buf[i].start);
      size_t j = r;
      while (j < R && allmiss[j]) j++;
      if (j < R) d = std::min(d, (double)(buf[j].start + buf[j].len - buf[r].start));
    }
    out.dist_to_mutation.push_back(d);
    if (e >= seqlen) break;               // the sweep stops at the row that reaches the locus end (smcsmc.cpp:353-356)
  }
  return out;
}

}  // namespace host
