// smmc_io.cpp -- data-format edges of the hot path (host code, no device): the historical-returns CSV the callers
// feed into the bootstrap table, and the per-trajectory output file the reference's Python tooling reads back.
//
// Reference: read_historical_returns (src/simulations.cpp:83-93, io::CSVReader<1> with
// read_header(io::ignore_extra_column, "returns")) and write_data_file (src/helpers.cpp:23-39, consumed by
// python/plot_returns.py:40).  The CSV parser of the reference is a third-party header (fast-cpp-csv-parser,
// un-vendored); what is restated here is its documented behaviour for this call: comma separator, a header line in
// which the column "returns" is located by exact name (other columns ignored), spaces/tabs trimmed around fields,
// a trailing '\r' dropped, one float per data row.  Blank lines are skipped.
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/smmc.h"

namespace smmc {
smmc_status set_error(smmc_status st, const char *fmt, ...);  // smmc_api.cu
}

namespace {

std::string trimmed(const std::string &s) {
  size_t a = 0, b = s.size();
  while (a < b && (s[a] == ' ' || s[a] == '\t')) a++;
  while (b > a && (s[b - 1] == ' ' || s[b - 1] == '\t')) b--;
  return s.substr(a, b - a);
}

std::vector<std::string> split_fields(const std::string &line) {
  std::vector<std::string> out;
  size_t start = 0;
  for (;;) {
    const size_t c = line.find(',', start);
    out.push_back(trimmed(line.substr(start, c == std::string::npos ? std::string::npos : c - start)));
    if (c == std::string::npos) break;
    start = c + 1;
  }
  return out;
}

bool next_line(std::ifstream &in, std::string &line) {
  if (!std::getline(in, line)) return false;
  if (!line.empty() && line.back() == '\r') line.pop_back();
  return true;
}

}  // namespace

extern "C" {

smmc_status smmc_read_returns_csv(const char *path, float *out, int32_t cap, int32_t *n_out) {
  if (!path || !n_out || cap < 0) return smmc::set_error(SMMC_ERR_INVALID_ARG, "read_returns_csv: bad arguments");
  *n_out = 0;
  std::ifstream in(path);
  if (!in) return smmc::set_error(SMMC_ERR_INVALID_ARG, "read_returns_csv: cannot open %s", path);
  std::string line;
  if (!next_line(in, line)) return smmc::set_error(SMMC_ERR_INVALID_ARG, "read_returns_csv: %s has no header line", path);
  const std::vector<std::string> header = split_fields(line);
  int col = -1;
  for (size_t i = 0; i < header.size(); i++)
    if (header[i] == "returns") {
      if (col >= 0) return smmc::set_error(SMMC_ERR_INVALID_ARG, "read_returns_csv: duplicated column \"returns\" in %s", path);
      col = (int)i;
    }
  if (col < 0) return smmc::set_error(SMMC_ERR_INVALID_ARG, "read_returns_csv: missing column \"returns\" in header of %s", path);
  int64_t n = 0, line_no = 1;
  while (next_line(in, line)) {
    line_no++;
    if (trimmed(line).empty()) continue;
    const std::vector<std::string> f = split_fields(line);
    if ((int)f.size() != (int)header.size())
      return smmc::set_error(SMMC_ERR_INVALID_ARG, "read_returns_csv: %s line %lld has %zu columns, header has %zu", path,
                             (long long)line_no, f.size(), header.size());
    const std::string &tok = f[(size_t)col];
    char *end = nullptr;
    errno = 0;
    const float v = strtof(tok.c_str(), &end);
    if (tok.empty() || end != tok.c_str() + tok.size())
      return smmc::set_error(SMMC_ERR_INVALID_ARG, "read_returns_csv: %s line %lld: \"%s\" is not a number", path,
                             (long long)line_no, tok.c_str());
    if (out && n < cap) out[n] = v;
    n++;
    if (n > 0x7fffffff) return smmc::set_error(SMMC_ERR_INVALID_ARG, "read_returns_csv: too many rows in %s", path);
  }
  *n_out = (int32_t)n;
  if (out && n > cap) return smmc::set_error(SMMC_ERR_INVALID_ARG, "read_returns_csv: %lld rows, capacity %d", (long long)n, cap);
  return SMMC_OK;
}

smmc_status smmc_write_data_file(const char *path, const float *returns, int32_t n_returns, const float *values,
                                 int32_t n_values) {
  if (!path || n_returns < 0 || n_values < 0 || (n_returns > 0 && !returns) || (n_values > 0 && !values))
    return smmc::set_error(SMMC_ERR_INVALID_ARG, "write_data_file: bad arguments");
  std::ofstream f(path);
  if (!f) return smmc::set_error(SMMC_ERR_INVALID_ARG, "write_data_file: cannot open %s for writing", path);
  // same stream formatting as the reference (operator<<(float), default precision 6)
  f << "Returns,,";
  for (int32_t i = 0; i < n_returns; i++) f << returns[i] << ",";
  f << "\nValues,";
  for (int32_t i = 0; i < n_values; i++) f << values[i] << ",";
  f.flush();
  if (!f) return smmc::set_error(SMMC_ERR_INVALID_ARG, "write_data_file: write to %s failed", path);
  return SMMC_OK;
}

}  // extern "C"
