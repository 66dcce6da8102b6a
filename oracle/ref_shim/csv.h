// Stand-in for ben-strasser/fast-cpp-csv-parser's "csv.h" (a FetchContent dependency of the reference,
// CMakeLists.txt:112-121, absent from /root/reference and unreachable without network).  Implements exactly the
// subset the reference uses in read_historical_returns (src/simulations.cpp:83-93):
//   io::CSVReader<1> in(path); in.read_header(io::ignore_extra_column, "returns"); while (in.read_row(x)) ...
// TEST INFRASTRUCTURE (oracle/_ref build only).
#pragma once
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace io {
enum ignore_column_t { ignore_no_column = 0, ignore_extra_column = 1, ignore_missing_column = 2 };

template <unsigned N>
class CSVReader {
  static_assert(N == 1, "shim supports a single selected column");

 public:
  explicit CSVReader(const std::string &path) : in_(path) {
    if (!in_) throw std::runtime_error("csv.h shim: cannot open " + path);
  }
  void read_header(ignore_column_t, const std::string &name) {
    std::string line;
    if (!std::getline(in_, line)) throw std::runtime_error("csv.h shim: empty file");
    std::stringstream ss(line);
    std::string cell;
    int i = 0;
    col_ = -1;
    while (std::getline(ss, cell, ',')) {
      while (!cell.empty() && (cell.back() == '\r' || cell.back() == ' ')) cell.pop_back();
      if (cell == name) col_ = i;
      ++i;
    }
    if (col_ < 0) throw std::runtime_error("csv.h shim: missing column " + name);
  }
  bool read_row(float &value) {
    std::string line;
    while (std::getline(in_, line)) {
      if (line.empty()) continue;
      std::stringstream ss(line);
      std::string cell;
      int i = 0;
      while (std::getline(ss, cell, ',')) {
        if (i == col_) {
          value = std::stof(cell);
          return true;
        }
        ++i;
      }
    }
    return false;
  }

 private:
  std::ifstream in_;
  int col_ = -1;
};
}  // namespace io
