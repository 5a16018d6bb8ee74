// STAND-IN for the third-party set-cover library the reference's include/spacer_ordering.h includes ("core/cft.hpp",
// https://github.com/c4v4/cft; an unpinned git submodule of the reference that is absent from this image).  A
// maintainer's build has the real library on its include path; this file only exists so that the reference's sources
// compile and run HERE.  It restates the interface src/spacer_ordering.cpp:270-314 uses (Instance{cols, costs, rows},
// fill_rows_from_cols, Environment, run -> sol.idxs) over a DETERMINISTIC unit-cost cover, so that two builds of the
// pipeline (reference CycleFinder / GPU drop-in) can be compared output for output:
//   greedy (most uncovered rows, lowest column index on ties), redundant columns dropped (highest index first), and
//   for instances of at most 24 columns an exact search for a smaller cover (iterative deepening over the size,
//   columns in index order), so that small instances get an optimal answer like the real solver's.
// The real library is a time-limited Lagrangian heuristic (the reference gives it 10 s, :303): its answers on large
// instances need not coincide with this one's.
#pragma once
#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <vector>
namespace cft {
struct Timer {
  void restart() {}
};
struct Environment {
  double time_limit = 0;
  int verbose = 0;
  Timer timer;
};
struct Instance {
  std::vector<std::vector<unsigned int>> cols;
  std::vector<double> costs;
  std::vector<std::vector<std::size_t>> rows;
};
struct Solution {
  std::vector<std::size_t> idxs;
  double cost = 0;
};
struct Result {
  Solution sol;
};
template <class C, class R>
inline void fill_rows_from_cols(const C &cols, std::size_t n_rows, R &rows) {
  rows.assign(n_rows, typename R::value_type());
  for (std::size_t j = 0; j < cols.size(); ++j)
    for (auto i : cols[j])
      if ((std::size_t)i < n_rows) rows[i].push_back(j);
}
namespace detail {
inline bool exact(const std::vector<std::vector<uint64_t>> &colbits, const std::vector<uint64_t> &covered, std::size_t from,
                  std::size_t left, std::vector<std::size_t> &pick) {
  bool all = true;
  for (uint64_t w : covered) all = all && w == ~0ull;
  if (all) return true;
  if (left == 0) return false;
  for (std::size_t j = from; j < colbits.size(); ++j) {
    std::vector<uint64_t> next = covered;
    bool gain = false;
    for (std::size_t w = 0; w < next.size(); ++w) {
      gain = gain || (colbits[j][w] & ~next[w]);
      next[w] |= colbits[j][w];
    }
    if (!gain) continue;
    pick.push_back(j);
    if (exact(colbits, next, j + 1, left - 1, pick)) return true;
    pick.pop_back();
  }
  return false;
}
}  // namespace detail
inline Result run(Environment &, Instance &inst) {
  const std::size_t n_rows = inst.rows.size(), n_cols = inst.cols.size();
  std::vector<char> covered(n_rows, 0);
  std::size_t n_covered = 0;
  std::vector<std::size_t> chosen;
  while (n_covered < n_rows) {  // greedy
    std::size_t best = n_cols, best_gain = 0;
    for (std::size_t j = 0; j < n_cols; ++j) {
      std::size_t gain = 0;
      std::vector<unsigned int> seen;
      for (auto i : inst.cols[j])
        if ((std::size_t)i < n_rows && !covered[i] && std::find(seen.begin(), seen.end(), i) == seen.end()) {
          seen.push_back(i);
          ++gain;
        }
      if (gain > best_gain) {
        best_gain = gain;
        best = j;
      }
    }
    if (best == n_cols) break;  // rows no column covers
    chosen.push_back(best);
    for (auto i : inst.cols[best])
      if ((std::size_t)i < n_rows && !covered[i]) {
        covered[i] = 1;
        ++n_covered;
      }
  }
  std::sort(chosen.begin(), chosen.end());
  for (std::size_t q = chosen.size(); q-- > 0;) {  // drop redundant columns, highest index first
    std::vector<int> cnt(n_rows, 0);
    for (std::size_t p = 0; p < chosen.size(); ++p)
      if (p != q)
        for (auto i : inst.cols[chosen[p]])
          if ((std::size_t)i < n_rows) cnt[i] = 1;
    bool needed = false;
    for (auto i : inst.cols[chosen[q]])
      if ((std::size_t)i < n_rows && !cnt[i]) needed = true;
    if (!needed) chosen.erase(chosen.begin() + (std::ptrdiff_t)q);
  }
  if (n_cols <= 24 && chosen.size() > 1) {  // small instance: is there a smaller cover?
    const std::size_t words = (n_rows + 63) / 64;
    std::vector<std::vector<uint64_t>> colbits(n_cols, std::vector<uint64_t>(words, 0));
    std::vector<uint64_t> start(words, 0);
    for (std::size_t j = 0; j < n_cols; ++j)
      for (auto i : inst.cols[j])
        if ((std::size_t)i < n_rows) colbits[j][i >> 6] |= 1ull << (i & 63);
    for (std::size_t i = 0; i < words * 64; ++i)  // rows beyond n_rows and rows no column covers count as covered
      if (i >= n_rows || !covered[i]) start[i >> 6] |= 1ull << (i & 63);
    for (std::size_t size = 1; size < chosen.size(); ++size) {
      std::vector<std::size_t> pick;
      if (detail::exact(colbits, start, 0, size, pick)) {
        chosen = pick;
        break;
      }
    }
  }
  Result r;
  r.sol.idxs = chosen;
  r.sol.cost = (double)chosen.size();
  return r;
}
}  // namespace cft
