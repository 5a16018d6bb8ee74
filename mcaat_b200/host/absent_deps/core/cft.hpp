// TEST INFRASTRUCTURE. Minimal stand-in for the un-vendored third-party set-cover library the
// reference's spacer_ordering.cpp includes ("core/cft.hpp"): just enough declarations for that
// translation unit to compile, so that its keep_crispr_regions_extended_by_k can be linked and run
// unmodified.  cft::run is never reached by the harness; it aborts if it ever is.
#pragma once
#include <cstddef>
#include <cstdlib>
#include <vector>
namespace cft {
struct Timer { void restart() {} };
struct Environment { double time_limit = 0; int verbose = 0; Timer timer; };
struct Instance {
    std::vector<std::vector<unsigned int>> cols;
    std::vector<double> costs;
    std::vector<std::vector<std::size_t>> rows;
};
struct Solution { std::vector<std::size_t> idxs; };
struct Result { Solution sol; };
template <class C, class R> inline void fill_rows_from_cols(const C &, std::size_t, R &) {}
inline Result run(Environment &, Instance &) { std::abort(); }
}  // namespace cft
