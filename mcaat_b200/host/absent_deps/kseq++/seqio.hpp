// TEST INFRASTRUCTURE. Stand-in for the un-vendored third-party header kseq++/seqio.hpp, which the reference's
// include/reads.h includes: only the declarations src/reads.cpp:3-18 needs to compile (a FASTQ stream whose
// read() returns records with a .seq member).  The harnesses never open a file through it — they feed
// sequences to the reference's get_read_from_sequence directly — so read() aborts if it is ever reached.
#pragma once
#include <cstdlib>
#include <string>
#include <vector>
namespace klibpp {
struct KSeq { std::string name, comment, seq, qual; };
struct SeqStreamIn {
  explicit SeqStreamIn(const char *) {}
  std::vector<KSeq> read() { std::abort(); }
};
}  // namespace klibpp
