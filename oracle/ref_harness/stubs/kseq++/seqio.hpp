// TEST INFRASTRUCTURE. Empty stand-in for the un-vendored third-party header kseq++/seqio.hpp, which
// the reference's include/reads.h includes.  Nothing of it is used by the functions the region
// harness calls, so no declarations are needed.
#pragma once
