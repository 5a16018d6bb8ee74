// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// Harness around the UNMODIFIED reference graph producer `SDBGBuild` (reference
// src/sdbg_build.cpp:3-13: buildlib + read2sdbg -k 23 -m 1).
//   build_sdbg "<r1.fq> [r2.fq]" <graph_dir> [threads=8] [host_mem_bytes=8e9]
// Writes <graph_dir>/graph.sdbg_info and graph.sdbg.<i>.
#include <cstdlib>
#include <string>
#include "sdbg_build.h"
#include "settings.h"

int main(int argc, char **argv) {
  if (argc < 3) return 1;
  Settings s;
  s.input_files = argv[1];
  s.graph_folder = argv[2];
  s.threads = argc > 3 ? atoi(argv[3]) : 8;
  s.ram = argc > 4 ? atof(argv[4]) : 8e9;  // forwarded verbatim as --host_mem (sdbg_build.cpp:202)
  SDBGBuild b(s);
  return 0;
}
