// Internal to mcaat_b200/host: what the drop-in translation units share.  A device graph made by the
// CycleFinder drop-in stays resident for as long as that CycleFinder object lives (reference src/main.cpp:536
// keeps it until the end of main), so the stages after it can run on the same HBM-resident graph.
#pragma once
#include <cstdint>

#include "sdbg/sdbg.h"  // the reference's vendored MEGAHIT header
#include "mcaat_gpu.h"

namespace mcaat_gpu_bridge {
[[noreturn]] void fail(const char *what, int rc);
// device graph of a live CycleFinder drop-in constructed on this SDBG, or nullptr
mg_graph *resident_graph(const SDBG *sdbg);
// extract every item through the public getters and upload (current invalid bits included)
mg_graph *upload(SDBG &sdbg);
// device invalid bits -> SetInvalidEdge on the host SDBG
void write_back_invalid(mg_graph *g, SDBG &sdbg);
}  // namespace mcaat_gpu_bridge
