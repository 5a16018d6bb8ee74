"""Host-side mirror of the reference's `Settings` / `CycleFinder` interface for the hot path
(reference include/settings.h:22-56, include/cycle_finder.h:25-85, src/cycle_finder.cpp).

Same names, argument meaning and error behaviour as the C++ class: constructing `CycleFinder`
runs the whole search (src/cycle_finder.cpp:131-138) and leaves `results` — one entry per searched
start node, in the reference's threads=1 insertion order, each value the raw cycle sequence — and
the pruned invalid bitmap.  All compute goes through libmcaat_gpu.so on a B200; nothing here
falls back to the CPU.
"""
import sys
from dataclasses import dataclass, field

import numpy as np

from . import gpu


@dataclass
class CycleFinderSettings:  # settings.h:33-38
    threshold_multiplicity: int = 20
    low_abundance: bool = True  # no effect on results, as in the reference (cycle_finder.cpp:391-393,439)
    cycle_max_length: int = 77
    cycle_min_length: int = 27


@dataclass
class Settings:  # settings.h:22-56 (the fields the hot path reads)
    threads: int = 1
    cycle_finder_settings: CycleFinderSettings = field(default_factory=CycleFinderSettings)
    sdbg: object = None  # a gpu.DeviceGraph (the device-resident counterpart of SDBG*)
    quiet: bool = False


class CycleFinder:
    """Drop-in for the reference class: `CycleFinder(settings)` then read `.results`."""

    def __init__(self, settings):
        self.settings = settings
        if settings.sdbg is None:  # cycle_finder.cpp:134-136
            raise RuntimeError("CycleFinder requires settings.sdbg to be set to a valid SDBG instance")
        if not isinstance(settings.sdbg, gpu.DeviceGraph):
            raise RuntimeError("settings.sdbg must be a device-resident graph (mcaat_b200.DeviceGraph)")
        self.results = {}       # start node -> [cycle ndarray, ...]   (insertion order = reference order)
        self.start_nodes_chunked = []  # [(ceil(log2 mult), ids ndarray)] descending key
        self.stats = {}
        self.total_cycles = self.FindApproximateCRISPRArrays()

    def _log(self, *a):
        if not self.settings.quiet:
            print(*a, file=sys.stdout, flush=True)

    # ---- pieces (cycle_finder.h public methods)
    def ChunkStartNodes(self):
        """cycle_finder.cpp:387-427 -> (buckets, total)"""
        s = self.settings.cycle_finder_settings
        buckets, info = self.settings.sdbg.select_starts(s.threshold_multiplicity, s.cycle_max_length)
        for key, ids in buckets:
            self._log(f"Chunked start nodes: multiplicity bucket (log2)={key}, nodes={ids.size}")
        self.stats.update(info)
        return buckets, sum(ids.size for _, ids in buckets)

    def DepthLevelSearch(self, start, target, limit):
        """cycle_finder.cpp:248-343; only target == start is supported (the reference's only call site)."""
        if start != target:
            raise NotImplementedError("DepthLevelSearch is only used with target == start on this path")
        return bool(self.settings.sdbg.dls([start], limit)[0][0])

    def FindCycleUtil(self, startnode):
        """cycle_finder.cpp:231-243 against the current `visited` state (marks visited like :217-222)."""
        s = self.settings.cycle_finder_settings
        searched, _ = self.settings.sdbg.find_cycles([startnode], s.cycle_min_length, s.cycle_max_length)
        return searched[0][1] if searched else []

    # ---- driver (cycle_finder.cpp:433-492)
    def FindApproximateCRISPRArrays(self):
        g = self.settings.sdbg
        s = self.settings.cycle_finder_settings
        st = g.prune()
        self.stats.update(st)
        self._log(f"Graph size: {g.n} nodes; gathered tips: {st['n_tips_before']}")
        self._log(f"Pre-filter: invalidated {st['n_mult1']} node(s) with multiplicity <= 1.")
        self._log(f"After pruning, tips: {st['n_tips_after']}, valid edges: {st['n_valid']}")
        self._log(f"Total nodes in graph: {g.n}")
        self._log(f"Starting cycle enumeration: max_len={s.cycle_max_length} min_len={s.cycle_min_length} "
                  f"threads={self.settings.threads}")
        buckets, total = self.ChunkStartNodes()
        self.start_nodes_chunked = buckets
        self._log(f"Start nodes found in chunks: {total}")
        starts = np.concatenate([ids for _, ids in buckets]) if buckets else np.zeros(0, np.uint64)
        searched, stats = g.find_cycles(starts, s.cycle_min_length, s.cycle_max_length)
        self.stats.update(stats)
        for start, cycles in searched:
            self.results[start] = cycles
        cumulative = 0
        for key, ids in buckets:
            in_bucket = sum(len(self.results[int(v)]) for v in ids if int(v) in self.results)
            cumulative += in_bucket
            self._log(f"Bucket log2_mult={key}: processed {ids.size} nodes, found {in_bucket} cycles "
                      f"(cumulative {cumulative})")
        self._log(f"Cycle enumeration completed: total cycles={cumulative}, result nodes={len(self.results)}")
        return cumulative
