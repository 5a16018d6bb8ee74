#!/usr/bin/env python3
"""bench.py — MCAAT cycle-search hot path on B200 (BASELINE.json metric: cycle-search start nodes/s
and cycle-find wall time).

A "step" is one full pass of the hot path over the resident graph: reset -> prune (tips,
multiplicity filter, peel) -> start-node selection (candidate scan + admission search) -> ordered
cycle enumeration, results back on the host (reference src/cycle_finder.cpp:433-492).
Workload: BASELINE.json configs[1] — synthetic 50-genome community (+5 strain copies), ~200M-item
SDBG built on the GPU by tools/synth_sdbg.py (pinned against the reference's own graph producer),
default cycle bounds 27-77, multiplicity threshold 20.

  python bench.py --gpus 1 --steps K --warmup W            our arm (CUDA, libmcaat_gpu.so)
  python bench.py --impl reference ...                     the reference's CPU CycleFinder (oracle/_ref)
  torchrun ... bench.py --gpus N ...                       N ranks: NCCL broadcast of the raw graph,
                                                           admission searches sharded over ranks
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "cycle_search_start_nodes_per_sec"
UNIT = "start_nodes/s"
THR, MIN_LEN, MAX_LEN = 20, 27, 77
# algorithmic bytes per unit of work
BYTES_STREAM_NODE = 2.875   # W 4 bit + last/tip/invalid 3 bit + uint16 multiplicity, per node per pass
BYTES_EXPANSION = 256.0     # 8 sectors per node expansion (layout-independent nominal)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows = []
        self.rows_before = 0
        self.proc = None
        self.idx = gpu_index

    def start(self, uuid=None):
        """In-process NVML polling (the library nvidia-smi itself reads; a query costs microseconds and takes no
        driver-wide lock, whereas every `nvidia-smi -lms` tick stalled kernel launches for milliseconds and showed up
        as one slow timed step).  Falls back to the nvidia-smi loop when pynvml is missing."""
        self.source = "nvml"
        try:
            import pynvml
            pynvml.nvmlInit()
            h = None
            if uuid:
                for cand in (uuid, "GPU-" + uuid):
                    try:
                        h = pynvml.nvmlDeviceGetHandleByUUID(cand.encode() if isinstance(cand, str) else cand)
                        break
                    except Exception:
                        h = None
            if h is None:
                h = pynvml.nvmlDeviceGetHandleByIndex(self.idx)
            mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
            reasons_fn = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            self._stop = threading.Event()

            def poll():
                while not self._stop.is_set():
                    try:
                        sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                        r = int(reasons_fn(h))
                        # row layout of the nvidia-smi query below: idx, sm, max, power, active, hw, hw_thermal, sw_thermal, sw_power
                        flag = lambda bit: "Active" if r & bit else "Not Active"
                        self.rows.append([str(self.idx), str(sm), str(mx), "", hex(r), flag(0x8), flag(0x40), flag(0x20), flag(0x4)])
                    except Exception:
                        pass
                    self._stop.wait(0.05)
            threading.Thread(target=poll, daemon=True).start()
            return
        except Exception:
            self.source = "nvidia-smi"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if getattr(self, "_stop", None):
            self._stop.set()
        if self.proc:
            self.proc.terminate()
        if len(self.rows) > self.rows_before:
            self.rows = self.rows[self.rows_before:]
        sm = [float(r[1]) for r in self.rows if len(r) > 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) > 8 and r[2].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) > 8:
                for nm, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm), source=getattr(self, "source", "nvidia-smi"))


def unpack_for_oracle(g):
    n = g["n"]
    w8 = g["w"].view(np.uint8)
    w = np.empty(w8.size * 2, np.uint8)
    w[0::2] = w8 & 15
    w[1::2] = w8 >> 4
    last = np.unpackbits(g["last"].view(np.uint8), bitorder="little")[:n]
    tip = np.unpackbits(g["tip"].view(np.uint8), bitorder="little")[:n]
    return w[:n], last, tip


def hot_path_step(g, rank=0, world=1, dist=None, torch=None):
    """One pass of the hot path through the C ABI.  With world > 1 the admission searches
    (DepthLevelSearch, the dominant phase) are sharded block-cyclically over the ranks and their
    pass flags all-gathered; prune / candidate scan / ordered enumeration run replicated (no
    collective inside a search)."""
    dbg = os.environ.get("BENCH_DEBUG")
    tt = [time.perf_counter()]
    g.reset()
    tt.append(time.perf_counter())
    st = g.prune()
    tt.append(time.perf_counter())
    cand = g.candidates(THR)
    tt.append(time.perf_counter())
    if world == 1:
        ok, exp = g.dls(cand, MAX_LEN)
    else:
        from mcaat_b200 import shard
        ok_m, exp = g.dls(shard.my_share(cand, rank, world), MAX_LEN)
        ok = shard.gather_verdicts(ok_m, cand.size, rank, world, dist, torch, "cuda")
        e = torch.tensor([exp], dtype=torch.int64, device="cuda")
        dist.all_reduce(e)
        exp = int(e.item())
    tt.append(time.perf_counter())
    buckets, _ = g.bucket_starts(cand[ok])
    starts = np.concatenate([ids for _, ids in buckets]) if buckets else np.zeros(0, np.uint64)
    tt.append(time.perf_counter())
    (ss, coff, noff, nodes), cs = g.find_cycles(starts, MIN_LEN, MAX_LEN, csr=True)
    tt.append(time.perf_counter())
    if dbg:
        print("[buckets] " + " ".join("%d:%d" % (int(k), len(ids)) for k, ids in buckets), file=sys.stderr)
        print("[step wall ms] reset %.1f prune %.1f candidates %.1f dls %.1f bucket %.1f find_cycles %.1f" %
              tuple((b - a) * 1e3 for a, b in zip(tt, tt[1:])), file=sys.stderr)
    res_bytes = 8 * int(nodes.size + ss.size)
    hot_path_step.last_cycle_nodes = nodes  # all cycles, flattened
    hot_path_step.last_result = (ss, coff, noff, nodes)
    stats = dict(st, n_candidates=int(cand.size), n_starts=int(starts.size), dls_expansions=int(exp), **cs)
    return stats, res_bytes


def cpu_baseline_port(sample_name, device):
    """Oracle port (oracle/mcaat_oracle.c, 1 thread) on the bounded sample workload."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O
    import synth_sdbg as S
    g, info = S.make_workload(sample_name, device=device)
    w, last, tip = unpack_for_oracle(g)
    og = O.OracleGraph.from_arrays(w, last, tip, g["mult"], g["f"], g["k"])
    t0 = time.perf_counter()
    r = og.run(THR, MIN_LEN, MAX_LEN)
    dt = time.perf_counter() - t0
    n_starts = int(r.start_list().size)
    if getattr(cpu_baseline_port, "with_regions", False):  # N4 on the same sample: literal restatement, 1 thread
        seeds = [p for _, ps in r.searched for p in ps]
        seeds = np.concatenate(seeds).astype(np.uint64) if seeds else np.zeros(0, np.uint64)
        t1 = time.perf_counter()
        gone, kept = og.keep_regions(seeds, 128)
        cpu_baseline_port.regions = dict(workload=sample_name, n_items=int(g["n"]), k=128, seed_ids=int(seeds.size),
                                         n_kept=kept, seconds=time.perf_counter() - t1, cores=1, kind="port")
    return dict(value=n_starts / dt, unit=UNIT, cores=1, kind="port",
                sample=f"workload '{sample_name}' (same recipe at 1/20 scale): {g['n']} items, {n_starts} start nodes, "
                       f"{dt:.2f} s for the whole search with oracle/mcaat_oracle.c on 1 thread",
                seconds=dt, n_items=int(g["n"]), n_starts=n_starts)


def reference_arm(args):
    """--impl reference: the reference's own CPU CycleFinder (oracle/_ref/ref_cyclefinder, compiled from the
    unmodified sources) with all host threads it can use (BASELINE.json quotes --threads 16), on the SAME graph as the
    B200 arm (args.workload).  The graph files are written once, loaded once; every repetition constructs the
    CycleFinder in a forked child (oracle/ref_harness/ref_cyclefinder_main.cpp --bench).  Repetitions stop when
    --warmup + --steps are done or the time budget (BENCH_REF_BUDGET_S, default 240 s) is used up; the number
    actually timed is printed."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O
    import synth_sdbg as S
    try:
        import torch
        device = "cuda" if torch.cuda.is_available() else "cpu"
    except Exception:
        device = "cpu"
    try:
        cores_avail = len(os.sched_getaffinity(0))  # what `nproc` reports; os.cpu_count() ignores the container's CPU set
    except AttributeError:
        cores_avail = os.cpu_count() or 1
    workload = args.workload if device == "cuda" or args.workload != "c2" else args.sample  # no GPU: the full graph cannot be generated
    g, info = S.make_workload(workload, device=device, want_labels=True)
    n_items = int(g["n"])
    graph_sha = graph_digest(g)
    budget = float(os.environ.get("BENCH_REF_BUDGET_S", "240"))
    rep_timeout = float(os.environ.get("BENCH_REF_TIMEOUT_S", "400"))
    exe = O.ref_binary("ref_cyclefinder")
    times, n_starts, kind, threads_used, why, load_s = [], 0, "reference", 0, "", None
    threads_requested = min(cores_avail, 16)
    if exe:
        tmp = tempfile.mkdtemp(prefix="mcaat_ref_")
        prefix = os.path.join(tmp, "graph")
        S.write_megahit_files(g, prefix)
        # With more than one thread the reference races on its result map and can corrupt it or never
        # return, so every repetition is bounded and the arm falls back to fewer threads instead of hanging.
        for threads in sorted({threads_requested, min(cores_avail, 4), 1}, reverse=True):
            p = subprocess.run([exe, "--bench", prefix, str(threads), str(THR), str(MIN_LEN), str(MAX_LEN),
                                str(args.warmup + args.steps), str(budget), str(rep_timeout)], capture_output=True, text=True)
            reps = [json.loads(ln[9:]) for ln in p.stdout.splitlines() if ln.startswith("REF_JSON ")]
            for ln in p.stdout.splitlines():
                if ln.startswith("REF_LOAD "):
                    load_s = json.loads(ln[9:])["load_s"]
            if reps:
                threads_used = threads
                n_starts = reps[-1]["start_nodes"]
                times = [r["ctor_s"] for r in reps]
                if len(times) > args.warmup:  # enough repetitions fitted: drop the warm-up ones
                    times = times[args.warmup:]
                break
            fails = [ln for ln in p.stdout.splitlines() if ln.startswith("REF_FAIL ")]
            why += f"{threads}t:{json.loads(fails[-1][9:])['why'] if fails else 'rc%d' % p.returncode};"
        shutil.rmtree(tmp, ignore_errors=True)
        if not times:
            exe = None
    if not exe:  # reference binary absent (or it failed at every thread count): the oracle port is the CPU implementation
        kind, threads_used = "port", 1
        w, last, tip = unpack_for_oracle(g)
        t_start = time.perf_counter()
        for i in range(args.warmup + args.steps):
            og = O.OracleGraph.from_arrays(w, last, tip, g["mult"], g["f"], g["k"])
            t0 = time.perf_counter()
            r = og.run(THR, MIN_LEN, MAX_LEN)
            times.append(time.perf_counter() - t0)
            n_starts = int(r.start_list().size)
            if time.perf_counter() - t_start > budget:
                break
    dt = float(np.mean(times))
    val = n_starts / dt
    sample = (f"whole workload '{workload}': {n_items} items, {n_starts} start nodes; CycleFinder constructor "
              f"{dt:.2f} s, {threads_used} threads, {len(times)} repetitions timed")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": config_dict(workload, n_items, args.gpus),
        "reps_timed": len(times), "rep_s": [round(t, 3) for t in times], "threads_requested": threads_requested,
        "threads_used": threads_used, "cores_available": cores_avail, "fallback_reason": why or None,
        "load_s": load_s, "n_starts": n_starts, "graph_sha256": graph_sha,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads_used, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def pinned(a, torch):
    """numpy array -> numpy view of a pinned (page-locked) copy"""
    t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).copy()).pin_memory()
    return t, t.numpy().view(a.dtype)


def config_dict(workload, n_items, world):
    """the `config` object of the JSON line — identical for both arms (--impl reference times the same graph)"""
    return {"workload": f"{workload}: {WORKLOAD_TEXT.get(workload, 'synthetic community')}, "
                        f"{n_items} SDBG items (k=23), built by tools/synth_sdbg.py",
            "n_items": int(n_items), "threshold_multiplicity": THR, "cycle_min_length": MIN_LEN,
            "cycle_max_length": MAX_LEN, "l2": "graph arrays + navigation cache (>2 GB) exceed the 126 MB L2; "
                                                "no explicit flush", "parallelism": f"dls-shard x{world}",
            "timing": "CUDA events on the library stream around each step (mg_step_begin/end), max over ranks"}


def graph_digest(g):
    """sha256 over the packed SDBG arrays (w nibbles, last, tip, multiplicities) and f[]: two runs timed the same input
    <=> equal digests (the generator samples on the device it runs on, so a CPU-built and a GPU-built graph differ)"""
    import hashlib
    h = hashlib.sha256()
    for key in ("w", "last", "tip", "mult", "f"):
        h.update(np.ascontiguousarray(g[key]).tobytes())
    return h.hexdigest()


def result_digest(ss, coff, noff, nodes, inv_words):
    """sha256 over the canonical result: invalid bitmap words, searched start ids in insertion order, cycle offsets,
    node offsets, node ids (all little-endian uint64) — equal results <=> equal digests"""
    import hashlib
    h = hashlib.sha256()
    for a in (inv_words, ss, coff, noff, nodes):
        h.update(np.ascontiguousarray(a, np.uint64).tobytes())
    return h.hexdigest()


WORKLOAD_TEXT = {
    "c2": "BASELINE.json configs[1], synthetic 50-genome community + 5 strain copies",
    "c4": "BASELINE.json configs[3], synthetic high-diversity community (250 genomes + 10 strain copies, abundance sigma 2)",
    "c1": "BASELINE.json configs[0], 5 x 2 Mb genomes with 10 planted arrays, error-free",
    "c2_sample": "BASELINE.json configs[1] recipe at 1/20 scale",
    "c2_errors": "BASELINE.json configs[1] item count with the error share of real data (12 genomes, 3.5 M substitution events)",
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--sample", default="c2_sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--ingest-check", action="store_true",
                    help="also time mg_graph_create_from_files on the sample workload written in MEGAHIT's on-disk format")
    ap.add_argument("--reads-check", action="store_true",
                    help="also time mg_map_reads (SURVEY 8(f) N3) on synthetic reads of the sample workload")
    ap.add_argument("--reads", type=int, default=400000, help="reads per file for --reads-check")
    ap.add_argument("--reads-full", action="store_true", help="--reads-check also on the full bench workload (array-built graph)")
    ap.add_argument("--sweep", action="store_true",
                    help="BASELINE.json configs[4]: after the timed steps, time the hot path on the same resident graph at "
                         "cycle bounds 27-77 / 20-100 x threshold_multiplicity 10 / 20 / 50 (1 warm-up + 2 timed passes each)")
    ap.add_argument("--no-dropin", action="store_true",
                    help="skip the leg that times the C++ drop-in's CycleFinder constructor (shim_cyclefinder) on the bench workload")
    ap.add_argument("--regions-check", action="store_true",
                    help="also time mg_keep_regions (SURVEY 8(f) N4, k = 128) on the cycles of the bench workload")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import synth_sdbg as S
    from mcaat_b200 import DeviceGraph, gpu

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL prints its version banner on stdout when the communicator is created; stdout must carry ONE JSON line,
        # so file descriptor 1 points at stderr until the first collective has run
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)

    # ---- workload: built on the GPU by rank 0; raw arrays broadcast over NCCL (NVLink) to the other ranks
    t_gen = time.perf_counter()
    graph_sha = None
    if rank == 0:
        graph, info = S.make_workload(args.workload, device="cuda")
        torch.cuda.empty_cache()
        graph_sha = graph_digest(graph)
    if world > 1:
        meta = [dict(n=graph["n"], k=graph["k"], f=[int(x) for x in graph["f"]], info=info,
                     sizes=[graph[k].size for k in ("w", "last", "tip", "mult")])] if rank == 0 else [None]
        dist.broadcast_object_list(meta, src=0)
        m = meta[0]
        dev = {}
        tb = time.perf_counter()
        for key, size, dt in zip(("w", "last", "tip", "mult"), m["sizes"], (torch.int64, torch.int64, torch.int64, torch.int16)):
            if rank == 0:
                t = torch.from_numpy(graph[key].view(np.int64 if dt == torch.int64 else np.int16)).cuda()
            else:
                t = torch.empty(size, dtype=dt, device="cuda")
            dist.broadcast(t.view(torch.uint8), src=0)  # NCCL has no 16-bit integer type: broadcast raw bytes
            dev[key] = t
        torch.cuda.synchronize()
        bcast_s = time.perf_counter() - tb
        packed_dev = dict(n=m["n"], k=m["k"], f=np.array(m["f"], np.int64), **dev)
        info = m["info"]
        g = DeviceGraph(packed_dev, device=local, memory=gpu.MG_MEM_DEVICE)
        n_items = m["n"]
    else:
        bcast_s = 0.0
        g = DeviceGraph(graph, device=local)
        n_items = graph["n"]
    gen_s = time.perf_counter() - t_gen

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up, then EXACTLY --steps timed steps.  nvidia-smi starts before the warm-up (its start-up
    #      takes driver locks for about a second and would stall a timed step); only the samples taken
    #      during the timed region are used.
    clocks = ClockSampler(local)
    if rank == 0 and not os.environ.get("BENCH_NO_CLOCKS"):
        try:
            uuid = str(torch.cuda.get_device_properties(local).uuid)
        except Exception:
            uuid = None
        clocks.start(uuid)
    for _ in range(args.warmup):
        stats, res_bytes = hot_path_step(g, rank, world, dist, torch)
    g.timers_reset()
    launches0 = gpu.launch_count()
    barrier()
    clocks.rows_before = len(clocks.rows)
    t0 = time.perf_counter()
    dev_ms, step_ms = 0.0, []
    for _ in range(args.steps):
        g.step_begin()
        stats, res_bytes = hot_path_step(g, rank, world, dist, torch)
        step_ms.append(g.step_end())
        dev_ms += step_ms[-1]
    barrier()
    wall = time.perf_counter() - t0
    launches = gpu.launch_count() - launches0
    timers = g.timers()
    clk = clocks.stop() if rank == 0 else None
    # max over ranks of the CUDA-event time of the K steps
    if world > 1:
        t = torch.tensor([dev_ms, wall * 1e3], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, wall_ms = float(t[0]), float(t[1])
        lt = torch.tensor([launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(lt)
        launches = int(lt.item())
    else:
        wall_ms = wall * 1e3
    ms_per_step = dev_ms / args.steps
    value = stats["n_starts"] / (ms_per_step / 1e3)
    # ---- what the step computed, as a digest: equal on every rank, and equal for every N
    result_sha = result_digest(*hot_path_step.last_result, g.invalid_words())
    graph_device_bytes = g.device_bytes()
    ranks_agree = True
    if world > 1:
        mine = torch.tensor(list(bytes.fromhex(result_sha)), dtype=torch.uint8, device="cuda")
        every = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        ranks_agree = all(bool(torch.equal(e, every[0])) for e in every)

    # ---- BASELINE.json configs[4]: search-explosion sweep on the resident graph (not part of the timed steps above)
    sweep = None
    if args.sweep:
        global THR, MIN_LEN, MAX_LEN
        saved, sweep = (THR, MIN_LEN, MAX_LEN), []
        for mn, mx in ((27, 77), (20, 100)):
            for thr in (10, 20, 50):
                THR, MIN_LEN, MAX_LEN = thr, mn, mx
                hot_path_step(g, rank, world, dist, torch)
                g.timers_reset()
                barrier()
                ms = []
                for _ in range(2):
                    g.step_begin()
                    st_s, _ = hot_path_step(g, rank, world, dist, torch)
                    ms.append(g.step_end())
                barrier()
                tm = g.timers()
                sweep.append(dict(threshold_multiplicity=thr, cycle_min_length=mn, cycle_max_length=mx,
                                  ms_per_step=round(sum(ms) / len(ms), 2), start_nodes_per_s=round(st_s["n_starts"] / (sum(ms) / len(ms) / 1e3), 1),
                                  phase_ms={k: round(v["ms"] / len(ms), 2) for k, v in tm.items() if v["launches"]},
                                  **{k: st_s[k] for k in ("n_candidates", "n_starts", "n_searched", "n_cycles", "n_bound_hits",
                                                          "n_cap_hits", "dls_expansions", "expansions")}))
                if rank == 0:
                    print("[sweep]", json.dumps(sweep[-1]), file=sys.stderr, flush=True)
        THR, MIN_LEN, MAX_LEN = saved

    # ---- end to end through the C ABI with HOST buffers: upload + index build + search + read-back inside the timed
    #      region.  N > 1: rank 0 holds the pinned host arrays; upload on rank 0, NCCL broadcast of the raw arrays,
    #      every rank builds its replica and runs the step (admission searches sharded), rank 0 reads the results back.
    e2e = None
    if not args.no_e2e:
        keep, host, h2d = [], None, 0
        if rank == 0:
            host = dict(n=graph["n"], k=graph["k"], f=graph["f"])
            keys = ["w", "last", "tip", "mult"]
            src = dict(graph)
            if world == 1:
                # multiplicities as the reference holds them in memory (libs/megahit/src/sdbg/sdbg.h:96-105): one byte per
                # item, sentinel 255 -> (id, value) table.  (N > 1 broadcasts the uint16 array the device keeps.)
                mult = np.asarray(graph["mult"])
                big = mult >= 255
                src["mult8"] = np.where(big, 255, mult).astype(np.uint8)
                src["large_ids"] = np.flatnonzero(big).astype(np.uint64)
                src["large_vals"] = np.ascontiguousarray(mult[big], np.uint16)
                keys = ["w", "last", "tip", "mult8", "large_ids", "large_vals"]
            for key in keys:
                t, v = pinned(src[key], torch)
                keep.append(t)
                host[key] = v
            h2d = sum(host[k].nbytes for k in keys)
            t_inv, inv_buf = pinned(np.zeros((graph["n"] + 63) // 64, np.uint64), torch)  # page-locked like the inputs
            keep.append(t_inv)
        g.close()
        if world > 1:
            del dev, packed_dev
            torch.cuda.empty_cache()
        e2e_t = []
        for i in range(1 + max(1, min(args.steps, 2))):
            barrier()
            ta = time.perf_counter()
            if world == 1:
                g2 = DeviceGraph(host, device=local)
            else:
                dev2 = {}
                for key, size, dt in zip(("w", "last", "tip", "mult"), m["sizes"], (torch.int64, torch.int64, torch.int64, torch.int16)):
                    if rank == 0:
                        t = torch.from_numpy(host[key].view(np.int64 if dt == torch.int64 else np.int16)).cuda(non_blocking=True)
                    else:
                        t = torch.empty(size, dtype=dt, device="cuda")
                    dist.broadcast(t.view(torch.uint8), src=0)
                    dev2[key] = t
                torch.cuda.synchronize()
                g2 = DeviceGraph(dict(n=m["n"], k=m["k"], f=np.array(m["f"], np.int64), **dev2), device=local, memory=gpu.MG_MEM_DEVICE)
            t_up = time.perf_counter()
            st2, rb = hot_path_step(g2, rank, world, dist, torch)
            t_hp = time.perf_counter()
            inv = g2.invalid_words(out=inv_buf) if rank == 0 else None
            barrier()
            tb_ = time.perf_counter()
            if os.environ.get("BENCH_DEBUG") and rank == 0:
                print("[e2e wall ms] upload+index %.1f hot path %.1f invalid read-back %.1f" %
                      ((t_up - ta) * 1e3, (t_hp - t_up) * 1e3, (tb_ - t_hp) * 1e3), file=sys.stderr)
            g2.close()
            if world > 1:
                del dev2
            if i > 0:
                e2e_t.append(tb_ - ta)
        dt = float(np.mean(e2e_t))
        if world > 1:
            tt_ = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt_, op=dist.ReduceOp.MAX)
            dt = float(tt_[0])
        if rank == 0:
            e2e = dict(value=st2["n_starts"] / dt, unit=UNIT, h2d_bytes_per_step=int(h2d),
                       d2h_bytes_per_step=int(inv.nbytes + rb), ms_per_step=dt * 1e3,
                       nccl_broadcast_bytes_per_step=int(h2d) if world > 1 else 0)

    # ---- the same hot path on the bounded sample the CPU arms are timed on, so that the two can be compared on
    #      identical input as well (start nodes/s is not scale-free: the sample has 1/20 of the items)
    sample_leg = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        gsm, _ = S.make_workload(args.sample, device="cuda")
        gq = DeviceGraph(gsm, device=local)
        for _ in range(2):
            st_s, _ = hot_path_step(gq)
        ts_ = []
        for _ in range(3):
            gq.step_begin()
            st_s, _ = hot_path_step(gq)
            ts_.append(gq.step_end())
        gq.close()
        del gsm
        sample_leg = dict(workload=args.sample, n_items=int(gq.n), n_starts=int(st_s["n_starts"]),
                          ms_per_step=float(np.mean(ts_)), value=st_s["n_starts"] / (float(np.mean(ts_)) / 1e3), unit=UNIT,
                          note="graph resident, CUDA events around the step; cpu_baseline and --impl reference run this same graph")
    # ---- SURVEY 8(d)'s T for the drop-in a maintainer links: wall time of the C++ `CycleFinder(settings)` constructor
    #      (mcaat_b200/host/cycle_finder_gpu.cpp behind the reference's header) on the bench workload, in a fresh process:
    #      graph onto the device (decoded from the graph files / extracted through the SDBG getters), search, results
    #      map, write-back of the invalid bits into the host SDBG.  SDBG::LoadFromFile (host) is outside, as in the CPU arm.
    dropin = None
    shim = os.path.join(ROOT, "mcaat_b200", "host", "_build", "shim_cyclefinder")
    if rank == 0 and world == 1 and not args.no_dropin and os.path.exists(shim):
        gs, _ = S.make_workload(args.workload, device="cuda", want_labels=True)
        tmp = tempfile.mkdtemp(prefix="mcaat_dropin_")
        prefix = os.path.join(tmp, "graph")
        S.write_megahit_files(gs, prefix)
        del gs
        dropin = {}
        for mode, key in (("1", "device_ingest"), ("0", "sdbg_getters")):
            best = None
            for _ in range(2):
                p = subprocess.run([shim, prefix, "-", str(min(16, os.cpu_count() or 1)), str(THR), str(MIN_LEN), str(MAX_LEN)],
                                   capture_output=True, text=True, env=dict(os.environ, MCAAT_GPU_INGEST_FILES=mode, MCAAT_GPU_TIMING="1", MCAAT_REF_COUNT_STARTS="1"))
                js = [json.loads(ln[9:]) for ln in p.stdout.splitlines() if ln.startswith("REF_JSON ")]
                tm = [json.loads(ln[16:]) for ln in p.stderr.splitlines() if ln.startswith("GPU_SHIM_TIMING ")]
                if p.returncode == 0 and js and (best is None or js[-1]["ctor_s"] < best["ctor_ms"] / 1e3):
                    best = dict(ctor_ms=js[-1]["ctor_s"] * 1e3, host_load_ms=js[-1]["load_s"] * 1e3, n_starts=js[-1]["start_nodes"],
                                cycles=js[-1]["cycles"], breakdown=tm[-1] if tm else None)
            dropin[key] = best
        dropin["note"] = ("wall time of the C++ CycleFinder constructor in a fresh process (CUDA context creation included), best of 2; "
                          "unit of the e2e ratio: n_starts / ctor")
        shutil.rmtree(tmp, ignore_errors=True)
    ingest = None
    if rank == 0 and args.ingest_check:  # on-disk SDBG decoded on the device
        gs, _ = S.make_workload(args.sample, device="cuda", want_labels=True)
        tmp = tempfile.mkdtemp(prefix="mcaat_ingest_")
        prefix = os.path.join(tmp, "graph")
        S.write_megahit_files(gs, prefix)
        nbytes = sum(os.path.getsize(os.path.join(tmp, f)) for f in os.listdir(tmp))
        ts_ = []
        for _ in range(3):
            torch.cuda.synchronize()
            ta = time.perf_counter()
            gi = DeviceGraph.from_files(prefix, device=local)
            ts_.append(time.perf_counter() - ta)
            n_i = gi.n
            gi.close()
        ingest = dict(workload=args.sample, n_items=int(n_i), file_bytes=int(nbytes), seconds=min(ts_),
                      items_per_s=n_i / min(ts_), note="files in page cache; read + H2D + device decode + index build")
    reads = None
    if rank == 0 and args.reads_check:  # read -> node ids, graph (with its k-mer index) resident
        def reads_leg(name, via_files):
            gs, _ = S.make_workload(name, device="cuda", want_labels=True)
            tmp = tempfile.mkdtemp(prefix="mcaat_reads_")
            prefix = os.path.join(tmp, "graph")
            if via_files:  # the index comes with the decoded graph files
                S.write_megahit_files(gs, prefix)
                gq = DeviceGraph.from_files(prefix, device=local)
            else:          # array-built graph + mg_kmer_index_create
                gq = DeviceGraph(gs, device=local)
                gq.set_kmer_index(gs["tip"], *S.kmer_index_arrays(gs))
            hot_path_step(gq)
            cyc = np.unique(hot_path_step.last_cycle_nodes).astype(np.uint64)
            chars, roff = S.sample_reads(name, args.reads, seed=7)
            keep_pinned, chars = pinned(chars, torch)  # page-locked, like the e2e leg's inputs
            best = None
            for _ in range(3):
                gq.timers_reset()
                torch.cuda.synchronize()
                ta = time.perf_counter()
                o1, i1 = gq.map_reads((chars, roff), cyc, reverse=False)  # "file 1"
                o2, i2 = gq.map_reads((chars, roff), cyc, reverse=True)   # "file 2": reverse-complemented first
                dt = time.perf_counter() - ta
                dev = gq.timers().get("map_reads", {}).get("ms")
                if best is None or dt < best[0]:
                    best = (dt, dev)
            n_r = 2 * args.reads
            kept = int((np.diff(o1.astype(np.int64)) > 0).sum() + (np.diff(o2.astype(np.int64)) > 0).sum())
            lookups = 2 * n_r + int(i1.size + i2.size) - 2 * kept
            out = dict(workload=name, n_items=int(gq.n), index="graph files" if via_files else "mg_kmer_index_create",
                       reads=n_r, read_len=150, cycle_nodes=int(cyc.size), kept=kept, kmer_lookups=lookups,
                       null_ids=int((i1 == np.uint64(2**64 - 1)).sum() + (i2 == np.uint64(2**64 - 1)).sum()),
                       wall_ms=best[0] * 1e3, device_ms=best[1], reads_per_s=n_r / best[0],
                       lookups_per_s_device=lookups / (best[1] / 1e3) if best[1] else None,
                       h2d_bytes=int(2 * chars.nbytes + 2 * roff.nbytes), d2h_bytes=int(8 * (i1.size + i2.size + o1.size + o2.size)),
                       note="wall = H2D of the reads (pinned) + lookups + D2H of the ids, through the C ABI; graph resident")
            if via_files and not args.no_cpu_baseline:  # the oracle's literal restatement on a bounded sample, 1 thread
                sys.path.insert(0, os.path.join(ROOT, "oracle"))
                import oracle_py as O
                og = O.OracleGraph.load(prefix)
                m = min(args.reads, 20000)
                sub = [bytes(chars[int(roff[i]):int(roff[i + 1])]) for i in range(m)]
                ta = time.perf_counter()
                want = og.map_reads(sub, cyc, reverse=False)
                dt = time.perf_counter() - ta
                ok = all(np.array_equal(w_, i1[int(o1[j]):int(o1[j + 1])]) for j, w_ in enumerate(want))
                out["cpu_baseline"] = dict(kind="port", cores=1, reads=m, seconds=dt, reads_per_s=m / dt, matches_gpu=bool(ok))
            gq.close()
            del gs
            shutil.rmtree(tmp, ignore_errors=True)
            return out
        reads = reads_leg(args.sample, True)
        if args.reads_full:
            reads["full_size"] = reads_leg(args.workload, False)
    regions = None
    if rank == 0 and args.regions_check:  # k-hop region extraction around the cycles, device resident
        gr = DeviceGraph(graph, device=local)
        hot_path_step(gr)
        seeds = hot_path_step.last_cycle_nodes.astype(np.uint64)
        inv_before = gr.invalid_words().copy()
        ts_, st_r = [], None
        for _ in range(3):
            gr.timers_reset()
            torch.cuda.synchronize()
            ta = time.perf_counter()
            st_r = gr.keep_regions(seeds, 128)
            ts_.append((time.perf_counter() - ta, gr.timers().get("keep_regions", {}).get("ms")))
            gr.reset()  # back to the loaded graph, then to the pruned state the stage starts from
            gr.prune()
        assert np.array_equal(gr.invalid_words(), inv_before)
        gr.close()
        best = min(ts_)
        regions = dict(workload=args.workload, k=128, seed_ids=int(seeds.size), n_kept=st_r["n_kept"],
                       n_invalidated=st_r["n_invalidated"], levels=st_r["levels"], wall_ms=best[0] * 1e3,
                       device_ms=best[1], items_per_s=n_items / best[0],
                       note="graph resident; wall includes the id upload and the counter read-backs")
    if rank != 0:
        return
    # ---- roofline of the dominant kernel family (CUDA-event time accumulated inside the library)
    peak, peak_kind = peaks()
    alg = {  # algorithmic bytes per step and family
        "collect_tips": n_items * BYTES_STREAM_NODE, "invalidate_mult1": n_items * BYTES_STREAM_NODE,
        "nav_cache": n_items * BYTES_STREAM_NODE, "candidates": n_items * BYTES_STREAM_NODE,
        "peel": stats["n_peeled"] * BYTES_EXPANSION, "dls": stats["dls_expansions"] / world * BYTES_EXPANSION,
        "enumerate": stats["expansions"] * BYTES_EXPANSION,
    }
    fam = {k: v for k, v in timers.items() if k in alg and v["launches"]}
    top = max(fam, key=lambda k: fam[k]["ms"])
    per_launch_ms = fam[top]["ms"] / fam[top]["launches"]
    launches_per_step = fam[top]["launches"] / args.steps
    achieved = alg[top] / launches_per_step / (per_launch_ms * 1e-3) / 1e9
    traffic = None  # DRAM bytes per launch of that kernel from the committed ncu --set full capture (same workload)
    if args.workload == "c2":
        try:
            with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as f:  # written from the ncu capture of THIS kernel version
                traffic = json.load(f).get(top, {}).get("bytes_per_launch")
        except Exception:
            traffic = None
    roofline = dict(bound="hbm", kernel=top, achieved=achieved, peak=peak, unit="GB/s", frac=achieved / peak,
                    traffic=traffic, peak_source=peak_kind, ms_per_launch=per_launch_ms,
                    algorithmic_bytes_per_launch=alg[top] / launches_per_step)
    # the same figure for every phase and for the whole step ("per phase and overall")
    roofline["by_phase"] = {k: dict(ms_per_step=round(v["ms"] / args.steps, 3),
                                    achieved=round(alg[k] / (v["ms"] / args.steps * 1e-3) / 1e9, 2),
                                    frac=alg[k] / (v["ms"] / args.steps * 1e-3) / 1e9 / peak) for k, v in fam.items()}
    roofline["overall"] = dict(achieved=sum(alg[k] for k in fam) / (ms_per_step * 1e-3) / 1e9,
                               frac=sum(alg[k] for k in fam) / (ms_per_step * 1e-3) / 1e9 / peak)
    phases = {k: round(v["ms"] / args.steps, 3) for k, v in timers.items() if v["launches"]}
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu_baseline_port.with_regions = bool(regions)
        cpu = cpu_baseline_port(args.sample, "cuda")
        if regions:
            regions["cpu_baseline"] = getattr(cpu_baseline_port, "regions", None)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": config_dict(args.workload, n_items, world),
        "result_sha": result_sha, "ranks_agree": bool(ranks_agree),
        "cycle_find_wall_ms": wall_ms / args.steps, "step_ms": [round(x, 2) for x in step_ms], "gpu_launches": int(launches), "clocks": clk,
        "counts": {k: stats[k] for k in ("n_tips_before", "n_mult1", "n_peeled", "n_valid", "n_tips_after", "n_candidates",
                                          "n_starts", "n_searched", "n_cycles", "n_skipped", "n_bound_hits", "n_cap_hits",
                                          "n_reruns", "n_rounds", "n_searches_run", "n_abandoned", "n_out_of_order", "dls_expansions", "expansions")},
        "phase_ms_per_step": phases,
        "rates": {"nodes_scanned_per_s": n_items / (ms_per_step / 1e3),
                  "candidates_per_s": stats["n_candidates"] / max(phases.get("dls", 0) / 1e3, 1e-9),
                  "edges_traversed_per_s": (stats["dls_expansions"] + stats["expansions"] + stats["n_peeled"]) / (ms_per_step / 1e3)},
        "roofline": roofline, "setup": {"generate_s": round(gen_s, 2), "nccl_broadcast_s": round(bcast_s, 4),
                                        "graph_device_bytes": graph_device_bytes, "graph_sha256": graph_sha},
    }
    if sweep:
        line["sweep"] = sweep
    if ingest:
        line["ingest"] = ingest
    if regions:
        line["regions"] = regions
    if reads:
        line["reads"] = reads
    if e2e:
        line["e2e"] = e2e
    if dropin:
        line["e2e_dropin"] = dropin
    if sample_leg:
        line["same_input_as_cpu"] = sample_leg
    if cpu:
        line["cpu_baseline"] = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
