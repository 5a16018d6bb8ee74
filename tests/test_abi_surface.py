"""CPU-side checks of the drop-in boundary: libmcaat_gpu.so loads, exports every symbol that
include/mcaat_gpu.h declares, and fails loudly (no CPU fallback) when no CUDA device exists."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import golden_util as G
from mcaat_b200 import gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_and_binding_list_the_same_symbols():
    hdr = open(os.path.join(ROOT, "include", "mcaat_gpu.h")).read()
    declared = set(re.findall(r"\b(mg_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(gpu.SYMBOLS)


def test_library_exports_every_declared_symbol():
    L = gpu.lib()
    for name in gpu.SYMBOLS:
        assert hasattr(L, name), name
    assert b"sm_100a" in L.mg_version()


def test_struct_layouts_match_header():
    # sizes of the ABI structs as the header lays them out (LP64)
    assert C.sizeof(gpu.GraphDesc) == 8 + 4 * 4 + 8 * 8 + 8 + 48
    assert C.sizeof(gpu.PruneStats) == 40
    assert C.sizeof(gpu.Starts) == 56
    assert C.sizeof(gpu.Cycles) == 120


def test_struct_sizes_against_the_c_compiler(tmp_path):
    """every struct of include/mcaat_gpu.h: sizeof as gcc lays it out == the ctypes mirror in mcaat_b200/gpu.py"""
    import os
    import shutil
    import subprocess
    cc = shutil.which("gcc") or shutil.which("cc")
    if not cc:
        pytest.skip("no C compiler")
    pairs = [("mg_graph_desc", gpu.GraphDesc), ("mg_prune_stats", gpu.PruneStats), ("mg_starts", gpu.Starts),
             ("mg_cycles", gpu.Cycles), ("mg_region_stats", gpu.RegionStats), ("mg_subgraph", gpu.Subgraph),
             ("mg_kmer_index_desc", gpu.KmerIndexDesc), ("mg_reads", gpu.Reads)]
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "mcaat_gpu.h"\nint main(void){' +
                   "".join(f'printf("%zu\\n", sizeof({c}));' for c, _ in pairs) + "return 0;}\n")
    exe = str(tmp_path / "sizes")
    subprocess.run([cc, "-std=c11", "-I", os.path.join(gpu.ROOT, "include"), str(src), "-o", exe], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout.split()
    for (cname, ct), size in zip(pairs, out):
        assert C.sizeof(ct) == int(size), cname


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_cuda(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback_without_a_device():
    packed = G.pack_raw(G.load_graph_npz("tiny"))
    with pytest.raises(gpu.MgError) as e:
        gpu.DeviceGraph(packed)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(gpu, "_lib", None)
    monkeypatch.setattr(gpu, "LIB_PATH", "/nonexistent/libmcaat_gpu.so")
    with pytest.raises(gpu.MgError):
        gpu.lib()


def test_cycle_finder_requires_a_graph():
    from mcaat_b200 import CycleFinder, Settings
    with pytest.raises(RuntimeError):  # reference src/cycle_finder.cpp:134-136
        CycleFinder(Settings())
