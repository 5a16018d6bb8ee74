"""world_size-2 `gloo` run of the sharded start-node selection (mcaat_b200/shard.py) on CPU: every rank runs
the admission searches of its share with the host build of the device search code (tests/emu) and the
gathered verdicts must equal the serial ones — i.e. the reference's start list."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _worker(rank, world, port, name, thr, mx, out_dir):
    import torch
    import torch.distributed as dist
    import emu_py as E
    import golden_util as G
    from mcaat_b200 import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = E.EmuGraph(G.pack_raw(G.load_graph_npz(name)))
    g.prune()
    cand = g.candidates(thr)
    mine = shard.my_share(cand, rank, world)
    ok_m = np.zeros(mine.size, np.uint8)
    for i, c in enumerate(mine):
        code = g.dls(c, mx)[0]
        if code == -1:
            code = g.dls_bitmap(c, mx)
        ok_m[i] = code
    ok = shard.gather_verdicts(ok_m, cand.size, rank, world, dist, torch, "cpu")
    np.save(os.path.join(out_dir, f"starts_{rank}.npy"), cand[ok])
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_selection_equals_reference_start_list(tmp_path, world):
    import torch.multiprocessing as mp
    import golden_util as G
    import oracle_py as O
    name, thr, mn, mx, dump = [c for c in G.cases() if c[0] == "repeatrich" and c[1] == 20][0]
    port = 29600 + os.getpid() % 300 + world
    mp.spawn(_worker, args=(world, port, name, thr, mx, str(tmp_path)), nprocs=world, join=True)
    want = np.sort(O.parse_dump(dump).start_list())
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), f"starts_{r}.npy"))
        assert np.array_equal(np.sort(got.astype(np.uint64)), want.astype(np.uint64))


def test_share_and_gather_shapes():
    from mcaat_b200 import shard
    cand = np.arange(10, dtype=np.uint64)
    parts = [shard.my_share(cand, r, 3) for r in range(3)]
    assert sorted(np.concatenate(parts).tolist()) == list(range(10))
    assert [p.size for p in parts] == [4, 3, 3]
