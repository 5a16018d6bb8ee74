"""Start-node selection sharded over ranks (one process per GPU).

The admission searches (reference src/cycle_finder.cpp:248-343, called per candidate at :411) are
independent of each other, so the ascending candidate list is dealt block-cyclically over the ranks,
each rank searches its share, and the verdict bytes are all-gathered; there is no collective inside a
search.  Everything else of the hot path runs replicated.
"""
import numpy as np


def my_share(cand, rank, world):
    """candidates this rank searches: positions rank, rank + world, ... of the ascending list"""
    return cand[rank::world]


def gather_verdicts(ok_mine, n_total, rank, world, dist, torch, device):
    """all-gather of the per-rank verdicts -> bool array aligned with the full candidate list"""
    per = (n_total + world - 1) // world
    buf = torch.zeros(max(per, 1), dtype=torch.uint8, device=device)
    mine = np.ascontiguousarray(ok_mine, dtype=np.uint8)
    if mine.size:
        buf[:mine.size] = torch.from_numpy(mine).to(device)
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf)
    ok = np.zeros(n_total, bool)
    for r in range(world):
        cnt = len(range(r, n_total, world))
        ok[r::world] = out[r].cpu().numpy()[:cnt].astype(bool)
    return ok
