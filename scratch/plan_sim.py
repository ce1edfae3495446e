"""Simulate band_plan_kernel (narrow cuts, nbcap) on CPU-oracle trees of the bench inputs."""
import sys; sys.path.insert(0, '.')
import numpy as np, torch
from oracle import tree_cpu
tree_cpu.build(ref=False)
from bench import make_inputs
from scratch.tree_stats_cpu_lib import level_starts

def plan(ls, nl, pmax, kdiv, nbcap=48, frac=4):
    kmax = max(4, -(-nl // kdiv))
    esc = 0
    while True:
        cuts = [0]; L0 = 0; over = False
        while L0 < nl:
            if len(cuts) - 1 >= nbcap: over = True; break
            hi = min(L0 + kmax, nl); lo = L0 + 1
            while lo < hi:
                mid = (lo + hi + 1) // 2
                if ls[mid] - ls[L0] <= pmax: lo = mid
                else: hi = mid - 1
            if lo < nl and frac:
                first = max(L0 + 1, lo - (lo - L0) // frac)
                best, bw = lo, ls[lo] - ls[lo - 1]
                for L1 in range(lo - 1, first - 1, -1):
                    w = ls[L1] - ls[L1 - 1]
                    if w < bw: bw, best = w, L1
                lo = best
            cuts.append(lo); L0 = lo
        if not over: return cuts, esc, pmax
        pmax += pmax // 4; kmax += kmax // 4 + 1; esc += 1

if __name__ == '__main__':
    trees = level_starts()
    for kdiv in (28, 32, 36, 40):
        for frac in (0, 4, 3):
            row = []
            for name, ls in trees:
                nl = len(ls) - 1
                cuts, esc, pm = plan(ls, nl, 2288, kdiv, frac=frac)
                nb = len(cuts) - 1
                T = sum(ls[c] - ls[c - 1] for c in cuts[1:])
                maxlev = max(b - a for a, b in zip(cuts[:-1], cuts[1:]))
                maxpos = max(ls[b] - ls[a] for a, b in zip(cuts[:-1], cuts[1:]))
                row.append(f"nb={nb} esc={esc} T={T} K={maxlev} P={maxpos}")
            print(kdiv, frac, ' | '.join(row))
