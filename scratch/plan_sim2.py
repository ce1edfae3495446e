import sys; sys.path.insert(0, '.')
from scratch.plan_sim import plan
from scratch.tree_stats_cpu_lib import level_starts
trees = level_starts()
for pmax in (3100,):
    for nbcap in (36,):
        for kdiv in (24, 26, 28, 30, 32):
            row = []
            for name, ls in trees:
                nl = len(ls) - 1
                cuts, esc, pm = plan(ls, nl, pmax, kdiv, nbcap=nbcap, frac=0)
                nb = len(cuts) - 1
                T = sum(ls[c] - ls[c - 1] for c in cuts[1:])
                maxlev = max(b - a for a, b in zip(cuts[:-1], cuts[1:]))
                row.append(f"{nb}/{esc}/{T}/{maxlev}")
            print(pmax, nbcap, kdiv, ' '.join(row))
