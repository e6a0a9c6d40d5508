"""The reference's README score table (README.md:44-50) cell by cell: GPU fused tournament at large N vs the C oracle
playing the same protocol on the CPU (N ~ 10^3; maximin-3 cells N = 200, sampled) vs the README's own 100-game cell.
Used by tests/test_gpu_readme_table.py and scripts/readme_table_report.py."""
import math

README = {   # black (row) vs white (column): wins / draws / losses of black, 100 games, first 10 plies random
    ("rand", "rand"): (48, 6, 46), ("rand", "greedy"): (38, 1, 61), ("rand", "maximin2"): (32, 4, 64), ("rand", "maximin3"): (13, 2, 85),
    ("greedy", "rand"): (61, 5, 34), ("greedy", "greedy"): (42, 4, 54), ("greedy", "maximin2"): (27, 3, 70), ("greedy", "maximin3"): (25, 1, 74),
    ("maximin2", "rand"): (72, 1, 27), ("maximin2", "greedy"): (67, 1, 32), ("maximin2", "maximin2"): (35, 1, 64), ("maximin2", "maximin3"): (33, 2, 65),
    ("maximin3", "rand"): (78, 3, 19), ("maximin3", "greedy"): (66, 4, 30), ("maximin3", "maximin2"): (63, 2, 35), ("maximin3", "maximin3"): (46, 1, 53),
}
# the Maximin-1 row / column repeats Greedy (README.md:54): 9 more cells of the 5x5 table, checked as identities
CHI2_DF2_P001 = 13.816        # chi-square, 2 degrees of freedom, p = 0.001


def cpu_games(black, white):
    return 200 if "maximin3" in (black, white) else 1000


def cpu_tally(black, white, n_cpu=None):
    """C oracle (the reference's algorithm, cell arrays + ray walks) playing the README protocol -> [W, D, L] of black"""
    from gymothelloenv_b200.tournament import parse_policy
    from oracle import oracle
    n_cpu = n_cpu or cpu_games(black, white)
    _, c = oracle.tournament(8, 7 + sum(ord(ch) for ch in black + '/' + white), n_cpu, black=parse_policy(black),
                             white=parse_policy(white), init_rand=10)
    return list(c)


def all_cpu_tallies():
    """every cell at once, one thread per cell (ctypes releases the GIL): ~25 s instead of ~3 min"""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle
    oracle.lib()
    cells = sorted(README, key=lambda c: -("maximin3" in c) - (c == ("maximin3", "maximin3")))
    with ThreadPoolExecutor(len(cells)) as ex:
        res = list(ex.map(lambda c: cpu_tally(*c), cells))
    return dict(zip(cells, res))


def cell(black, white, cpu=None, n_gpu=None, seed=2025):
    import torch
    from gymothelloenv_b200.tournament import rollout
    deep = "maximin3" in (black, white)
    n_gpu = n_gpu or ((1 << 18) if deep else (1 << 20))
    tally, _ = rollout(n_gpu, black, white, board_size=8, initial_rand_steps=10, seed=seed)
    torch.cuda.synchronize()
    g = [int(x) for x in tally.tolist()]
    p = [g[k] / n_gpu for k in range(3)]
    c = cpu if cpu is not None else cpu_tally(black, white)
    n_cpu = sum(c)
    chi2 = sum((c[k] - n_cpu * p[k]) ** 2 / max(n_cpu * p[k], 1e-9) for k in range(3))
    rd = README[(black, white)]
    z = [(rd[k] / 100.0 - p[k]) / math.sqrt(max(p[k] * (1 - p[k]), 1e-6) / 100.0) for k in range(3)]
    return {"black": black, "white": white, "gpu_games": n_gpu, "gpu_wdl_pct": [round(100 * x, 3) for x in p],
            "plies_per_game": g[3] / n_gpu, "cpu_oracle_games": n_cpu, "cpu_oracle_wdl": list(c),
            "cpu_sampled": deep, "chi2_df2": chi2, "readme_wdl": list(rd), "readme_z": [round(x, 2) for x in z]}
