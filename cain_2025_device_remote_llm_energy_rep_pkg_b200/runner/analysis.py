"""The statistics of the reference's analysis notebook in Python, so that a `run_table.csv` written on
a B200 box gets the same tables without R (`data-analysis/analysis-visualization.ipynb`; cell numbers below are the
notebook's code cells):

  * subsets per (method, length) with the IQR outlier filter applied column after column      (cells 11, 13)
  * Shapiro-Wilk on the energy column of every subset                                         (cell 33)
  * Wilcoxon rank-sum test on-device vs remote per length, Cliff's delta with its interval    (cell 37)
  * Spearman correlation of energy with time / CPU / GPU / memory per subset                  (cell 42)

numpy + scipy only.  `tests/test_analysis.py` pins the port to the numbers printed in the notebook for the reference's
own `data-analysis/run_table.csv`.  The energy column is a parameter: the reference analyses `energy_usage_J`
(powermetrics, added post hoc, RunnerConfig.py:253); a B200 table has `nvml__energy_J`.
"""
from __future__ import annotations

import csv
import math
from typing import Dict, List, Sequence

import numpy as np
from scipy import stats

ON_DEVICE, REMOTE = "on_device", "remote"
LENGTH_MAP = {"short": 100, "medium": 500, "long": 1000}                     # cell 6
ENERGY, TIME, CPU, GPU, MEMORY = "energy_usage_J", "execution_time", "cpu_usage", "gpu_usage", "memory_usage"


def load_run_table(path: str) -> List[dict]:
    with open(path, newline="") as f:
        rows = list(csv.DictReader(f))
    for r in rows:
        for k, v in list(r.items()):
            if k in ("__run_id", "__done", "model", "method", "topic"):
                continue
            try:
                r[k] = float(v)
            except (TypeError, ValueError):
                r[k] = float("nan")
    return rows


def remove_outliers(rows: List[dict], columns: Sequence[str]) -> List[dict]:
    """cell 11: for each column in turn, keep rows inside [Q1 - 1.5 IQR, Q3 + 1.5 IQR] of what is left (R quantile type 7)."""
    kept = list(rows)
    for c in columns:
        v = np.array([r[c] for r in kept], dtype=np.float64)
        if not len(v):
            break
        q1, q3 = np.nanquantile(v, 0.25), np.nanquantile(v, 0.75)
        lo, hi = q1 - 1.5 * (q3 - q1), q3 + 1.5 * (q3 - q1)
        kept = [r for r, x in zip(kept, v) if lo <= x <= hi]       # NaN compares false and is dropped, as dplyr::filter does
    return kept


def subsets(rows: List[dict], metrics: Sequence[str]) -> Dict[str, List[dict]]:
    """cell 13: {"on_device_short": [...], ...}, outliers removed over `metrics`."""
    out = {}
    for method in (ON_DEVICE, REMOTE):
        for label, n in LENGTH_MAP.items():
            out[f"{method}_{label}"] = remove_outliers([r for r in rows if r["method"] == method and r["length"] == n], metrics)
    return out


def shapiro_table(sub: Dict[str, List[dict]], energy: str = ENERGY) -> Dict[str, tuple]:
    """cell 33: (W, p) of the energy column per subset."""
    return {k: tuple(map(float, stats.shapiro([r[energy] for r in v]))) for k, v in sub.items()}


def cliffs_delta(x: Sequence[float], y: Sequence[float], conf_level: float = 0.95):
    """effsize::cliff.delta defaults (unbiased variance, Student quantile, the asymmetric interval of Feng & Cliff):
    returns (delta, lower, upper, magnitude)."""
    x, y = np.asarray(x, dtype=np.float64), np.asarray(y, dtype=np.float64)
    n1, n2 = len(x), len(y)
    dom = np.sign(x[:, None] - y[None, :])
    d = float(dom.mean())
    di, dj = dom.mean(axis=1), dom.mean(axis=0)
    ssr, ssc, sst = float(((di - d) ** 2).sum()), float(((dj - d) ** 2).sum()), float(((dom - d) ** 2).sum())
    sd = math.sqrt(max(0.0, (n2 ** 2 * ssr + n1 ** 2 * ssc - sst) / (n1 * n2 * (n1 - 1) * (n2 - 1))))
    z = float(stats.t.ppf(1 - (1 - conf_level) / 2, n1 + n2 - 2))
    root = z * sd * math.sqrt(max(0.0, 1 - 2 * d * d + d ** 4 + (z * sd) ** 2))
    den = 1 - d * d + (z * sd) ** 2
    if den == 0.0:      # complete separation with no spread: the interval collapses onto the estimate
        return d, d, d, "Large"
    lo, hi = (d - d ** 3 - root) / den, (d - d ** 3 + root) / den
    a = abs(d)
    mag = "Negligible" if a < 0.147 else "Small" if a < 0.33 else "Medium" if a < 0.474 else "Large"
    return d, lo, hi, mag


def h1_table(sub: Dict[str, List[dict]], energy: str = ENERGY) -> Dict[str, dict]:
    """cell 37: per length, wilcox.test(on_device, remote) (R's W = the Mann-Whitney U of the first sample) and Cliff's delta."""
    out = {}
    for label in LENGTH_MAP:
        a = [r[energy] for r in sub[f"{ON_DEVICE}_{label}"]]
        b = [r[energy] for r in sub[f"{REMOTE}_{label}"]]
        u = stats.mannwhitneyu(a, b, alternative="two-sided", method="asymptotic", use_continuity=True)
        d, lo, hi, mag = cliffs_delta(a, b)
        out[label] = {"W": float(u.statistic), "p": float(u.pvalue), "delta": d, "lower": lo, "upper": hi, "magnitude": mag}
    return out


def spearman_table(sub: Dict[str, List[dict]], energy: str = ENERGY, metrics: Sequence[str] = (TIME, CPU, GPU, MEMORY)) -> Dict[str, dict]:
    """cell 42: cor.test(energy, metric, method = "spearman") per subset -> (rho, p)."""
    out = {}
    for k, v in sub.items():
        e = [r[energy] for r in v]
        out[k] = {}
        for m in metrics:
            col = [r[m] for r in v]
            if len(set(col)) < 2 or len(set(e)) < 2:
                out[k][m] = (float("nan"), float("nan"))
                continue
            res = stats.spearmanr(e, col)
            out[k][m] = (float(res.statistic), float(res.pvalue))
    return out


def analyse(path: str, energy: str = ENERGY, metrics: Sequence[str] = (ENERGY, TIME, CPU, GPU, MEMORY)) -> dict:
    rows = load_run_table(path)
    metrics = [energy if m == ENERGY else m for m in metrics]
    metrics = [m for m in metrics if rows and m in rows[0] and not all(math.isnan(r[m]) for r in rows)]   # a B200 table has no CPU / RAM columns filled
    sub = subsets(rows, metrics)
    res = {"n": {k: len(v) for k, v in sub.items()}}
    if all(len(v) >= 3 for v in sub.values()):
        res["shapiro"] = shapiro_table(sub, energy)
        res["h1"] = h1_table(sub, energy)
        res["spearman"] = spearman_table(sub, energy, [m for m in metrics if m != energy])
    return res


if __name__ == "__main__":
    import argparse
    import json
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("run_table")
    ap.add_argument("--energy", default=ENERGY, help="energy column (nvml__energy_J for a B200 table)")
    a = ap.parse_args()
    print(json.dumps(analyse(a.run_table, a.energy), indent=1))
