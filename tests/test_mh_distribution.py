"""Metropolis-Hastings mode against the REAL reference, distributionally.

The reference keys nothing in this mode (one global generator, config/globals.h:26), so bit parity is defined against
the keyed restatement (oracle/wharf_oracle.c; tests/test_gpu_parity.py::test_keyed_mh_mode_vs_oracle).  What pins the
restatement itself to the reference is the law of the walks: tests/golden/mh_stats.json holds, for 20 generator seeds
of the unmodified reference on cora (10 x 80, five model/init configurations), the fractions of node2vec's three step
classes (return to prev / common neighbour of prev and cur / away) and the visit shares of the 64 most visited
vertices (tests/golden/make_mh_stats.py).  The same statistics of 20 keyed seeds must agree:
    class fractions   |mean_ref - mean_ours| <= 4.5 * sqrt(var_ref / 20 + var_ours / 20) + 2e-4
    visit shares      the same bound with 5 standard errors + 2e-5, for each of the 64 vertices
The statistic discriminates: in the reference's own numbers the return fraction for p = 0.5, q = 2 is 0.749 (WEIGHT),
0.424 (RANDOM), 0.558 (BURNIN) with standard errors below 0.001, so a restatement that treated the stored sampler
differently (e.g. let it advance instead of staying at its init value) cannot pass.
"""
import importlib.util
import json
from pathlib import Path

import numpy as np
import pytest

HERE = Path(__file__).resolve().parent
SEEDS = 20


def _stats_module():
    spec = importlib.util.spec_from_file_location("make_mh_stats", HERE / "golden" / "make_mh_stats.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="module")
def mh_gold():
    return json.loads((HERE / "golden" / "mh_stats.json").read_text())


def _compare(cfg, ours, shares):
    ref = cfg["runs"]
    for k in ("return", "common", "away"):
        a, b = np.array([r[k] for r in ref]), np.array([r[k] for r in ours])
        bound = 4.5 * np.sqrt(a.var(ddof=1) / len(a) + b.var(ddof=1) / len(b)) + 2e-4
        assert abs(a.mean() - b.mean()) <= bound, (cfg["model"], cfg["init"], k, a.mean(), b.mean(), bound)
    top = np.array(cfg["top_vertices"])
    mine = np.array(shares)[:, top]
    ref_mean, ref_std = np.array(cfg["top_share_mean"]), np.array(cfg["top_share_std"])
    bound = 5.0 * np.sqrt(ref_std ** 2 / SEEDS + mine.var(axis=0, ddof=1) / SEEDS) + 2e-5
    worst = np.abs(mine.mean(axis=0) - ref_mean) - bound
    assert (worst <= 0).all(), (cfg["model"], cfg["init"], int(top[np.argmax(worst)]), float(worst.max()))


INIT = {"weight": 2, "random": 0, "burnin": 1}


@pytest.mark.parametrize("index", range(5))
def test_keyed_restatement_has_the_reference_law(mh_gold, cora, wo, index):
    """CPU: the keyed plain-C restatement vs the unmodified reference."""
    ms = _stats_module()
    cfg = mh_gold["configs"][index]
    n, off, adj = cora
    model = wo.NODE2VEC if cfg["model"] == "node2vec" else wo.DEEPWALK
    ours, shares = [], []
    for seed in range(SEEDS):
        o = wo.Oracle(n, off, adj, walks_per_vertex=mh_gold["wpv"], walk_length=mh_gold["len"], model=model, p=cfg["p"], q=cfg["q"],
                      deterministic=False, init_strategy=INIT[cfg["init"]], seed=500 + seed)
        o.generate()
        st = ms.transition_stats(o.walks(), off, adj)
        shares.append(st.pop("visit_share"))
        ours.append(st)
        o.close()
    _compare(cfg, ours, shares)


@pytest.mark.gpu
@pytest.mark.parametrize("index", range(5))
def test_engine_has_the_reference_law(mh_gold, cora, wharf, index):
    """GPU: the CUDA engine (through the C ABI) vs the unmodified reference."""
    ms = _stats_module()
    cfg = mh_gold["configs"][index]
    n, off, adj = cora
    model = wharf.NODE2VEC if cfg["model"] == "node2vec" else wharf.DEEPWALK
    ours, shares = [], []
    for seed in range(SEEDS):
        c = wharf.Config(walks_per_vertex=mh_gold["wpv"], walk_length=mh_gold["len"], random_walk_model=model, paramP=cfg["p"], paramQ=cfg["q"],
                         deterministic_mode=False, sampler_init_strategy=INIT[cfg["init"]], seed=900 + seed)
        with wharf.WharfMH(n, int(adj.shape[0]), off, adj, config=c) as e:
            e.generate_initial_random_walks()
            st = ms.transition_stats(e.walks(), off, adj)
        shares.append(st.pop("visit_share"))
        ours.append(st)
    _compare(cfg, ours, shares)
