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


# ---- MetropolisHastingsSampler::sample as a chain: the reference's own fixture, tests/sampler.cpp:86-127 ----------------
# golden.json["sampler"] holds what the UNMODIFIED reference's sampler returned (oracle/_ref `sampler` command: 4 states of the
# six-vertex graph, 20 generator seeds x 10 000 samples each).  A chain's stored vertex advances, so its law is the model's
# weight distribution over N(cur); the statistic is Pearson's chi-square of 20 keyed chains against the reference's pooled
# frequencies.  Consecutive MH samples are positively correlated (a rejected proposal repeats the stored vertex), which
# inflates the statistic by the chain's integrated autocorrelation time; the reference's own 20 runs measure that factor.

def _pooled(counts_by_run, n=6):
    tot = np.zeros(n)
    for run in counts_by_run:
        for k, v in run.items():
            tot[int(k)] += v
    return tot


def _chi2(obs, expect_freq):
    keep = expect_freq > 0
    expect = expect_freq[keep] * obs.sum()
    return float(((obs[keep] - expect) ** 2 / expect).sum()), int(keep.sum()) - 1


def _reference_inflation(entry):
    """Mean chi-square per degree of freedom of the reference's own single runs against its pooled frequencies."""
    pooled = _pooled([r["counts"] for r in entry["runs"]])
    freq = pooled / pooled.sum()
    stats = [_chi2(_pooled([r["counts"]]), freq) for r in entry["runs"]]
    return max(1.0, float(np.mean([c / max(dof, 1) for c, dof in stats])))


def _chain_law_ok(entry, counts_of_seed):
    pooled_ref = _pooled([r["counts"] for r in entry["runs"]])
    freq = pooled_ref / pooled_ref.sum()
    ours = np.zeros(6)
    for seed in range(20):
        ours += counts_of_seed(seed)
    assert ours.sum() == 20 * entry["iters"]
    assert ((ours > 0) == (pooled_ref > 0)).all()                      # the same support: N(cur)
    chi2, dof = _chi2(ours, freq)
    # both pools are samples (the reference's frequencies carry the same noise): the difference has twice the variance
    limit = 2.0 * _reference_inflation(entry) * (dof + 6.0 * np.sqrt(2.0 * dof))
    assert chi2 < limit, (chi2, limit, ours / ours.sum(), freq)


@pytest.mark.parametrize("index", range(4))
def test_sampler_chain_of_the_restatement_has_the_reference_law(golden, wo, index):
    from conftest import SIX_EDGES, SIX_OFFSETS
    entry = golden["sampler"][index]
    model = wo.NODE2VEC if entry["model"] == "node2vec" else wo.DEEPWALK

    def counts(seed):
        o = wo.Oracle(6, SIX_OFFSETS, SIX_EDGES, walks_per_vertex=2, walk_length=8, model=model, p=entry["p"], q=entry["q"],
                      deterministic=False, seed=seed)
        init, c = o.sampler_chain(entry["cur"], entry["prev"], entry["iters"])
        row_ptr, col = o.csr()
        nb = col[row_ptr[entry["cur"]]:row_ptr[entry["cur"] + 1]]
        assert init in nb.tolist() and init == o.sampler_init_value(entry["cur"], entry["prev"])
        out = np.zeros(6)
        out[nb] = c
        return out

    _chain_law_ok(entry, counts)


@pytest.mark.gpu
@pytest.mark.parametrize("index", range(4))
def test_sampler_chain_on_the_device(golden, wo, wharf, index):
    """wmh_sampler_chain: bit-exact against the restatement for every seed, and the reference's law (chi-square)."""
    from conftest import SIX_EDGES, SIX_OFFSETS
    entry = golden["sampler"][index]
    model = wharf.NODE2VEC if entry["model"] == "node2vec" else wharf.DEEPWALK
    omodel = wo.NODE2VEC if entry["model"] == "node2vec" else wo.DEEPWALK
    nb = SIX_EDGES[int(SIX_OFFSETS[entry["cur"]]):int(SIX_OFFSETS[entry["cur"] + 1]) if entry["cur"] + 1 < 6 else len(SIX_EDGES)]

    def counts(seed):
        cfg = wharf.Config(walks_per_vertex=2, walk_length=8, random_walk_model=model, paramP=entry["p"], paramQ=entry["q"],
                           deterministic_mode=False, seed=seed)
        with wharf.WharfMH(6, len(SIX_EDGES), SIX_OFFSETS, SIX_EDGES, config=cfg) as e:
            init, c = e.sampler_chain(entry["cur"], entry["prev"], entry["iters"])
            if seed < 4:                                                # also with the interleaved row layout of a generated MH corpus
                e.generate_initial_random_walks()
                init2, c2 = e.sampler_chain(entry["cur"], entry["prev"], entry["iters"])
                assert init2 == init and np.array_equal(c2, c)
        o = wo.Oracle(6, SIX_OFFSETS, SIX_EDGES, walks_per_vertex=2, walk_length=8, model=omodel, p=entry["p"], q=entry["q"],
                      deterministic=False, seed=seed)
        oinit, oc = o.sampler_chain(entry["cur"], entry["prev"], entry["iters"])
        assert init == oinit and np.array_equal(c, oc)                  # keyed draws: the same chain, bit for bit
        out = np.zeros(6)
        out[np.asarray(nb)] = c
        return out

    _chain_law_ok(entry, counts)
