"""SURVEY 8(f) row N1 on the CPU: the restatement of Sampler.edge_partitioning
(oracle/sampler_oracle.py) against the golden vectors produced by the unmodified
reference method (tests/golden/epart.npz, made by tests/golden/make_epart_golden.py)."""
import os

import numpy as np

from oracle import sampler_oracle as SO

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "epart.npz"))


def golden_case(name):
    L = int(GOLD[f"{name}.layers"])
    blocks = [(GOLD[f"{name}.{l}.src"], GOLD[f"{name}.{l}.dst"]) for l in range(L)]
    want = []
    for l in range(L):
        parts = []
        for g in range(4):
            keys = [int(k) for k in GOLD[f"{name}.{l}.{g}.shuffle_keys"]]
            parts.append({"src": GOLD[f"{name}.{l}.{g}.src"], "dest": GOLD[f"{name}.{l}.{g}.dest"],
                          "owned_nds": GOLD[f"{name}.{l}.{g}.owned_nds"],
                          "shuffle_nds": {r: GOLD[f"{name}.{l}.{g}.shuffle_nds.{r}"] for r in keys}, "device": g})
        want.append(parts)
    return GOLD[f"{name}.owner"], GOLD[f"{name}.mask"], blocks, want


def same_partitioning(got, want):
    assert len(got) == len(want)
    for l, (gl, wl) in enumerate(zip(got, want)):
        for g in range(4):
            a, b = gl[g], wl[g]
            for k in ("src", "dest", "owned_nds"):
                assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), (l, g, k)
            assert sorted(a["shuffle_nds"]) == sorted(b["shuffle_nds"]), (l, g, "shuffle keys")
            for r in b["shuffle_nds"]:
                assert np.array_equal(np.asarray(a["shuffle_nds"][r]), np.asarray(b["shuffle_nds"][r])), (l, g, r)


def test_restatement_matches_the_reference_method_on_every_golden_case():
    for name in GOLD["cases"]:
        owner, mask, blocks, want = golden_case(str(name))
        same_partitioning(SO.edge_partitioning(blocks, owner, mask), want)


def test_golden_cases_cover_the_last_layer_branches():
    # a part with extra edges (shuffle_nds filled), one without (shuffle_nds == {}), a one-layer case
    owner, mask, blocks, want = golden_case("a")
    assert any(p["shuffle_nds"] for p in want[-1])
    owner, mask, blocks, want = golden_case("c")
    assert len(want) == 1
    owner, mask, blocks, want = golden_case("b")  # nothing cached beyond the own partition
    assert all(len(p["shuffle_nds"]) in (0, 3) for p in want[-1])
