"""Noise-floor calibration against an extended-precision evaluation (SURVEY.md 8(c) oracle plan; model in
tests/_truth_model.py): how far is each fp64 evaluation of the reference's arithmetic - oracle A (literal
dense), oracle B (structured), the GPU path - from the same arithmetic carried out with 64 mantissa bits?

What it pins down: (1) the oracles' association decisions are not artefacts of fp64 rounding (identical in
long double, and no decision sits near the threshold); (2) the 1e-9 bar holds for the state and norm-wise for
Sigma; (3) on the observed block the reference's own fp64 arithmetic is only ~1e-8 accurate (its 1e7 prior
cancels at every first sighting), which is why that block is held to 1e-6 and not to 1e-9 anywhere in tests/."""
import numpy as np
import pytest

from _harness import run_gpu, run_literal, run_struct, synth
from _truth_model import run_truth

LD = np.longdouble


def _cases():
    return [("config1 known", True, synth.circle_drive(n_scans=120, seed=0)),
            ("config2 unknown", False, synth.circle_drive(n_scans=120, seed=1, drop_out_of_range=True, shuffle=True))]


def _errors(X, T, count):
    k = 3 + 2 * count
    es = float(np.linalg.norm(X[1].astype(LD) - T[1]) / np.linalg.norm(T[1]))
    eS = float(np.linalg.norm(X[2].astype(LD) - T[2]) / np.linalg.norm(T[2]))
    ea = float(np.linalg.norm(X[2][:k, :k].astype(LD) - T[2][:k, :k]) / np.linalg.norm(T[2][:k, :k]))
    return es, eS, ea


def _same_assoc(a, b):
    return [list(map(int, x)) for x in a] == [list(map(int, x)) for x in b]


@pytest.mark.parametrize("case", [0, 1])
def test_oracles_against_long_double(case):
    name, known, scans = _cases()[case]
    T = run_truth(scans, known)
    assert T[3] == 4
    if not known:
        assert min(T[4]) > 0.5            # every accept / reject decision is far from the threshold
    for X in (run_literal(scans, known), run_struct(scans, known)):
        assert _same_assoc(X[0], T[0])
        es, eS, ea = _errors(X, T, 4)
        assert es <= 1e-9 and eS <= 1e-9
        assert 1e-10 < ea <= 1e-6         # the cancellation floor of the reference's own arithmetic


@pytest.mark.gpu
@pytest.mark.parametrize("case", [0, 1])
def test_gpu_against_long_double(case):
    name, known, scans = _cases()[case]
    T = run_truth(scans, known)
    G = run_gpu(scans, known)
    A = run_literal(scans, known)
    assert _same_assoc(G[0], T[0])
    gs, gS, ga = _errors(G, T, 4)
    _, _, aa = _errors(A, T, 4)
    assert gs <= 1e-9 and gS <= 1e-9
    assert ga <= 1e-6 and ga <= 20.0 * aa      # no worse than the reference's arithmetic itself, to within its scatter
