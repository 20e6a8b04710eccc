"""Oracle A (dense literal) vs oracle B (structured C) vs the numpy model of the GPU algorithm,
on the config-1 (known association) and config-2-shaped (unknown association) traces.
Tolerances: association identical; state and whole-Sigma Frobenius <= 1e-9 relative (the
north-star criterion, norm-wise); the active block is reported against the 1e7-prior
cancellation floor: <= 1e-6 relative."""
import numpy as np
import pytest

from _harness import active_block, rel_fro, run_literal, run_model, run_struct, synth


@pytest.mark.parametrize("known", [True, False])
def test_three_way_agreement(known):
    scans = synth.circle_drive(n_scans=80, seed=0) if known else \
        synth.circle_drive(n_scans=80, seed=1, drop_out_of_range=True, shuffle=True)
    aL, sL, SL, cL = run_literal(scans, known)
    aS, sS, SS, cS = run_struct(scans, known)
    aM, sM, SM, cM = run_model(scans, known)
    assert aL == aS == aM
    if not known:
        assert cL == cS == cM == 4
    for s, S in ((sS, SS), (sM, SM)):
        assert rel_fro(s, sL) <= 1e-9
        assert rel_fro(S, SL) <= 1e-9
        assert rel_fro(active_block(S, 4), active_block(SL, 4)) <= 1e-6


def test_more_landmarks_unknown():
    field = synth.scattered_field(12, seed=5, extent=3.0, min_sep=0.5)
    scans = synth.circle_drive(n_scans=40, seed=2, landmarks=field, drop_out_of_range=True, shuffle=True, max_range=1.6)
    scans = [s for s in scans if len(s.xy)]
    aL, sL, SL, cL = run_literal(scans, False)
    aS, sS, SS, cS = run_struct(scans, False)
    aM, sM, SM, cM = run_model(scans, False)
    assert aL == aS == aM and cL == cS == cM
    assert rel_fro(sS, sL) <= 1e-9 and rel_fro(sM, sL) <= 1e-9
    assert rel_fro(SS, SL) <= 1e-9 and rel_fro(SM, SL) <= 1e-9
