"""The warp-uniform device logic of csrc/go_bb.cuh, compiled for the host with a 32-row array
standing in for the warp, against the reference's golden planes and against the C oracle on
fresh random playouts (CPU only; the same comparisons run on the real kernels under -m gpu)."""
import numpy as np
import pytest

from oracle import go_oracle
from tests import golden_util as gu
from tests import hostsim

ALL = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size",
       "liberties_after", "ladder_capture", "ladder_escape", "sensibleness", "zeros", "legal"]


def _explain(got, want):
    bad = np.argwhere(got != want)
    return "mismatches %d, first (pos, plane, x, y): %s" % (len(bad), bad[:6].tolist())


@pytest.mark.parametrize("name", gu.FEATURE_FILES)
def test_device_logic_matches_reference_golden(name):
    g = gu.load_features(name)
    got, _ = hostsim.features49(g["board"], g["age"], g["ko"], g["player"])
    assert np.array_equal(got, g["planes"]), _explain(got, g["planes"])


@pytest.mark.parametrize("size,n,seed", [(19, 1500, 11), (13, 400, 12), (9, 400, 13), (6, 300, 14)])
def test_device_logic_matches_oracle_on_playouts(size, n, seed):
    b, a, ko, pl = go_oracle.random_playouts(n, size=size, seed=seed, max_ply=size * size + 40, stride=5)
    want = go_oracle.features(b, a, ko, pl, ALL)
    got, stats = hostsim.features49(b, a, ko, pl)
    assert np.array_equal(got, want), _explain(got, want)


def test_device_logic_with_superko_matches_reference_golden():
    """sk_forbidden_moves + ladder_search_sk (csrc/go_bb.cuh) on the positions recorded from reference
    games played with enforce_superko=True."""
    feats, sets = gu.load_superko()
    assert feats == ALL
    for key, g in sets.items():
        size = g["board"].shape[-1]
        got = hostsim.features49_superko(g["board"], g["age"], g["ko"], g["player"], g["sk"], go_oracle.zobrist_table(size))
        assert np.array_equal(got, g["planes"]), (key, _explain(got, g["planes"]))
