"""Writes tests/golden/minimodel_forward.npz -- TEST INFRASTRUCTURE ONLY.

The reference's own network fixture (tests/test_data/minimodel.json + hdf5/random_minimodel_weights.hdf5,
whose arrays tests/golden/minimodel.npz holds and tests/test_formats.py checks against the HDF5 file)
evaluated on positions of the reference's SGFs (tests/golden/features_sgf.npz, planes recorded from the
reference's Preprocess) by the scipy restatement oracle/cnn_oracle2.py, after checking that the torch
restatement oracle/cnn_oracle.py gives the same numbers.  Keras itself cannot run here, so this pins
the two restatements to each other and the device path to both -- not to Keras.

    python -m oracle.make_golden_minimodel_forward
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import cnn_oracle, cnn_oracle2  # noqa: E402
from tests import golden_util as gu  # noqa: E402


def main():
    z = np.load(os.path.join(gu.GOLDEN, "minimodel.npz"))
    feats = [str(s) for s in z["feature_list"]]
    convs = [tuple(int(v) for v in c) for c in z["convs"]]
    weights = [z["w%02d" % i] for i in range(13)]
    g = gu.load_features("features_sgf.npz")
    # BASELINE config 1's positions: plies 20 / 41 / 101 / 161 of the five Lee Sedol games, plus every 7th position
    names = [str(s) for s in g["extra"]["names"]]
    game, ply = g["extra"]["game"], g["extra"]["ply"]
    pick = [i for i in range(len(ply)) if "Lee-Sedol" in names[int(game[i])] and int(ply[i]) in (20, 41, 101, 161)]
    pick = sorted(set(pick) | set(range(0, len(ply), 7)))
    planes = gu.select_planes(g["planes"], feats)[pick].astype(np.float32)
    p2 = cnn_oracle2.forward(weights, planes, convs[:-1])
    p1 = cnn_oracle.forward(weights, planes, convs[:-1])
    d = float(np.abs(p1 - p2).max())
    assert d <= 1e-7, d
    out = os.path.join(gu.GOLDEN, "minimodel_forward.npz")
    np.savez_compressed(out, index=np.asarray(pick, np.int32), game=game[pick], ply=ply[pick], probs=p2,
                        feature_list=np.asarray(feats), oracle_max_abs_difference=np.float64(d))
    print("wrote", out, "positions", len(pick), "max |torch - scipy| =", d)


if __name__ == "__main__":
    main()
