"""Turns the reference's own HDF5 fixtures into .npz fixtures that travel with the repository.

    python oracle/make_golden_hdf5.py            (needs /root/reference; run once, outputs committed)

  tests/golden/hdf5_lee_sedol_features.npz   from tests/test_data/hdf5/alphago-vs-lee-sedol-features.hdf5:
        the `states` (1033, 12, 19, 19) / `actions` datasets and `file_offsets` the reference's
        GameConverter wrote for its five AlphaGo - Lee Sedol records with features
        board,ones,turns_since (used by tests/test_supervised_policy_trainer.py:9).  A golden
        vector produced by the reference's authors, not by this repository.
  tests/golden/minimodel.npz                  from tests/test_data/minimodel.json + hdf5/random_minimodel_weights.hdf5:
        the architecture fields and the Keras weight arrays in layer order.

The HDF5 files are read with rocalphago_b200/hdf5_lite.py (no HDF5 library on this image) -- which
makes these fixtures a test of that reader as well: tests/test_formats.py checks the arrays against
the oracle's features of the same positions.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rocalphago_b200 import hdf5_lite  # noqa: E402

REF = "/root/reference/tests/test_data"
OUT = os.path.join(ROOT, "tests", "golden")


def main():
    f = hdf5_lite.File(os.path.join(REF, "hdf5", "alphago-vs-lee-sedol-features.hdf5"))
    states, actions = f["states"][...], f["actions"][...]
    keys = f["file_offsets"].keys()
    offsets = np.array([f["file_offsets"][k][...] for k in keys], np.int64)
    assert states.max() <= 1
    np.savez_compressed(os.path.join(OUT, "hdf5_lee_sedol_features.npz"),
                        states_packed=np.packbits(states.reshape(-1)), states_shape=np.array(states.shape),
                        actions=actions, file_keys=np.array(keys), file_offsets=offsets,
                        feature_list=np.array(["board", "ones", "turns_since"]))
    print("states", states.shape, "actions", actions.shape, "games", len(keys))

    spec = json.load(open(os.path.join(REF, "minimodel.json")))
    w = hdf5_lite.File(os.path.join(REF, "hdf5", "random_minimodel_weights.hdf5"))
    arrays, names = [], []
    for layer in w.attrs["layer_names"]:
        g = w[layer.decode()]
        for wn in np.atleast_1d(g.attrs["weight_names"]):
            if isinstance(wn, bytes):
                arrays.append(g[wn.decode()][...])
                names.append(wn.decode())
    km = json.loads(spec["keras_model"])
    convs = [(l["config"]["nb_filter"], l["config"]["nb_row"]) for l in km["config"] if l["class_name"] == "Convolution2D"]
    np.savez_compressed(os.path.join(OUT, "minimodel.npz"), feature_list=np.array(spec["feature_list"]),
                        net_class=np.array(spec["class"]), convs=np.array(convs),
                        input_shape=np.array(km["config"][0]["config"]["batch_input_shape"][1:]),
                        weight_names=np.array(names), **{"w%02d" % i: a for i, a in enumerate(arrays)})
    print("minimodel", spec["class"], spec["feature_list"], convs, [a.shape for a in arrays])


if __name__ == "__main__":
    main()
