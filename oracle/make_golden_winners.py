"""GameState.get_winner() (go.py:482-506) of the golden random-playout positions, by the unmodified
reference:  python oracle/make_golden_winners.py  ->  tests/golden/winners.npz"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402


def main():
    go, _ = ref_shim.load()
    out = {}
    for size in (19, 9, 5):
        z = np.load(os.path.join(ROOT, "tests", "golden", "features_playouts_%d.npz" % size))
        rows = list(range(0, len(z["board"]), 3))
        for komi in (7.5, 0.0):
            w = []
            for i in rows:
                st = ref_shim.state_from_arrays(go, z["board"][i], z["age"][i], int(z["ko"][i]), int(z["player"][i]))
                st.komi = komi
                w.append(st.get_winner())
            out["winner_%d_%g" % (size, komi)] = np.asarray(w, np.int32)
        out["rows_%d" % size] = np.asarray(rows, np.int32)
        print(size, len(rows), np.bincount(np.asarray(w) + 1, minlength=3))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "winners.npz"), **out)


if __name__ == "__main__":
    main()
