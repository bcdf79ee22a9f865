"""Golden vectors for the `maximum` argument of the reference's feature functions
(preprocessing.py:25-169): get_turns_since / get_liberties / get_capture_size /
get_self_atari_size / get_liberties_after called directly with maximum in {1, 3, 5, 8} on every
eighth position of tests/golden/features_sgf.npz, by the unmodified reference (oracle/ref_shim.py).

    python oracle/make_golden_maximum.py      ->  tests/golden/feature_functions_maximum.npz
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

NAMES = ["turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after"]
MAXIMA = [1, 3, 5, 8]


def main():
    go, prep = ref_shim.load()
    z = np.load(os.path.join(ROOT, "tests", "golden", "features_sgf.npz"))
    rows = list(range(0, len(z["board"]), 8))
    out = {"rows": np.asarray(rows, np.int32), "maxima": np.asarray(MAXIMA, np.int32), "names": np.asarray(NAMES)}
    for name in NAMES:
        fn = getattr(prep, "get_" + name)
        for m in MAXIMA:
            planes = []
            for i in rows:
                st = ref_shim.state_from_arrays(go, z["board"][i], z["age"][i], int(z["ko"][i]), int(z["player"][i]))
                planes.append(fn(st, maximum=m).astype(np.uint8))
            out["%s_%d" % (name, m)] = np.packbits(np.stack(planes).reshape(-1))
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "feature_functions_maximum.npz"), **out)
    print("positions", len(rows), "arrays", len(out) - 3)


if __name__ == "__main__":
    main()
