"""Loaders for tests/golden/*.npz (outputs of the reference; see oracle/make_golden.py)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FEATURE_FILES = ["features_fixture_boards.npz", "features_sgf.npz", "features_playouts_19.npz",
                 "features_playouts_13.npz", "features_playouts_9.npz", "features_playouts_7.npz",
                 "features_playouts_5.npz"]
PLANE_OFFSETS = {"board": 0, "ones": 3, "turns_since": 4, "liberties": 12, "capture_size": 20,
                 "self_atari_size": 28, "liberties_after": 36, "ladder_capture": 44,
                 "ladder_escape": 45, "sensibleness": 46, "zeros": 47, "legal": 48}
PLANE_SIZES = {"board": 3, "ones": 1, "turns_since": 8, "liberties": 8, "capture_size": 8,
               "self_atari_size": 8, "liberties_after": 8, "ladder_capture": 1, "ladder_escape": 1,
               "sensibleness": 1, "zeros": 1, "legal": 1}


def load_features(name):
    z = np.load(os.path.join(GOLDEN, name))
    shape = tuple(int(v) for v in z["planes_shape"])
    planes = np.unpackbits(z["planes_packed"])[:int(np.prod(shape))].reshape(shape)
    return dict(size=int(z["size"]), board=z["board"], age=z["age"], ko=z["ko"], player=z["player"],
                planes=planes, feature_list=[str(s) for s in z["feature_list"]],
                legal_moves=z["legal_moves"], n_legal=z["n_legal"], n_noneye=z["n_noneye"],
                extra={k: z[k] for k in z.files if k in ("game", "ply", "names")})


def select_planes(planes, feature_list):
    """Re-orders the golden 49 planes (48 default + legal) into an arbitrary feature list."""
    idx = []
    for f in feature_list:
        f = f.lower()
        idx += list(range(PLANE_OFFSETS[f], PLANE_OFFSETS[f] + PLANE_SIZES[f]))
    return planes[:, idx]


def load_kat():
    z = np.load(os.path.join(GOLDEN, "kat_reference_tests.npz"))
    return {k: z[k] for k in z.files}


def load_replay():
    z = np.load(os.path.join(GOLDEN, "replay_sgf.npz"))
    return {k: z[k] for k in z.files}


def load_superko():
    """tests/golden/features_superko.npz (oracle/make_golden_superko.py): per set of positions recorded
    from reference games played with enforce_superko=True -> dict of arrays + the reference's planes."""
    z = np.load(os.path.join(GOLDEN, "features_superko.npz"))
    feats = [str(f) for f in z["feature_list"]]
    out = {}
    for key in [str(k) for k in z["sets"]]:
        g = {name: z["%s_%s" % (name, key)] for name in ("board", "age", "ko", "player", "hash", "prev", "prev_off", "masks",
                                                        "hist_len", "has_handicap", "moves")}
        shape = tuple(int(v) for v in z["planes_shape_%s" % key])
        g["planes"] = np.unpackbits(z["planes_%s" % key])[:int(np.prod(shape))].reshape(shape)
        g["sk"] = dict(enforce=np.ones(shape[0], np.int32), has_handicap=g["has_handicap"], hist_len=g["hist_len"],
                       hash=g["hash"], prev=g["prev"], prev_off=g["prev_off"], masks=g["masks"])
        out[key] = g
    return feats, out
