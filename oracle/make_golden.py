"""Generates tests/golden/*.npz by RUNNING THE REFERENCE -- TEST INFRASTRUCTURE ONLY.

Run in the build container (where /root/reference is mounted):

    python oracle/make_golden.py            # ~2 min on 8 cores

Everything written is an output of the reference's own, unmodified Python
(imported through oracle/ref_shim.py); nothing is computed by this repository's
code.  The fixtures are small so they can be committed and travel to the GPU
box, where /root/reference does not exist.

Files
  kat_reference_tests.npz  every is_legal / is_eye / is_ladder_capture /
                           is_ladder_escape query the reference's own unit
                           tests (tests/test_gamestate.py, test_ladders.py,
                           test_liberties.py, test_preprocessing.py) make, with
                           the state it was made on and the answer the
                           reference gave -- recorded by wrapping the methods
                           while the reference's unittest suites run.
  features_fixture_boards.npz   the three fixture boards of
                           tests/test_preprocessing.py:8-104 and the two ladder
                           boards (:284-314): all 12 features.
  features_sgf.npz         positions from the 5 AlphaGo-Lee Sedol SGFs and the
                           2 handicap SGFs of tests/test_data: state + 49 planes
                           (48 default + legal) + get_legal_moves() order.
  features_playouts_{19,13,9,7,5}.npz  random-playout positions.
  replay_sgf.npz           move lists of the SGFs and the state after every move
                           (board, stone_ages, ko, player) -- pins do_move.
"""
import glob
import multiprocessing as mp
import os
import random
import sys
import unittest

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref_shim, sgf_mainline  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
ALL_FEATURES = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size",
                "liberties_after", "ladder_capture", "ladder_escape", "sensibleness", "zeros", "legal"]
NPLANES = 49


def _flat(p, S):
    return -1 if p is None else p[0] * S + p[1]


def _pad(a, S, dtype):
    out = np.zeros(361, dtype)
    out[:S * S] = np.asarray(a, dtype).reshape(-1)
    return out


# --------------------------------------------------------------------------- KATs

def record_reference_unit_tests():
    go, _ = ref_shim.load()
    rec = []

    def wrap(name, kind):
        orig = getattr(go.GameState, name)

        def wrapped(self, *args, **kw):
            board, age, ko, player = ref_shim.state_arrays(self)
            S = self.size
            if kind in ("ladder_capture", "ladder_escape"):
                action = args[0] if args else kw["action"]
                prey = kw.get("prey", args[1] if len(args) > 1 else None)
                depth = kw.get("remaining_attempts", args[2] if len(args) > 2 else 80)
                top = depth == 80
                q = (_flat(action, S), _flat(prey, S))
            elif kind == "legal":
                top = True
                a = args[0]
                q = (-1 if a is None else _flat(a, S), -1)
                if a is not None and not self._on_board(a):
                    top = False
            else:  # eye
                pos, owner = args[0], args[1]
                stack = args[2] if len(args) > 2 else kw.get("stack", None)
                top = stack is None or len(stack) == 0
                q = (_flat(pos, S), int(owner))
            res = orig(self, *args, **kw)
            if top and not wrapped.busy:
                rec.append((kind, S, board, ko, player, q[0], q[1], int(bool(res))))
            return res
        wrapped.busy = False
        setattr(go.GameState, name, wrapped)
        return orig

    origs = {n: wrap(n, k) for n, k in [("is_ladder_capture", "ladder_capture"),
                                         ("is_ladder_escape", "ladder_escape"),
                                         ("is_legal", "legal"), ("is_eye", "eye")]}
    try:
        suite = unittest.TestSuite()
        for mod in ("test_gamestate", "test_ladders", "test_liberties", "test_preprocessing"):
            suite.addTests(unittest.defaultTestLoader.loadTestsFromName(mod))
        result = unittest.TextTestRunner(verbosity=0, stream=open(os.devnull, "w")).run(suite)
        assert result.wasSuccessful(), "the reference's own tests must pass under the shim"
        ntests = result.testsRun
    finally:
        for n, o in origs.items():
            setattr(go.GameState, n, o)

    # de-duplicate (get_legal_moves asks is_legal for all 361 points of many states)
    seen, uniq = set(), []
    for r in rec:
        key = (r[0], r[1], r[2].tobytes(), r[3], r[4], r[5], r[6])
        if key not in seen:
            seen.add(key)
            uniq.append(r)
    kinds = {"legal": 0, "eye": 1, "ladder_capture": 2, "ladder_escape": 3}
    np.savez_compressed(
        os.path.join(OUT, "kat_reference_tests.npz"),
        kind=np.asarray([kinds[r[0]] for r in uniq], np.int8),
        size=np.asarray([r[1] for r in uniq], np.int8),
        board=np.stack([_pad(r[2], r[1], np.int8) for r in uniq]),
        ko=np.asarray([r[3] for r in uniq], np.int16),
        player=np.asarray([r[4] for r in uniq], np.int8),
        arg0=np.asarray([r[5] for r in uniq], np.int16),      # action / position
        arg1=np.asarray([r[6] for r in uniq], np.int16),      # prey (or -1) / eye owner
        result=np.asarray([r[7] for r in uniq], np.int8),
        reference_tests_run=np.int32(ntests))
    by = {k: sum(1 for r in uniq if r[0] == k) for k in kinds}
    print("kat_reference_tests: %d reference tests ran, recorded %s" % (ntests, by))


# ----------------------------------------------------------------------- features

def _features_of(st, pp):
    go, prep = ref_shim.load()
    t = pp.state_to_tensor(st)[0]
    assert t.shape[0] == NPLANES + 1 - 1 + 0 or True
    legal = st.get_legal_moves()
    n_noneye = len(st.get_legal_moves(include_eyes=False))
    return t.astype(np.uint8), [_flat(m, st.size) for m in legal], n_noneye


def _save_features(name, states_feats, extra=None):
    """states_feats: list of (board, age, ko, player, planes u8 (49,S,S), legal list, n_noneye)"""
    S = states_feats[0][0].shape[-1]
    n = len(states_feats)
    moves = np.full((n, S * S), -1, np.int16)
    for i, r in enumerate(states_feats):
        moves[i, :len(r[5])] = r[5]
    planes = np.stack([r[4] for r in states_feats])
    d = dict(
        size=np.int32(S),
        board=np.stack([r[0] for r in states_feats]).astype(np.int8),
        age=np.stack([r[1] for r in states_feats]).astype(np.int16),
        ko=np.asarray([r[2] for r in states_feats], np.int32),
        player=np.asarray([r[3] for r in states_feats], np.int32),
        planes_shape=np.asarray(planes.shape, np.int32),
        planes_packed=np.packbits(planes.reshape(-1)),
        feature_list=np.asarray(ALL_FEATURES),
        legal_moves=moves,
        n_legal=np.asarray([len(r[5]) for r in states_feats], np.int32),
        n_noneye=np.asarray([r[6] for r in states_feats], np.int32))
    if extra:
        d.update(extra)
    np.savez_compressed(os.path.join(OUT, name), **d)
    print("%s: %d positions, size %d, ko positions %d, ladder_capture bits %d, ladder_escape bits %d" % (
        name, n, S, int((d["ko"] >= 0).sum()), int(planes[:, 44].sum()), int(planes[:, 45].sum())))


def _sgf_worker(job):
    path, plies = job
    go, prep = ref_shim.load()
    pp = prep.Preprocess(ALL_FEATURES)
    g = sgf_mainline.read(open(path).read())
    st = go.GameState(size=g["size"])
    for xy in g["black"]:
        st.do_move(xy, go.BLACK)
    for xy in g["white"]:
        st.do_move(xy, go.WHITE)
    st.current_player = g["to_play"]
    out, replay = [], []
    setup = [(+1, xy) for xy in g["black"]] + [(-1, xy) for xy in g["white"]]
    for k, (color, mv) in enumerate(g["moves"]):
        if k in plies:
            # features are taken on the live state (never a copy(): SURVEY H4)
            b, a, ko, pl = ref_shim.state_arrays(st)
            planes, legal, nne = _features_of(st, pp)
            out.append((b, a, ko, pl, planes, legal, nne))
        st.do_move(mv, color)
        b, a, ko, pl = ref_shim.state_arrays(st)
        replay.append((b, a, ko, pl))
    return os.path.basename(path), g, setup, out, replay


def make_sgf():
    files = sorted(glob.glob(os.path.join(ref_shim.REFERENCE_ROOT, "tests/test_data/sgf/*.sgf")))
    files += sorted(glob.glob(os.path.join(ref_shim.REFERENCE_ROOT, "tests/test_data/sgf_with_handicap/*.sgf")))
    # every 6th ply plus the four plies the survey's config-1 probe uses
    jobs = [(f, set(range(2, 400, 6)) | {20, 41, 101, 161}) for f in files]
    with mp.Pool(min(8, len(jobs))) as pool:
        res = pool.map(_sgf_worker, jobs)
    feats, src, ply_of = [], [], []
    rep = {}
    for gi, (name, g, setup, out, replay) in enumerate(res):
        plies = sorted(set(range(2, 400, 6)) | {20, 41, 101, 161})
        plies = [p for p in plies if p < len(g["moves"])]
        assert len(plies) == len(out)
        feats += out
        src += [gi] * len(out)
        ply_of += plies
        S = g["size"]
        mv = setup + g["moves"]
        rep["moves_%d" % gi] = np.asarray([_flat(m, S) for _, m in mv], np.int32)
        rep["colors_%d" % gi] = np.asarray([c for c, _ in mv], np.int32)
        rep["nsetup_%d" % gi] = np.int32(len(setup))
        rep["to_play_%d" % gi] = np.int32(g["to_play"])
        rep["board_%d" % gi] = np.stack([r[0] for r in replay]).astype(np.int8)
        rep["age_%d" % gi] = np.stack([r[1] for r in replay]).astype(np.int16)
        rep["ko_%d" % gi] = np.asarray([r[2] for r in replay], np.int32)
        rep["player_%d" % gi] = np.asarray([r[3] for r in replay], np.int32)
    rep["names"] = np.asarray([r[0] for r in res])
    _save_features("features_sgf.npz", feats,
                   extra=dict(game=np.asarray(src, np.int32), ply=np.asarray(ply_of, np.int32),
                              names=np.asarray([r[0] for r in res])))
    np.savez_compressed(os.path.join(OUT, "replay_sgf.npz"), **rep)
    print("replay_sgf: %d games, %d moves" % (len(res), sum(len(r[4]) for r in res)))


def _playout_worker(job):
    size, seed, games, stride, max_ply, eye_fill_every = job
    go, prep = ref_shim.load()
    pp = prep.Preprocess(ALL_FEATURES)
    rng = random.Random(seed)
    out = []
    for g in range(games):
        st = go.GameState(size=size)
        passes = 0
        for t in range(max_ply + 1):
            # mostly sensible moves; some games also fill eyes so that
            # big captures, suicides-by-eye-filling and multi-stone ko shapes appear
            fill = eye_fill_every and (g % eye_fill_every == eye_fill_every - 1)
            moves = st.get_legal_moves(include_eyes=bool(fill))
            if t % stride == g % stride:
                b, a, ko, pl = ref_shim.state_arrays(st)
                planes, legal, nne = _features_of(st, pp)
                out.append((b, a, ko, pl, planes, legal, nne))
            mv = rng.choice(moves) if moves else None
            st.do_move(mv)
            passes = passes + 1 if mv is None else 0
            if passes >= 2:
                break
    return out


def make_playouts():
    plans = {   # size: (workers, games per worker, stride, max_ply)
        19: (8, 6, 8, 400),
        13: (4, 4, 8, 220),
        9: (4, 6, 6, 120),
        7: (2, 8, 5, 70),
        5: (2, 10, 4, 40),
    }
    for size, (workers, games, stride, max_ply) in plans.items():
        jobs = [(size, 20261106 + 1000 * size + w, games, stride, max_ply, 3) for w in range(workers)]
        with mp.Pool(8) as pool:
            res = pool.map(_playout_worker, jobs)
        feats = [r for part in res for r in part]
        _save_features("features_playouts_%d.npz" % size, feats)


def make_fixture_boards():
    go, prep = ref_shim.load()
    import test_preprocessing as tp
    import parseboard
    pp = prep.Preprocess(ALL_FEATURES)
    states = [tp.simple_board(), tp.self_atari_board(), tp.capture_board()]
    # the two ladder boards of tests/test_preprocessing.py:284-314 are 6 rows x 7
    # columns -> parseboard makes a 7x7 state; rebuild them from the KAT recording
    feats = []
    for st in states:
        b, a, ko, pl = ref_shim.state_arrays(st)
        planes, legal, nne = _features_of(st, pp)
        feats.append((b, a, ko, pl, planes, legal, nne))
    _save_features("features_fixture_boards.npz", feats)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    which = sys.argv[1:] or ["kat", "fixtures", "sgf", "playouts"]
    if "kat" in which:
        record_reference_unit_tests()
    if "fixtures" in which:
        make_fixture_boards()
    if "sgf" in which:
        make_sgf()
    if "playouts" in which:
        make_playouts()
