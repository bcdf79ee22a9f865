"""Golden vectors for states that enforce positional superko (go.py:242-265, 282-283), made by the
unmodified reference (oracle/ref_shim.py).

    python oracle/make_golden_superko.py     ->  tests/golden/features_superko.npz

Games on 2x2 ... 9x9 boards (tiny boards because positions recur there all the time; some with
handicap stones) are played with GameState(size,
enforce_superko=True), every move drawn uniformly from get_legal_moves() -- own eyes included, so that
groups keep dying and positions recur.  Every `stride`-th position is recorded with what
is_positional_superko reads (current_hash, previous_hashes, the points of history[0::2], history[1::2]
and handicaps) and with the reference's planes (48 default + legal) and legal-move list.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

FEATURES = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after",
            "ladder_capture", "ladder_escape", "sensibleness", "zeros", "legal"]


def record(st):
    S = st.size
    masks = np.zeros((3, S * S), np.uint8)
    for i, mv in enumerate(st.history):
        if mv is not None:
            masks[i & 1, mv[0] * S + mv[1]] = 1
    for mv in st.handicaps:
        masks[2, mv[0] * S + mv[1]] = 1
    board, age, ko, player = ref_shim.state_arrays(st)
    return dict(board=board, age=age, ko=ko, player=player, hash=np.uint64(st.current_hash),
                prev=np.array(sorted(int(h) for h in st.previous_hashes), np.uint64), masks=masks,
                hist_len=len(st.history), has_handicap=int(len(st.handicaps) > 0))


def main():
    go, prep = ref_shim.load()
    pp = prep.Preprocess(FEATURES)
    rng = np.random.RandomState(20261107)
    plans = [(2, 60, 40, 2, 0), (3, 60, 120, 3, 0), (4, 40, 160, 3, 0), (5, 20, 120, 5, 0), (7, 8, 220, 7, 2), (9, 4, 300, 11, 3)]
    out = {}
    kat = [(0, 3), (0, 4), (1, 3), (1, 4), (2, 3), (2, 4), (2, 2), (3, 4), (2, 1), (3, 3), (3, 1), (3, 2), (3, 0),
           (4, 2), (1, 1), (4, 1), (8, 0), (4, 0), (8, 1), (0, 2), (8, 2), (0, 1), (8, 3), (1, 0), (8, 4), (2, 0), (0, 0)]
    plans.append((9, 1, len(kat) + 1, 1, "kat"))     # the reference's own scenario, tests/test_gamestate.py:60-77, every ply
    for size, games, plies, stride, handi in plans:
        recs, planes, moves_all, blocked = [], [], [], 0
        for gi in range(games):
            st = go.GameState(size, enforce_superko=True)
            if handi and handi != "kat":
                pts = [(1, 1), (size - 2, size - 2), (1, size - 2)][:handi]
                st.place_handicaps(pts)
                st.current_player = go.WHITE
            for t in range(plies):
                legal = st.get_legal_moves()
                if not legal:                          # nothing but a pass is left: pass, stop at the second one
                    if len(st.history) > 0 and st.history[-1] is None:
                        break
                    st.do_move(None)
                    continue
                if t % stride == gi % stride:
                    r = record(st)
                    planes.append(pp.state_to_tensor(st)[0].astype(np.uint8))
                    moves_all.append([m[0] * size + m[1] for m in legal])
                    st.enforce_superko = False
                    st._GameState__legal_move_cache = None
                    blocked += len(st.get_legal_moves()) - len(legal)
                    st.enforce_superko = True
                    st._GameState__legal_move_cache = None
                    recs.append(r)
                if handi == "kat":
                    if t == len(kat):
                        assert not st.is_legal((1, 0))
                        break
                    st.do_move(kat[t])
                else:
                    st.do_move(legal[rng.randint(len(legal))])
        n = len(recs)
        prev_off = np.zeros(n + 1, np.int32)
        prev_off[1:] = np.cumsum([len(r["prev"]) for r in recs])
        mv = np.full((n, size * size), -1, np.int16)
        for i, m in enumerate(moves_all):
            mv[i, :len(m)] = m
        P = np.stack(planes)
        out.update({
            "board_%d_%s" % (size, handi): np.stack([r["board"] for r in recs]),
            "age_%d_%s" % (size, handi): np.stack([r["age"] for r in recs]),
            "ko_%d_%s" % (size, handi): np.array([r["ko"] for r in recs], np.int32),
            "player_%d_%s" % (size, handi): np.array([r["player"] for r in recs], np.int32),
            "hash_%d_%s" % (size, handi): np.array([r["hash"] for r in recs], np.uint64),
            "prev_%d_%s" % (size, handi): np.concatenate([r["prev"] for r in recs]) if n else np.zeros(0, np.uint64),
            "prev_off_%d_%s" % (size, handi): prev_off,
            "masks_%d_%s" % (size, handi): np.stack([r["masks"] for r in recs]),
            "hist_len_%d_%s" % (size, handi): np.array([r["hist_len"] for r in recs], np.int32),
            "has_handicap_%d_%s" % (size, handi): np.array([r["has_handicap"] for r in recs], np.int32),
            "planes_shape_%d_%s" % (size, handi): np.array(P.shape, np.int32),
            "planes_%d_%s" % (size, handi): np.packbits(P.reshape(-1)),
            "moves_%d_%s" % (size, handi): mv,
        })
        print("size %d handicap %s: %d positions, %d moves forbidden by superko alone, ladder bits %d/%d" % (
            size, handi, n, blocked, int(P[:, 44].sum()), int(P[:, 45].sum())))
    out["sets"] = np.array(["%d_%s" % (s, h) for s, _, _, _, h in plans])
    out["feature_list"] = np.array(FEATURES)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "features_superko.npz"), **out)


if __name__ == "__main__":
    main()
