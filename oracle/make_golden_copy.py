"""Golden vectors for positions as MCTS sees them: GameState.copy() does not copy stone_ages
(go.py:186-217), so on a copy every stone placed before the copy has age -1 -- it is in no
turns_since plane and do_move never ages it (go.py:533) -- while stones played on the copy age normally.

    python oracle/make_golden_copy.py    ->  tests/golden/features_copied_states.npz

Every 5th position of tests/golden/features_sgf.npz is rebuilt in the unmodified reference, copied, and
0..4 random legal moves are played on the copy; recorded: the copy's fields and the reference's planes.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

FEATURES = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after",
            "ladder_capture", "ladder_escape", "sensibleness", "zeros", "legal"]


def main():
    go, prep = ref_shim.load()
    pp = prep.Preprocess(FEATURES)
    z = np.load(os.path.join(ROOT, "tests", "golden", "features_sgf.npz"))
    rng = np.random.RandomState(5)
    recs, planes, played = [], [], []
    for k, i in enumerate(range(0, len(z["board"]), 5)):
        st = ref_shim.state_from_arrays(go, z["board"][i], z["age"][i], int(z["ko"][i]), int(z["player"][i]))
        c = st.copy()
        mv = []
        for _ in range(k % 5):
            legal = c.get_legal_moves(include_eyes=False)
            if not legal:
                break
            m = legal[rng.randint(len(legal))]
            c.do_move(m)
            mv.append(m[0] * 19 + m[1])
        recs.append(ref_shim.state_arrays(c))
        planes.append(pp.state_to_tensor(c)[0].astype(np.uint8))
        played.append(mv + [-1] * (4 - len(mv)))
    P = np.stack(planes)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "features_copied_states.npz"),
                        board=np.stack([r[0] for r in recs]), age=np.stack([r[1] for r in recs]),
                        ko=np.array([r[2] for r in recs], np.int32), player=np.array([r[3] for r in recs], np.int32),
                        planes_shape=np.array(P.shape, np.int32), planes=np.packbits(P.reshape(-1)),
                        source_rows=np.arange(0, len(z["board"]), 5, dtype=np.int32), played=np.array(played, np.int32),
                        feature_list=np.array(FEATURES))
    ages = np.stack([r[1] for r in recs]); boards = np.stack([r[0] for r in recs])
    print(len(recs), "positions; stones without an age:", int(((ages < 0) & (boards != 0)).sum()), "with:", int(((ages >= 0) & (boards != 0)).sum()),
          "turns_since bits:", int(P[:, 4:12].sum()))


if __name__ == "__main__":
    main()
