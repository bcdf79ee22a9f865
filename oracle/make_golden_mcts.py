"""Visit counts of the reference's MCTS (AlphaGo/mcts.py) for tests/test_gpu_mcts.py.

    python oracle/make_golden_mcts.py      ->  tests/golden/mcts_reference.npz

mcts.py is Python 2 (dict.iteritems); its source is read from /root/reference, `iteritems` is replaced
by `items` IN MEMORY and the module is executed -- no file of the reference is copied or changed.  The
search runs on the reference's GameState with policy / value functions that are the CPU oracles
(reference feature planes + oracle/cnn_oracle.py forward) of fixed small networks; the test rebuilds the
same networks on the device (dtype fp16x3) and must arrive at the same visit counts with batch=1.
"""
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import cnn_oracle, ref_shim  # noqa: E402

FEATS = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size", "liberties_after",
         "ladder_capture", "ladder_escape", "sensibleness", "zeros"]
SIZE, LAYERS, FILTERS = 9, 3, 32
CASES = [dict(name="open", moves=[(2, 2), (6, 6), (2, 6), (6, 2), (4, 4)], n_playout=60, depth=4, c_puct=5),
         dict(name="fight", moves=[(4, 4), (4, 5), (5, 5), (5, 4), (3, 5), (4, 6), (3, 4), (5, 6), (6, 5)], n_playout=80, depth=6, c_puct=2)]


def main():
    go, prep = ref_shim.load()
    src = open(os.path.join(ref_shim.REFERENCE_ROOT, "AlphaGo", "mcts.py")).read().replace("iteritems", "items")
    mod = types.ModuleType("ref_mcts")
    exec(compile(src, "AlphaGo/mcts.py (iteritems -> items)", "exec"), mod.__dict__)
    pp = prep.Preprocess(FEATS)
    ppv = prep.Preprocess(FEATS)                 # the value net takes the 48 planes + colour (build-defined)
    specs = cnn_oracle.layer_specs(board=SIZE, input_dim=48, layers=LAYERS, filters_per_layer=FILTERS)
    pw = cnn_oracle.uniform_weights(specs, 48, SIZE, seed=101)
    pw[-3] = pw[-3] * 60.0                       # a policy with opinions, so that the search is not all ties
    vspecs = cnn_oracle.layer_specs(board=SIZE, input_dim=49, layers=LAYERS, filters_per_layer=FILTERS)
    vw = cnn_oracle.uniform_weights(vspecs, 49, SIZE, head="value", dense_units=64, seed=202)
    vw[-4] = vw[-4] * 40.0
    calls = [0]

    def policy_fn(state):
        calls[0] += 1
        moves = state.get_legal_moves()
        if not moves:
            return []
        x = pp.state_to_tensor(state).astype(np.float32)
        out = cnn_oracle.forward(pw, x, specs)[0]
        p = np.array([out[m[0] * SIZE + m[1]] for m in moves], np.float32)
        p = p / p.sum()
        return list(zip(moves, p))

    def value_fn(state):
        x = ppv.state_to_tensor(state).astype(np.float32)
        colour = np.full((1, 1, SIZE, SIZE), 1.0 if state.current_player == go.BLACK else 0.0, np.float32)
        return float(cnn_oracle.forward(vw, np.concatenate([x, colour], 1), vspecs, head="value")[0])

    out = {"size": np.int32(SIZE), "layers": np.int32(LAYERS), "filters": np.int32(FILTERS), "names": np.array([c["name"] for c in CASES])}
    for i, w in enumerate(pw):
        out["pw%02d" % i] = w
    for i, w in enumerate(vw):
        out["vw%02d" % i] = w
    for c in CASES:
        st = go.GameState(SIZE)
        for mv in c["moves"]:
            st.do_move(mv)
        search = mod.MCTS(value_fn, policy_fn, None, lmbda=0, c_puct=c["c_puct"], playout_depth=c["depth"], n_playout=c["n_playout"])
        values = []
        orig_value = search._value
        search._value = lambda s_, _f=orig_value, _v=values: (_v.append(_f(s_)), _v[-1])[1]
        trace = []
        orig_select = mod.TreeNode.select

        def logged_select(self, _o=orig_select, _t=trace):
            a, nd = _o(self)
            _t.append(a[0] * SIZE + a[1])
            return a, nd
        mod.TreeNode.select = logged_select
        search._value = lambda s_, _f=orig_value, _v=values, _t=trace: (_t.append(-1), _v.append(_f(s_)), _v[-1])[2]
        move = search.get_move(st)
        mod.TreeNode.select = orig_select
        out["%s_trace" % c["name"]] = np.array(trace, np.int32)
        out["%s_leaf_values" % c["name"]] = np.array(values, np.float64)
        out["%s_priors" % c["name"]] = np.array([k._P for k in search._root._children.values()], np.float64)
        kids = search._root._children
        acts = np.array([a[0] * SIZE + a[1] for a in kids], np.int32)
        out["%s_actions" % c["name"]] = acts
        out["%s_visits" % c["name"]] = np.array([k._n_visits for k in kids.values()], np.int32)
        out["%s_q" % c["name"]] = np.array([k._Q for k in kids.values()], np.float64)
        out["%s_move" % c["name"]] = np.int32(move[0] * SIZE + move[1])
        out["%s_setup" % c["name"]] = np.array([m[0] * SIZE + m[1] for m in c["moves"]], np.int32)
        out["%s_params" % c["name"]] = np.array([c["n_playout"], c["depth"], c["c_puct"]], np.int32)
        # the tree after re-rooting: second search from the position after the chosen move
        search.update_with_move(move)
        st.do_move(move)
        move2 = search.get_move(st)
        kids = search._root._children
        out["%s_actions2" % c["name"]] = np.array([a[0] * SIZE + a[1] for a in kids], np.int32)
        out["%s_visits2" % c["name"]] = np.array([k._n_visits for k in kids.values()], np.int32)
        out["%s_move2" % c["name"]] = np.int32(move2[0] * SIZE + move2[1])
        v = out["%s_visits" % c["name"]]
        print(c["name"], "move", move, "visits max", v.max(), "children visited", int((v > 0).sum()), "policy calls", calls[0])
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "mcts_reference.npz"), **out)


if __name__ == "__main__":
    main()
