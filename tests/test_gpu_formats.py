"""SURVEY 8f-2 on the device: SGF records -> states/actions through the batched replay and the
feature kernels, checked against the feature file the reference's authors stored for the AlphaGo -
Lee Sedol games, against the reference's replay of its seven test records, and against the 48-plane
golden vectors; and the reference's minimodel (Keras JSON + HDF5 weights) evaluated on the device."""
import os
import warnings

import numpy as np
import pytest
import torch

from oracle import cnn_oracle, go_oracle
from rocalphago_b200 import BatchedGameState, hdf5_lite, sgf
from rocalphago_b200.game_converter import GameConverter, IllegalMove, SizeMismatchError, replay_games, run_game_converter
from rocalphago_b200.models import CNNPolicy
from tests import golden_util as gu
from tests.test_formats import _lee_sedol, _replayed_games

pytestmark = pytest.mark.gpu
LETTERS = "abcdefghijklmnopqrstuvwxyz"


def _game_from_replay(g):
    ns = int(g["nsetup"])
    pts = [None if m < 0 else (int(m) // 19, int(m) % 19) for m in g["moves"]]
    cols = [int(c) for c in g["colors"]]
    return sgf.Game(19, [p for p, c in zip(pts[:ns], cols[:ns]) if c > 0], [p for p, c in zip(pts[:ns], cols[:ns]) if c < 0],
                    int(g["to_play"]), list(zip(cols[ns:], pts[ns:])))


def _sgf_text(game):
    pt = lambda p: "" if p is None else LETTERS[p[0]] + LETTERS[p[1]]  # noqa: E731
    s = "(;GM[1]FF[4]SZ[%d]" % game.size
    if game.add_black:
        s += "AB" + "".join("[%s]" % pt(p) for p in game.add_black)
    if game.add_white:
        s += "AW" + "".join("[%s]" % pt(p) for p in game.add_white)
    s += "PL[%s]" % ("B" if game.first_player > 0 else "W")
    for c, p in game.moves:
        s += ";%s[%s]" % ("B" if c > 0 else "W", pt(p))
    return s + ")"


def test_lockstep_replay_matches_the_references_replay():
    """All seven records (two with set-up stones, different lengths) replayed together."""
    ref = _replayed_games()
    names = sorted(ref)
    games = [_game_from_replay(ref[n]) for n in names]
    packed, gi, ply, moves, illegal_at = replay_games(games, 19)
    assert (illegal_at == -1).all() and len(gi) == sum(len(g.moves) for g in games)
    board, age, ko, player = BatchedGameState(packed, 19).to_arrays()
    for k, name in enumerate(names):
        g, rows = ref[name], np.flatnonzero(gi == k)
        ns = int(g["nsetup"])
        assert np.array_equal(ply[rows], np.arange(len(rows))) and np.array_equal(moves[rows], g["moves"][ns:])
        # position before move t (t >= 1) is the reference's state after move t-1
        assert np.array_equal(board[rows[1:]], g["board"][:-1]), name
        assert np.array_equal(age[rows[1:]], np.minimum(g["age"][:-1], 7)), name
        assert np.array_equal(ko[rows[1:]], g["ko"][:-1]) and np.array_equal(player[rows[1:]], g["player"][:-1]), name
        # position before the first move: the set-up stones, replayed by the oracle
        _, b0, a0, k0, _ = go_oracle.replay(19, g["moves"][:ns], g["colors"][:ns]) if ns else (0, np.zeros((1, 19, 19), np.int8), np.full((1, 19, 19), -1, np.int16), [-1], None)
        assert np.array_equal(board[rows[0]], b0[0]) and np.array_equal(age[rows[0]], np.minimum(a0[0], 7)) and ko[rows[0]] == k0[0]
        assert player[rows[0]] == int(g["to_play"])


def test_converter_reproduces_the_references_stored_feature_file(tmp_path):
    """sgfs_to_hdf5 on the five Lee Sedol records == tests/test_data/hdf5/alphago-vs-lee-sedol-features.hdf5."""
    states, actions, keys, offsets = _lee_sedol()
    ref = _replayed_games()
    files = []
    for key in keys:
        name = key.split(":")[-1]
        path = tmp_path / "sgf" / name
        path.parent.mkdir(exist_ok=True)
        path.write_text(_sgf_text(_game_from_replay(ref[name])))
        files.append(str(path))
    out = str(tmp_path / "features.hdf5")
    GameConverter(["board", "ones", "turns_since"], batch_games=2).sgfs_to_hdf5(files, out, bd_size=19)
    assert not os.path.exists(str(tmp_path / ".tmp.features.hdf5")) and not os.path.exists(str(tmp_path / ".tmp.features.hdf5.states"))
    f = hdf5_lite.File(out)
    assert f["states"].shape == states.shape and f["states"].dtype == np.uint8
    assert np.array_equal(f["states"][...], states)
    assert np.array_equal(f["actions"][...], actions)
    assert f["features"][()] == b"board,ones,turns_since"
    for path, (start, count) in zip(files, offsets):
        assert f["file_offsets"][path.replace("/", ":")][...].tolist() == [start, count]


def test_converter_planes_match_the_48_plane_golden_vectors():
    z = gu.load_features("features_sgf.npz")
    ref = _replayed_games()
    names = [str(n) for n in z["extra"]["names"]]
    games = [_game_from_replay(ref[n]) for n in names]
    packed, gi, ply, moves, _ = replay_games(games, 19)
    conv = GameConverter(go_oracle.DEFAULT_FEATURES)
    index = {(int(g), int(p)): i for i, (g, p) in enumerate(zip(gi, ply))}
    rows, want = [], []
    for i, (g, p) in enumerate(zip(z["extra"]["game"], z["extra"]["ply"])):
        if (int(g), int(p)) in index:                          # the position after the last move has no row
            rows.append(index[(int(g), int(p))])
            want.append(i)
    assert len(rows) > 200
    got = conv._planes(packed[torch.as_tensor(rows, device=packed.device)], 19)
    assert np.array_equal(got, z["planes"][want][:, :48])


def test_convert_game_semantics(tmp_path):
    conv = GameConverter(["board", "ones"])
    ok = tmp_path / "ok.sgf"
    ok.write_text("(;SZ[9];B[aa];W[];B[bb];W[tt];B[cc])")       # passes are not yielded (game_converter.py:37)
    pairs = list(conv.convert_game(str(ok), 9))
    assert [m for _, m in pairs] == [(0, 0), (1, 1), (2, 2)]
    assert pairs[0][0].shape == (1, 4, 9, 9) and pairs[0][0].dtype == np.uint8
    assert pairs[0][0][0, 2].sum() == 81                        # empty board before the first move
    assert pairs[1][0][0, 0, 0, 0] == 1 and pairs[1][0][0, 1].sum() == 0    # after W's pass it is B's turn again: own stone at (0, 0)
    with pytest.raises(SizeMismatchError):
        list(conv.convert_game(str(ok), 19))
    bad = tmp_path / "bad.sgf"
    bad.write_text("(;SZ[9];B[aa];W[bb];B[bb];W[cc])")          # third move on an occupied point
    got = []
    with pytest.raises(IllegalMove):
        for pair in conv.convert_game(str(bad), 9):
            got.append(pair[1])
    assert got == [(0, 0), (1, 1), (1, 1)]       # the pair at the illegal ply is yielded before do_move raises (util.py:124-126)
    # the file-level converter keeps the pairs up to the illegal move, warns, and skips broken files
    broken = tmp_path / "broken.sgf"
    broken.write_text("(;SZ[9];B[aa")
    small = tmp_path / "small.sgf"
    small.write_text("(;SZ[5];B[aa])")
    out = str(tmp_path / "o.h5")
    with warnings.catch_warnings(record=True) as caught:
        warnings.simplefilter("always")
        conv.sgfs_to_hdf5([str(ok), str(bad), str(broken), str(small)], out, bd_size=9)
    text = " | ".join(str(w.message) for w in caught)
    assert "Illegal Move encountered" in text and "Could not parse" in text and "wrong board size" in text
    f = hdf5_lite.File(out)
    assert f["states"].shape == (6, 4, 9, 9) and f["actions"][...].tolist() == [[0, 0], [1, 1], [2, 2], [0, 0], [1, 1], [1, 1]]
    assert f["file_offsets"][str(ok).replace("/", ":")][...].tolist() == [0, 3]
    assert f["file_offsets"][str(bad).replace("/", ":")][...].tolist() == [3, 3]
    # a coordinate past the board ("fz" on 9x9 would alias another point if flattened blindly) is an illegal move
    off = tmp_path / "off.sgf"
    off.write_text("(;SZ[9];B[aa];W[az];B[cc])")
    got = []
    with pytest.raises(IllegalMove):
        for pair in conv.convert_game(str(off), 9):
            got.append(pair[1])
    assert got[0] == (0, 0) and len(got) == 2
    assert len(f["file_offsets"]) == 2


def test_command_line_converter(tmp_path):
    """tests/test_game_converter.py:16-29 of the reference: a directory, and a recursive walk."""
    ref = _replayed_games()
    (tmp_path / "a" / "b").mkdir(parents=True)
    names = sorted(ref)
    for i, n in enumerate(names):
        (tmp_path / "a" / ("b" if i % 2 else "") / n).write_text(_sgf_text(_game_from_replay(ref[n])))
    out = str(tmp_path / "out.h5")
    run_game_converter(["--features", "board,ones,turns_since", "--outfile", out, "--directory", str(tmp_path / "a")])
    top = sum(len(ref[n]["moves"]) - int(ref[n]["nsetup"]) for i, n in enumerate(names) if i % 2 == 0)
    assert hdf5_lite.File(out)["states"].shape == (top, 12, 19, 19)
    run_game_converter(["--features", "board,ones,turns_since", "--outfile", out, "--directory", str(tmp_path), "--recurse"])
    total = sum(len(ref[n]["moves"]) - int(ref[n]["nsetup"]) for n in names)
    f = hdf5_lite.File(out)
    assert f["states"].shape == (total, 12, 19, 19) and len(f["file_offsets"]) == 7


def test_references_minimodel_on_the_device():
    """The reference's minimodel (Keras 1.1.2 JSON, HDF5 weights; fixture tests/golden/minimodel.npz)
    evaluated on the Lee Sedol positions: within north_star's tolerances of the CPU restatement, and
    of the committed outputs tests/golden/minimodel_forward.npz (oracle/make_golden_minimodel_forward.py:
    the scipy restatement, equal to the torch one to 5e-10) in every precision mode."""
    z = np.load(os.path.join(gu.GOLDEN, "minimodel.npz"))
    feats = [str(s) for s in z["feature_list"]]
    convs = [tuple(int(v) for v in c) for c in z["convs"]]
    weights = [z["w%02d" % i] for i in range(13)]
    kw = {"filter_width_%d" % (i + 1): w for i, (_, w) in enumerate(convs[:-1])}
    net = CNNPolicy(feats, board=19, layers=len(convs) - 1, filters_per_layer=convs[0][0], dtype="fp16", **kw)
    net.model.set_weights(weights)
    states, _, _, _ = _lee_sedol()
    x = states[::3].astype(np.float32)
    want = cnn_oracle.forward(weights, x, convs[:-1], head="policy")
    got = net.forward(x)
    assert got.shape == want.shape == (len(x), 361)
    assert np.abs(got - want).max() <= 2e-3
    assert (got.argmax(1) == want.argmax(1)).mean() >= 0.995
    gold = np.load(os.path.join(gu.GOLDEN, "minimodel_forward.npz"))
    g = gu.load_features("features_sgf.npz")
    idx = gold["index"]
    planes = gu.select_planes(g["planes"], feats)[idx].astype(np.float32)
    batch = BatchedGameState.from_arrays(g["board"][idx], g["age"][idx], g["ko"][idx], g["player"][idx])
    for dtype in ("fp16c32", "fp16", "bf16", "fp16x3", "fp16c64"):
        net = CNNPolicy(feats, board=19, layers=len(convs) - 1, filters_per_layer=convs[0][0], dtype=dtype, **kw)
        net.model.set_weights(weights)
        for got in (net.forward(planes), net.model.eval_batch_device(batch, net.preprocessor.plane_src).cpu().numpy()):
            assert np.abs(got - gold["probs"]).max() <= (2e-3 if dtype == "bf16" else 2e-5), dtype
            assert (got.argmax(1) == gold["probs"].argmax(1)).mean() >= (0.9 if dtype == "bf16" else 0.98), dtype
