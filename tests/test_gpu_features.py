"""Parity of the CUDA feature path (through the C ABI) with the reference: bit-exact planes
against the golden fixtures recorded from the reference, and against the C oracle on fresh
random playouts at sizes the fixtures cannot hold.  Run on the B200 box: pytest -m gpu."""
import os
import numpy as np
import pytest
import torch

from oracle import go_oracle
from rocalphago_b200 import BatchedGameState, Preprocess, PositionView, _cabi
from rocalphago_b200 import preprocessing as P
from tests import golden_util as gu

pytestmark = pytest.mark.gpu

ALL = ["board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size",
       "liberties_after", "ladder_capture", "ladder_escape", "sensibleness", "zeros", "legal"]


def _explain(got, want):
    bad = np.argwhere(got != want)
    planes = sorted(set(bad[:, 1].tolist()))
    return "mismatches %d in planes %s, first (pos, plane, x, y): %s" % (len(bad), planes, bad[:6].tolist())


def _gpu_planes(board, age, ko, player, feature_list=ALL, layout=P.LAYOUT_NCHW_U8):
    batch = BatchedGameState.from_arrays(board, age, ko, player)
    return Preprocess(feature_list).to_device(batch, layout).cpu().numpy()


@pytest.mark.parametrize("name", gu.FEATURE_FILES)
def test_planes_bit_exact_vs_reference_golden(name):
    g = gu.load_features(name)
    got = _gpu_planes(g["board"], g["age"], g["ko"], g["player"])
    assert got.dtype == np.uint8 and got.shape == g["planes"].shape
    assert np.array_equal(got, g["planes"]), _explain(got, g["planes"])


@pytest.mark.parametrize("size,n,seed", [(19, 12000, 101), (13, 2000, 102), (9, 2000, 103), (7, 1000, 104),
                                         (6, 500, 105), (5, 500, 106), (2, 20, 107), (17, 600, 108), (15, 601, 109),
                                         (16, 333, 110)])
def test_planes_bit_exact_vs_oracle_on_playouts(size, n, seed):
    b, a, ko, pl = go_oracle.random_playouts(n, size=size, seed=seed, max_ply=size * size + 40, stride=5)
    want = go_oracle.features(b, a, ko, pl, ALL)
    got = _gpu_planes(b, a, ko, pl)
    assert np.array_equal(got, want), _explain(got, want)


@pytest.mark.parametrize("size", [19, 16])
def test_flat_u8_emit_with_plane_lists_that_break_16_byte_alignment(size):
    """The uint8 emit for boards of 15+ lines treats the batch's output as one flat byte stream cut into
    16-byte chunks: with F*S*S not a multiple of 16 the chunks straddle rows, planes AND positions, and
    the stream ends in a partial chunk."""
    n = 37
    b, a, ko, pl = go_oracle.random_playouts(n, size=size, seed=211, max_ply=size * size, stride=9)
    for fl in (["ones"], ["board"], ["ladder_capture", "board", "zeros", "sensibleness", "ones"],
               ["liberties", "ones", "turns_since", "capture_size", "legal", "board", "board"]):
        want = go_oracle.features(b, a, ko, pl, fl)
        for k in (n, 1, 2, 5):
            got = _gpu_planes(b[:k], a[:k], ko[:k], pl[:k], fl)
            assert np.array_equal(got, want[:k]), (fl, k)


def test_feature_subsets_orders_repeats_and_float_layout():
    g = gu.load_features("features_playouts_9.npz")
    for fl in (["board", "sensibleness", "capture_size"], ["legal"], ["zeros", "ones", "board", "board"],
               ["LADDER_ESCAPE", "Liberties"], ["ladder_capture"]):
        got = _gpu_planes(g["board"], g["age"], g["ko"], g["player"], fl)
        assert np.array_equal(got, gu.select_planes(g["planes"], fl)), fl
    got = _gpu_planes(g["board"], g["age"], g["ko"], g["player"], P.DEFAULT_FEATURES, P.LAYOUT_NCHW_F32)
    assert got.dtype == np.float32
    assert np.array_equal(got, gu.select_planes(g["planes"], P.DEFAULT_FEATURES).astype(np.float32))


@pytest.mark.parametrize("dtype", [P.DTYPE_BF16, P.DTYPE_FP16])
@pytest.mark.parametrize("size,channels", [(19, 64), (19, 128), (9, 64)])
def test_conv_input_layout(size, channels, dtype):
    """Dense pixel rows (NHWC): row b*S*S + x*S + y carries the planes of point (x, y) as 1.0/0.0,
    channels past the listed planes are zero."""
    g = gu.load_features("features_playouts_%d.npz" % size)
    n = 37
    batch = BatchedGameState.from_arrays(g["board"][:n], g["age"][:n], g["ko"][:n], g["player"][:n])
    pp = Preprocess(P.DEFAULT_FEATURES)
    out = pp.to_device(batch, P.LAYOUT_CONV_IN, conv_dtype=dtype, conv_channels=channels)
    rows = out.float().cpu().numpy()
    planes = gu.select_planes(g["planes"][:n], P.DEFAULT_FEATURES)            # (n,48,S,S)
    want = np.zeros((n * size * size, channels), np.float32)
    want[:, :48] = planes.transpose(0, 2, 3, 1).reshape(n * size * size, 48)
    assert rows.shape[0] % 256 == 0 and rows.shape[0] >= n * size * size and rows.shape[1] == channels
    assert np.array_equal(rows[:n * size * size], want)


def test_state_to_tensor_keeps_reference_signature():
    g = gu.load_features("features_fixture_boards.npz")
    pp = Preprocess()
    for i in range(len(g["ko"])):
        st = PositionView(g["board"][i], g["age"][i], int(g["ko"][i]), int(g["player"][i]))
        t = pp.state_to_tensor(st)
        assert t.shape == (1, 48, 7, 7) and t.dtype == np.float64
        assert np.array_equal(t[0], gu.select_planes(g["planes"][i:i + 1], P.DEFAULT_FEATURES)[0])
    assert Preprocess(["board"]).states_to_tensor([]).shape[0] == 0
    a = PositionView(np.zeros((9, 9)), -np.ones((9, 9)), None, 1)
    b = PositionView(np.zeros((7, 7)), -np.ones((7, 7)), None, 1)
    with pytest.raises(ValueError, match="same size"):
        pp.states_to_tensor([a, b])


def test_legal_move_lists_follow_reference_order():
    g = gu.load_features("features_sgf.npz")
    batch = BatchedGameState.from_arrays(g["board"], g["age"], g["ko"], g["player"])
    pp = Preprocess()
    ws = pp.analyze(batch)
    lists = pp.legal_move_lists(ws, len(batch), 19)
    for i, lst in enumerate(lists):
        want = g["legal_moves"][i, :g["n_legal"][i]]
        assert [x * 19 + y for (x, y) in lst] == want.tolist()


def test_do_move_matches_oracle():
    for size, seed in ((19, 201), (9, 202), (5, 203)):
        n = 3000
        b, a, ko, pl, mv = go_oracle.random_playouts(n, size=size, seed=seed, max_ply=size * size + 40,
                                                     stride=3, with_moves=True)
        rng = np.random.RandomState(seed)
        # a third of the moves are replaced by arbitrary points (occupied, suicide, ko ...) or passes
        rnd = rng.randint(-1, size * size, size=n).astype(np.int32)
        moves = np.where(rng.rand(n) < 0.33, rnd, mv).astype(np.int32)
        koidx = np.flatnonzero(ko >= 0)
        moves[koidx[::2]] = ko[koidx[::2]]                  # retake the ko: must be refused
        wb, wa, wko, wpl, wst = go_oracle.do_move_batch(b, a, ko, pl, moves)
        batch = BatchedGameState.from_arrays(b, a, ko, pl)
        st = batch.do_move(torch.from_numpy(moves)).cpu().numpy()
        gb, ga, gko, gpl = batch.to_arrays()
        assert np.array_equal(st, wst)
        assert (st == -1).sum() > 10 and (st == 0).sum() > n // 2
        assert np.array_equal(gb, wb) and np.array_equal(gko, wko) and np.array_equal(gpl, wpl)
        assert np.array_equal(ga, np.minimum(wa, 7))


def test_device_playouts_replay_on_oracle():
    """Positions the device playout generator records equal the oracle's replay of the very
    moves it logged (board, saturated ages, ko, side to move)."""
    n, max_ply, stride = 600, 400, 8
    batch, log = BatchedGameState.random_playouts(n, seed=31, max_ply=max_ply, stride=stride, with_log=True)
    gb, ga, gko, gpl = batch.to_arrays()
    log = log.cpu().numpy()
    quota = max_ply // stride + 1
    checked = 0
    for w in range(log.shape[0]):
        seq = log[w]
        slot = w * quota
        games = np.flatnonzero(seq == -2)
        for gi, start in enumerate(games):
            end = games[gi + 1] if gi + 1 < len(games) else len(seq)
            moves = seq[start + 1:end]
            moves = moves[moves > -2]
            played, rb, ra, rko, rpl = go_oracle.replay(19, moves, None, every=True)
            assert played == len(moves)
            # iteration t of a game records the position before ply t (when t matches the
            # sampling rule and the warp still has a free slot), then plays and logs move t
            for t in range(len(moves)):
                if t % stride != (w + gi) % stride or slot >= min(n, (w + 1) * quota):
                    continue
                if t == 0:
                    assert not gb[slot].any() and gpl[slot] == 1 and gko[slot] == -1
                else:
                    assert np.array_equal(gb[slot], rb[t - 1]), (w, gi, t)
                    assert np.array_equal(ga[slot], np.minimum(ra[t - 1], 7))
                    assert gko[slot] == rko[t - 1] and gpl[slot] == rpl[t - 1]
                slot += 1
                checked += 1
        assert slot == min(n, (w + 1) * quota)
    assert checked == n
    stones = (gb != 0).sum(axis=(1, 2))
    assert stones.max() > 200 and stones.min() < 10          # all game phases are covered


@pytest.mark.parametrize("handover,helpers", [(0, 32), (3, 32), (96, 32), (-1, 32), (0, 0), (0, 1), (3, -1)])
def test_ladder_planes_do_not_depend_on_reply_handover(handover, helpers):
    """Long ladder searches hand replies over to idle warps after `handover` tried moves
    (0: from the start, -1: never); `helpers` CTAs keep a warp waiting for such replies (0: nobody
    helps and every search takes its offers back, -1: every CTA).  The planes must be the
    reference's either way."""
    from rocalphago_b200 import _cabi
    L = _cabi.lib()
    g = gu.load_features("features_playouts_19.npz")
    b, a, ko, pl = go_oracle.random_playouts(4000, size=19, seed=301, max_ply=400, stride=5)
    want = go_oracle.features(b, a, ko, pl, ["ladder_capture", "ladder_escape"])
    old = L.rag_set_ladder_handover(handover)
    old_helpers = L.rag_set_ladder_helpers(helpers)
    try:
        for rep in range(2):
            got = _gpu_planes(g["board"], g["age"], g["ko"], g["player"], ["ladder_capture", "ladder_escape"])
            assert np.array_equal(got, g["planes"][:, 44:46]), _explain(got, g["planes"][:, 44:46])
            got = _gpu_planes(b, a, ko, pl, ["ladder_capture", "ladder_escape"])
            assert np.array_equal(got, want), _explain(got, want)
        small = gu.load_features("features_playouts_7.npz")
        got = _gpu_planes(small["board"], small["age"], small["ko"], small["player"], ["ladder_escape", "ladder_capture"])
        assert np.array_equal(got, small["planes"][:, [45, 44]])
    finally:
        L.rag_set_ladder_handover(old)
        L.rag_set_ladder_helpers(old_helpers)


def test_module_level_feature_functions_and_their_maximum_argument():
    """get_board ... get_legal (preprocessing.py:14-205) keep their names and signatures; the
    `maximum` argument folds the saturating planes exactly as the reference does (golden vectors made
    by calling the reference's functions with maximum in {1, 3, 5, 8}, oracle/make_golden_maximum.py)."""
    from rocalphago_b200 import preprocessing as P
    g = gu.load_features("features_sgf.npz")
    z = np.load(os.path.join(gu.GOLDEN, "feature_functions_maximum.npz"))
    rows = [int(r) for r in z["rows"]]
    views = [PositionView(g["board"][i], g["age"][i], int(g["ko"][i]), int(g["player"][i])) for i in rows]
    for name in [str(s) for s in z["names"]]:
        fn = getattr(P, "get_" + name)
        assert P.FEATURES[name]["function"] is fn
        for m in [int(v) for v in z["maxima"]]:
            want = np.unpackbits(z["%s_%d" % (name, m)])[:len(rows) * m * 361].reshape(len(rows), m, 19, 19)
            for k, v in enumerate(views[::6]):
                got = fn(v, maximum=m)
                assert got.shape == (m, 19, 19) and got.dtype == np.float64
                assert np.array_equal(got, want[6 * k]), (name, m, k)
        with pytest.raises(ValueError):
            fn(views[0], maximum=9)
    v, i = views[3], rows[3]
    for name, lo in (("board", 0), ("ladder_capture", 44), ("ladder_escape", 45), ("sensibleness", 46), ("legal", 48)):
        got = getattr(P, "get_" + name)(v)
        assert np.array_equal(got, g["planes"][i, lo:lo + P.FEATURES[name]["size"]]), name
    assert np.array_equal(P.FEATURES["ones"]["function"](v), np.ones((1, 19, 19)))
    assert np.array_equal(P.FEATURES["zeros"]["function"](v), np.zeros((1, 19, 19)))
    assert P.Preprocess(["board", "liberties"]).processors == [P.get_board, P.get_liberties]


def test_copied_states_without_ages_and_forget_ages():
    """Stones with stone_ages -1 (what GameState.copy() leaves, go.py:186-217): planes against the
    reference's, pack/unpack round trip, and do_move keeping such stones ageless while new ones age."""
    z = np.load(os.path.join(gu.GOLDEN, "features_copied_states.npz"))
    shape = tuple(int(v) for v in z["planes_shape"])
    want = np.unpackbits(z["planes"])[:int(np.prod(shape))].reshape(shape)
    feats = [str(f) for f in z["feature_list"]]
    batch = BatchedGameState.from_arrays(z["board"], z["age"], z["ko"], z["player"])
    assert np.array_equal(Preprocess(feats).to_device(batch).cpu().numpy(), want)
    b, a, ko, pl = batch.to_arrays()
    assert np.array_equal(a, np.where(z["age"] < 0, -1, np.minimum(z["age"], 7))) and np.array_equal(b, z["board"])
    # forget_ages on the original positions + the recorded moves == the reference's copy() + do_move
    g = gu.load_features("features_sgf.npz")
    rows = z["source_rows"]
    orig = BatchedGameState.from_arrays(g["board"][rows], g["age"][rows], g["ko"][rows], g["player"][rows]).forget_ages()
    for step in range(4):
        mv = z["played"][:, step]
        keep = orig.packed.clone()
        orig.do_move(np.where(mv >= 0, mv, -1))
        orig.packed = torch.where(torch.as_tensor(mv >= 0, device=orig.device)[:, None], orig.packed, keep)
    assert np.array_equal(Preprocess(feats).to_device(orig).cpu().numpy(), want)


def test_ladder_kernel_stress_all_replies_handed_over_with_one_helper():
    """Liveness of the reply hand-over under its worst settings -- every search hands its replies over from
    the first move (rag_set_ladder_handover(0)) and a single CTA keeps a helper warp -- on 100,000
    positions: the kernel finishes, the watchdog does not fire, and the planes equal those of the default
    settings (hand-over never changes results)."""
    L = _cabi.lib()
    n = 100000
    batch = BatchedGameState.random_playouts(n, seed=7007)
    pp = Preprocess(["ladder_capture", "ladder_escape"])
    want = pp.to_device(batch).clone()
    P.check_workspace(pp._ws.get(n, batch.device))
    old_h, old_c = L.rag_set_ladder_handover(0), L.rag_set_ladder_helpers(1)
    try:
        got = pp.to_device(batch)
        P.check_workspace(pp._ws.get(n, batch.device))
    finally:
        L.rag_set_ladder_handover(old_h if old_h < (1 << 29) else -1)
        L.rag_set_ladder_helpers(old_c)
    assert int(want.sum()) > 1000
    assert torch.equal(got, want)


def test_ladder_watchdog_reports_instead_of_killing_the_context():
    """With an absurdly short watchdog, warps that wait for handed-over replies give up: the kernel still
    ends, rag_features_status (raised by the Python layer where it synchronises) says so, and the next
    call with the normal watchdog is clean and exact again."""
    L = _cabi.lib()
    b, a, ko, pl = go_oracle.random_playouts(3000, size=19, seed=808, max_ply=380, stride=3)
    batch = BatchedGameState.from_arrays(b, a, ko, pl)
    pp = Preprocess(["ladder_capture", "ladder_escape"])
    old_w, old_h = L.rag_set_ladder_watchdog(1), L.rag_set_ladder_handover(0)
    try:
        with pytest.raises(_cabi.RagError, match="watchdog"):
            pp.states_to_tensor(batch)
    finally:
        L.rag_set_ladder_watchdog(old_w)
        L.rag_set_ladder_handover(old_h if old_h < (1 << 29) else -1)
    got = pp.states_to_tensor(batch)                         # the context is alive, the flag was per call
    assert np.array_equal(got, go_oracle.features(b, a, ko, pl, ["ladder_capture", "ladder_escape"]))


def test_a_ladder_search_the_reference_cannot_finish_is_cut_off():
    """Device-generated batch 21805311 (found by the 4-GPU feature sweep) holds a position whose ladder
    recursion is astronomically large: the reference's unbounded recursion (go.py:329-463) and the C
    oracle do not return on it.  The device search gives up after 65,536 tried moves along one line,
    says so (rag_features_ladder_cutoffs), and the batch still finishes in well under a second."""
    import time
    L = _cabi.lib()
    n = 32768
    batch = BatchedGameState.random_playouts(n, seed=21805311)
    pp = Preprocess(["ladder_capture", "ladder_escape"])
    torch.cuda.synchronize()
    t0 = time.time()
    planes = pp.to_device(batch)
    torch.cuda.synchronize()
    took = time.time() - t0
    ws = pp._ws.get(n, batch.device)
    P.check_workspace(ws)                                     # the watchdog did not fire
    cut = L.rag_features_ladder_cutoffs(ws.data_ptr(), None)
    assert 1 <= cut <= 8, cut
    assert took < 5.0, took
    assert int(planes.sum()) > 1000
    # an ordinary batch cuts nothing off
    other = BatchedGameState.random_playouts(4096, seed=5)
    pp.to_device(other)
    assert L.rag_features_ladder_cutoffs(pp._ws.get(4096, other.device).data_ptr(), None) == 0
