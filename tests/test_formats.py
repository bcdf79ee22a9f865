"""On-disk formats either side of the hot path: the HDF5 subset reader/writer, the
SGF main-line reader, and -- the reference's own golden file -- the feature planes its authors'
GameConverter stored for the five AlphaGo - Lee Sedol games, against the oracle and the device."""
import glob
import json
import os

import numpy as np
import pytest

from oracle import go_oracle
from rocalphago_b200 import hdf5_lite, sgf
from tests import golden_util as gu

REF_DATA = "/root/reference/tests/test_data"
needs_reference = pytest.mark.skipif(not os.path.isdir(REF_DATA), reason="the reference checkout is not on this machine")


def _lee_sedol():
    z = np.load(os.path.join(gu.GOLDEN, "hdf5_lee_sedol_features.npz"))
    shape = tuple(int(v) for v in z["states_shape"])
    states = np.unpackbits(z["states_packed"])[:int(np.prod(shape))].reshape(shape)
    return states, z["actions"], [str(k) for k in z["file_keys"]], z["file_offsets"]


def _replayed_games():
    """name -> (moves, colours, n_setup, first player, states after every move) from the reference's replay."""
    z = np.load(os.path.join(gu.GOLDEN, "replay_sgf.npz"))
    out = {}
    for i, name in enumerate(z["names"]):
        out[str(name)] = {k: z["%s_%d" % (k, i)] for k in ("moves", "colors", "nsetup", "to_play", "board", "age", "ko", "player")}
    return out


def _positions_before_moves(g):
    """States before each move of a game without set-up stones, from the states after each move."""
    assert int(g["nsetup"]) == 0
    n = len(g["board"])
    board = np.concatenate([np.zeros((1, 19, 19), np.int8), g["board"][:-1]])
    age = np.concatenate([np.full((1, 19, 19), -1, np.int16), g["age"][:-1]])
    ko = np.concatenate([[-1], g["ko"][:-1]]).astype(np.int32)
    player = np.concatenate([[int(g["to_play"])], g["player"][:-1]]).astype(np.int32)
    return board[:n], age[:n], ko, player


# ---- the reference's golden HDF5 vs the oracle ----------------------------------------------------

def test_oracle_reproduces_the_references_stored_feature_file():
    states, actions, keys, offsets = _lee_sedol()
    games = _replayed_games()
    assert states.shape == (1033, 12, 19, 19) and len(keys) == 5
    for key, (start, count) in zip(keys, offsets):
        g = games[key.split(":")[-1]]
        board, age, ko, player = _positions_before_moves(g)
        assert count == len(board)                              # no passes in these records
        got = go_oracle.features(board, age, ko, player, ["board", "ones", "turns_since"])
        assert np.array_equal(got, states[start:start + count]), key
        mv = g["moves"]
        assert np.array_equal(np.stack([mv // 19, mv % 19], 1), actions[start:start + count]), key


# ---- HDF5 subset ------------------------------------------------------------------------------------

def test_lzf_decoder_known_answers():
    assert hdf5_lite.lzf_decompress(bytes([4]) + b"hello", 5) == b"hello"
    # literal "ab", then a back reference of length 6 at distance 2 (overlapping: abababab)
    assert hdf5_lite.lzf_decompress(bytes([1]) + b"ab" + bytes([(4 << 5) | 0, 1]), 8) == b"abababab"
    # long match: length 7+3+2 = 12 at distance 1
    assert hdf5_lite.lzf_decompress(bytes([0]) + b"z" + bytes([(7 << 5) | 0, 3, 0]), 13) == b"z" * 13
    with pytest.raises(hdf5_lite.Hdf5Error):
        hdf5_lite.lzf_decompress(bytes([(1 << 5) | 0, 5]), 3)      # reference before the start
    with pytest.raises(hdf5_lite.Hdf5Error):
        hdf5_lite.lzf_decompress(bytes([4]) + b"hello", 6)         # short stream


def test_hdf5_writer_reader_round_trip(tmp_path):
    rng = np.random.RandomState(3)
    w = hdf5_lite.Writer()
    w.attrs["layer_names"] = np.array([b"convolution2d_1", b"bias_1", b"activation_1"])
    g = w.create_group("convolution2d_1")
    g.attrs["weight_names"] = np.array([b"convolution2d_1_W", b"convolution2d_1_b"])
    W = rng.rand(8, 3, 5, 5).astype(np.float32)
    g["convolution2d_1_W"], g["convolution2d_1_b"] = W, np.arange(8, dtype=np.float32)
    w.create_group("activation_1").attrs["weight_names"] = np.zeros(0, "S1")
    b = w.create_group("bias_1")
    b["param_0"] = np.zeros(361, np.float32)
    w["features"] = np.bytes_("board,ones,turns_since")
    w["states"] = rng.randint(0, 2, size=(7, 12, 19, 19)).astype(np.uint8)
    w["i64"], w["f64"], w["f16"], w["none"] = np.array([-5, 2 ** 40]), np.array(2.5), np.array([1.5], np.float16), np.zeros((0, 2), np.uint8)
    fo = w.create_group("file_offsets")
    for i in range(700):                                        # more entries than a default symbol node holds
        fo["dir:game%04d.sgf" % i] = np.array([i * 10, 10])
    path = str(tmp_path / "x.hdf5")
    w.save(path)
    assert open(path, "rb").read() == w.tobytes()
    f = hdf5_lite.File(path)
    assert f.keys() == sorted(["convolution2d_1", "activation_1", "bias_1", "features", "states", "i64", "f64", "f16", "none", "file_offsets"])
    assert [n.decode() for n in f.attrs["layer_names"]] == ["convolution2d_1", "bias_1", "activation_1"]
    assert np.array_equal(f["convolution2d_1/convolution2d_1_W"][...], W)
    assert f["convolution2d_1"].attrs["weight_names"].tolist() == [b"convolution2d_1_W", b"convolution2d_1_b"]
    assert len(f["activation_1"].attrs["weight_names"]) == 0 and f["activation_1"].keys() == []
    assert f["features"][()] == b"board,ones,turns_since"
    assert np.array_equal(f["states"][...], w["states"].data) and f["states"].shape == (7, 12, 19, 19)
    assert f["i64"][...].tolist() == [-5, 2 ** 40] and f["f64"][()] == 2.5 and f["f16"][...].tolist() == [1.5]
    assert f["none"].shape == (0, 2) and f["none"][...].shape == (0, 2)
    assert len(f["file_offsets"]) == 700 and f["file_offsets"]["dir:game0699.sgf"][...].tolist() == [6990, 10]
    assert "dir:game0123.sgf" in f["file_offsets"] and "nope" not in f["file_offsets"]
    with pytest.raises(KeyError):
        f["missing"]
    with pytest.raises(hdf5_lite.Hdf5Error):
        hdf5_lite.File(b"not an hdf5 file at all")


@needs_reference
def test_reads_the_references_weight_file_and_feature_file():
    """Files written by h5py / libhdf5 (old-style groups, contiguous float datasets, string-array
    attributes; chunked LZF-compressed uint8 datasets)."""
    f = hdf5_lite.File(os.path.join(REF_DATA, "hdf5", "random_minimodel_weights.hdf5"))
    z = np.load(os.path.join(gu.GOLDEN, "minimodel.npz"))
    names = [n.decode() for n in f.attrs["layer_names"]]
    assert names == ["convolution2d_%d" % i for i in range(1, 7)] + ["flatten_1", "bias_1", "activation_1"]
    assert f["convolution2d_1"]["convolution2d_1_W"].shape == (16, 12, 5, 5)
    assert np.array_equal(f["bias_1"]["param_0"][...], z["w12"])
    assert np.array_equal(f["convolution2d_3/convolution2d_3_W"][...], z["w04"])
    g = hdf5_lite.File(os.path.join(REF_DATA, "hdf5", "alphago-vs-lee-sedol-features.hdf5"))
    states, actions, keys, offsets = _lee_sedol()
    assert g["states"].shape == (1033, 12, 19, 19) and g["states"].dtype == np.uint8
    assert np.array_equal(g["states"][...], states) and np.array_equal(g["actions"][...], actions)
    assert g["file_offsets"].keys() == keys
    assert np.array_equal(np.array([g["file_offsets"][k][...] for k in keys]), offsets)


# ---- SGF ------------------------------------------------------------------------------------------

def test_sgf_main_line():
    g = sgf.read("(;GM[1]SZ[9];B[aa](;W[bb];B[cc](;W[ee])(;W[ff]))(;W[dd];B[gg]))")
    assert g.size == 9 and g.moves == [(1, (0, 0)), (-1, (1, 1)), (1, (2, 2)), (-1, (4, 4))]
    g = sgf.read("(;SZ[9]C[a \\] ( ; b];B[aa];W[];B[tt])")
    assert g.moves == [(1, (0, 0)), (-1, None), (1, None)]
    g = sgf.read("(;AB[aa][bb]AW[cc]PL[W];W[dd])")
    assert (g.size, g.add_black, g.add_white, g.first_player, g.moves) == (19, [(0, 0), (1, 1)], [(2, 2)], -1, [(-1, (3, 3))])
    g = sgf.read("(;B[ab]W[cd])")                              # root only: no moves (util.py:113 iterates game.rest)
    assert g.moves == []
    with pytest.raises(sgf.ParseException):
        sgf.read("no game here")
    with pytest.raises(sgf.ParseException):
        sgf.read("(;SZ[9];B[aa")


@needs_reference
def test_sgf_reader_agrees_with_the_references_replay_of_its_test_games():
    games = _replayed_games()
    files = glob.glob(os.path.join(REF_DATA, "sgf*", "*.sgf"))
    assert len(files) == 7
    for path in files:
        g = sgf.read(open(path).read())
        ref = games[os.path.basename(path)]
        assert g.size == 19 and g.first_player == int(ref["to_play"])
        assert len(g.add_black) + len(g.add_white) == int(ref["nsetup"])
        flat = [(-1 if p is None else p[0] * 19 + p[1]) for _, p in g.moves]
        setup = [p[0] * 19 + p[1] for p in g.add_black + g.add_white]
        assert setup + flat == ref["moves"].tolist()
        assert [1] * len(g.add_black) + [-1] * len(g.add_white) + [c for c, _ in g.moves] == ref["colors"].tolist()


# ---- model files ----------------------------------------------------------------------------------

@needs_reference
def test_load_model_reads_the_references_minimodel(tmp_path):
    """NeuralNetBase.load_model on tests/test_data/minimodel.json: Keras 1.1.2 Sequential JSON plus
    an HDF5 weight file, resolved against the reference's root as the reference does."""
    from rocalphago_b200.models import NeuralNetBase
    net = NeuralNetBase.load_model(os.path.join(REF_DATA, "minimodel.json"))
    z = np.load(os.path.join(gu.GOLDEN, "minimodel.npz"))
    assert type(net).__name__ == str(z["net_class"]) == "CNNPolicy"
    assert net.preprocessor.feature_list == [str(s) for s in z["feature_list"]]
    m = net.model
    assert (m.input_dim, m.board, m.board) == tuple(z["input_shape"])
    assert m.specs == [tuple(int(v) for v in c) for c in z["convs"][:-1]] and m.head == "policy"
    ws = m.get_weights()
    assert len(ws) == 13 and all(np.array_equal(w, z["w%02d" % i]) for i, w in enumerate(ws))
    # save_model with an .hdf5 weights file writes the Keras layout back; load_model reads it again
    js, wf = str(tmp_path / "m.json"), str(tmp_path / "weights.hdf5")
    net.save_model(js, wf)
    f = hdf5_lite.File(wf)
    assert [n.decode() for n in f.attrs["layer_names"]] == ["convolution2d_%d" % i for i in range(1, 7)] + ["flatten_1", "bias_1", "activation_1"]
    assert f["convolution2d_6"].attrs["weight_names"].tolist() == [b"convolution2d_6_W", b"convolution2d_6_b"]
    again = NeuralNetBase.load_model(js)
    assert all(np.array_equal(a, b) for a, b in zip(ws, again.model.get_weights()))
    assert json.load(open(js))["class"] == "CNNPolicy"


def test_writer_handles_groups_with_more_children_than_one_symbol_table_node(tmp_path):
    """sgfs_to_hdf5 makes one `file_offsets` entry per game; a KGS-sized corpus has more games than the
    65,534 entries one symbol-table node can hold.  The writer then spreads the group over several nodes
    under its B-tree node; the reader (which also reads h5py's files) finds every entry."""
    w = hdf5_lite.Writer()
    g = w.create_group("file_offsets")
    n = 70000
    for i in range(n):
        g["game_%06d.sgf" % i] = np.array([i, i % 7], np.int64)
    w["features"] = np.bytes_("board,ones")
    path = str(tmp_path / "many.h5")
    w.save(path)
    f = hdf5_lite.File(path)
    assert len(f["file_offsets"]) == n
    for i in (0, 1, 65533, 65534, 65535, n - 1):
        assert f["file_offsets"]["game_%06d.sgf" % i][...].tolist() == [i, i % 7]
    assert sorted(f["file_offsets"])[:2] == ["game_000000.sgf", "game_000001.sgf"]


def test_functional_model_json_of_a_resnet_policy_round_trips():
    """A reference-built ResnetPolicy is a Keras functional `Model` (policy.py:271); nn_util.py:96-101 saves
    `model.to_json()`: InputLayer, layers with inbound_nodes, Merge(mode='sum') closing each residual unit.
    Network.from_json reads that form (and to_json writes it), including uneven n_skip_K."""
    import json
    from rocalphago_b200.models.nn_util import Network
    from rocalphago_b200.models.policy import ResnetPolicy
    net = ResnetPolicy.create_network(input_dim=12, board=9, layers=7, filters_per_layer=16, n_skip_1=2, n_skip_3=1,
                                      n_skip_4=3, seed=1)
    text = net.to_json()
    km = json.loads(text)
    assert km["class_name"] == "Model" and km["config"]["input_layers"] == [["input_1", 0, 0]]
    classes = [l["class_name"] for l in km["config"]["layers"]]
    assert classes.count("Merge") == sum(net.unit_end) - 1 and classes.count("BatchNormalization") == len(net.specs) - 1
    merge = [l for l in km["config"]["layers"] if l["class_name"] == "Merge"][0]
    assert [s[0] for s in merge["inbound_nodes"][0]] == ["convolution2d_1", "convolution2d_3"] and merge["config"]["mode"] == "sum"
    back = Network.from_json(text)
    assert back.specs == net.specs and back.unit_end == net.unit_end and (back.board, back.input_dim) == (9, 12)
    assert [w.shape for w in back.get_weights()] == [w.shape for w in net.get_weights()]
    # a graph this path does not run is refused with a reason
    km["config"]["layers"][2]["config"]["axis"] = 1
    with pytest.raises(ValueError, match="BatchNormalization"):
        Network.from_json(json.dumps(km))
