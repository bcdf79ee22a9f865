"""SGF game records -> (states, actions) training data, on the device.

Drop-in for AlphaGo/preprocessing/game_converter.py: `GameConverter(features)`,
`.convert_game(file_name, bd_size)`, `.sgfs_to_hdf5(sgf_files, hdf5_file, bd_size, ignore_errors,
verbose)` and `run_game_converter(argv)` keep their names, arguments, warnings and output layout
(game_converter.py:16-141):

    states        uint8 (n, n_features, S, S)     the position *before* each non-pass move
    actions       uint8 (n, 2)                    (x, y) of the move played there
    file_offsets  group: file name with '/' -> ':'  ->  [first index, count]
    features      the comma-separated feature list

What differs is how the work is done: the games of a batch are replayed in lock step by
`rag_do_move` (one warp per game), every position before a move is kept in HBM, and the planes of
all of them come from one pass of the feature kernels.  The HDF5 file is written by
`hdf5_lite.Writer` (contiguous, uncompressed datasets in the layout libhdf5 1.8 gives a new file; read
back here by `hdf5_lite.File`, which also reads the reference's chunked LZF files -- no libhdf5 / h5py is
on this image to cross-check the writer with).
"""
import os
import warnings

import numpy as np
import torch

from . import hdf5_lite, sgf
from .preprocessing import Preprocess, LAYOUT_NCHW_U8
from .state import BatchedGameState, STATE_WORDS, pack_arrays

PLAYER_WORD = 5 * 32 + 1                # meta plane, word 1 (include/rocalphago_b200.h: packed game states)


class SizeMismatchError(Exception):
    pass


class IllegalMove(Exception):            # go.py:584
    pass


def _flat(pt, size):
    """Flat point x*S+y; a coordinate off the board (an SGF letter past the board size) becomes S*S,
    which no kernel accepts: the move is reported illegal, as GameState.is_legal does (go.py:267-270)."""
    if pt is None:
        return -1
    x, y = pt
    return x * size + y if 0 <= x < size and 0 <= y < size else size * size


def replay_games(games, size, device=None):
    """Replays parsed games (sgf.Game) in lock step on the device.

    Returns (packed, game_index, ply, moves, illegal_at): `packed` (m, 192) int32 on the device
    holds the position before every main-line move that was reached (m = sum of usable plies),
    `game_index`/`ply` say where each row comes from, `moves` the flat move played there (-1 =
    pass), and `illegal_at[g]` is the ply at which game g hit an illegal move (that ply's pair is kept, as
    the reference's generator yields it before do_move raises; later plies are dropped,
    game_converter.py:117-119) or -1."""
    dev = torch.device(device or "cuda")
    G = len(games)
    if G == 0:
        return (torch.empty((0, STATE_WORDS), dtype=torch.int32, device=dev), np.zeros(0, np.int64),
                np.zeros(0, np.int64), np.zeros(0, np.int32), np.zeros(0, np.int64))
    # set-up stones are played with do_move(stone, colour) before the first node (util.py:37-47)
    setup = [[(sgf.BLACK, p) for p in g.add_black] + [(sgf.WHITE, p) for p in g.add_white] for g in games]
    n_setup = np.array([len(s) for s in setup])
    n_moves = np.array([len(g.moves) for g in games])
    T0, T1 = int(n_setup.max()), int(n_moves.max()) if G else 0
    mv = np.full((G, T0 + T1), -1, np.int32)
    col = np.ones((G, T0 + T1), np.int32)
    for i, g in enumerate(games):
        # right-align the set-up so that every game's first real move is at step T0
        seq = setup[i]
        for j, (c, p) in enumerate(seq):
            mv[i, T0 - len(seq) + j], col[i, T0 - len(seq) + j] = _flat(p, size), c
        for j, (c, p) in enumerate(g.moves):
            mv[i, T0 + j], col[i, T0 + j] = _flat(p, size), c
    first = np.array([g.first_player for g in games], np.int32)

    empty = pack_arrays(np.zeros((G, size, size), np.int8), np.full((G, size, size), -1, np.int16),
                        np.full(G, -1, np.int32), np.ones(G, np.int32))
    batch = BatchedGameState(empty.to(dev), size)
    d_mv, d_col = torch.as_tensor(mv).to(dev), torch.as_tensor(col).to(dev)
    d_first = torch.as_tensor(first).to(dev)
    d_nsetup, d_nmoves = torch.as_tensor(n_setup).to(dev), torch.as_tensor(n_moves).to(dev)
    snaps = torch.empty((T1, G, STATE_WORDS), dtype=torch.int32, device=dev)
    never = T0 + T1 + 1
    bad_at = torch.full((G,), never, dtype=torch.int64, device=dev)
    for t in range(T0 + T1):
        if t >= T0:
            if t == T0:
                batch.packed[:, PLAYER_WORD] = d_first                     # 'PL' (util.py:48-49)
            snaps[t - T0].copy_(batch.packed)
            active = (d_nmoves > t - T0)
        else:
            active = (d_nsetup >= T0 - t)
        before = batch.packed[:, PLAYER_WORD].clone()
        batch.packed[:, PLAYER_WORD] = torch.where(active, d_col[:, t], before)   # do_move(action, color): go.py:526-528
        # games that are not at this step (already over / set-up not started yet) play nothing
        # that matters: their state is never looked at again, or is overwritten below
        frozen = batch.packed.clone() if t < T0 else None
        status = batch.do_move(d_mv[:, t])
        illegal = (status < 0) & active
        batch.packed[:, PLAYER_WORD] = torch.where(illegal, before, batch.packed[:, PLAYER_WORD])
        bad_at = torch.where(illegal & (bad_at == never), torch.full_like(bad_at, t), bad_at)
        if frozen is not None:
            batch.packed.copy_(torch.where(active[:, None], batch.packed, frozen))
    bad = bad_at.cpu().numpy()
    if (bad < T0).any():
        i = int(np.argmax(bad < T0))
        raise IllegalMove("set-up stone of game %d cannot be placed" % i)
    # sgf_iter_states (util.py:124-126) yields (state, move) BEFORE do_move raises, so the reference stores the
    # pair whose move is illegal and drops only what follows
    usable = np.minimum(n_moves, bad - T0 + 1)
    game_index = np.repeat(np.arange(G), usable)
    ply = np.concatenate([np.arange(u) for u in usable]) if G else np.zeros(0, np.int64)
    rows = torch.as_tensor(ply * G + game_index, device=dev)
    packed = snaps.reshape(T1 * G, STATE_WORDS)[rows] if len(ply) else snaps.reshape(-1, STATE_WORDS)[:0]
    moves = mv[game_index, T0 + ply] if len(ply) else np.zeros(0, np.int32)
    illegal_at = np.where(bad < never, bad - T0, -1)
    return packed, game_index, ply, moves, illegal_at


class GameConverter(object):

    def __init__(self, features, batch_games=256, batch_positions=16384):
        self.feature_processor = Preprocess(features)
        self.n_features = self.feature_processor.output_dim
        self.batch_games = int(batch_games)
        self.batch_positions = int(batch_positions)

    # ---- one game (reference API) ------------------------------------------------------------
    def _read(self, file_name, bd_size):
        with open(file_name, "r") as fh:
            game = sgf.read(fh.read())
        if game.size != bd_size:
            raise SizeMismatchError()
        return game

    def _planes(self, packed, size):
        out = []
        for i in range(0, packed.shape[0], self.batch_positions):
            part = BatchedGameState(packed[i:i + self.batch_positions].contiguous(), size)
            out.append(self.feature_processor.to_device(part, LAYOUT_NCHW_U8).cpu().numpy())
        return np.concatenate(out) if out else np.zeros((0, self.n_features, size, size), np.uint8)

    def convert_game(self, file_name, bd_size):
        """Yields (nn_input, move) for every non-pass move of the main line: nn_input is the
        (1, F, S, S) uint8 feature tensor of the position before the move, move its (x, y).
        Raises SizeMismatchError, sgf.ParseException, or -- after the pairs before it --
        IllegalMove, like game_converter.py:21-40."""
        game = self._read(file_name, bd_size)
        packed, _, _, moves, illegal_at = replay_games([game], bd_size)
        keep = np.flatnonzero(moves >= 0)
        planes = self._planes(packed[torch.as_tensor(keep, device=packed.device)], bd_size)
        for k, i in enumerate(keep):
            yield planes[k:k + 1], tuple(game.moves[i][1])       # the point as the file gives it (also when it is off the board)
        if illegal_at[0] >= 0:
            raise IllegalMove("move %d of %s" % (illegal_at[0], file_name))

    # ---- many games ----------------------------------------------------------------------------
    def sgfs_to_hdf5(self, sgf_files, hdf5_file, bd_size=19, ignore_errors=True, verbose=False):
        """Converts all files in the iterable `sgf_files` into `hdf5_file` (layout in the module
        docstring; game_converter.py:42-157).  Unparsable games, games of another size and the tail
        of games with an illegal move are skipped with the reference's warnings."""
        out_dir = os.path.dirname(hdf5_file)
        tmp_file = os.path.join(out_dir, ".tmp." + os.path.basename(hdf5_file))
        raw_file = tmp_file + ".states"
        actions, offsets, n_total = [], [], 0
        try:
            with open(raw_file, "wb") as raw:
                pending = []

                def flush():
                    nonlocal n_total
                    if not pending:
                        return
                    names, games = zip(*pending)
                    del pending[:]
                    packed, gi, ply, moves, illegal_at = replay_games(list(games), bd_size)
                    keep = moves >= 0
                    rows = torch.as_tensor(np.flatnonzero(keep), device=packed.device)
                    planes = self._planes(packed[rows], bd_size)
                    raw.write(planes.tobytes())
                    mk, gk = moves[keep], gi[keep]
                    act = np.stack([mk // bd_size, mk % bd_size], axis=1)
                    for k in np.flatnonzero(mk >= bd_size * bd_size):          # an off-board point: stored as the file gives it
                        act[k] = games[gk[k]].moves[ply[keep][k]][1]
                    actions.append(act.astype(np.uint8))
                    counts = np.bincount(gk, minlength=len(games))
                    for g, name in enumerate(names):
                        if illegal_at[g] >= 0:
                            warnings.warn("Illegal Move encountered in %s\n"
                                          "\tdropping the remainder of the game" % name)
                        if counts[g] > 0:
                            offsets.append((name.replace("/", ":"), n_total, int(counts[g])))
                            if verbose:
                                print("\t%d state/action pairs extracted from %s" % (counts[g], name))
                        elif verbose:
                            print("\t-no usable data- in %s" % name)
                        n_total += int(counts[g])

                for file_name in sgf_files:
                    if verbose:
                        print(file_name)
                    try:
                        pending.append((file_name, self._read(file_name, bd_size)))
                    except sgf.ParseException:
                        warnings.warn("Could not parse %s\n\tdropping game" % file_name)
                    except SizeMismatchError:
                        warnings.warn("Skipping %s; wrong board size" % file_name)
                    except Exception as e:
                        if ignore_errors:
                            warnings.warn("Unkown exception with file %s\n\t%s" % (file_name, e), stacklevel=2)
                        else:
                            raise
                    if len(pending) >= self.batch_games:
                        flush()
                flush()

            w = hdf5_lite.Writer()
            shape = (n_total, self.n_features, bd_size, bd_size)
            w["states"] = (np.memmap(raw_file, dtype=np.uint8, mode="r", shape=shape) if n_total
                           else np.zeros(shape, np.uint8))
            w["actions"] = np.concatenate(actions) if actions else np.zeros((0, 2), np.uint8)
            fo = w.create_group("file_offsets")
            for key, start, count in offsets:
                fo[key] = np.array([start, count], np.int64)
            w["features"] = np.bytes_(",".join(self.feature_processor.feature_list))
            w.save(tmp_file)
        except BaseException:
            print("sgfs_to_hdf5 failed")
            for f in (tmp_file, raw_file):
                if os.path.exists(f):
                    os.remove(f)
            raise
        os.remove(raw_file)
        if verbose:
            print("finished. renaming %s to %s" % (tmp_file, hdf5_file))
        os.rename(tmp_file, hdf5_file)


def run_game_converter(cmd_line_args=None):
    """Command line of the reference's converter (game_converter.py:160-226): same options."""
    import argparse
    import sys

    parser = argparse.ArgumentParser(description="Prepare SGF Go game files for training the neural network model.")
    parser.add_argument("--features", "-f", default="all", help="Comma-separated list of features to compute and store or 'all'")
    parser.add_argument("--outfile", "-o", required=True, help="Destination to write data (hdf5 file)")
    parser.add_argument("--recurse", "-R", default=False, action="store_true", help="Recurse through directories searching for SGF files")
    parser.add_argument("--directory", "-d", default=None, help="Directory containing SGF files to process; without it, file names are read from stdin")
    parser.add_argument("--size", "-s", type=int, default=19, help="Size of the game board. SGFs not matching this are discarded with a warning")
    parser.add_argument("--verbose", "-v", default=False, action="store_true", help="Turn on verbose mode")
    args = parser.parse_args(cmd_line_args)

    if args.features.lower() == "all":
        from .preprocessing import DEFAULT_FEATURES
        feature_list = list(DEFAULT_FEATURES)
    else:
        feature_list = args.features.split(",")
    if args.verbose:
        print("using features", feature_list)
    converter = GameConverter(feature_list)

    def is_sgf(name):
        return name.strip()[-4:] == ".sgf"

    def walk(root):
        for dirpath, _, names in sorted(os.walk(root)):
            for name in sorted(names):
                if is_sgf(name):
                    yield os.path.join(dirpath, name)

    def listing(root):
        for name in sorted(os.listdir(root)):
            if is_sgf(name):
                yield os.path.join(root, name)

    if args.directory:
        files = walk(args.directory) if args.recurse else listing(args.directory)
    else:
        files = (line.strip() for line in sys.stdin if is_sgf(line))
    converter.sgfs_to_hdf5(files, args.outfile, bd_size=args.size, verbose=args.verbose)


if __name__ == "__main__":
    run_game_converter()
