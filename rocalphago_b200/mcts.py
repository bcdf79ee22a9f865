"""Monte Carlo tree search whose leaf evaluations are batched on the device.

The search rule is AlphaGo/mcts.py's, statement for statement:

  * a playout walks from the root for `playout_depth` plies; a node met as a leaf is expanded with the
    policy network's distribution over get_legal_moves() (mcts.py:151-158), then the child of largest
    Q + u is taken (mcts.py:40-46, 160-161);
  * the state reached is scored `(1 - lmbda) * value(state) + lmbda * rollout` (mcts.py:167-169) and the
    SAME number is added to every node of the walk, root first: n += 1, Q += (v - Q) / n,
    u = c_puct * P * sqrt(parent.n) / (1 + n) (mcts.py:60-80) -- no sign flip between plies, as written there;
  * get_move returns the most visited child of the root (mcts.py:203-206); update_with_move re-roots.

What differs is the schedule.  The reference runs one playout at a time, each evaluation a network call
on one position.  Here `batch` playouts advance in lock step: at every ply the positions that sit on
unexpanded nodes go through the policy network in ONE batch (features + forward on the device,
CNNPolicy.eval_arrays), all moves are played by one rag_do_move launch, and the leaves are scored by one
value-network batch.  While a wave is in flight each playout leaves a virtual visit on the nodes it
took (`virtual_loss` subtracted from their score), so that the wave spreads over the tree; the visits
become real ones at back-up.  With batch=1 there is no virtual visit and the search is the reference's,
playout for playout (tests/test_gpu_mcts.py compares visit counts with the reference's own code).

`forget_ages=True` (default) reproduces one more thing the reference does: every playout starts from
`state.copy()`, and GameState.copy() does not copy stone_ages (go.py:186-217), so the networks see the
stones of the root position without any turns_since bit.  Pass False to search with the real ages.

The tree itself is native host code (csrc/tree.cpp behind rag_tree_*): a wave costs a handful of calls per ply
instead of a Python loop over its playouts; `TreeNode` below is a read-only view of one of its nodes.

Rollouts (lmbda > 0) use uniformly random sensible moves to the end of the game on the device
(`rollout_limit` plies) and area scoring as go.py:482-506; the reference's greedy-policy rollout
(one network call per ply) is not offered.
"""
import ctypes

import numpy as np
import torch

from . import _cabi
from .state import BatchedGameState, STATE_WORDS


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class TreeNode(object):
    """Read-only view of one node of the native tree and of its children's statistics (mcts.py:11-92 keeps one
    object per child): `actions` (flat moves x*S+y), `P`, `Q`, `u`, `n`, `virtual` as arrays over the children,
    `children` (views, None where a child was never walked into), `visits` / `q` of the node itself."""

    def __init__(self, search, node):
        self._search, self._node = search, int(node)

    def _arrays(self):
        L, h = _cabi.lib(), self._search._tree
        k = _cabi_check_count(L.rag_tree_node_children(h, self._node))
        out = (np.empty(k, np.int32), np.empty(k), np.empty(k), np.empty(k), np.empty(k, np.int64), np.empty(k, np.int64),
               np.empty(k, np.int32))
        _cabi.check(L.rag_tree_node_arrays(h, self._node, *[_p(a) for a in out]))
        return out

    def _stats(self):
        L, h = _cabi.lib(), self._search._tree
        visits, q, parent, index = ctypes.c_longlong(), ctypes.c_double(), ctypes.c_int32(), ctypes.c_int32()
        _cabi.check(L.rag_tree_node_stats(h, self._node, ctypes.byref(visits), ctypes.byref(q), ctypes.byref(parent), ctypes.byref(index)))
        return int(visits.value), float(q.value), int(parent.value), int(index.value)

    actions = property(lambda self: self._arrays()[0] if not self.is_leaf() else None)
    P = property(lambda self: self._arrays()[1])
    Q = property(lambda self: self._arrays()[2])
    u = property(lambda self: self._arrays()[3])
    n = property(lambda self: self._arrays()[4])
    virtual = property(lambda self: self._arrays()[5])
    visits = property(lambda self: self._stats()[0])
    q = property(lambda self: self._stats()[1])
    index = property(lambda self: self._stats()[3])

    @property
    def parent(self):
        p = self._stats()[2]
        return None if p < 0 else TreeNode(self._search, p)

    @property
    def children(self):
        return [None if c < 0 else TreeNode(self._search, c) for c in self._arrays()[6].tolist()]

    def is_leaf(self):
        return _cabi_check_count(_cabi.lib().rag_tree_node_children(self._search._tree, self._node)) == 0

    def is_root(self):
        return self._stats()[2] < 0

    def n_visits(self):
        return self.visits


def _cabi_check_count(rc):
    if rc < 0:
        _cabi.check(rc)
    return rc


class BatchedMCTS(object):
    """mcts.MCTS with batched leaf evaluation.

        search = BatchedMCTS(policy, value, c_puct=5, playout_depth=20, n_playout=1600, batch=64)
        move = search.get_move(state)            # (x, y) or None (pass) -- state: GameState-like or BatchedGameState of 1
        search.update_with_move(move)
    """

    def __init__(self, policy, value=None, lmbda=0.0, c_puct=5, rollout_limit=500, playout_depth=20, n_playout=10000,
                 batch=64, virtual_loss=1.0, seed=0, forget_ages=True):
        if not 0.0 <= lmbda <= 1.0:
            raise ValueError("lmbda must be in [0, 1]")
        if value is None and lmbda < 1.0:
            raise ValueError("a value network is needed unless lmbda == 1")
        self._policy, self._value = policy, value
        self._lmbda, self._c_puct = float(lmbda), float(c_puct)
        self._rollout_limit, self._L, self._n_playout = int(rollout_limit), int(playout_depth), int(n_playout)
        self._batch, self._vl = max(1, int(batch)), float(virtual_loss)
        self._seed = int(seed)
        self._forget_ages = bool(forget_ages)
        self._tree = ctypes.c_void_p()
        _cabi.check(_cabi.lib().rag_tree_create(ctypes.c_double(self._c_puct), ctypes.byref(self._tree)))
        self._size = None
        self.evaluations = 0            # positions sent through the policy network (for reports)

    def __del__(self):
        try:
            if self._tree:
                _cabi.lib().rag_tree_destroy(self._tree)
        except Exception:  # noqa: BLE001
            pass

    @property
    def _root(self):
        return TreeNode(self, _cabi.lib().rag_tree_root(self._tree))

    # ------------------------------------------------------------------ one wave of playouts
    def _wave(self, root_packed, size, count):
        L, h, dev = _cabi.lib(), self._tree, root_packed.device
        states = BatchedGameState(root_packed.expand(count, STATE_WORDS).clone(), size)       # a fresh copy even for count == 1
        if self._forget_ages:
            states.forget_ages()         # every playout starts from state.copy() (mcts.py:201), which drops stone_ages
        _cabi.check(L.rag_tree_wave_begin(h, count, ctypes.c_double(self._vl)))
        rows, moves = np.empty(count, np.int32), np.empty(count, np.int32)
        for _ in range(self._L):
            # expand the leaves among the nodes the live playouts stand on -- each node once, in one policy batch
            k = _cabi_check_count(L.rag_tree_wave_leaves(h, _p(rows)))
            if k:
                sub = BatchedGameState(states.packed[torch.as_tensor(rows[:k].astype(np.int64), device=dev)].contiguous(), size)
                _, flat, counts, probs = self._policy.eval_arrays(sub)
                self.evaluations += k
                flat = np.ascontiguousarray(flat, np.int64)
                counts = np.ascontiguousarray(counts, np.int64)
                probs = np.ascontiguousarray(probs, np.float32)
                _cabi.check(L.rag_tree_wave_expand(h, k, _p(rows), _p(counts), _p(flat), _p(probs)))
            # every live playout takes the child of largest Q + u; one without children is at the end of the game (mcts.py:154-156)
            alive = _cabi_check_count(L.rag_tree_wave_step(h, _p(moves)))
            if alive == 0:
                break
            moving = torch.as_tensor(moves >= 0, device=dev)
            before = states.packed.clone()
            states.do_move(np.where(moves >= 0, moves, -1))
            states.packed = torch.where(moving[:, None], states.packed, before)     # the others keep their position
        # leaf values: (1 - lmbda) * v + lmbda * z  (mcts.py:163-169)
        v = np.zeros(count)
        if self._lmbda < 1.0:
            v = np.asarray(self._value.batch_eval_state(states), np.float64)
        if self._lmbda > 0.0:
            z = self._rollouts(states)
            v = (1.0 - self._lmbda) * v + self._lmbda * z
        v = np.ascontiguousarray(v, np.float64)
        _cabi.check(L.rag_tree_wave_backup(h, _p(v)))

    def _rollouts(self, states):
        """+1 / -1 / 0 for the player to move at each leaf after a random playout to the end (mcts.py:171-189,
        with uniformly random sensible moves instead of the greedy rollout policy)."""
        from . import scoring
        self._seed += 1
        return scoring.random_rollout_results(states, self._rollout_limit, self._seed)

    # ------------------------------------------------------------------ reference API
    def get_move(self, state):
        """Runs n_playout playouts (in waves of `batch`) and returns the most visited move (mcts.py:191-206)."""
        batch = state if isinstance(state, BatchedGameState) else BatchedGameState.from_states([state])
        if len(batch) != 1:
            raise ValueError("get_move takes one position")
        self._size = batch.size
        done = 0
        while done < self._n_playout:
            count = min(self._batch, self._n_playout - done)
            self._wave(batch.packed[:1], batch.size, count)
            done += count
        root = self._root
        if root.is_leaf():
            return None
        a = int(root.actions[int(np.argmax(root.n))])
        return (a // batch.size, a % batch.size)

    def visit_counts(self, size):
        """{(x, y): visits} of the root's children."""
        r = self._root
        if r.is_leaf():
            return {}
        return {(int(a) // size, int(a) % size): int(n) for a, n in zip(r.actions, r.n)}

    def update_with_move(self, last_move, size=None):
        """Steps the root to the child for `last_move`, keeping its subtree (mcts.py:208-217)."""
        L = _cabi.lib()
        if last_move is None:
            _cabi.check(L.rag_tree_reset(self._tree))
            return
        if np.isscalar(last_move):
            flat = int(last_move)
        else:
            s = size if size is not None else self._size
            if s is None:
                raise ValueError("pass size= with an (x, y) move")
            flat = last_move[0] * s + last_move[1]
        _cabi_check_count(L.rag_tree_advance(self._tree, flat))


class ParallelMCTS(object):
    """Several searches at once (the reference's `ParallelMCTS`, mcts.py:219-220, is an empty subclass): one
    BatchedMCTS tree per position, all advanced in lock step, so that every ply costs ONE policy batch over the
    unexpanded nodes of all trees, one rag_do_move launch and, at the end of the wave, one value batch -- the
    leaf-evaluation pattern of BASELINE config 5.  Each tree ends up exactly as if it had been searched alone
    with the same settings (the device evaluates every position independently of its batch).

        par = ParallelMCTS(policy, value, games=16, n_playout=1600, batch=64)
        moves = par.get_moves(states)                  # list of (x, y) / None, one per state
        par.update_with_moves(moves)
    """

    def __init__(self, policy, value=None, games=16, **kw):
        self.searches = [BatchedMCTS(policy, value, **kw) for _ in range(int(games))]
        self._policy, self._value = policy, value

    def _wave(self, roots, size, count):
        L = _cabi.lib()
        S = self.searches[:len(roots)]
        G, dev = len(S), roots.device
        first = S[0]
        states = BatchedGameState(roots[:, None, :].expand(G, count, STATE_WORDS).reshape(G * count, STATE_WORDS).clone(), size)
        if first._forget_ages:
            states.forget_ages()
        for s in S:
            _cabi.check(L.rag_tree_wave_begin(s._tree, count, ctypes.c_double(s._vl)))
        rows, moves = np.empty(count, np.int32), np.empty(G * count, np.int32)
        for _ in range(first._L):
            need, per_tree = [], []
            for g, s in enumerate(S):
                k = _cabi_check_count(L.rag_tree_wave_leaves(s._tree, _p(rows)))
                per_tree.append(rows[:k].copy())
                need.append(rows[:k].astype(np.int64) + g * count)
            need = np.concatenate(need)
            if len(need):
                sub = BatchedGameState(states.packed[torch.as_tensor(need, device=dev)].contiguous(), size)
                _, flat, counts, probs = self._policy.eval_arrays(sub)
                flat, counts = np.ascontiguousarray(flat, np.int64), np.ascontiguousarray(counts, np.int64)
                probs = np.ascontiguousarray(probs, np.float32)
                ends = np.concatenate([[0], np.cumsum(counts)])
                at = 0
                for g, s in enumerate(S):
                    k = len(per_tree[g])
                    if k:
                        a, b = int(ends[at]), int(ends[at + k])
                        r = np.ascontiguousarray(per_tree[g])
                        c, f, pr = np.ascontiguousarray(counts[at:at + k]), np.ascontiguousarray(flat[a:b]), np.ascontiguousarray(probs[a:b])
                        _cabi.check(L.rag_tree_wave_expand(s._tree, k, _p(r), _p(c), _p(f), _p(pr)))
                        s.evaluations += k
                        at += k
            alive = 0
            for g, s in enumerate(S):
                mv = np.empty(count, np.int32)
                alive += _cabi_check_count(L.rag_tree_wave_step(s._tree, _p(mv)))
                moves[g * count:(g + 1) * count] = mv
            if alive == 0:
                break
            moving = torch.as_tensor(moves >= 0, device=dev)
            before = states.packed.clone()
            states.do_move(np.where(moves >= 0, moves, -1))
            states.packed = torch.where(moving[:, None], states.packed, before)
        v = np.zeros(G * count)
        if first._lmbda < 1.0:
            v = np.asarray(self._value.batch_eval_state(states), np.float64)
        if first._lmbda > 0.0:
            v = (1.0 - first._lmbda) * v + first._lmbda * first._rollouts(states)
        for g, s in enumerate(S):
            _cabi.check(L.rag_tree_wave_backup(s._tree, _p(np.ascontiguousarray(v[g * count:(g + 1) * count], np.float64))))

    def get_moves(self, states):
        """Runs n_playout playouts in every tree (waves of `batch` per tree) and returns the most visited move of each."""
        batch = states if isinstance(states, BatchedGameState) else BatchedGameState.from_states(states)
        G = len(batch)
        if G > len(self.searches):
            raise ValueError("%d positions for %d trees" % (G, len(self.searches)))
        first = self.searches[0]
        done = 0
        while done < first._n_playout:
            count = min(first._batch, first._n_playout - done)
            self._wave(batch.packed, batch.size, count)
            done += count
        out = []
        for s in self.searches[:G]:
            s._size = batch.size
            root = s._root
            if root.is_leaf():
                out.append(None)
            else:
                a = int(root.actions[int(np.argmax(root.n))])
                out.append((a // batch.size, a % batch.size))
        return out

    def update_with_moves(self, moves, size=None):
        for s, m in zip(self.searches, moves):
            s.update_with_move(m, size=size)


def _network_of(fn, what):
    """The CNNPolicy / CNNValue behind a reference-style callable (the object itself, or its bound
    eval_state / __call__)."""
    net = getattr(fn, "__self__", fn)
    if not hasattr(net, "model") or not hasattr(net, "preprocessor"):
        raise TypeError("%s must be a network of this package (or its eval_state method): the device search cannot "
                        "call back into arbitrary Python functions" % what)
    return net


class MCTS(BatchedMCTS):
    """AlphaGo/mcts.py:109-217 under its own name and constructor:
        MCTS(value_fn, policy_fn, rollout_policy_fn, lmbda=0.5, c_puct=5, rollout_limit=500, playout_depth=20, n_playout=10000)
    `value_fn` / `policy_fn` are this package's CNNValue / CNNPolicy (or their bound `eval_state`);
    `rollout_policy_fn` is accepted for the signature and not called -- rollouts (lmbda > 0) are uniformly
    random sensible-move playouts on the device.  Playouts run one at a time, as in the reference
    (`batch=1`: the same visit counts, playout for playout); pass batch=64 or more for throughput."""

    def __init__(self, value_fn, policy_fn, rollout_policy_fn=None, lmbda=0.5, c_puct=5, rollout_limit=500,
                 playout_depth=20, n_playout=10000, batch=1, **kw):
        value = None if value_fn is None else _network_of(value_fn, "value_fn")
        BatchedMCTS.__init__(self, _network_of(policy_fn, "policy_fn"), value, lmbda=lmbda, c_puct=c_puct,
                             rollout_limit=rollout_limit, playout_depth=playout_depth, n_playout=n_playout,
                             batch=batch, **kw)
        self._rollout_policy_fn = rollout_policy_fn
