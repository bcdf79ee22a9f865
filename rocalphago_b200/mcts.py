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

Rollouts (lmbda > 0) use uniformly random sensible moves to the end of the game on the device
(`rollout_limit` plies) and area scoring as go.py:482-506; the reference's greedy-policy rollout
(one network call per ply) is not offered.
"""
import numpy as np
import torch

from .state import BatchedGameState, STATE_WORDS


class TreeNode(object):
    """A node and its children's statistics as arrays (mcts.py:11-92 keeps one object per child)."""

    __slots__ = ("parent", "index", "actions", "P", "Q", "u", "n", "virtual", "children", "visits", "q")

    def __init__(self, parent, index):
        self.parent, self.index = parent, index          # `index`: which child of `parent` this node is
        self.actions = None                              # flat moves x*S+y of the children, None while a leaf
        self.P = self.Q = self.u = self.n = self.virtual = None
        self.children = None                             # TreeNode or None per child, made when first walked into
        self.visits, self.q = 0, 0.0                     # the root's own counters (a child's live in its parent's arrays)

    def is_leaf(self):
        return self.actions is None

    def is_root(self):
        return self.parent is None

    def expand(self, actions, priors):
        """mcts.py:26-38: one child per (action, prior); u starts as the prior (mcts.py:22)."""
        k = len(actions)
        self.actions = np.asarray(actions, np.int32)
        self.P = np.asarray(priors, np.float64)
        self.u = self.P.copy()
        self.Q, self.n, self.virtual = np.zeros(k), np.zeros(k, np.int64), np.zeros(k, np.int64)
        self.children = [None] * k

    def select(self, virtual_loss):
        """mcts.py:40-46: the child of largest Q + u (the first one on ties, as max() over the
        children in insertion order gives); virtual visits of the wave in flight count against it."""
        score = self.Q + self.u
        if virtual_loss:
            score = score - virtual_loss * self.virtual
        return int(np.argmax(score))

    def child(self, i):
        if self.children[i] is None:
            self.children[i] = TreeNode(self, i)
        return self.children[i]

    def n_visits(self):
        return self.visits if self.parent is None else int(self.parent.n[self.index])


def _update_path(path, leaf_value, c_puct, root):
    """update_recursive (mcts.py:72-80) along `path` = [(node, child index), ...] from the root down."""
    root.visits += 1
    root.q += (leaf_value - root.q) / root.visits
    for node, i in path:
        node.n[i] += 1
        node.Q[i] += (leaf_value - node.Q[i]) / node.n[i]
        parent_visits = node.visits if node.parent is None else node.parent.n[node.index]
        node.u[i] = c_puct * node.P[i] * np.sqrt(parent_visits) / (1 + node.n[i])


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
        self._root = TreeNode(None, 0)
        self.evaluations = 0            # positions sent through the policy network (for reports)

    # ------------------------------------------------------------------ one wave of playouts
    def _wave(self, root_packed, size, count):
        dev = root_packed.device
        states = BatchedGameState(root_packed.expand(count, STATE_WORDS).clone(), size)       # a fresh copy even for count == 1
        if self._forget_ages:
            states.forget_ages()         # every playout starts from state.copy() (mcts.py:201), which drops stone_ages
        nodes = [self._root] * count
        paths = [[] for _ in range(count)]
        alive = np.ones(count, bool)
        vl = self._vl if count > 1 else 0.0
        for _ in range(self._L):
            # expand the leaves among the nodes the live playouts stand on -- each node once
            need = {}
            for j in range(count):
                if alive[j] and nodes[j].is_leaf():
                    need.setdefault(id(nodes[j]), j)
            if need:
                rows = sorted(need.values())
                sub = BatchedGameState(states.packed[torch.as_tensor(rows, device=dev)].contiguous(), size)
                _, flat, counts, probs = self._policy.eval_arrays(sub)
                self.evaluations += len(rows)
                ends = np.cumsum(counts)
                for k, j in enumerate(rows):
                    a, b = ends[k] - counts[k], ends[k]
                    if b > a:
                        nodes[j].expand(flat[a:b], probs[a:b])
            moves = np.full(count, -2, np.int32)               # -2: this playout does not move
            for j in range(count):
                if not alive[j]:
                    continue
                node = nodes[j]
                if node.is_leaf():                             # no legal move: end of game (mcts.py:154-156)
                    alive[j] = False
                    continue
                i = node.select(vl)
                node.virtual[i] += 1
                paths[j].append((node, i))
                moves[j] = node.actions[i]
                nodes[j] = node.child(i)
            if not alive.any():
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
        for j in range(count):
            for node, i in paths[j]:
                node.virtual[i] -= 1
            _update_path(paths[j], float(v[j]), self._c_puct, self._root)

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
        r = self._root
        if last_move is not None and not r.is_leaf():
            s = size if size is not None else None
            flat = last_move if np.isscalar(last_move) else None
            if flat is None:
                if s is None:
                    raise ValueError("pass size= with an (x, y) move")
                flat = last_move[0] * s + last_move[1]
            hit = np.flatnonzero(r.actions == flat)
            if len(hit):
                child = r.child(int(hit[0]))
                child.visits, child.q = int(r.n[hit[0]]), float(r.Q[hit[0]])
                child.parent = None
                self._root = child
                return
        self._root = TreeNode(None, 0)


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
