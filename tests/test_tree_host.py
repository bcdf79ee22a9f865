"""The native search tree (csrc/tree.cpp behind rag_tree_*) against a pure-Python restatement of the reference's
TreeNode / _playout rules (AlphaGo/mcts.py:11-92, 137-169, 208-217) -- host code only, no GPU: with one playout
per wave every count and every float must be identical; with many playouts per wave the tree must stay consistent."""
import ctypes
import zlib

import numpy as np

from rocalphago_b200 import _cabi


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _seed(path, salt):
    return zlib.crc32(np.asarray(list(path) + [salt], np.int64).tobytes()) & 0x7FFFFFFF


def _policy(path):
    """A deterministic stand-in for the policy network: children and priors as a function of the moves so far."""
    r = np.random.RandomState(_seed(path, 1))
    k = 0 if (len(path) >= 2 and r.rand() < 0.08) else int(r.randint(2, 9))        # now and then: no legal move
    pr = r.rand(k).astype(np.float32)
    return [int(a) for a in r.permutation(50)[:k]], (pr / pr.sum() if k else pr)


def _value(path):
    return float(np.random.RandomState(_seed(path, 2)).uniform(-1, 1))


class _Node(object):
    def __init__(self, parent, index):
        self.parent, self.index, self.kids, self.visits, self.q = parent, index, None, 0, 0.0


def _python_playout(root, depth, c_puct):
    node, path, acts = root, [], []
    for _ in range(depth):
        if node.kids is None:
            a, pr = _policy(acts)
            if len(a):
                node.acts, node.P = a, np.asarray(pr, np.float64)
                node.u, node.Q, node.n = node.P.copy(), np.zeros(len(a)), np.zeros(len(a), np.int64)
                node.kids = [None] * len(a)
        if node.kids is None:
            break                                                    # end of game (mcts.py:154-156)
        i = int(np.argmax(node.Q + node.u))
        path.append((node, i))
        acts.append(node.acts[i])
        if node.kids[i] is None:
            node.kids[i] = _Node(node, i)
        node = node.kids[i]
    v = _value(acts)
    root.visits += 1
    root.q += (v - root.q) / root.visits
    for nd, i in path:
        nd.n[i] += 1
        nd.Q[i] += (v - nd.Q[i]) / nd.n[i]
        pv = nd.visits if nd.parent is None else nd.parent.n[nd.index]
        nd.u[i] = c_puct * nd.P[i] * np.sqrt(pv) / (1 + nd.n[i])


def _native_wave(L, h, count, depth, vl, prefix=()):
    rows, mv = np.empty(count, np.int32), np.empty(count, np.int32)
    L.rag_tree_wave_begin(h, count, ctypes.c_double(vl))
    paths = [list(prefix) for _ in range(count)]
    for _ in range(depth):
        k = L.rag_tree_wave_leaves(h, _p(rows))
        if k:
            counts, moves, priors = [], [], []
            for j in rows[:k]:
                a, pr = _policy(paths[j])
                counts.append(len(a)); moves += a; priors += pr.tolist()
            L.rag_tree_wave_expand(h, k, _p(rows), _p(np.asarray(counts, np.int64)), _p(np.asarray(moves, np.int64)),
                                   _p(np.asarray(priors, np.float32)))
        if L.rag_tree_wave_step(h, _p(mv)) == 0:
            break
        for j in range(count):
            if mv[j] >= 0:
                paths[j].append(int(mv[j]))
    L.rag_tree_wave_backup(h, _p(np.asarray([_value(pp) for pp in paths], np.float64)))


def _arrays(L, h, node):
    k = L.rag_tree_node_children(h, node)
    out = (np.empty(k, np.int32), np.empty(k), np.empty(k), np.empty(k), np.empty(k, np.int64), np.empty(k, np.int64), np.empty(k, np.int32))
    assert L.rag_tree_node_arrays(h, node, *[_p(a) for a in out]) == 0
    return out


def _compare(L, h, node, py):
    acts, P, Q, u, n, virt, child = _arrays(L, h, node)
    if py.kids is None:
        assert len(acts) == 0
        return 1
    assert acts.tolist() == py.acts and np.array_equal(P, py.P) and np.array_equal(n, py.n)
    assert np.array_equal(Q, py.Q) and np.array_equal(u, py.u) and (virt == 0).all()          # bit for bit
    seen = 1
    for i, kid in enumerate(py.kids):
        assert (child[i] >= 0) == (kid is not None)
        if kid is not None:
            seen += _compare(L, h, int(child[i]), kid)
    return seen


def test_one_playout_per_wave_is_the_references_search_bit_for_bit():
    L = _cabi.lib()
    for c_puct, depth, playouts in ((5.0, 4, 400), (1.0, 7, 300), (0.3, 3, 200)):
        h = ctypes.c_void_p()
        assert L.rag_tree_create(ctypes.c_double(c_puct), ctypes.byref(h)) == 0
        root = _Node(None, 0)
        for _ in range(playouts):
            _python_playout(root, depth, c_puct)
            _native_wave(L, h, 1, depth, 1.0)
        assert _compare(L, h, L.rag_tree_root(h), root) >= 4
        visits, q = ctypes.c_longlong(), ctypes.c_double()
        L.rag_tree_node_stats(h, L.rag_tree_root(h), ctypes.byref(visits), ctypes.byref(q), None, None)
        assert visits.value == root.visits == playouts and q.value == root.q
        # update_with_move: the most visited child becomes the root and keeps its subtree and counters
        best = int(np.argmax(root.n))
        before = L.rag_tree_size(h)
        assert L.rag_tree_advance(h, root.acts[best]) == 1 and L.rag_tree_size(h) < before
        kid = root.kids[best]
        _compare(L, h, L.rag_tree_root(h), kid)
        L.rag_tree_node_stats(h, L.rag_tree_root(h), ctypes.byref(visits), ctypes.byref(q), None, None)
        assert visits.value == root.n[best] and q.value == root.Q[best]
        assert L.rag_tree_advance(h, 12345) == 0 and L.rag_tree_size(h) == 1               # unknown move: a new tree
        L.rag_tree_destroy(h)


def test_waves_of_many_playouts_keep_the_tree_consistent():
    L = _cabi.lib()
    h = ctypes.c_void_p()
    L.rag_tree_create(ctypes.c_double(2.0), ctypes.byref(h))
    for _ in range(20):
        _native_wave(L, h, 64, 6, 1.0)
    root = L.rag_tree_root(h)
    visits = ctypes.c_longlong()
    L.rag_tree_node_stats(h, root, ctypes.byref(visits), None, None, None)
    assert visits.value == 20 * 64

    def check(node, budget):
        acts, P, Q, u, n, virt, child = _arrays(L, h, node)
        assert (virt == 0).all() and (n >= 0).all() and int(n.sum()) <= budget
        assert len(set(acts.tolist())) == len(acts)
        for i, c in enumerate(child):
            if c >= 0:
                check(int(c), int(n[i]))
    check(root, 20 * 64)
    acts, P, Q, u, n, virt, child = _arrays(L, h, root)
    assert int(n.sum()) == 20 * 64 and (n > 0).sum() >= 2                  # virtual visits spread a wave over the children
    L.rag_tree_destroy(h)
