"""CNNPolicy on the B200 path (drop-in for AlphaGo/models/policy.py:9-138)."""
import collections.abc
import ctypes

import numpy as np
import torch

from .. import _cabi
from ..preprocessing import _points, check_workspace
from ..state import BatchedGameState, _ptr, _stream
from .nn_util import NeuralNetBase, Network, neuralnet


# What CNNPolicy computes in unless told otherwise: fp16 tensor-core operands with the 32 most energetic
# input channels of every layer after the first corrected by their fp16 remainders (include/
# rocalphago_b200.h RAG_DTYPE_FP16C32).  Plain 'fp16' is 1.3x faster but sits at north_star's 99.5 %
# top-1 gate on random-init weights (99.4-99.8 % by seed); this mode holds it (99.6-99.9 %).
DEFAULT_POLICY_DTYPE = "fp16c32"


class MoveList(collections.abc.Sequence):
    """One state's [((x, y), probability), ...] (policy.py:15-25), built from the batch's arrays the first
    time it is looked at.  Behaves like the reference's list for what its callers do with it -- len(),
    indexing, slicing, iteration, zip(*dist) (ai.py:118-124), == with a list -- without a million tuples
    being made for a 4,096-state batch nobody reads in full."""
    __slots__ = ("_dist", "_i", "_items")

    def __init__(self, dist, i):
        self._dist, self._i, self._items = dist, i, None

    def to_list(self):
        if self._items is None:
            d, i = self._dist, self._i
            c = int(d.counts[i])
            pts = d.points
            self._items = list(zip([pts[p] for p in d.moves[i, :c].tolist()], d.probs[i, :c].tolist()))
        return self._items

    def __len__(self):
        return int(self._dist.counts[self._i])

    def __getitem__(self, k):
        return self.to_list()[k]

    def __iter__(self):
        return iter(self.to_list())

    def __eq__(self, other):
        if isinstance(other, MoveList):
            other = other.to_list()
        return self.to_list() == other if isinstance(other, list) else NotImplemented

    __hash__ = None

    def __repr__(self):
        return repr(self.to_list())


class MoveDistributions(object):
    """The arrays behind a batch_eval_state result: `.moves` int16 (n, S*S) flat points x*S+y in
    get_legal_moves() order, `.probs` float32 (n, S*S) renormalised over them, `.counts` int32 (n,);
    entries past a state's count are undefined.  `as_list()` gives the reference-shaped view."""

    def __init__(self, moves, probs, counts, size):
        self.moves, self.probs, self.counts, self.size = moves, probs, counts, int(size)
        self.points = _points(self.size)

    def __len__(self):
        return int(self.counts.shape[0])

    def as_list(self):
        return [MoveList(self, i) for i in range(len(self))]


def flatten_idx(position, size):
    (x, y) = position
    return x * size + y            # AlphaGo/util.py:11-13


@neuralnet
class CNNPolicy(NeuralNetBase):
    """uses a convolutional neural network to evaluate the state of the game
    and compute a probability distribution over the next action"""

    def _select_moves_and_normalize(self, nn_output, moves, size):
        """normalize a distribution over the given list of moves and return a list of
        (move, prob) tuples (policy.py:15-25; a list, as Python 2's zip gave)"""
        if len(moves) == 0:
            return []
        move_indices = [flatten_idx(m, size) for m in moves]
        distribution = nn_output[move_indices]
        distribution = distribution / distribution.sum()
        return list(zip(moves, distribution))

    def eval_arrays(self, batch):
        """batch_eval_state of a BatchedGameState as flat arrays: (state index, flat move x*S+y,
        moves per state, probability renormalised over each state's legal moves), moves in
        get_legal_moves() order.  What batched callers (players, tree search) consume."""
        dist = self.eval_distributions(batch)
        inside = np.arange(batch.size * batch.size)[None, :] < dist.counts[:, None]
        which = np.nonzero(inside)[0]
        return which, dist.moves[inside].astype(np.int64), dist.counts.astype(np.int64), dist.probs[inside]

    def eval_distributions(self, batch):
        """batch_eval_state of a BatchedGameState with the move lists left on the device path: features,
        forward, then rag_move_distributions (get_legal_moves() order and renormalisation per state) and
        one copy of three arrays to the host.  Returns a MoveDistributions."""
        n, P = len(batch), batch.size * batch.size
        probs = self.model.eval_batch_device(batch, self.preprocessor.plane_src)
        dev = batch.device
        moves = torch.empty((n, P), dtype=torch.int16, device=dev)
        out = torch.empty((n, P), dtype=torch.float32, device=dev)
        counts = torch.empty((n,), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            _cabi.check(_cabi.lib().rag_move_distributions(_ptr(probs), _ptr(self.model.features_workspace()), n, batch.size, 1,
                                                           _ptr(moves), _ptr(out), _ptr(counts), _stream()))
        check_workspace(self.model.features_workspace())
        return MoveDistributions(moves.cpu().numpy(), out.cpu().numpy(), counts.cpu().numpy(), batch.size)

    def batch_eval_state(self, states, moves_lists=None):
        """Given a list of states (or a BatchedGameState), evaluates them all at once.
        Returns a parallel list of move distributions as in eval_state (policy.py:27-52).

        States are packed once, then features, convolutions and softmax run on the device with
        no host round trip (rag_eval_batch); the default move lists -- get_legal_moves() order,
        non-eyes then own eyes (go.py:465-480) -- come from the same analysis."""
        n_states = len(states)
        if n_states == 0:
            return []
        if isinstance(states, BatchedGameState):
            batch, state_size = states, states.size
        else:
            state_size = states[0].size
            if not all([st.size == state_size for st in states]):
                raise ValueError("all states must have the same size")
            batch = BatchedGameState.from_states(states)
        if not moves_lists:
            # all legal moves of every state, selected and renormalised on the device; a list with one
            # MoveList per state, whose tuples are made when the caller looks at that state
            return self.eval_distributions(batch).as_list()
        probs = self.model.eval_batch_device(batch, self.preprocessor.plane_src)
        network_output = probs.cpu().numpy()
        results = [None] * n_states
        for i in range(n_states):
            results[i] = self._select_moves_and_normalize(network_output[i], moves_lists[i], state_size)
        return results

    def eval_state(self, state, moves=None):
        """Given a GameState object, returns a list of (action, probability) pairs
        according to the network outputs (policy.py:54-64)"""
        dist = self.batch_eval_state([state], [moves] if moves else None)[0]
        return dist.to_list() if isinstance(dist, MoveList) else dist

    @staticmethod
    def create_network(**kwargs):
        """construct a convolutional neural network (policy.py:66-138).

        Keyword Arguments (same names as the reference):
        - input_dim, board (19), filters_per_layer (128), filters_per_layer_K, layers (12),
          filter_width_K (3; first layer 5)
        and, for this implementation: dtype -- 'fp16c32' (default: fp16 operands, the 32 most energetic
        input channels of layers 2.. corrected by their fp16 remainders), 'fp16c64', 'fp16', 'bf16', or
        'fp16x3' (every operand split into hi + lo parts: fp32-like results at a third of fp16's speed),
        kernel_flip (True = Theano convolution semantics), seed (None = fresh random weights).
        """
        defaults = {
            "board": 19,
            "filters_per_layer": 128,
            "layers": 12,
            "filter_width_1": 5
        }
        params = defaults
        params.update(kwargs)
        specs = []
        for i in range(1, params["layers"] + 1):
            width = params.get("filter_width_%d" % i, 3)
            nb = params.get("filters_per_layer_%d" % i, params["filters_per_layer"])
            specs.append((nb, width))
        return Network(params["board"], params["input_dim"], specs, head="policy",
                       dtype=params.get("dtype", DEFAULT_POLICY_DTYPE), kernel_flip=params.get("kernel_flip", True),
                       seed=params.get("seed"))


@neuralnet
class ResnetPolicy(CNNPolicy):
    """Residual network architecture as per He at al. 2015 (policy.py:141-271)"""

    @staticmethod
    def create_network(**kwargs):
        """Same keyword arguments as the reference: input_dim, board (19), filters_per_layer (128),
        layers (20), filter_width_K (3; filter_width_1 = 5), n_skip_K (1): the first convolution is
        linear; then units of n_skip_K x [BatchNormalization -> ReLU -> convolution] whose output is
        added to the unit's input; a last ReLU, the 1x1 convolution, Flatten, Bias, softmax.
        Faithful to two quirks of the reference: filter_width_K inside the units is looked up with
        the unit's own layer index, so the first unit's first convolution is again filter_width_1
        wide (policy.py:229-231 with K = 1); BatchNormalization() is built with Keras-1 defaults,
        i.e. it normalises over the LAST axis (the y coordinate), not over channels."""
        params = {"board": 19, "filters_per_layer": 128, "layers": 20, "filter_width_1": 5}
        params.update(kwargs)
        specs = [(params["filters_per_layer"], params["filter_width_1"])]
        unit_end = [1]
        layer = 1
        while layer < params["layers"]:
            n_skip = params.get("n_skip_%d" % layer, 1)
            for i in range(n_skip):
                specs.append((params["filters_per_layer"], params.get("filter_width_%d" % (layer + i), 3)))
                unit_end.append(1 if i == n_skip - 1 else 0)
            layer += n_skip
        return Network(params["board"], params["input_dim"], specs, head="policy", dtype=params.get("dtype", "fp16"),
                       kernel_flip=params.get("kernel_flip", True), seed=params.get("seed"), unit_end=unit_end)
