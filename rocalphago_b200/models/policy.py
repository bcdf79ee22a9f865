"""CNNPolicy on the B200 path (drop-in for AlphaGo/models/policy.py:9-138)."""
import numpy as np

from ..preprocessing import _points
from ..state import BatchedGameState
from .nn_util import NeuralNetBase, Network, neuralnet


# What CNNPolicy computes in unless told otherwise: fp16 tensor-core operands with the 32 most energetic
# input channels of every layer after the first corrected by their fp16 remainders (include/
# rocalphago_b200.h RAG_DTYPE_FP16C32).  Plain 'fp16' is 1.3x faster but sits at north_star's 99.5 %
# top-1 gate on random-init weights (99.4-99.8 % by seed); this mode holds it (99.6-99.9 %).
DEFAULT_POLICY_DTYPE = "fp16c32"


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
        n = len(batch)
        probs = self.model.eval_batch_device(batch, self.preprocessor.plane_src)
        which, flat, counts = self.preprocessor.legal_move_indices(self.model.features_workspace(), n, batch.size)
        picked = probs.cpu().numpy()[which, flat]
        # per-state sums: states without a legal move (counts == 0) own no entry of `picked`
        sums = np.bincount(which, weights=picked, minlength=n).astype(picked.dtype)
        return which, flat, counts, picked / np.repeat(sums, counts)

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
            # all legal moves of every state: select and renormalise in one vectorised pass
            _, flat, counts, picked = self.eval_arrays(batch)
            picked = picked.tolist()
            points = _points(state_size)
            moves = [points[p] for p in flat.tolist()]
            ends = np.cumsum(counts).tolist()
            return [list(zip(moves[e - c:e], picked[e - c:e])) for e, c in zip(ends, counts.tolist())]
        probs = self.model.eval_batch_device(batch, self.preprocessor.plane_src)
        network_output = probs.cpu().numpy()
        results = [None] * n_states
        for i in range(n_states):
            results[i] = self._select_moves_and_normalize(network_output[i], moves_lists[i], state_size)
        return results

    def eval_state(self, state, moves=None):
        """Given a GameState object, returns a list of (action, probability) pairs
        according to the network outputs (policy.py:54-64)"""
        return self.batch_eval_state([state], [moves] if moves else None)[0]

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
