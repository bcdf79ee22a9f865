"""Value network on the B200 path.

The reference has only `value_trainer.model` (AlphaGo/models/value.py:12-28): 49 input planes,
12 convolutions of K=152 filters, a 1x1 convolution, Flatten, Dense(256) with NO activation and
Dense(1, tanh); no wrapper class and no 49th feature exist in that fork.
`CNNValue` gives that stack the same NeuralNetBase surface as CNNPolicy and defines the 49th
plane as `color` (all ones iff BLACK is to move), a definition of this build, not of the reference.
"""
from ..state import BatchedGameState
from ..preprocessing import DEFAULT_FEATURES
from .nn_util import NeuralNetBase, Network, neuralnet

K = 152                        # value.py:7
DEFAULT_VALUE_FEATURES = DEFAULT_FEATURES + ["color"]


@neuralnet
class CNNValue(NeuralNetBase):
    """state -> float in [-1, 1], the callable MCTS consumes as `value_fn` (mcts.py:112-113,164)"""

    def batch_eval_state(self, states):
        n_states = len(states)
        if n_states == 0:
            return []
        if isinstance(states, BatchedGameState):
            batch = states
        else:
            size = states[0].size
            if not all([st.size == size for st in states]):
                raise ValueError("all states must have the same size")
            batch = BatchedGameState.from_states(states)
        out = self.model.eval_batch_device(batch, self.preprocessor.plane_src)
        return [float(v) for v in out.cpu().numpy()]

    def eval_state(self, state):
        return self.batch_eval_state([state])[0]

    __call__ = eval_state

    @staticmethod
    def create_network(**kwargs):
        """value.py:14-28 with the reference's constants as defaults: board 19, 12 k x k layers
        (5x5 then 3x3) of `filters_per_layer` = 152 filters, Dense(256) linear, Dense(1) tanh."""
        params = {"board": 19, "filters_per_layer": K, "layers": 12, "filter_width_1": 5, "dense_units": 256}
        params.update(kwargs)
        specs = []
        for i in range(1, params["layers"] + 1):
            width = params.get("filter_width_%d" % i, 3)
            nb = params.get("filters_per_layer_%d" % i, params["filters_per_layer"])
            specs.append((nb, width))
        return Network(params["board"], params["input_dim"], specs, head="value",
                       dense_units=params["dense_units"], dtype=params.get("dtype", "fp16"),
                       kernel_flip=params.get("kernel_flip", True), seed=params.get("seed"))


class value_trainer(object):
    """Name kept from value.py:12: `.model` is the 49-plane, K=152 stack."""

    def __init__(self):
        self.model = CNNValue.create_network(input_dim=49)
