from .nn_util import NeuralNetBase, Network, Bias, neuralnet  # noqa: F401
from .policy import CNNPolicy, ResnetPolicy  # noqa: F401
from .value import CNNValue, value_trainer  # noqa: F401
