"""B200-native batched position evaluation for RocAlphaGo.

Drop-in for `AlphaGo/preprocessing/preprocessing.py` and `AlphaGo/models/{policy,value,nn_util}.py`
of hiropppe/RocAlphaGo: same class and function names, same feature and move ordering, with
the work done by hand-written sm_100a CUDA kernels behind a C ABI (include/rocalphago_b200.h).
"""
from .state import BatchedGameState, PositionView  # noqa: F401
from .preprocessing import Preprocess, FEATURES, DEFAULT_FEATURES  # noqa: F401
