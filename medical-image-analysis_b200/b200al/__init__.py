"""b200al - B200-native active-learning acquisition (drop-in for the
`src/activelearning` selectors of trnKhanh/medical-image-analysis).

Everything that computes runs in libb200al.so (hand-written sm_100a CUDA behind
the C ABI in include/b200al.h).  Importing this package does not need a GPU;
creating an Engine or calling a selector does, and fails loudly otherwise.
"""
from . import _lib, dist  # noqa: F401
from .engine import Engine, conv_flops_per_sample, validate_state_dict  # noqa: F401
from .selectors import (  # noqa: F401
    SELECTORS,
    ActiveSelector,
    B200BADGESelector,
    B200ConfidenceSelector,
    B200CoresetSelector,
    B200DiceUncertaintySelector,
    B200EntropySelector,
    B200HybridSelector,
    B200KMeanSelector,
    B200MarginSelector,
    B200MCDropoutEntropySelector,
    PoolSource,
)
from .integration import install_into_al_trainer  # noqa: F401
