"""helm_b200 -- B200-native batched TFHE bootstrapping engine for HELM's gate/LUT evaluation hot path."""
from .engine import Engine, Wires, EngineError, make_gates, GATE_DTYPE, LIB_PATH  # noqa: F401
