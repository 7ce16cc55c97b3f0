"""cgnn_b200 - B200-native engine for the CGNN hot path."""
from .utils.Settings import SETTINGS
from .engine import Engine, EngineError

__all__ = ["SETTINGS", "Engine", "EngineError"]
