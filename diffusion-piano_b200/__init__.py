"""B200-native batched RoboPianist environment step."""
__version__ = "0.1.0"
