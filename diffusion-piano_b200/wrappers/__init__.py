"""Environment wrappers (reference robopianist/wrappers/)."""
from .evaluation import MidiEvaluationWrapper

__all__ = ["MidiEvaluationWrapper"]
