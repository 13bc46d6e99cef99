"""Minimal dm_env look-alikes (dm_env itself is a third-party package the reference gets via
dm_control; it is not importable here). Same field names and semantics as dm_env.specs.Array /
BoundedArray and dm_env.TimeStep / StepType, which is all the reference's callers touch
(parallelized_base_v2.py:56-60,82-86; suite_test.py:37-52)."""

from __future__ import annotations

import enum
from typing import Any, NamedTuple

import numpy as np


class StepType(enum.IntEnum):
    FIRST = 0
    MID = 1
    LAST = 2


class TimeStep(NamedTuple):
    step_type: Any
    reward: Any
    discount: Any
    observation: Any

    def first(self) -> bool:
        return self.step_type == StepType.FIRST

    def mid(self) -> bool:
        return self.step_type == StepType.MID

    def last(self) -> bool:
        return self.step_type == StepType.LAST


class Array:
    def __init__(self, shape, dtype, name=None):
        self.shape = tuple(int(s) for s in shape)
        self.dtype = np.dtype(dtype)
        self.name = name

    def validate(self, value):
        value = np.asarray(value)
        if value.shape != self.shape:
            raise ValueError(f"Expected shape {self.shape} but found {value.shape}")
        if value.dtype != self.dtype:
            raise ValueError(f"Expected dtype {self.dtype} but found {value.dtype}")
        return value

    def generate_value(self):
        return np.zeros(self.shape, self.dtype)

    def __repr__(self):
        return f"Array(shape={self.shape}, dtype={self.dtype}, name={self.name!r})"


class BoundedArray(Array):
    def __init__(self, shape, dtype, minimum, maximum, name=None):
        super().__init__(shape, dtype, name)
        self.minimum = np.broadcast_to(np.asarray(minimum, self.dtype), self.shape).copy()
        self.maximum = np.broadcast_to(np.asarray(maximum, self.dtype), self.shape).copy()

    def validate(self, value):
        value = super().validate(value)
        if (value < self.minimum).any() or (value > self.maximum).any():
            raise ValueError("Values were not all within bounds")
        return value

    def generate_value(self):
        return self.minimum.copy()

    def __repr__(self):
        return f"BoundedArray(shape={self.shape}, dtype={self.dtype}, name={self.name!r})"
