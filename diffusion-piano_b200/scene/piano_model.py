"""The 88-key piano as plain tables (exact: every number is in the reference tree).

Restates what `robopianist/models/piano/piano_mjcf.py:25-402` builds from
`piano_constants.py:22-85`: one box body + hinge per key, sorted by key number, plus
the static base box. Derived, not copied: the key layout is generated from the
octave pattern instead of the reference's index lists (asserted equal in tests).
"""

from __future__ import annotations

from dataclasses import dataclass
from math import atan, pi
from typing import List

NUM_KEYS = 88
NUM_WHITE_KEYS = 52
WHITE_KEY_WIDTH = 0.0225
WHITE_KEY_LENGTH = 0.15
WHITE_KEY_HEIGHT = WHITE_KEY_WIDTH
SPACING = 0.001
BLACK_KEY_WIDTH = 0.01
BLACK_KEY_LENGTH = 0.09
BLACK_KEY_HEIGHT = 0.018
PIANO_LENGTH = NUM_WHITE_KEYS * WHITE_KEY_WIDTH + (NUM_WHITE_KEYS - 1) * SPACING  # 1.221

WHITE_X = 0.0
WHITE_Z = WHITE_KEY_HEIGHT / 2
BLACK_X = -WHITE_KEY_LENGTH / 2 + BLACK_KEY_LENGTH / 2  # -0.03
BLACK_Z = WHITE_KEY_HEIGHT + 0.0125 - BLACK_KEY_HEIGHT / 2  # 0.026

BASE_SIZE = (0.1 / 2, PIANO_LENGTH / 2, 0.04 / 2)
BASE_POS = (-WHITE_KEY_LENGTH / 2 - 0.5 * 0.1 - 0.002, 0.0, 0.04 / 2)

WHITE_MAX_ANGLE = atan(0.01 / WHITE_KEY_LENGTH)
BLACK_MAX_ANGLE = atan(0.008 / BLACK_KEY_LENGTH)
WHITE_MASS, BLACK_MASS = 0.04, 0.02
SPRINGREF = -1 * pi / 180
STIFFNESS = 2.0
DAMPING = 0.05
ARMATURE = 0.001

KEY_THRESHOLD = 0.00872665  # piano.py:31
SUSTAIN_THRESHOLD = 0.5  # piano.py:32

_BLACK_IN_OCTAVE = (0, 1, 0, 0, 1, 0, 1, 0, 0, 1, 0, 1)  # key % 12, A0 = key 0


def is_key_black(key: int) -> bool:
    return bool(_BLACK_IN_OCTAVE[key % 12])


@dataclass(frozen=True)
class Key:
    number: int
    black: bool
    pos: tuple  # body position (world; piano is attached at the origin)
    half_size: tuple  # box half extents
    mass: float
    joint_pos: tuple  # hinge anchor in body frame
    max_angle: float


def build_keys() -> List[Key]:
    keys: List[Key] = []
    pitch = WHITE_KEY_WIDTH + SPACING
    white_index = 0
    for k in range(NUM_KEYS):
        if not is_key_black(k):
            y = -PIANO_LENGTH * 0.5 + WHITE_KEY_WIDTH * 0.5 + white_index * pitch
            keys.append(Key(k, False, (WHITE_X, y, WHITE_Z),
                            (WHITE_KEY_LENGTH / 2, WHITE_KEY_WIDTH / 2, WHITE_KEY_HEIGHT / 2),
                            WHITE_MASS, (-WHITE_KEY_LENGTH / 2, 0.0, 0.0), WHITE_MAX_ANGLE))
            white_index += 1
        else:
            # a black key sits on the boundary after the `white_index`-th white key.
            # piano_mjcf.py:249-251 (lone key), :290-295 (twins), :339-344 (triplets)
            if k == 1:
                y = WHITE_KEY_WIDTH + 0.5 * (-PIANO_LENGTH + SPACING)
            else:
                y = -PIANO_LENGTH * 0.5 + white_index * pitch
            keys.append(Key(k, True, (BLACK_X, y, BLACK_Z),
                            (BLACK_KEY_LENGTH / 2, BLACK_KEY_WIDTH / 2, BLACK_KEY_HEIGHT / 2),
                            BLACK_MASS, (-BLACK_KEY_LENGTH / 2, 0.0, 0.0), BLACK_MAX_ANGLE))
    return keys
