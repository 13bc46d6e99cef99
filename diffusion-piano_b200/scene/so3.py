"""Tiny fp64 quaternion helpers for the host-side model compiler (w, x, y, z order)."""

import numpy as np


def qnorm(q):
    q = np.asarray(q, np.float64)
    return q / np.linalg.norm(q)


def qmul(a, b):
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([
        aw * bw - ax * bx - ay * by - az * bz,
        aw * bx + ax * bw + ay * bz - az * by,
        aw * by - ax * bz + ay * bw + az * bx,
        aw * bz + ax * by - ay * bx + az * bw,
    ])


def qmat(q):
    w, x, y, z = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)],
    ])


def qrot(q, v):
    return qmat(q) @ np.asarray(v, np.float64)


def axis_angle(axis, angle):
    axis = np.asarray(axis, np.float64)
    s = np.sin(angle / 2)
    return np.array([np.cos(angle / 2), *(axis * s)])


def mirror_vec(v):
    return np.array([-v[0], v[1], v[2]], np.float64)


def mirror_quat(q):
    return np.array([q[0], q[1], -q[2], -q[3]], np.float64)


def mirror_axis(a):
    return np.array([a[0], -a[1], -a[2]], np.float64)
