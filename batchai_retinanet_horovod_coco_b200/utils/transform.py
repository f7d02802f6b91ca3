"""Random affine transforms (keras-retinanet ``utils/transform.py``; bound at reference
``train.py:54,180-193``).  3x3 homogeneous matrices composed from rotation,
translation, shear, scaling and flips; ranges come from the caller."""

from __future__ import annotations

import numpy as np

DEFAULT_PRNG = np.random


def colvec(*args):
    return np.array([args]).T


def transform_aabb(transform, aabb):
    """Axis-aligned bounding box of the transformed box (x1, y1, x2, y2)."""
    x1, y1, x2, y2 = aabb
    points = transform.dot([[x1, x2, x1, x2], [y1, y2, y2, y1], [1, 1, 1, 1]])
    mn, mx = points.min(axis=1), points.max(axis=1)
    return [mn[0], mn[1], mx[0], mx[1]]


def _random_vector(mn, mx, prng=DEFAULT_PRNG):
    mn, mx = np.array(mn), np.array(mx)
    assert mn.shape == mx.shape and len(mn.shape) == 1
    return prng.uniform(mn, mx)


def rotation(angle):
    return np.array([[np.cos(angle), -np.sin(angle), 0], [np.sin(angle), np.cos(angle), 0], [0, 0, 1]])


def random_rotation(mn, mx, prng=DEFAULT_PRNG):
    return rotation(prng.uniform(mn, mx))


def translation(t):
    return np.array([[1, 0, t[0]], [0, 1, t[1]], [0, 0, 1]])


def random_translation(mn, mx, prng=DEFAULT_PRNG):
    return translation(_random_vector(mn, mx, prng))


def shear(angle):
    return np.array([[1, -np.sin(angle), 0], [0, np.cos(angle), 0], [0, 0, 1]])


def random_shear(mn, mx, prng=DEFAULT_PRNG):
    return shear(prng.uniform(mn, mx))


def scaling(factor):
    return np.array([[factor[0], 0, 0], [0, factor[1], 0], [0, 0, 1]])


def random_scaling(mn, mx, prng=DEFAULT_PRNG):
    return scaling(_random_vector(mn, mx, prng))


def random_flip(flip_x_chance, flip_y_chance, prng=DEFAULT_PRNG):
    fx = prng.uniform(0, 1) < flip_x_chance
    fy = prng.uniform(0, 1) < flip_y_chance
    return scaling((1 - 2 * fx, 1 - 2 * fy))  # True -> -1


def change_transform_origin(transform, center):
    center = np.array(center)
    return np.linalg.multi_dot([translation(center), transform, translation(-center)])


def random_transform(min_rotation=0, max_rotation=0, min_translation=(0, 0), max_translation=(0, 0), min_shear=0,
                     max_shear=0, min_scaling=(1, 1), max_scaling=(1, 1), flip_x_chance=0, flip_y_chance=0,
                     prng=DEFAULT_PRNG):
    return np.linalg.multi_dot([
        random_rotation(min_rotation, max_rotation, prng),
        random_translation(min_translation, max_translation, prng),
        random_shear(min_shear, max_shear, prng),
        random_scaling(min_scaling, max_scaling, prng),
        random_flip(flip_x_chance, flip_y_chance, prng)])


def random_transform_generator(prng=None, **kwargs):
    """Infinite generator of random transforms; a private ``RandomState`` when ``prng`` is None."""
    if prng is None:
        prng = np.random.RandomState()
    while True:
        yield random_transform(prng=prng, **kwargs)
