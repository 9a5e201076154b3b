"""Host-side data transforms with the reference's surface
(``gpax.transforms``: ``fit`` / ``forward`` / ``reverse`` with
``transforms_as in {"mean","std"}``; gpax/transforms.py:79-226).

These are O(n*d) element-wise maps applied once on the host before the device
path runs; hyper-priors live in transformed space
(gpax/kernels.py:4-9), so they must agree with the reference including its
epsilon asymmetries (gpax/transforms.py:250-274, 301-317).
"""
from __future__ import annotations

import numpy as np

EPSILON = 1e-8          # gpax/transforms.py:58
T_LO, T_HI = -1.0, 1.0  # target interval of ScaleTransform


def _as_2d(a):
    a = np.asarray(a, dtype=np.float64)
    return a.reshape(-1, 1) if a.ndim == 1 else a


class Transform:
    """Base class: one-shot ``fit`` then ``forward``/``reverse``.  Same error
    behaviour as the reference: a second ``fit`` or a transform before ``fit``
    raises ``RuntimeError``; an unknown ``transforms_as`` raises ``ValueError``;
    ``None`` passes through."""

    def __init__(self):
        self._is_fit = False

    # subclasses implement these four maps
    def _fit(self, x):
        raise NotImplementedError

    def _apply(self, x, direction, kind):
        raise NotImplementedError

    def fit(self, x):
        if self._is_fit:
            raise RuntimeError("Cannot call fit on a Transform more than once")
        self._is_fit = True
        if x is not None:
            self._fit(_as_2d(x))

    def _run(self, x, direction, transforms_as):
        if not self._is_fit:
            raise RuntimeError("Cannot execute transform before fit is called")
        if x is None:
            return None
        if transforms_as not in ("mean", "std"):
            raise ValueError(f"arr has invalid transform type {transforms_as}")
        return self._apply(_as_2d(x), direction, transforms_as)

    def forward(self, x, transforms_as="mean"):
        return self._run(x, "forward", transforms_as)

    def reverse(self, x, transforms_as="mean"):
        return self._run(x, "reverse", transforms_as)


class IdentityTransform(Transform):
    def _fit(self, x):
        pass

    def _apply(self, x, direction, kind):
        return x


class ScaleTransform(Transform):
    """Per-column affine map of the fitted [min, max] onto [-1, 1]."""

    def __init__(self):
        super().__init__()
        self._minima = None
        self._maxima = None

    def _fit(self, x):
        self._minima = x.min(axis=0, keepdims=True)
        self._maxima = x.max(axis=0, keepdims=True)

    def _apply(self, x, direction, kind):
        span_r = self._maxima - self._minima + EPSILON
        if direction == "forward":
            span_t = T_HI - T_LO
            if kind == "mean":
                return (x - self._minima) / span_r * span_t + T_LO
            return x / span_r * span_t
        span_t = T_HI - T_LO + EPSILON
        if kind == "mean":
            return (x * span_r + self._minima * span_t - T_LO * span_r) / span_t
        return x * span_r / span_t


class NormalizeTransform(Transform):
    """Per-column z-scoring; note ``forward(std)`` divides by ``std`` with no
    epsilon while ``forward(mean)`` divides by ``std + eps``."""

    def __init__(self):
        super().__init__()
        self._mean = None
        self._std = None

    def _fit(self, x):
        self._mean = x.mean(axis=0, keepdims=True)
        self._std = x.std(axis=0, keepdims=True)

    def _apply(self, x, direction, kind):
        if direction == "forward":
            if kind == "mean":
                return (x - self._mean) / (self._std + EPSILON)
            return x / self._std
        if kind == "mean":
            return x * (self._std + EPSILON) + self._mean
        return x * self._std
