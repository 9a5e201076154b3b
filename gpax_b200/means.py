"""Mean functions with the reference's surface (``gpax.means``: ``Mean``,
``ConstantMean``; gpax/means.py:35-78).  Host-side O(n) work: the device path
receives the mean either as a per-sample constant (``m_const`` of
``gpx_predict``) or as a precomputed ``m_new`` array.
"""
from __future__ import annotations

import numpy as np

from .kernels import Prior


class Mean(Prior):
    """Base class; subclasses define ``_mean_function(x, **params) -> [N]``.
    Learnable parameters are ``Parameter`` attributes prefixed ``m_``."""

    attribute_prefix = "m_"

    def _mean_function(self, x, **params):
        raise NotImplementedError

    def sample_prior(self, x, rng=None):
        p = {k: v[0] for k, v in self.sample_parameters(rng, 1).items()}
        return self.__call__(x, **p)

    def __call__(self, x, **params):
        return np.squeeze(self._mean_function(np.asarray(x, float), **params))


class ConstantMean(Mean):
    """Returns ``mean`` at every point (default 1.0, gpax/means.py:75-78)."""

    def __init__(self, mean=1.0):
        if not isinstance(mean, (int, float)):
            raise TypeError("mean must be a float or int")
        self.mean = mean

    def _mean_function(self, x):
        return np.ones(x.shape[0]) * self.mean
