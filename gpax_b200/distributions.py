"""Minimal prior distributions for the hyper-parameter declarations.

The reference declares priors with ``numpyro.distributions`` objects
(gpax/kernels.py:69-75,101-107); NumPyro is a third-party dependency that is
absent here, so the handful of families the reference's defaults and notebooks
use are restated (density, sampler, mean, support).  Host-side and tiny: these
never touch the device path.
"""
from __future__ import annotations

import math

import numpy as np


class Distribution:
    """``support`` is one of "real", "positive", ("interval", lo, hi)."""
    support = "real"

    def sample(self, rng: np.random.Generator, shape=()):
        raise NotImplementedError

    def log_prob(self, x):
        raise NotImplementedError

    @property
    def mean(self):
        raise NotImplementedError

    # d log_prob / dx (for the HMC driver)
    def grad_log_prob(self, x):
        raise NotImplementedError


class Normal(Distribution):
    def __init__(self, loc=0.0, scale=1.0):
        self.loc, self.scale = float(loc), float(scale)

    def sample(self, rng, shape=()):
        return rng.normal(self.loc, self.scale, shape)

    def log_prob(self, x):
        z = (np.asarray(x) - self.loc) / self.scale
        return -0.5 * z * z - math.log(self.scale) - 0.5 * math.log(2 * math.pi)

    def grad_log_prob(self, x):
        return -(np.asarray(x) - self.loc) / self.scale**2

    @property
    def mean(self):
        return self.loc


class LogNormal(Distribution):
    support = "positive"

    def __init__(self, loc=0.0, scale=1.0):
        self.loc, self.scale = float(loc), float(scale)

    def sample(self, rng, shape=()):
        return rng.lognormal(self.loc, self.scale, shape)

    def log_prob(self, x):
        x = np.asarray(x)
        z = (np.log(x) - self.loc) / self.scale
        return -0.5 * z * z - np.log(x) - math.log(self.scale) - 0.5 * math.log(2 * math.pi)

    def grad_log_prob(self, x):
        x = np.asarray(x)
        return -((np.log(x) - self.loc) / self.scale**2 + 1.0) / x

    @property
    def mean(self):
        return math.exp(self.loc + 0.5 * self.scale**2)


class HalfNormal(Distribution):
    support = "positive"

    def __init__(self, scale=1.0):
        self.scale = float(scale)

    def sample(self, rng, shape=()):
        return np.abs(rng.normal(0.0, self.scale, shape))

    def log_prob(self, x):
        z = np.asarray(x) / self.scale
        return -0.5 * z * z + 0.5 * math.log(2 / math.pi) - math.log(self.scale)

    def grad_log_prob(self, x):
        return -np.asarray(x) / self.scale**2

    @property
    def mean(self):
        return self.scale * math.sqrt(2 / math.pi)


class Uniform(Distribution):
    def __init__(self, low=0.0, high=1.0):
        self.low, self.high = float(low), float(high)
        self.support = ("interval", self.low, self.high)

    def sample(self, rng, shape=()):
        return rng.uniform(self.low, self.high, shape)

    def log_prob(self, x):
        x = np.asarray(x)
        inside = (x >= self.low) & (x <= self.high)
        return np.where(inside, -math.log(self.high - self.low), -np.inf)

    def grad_log_prob(self, x):
        return np.zeros_like(np.asarray(x, float))

    @property
    def mean(self):
        return 0.5 * (self.low + self.high)


class Gamma(Distribution):
    support = "positive"

    def __init__(self, concentration, rate=1.0):
        self.a, self.b = float(concentration), float(rate)

    def sample(self, rng, shape=()):
        return rng.gamma(self.a, 1.0 / self.b, shape)

    def log_prob(self, x):
        x = np.asarray(x)
        return self.a * math.log(self.b) - math.lgamma(self.a) + (self.a - 1) * np.log(x) - self.b * x

    def grad_log_prob(self, x):
        return (self.a - 1) / np.asarray(x) - self.b

    @property
    def mean(self):
        return self.a / self.b
