"""Acquisition functions with the reference's surface (``gpax.acquisition``:
``AcquisitionFunction``, ``UpperConfidenceBound``, ``ExpectedImprovement``;
gpax/acquisition.py:40-310) on the device hot path.

For ``q == 1`` the analytic value and the optimiser's argmax come out of one
fused device pass (``gpx_predict`` -> ``gpx_acq_reduce`` -> ``gpx_argmax_finalize``)
over ALL candidates; the reference's ``batch_threshold`` chunking
(gpax/acquisition.py:203-209) exists only to bound its ``[M, M]`` covariance and
is accepted but unused on that path.  ``q > 1`` / ``force_monte_carlo`` follow
the reference's Monte-Carlo integrands over joint draws (row N1, host NumPy on
``model.sample(...).y`` chunked by ``batch_threshold``).
"""
from __future__ import annotations

import math
from typing import Callable

import numpy as np
from scipy.stats import qmc

from . import _cabi, state


def x_shape_err_msg(x, expected):
    return f"x has shape {x.shape} but expected shape {expected}"


class AcquisitionFunction:
    """Fields and validators of gpax/acquisition.py:41-64."""

    def __init__(self, bounds, q=1, penalty_function=None, penalty_strength=0.0,
                 batch_threshold=100, force_monte_carlo=False, fast=True):
        if not isinstance(bounds, (np.ndarray, list, tuple)):
            raise TypeError("bounds must be array-like of shape (2, d)")
        if not isinstance(q, int) or isinstance(q, bool):
            raise TypeError("q must be an int")
        if q < 1:
            raise ValueError("q must be >= 1")
        penalty_function = (lambda _: 0.0) if penalty_function is None else penalty_function
        if not isinstance(penalty_function, Callable):
            raise TypeError("penalty_function must be callable")
        if not isinstance(penalty_strength, (float, int)) or isinstance(penalty_strength, bool):
            raise TypeError("penalty_strength must be a float or int")
        if penalty_strength < 0.0:
            raise ValueError("penalty_strength must be >= 0")
        if not isinstance(batch_threshold, int) or isinstance(batch_threshold, bool):
            raise TypeError("batch_threshold must be an int")
        if batch_threshold <= 0:
            raise ValueError("batch_threshold must be > 0")
        if not isinstance(force_monte_carlo, bool) or not isinstance(fast, bool):
            raise TypeError("force_monte_carlo and fast must be bool")
        self.bounds = np.array(bounds, dtype=float)
        assert self.bounds.ndim == 2
        assert self.bounds.shape[0] == 2
        self.q, self.penalty_function, self.penalty_strength = q, penalty_function, penalty_strength
        self.batch_threshold, self.force_monte_carlo, self.fast = batch_threshold, force_monte_carlo, fast

    @property
    def name(self):
        return f"{self.__class__.__name__}(q={self.q})"

    # -- device (analytic, q == 1) -------------------------------------------
    def _device_spec(self, model):
        """``(acq id, beta, y_max)`` for ``gpx_acq_reduce``."""
        raise NotImplementedError(
            f"Acquisition function {self.__class__.__name__} does not have an "
            "analytic acquisition function defined")

    def _analytic(self, model, x):
        acq, beta, y_max = self._device_spec(model)
        return model._acquire(x, self.fast, acq, beta=beta, y_max=y_max)[2]

    def _analytic_evaluation(self, model, x):
        if self.q > 1:
            raise ValueError("Cannot evaluate analytic expression for q!=1")
        if x.ndim == 3 and x.shape[1] == 1:
            x = x.reshape(x.shape[0], x.shape[-1])
        if x.ndim != 2:
            raise ValueError(x_shape_err_msg(x, ("N", "d")))
        return self._analytic(model, x)

    # -- Monte Carlo (q > 1) ---------------------------------------------------
    def _get_integrand_samples(self, model, x):
        N, d = x.shape[0], x.shape[2]
        result = model.sample(x.reshape(-1, d), fast=self.fast)
        return result.y.reshape(-1, N, self.q)

    def _integrand(self, model, x):
        raise NotImplementedError

    def _monte_carlo_evaluation(self, model, x):
        if x.ndim == 2 and self.q == 1:
            x = x.reshape(x.shape[0], 1, x.shape[-1])
        if x.ndim != 3 or x.shape[1] != self.q:
            raise ValueError(x_shape_err_msg(x, ("N", "q", "d")))
        return self._integrand(model, x).mean(axis=0)

    def __call__(self, model, x):
        x = np.array(x, dtype=float)
        if self.force_monte_carlo or self.q > 1:
            return self._monte_carlo_evaluation(model, x)
        return self._analytic_evaluation(model, x)

    # -- optimiser -------------------------------------------------------------
    def _halton(self, n_samples):
        key, _ = state.get_rng_key()
        halton = qmc.Halton(d=self.bounds.shape[1] * self.q, seed=key)
        samples = halton.random(n=n_samples)
        l_bounds = np.tile(np.atleast_1d(self.bounds[0, :].squeeze()), self.q)
        u_bounds = np.tile(np.atleast_1d(self.bounds[1, :].squeeze()), self.q)
        if state.verbose > 1:
            print(f"qmc discrepancy (sample quality index) = {qmc.discrepancy(samples):.02e}")
        samples = qmc.scale(samples, l_bounds, u_bounds)
        return samples.reshape(n_samples, self.q, -1)

    def _bounds_tiled(self):
        l_bounds = np.tile(np.atleast_1d(self.bounds[0, :].squeeze()), self.q)
        u_bounds = np.tile(np.atleast_1d(self.bounds[1, :].squeeze()), self.q)
        return l_bounds, u_bounds

    def _optimize_halton_device(self, model, n_samples):
        """Analytic q == 1 path with the candidates generated ON the device (``gpx_halton``:
        SciPy's scrambled Halton points bit for bit, scaled to the bounds and pushed through the
        model's ScaleTransform in the same kernel) and swept by the fused ``gpx_predict_acq``;
        only the winner's coordinates and value return to the host.  Returns None when the model
        or its input transform needs the host path."""
        from . import engine as E
        from .transforms import EPSILON, IdentityTransform, ScaleTransform

        t = getattr(model, "input_transform", None)
        if model.x is None or not isinstance(t, (IdentityTransform, ScaleTransform)):
            return None
        key, _ = state.get_rng_key()
        halton = qmc.Halton(d=self.bounds.shape[1] * self.q, seed=key)      # draws the permutations
        if not hasattr(halton, "_permutations") or not hasattr(halton, "num_generated"):
            return None
        lb, ub = self._bounds_tiled()
        if isinstance(t, ScaleTransform):
            t_min, t_span = t._minima.reshape(-1), (t._maxima - t._minima + EPSILON).reshape(-1)
        else:
            t_min = t_span = None
        raw, xt = E.halton(halton, n_samples, lb, ub, t_min, t_span)
        if state.verbose > 1:
            unit = (raw.cpu().numpy() - lb) / (ub - lb)
            print(f"qmc discrepancy (sample quality index) = {qmc.discrepancy(unit):.02e}")
        acq, beta, y_max = self._device_spec(model)
        _, _, _, val, idx = model._acquire(None, self.fast, acq, beta=beta, y_max=y_max,
                                           want_arrays=False, x_new_t_dev=xt)
        return raw[idx].cpu().numpy().reshape(self.q, -1), val

    def _optimize_halton(self, model, n_samples):
        """gpax/acquisition.py:166-216: best of ``n_samples`` Halton points.

        Contract difference, flagged: the reference stacks the per-chunk values with
        ``jnp.array(vals)`` (gpax/acquisition.py:211), which only works when
        ``n_samples % batch_threshold == 0`` (ragged chunks raise there).  Here the analytic
        path evaluates all candidates in one fused device sweep, so any ``n_samples`` is
        accepted; for multiples of ``batch_threshold`` the result is the reference's (same
        Halton points, same first-maximum rule)."""
        if not (self.force_monte_carlo or self.q > 1) and hasattr(model, "_acquire"):
            saved = state.rng_key
            out = self._optimize_halton_device(model, n_samples)
            if out is not None:
                return out
            state.rng_key = saved          # host path below re-draws the same key
        samples = self._halton(n_samples)
        if self.force_monte_carlo or self.q > 1:
            vals = np.concatenate([
                np.atleast_1d(self.__call__(model, samples[i:i + self.batch_threshold]))
                for i in range(0, n_samples, self.batch_threshold)])
            best = int(np.argmax(vals))
            return samples[best, ...], float(vals[best])
        acq, beta, y_max = self._device_spec(model)
        _, _, _, val, idx = model._acquire(samples[:, 0, :], self.fast, acq, beta=beta,
                                           y_max=y_max, want_arrays=False)
        return samples[idx, ...], val

    def optimize(self, model, n, method="halton"):
        method = method.lower()
        if method == "halton":
            return self._optimize_halton(model, n)
        raise ValueError(f"Unknown optimizer {method}")


class UpperConfidenceBound(AcquisitionFunction):
    """``mu + sqrt(beta + 1e-12) * sd``; ``beta == inf`` -> ``sd``; ``beta == 0`` ->
    ``mu`` (gpax/acquisition.py:244-275)."""

    def __init__(self, bounds, q=1, penalty_function=None, penalty_strength=0.0,
                 batch_threshold=100, force_monte_carlo=False, fast=True, beta=10.0):
        # attrs field order of the reference: the base-class fields, then beta
        # (gpax/acquisition.py:44-54,253)
        super().__init__(bounds, q, penalty_function, penalty_strength, batch_threshold,
                         force_monte_carlo, fast)
        if not isinstance(beta, (int, float)) or isinstance(beta, bool):
            raise TypeError("beta must be an int or float")
        if beta < 0.0:
            raise ValueError("beta must be >= 0")
        self.beta = beta

    @property
    def name(self):
        return f"{self.__class__.__name__}(beta={self.beta:.02f}, q={self.q})"

    def _device_spec(self, model):
        if math.isinf(self.beta):
            return _cabi.ACQ_UE, 0.0, 0.0
        if self.beta == 0.0:
            return _cabi.ACQ_MEAN, 0.0, 0.0
        return _cabi.ACQ_UCB, float(self.beta), 0.0

    def _integrand(self, model, x):
        samples = self._get_integrand_samples(model, x)
        mean = samples.mean(axis=0, keepdims=True)
        if math.isinf(self.beta):
            return np.abs(samples - mean).max(axis=-1)
        beta_prime = np.sqrt(self.beta * np.pi / 2.0)
        return (mean + beta_prime * np.abs(samples - mean)).max(axis=-1)


class ExpectedImprovement(AcquisitionFunction):
    """``sd * pdf(u) + (mu - y_max) * cdf(u)``, ``u = (mu - y_max) / sd``
    (gpax/acquisition.py:278-310)."""

    def _get_y_max(self, model):
        if model.y is not None:
            return float(np.max(model.y))
        acqf = UpperConfidenceBound(beta=0.0, q=1, bounds=self.bounds, fast=True)
        x_star, _ = acqf.optimize(model, n=1000, method="Halton")
        y_max, _ = model.predict(x_star, fast=True)
        return float(np.squeeze(y_max))

    def _device_spec(self, model):
        return _cabi.ACQ_EI, 0.0, self._get_y_max(model)

    def _integrand(self, model, x):
        samples = self._get_integrand_samples(model, x)
        y_max = self._get_y_max(model)
        f_max_over_q = samples.max(axis=-1)
        return (f_max_over_q - y_max) * (f_max_over_q > 0)


class UncertaintyExploration(UpperConfidenceBound):
    """Pure exploration: UCB with ``beta = inf``."""

    def __init__(self, bounds, q=1, penalty_function=None, penalty_strength=0.0,
                 batch_threshold=100, force_monte_carlo=False, fast=True):
        super().__init__(bounds, q, penalty_function, penalty_strength, batch_threshold,
                         force_monte_carlo, fast, beta=float("inf"))


EI = ExpectedImprovement
UCB = UpperConfidenceBound
UE = UncertaintyExploration
