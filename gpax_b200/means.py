"""Mean functions with the reference's surface (``gpax.means``: ``Mean``,
``ConstantMean``; gpax/means.py:28-78).  A mean function is user Python code, so it
runs on the host: the device path receives its values, per hyper-parameter sample,
as ``m_train[S, n]`` (folded into the residual ``y - m``) and ``m_new[S, M]``
(``gpx_predict`` / ``gpx_predict_acq``), or as one constant.

Learnable parameters are ``Parameter`` attributes whose names start with ``m_``
(gpax/means.py:8-13, examples/04_Structured_GP.ipynb cell 10): they are sampled
by ``fit`` together with the kernel hyper-parameters, the marginal-likelihood
gradient reaching them through ``d MLL / d m = alpha`` (``gpx_mll``) and the
Jacobian of the mean (``jacobian`` below).
"""
from __future__ import annotations

import numpy as np

from .kernels import Prior


class Mean(Prior):
    """Base class; subclasses define ``_mean_function(x, **params) -> [N]`` where
    ``params`` holds one value per ``m_*`` parameter (a scalar, or ``[plate]``)."""

    attribute_prefix = "m_"
    #: relative step of the central differences used when ``_mean_jacobian`` is not defined
    fd_step = 1.0e-6

    def _mean_function(self, x, **params):
        raise NotImplementedError

    def _mean_jacobian(self, x, **params):
        """Optional analytic Jacobian: ``{name: d mean / d name}`` with entries ``[N]``
        (``[N, plate]`` for plated parameters).  Return ``None`` to use central differences."""
        return None

    def sample_prior(self, x, rng=None):
        p = {k: v[0] for k, v in self.sample_parameters(rng, 1).items()}
        return self.__call__(x, **p)

    def __call__(self, x, **params):
        # the mean function should always be 1-d (gpax/means.py:61-63)
        return np.squeeze(self._mean_function(np.asarray(x, float), **params))

    # -- batched evaluation over hyper-parameter samples -----------------------
    def learnable(self):
        """Names of the ``m_*`` parameters that are distributions (sample sites)."""
        return [k for k, p in self._params.items() if not p.is_constant]

    def _params_at(self, hp, s):
        """Per-sample keyword arguments: sampled ``m_*`` values of sample ``s``; constants
        at their value (the reference merges ``mean.params`` under the samples, gpax/gp.py:328)."""
        out = {}
        for k, p in self._params.items():
            if k in hp:
                v = np.asarray(hp[k], float)
                v = v.reshape(-1)[0] if v.size == 1 else v[s]
                out[k] = float(v) if np.ndim(v) == 0 else np.asarray(v, float)
            elif p.is_constant:
                out[k] = float(p.value)
            else:
                raise ValueError(f"no value for the mean parameter {k!r}: fit() the model or pass it in hp")
        return out

    def evaluate(self, x, hp, S):
        """``[S, N]`` values (one row per sample), or ``[1, N]`` when the mean has no
        sampled parameter (every sample shares it)."""
        x = np.asarray(x, float)
        N = x.shape[0]
        if not any(k in hp for k in self._params):
            v = np.broadcast_to(np.asarray(self.__call__(x, **self._params_at(hp, 0)), float), (N,))
            return v.reshape(1, N).copy()
        out = np.empty((S, N))
        for s in range(S):
            out[s] = np.broadcast_to(np.asarray(self.__call__(x, **self._params_at(hp, s)), float), (N,))
        return out

    def jacobian(self, x, **params):
        """``{name: [N, plate]}`` for the learnable parameters at one parameter point."""
        x = np.asarray(x, float)
        N = x.shape[0]
        names = self.learnable()
        an = self._mean_jacobian(x, **params)
        if an is not None:
            return {k: np.asarray(an[k], float).reshape(N, -1) for k in names}
        out = {}
        for k in names:
            scalar = np.ndim(params[k]) == 0
            v0 = np.atleast_1d(np.asarray(params[k], float))
            J = np.empty((N, v0.size))
            for q in range(v0.size):
                h = self.fd_step * max(1.0, abs(v0[q]))
                hi, lo = v0.copy(), v0.copy()
                hi[q] += h
                lo[q] -= h
                f1 = np.broadcast_to(self.__call__(x, **{**params, k: float(hi[0]) if scalar else hi}), (N,))
                f0 = np.broadcast_to(self.__call__(x, **{**params, k: float(lo[0]) if scalar else lo}), (N,))
                J[:, q] = (f1 - f0) / (2.0 * h)
            out[k] = J
        return out


class ConstantMean(Mean):
    """Returns ``mean`` at every point (default 1.0, gpax/means.py:75-78)."""

    def __init__(self, mean=1.0):
        if not isinstance(mean, (int, float)):
            raise TypeError("mean must be a float or int")
        self.mean = mean

    def _mean_function(self, x):
        return np.ones(x.shape[0]) * self.mean
