"""Covariance-function objects with the reference's surface
(``gpax.kernels``: ``Parameter``, ``Kernel``, ``ScaleKernel``, ``RBFKernel``,
``MaternKernel``, ``PeriodicKernel``; gpax/kernels.py:56-168 and
gpax/utils/prior_utils.py:95-185).

A kernel object is a declaration (family + hyper-priors); evaluating it runs
``gpx_gram`` on the device (``include/gpax_b200.h``).  Unlike the reference,
hyper-parameters may carry a leading sample axis: scalars give ``[n, m]``,
arrays of ``S`` samples give ``[S, n, m]`` in one launch.
"""
from __future__ import annotations

import numpy as np

from . import distributions as dist

_PREFIX_ERR = "{} must be a Parameter"


class Parameter:
    """A constant or a prior distribution, with ``plate`` independent copies
    (gpax/utils/prior_utils.py:95-118)."""

    def __init__(self, value, plate: int = 1):
        if not isinstance(plate, int) or isinstance(plate, bool):
            raise TypeError("plate must be an int")
        if plate < 1:
            raise ValueError("plate must be >= 1")
        if not isinstance(value, (dist.Distribution, int, float)):
            raise TypeError("value must be a Distribution, int or float")
        self.value = value
        self.plate = plate

    @property
    def is_constant(self):
        return not isinstance(self.value, dist.Distribution)

    def draw(self, rng, num):
        """``num`` prior draws, shape ``[num]`` (plate 1) or ``[num, plate]``."""
        shape = (num,) if self.plate == 1 else (num, self.plate)
        if self.is_constant:
            return np.full(shape, float(self.value))
        return np.asarray(self.value.sample(rng, shape), float)

    def __repr__(self):
        return f"Parameter({self.value!r}, plate={self.plate})"


class Prior:
    """Anything holding ``Parameter`` attributes whose names start with
    ``attribute_prefix`` (gpax/utils/prior_utils.py:121-185)."""

    attribute_prefix = ""

    @property
    def _params(self):
        """``Parameter`` attributes whose name starts with the prefix, declared either on
        the instance or on the class (the reference's attrs ``field(default=Parameter(...))``
        pattern, examples/04_Structured_GP.ipynb cell 10), in declaration order."""
        out = {}
        for klass in reversed(type(self).__mro__):
            for key, obj in vars(klass).items():
                if key.startswith(self.attribute_prefix) and isinstance(obj, Parameter):
                    out[key] = obj
        for key, obj in vars(self).items():
            if key.startswith(self.attribute_prefix) and isinstance(obj, Parameter):
                out[key] = obj
        return {k: getattr(self, k) for k in out if isinstance(getattr(self, k), Parameter)}

    @property
    def params(self):
        return {k: v.value for k, v in self._params.items()}

    def sample_parameters(self, rng=None, num_samples=1):
        """Prior draws of every parameter: dict of ``[num_samples(, plate)]``.
        (The reference returns NumPyro sample sites; here these are values.)"""
        rng = np.random.default_rng() if rng is None else rng
        return {k: v.draw(rng, num_samples) for k, v in self._params.items()}

    def get_sample_parameter_means(self):
        return {k: (float(v.value) if v.is_constant else v.value.mean)
                for k, v in self._params.items()}


def _param(obj, name):
    if not isinstance(obj, Parameter):
        raise TypeError(_PREFIX_ERR.format(name))
    return obj


class Kernel(Prior):
    """Base: ``k_noise`` (HalfNormal(0.01)) and ``k_jitter`` (1e-6 constant),
    gpax/kernels.py:69-75."""

    attribute_prefix = "k_"
    kind = None

    def __init__(self, k_noise=None, k_jitter=None):
        self.k_noise = _param(Parameter(dist.HalfNormal(0.01), 1) if k_noise is None else k_noise,
                              "k_noise")
        self.k_jitter = _param(Parameter(1.0e-6, 1) if k_jitter is None else k_jitter, "k_jitter")

    def _kernel_function(self, x1, x2, **params):
        raise NotImplementedError

    def sample_prior(self, x1, x2, rng=None):
        p = {k: v[0] for k, v in self.sample_parameters(rng, 1).items()}
        return self.__call__(x1, x2, **p)

    def __call__(self, x1, x2, **params):
        return self._kernel_function(x1, x2, **params)

    # -- device evaluation --------------------------------------------------
    def _gram(self, x1, x2, k_scale, k_length, k_noise, k_jitter, apply_noise, k_period=None,
              noise_is_vector=None):
        from . import engine as E   # deferred: importing the declarations needs no GPU

        dev = E._require_cuda()
        x1 = np.asarray(x1, float)
        x2 = np.asarray(x2, float)
        x1 = x1.reshape(len(x1), -1)
        x2 = x2.reshape(len(x2), -1)
        if x1.shape[1] != x2.shape[1]:
            raise ValueError(f"x1 and x2 differ in dimension: {x1.shape} vs {x2.shape}")
        # size-1 arrays are scalars, exactly as NumPy/JAX broadcasting treats them in the
        # reference's formulas; only k_scale with more than one entry opens a sample axis
        ks = np.asarray(k_scale, float)
        if ks.size == 1:
            ks = ks.reshape(())
        if ks.ndim > 1:
            raise ValueError(f"k_scale must be a scalar or [S], got shape {ks.shape}")
        batched = ks.ndim == 1
        hp = {"k_scale": np.atleast_1d(ks)}
        S = hp["k_scale"].shape[0]
        kj = np.asarray(k_jitter, float)
        hp["k_jitter"] = kj.reshape(()) if kj.size == 1 else kj
        kl = np.asarray(k_length, float)
        hp["k_length"] = kl.reshape(S, -1) if batched else kl.reshape(1, -1)
        if k_period is not None:
            kp_ = np.asarray(k_period, float)
            hp["k_period"] = kp_.reshape(S, -1) if batched else kp_.reshape(1, -1)
        # the reference adds the diagonal only for equal shapes (gpax/kernels.py:124,145,166)
        add_diag = bool(apply_noise) and x1.shape == x2.shape
        kn = np.asarray(k_noise, float)
        if kn.size == 1:
            kn = kn.reshape(())
        noise_vec = None
        n1 = x1.shape[0]
        if noise_is_vector is None:
            # un-batched call (the reference's own signature): a length-n vector is per-point
            # noise (y_std**2, gpax/gp.py:289-290).  Batched call: [S] is per-sample; when S == n
            # the two readings collide, so the caller has to say which one it means.
            if kn.ndim == 1 and batched and kn.shape[0] == S and S == n1 and add_diag:
                raise ValueError(f"k_noise of length {S} is ambiguous (S == n == {S}): pass "
                                 "noise_is_vector=True (per point) or False (per sample)")
            noise_is_vector = kn.ndim == 1 and not (batched and kn.shape[0] == S)
        if kn.ndim == 0 or not noise_is_vector:
            if kn.ndim == 1 and kn.shape[0] != S:
                raise ValueError(f"per-sample k_noise has {kn.shape[0]} entries for {S} samples")
            hp["k_noise"] = kn
        elif kn.ndim == 1 and kn.shape[0] == n1:
            noise_vec = kn              # heteroscedastic y_std**2 (gpax/gp.py:289-290)
        else:
            raise ValueError(f"k_noise of shape {kn.shape} matches neither the sample axis "
                             f"({S}) nor the number of points ({n1})")
        hpd = E.HyperParams(self.kind, x1.shape[1], hp, dev)
        d1, d2 = E._dev(x1, dev), E._dev(x2, dev)
        nz = None if noise_vec is None else E._dev(noise_vec, dev)
        out = E.gram(hpd, d1, d2, add_diag=add_diag, noise=nz, noise_is_vector=nz is not None)
        k = out[:, :, : x2.shape[0]].cpu().numpy()
        return k if batched else k[0]


class ScaleKernel(Kernel):
    """Adds ``k_scale`` and ``k_length`` (LogNormal(0,1) each, gpax/kernels.py:101-107)."""

    def __init__(self, k_scale=None, k_length=None, k_noise=None, k_jitter=None):
        super().__init__(k_noise, k_jitter)
        self.k_scale = _param(Parameter(dist.LogNormal(0.0, 1.0), 1) if k_scale is None
                              else k_scale, "k_scale")
        self.k_length = _param(Parameter(dist.LogNormal(0.0, 1.0), 1) if k_length is None
                               else k_length, "k_length")

    def _kernel_function(self, x1, x2, k_scale, k_length, k_noise=0.0, k_jitter=1.0e-6,
                         apply_noise=True, noise_is_vector=None):
        return self._gram(x1, x2, k_scale, k_length, k_noise, k_jitter, apply_noise,
                          noise_is_vector=noise_is_vector)


class RBFKernel(ScaleKernel):
    """``k_scale * exp(-r2/2)`` (gpax/kernels.py:110-126)."""
    kind = "rbf"


class MaternKernel(ScaleKernel):
    """Matern nu=5/2 (gpax/kernels.py:129-147)."""
    kind = "matern52"


class PeriodicKernel(ScaleKernel):
    """``k_scale * exp(-2 * sum_d sin^2(pi*(x-z)/p) / l^2)`` (gpax/kernels.py:150-168).
    The reference's call signature requires ``k_period`` but declares no prior
    for it; here it is a declared Parameter (LogNormal(0,1)) so the family is
    usable with ``fit``."""
    kind = "periodic"

    def __init__(self, k_scale=None, k_length=None, k_period=None, k_noise=None, k_jitter=None):
        super().__init__(k_scale, k_length, k_noise, k_jitter)
        self.k_period = _param(Parameter(dist.LogNormal(0.0, 1.0), 1) if k_period is None
                               else k_period, "k_period")

    def _kernel_function(self, x1, x2, k_scale, k_length, k_period, k_noise=0.0,
                         k_jitter=1.0e-6, apply_noise=True, noise_is_vector=None):
        return self._gram(x1, x2, k_scale, k_length, k_noise, k_jitter, apply_noise, k_period,
                          noise_is_vector=noise_is_vector)
