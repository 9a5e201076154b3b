"""``ExactGP`` with the reference's call surface (gpax/gp.py:98-131 fields,
``fit`` :513, ``sample`` :395, ``predict`` :486, ``GPSample`` :52-93), running
on the device hot path of ``include/gpax_b200.h``.

Differences that follow from the scope:

* ``predict(x_new, fast=False)`` returns the RNG-free mixture moments over the
  hyper-parameter samples -- the ``gp_samples -> infinity`` limit of the
  reference's ``np.mean/np.std(y, axis=(0, 1))`` (gpax/gp.py:76-90) -- computed
  by ``gpx_predict`` + ``gpx_acq_reduce``; no ``[M, M]`` covariance is formed.
* ``sample`` additionally returns joint draws ``y [hp, gp, M]`` like the
  reference.  Drawing needs the full ``[M, M]`` posterior covariance and its
  Cholesky factor (row N1, "next"); until that has its own kernels it is
  assembled on the device from ``gpx_gram`` outputs with cuBLAS/cuSOLVER calls
  through torch and is limited to modest ``M``.  Pass ``draw=False`` to skip.
* ``fit`` runs the samplers of ``gpax_b200.infer`` (NUTS by default) over the kernel
  AND mean (``m_*``) parameters.
* ``use_cholesky``: the device path always works from the Cholesky factor.  The
  reference's DEFAULT is the ``jnp.linalg.inv`` branch (gpax/gp.py:129,303-310),
  which computes the same quantities; the two agree to ~cond(K)*eps (DESIGN.md
  section 2), so the flag is accepted with either value and recorded in
  ``metadata["solver"]``.
* Learnable mean parameters of an UN-fit model are drawn from their priors next to
  the kernel parameters (the reference passes the Distribution objects themselves
  to the mean function there, gpax/gp.py:328, which only works for constants).
"""
from __future__ import annotations

import numpy as np

from . import state
from .kernels import Kernel
from .means import ConstantMean, Mean
from .transforms import IdentityTransform, NormalizeTransform, ScaleTransform, Transform

_ARRAY_OR_NONE = (np.ndarray, type(None))
_MAX_DRAW_ELEMS = 1 << 28        # S*M*M doubles assembled per chunk for joint draws


def _check(name, value, types):
    if not isinstance(value, types):
        raise TypeError(f"'{name}' must be {types} (got {value!r} of type {type(value)})")


def _check_pos_int(name, value):
    if not isinstance(value, int) or isinstance(value, bool):
        raise TypeError(f"'{name}' must be int (got {type(value)})")
    if value <= 0:
        raise ValueError(f"'{name}' must be > 0: {value}")


class GPSample:
    """Container for results of ``sample`` (gpax/gp.py:52-93)."""

    def __init__(self, hp=None, y=None, mu=None, sd=None):
        self._hp, self._y, self._mu, self._sd = hp, y, mu, sd
        self._cache = {}

    @property
    def y(self):
        """Joint draws, shape ``(hp, gp, M)``; ``None`` when ``draw=False``."""
        return self._y

    @property
    def hp(self):
        return self._hp

    @property
    def mu_per_hp(self):
        """Posterior mean for every hyper-parameter sample, ``(hp, M)``."""
        return self._mu

    @property
    def sd_per_hp(self):
        return self._sd

    def _moments(self):
        if "m" not in self._cache:
            mu, sd = np.asarray(self._mu), np.asarray(self._sd)
            if mu.shape[0] == 1:
                self._cache["m"] = (mu.squeeze(), sd.squeeze())
            elif self._y is not None:           # the reference's Monte-Carlo estimate
                self._cache["m"] = (np.mean(self._y, axis=(0, 1)).squeeze(),
                                    np.std(self._y, axis=(0, 1)).squeeze())
            else:                               # its RNG-free limit
                m = mu.mean(axis=0)
                v = (sd**2).mean(axis=0) + mu.var(axis=0)      # law of total variance, two-pass
                self._cache["m"] = (m.squeeze(), np.sqrt(v).squeeze())
        return self._cache["m"]

    @property
    def mu(self):
        return self._moments()[0]

    @property
    def sd(self):
        return self._moments()[1]

    @property
    def ci(self):
        return self.mu - 2.0 * self.sd, self.mu + 2.0 * self.sd


def _affine_of(transform):
    """``reverse`` of an output transform as (mu_scale, mu_shift, sd_scale) so it
    can be folded into ``gpx_acq_reduce`` (gpax/transforms.py:312-317)."""
    from .transforms import EPSILON, T_HI, T_LO

    if isinstance(transform, IdentityTransform):
        return 1.0, 0.0, 1.0
    if isinstance(transform, NormalizeTransform):
        std, mean = float(transform._std.squeeze()), float(transform._mean.squeeze())
        return std + EPSILON, mean, std
    if isinstance(transform, ScaleTransform):
        span_r = float((transform._maxima - transform._minima).squeeze()) + EPSILON
        span_t = T_HI - T_LO + EPSILON
        mn = float(transform._minima.squeeze())
        return span_r / span_t, (mn * span_t - T_LO * span_r) / span_t, span_r / span_t
    z = np.zeros((1, 1))
    shift = float(np.squeeze(transform.reverse(z, transforms_as="mean")))
    scale = float(np.squeeze(transform.reverse(z + 1.0, transforms_as="mean"))) - shift
    return scale, shift, float(np.squeeze(transform.reverse(z + 1.0, transforms_as="std")))


class GaussianProcess:
    """Fields and validators of gpax/gp.py:98-131."""

    def __init__(self, kernel, x=None, y=None, y_std=None, observation_noise=False,
                 hp_samples=100, gp_samples=10, input_transform="default",
                 output_transform="default", metadata=None, use_cholesky=False, mean=None):
        _check("kernel", kernel, Kernel)
        _check("x", x, _ARRAY_OR_NONE)
        _check("y", y, _ARRAY_OR_NONE)
        if not isinstance(y_std, _ARRAY_OR_NONE + (float, int)) or isinstance(y_std, bool):
            raise TypeError(f"'y_std' must be an array, float, int or None (got {type(y_std)})")
        _check("observation_noise", observation_noise, bool)
        _check_pos_int("hp_samples", hp_samples)
        _check_pos_int("gp_samples", gp_samples)
        _check("use_cholesky", use_cholesky, bool)
        if input_transform == "default":
            input_transform = ScaleTransform()
        if output_transform == "default":
            output_transform = NormalizeTransform()
        _check("input_transform", input_transform, (Transform, type(None)))
        _check("output_transform", output_transform, (Transform, type(None)))
        mean = ConstantMean() if mean is None else mean
        _check("mean", mean, Mean)

        self.kernel, self.x, self.y, self.y_std = kernel, x, y, y_std
        self.observation_noise = observation_noise
        self.hp_samples, self.gp_samples = hp_samples, gp_samples
        self.metadata = {} if metadata is None else metadata
        self.use_cholesky, self.mean = use_cholesky, mean
        self.metadata.setdefault(
            "solver", "device Cholesky (gpx_potrf_batched); use_cholesky=%s accepted, the reference's "
                      "inv branch agrees to ~cond(K)*eps" % use_cholesky)
        self._is_fit = False
        self._posterior_cache = None

        # gpax/gp.py:170-190
        if self.x is None and self.y_std is not None:
            raise ValueError("y_std cannot be set if no x or y data is given")
        if self.y is not None and isinstance(self.y_std, (float, int)):
            self.y_std = np.ones(self.y.shape) * self.y_std
        if (self.y is None) ^ (self.x is None):
            raise ValueError("x and y must either both be None or not")
        if input_transform is None or self.x is None:
            input_transform = IdentityTransform()
        if output_transform is None or self.y is None:
            output_transform = IdentityTransform()
        self.input_transform, self.output_transform = input_transform, output_transform
        self.input_transform.fit(self.x)
        self.output_transform.fit(self.y)

    # -- transformed views (gpax/gp.py:193-208) -------------------------------
    @property
    def x_transformed(self):
        return self.input_transform.forward(self.x, transforms_as="mean")

    @property
    def y_transformed(self):
        y = self.output_transform.forward(self.y, transforms_as="mean")
        return np.atleast_1d(y.squeeze())

    @property
    def y_std_transformed(self):
        y_std = self.output_transform.forward(self.y_std, transforms_as="std")
        return None if y_std is None else np.atleast_1d(y_std.squeeze())

    # -- hyper-parameter sources ----------------------------------------------
    @staticmethod
    def _median(hp):
        # jnp.median(value, keepdims=True) over ALL axes (gpax/gp.py:345-349)
        return {k: np.median(np.asarray(v, float), keepdims=True) for k, v in hp.items()}

    def _get_hp_samples_from_prior(self, rng, fast):
        hp = self.kernel.sample_parameters(rng, self.hp_samples)
        for k in self.mean.learnable():          # see the module docstring
            hp[k] = self.mean._params[k].draw(rng, self.hp_samples)
        return (self._median(hp), 1) if fast else (hp, self.hp_samples)

    def _get_hp_samples_from_posterior(self, fast):
        raise NotImplementedError

    def _hp_for(self, rng, fast):
        if self.x is None or not self._is_fit:
            return self._get_hp_samples_from_prior(rng, fast)
        return self._get_hp_samples_from_posterior(fast)

    # -- device moments ------------------------------------------------------
    @staticmethod
    def _n_samples(hp):
        return len(np.atleast_1d(hp["k_scale"]))

    def _mean_values(self, x, hp):
        """Mean function per hyper-parameter sample (gpax/gp.py:266-278: the ``m_*`` entries of
        the sample dict go to the mean, the ``k_*`` entries to the kernel): ``[S, N]``, or
        ``[1, N]`` when no mean parameter is sampled."""
        return self.mean.evaluate(x, hp, self._n_samples(hp))

    def _m_new_kw(self, x_new_t, hp):
        """Mean at the new points as the argument ``gpx_predict`` wants: one constant, or [S, M]."""
        m_new = self._mean_values(x_new_t, hp)
        if m_new.shape[0] == 1 and np.all(m_new == m_new[0, 0]):
            return {"m_const": float(m_new[0, 0])}
        S = self._n_samples(hp)
        return {"m_new": np.ascontiguousarray(np.broadcast_to(m_new, (S, m_new.shape[1])))}

    def _posterior(self, hp):
        """Cached device factorisation for this exact set of samples."""
        from . import engine as E

        key = tuple((k, np.asarray(v).tobytes()) for k, v in sorted(hp.items()))
        if self._posterior_cache is None or self._posterior_cache[0] != key:
            post = E.Posterior(self.kernel.kind, self.x_transformed, self.y_transformed, hp,
                               m_train=self._mean_values(self.x_transformed, hp),
                               y_std=self.y_std_transformed)
            bad = post.info.cpu().numpy()
            if bad.any():
                raise np.linalg.LinAlgError(
                    f"covariance not positive definite for {int((bad != 0).sum())} of "
                    f"{len(bad)} hyper-parameter samples")
            self._posterior_cache = (key, post)
        return self._posterior_cache[1]

    def _moments_device(self, x_new_t, hp):
        """``(mu [S, M], var [S, M])`` device tensors in transformed space."""
        import torch

        from . import engine as E

        x_new_2d = np.asarray(x_new_t, float).reshape(len(x_new_t), -1)
        if self.x is None:
            # unconditioned prior (gpax/gp.py:211-243): stationary kernel, diagonal only
            dev = E._require_cuda()
            S = self._n_samples(hp)
            m_new = self._mean_values(x_new_t, hp)
            noise = np.asarray(hp["k_noise"], float).reshape(-1) if self.observation_noise else 0.0
            var = np.asarray(hp["k_scale"], float).reshape(-1) + noise \
                + np.asarray(hp["k_jitter"], float).reshape(-1)
            mu = E._dev(np.broadcast_to(m_new, (S, m_new.shape[1])).copy(), dev)
            return mu, E._dev(np.broadcast_to(var[:, None], mu.shape).copy(), dev), None
        post = self._posterior(hp)
        mu, var = post.predict(x_new_2d, observation_noise=self.observation_noise,
                               **self._m_new_kw(x_new_t, hp))
        return mu, var, post

    def _draw(self, rng_seed, x_new_t, hp, mu, post):
        """Joint draws ``y [S, G, M]`` from ``N(mu_s, cov_s)`` (gpax/gp.py:332-338) by ``gpx_draw``:
        covariance, its Cholesky factor and JAX-compatible threefry normals all on the device.
        ``state.threefry_partitionable`` selects JAX's counter layout (see ``gpx_random_normal``)."""
        from . import engine as E

        S, M = mu.shape
        G = self.gp_samples
        if S * M * M > _MAX_DRAW_ELEMS * 8:
            raise ValueError(f"joint draws need {S}x{M}x{M} covariances; lower M or hp_samples, "
                             "or call sample(..., draw=False) / predict()")
        dev = mu.device
        xn = E._dev(np.asarray(x_new_t, float).reshape(M, -1), dev)
        y = None
        chunk = max(1, _MAX_DRAW_ELEMS // max(1, M * max(M, 0 if post is None else post.n)))
        if chunk < S:
            raise ValueError(f"joint draws of {S} samples at M={M} exceed the draw workspace; lower M "
                             "or hp_samples, or call sample(..., draw=False) / predict()")
        hpd = post.hpd if post is not None else E.HyperParams(self.kernel.kind, xn.shape[1], hp, dev)
        y, info = E.draw(hpd, None if post is None else post.x, xn, None if post is None else post.W,
                         mu, G, rng_seed, diag_noise=hpd.noise if self.observation_noise else None,
                         partitionable=state.threefry_partitionable)
        if int(info.abs().max()) != 0:
            raise ValueError("posterior covariance at x_new is not positive definite "
                             "(add observation noise or jitter, or use fewer / more distinct points)")
        return y

    # -- public API ----------------------------------------------------------
    def sample(self, x_new, fast=False, draw=True):
        """gpax/gp.py:395-461.  Returns ``GPSample`` in un-transformed units."""
        x_new_t = self.input_transform.forward(np.asarray(x_new, float), transforms_as="mean")
        key, rng = state.get_rng_key()
        hp, _ = self._hp_for(rng, fast)
        mu, var, post = self._moments_device(x_new_t, hp)
        sd = var.sqrt()          # un-clamped like gpax/gp.py:338: a negative variance gives NaN
        y = None
        if draw:
            y = self._draw(key, x_new_t, hp, mu, post).cpu().numpy()
            y = self.output_transform.reverse(y.reshape(-1, 1), transforms_as="mean").reshape(y.shape)
        rev = self.output_transform.reverse
        S, M = mu.shape
        mu_h = rev(mu.cpu().numpy().reshape(-1, 1), transforms_as="mean").reshape(S, M)
        sd_h = rev(sd.cpu().numpy().reshape(-1, 1), transforms_as="std").reshape(S, M)
        hp = {k: np.atleast_1d(v) for k, v in hp.items()}
        return GPSample(hp=hp, y=y, mu=mu_h, sd=sd_h)

    def _acquire(self, x_new, fast, acq, beta=10.0, y_max=0.0, want_arrays=True, x_new_t_dev=None):
        """Fused moments -> reverse transform -> acquisition -> argmax on the device
        (gpax/gp.py:76-90 + gpax/acquisition.py:211-216,260-266,291-298).
        Returns ``(mu, sd, acq values, best value, best index)``.  ``x_new_t_dev``: candidates
        that already live on the device in kernel (transformed) space (``gpx_halton``); only a
        conditioned model takes them, and only a non-constant mean brings them to the host."""
        from . import engine as E

        _, rng = state.get_rng_key()
        hp, _ = self._hp_for(rng, fast)
        a, b, c = _affine_of(self.output_transform)
        if x_new_t_dev is not None:
            if self.x is None:
                raise ValueError("device-resident candidates need a conditioned model")
            post = self._posterior(hp)
            if isinstance(self.mean, ConstantMean):
                kw = {"m_const": float(self.mean.mean)}
            else:                      # user Python: evaluate it where it can run
                kw = self._m_new_kw(x_new_t_dev.cpu().numpy(), hp)
            o_mu, o_sd, o_acq, bv, bi = post.predict_acq(
                x_new_t_dev, acq, beta=beta, y_max=y_max, mu_scale=a, mu_shift=b, sd_scale=c,
                observation_noise=self.observation_noise, want_arrays=want_arrays, **kw)
            host = [None if t is None else t.cpu().numpy() for t in (o_mu, o_sd, o_acq)]
            return host[0], host[1], host[2], float(bv.item()), int(bi.item())
        x_new_t = self.input_transform.forward(np.asarray(x_new, float), transforms_as="mean")
        if self.x is not None:
            # conditioned model: one fused gpx_predict_acq sweep, no [S, M] moment buffers
            post = self._posterior(hp)
            kw = self._m_new_kw(x_new_t, hp)
            x_new_t = np.asarray(x_new_t, float).reshape(len(x_new_t), -1)
            o_mu, o_sd, o_acq, bv, bi = post.predict_acq(
                x_new_t, acq, beta=beta, y_max=y_max, mu_scale=a, mu_shift=b, sd_scale=c,
                observation_noise=self.observation_noise, want_arrays=want_arrays, **kw)
        else:
            mu, var, _ = self._moments_device(x_new_t, hp)
            o_mu, o_sd, o_acq, pv, pi = E.acq_reduce(mu, var, acq, beta=beta, y_max=y_max,
                                                     mu_scale=a, mu_shift=b, sd_scale=c,
                                                     want_arrays=want_arrays)
            bv, bi = E.argmax_finalize(pv, pi)
        host = [None if t is None else t.cpu().numpy() for t in (o_mu, o_sd, o_acq)]
        return host[0], host[1], host[2], float(bv.item()), int(bi.item())

    def predict(self, x_new, fast=False):
        """gpax/gp.py:486-503: ``(mu, sd)`` at ``x_new`` in un-transformed units."""
        from ._cabi import ACQ_MEAN

        mu, sd, _, _, _ = self._acquire(x_new, fast, ACQ_MEAN)
        return mu.squeeze(), sd.squeeze()

    def fit(self):
        raise NotImplementedError


class ExactGP(GaussianProcess):
    """gpax/gp.py:506-545.  ``mcmc`` holds the sampler output after ``fit``."""

    def __init__(self, kernel, x=None, y=None, y_std=None, observation_noise=False,
                 hp_samples=100, gp_samples=10, input_transform="default",
                 output_transform="default", metadata=None, use_cholesky=False, mean=None,
                 mcmc=None, num_warmup=2000, num_chains=1, chain_method="sequential",
                 mcmc_run_kwargs=None):
        super().__init__(kernel, x, y, y_std, observation_noise, hp_samples, gp_samples,
                         input_transform, output_transform, metadata, use_cholesky, mean)
        _check_pos_int("num_warmup", num_warmup)
        _check_pos_int("num_chains", num_chains)
        _check("chain_method", chain_method, str)
        if chain_method not in ("sequential", "vectorized", "parallel"):
            raise ValueError("chain_method must be 'sequential', 'vectorized' or 'parallel'")
        self.num_warmup, self.num_chains = num_warmup, num_chains
        # NumPyro's options (gpax/gp.py:509,521): "sequential" (the reference default) runs the
        # chains one after another, "vectorized" / "parallel" advance them in lock-step as one
        # batch of factorisations -- the fast choice on the device
        self.chain_method = chain_method
        self.mcmc_run_kwargs = {} if mcmc_run_kwargs is None else dict(mcmc_run_kwargs)
        self.mcmc = mcmc
        if mcmc is not None:                     # a model re-created around existing sampler output
            self._is_fit = True

    def fit(self):
        """NUTS over the kernel and mean hyper-parameters (gpax/gp.py:512-534).  As in the
        reference's ``_gp_prior`` (gpax/gp.py:133-151) the likelihood uses the sampled
        ``k_noise``; ``y_std`` only enters at prediction time.  ``mcmc_run_kwargs`` reaches the
        sampler (``sampler="nuts"|"hmc"``, ``target_accept``, ``max_tree_depth``, ...)."""
        from . import infer

        if self.x is None:
            raise ValueError("fit() needs data: x and y are None")
        key, rng = state.get_rng_key()
        ld = infer.LogDensity(self.kernel, self.mean, self.x_transformed, self.y_transformed)
        draws, info = infer.run(ld, self.num_warmup, self.hp_samples, self.num_chains, rng,
                                chain_method=self.chain_method, **self.mcmc_run_kwargs)
        flat = draws.reshape(-1, draws.shape[-1])          # group_by_chain=False
        samples = ld.constrained(flat)
        for name, p in self.kernel._params.items():        # constants are not sample sites
            if p.is_constant:
                samples.pop(name, None)
        self.mcmc = {"samples": samples, **info}
        self.metadata["fit_key"] = np.array(key)
        self._is_fit = True
        self._posterior_cache = None

    def _get_hp_samples_from_posterior(self, fast):
        hp = dict(self.mcmc["samples"])
        n = len(next(iter(hp.values())))
        for name, p in self.kernel._params.items():
            if p.is_constant:
                hp[name] = np.full(n, float(p.value))
        if fast:
            hp = self._median(hp)
        return hp, len(next(iter(hp.values())))
