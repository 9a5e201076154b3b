"""Process-wide RNG key and verbosity, same surface as ``gpax.state``
(gpax/state.py:40-126): ``set_rng_key`` / ``get_rng_key`` / ``set_verbosity``.

``get_rng_key`` advances the integer key with the reference's chain
(``np.random.seed(key); np.random.randint(0, 2**32-1)``, gpax/state.py:117-126)
so seeded NumPy/SciPy consumers (the Halton sampler, gpax/acquisition.py:186)
see the same seeds.  The second element of the returned pair is a
``numpy.random.Generator`` where the reference returns a JAX key (JAX is not a
dependency of this package).
"""
from __future__ import annotations

import numpy as np

_KEY_MAX = 2**32 - 1
_VERBOSITY = {"silent": 0, "normal": 1, "debug": 2}

rng_key = 0
verbose = 1
#: counter layout of the threefry draws (``gpx_random_normal``): False = jax_threefry_partitionable
#: off (default of the reference's minimum jax 0.4.31), True = the default from jax 0.5.0
threefry_partitionable = False


def set_rng_key(key: int) -> None:
    global rng_key
    if not isinstance(key, (int, np.integer)) or key < 0 or key >= _KEY_MAX:
        raise ValueError(f"rng key must be an integer in [0, {_KEY_MAX})")
    rng_key = int(key)


def get_rng_key():
    """Returns ``(int_key, numpy Generator)`` and advances the state."""
    global rng_key
    np.random.seed(rng_key)
    new_key = int(np.random.randint(low=0, high=_KEY_MAX))
    rng_key = new_key
    return new_key, np.random.default_rng(new_key)


def set_verbosity(mode) -> None:
    global verbose
    if isinstance(mode, str):
        if mode not in _VERBOSITY:
            raise ValueError(f"Unknown verbosity {mode}; choose from {sorted(_VERBOSITY)}")
        verbose = _VERBOSITY[mode]
    else:
        verbose = int(mode)
