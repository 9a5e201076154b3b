"""gpax_b200 -- B200-native (sm_100a) exact-GP hot path with the gpax call surface.

Public surface mirrors the reference for the hot path only:
``gpax_b200.kernels`` (RBFKernel / MaternKernel / PeriodicKernel),
``gpax_b200.gp.ExactGP`` (fit / predict / sample) and ``gpax_b200.acquisition``
(ExpectedImprovement / UpperConfidenceBound, aliases EI / UCB / UE).
All numerics run in ``libgpax_b200.so`` (C ABI in ``include/gpax_b200.h``).
"""
__version__ = "0.1.0"
