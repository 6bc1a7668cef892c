"""bayesbridge_b200: B200-native (sm_100a) Gibbs sweep for Bayesian bridge regression.

The product is libbayesbridge_b200.so (CUDA kernels + C ABI, include/bayesbridge_b200.h);
this package is the Python-side mirror of the reference's R front end plus helpers for
synthetic data, ESS and multi-GPU sharding.
"""
from .api import (Chains, Model, bridge_batched, bridge_batched_summary, bridge_reg, bridge_reg_stb, bridge_reg_tri, make_hyper)  # noqa: F401
from . import _lib  # noqa: F401

__all__ = ["bridge_reg_tri", "bridge_reg_stb", "bridge_reg", "bridge_batched", "bridge_batched_summary", "Model", "Chains", "make_hyper"]
