"""matFact on B200: the factorization loop of vladstojna/recommender-system as
hand-written sm_100a CUDA behind a C ABI (include/matfact_b200.h).

Only what the hot path needs lives here: ``csrc/`` (kernels + C ABI),
``host/`` (plain-C pattern builder and the drop-in ``matFact`` program),
``api`` (ctypes mirror of the reference's two calls at the seam), ``io``
(the reference's file format) and ``synth`` (stand-ins for the missing inputs).
"""
from . import io  # noqa: F401

__all__ = ["io", "api", "synth"]
