"""aligator-swk_b200 — B200-native batched backend for aligator's `gar` LQ sub-problem solver.

The directory name carries a hyphen (repo convention); import it as ``aligator_swk_b200``
through ``garb200_pkg.load()`` at the repo root.  Sub-modules:

* ``layout`` — packed record layouts (mirror of include/garb200_layout.h)
* ``synth``  — synthetic random LQ problems (host, numpy)
* ``capi``   — ctypes binding of the C-ABI shared library (libgarb200.so); fails loudly if the
               CUDA library is missing — there is NO CPU fallback
* ``gar``    — host-side mirror of the reference's gar API (LqrKnot, LqrProblem,
               ProximalRiccatiSolver, ParallelRiccatiSolver, lqrComputeKktError, ...)
"""
__all__ = ["layout", "synth", "capi", "gar"]
