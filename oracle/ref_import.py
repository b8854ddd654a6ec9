"""TEST INFRASTRUCTURE ONLY.  Makes the *real* reference importable in the build
container (``import tracer`` from /root/reference) so that
tests/golden/make_golden.py can record golden vectors and tests can cross-check
the NumPy oracle.  /root/reference does not exist on the GPU box: nothing under
tracer_b200/, bench.py or the `-m gpu` tests may import this module.
"""
import builtins
import os
import sys

REFERENCE_ROOT = os.environ.get("TRACER_REFERENCE", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "tracer"))


def enable():
    """Put the reference and the shapely stub on sys.path; add the Python-2 shims
    the reference still needs (xrange: sphere_surface.py:198)."""
    if not available():
        raise ImportError("reference tree not present at %s" % REFERENCE_ROOT)
    shim = os.path.join(os.path.dirname(os.path.abspath(__file__)), "refshim")
    # appended, not prepended: the reference has top-level packages named like ours
    # (tests/), and child processes spawned by the tests inherit sys.path
    for p in (shim, REFERENCE_ROOT):
        if p not in sys.path:
            sys.path.append(p)
    if not hasattr(builtins, "xrange"):
        builtins.xrange = range
    import tracer  # noqa: F401
    return tracer
