"""The scene builders live in the package (tracer_b200/models/scenes.py); the tests keep this name."""
from tracer_b200.models.scenes import *  # noqa: F401,F403
from tracer_b200.models.scenes import NS, HSTAT_CSV  # noqa: F401
