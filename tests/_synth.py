"""Synthetic inputs for the tests: the package's seeded generators."""
from lxy_b200.synth import *  # noqa: F401,F403
