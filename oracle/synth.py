"""Compatibility shim: the synthetic workload generators live in ``workloads/synth.py`` (they
are input generators, not part of the oracle's arithmetic)."""

from workloads.synth import *  # noqa: F401,F403
from workloads.synth import ANCHORS, CLASSES, FANOUT_PROBS, FANOUT_VALUES  # noqa: F401
