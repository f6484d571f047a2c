"""Synthetic inputs live in tools/synth.py (bench.py uses them too and must not import oracle/ outside its
CPU-baseline leg); re-exported here for the tests and the oracle harness."""
from tools.synth import *  # noqa: F401,F403
from tools.synth import SHAPES, make_cloud, make_clouds, make_patches, normalize_patches  # noqa: F401
