"""Synthetic workloads of BASELINE.json (scene generators, body motion, default parameters).  Plain numpy: neither
the product library nor the oracle is imported here, so both arms of bench.py and the tests share one source of inputs."""
from . import scene  # noqa: F401
