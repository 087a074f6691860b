"""Round 2: single-garment ms/frame of every launch shape at the three sizes of the metric (GPU box only)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

a = argparse.Namespace(n=224)
print(json.dumps(bench.single_garment_ours(a, 0, frames=40, warmup=8), indent=1))
