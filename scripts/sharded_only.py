"""BASELINE config 5 alone: one batch-32-shaped proof on a multi-device context over all visible GPUs
(bench.sharded_block), one JSON line. Usage: python scripts/sharded_only.py [n_gpus]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import torch

import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
args = argparse.Namespace(steps=8)
print(json.dumps(bench.sharded_block(args, n, list(range(n)))), flush=True)
