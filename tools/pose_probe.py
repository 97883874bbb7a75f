"""Runs only the batched single-pose kernel (k_pose_lm, 4096 poses x 512 correspondences) through bench.py's own pose leg and
prints its JSON block — the command to put under `ncu -k regex:k_pose_lm` (the kernel is on-chip bound; no HBM fraction is
claimed for it).  GPU required."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba
import bench

args = argparse.Namespace(seed=0)
print(json.dumps(bench.bench_pose(bench.Dist(), b200ba, args)))
