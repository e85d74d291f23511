#!/usr/bin/env python
"""RANSAC leg of bench.py alone (cfg2: 100k hypotheses on the cfg1 scene)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import calib_b200  # noqa
from calib_b200 import calib_ransac as cr
out = cr.bench_leg(0, 1, torch.device("cuda", 0), steps=3, no_cpu=True)     # the CPU leg lives in bench.py
print(json.dumps(out))
