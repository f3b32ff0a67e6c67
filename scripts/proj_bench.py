#!/usr/bin/env python
"""Projection leg alone (bench_proj.run) on cuda:0: python scripts/proj_bench.py [c4|c4small]"""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench_proj
pkg = importlib.import_module("ithaca-fv_b200")
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
pod = pkg.Pod(0)
out, _, _, _ = bench_proj.run(pod, torch, dev, sys.argv[1] if len(sys.argv) > 1 else "c4", hbm_peak=6537.0, dmma_peak=pkg.dmma_peak(0))
print(json.dumps(out))
