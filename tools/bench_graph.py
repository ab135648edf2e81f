#!/usr/bin/env python
"""CUDA-graph round-robin figure of a small-batch workload (bench.run_graph_round_robin) on its own.
usage: python tools/bench_graph.py c4 24   -> one JSON line (arm flags come from the MCOV_DEC_* environment)"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench

wl = sys.argv[1] if len(sys.argv) > 1 else "c4"
replicas = int(sys.argv[2]) if len(sys.argv) > 2 else 24
out = bench.run_graph_round_robin(torch, torch.device("cuda:0"), wl, replicas)
out.update(workload=wl, bench="graph_round_robin",
           flags=" ".join("%s=%s" % kv for kv in sorted(os.environ.items()) if kv[0].startswith("MCOV_")))
print(json.dumps(out))
