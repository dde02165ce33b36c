"""cfg1 latency: one pair per call through the drop-in module (device us / wall us per call), plus the launch list hint.
usage: python tools/bench_cfg1.py"""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
dev = torch.device("cuda", 0)
print(json.dumps(bench.leg_cfg1(dev, bench.peaks()[0], False)))
