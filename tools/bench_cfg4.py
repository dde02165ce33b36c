"""cfg4 (64 OpenGG patches per launch) at w 64 / 128 / 256: ms per forward (+ winner map) + backward.  usage: python tools/bench_cfg4.py"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
d = bench.leg_cfg4(torch.device("cuda", 0), bench.peaks()[0], False)
print({k: round(v["ms_per_step"], 3) for k, v in d.items() if isinstance(v, dict) and "ms_per_step" in v})
