#!/usr/bin/env python
"""One configuration leg of bench.py on its own (for ncu): tools/leg.py config1|config3 <cnst>|config5 [steps]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gr-dtl_b200"))
import numpy as np
import bench
import dtl_b200


class A:
    steps = 5
    sustain_seconds = 0


ctx = bench.Ctx(A())
which = sys.argv[1]
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
if which == "config1":
    leg = bench.leg_config(ctx, dtl_b200, np, "config 1", int(sys.argv[2]) if len(sys.argv) > 2 else 4, False, 1024, 32, steps)
elif which == "config3":
    leg = bench.leg_config(ctx, dtl_b200, np, "config 3", int(sys.argv[2]), True, 1024, 16, steps)
else:
    leg = bench.leg_config(ctx, dtl_b200, np, "config 5", 6, True, 256, 32, steps, fft_len=1024, frame_len=5)
print(json.dumps(leg))
