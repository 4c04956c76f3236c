"""Runs only the fused equaliser/demapper on the bench workload (for ncu captures and ablations)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gr-dtl_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench, dtl_b200
from dtl_b200 import synth
sys.argv = ["bench.py"]
a = bench.parse_args()
S, F = int(os.environ.get("EQ_S", "256")), 32
rx = dtl_b200.ofdm_receiver(bench.product_config(a))
b = synth.make_batch(rx, S, F, 4, fec_id=1, seed=5, snr_db=22.0)
freq = rx.demod(b["iq"]); rx.sync()
rx.profile(True)
for i in range(4):
    rx.equalize(freq, b["taps"], b["meta"], S, F, want=("res", "llr"))
rx.sync()
print("eq ms/launch", rx.stage_times()["eq_demap"][0] / 4)
