"""Where the one-frame call's time goes: S = 1, F = 1 through dtl_b200_rx with host buffers, with device buffers + sync,
and the kernels alone (CUDA events)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gr-dtl_b200"))
import numpy as np, torch
import bench, dtl_b200
from dtl_b200 import synth
rx = dtl_b200.ofdm_receiver(bench.product_config(4), device=0)
b = synth.make_batch(rx, 1, 64, 4, fec_id=1, seed=3, snr_db=22.0)
iq, taps, meta = b["iq"], b["taps"], b["meta"]
h_iq, h_taps, h_meta = iq.cpu().numpy(), taps.cpu().numpy(), b["meta_np"]
rx.reset(1)
def med(fn, n=300):
    ts = []
    for i in range(n):
        t0 = time.perf_counter(); fn(i % 64); ts.append((time.perf_counter() - t0) * 1e6)
    return float(np.median(ts[20:]))
host = med(lambda j: rx.rx(h_iq[j:j + 1], h_taps[j:j + 1], h_meta[j:j + 1], 1, 1))
out = bench.new_out(rx, 1, 1, lambda s, d: rx._new(dtl_b200.MEM_DEVICE, s, np.dtype(d)), np)
def dev(j):
    rx.rx(iq[j:j + 1], taps[j:j + 1], meta[j:j + 1], 1, 1, out=out); rx.sync()
devt = med(dev)
def enq(j):
    rx.rx(iq[j:j + 1], taps[j:j + 1], meta[j:j + 1], 1, 1, out=out)
enqt = med(enq); rx.sync()
st = rx.torch_stream(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
rx.sync(); e0.record(st)
for j in range(200): enq(j % 64)
e1.record(st); rx.sync()
print("host-buffer call %.1f us | device buffers + sync %.1f us | enqueue only %.1f us | GPU time per call back to back %.1f us"
      % (host, devt, enqt, e0.elapsed_time(e1) * 1e3 / 200))
