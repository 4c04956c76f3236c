#!/usr/bin/env python
"""bench.py -- throughput of the adaptive-OFDM receive hot path (BASELINE.json metric: RX Msamples/s and decoded
Mbit/s at 1/2/4/8 B200, % of the HBM roofline, beside the host CPU).

  python bench.py [--gpus N] [--steps K] [--warmup W]            this repo's CUDA path (libdtl_b200.so)
  python bench.py --impl reference [...]                          the reference's CPU algorithm (oracle port) timed
                                                                  on the host cores, same metric / config

One step = one pass of the whole receive path (CP strip + FFT -> frame equaliser + serialiser + soft demapper ->
TB reassembly + LDPC BP decode + CRC-32 + byte packing) over one batch of `streams` independent streams x `frames`
consecutive frames each (a fresh set of receivers per step: per-stream state is re-initialised on the stream).
Workload = BASELINE.json configs[1]'s PHY (64 carriers, cp 16, 20 payload symbols/frame, LDPC n=300 k=152, 16-QAM
at 22 dB) at configs[3]'s per-GPU multiplicity (1024 streams per B200); streams are independent, so N GPUs run N
disjoint shards with no collective on the data path (weak scaling), per-GPU statistics are summed on the host side.

`value`   : IQ samples consumed per second (incl. CP, sync and header symbols), inputs resident in HBM, timed with
            CUDA events on the library's own stream, max over ranks.
`e2e`     : the same through the C ABI with HOST (page-locked) buffers: host->device copies of IQ/taps/tags and the
            device->host read of the decoded bytes + per-frame / per-TB records are inside the timed region.
`roofline`: the dominant HBM-bound kernel (fused equaliser + serialiser + soft demapper), algorithmic bytes per
            frame x frames per launch / its CUDA-event time, against MEASURED_PEAKS.json.
`cpu_baseline`: the oracle (CPU restatement of the reference algorithm; the reference itself needs GNU Radio 3.10,
            absent here) on a bounded sample of the same workload, all host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "gr-dtl_b200"))

ALIST = os.path.join(ROOT, "tests", "golden", "n_0300_k_0152_gap_03.alist")   # examples/n_0300_k_0152_gap_03.alist
METRIC, UNIT = "rx_msamples_per_s", "Msamples/s"
CNST_NAMES = {1: "BPSK", 2: "QPSK", 3: "8PSK", 4: "16QAM", 5: "64QAM", 6: "256QAM"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--streams", type=int, default=1024, help="streams per GPU")
    ap.add_argument("--frames", type=int, default=32, help="frames per stream per step")
    ap.add_argument("--cnst", type=int, default=4)
    ap.add_argument("--snr", type=float, default=22.0)
    ap.add_argument("--ldpc-mode", default="reference", choices=["reference", "layered"])
    ap.add_argument("--ref-streams", type=int, default=64, help="streams of the CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-from-iq", action="store_true")
    return ap.parse_args()


def workload_name(a, streams):
    return ("ofdm_adaptive RX, 64-carrier PHY (cp 16, 20 syms/frame), %s @ %g dB + LDPC n=300 k=152 (config_fec), "
            "%d streams x %d frames per GPU per step" % (CNST_NAMES[a.cnst], a.snr, streams, a.frames))


def n_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ---- clocks during the timed region (pynvml; same fields as the nvidia-smi line of B200_PROFILING.md) -----------
class ClockSampler(threading.Thread):
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.sm_max, self.ok = index, False, [], set(), None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        while self.ok and not self.stop_flag:
            try:
                self.sm.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                r = int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.05)

    def result(self):
        self.stop_flag = True
        if not self.ok or not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": [], "samples": 0}
        s = sorted(self.sm)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons), "samples": len(s)}


# ---- the CPU arm: oracle port of the reference algorithm (the ONLY place bench.py touches oracle/) ----------------
def cpu_run(a, steps, warmup, max_seconds=None):
    """Times `steps` passes of the oracle's receive path over a bounded sample (ref_streams x frames) with all
    host cores.  Returns (samples/s, decoded bit/s, ms_per_step, sample description, cores, steps done)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as orc
    th = {1: sys.float_info.min, 2: 13, 3: 18, 4: 23, 5: 28, 6: 33}
    cn = tuple(sorted({1, 2, 3, 4, a.cnst}))
    mcs = [(th[c], (c, 1)) for c in cn]
    mcs[0] = (sys.float_info.min, mcs[0][1])
    cfg = orc.Config(fft_len=64, frame_len=20, alists=[ALIST], mcs=mcs, initial_mcs=cn.index(a.cnst), constellations=cn)
    S, F, cores = a.ref_streams, a.frames, n_cores()
    t = cfg.tx_batch(S, F, seed=0xD71, cnst=a.cnst, fec_id=1, snr_db=a.snr, nthreads=cores)
    secs, nbytes, done = [], 0, 0
    t_begin = time.perf_counter()
    for i in range(warmup + steps):
        r = cfg.rx_batch(t["iq"], t["taps"], t["meta"], S, F, nthreads=cores, want_llr=False, want_syms=False)
        if i >= warmup:
            secs.append(r["seconds"])
            nbytes = sum(len(b) for b in r["bytes"])
            done += 1
            if max_seconds and time.perf_counter() - t_begin > max_seconds:
                break
    per = sum(secs) / len(secs)
    samples = S * F * cfg.frame_samples
    sample = "%d streams x %d frames (%d IQ samples) per step, %d timed steps" % (S, F, samples, done)
    return samples / per, nbytes * 8 / per, per * 1e3, sample, cores, done


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sps, bps, ms, sample, cores, done = cpu_run(a, a.steps, a.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": sps / 1e6, "unit": UNIT, "n_gpus": a.gpus, "steps": done,
        "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (equaliser/demapper) + f64 (LDPC sum-product)", "data": "synthetic",
        "decoded_mbit_per_s": bps / 1e6,
        "config": {"workload": workload_name(a, a.streams), "sample": sample},
        "cpu_baseline": {"value": sps / 1e6, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": sps / 1e6, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "oracle port of the reference algorithm (GNU Radio 3.10 absent: the reference itself cannot be built)",
    }
    print(json.dumps(line), flush=True)


# ---- the GPU arm ------------------------------------------------------------------------------------------------
def product_config(a):
    import dtl_b200
    th = {1: sys.float_info.min, 2: 13, 3: 18, 4: 23, 5: 28, 6: 33}           # examples/config.json:3-6 thresholds (+5 dB per extension row)
    cn = tuple(sorted({1, 2, 3, 4, a.cnst}))
    mcs = [(th[c], (dtl_b200.constellation_type_t(c), "fec_1")) for c in cn]
    mcs[0] = (sys.float_info.min, mcs[0][1])
    return dtl_b200.ofdm_adaptive_rx_config(fec=True, fec_codes=(("fec_1", ALIST),), mcs=tuple(mcs),
                                            initial_mcs_id=cn.index(a.cnst))


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(kernel):
    """dram bytes per frame of `kernel` from the committed ncu --set full capture (profiles/traffic.json)"""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return float(json.load(f)[kernel]["dram_bytes_per_frame"])
    except Exception:
        return None


def run_b200(a):
    import numpy as np
    import torch
    import dtl_b200
    from dtl_b200 import synth, frame_result_dtype, tb_result_dtype

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and a.gpus > 1:                  # not under torchrun: launch ourselves the way the driver does
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(a.gpus),
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- libdtl_b200 has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda:%d" % local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"              # keep NCCL's version banner off stdout: one JSON line only
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def reduce(x, op):
        if dist is None:
            return x
        t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
        return float(t.item())

    S, F = a.streams, a.frames
    mode = dtl_b200.LDPC_LAYERED if a.ldpc_mode == "layered" else dtl_b200.LDPC_REFERENCE
    rx = dtl_b200.ofdm_receiver(product_config(a), device=local, ldpc_mode=mode)
    N, L, cap, fs = rx.fft_len, rx.frame_length, rx.frame_capacity, rx.frame_samples
    bps = dtl_b200.get_bits_per_symbol(a.cnst)
    batch = synth.make_batch(rx, S, F, a.cnst, fec_id=1, seed=0xD71 + 1000 * 4 + rank, snr_db=a.snr)
    iq, taps, meta = batch["iq"], batch["taps"], batch["meta"]
    bytes_cap, tb_cap = (rx.bytes_stride + 8) * (F + 1), 2 * F + 1

    def new_out(alloc):
        return dict(res=alloc((S * F,), frame_result_dtype), bytes=alloc((S, bytes_cap), np.uint8),
                    nbytes=alloc((S,), np.int32), tbs=alloc((S, tb_cap), tb_result_dtype), ntbs=alloc((S,), np.int32))

    out = new_out(lambda shape, dt: rx._new(dtl_b200.MEM_DEVICE, shape, np.dtype(dt)))
    stream = rx.torch_stream()

    def step():
        rx.reset(S)
        rx.rx(iq, taps, meta, S, F, out=out)

    for _ in range(max(a.warmup, 3)):
        step()
    rx.sync()
    # ---- the decoded bytes of the last warm-up step must be the transmitted payload (CRC-passing TBs, in order)
    nby = out["nbytes"].cpu().numpy()
    got = out["bytes"].cpu().numpy()
    ntb = out["ntbs"].cpu().numpy()
    tbs = out["tbs"].cpu().numpy().view(tb_result_dtype).reshape(S, tb_cap)
    crc_ok = sum(int(tbs[s, :ntb[s]]["crc_ok"].sum()) for s in range(S))
    n_tbs = int(ntb.sum())
    pm = batch["tb_bytes"]
    mism = 0
    for s in range(S):
        ok = tbs[s, :ntb[s]]["crc_ok"].astype(bool)
        if ntb[s] != batch["n_done"]:
            mism += 1
            continue
        want = batch["payload"][s].reshape(-1, pm)[ok].reshape(-1)
        mism += int(not np.array_equal(got[s, :nby[s]], want))
    if mism:
        raise SystemExit("bench.py: decoded payload differs from the transmitted payload on %d streams" % mism)
    decoded_bits = float(nby.sum()) * 8.0

    # ---- device-resident timing: CUDA events on the library's stream --------------------------------------------
    # Two passes of the same K steps.  (1) The timed region proper, stage profiling OFF: the library then runs its small
    # latency-bound kernels (TB planning, SNR tags, per-stream feedback state) on its side stream beside the heavy ones.
    # (2) The same steps with stage profiling ON (everything serial on the main stream, one CUDA-event pair per stage) for
    # the per-stage table and the roofline of the dominant HBM-bound kernel.
    rx.profile(False)
    l0 = rx.launch_count
    sampler = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.start()
    e0.record(stream)
    for _ in range(a.steps):
        step()
    e1.record(stream)
    rx.sync()
    barrier()
    clocks = sampler.result()
    ms_total = reduce(e0.elapsed_time(e1), "MAX")
    launches = rx.launch_count - l0
    rx.profile(True)
    rx.stage_times(reset=True)
    e4, e5 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e4.record(stream)
    for _ in range(a.steps):
        step()
    e5.record(stream)
    rx.sync()
    barrier()
    ms_serial = reduce(e4.elapsed_time(e5), "MAX") / a.steps
    stages = rx.stage_times(reset=True)
    rx.profile(False)
    ms_step = ms_total / a.steps
    samples_all = reduce(S * F * fs, "SUM")
    bits_all = reduce(decoded_bits, "SUM")
    value = samples_all / (ms_step * 1e-3) / 1e6

    # ---- end to end through the C ABI with HOST buffers (H2D of the inputs + D2H of the results inside the timing)
    e2e = None
    if not a.no_e2e:
        h_iq = rx.host_alloc((S * F, fs), np.complex64)
        h_taps = rx.host_alloc((S * F, N), np.complex64)
        h_meta = rx.host_alloc((S * F,), dtl_b200.frame_meta_dtype)
        h_iq[:] = iq.cpu().numpy()
        h_taps[:] = taps.cpu().numpy()
        h_meta[:] = batch["meta_np"]
        h_out = new_out(rx.host_alloc)

        def step_host():
            rx.reset(S)
            rx.rx(h_iq, h_taps, h_meta, S, F, out=h_out)

        for _ in range(3):
            step_host()
        if not np.array_equal(h_out["nbytes"], nby) or not np.array_equal(h_out["bytes"][0, :nby[0]], got[0, :nby[0]]):
            raise SystemExit("bench.py: host-buffer path and device-buffer path disagree")
        k_e2e = max(3, min(a.steps, 10))
        barrier()
        t0 = time.perf_counter()
        for _ in range(k_e2e):
            step_host()
        rx.sync()
        t1 = time.perf_counter()
        barrier()
        ms_e2e = reduce((t1 - t0) * 1e3, "MAX") / k_e2e
        h2d = h_iq.nbytes + h_taps.nbytes + h_meta.nbytes
        d2h = sum(v.nbytes for v in h_out.values())
        e2e = {"value": samples_all / (ms_e2e * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e, "steps": k_e2e,
               "decoded_mbit_per_s": bits_all / (ms_e2e * 1e-3) / 1e6,
               "api": "dtl_b200_rx(mem=DTL_MEM_HOST) via dtl_b200.ofdm_receiver.rx, page-locked host buffers"}

    # ---- the receive path from IQ alone (header branch: channel estimation + header parsing on the GPU, then the same
    # payload branch on the tags it produced); device-resident, same clock as `value`
    from_iq = None
    if not a.no_from_iq:
        out2 = new_out(lambda shape, dt: rx._new(dtl_b200.MEM_DEVICE, shape, np.dtype(dt)))
        out2["meta"] = rx._new(dtl_b200.MEM_DEVICE, (S * F,), dtl_b200.frame_meta_dtype)
        out2["hres"] = rx._new(dtl_b200.MEM_DEVICE, (S * F,), dtl_b200.header_result_dtype)

        def step_iq():
            rx.reset(S)
            rx.rx_iq(iq, S, F, out=out2)

        for _ in range(3):
            step_iq()
        rx.sync()
        hdr_ok = int(out2["hres"].cpu().numpy().view(dtl_b200.header_result_dtype)["crc_ok"].sum())
        nby2 = out2["nbytes"].cpu().numpy()
        e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e2.record(stream)
        for _ in range(a.steps):
            step_iq()
        e3.record(stream)
        rx.sync()
        barrier()
        ms_iq = reduce(e2.elapsed_time(e3), "MAX") / a.steps
        rx.stage_times(reset=True)
        rx.profile(True)                          # the header branch on its own: a few profiled steps
        for _ in range(3):
            step_iq()
        rx.sync()
        hdr_ms = rx.stage_times(reset=True)["header"][0] / 3
        rx.profile(False)
        from_iq = {"value": samples_all / (ms_iq * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_iq,
                   "header_branch_ms_per_step": hdr_ms, "headers_parsed": hdr_ok, "frames": S * F,
                   "decoded_mbit_per_s": reduce(float(nby2.sum()) * 8.0, "SUM") / (ms_iq * 1e-3) / 1e6,
                   "api": "dtl_b200_rx_iq(mem=DTL_MEM_DEVICE): tags and channel taps come from the frames' own sync words and "
                          "headers instead of the caller"}

    # ---- TX FEC framer: user bytes -> chunks + tags for the same S x F frames, device-resident
    tx_framer = None
    fec_id, cap = 1, rx.frame_capacity
    if not a.no_from_iq:
        pay = 8 * F * ((cap * bps // 8) + 64)
        gen = torch.Generator(device=dev)
        gen.manual_seed(99 + rank)
        t_data = torch.randint(0, 256, (S, pay), generator=gen, device=dev, dtype=torch.uint8)
        t_c = torch.full((S * F,), int(a.cnst), device=dev, dtype=torch.int32)
        t_f = torch.full((S * F,), int(fec_id), device=dev, dtype=torch.int32)
        torch.cuda.synchronize(dev)
        for _ in range(3):
            rx.reset(S)
            fo = rx.fec_frame(t_data, t_c, t_f, S, F)
        rx.sync()
        used = float(fo["consumed"].sum().item())
        e6, e7 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e6.record(stream)
        for _ in range(a.steps):
            rx.reset(S)
            fo = rx.fec_frame(t_data, t_c, t_f, S, F)
        e7.record(stream)
        rx.sync()
        barrier()
        ms_tx = reduce(e6.elapsed_time(e7), "MAX") / a.steps
        tx_framer = {"ms_per_step": ms_tx, "frames": S * F, "payload_mbit_per_s": reduce(used * 8.0, "SUM") / (ms_tx * 1e-3) / 1e6,
                     "api": "dtl_b200_fec_frame(mem=DTL_MEM_DEVICE): CRC-32, TB split, LDPC encode, chunk packing, tags"}
        del t_data, fo

    if rank != 0:
        rx.close()
        if dist is not None:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant HBM-bound kernel + per-stage table ----------------------------------------------
    peak, peak_src = peaks()
    frames = S * F
    # FFT: the cyclic prefix is skipped in the load indexing (whole 128-byte groups never touched), so the bytes the stage
    # must move are N samples in + N bins out per symbol; BASELINE.md section 4's earlier 8*L*(2N+cp) counted the prefix.
    alg = {"fft": 8 * L * 2 * N, "eq_demap": 8 * N * (L + 2) + 4 * cap * bps + 32}
    stage_tab = {}
    for name, (ms, cnt) in stages.items():
        if cnt:
            per = ms / a.steps
            row = {"ms_per_step": per, "share": per / ms_serial, "launches_per_step": cnt / a.steps}
            if name in alg:
                row["alg_bytes_per_frame"] = alg[name]
                row["achieved_gbs"] = alg[name] * frames / (per * 1e-3) / 1e9
                row["frac_of_hbm_peak"] = row["achieved_gbs"] / peak
                if name == "fft":
                    row["survey_bytes_per_frame"] = 8 * L * (2 * N + rx.cp_len)     # SURVEY.md 8(d) counts the prefix as read
            stage_tab[name] = row
    eq_ms = stages["eq_demap"][0] / max(1, stages["eq_demap"][1])
    achieved = alg["eq_demap"] * frames / (eq_ms * 1e-3) / 1e9
    traffic_pf = ncu_traffic("eq_fast_kernel")
    roofline = {"kernel": "eq_fast_kernel + snr_tags_kernel (frame equaliser + serialiser + soft demapper; whole stage timed)", "bound": "hbm",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None if traffic_pf is None else traffic_pf * frames,
                "alg_bytes_per_launch": alg["eq_demap"] * frames, "ms_per_launch": eq_ms, "peak_source": peak_src}
    cpu = None
    if not a.no_cpu_baseline:
        sps, cbps, cms, sample, cores, done = cpu_run(a, 8, 1, max_seconds=25.0)
        cpu = {"value": sps / 1e6, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
               "decoded_mbit_per_s": cbps / 1e6, "ms_per_step": cms}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (FFT/equaliser/demapper) + %s" % ("f64 (LDPC sum-product, reference arithmetic)"
                                                         if a.ldpc_mode == "reference" else "f32 (LDPC layered min-sum)"),
        "data": "synthetic", "decoded_mbit_per_s": bits_all / (ms_step * 1e-3) / 1e6,
        "config": {"workload": workload_name(a, S), "ldpc_mode": a.ldpc_mode, "l2": "inputs larger than L2 (%.0f MB IQ + "
                   "%.0f MB LLR scratch per step vs 126 MB L2)" % (iq.numel() * 8 / 1e6, frames * rx.llr_stride * 4 / 1e6),
                   "tbs_per_step_per_gpu": n_tbs, "crc_ok_tbs": crc_ok, "verified": "decoded bytes == transmitted payload"},
        "e2e": e2e, "rx_from_iq": from_iq, "tx_fec_framer": tx_framer, "gpu_launches": int(launches), "ms_per_step_serial": ms_serial, "clocks": clocks, "roofline": roofline, "stages": stage_tab,
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    rx.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
