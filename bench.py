#!/usr/bin/env python
"""bench.py -- throughput of the adaptive-OFDM receive hot path (BASELINE.json metric: RX Msamples/s and decoded
Mbit/s at 1/2/4/8 B200, % of the HBM roofline, beside the host CPU).

  python bench.py [--gpus N] [--steps K] [--warmup W]            this repo's CUDA path (libdtl_b200.so)
  python bench.py --impl reference [...]                          the reference's CPU algorithm (oracle port) timed
                                                                  on the host cores, same metric / config

One step = one pass of the whole receive path (CP strip + FFT -> frame equaliser + serialiser + soft demapper ->
TB reassembly + LDPC BP decode + CRC-32 + byte packing) over one batch of `streams` independent streams x `frames`
consecutive frames each (a fresh set of receivers per step: per-stream state is re-initialised on the stream).
Headline workload = BASELINE.json configs[1]'s PHY (64 carriers, cp 16, 20 payload symbols/frame, LDPC n=300 k=152,
16-QAM at 22 dB, reference-arithmetic decoder) at configs[3]'s per-GPU multiplicity (1024 streams per B200); streams
are independent, so N GPUs run N disjoint shards with no collective on the data path (weak scaling).

Keys of the JSON line (one line, rank 0):
`value`        IQ samples consumed per second (incl. CP, sync and header symbols), inputs resident in HBM, CUDA events on
               the library's own stream, EXACTLY --steps steps, max over ranks.  `sustained`: the same over >= 1 s.
`e2e`          the same through the C ABI with HOST (page-locked) buffers: H2D of IQ/taps/tags and D2H of the decoded
               bytes + per-frame / per-TB records inside the timed region.  `e2e_from_iq`: dtl_b200_rx_iq with host IQ only.
`roofline`     the dominant HBM-bound kernel (fused equaliser + serialiser + soft demapper): algorithmic bytes per frame
                x frames per launch / its CUDA-event time, against MEASURED_PEAKS.json.
`roofline_ldpc`  the step-dominant stage (not HBM-bound): decoded bit/s, codewords/s, BP iterations/s from the stage clock;
               FP64-pipe and shared-memory figures from the committed ncu capture (profiles/traffic.json).
`layered`      the same workload with the layered normalised min-sum decoder (DTL_LDPC_LAYERED) + TB-level disagreement.
`configs`      BASELINE.json's other configurations: config 1 (no-FEC byte path), config 3 (per constellation BPSK..256QAM),
               config 5 (1024-point FFT, 256QAM + LDPC), each with its stage times and roofline fractions.
`latency`      S = 1, F = 1, DTL_MEM_HOST: what one work()-sized call costs (median microseconds).
`cpu_baseline` the oracle (CPU restatement of the reference algorithm; the reference itself needs GNU Radio 3.10, absent
               here) on a bounded sample of the same workload: all host cores and one thread, built -O3 -march=native.
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
BPS = {1: 1, 2: 2, 3: 3, 4: 4, 5: 6, 6: 8}
# SNR at which (nearly) every block of the constellation decodes (tests/test_gpu_configs.py: SNR_OK)
SNR_OK = {1: 16.0, 2: 22.0, 3: 27.0, 4: 31.0, 5: 46.0, 6: 54.0}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
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
    ap.add_argument("--no-configs", action="store_true", help="skip the config 1 / 3 / 5, layered and latency legs")
    ap.add_argument("--sustain-seconds", type=float, default=1.0)
    return ap.parse_args()


def workload_name(a, streams):
    return ("ofdm_adaptive RX, 64-carrier PHY (cp 16, 20 syms/frame), %s @ %g dB + LDPC n=300 k=152 (config_fec), "
            "%d streams x %d frames per GPU per step" % (CNST_NAMES[a.cnst], a.snr, streams, a.frames))


def config_dict(a, streams):
    """identical on both arms (the driver compares them)"""
    return {"workload": workload_name(a, streams), "ldpc_mode": a.ldpc_mode,
            "l2": "inputs larger than L2 (%.0f MB IQ + %.0f MB LLR scratch per step vs 126 MB L2)"
                  % (streams * a.frames * 1920 * 8 / 1e6, streams * a.frames * 960 * 4 * 4 / 1e6)}


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
            time.sleep(0.02)

    def result(self):
        self.stop_flag = True
        if not self.ok or not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max, "reasons": [], "samples": 0}
        s = sorted(self.sm)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons), "samples": len(s)}


# ---- the CPU arm: oracle port of the reference algorithm (the ONLY place bench.py touches oracle/) ----------------
def cpu_run(a, steps, warmup, max_seconds=None, threads=None, streams=None):
    """Times `steps` passes of the oracle's receive path over a bounded sample (streams x frames) with `threads` host
    threads (default: all cores).  Median over the timed steps.  Returns a dict."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as orc
    build = orc.build_native()                      # -O3 -march=native on THIS host
    th = {1: sys.float_info.min, 2: 13, 3: 18, 4: 23, 5: 28, 6: 33}
    cn = tuple(sorted({1, 2, 3, 4, a.cnst}))
    mcs = [(th[c], (c, 1)) for c in cn]
    mcs[0] = (sys.float_info.min, mcs[0][1])
    cfg = orc.Config(fft_len=64, frame_len=20, alists=[ALIST], mcs=mcs, initial_mcs=cn.index(a.cnst), constellations=cn)
    cores = n_cores()
    threads = threads or cores
    S, F = streams or a.ref_streams, a.frames
    t = cfg.tx_batch(S, F, seed=0xD71, cnst=a.cnst, fec_id=1, snr_db=a.snr, nthreads=cores)
    secs, nbytes = [], 0
    t_begin = time.perf_counter()
    for i in range(warmup + steps):
        r = cfg.rx_batch(t["iq"], t["taps"], t["meta"], S, F, nthreads=threads, want_llr=False, want_syms=False)
        if i >= warmup:
            secs.append(r["seconds"])
            nbytes = sum(len(b) for b in r["bytes"])
            if max_seconds and len(secs) >= 3 and time.perf_counter() - t_begin > max_seconds:
                break
    secs.sort()
    per = secs[len(secs) // 2]
    samples = S * F * cfg.frame_samples
    return {"value": samples / per / 1e6, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "%d streams x %d frames (%d IQ samples) per step, median of %d timed steps" % (S, F, samples, len(secs)),
            "decoded_mbit_per_s": nbytes * 8 / per / 1e6, "ms_per_step": per * 1e3, "steps": len(secs), "build": build,
            "host_cores": cores}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(3, a.steps)
    cpu = cpu_run(a, steps, a.warmup, max_seconds=150.0)
    one = cpu_run(a, 3, 1, max_seconds=30.0, threads=1, streams=max(1, a.ref_streams // 16))
    cpu["single_thread"] = {k: one[k] for k in ("value", "unit", "cores", "sample", "ms_per_step", "decoded_mbit_per_s")}
    line = {
        "impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": cpu["steps"],
        "warmup": a.warmup, "ms_per_step": cpu["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (equaliser/demapper) + f64 (LDPC sum-product)", "data": "synthetic",
        "decoded_mbit_per_s": cpu["decoded_mbit_per_s"],
        "config": config_dict(a, a.streams),
        "cpu_baseline": cpu,
        "e2e": {"value": cpu["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "oracle port of the reference algorithm on a bounded sample of the workload (rate-normalised); "
                "GNU Radio 3.10 is absent, the reference itself cannot be built",
    }
    print(json.dumps(line), flush=True)


# ---- the GPU arm ------------------------------------------------------------------------------------------------
def product_config(cnst, fec=True, fft_len=64, frame_len=20, lut="config"):
    """ofdm_adaptive_rx_config for one leg.  lut "config": examples/config.json:3-6 thresholds (13/18/23 dB; +5 dB per extension
    row); "config_fec": examples/config_fec.json:4-7 (11/16/21 dB)."""
    import dtl_b200
    th = ({1: sys.float_info.min, 2: 13, 3: 18, 4: 23, 5: 28, 6: 33} if lut == "config" else
          {1: sys.float_info.min, 2: 11, 3: 16, 4: 21, 5: 26, 6: 31})
    cn = tuple(sorted({1, 2, 3, 4, cnst}))
    mcs = [(th[c], (dtl_b200.constellation_type_t(c), "fec_1" if fec else "no_fec")) for c in cn]
    mcs[0] = (sys.float_info.min, mcs[0][1])
    kw = {}
    if fft_len != 64:
        occ, pil = dtl_b200.scaled_phy_carriers(fft_len)
        kw = dict(fft_len=fft_len, cp_len=fft_len // 4, occupied_carriers=occ, pilot_carriers=pil)
    return dtl_b200.ofdm_adaptive_rx_config(frame_length=frame_len, fec=fec, fec_codes=(("fec_1", ALIST),) if fec else (),
                                            mcs=tuple(mcs), initial_mcs_id=cn.index(cnst), **kw)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_facts():
    """figures that only a profiler sees, from the committed ncu --set full captures (profiles/traffic.json)"""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def alg_bytes(N, L, cap, bps, fec):
    """algorithmic HBM bytes per frame (DESIGN.md section 4 / BASELINE.md section 4)"""
    fft = 8 * L * 2 * N
    eq = 8 * N * (L + 2) + 4 * cap * bps + 32 if fec else 8 * N * (L + 1) + cap * bps // 8 + 32
    return {"fft": fft, "eq_demap": eq}


def numa_bind(torch, local):
    """Pin this rank's threads to the CPUs NVML names as closest to its GPU (the socket / NUMA node its PCIe root hangs off),
    so that the page-locked staging buffers of the e2e legs are allocated on that node and the feeding thread runs there."""
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(local).uuid)
        uuid = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
        before = len(os.sched_getaffinity(0))
        pynvml.nvmlDeviceSetCpuAffinity(h)
        cpus = sorted(os.sched_getaffinity(0))
        node = None
        try:
            bus = pynvml.nvmlDeviceGetPciInfo(h).busId
            bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
            with open("/sys/bus/pci/devices/%s/numa_node" % bus[-12:]) as f:
                node = int(f.read().strip())
        except Exception:
            pass
        return {"bound": True, "cpus": len(cpus), "cpus_before": before, "first_cpu": cpus[0], "node": node}
    except Exception as e:                                   # no NVML / not permitted: run unpinned and say so
        return {"bound": False, "why": str(e)[:80]}


def h2d_ceiling(ctx, mb=256, reps=12):
    """What the box gives every GPU when ALL ranks copy page-locked host memory to their device at the same time (the e2e
    legs cannot beat this): barrier, `reps` copies of `mb` MB per rank -- once as one copy on one stream, once as two halves on
    two streams (what the chunked host path does) -- slowest rank; the better of the two layouts is the ceiling."""
    torch = ctx.torch
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    h.fill_(1)
    d = torch.empty(n, dtype=torch.uint8, device=ctx.dev)
    d.copy_(h, non_blocking=True)
    side = torch.cuda.Stream(device=ctx.dev)
    half = n // 2

    def run(two):
        """per-copy times between events; the FASTEST copy of the rank counts (a ceiling is the best the link does), the slowest
        rank's figure is reported"""
        ctx.barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
        ev[0].record()
        for r in range(reps):
            if two:
                side.wait_stream(torch.cuda.current_stream(ctx.dev))
                with torch.cuda.stream(side):
                    d[half:].copy_(h[half:], non_blocking=True)
                d[:half].copy_(h[:half], non_blocking=True)
                torch.cuda.current_stream(ctx.dev).wait_stream(side)
            else:
                d.copy_(h, non_blocking=True)
            ev[r + 1].record()
        torch.cuda.synchronize(ctx.dev)
        ms = min(ev[r].elapsed_time(ev[r + 1]) for r in range(reps))
        return ctx.reduce(ms, "MAX"), ctx.reduce(ms, "SUM") / ctx.world, ctx.reduce(ev[0].elapsed_time(ev[reps]) / reps, "MAX")

    w1, m1, a1 = run(False)
    w2, m2, a2 = run(True)
    worst, mean = (w1, m1) if w1 <= w2 else (w2, m2)
    ctx.barrier()
    del h, d
    gbs = lambda ms: n / (ms * 1e-3) / 1e9
    return {"h2d_gbs_per_gpu": gbs(worst), "h2d_gbs_per_gpu_mean": gbs(mean), "one_stream_gbs": gbs(w1), "two_streams_gbs": gbs(w2),
            "h2d_gbs_per_gpu_sustained": gbs(min(a1, a2)),     # all `reps` copies back to back, slowest rank: what a steady feed sees
            "concurrent_gpus": ctx.world, "mb_per_copy": mb, "copies": reps,
            "note": "all ranks copying at once from page-locked memory allocated after NUMA binding; fastest copy of the slowest rank; better of one copy on one stream / two halves on two streams"}


class Ctx:
    """what every leg needs: device, barrier / reductions over the ranks, the event-timed loop"""

    def __init__(self, a):
        import torch
        self.torch, self.a = torch, a
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda:%d" % self.local)
        self.cpus_all = os.sched_getaffinity(0)
        self.numa = numa_bind(torch, self.local)       # before any page-locked buffer exists: first touch is NUMA-local then
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
                os.environ["NCCL_DEBUG"] = "WARN"          # keep NCCL's banner off stdout: one JSON line only
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        self.peak, self.peak_src = peaks()

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize(self.dev)

    def reduce(self, x, op):
        if self.dist is None:
            return x
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=getattr(self.dist.ReduceOp, op))
        return float(t.item())

    def timed(self, rx, step, steps, clocks=False):
        """`steps` calls of step() between two CUDA events on the library's stream, barrier + synchronize on both sides;
        returns (ms per step, max over ranks[, clocks])"""
        stream = rx.torch_stream()
        e0, e1 = self.torch.cuda.Event(enable_timing=True), self.torch.cuda.Event(enable_timing=True)
        sampler = ClockSampler(self.local) if clocks else None
        self.barrier()
        if sampler:
            sampler.start()
        e0.record(stream)
        for _ in range(steps):
            step()
        e1.record(stream)
        rx.sync()
        self.barrier()
        ms = self.reduce(e0.elapsed_time(e1), "MAX") / steps
        return (ms, sampler.result()) if sampler else ms


def new_out(rx, S, F, alloc, np):
    from dtl_b200 import frame_result_dtype, tb_result_dtype
    bytes_cap, tb_cap = (rx.bytes_stride + 8) * (F + 1), 2 * F + 1
    return dict(res=alloc((S * F,), frame_result_dtype), bytes=alloc((S, bytes_cap), np.uint8),
                nbytes=alloc((S,), np.int32), tbs=alloc((S, tb_cap), tb_result_dtype), ntbs=alloc((S,), np.int32),
                frame_crc=alloc((S * F,), np.int32))


def tb_table(out, S, np):
    """(ok flags per TB concatenated over the streams, sum of avg_it, TB count) of a device-side result set"""
    from dtl_b200 import tb_result_dtype
    ntb = out["ntbs"].cpu().numpy()
    tbs = out["tbs"].cpu().numpy().view(tb_result_dtype).reshape(S, -1)
    ok = np.concatenate([tbs[s, :ntb[s]]["crc_ok"] for s in range(S)]).astype(bool)
    its = int(sum(int(tbs[s, :ntb[s]]["avg_it"].sum()) for s in range(S)))
    return ok, its, int(ntb.sum())


def check_payload(out, batch, S, fec, np):
    """the decoded bytes must be the transmitted payload (CRC-passing TBs / frames, in order); returns decoded bits"""
    from dtl_b200 import tb_result_dtype
    nby, got = out["nbytes"].cpu().numpy(), out["bytes"].cpu().numpy()
    mism = 0
    if fec:
        ntb = out["ntbs"].cpu().numpy()
        tbs = out["tbs"].cpu().numpy().view(tb_result_dtype).reshape(S, -1)
        pm = batch["tb_bytes"]
        for s in range(S):
            ok = tbs[s, :ntb[s]]["crc_ok"].astype(bool)
            if ntb[s] != batch["n_done"]:
                mism += 1
                continue
            want = batch["payload"][s].reshape(-1, pm)[ok].reshape(-1)
            mism += int(not np.array_equal(got[s, :nby[s]], want))
        bits = float(nby.sum()) * 8.0
    else:
        crc = out["frame_crc"].cpu().numpy().reshape(S, -1)
        per = batch["payload"][0].size // crc.shape[1]
        good = 0
        for s in range(S):
            g = got[s, :nby[s]].reshape(crc.shape[1], per)
            w = batch["payload"][s].reshape(crc.shape[1], per)
            okf = crc[s] == 1
            mism += int(not np.array_equal(g[okf], w[okf]))
            good += int(okf.sum())
        bits = good * per * 8.0
    if mism:
        raise SystemExit("bench.py: decoded payload differs from the transmitted payload on %d streams" % mism)
    return bits


def stage_table(ctx, rx, step, steps, alg, frames):
    """the same steps with stage profiling ON (serial on the main stream, one CUDA-event pair per stage)"""
    rx.profile(True)
    rx.stage_times(reset=True)
    ms_serial = ctx.timed(rx, step, steps)
    stages = rx.stage_times(reset=True)
    rx.profile(False)
    tab = {}
    for name, (ms, cnt) in stages.items():
        if cnt:
            per = ms / steps
            row = {"ms_per_step": per, "share": per / ms_serial, "launches_per_step": cnt / steps}
            if name in alg:
                row["alg_bytes_per_frame"] = alg[name]
                row["achieved_gbs"] = alg[name] * frames / (per * 1e-3) / 1e9
                row["frac_of_hbm_peak"] = row["achieved_gbs"] / ctx.peak
            tab[name] = row
    return tab, ms_serial


def leg_config(ctx, dtl, np, name, cnst, fec, S, F, steps, fft_len=64, frame_len=20, snr=None, mode=None):
    """one of BASELINE.json's other configurations: device-resident value + stage table with roofline fractions"""
    from dtl_b200 import synth
    mode = dtl.LDPC_REFERENCE if mode is None else mode
    rx = dtl.ofdm_receiver(product_config(cnst, fec, fft_len, frame_len), device=ctx.local, ldpc_mode=mode)
    snr = SNR_OK[cnst] if snr is None else snr
    b = synth.make_batch(rx, S, F, cnst, fec_id=1 if fec else 0, seed=0xD71 + 31 * cnst + fft_len + ctx.rank, snr_db=snr)
    out = new_out(rx, S, F, lambda shape, dt: rx._new(dtl.MEM_DEVICE, shape, np.dtype(dt)), np)

    def step():
        rx.reset(S)
        rx.rx(b["iq"], b["taps"], b["meta"], S, F, out=out)

    for _ in range(3):
        step()
    rx.sync()
    bits = check_payload(out, b, S, fec, np)
    ms = ctx.timed(rx, step, steps)
    alg = alg_bytes(rx.fft_len, rx.frame_length, rx.frame_capacity, BPS[cnst], fec)
    tab, ms_serial = stage_table(ctx, rx, step, max(3, steps // 2), alg, S * F)
    samples = ctx.reduce(S * F * rx.frame_samples, "SUM")
    leg = {"workload": "%s: %d-point FFT, %d syms/frame, %s @ %g dB, %s, %d streams x %d frames per GPU per step"
                       % (name, rx.fft_len, rx.frame_length, CNST_NAMES[cnst], snr, "LDPC n=300 k=152" if fec else "no FEC", S, F),
           "value": samples / (ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms, "ms_per_step_serial": ms_serial,
           "decoded_mbit_per_s": ctx.reduce(bits, "SUM") / (ms * 1e-3) / 1e6, "steps": steps, "stages": tab,
           "eq_frac_of_hbm_peak": tab.get("eq_demap", {}).get("frac_of_hbm_peak"),
           "fft_frac_of_hbm_peak": tab.get("fft", {}).get("frac_of_hbm_peak")}
    rx.close()
    del b, out
    ctx.torch.cuda.empty_cache()
    return leg


def run_b200(a):
    import numpy as np
    import torch
    import dtl_b200
    from dtl_b200 import synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 and a.gpus > 1:                  # not under torchrun: launch ourselves the way the driver does
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(a.gpus),
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- libdtl_b200 has no CPU path (use --impl reference for the CPU arm)")
    ctx = Ctx(a)
    rank, local, dev = ctx.rank, ctx.local, ctx.dev
    S, F = a.streams, a.frames
    mode = dtl_b200.LDPC_LAYERED if a.ldpc_mode == "layered" else dtl_b200.LDPC_REFERENCE
    rx = dtl_b200.ofdm_receiver(product_config(a.cnst), device=local, ldpc_mode=mode)
    N, L, cap, fs = rx.fft_len, rx.frame_length, rx.frame_capacity, rx.frame_samples
    bps = BPS[a.cnst]
    batch = synth.make_batch(rx, S, F, a.cnst, fec_id=1, seed=0xD71 + 1000 * 4 + rank, snr_db=a.snr)
    iq, taps, meta = batch["iq"], batch["taps"], batch["meta"]
    dev_alloc = lambda shape, dt: rx._new(dtl_b200.MEM_DEVICE, shape, np.dtype(dt))
    out = new_out(rx, S, F, dev_alloc, np)

    def step():
        rx.reset(S)
        rx.rx(iq, taps, meta, S, F, out=out)

    for _ in range(max(a.warmup, 3)):
        step()
    rx.sync()
    decoded_bits = check_payload(out, batch, S, True, np)
    ok_ref, its_ref, n_tbs = tb_table(out, S, np)
    nby = out["nbytes"].cpu().numpy()
    got0 = out["bytes"].cpu().numpy()

    # ---- device-resident timing: CUDA events on the library's stream --------------------------------------------
    # (1) The timed region proper, EXACTLY --steps steps, stage profiling OFF: the library then runs its small latency-bound
    # kernels on its side stream beside the heavy ones.  (2) The same loop for >= --sustain-seconds (power / clock behaviour
    # of the FP64-heavy decoder over a window the clock sampler sees).  (3) Stage profiling ON (serial) for the per-stage table.
    rx.profile(False)
    l0 = rx.launch_count
    ms_step, clocks = ctx.timed(rx, step, a.steps, clocks=True)
    launches = rx.launch_count - l0
    n_sus = max(a.steps, int(a.sustain_seconds * 1e3 / ms_step) + 1)
    ms_sus, clocks_sus = ctx.timed(rx, step, n_sus, clocks=True)
    alg = alg_bytes(N, L, cap, bps, True)
    stage_tab, ms_serial = stage_table(ctx, rx, step, min(a.steps, 20), alg, S * F)
    samples_all = ctx.reduce(S * F * fs, "SUM")
    bits_all = ctx.reduce(decoded_bits, "SUM")
    value = samples_all / (ms_step * 1e-3) / 1e6
    sustained = {"value": samples_all / (ms_sus * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_sus, "steps": n_sus,
                 "seconds": ms_sus * n_sus * 1e-3, "clocks": clocks_sus}

    # ---- end to end through the C ABI with HOST buffers (H2D of the inputs + D2H of the results inside the timing)
    e2e = e2e_iq = ceiling = None
    if not a.no_e2e:
        ceiling = h2d_ceiling(ctx)
        h_iq = rx.host_alloc((S * F, fs), np.complex64)
        h_taps = rx.host_alloc((S * F, N), np.complex64)
        h_meta = rx.host_alloc((S * F,), dtl_b200.frame_meta_dtype)
        h_iq[:] = iq.cpu().numpy()
        h_taps[:] = taps.cpu().numpy()
        h_meta[:] = batch["meta_np"]
        h_out = new_out(rx, S, F, rx.host_alloc, np)

        def step_host():
            rx.reset(S)
            rx.rx(h_iq, h_taps, h_meta, S, F, out=h_out)

        for _ in range(3):
            step_host()
        if not np.array_equal(h_out["nbytes"], nby) or not np.array_equal(h_out["bytes"][0, :nby[0]], got0[0, :nby[0]]):
            raise SystemExit("bench.py: host-buffer path and device-buffer path disagree")
        k_e2e = max(3, min(a.steps, 10))

        def wall(fn):
            ctx.barrier()
            t0 = time.perf_counter()
            for _ in range(k_e2e):
                fn()
            rx.sync()
            t1 = time.perf_counter()
            ctx.barrier()
            return ctx.reduce((t1 - t0) * 1e3, "MAX") / k_e2e

        ms_e2e = wall(step_host)
        h2d = h_iq.nbytes + h_taps.nbytes + h_meta.nbytes
        d2h = sum(v.nbytes for v in h_out.values())
        e2e = {"value": samples_all / (ms_e2e * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e, "steps": k_e2e,
               "h2d_gbs": h2d / (ms_e2e * 1e-3) / 1e9, "h2d_gbs_per_gpu_ceiling": ceiling["h2d_gbs_per_gpu"], "h2d_gbs_per_gpu_sustained": ceiling["h2d_gbs_per_gpu_sustained"],
               "decoded_mbit_per_s": bits_all / (ms_e2e * 1e-3) / 1e6,
               "api": "dtl_b200_rx(mem=DTL_MEM_HOST) via dtl_b200.ofdm_receiver.rx, page-locked host buffers"}
        # the same from IQ alone: no taps / tags uploaded, the header branch derives them on the GPU
        h_out2 = new_out(rx, S, F, rx.host_alloc, np)
        h_out2["meta"] = rx.host_alloc((S * F,), dtl_b200.frame_meta_dtype)
        h_out2["hres"] = rx.host_alloc((S * F,), dtl_b200.header_result_dtype)

        def step_host_iq():
            rx.reset(S)
            rx.rx_iq(h_iq, S, F, out=h_out2)

        for _ in range(2):
            step_host_iq()
        ms_e2e_iq = wall(step_host_iq)
        d2h2 = sum(v.nbytes for v in h_out2.values())
        e2e_iq = {"value": samples_all / (ms_e2e_iq * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(h_iq.nbytes),
                  "d2h_bytes_per_step": int(d2h2), "ms_per_step": ms_e2e_iq, "steps": k_e2e,
                  "h2d_gbs": h_iq.nbytes / (ms_e2e_iq * 1e-3) / 1e9, "h2d_gbs_per_gpu_ceiling": ceiling["h2d_gbs_per_gpu"], "h2d_gbs_per_gpu_sustained": ceiling["h2d_gbs_per_gpu_sustained"],
                  "decoded_mbit_per_s": ctx.reduce(float(h_out2["nbytes"].sum()) * 8.0, "SUM") / (ms_e2e_iq * 1e-3) / 1e6,
                  "api": "dtl_b200_rx_iq(mem=DTL_MEM_HOST): host IQ only; channel taps and stream tags come from the frames' own "
                         "sync words and headers on the GPU"}

    # ---- the receive path from IQ alone, device-resident, same clock as `value`
    from_iq = None
    if not a.no_from_iq:
        out2 = new_out(rx, S, F, dev_alloc, np)
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
        ms_iq = ctx.timed(rx, step_iq, min(a.steps, 20))
        rx.stage_times(reset=True)
        rx.profile(True)                          # the header branch on its own: a few profiled steps
        for _ in range(3):
            step_iq()
        rx.sync()
        hdr_ms = rx.stage_times(reset=True)["header"][0] / 3
        rx.profile(False)
        from_iq = {"value": samples_all / (ms_iq * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_iq,
                   "header_branch_ms_per_step": hdr_ms, "headers_parsed": hdr_ok, "frames": S * F,
                   "decoded_mbit_per_s": ctx.reduce(float(nby2.sum()) * 8.0, "SUM") / (ms_iq * 1e-3) / 1e6,
                   "api": "dtl_b200_rx_iq(mem=DTL_MEM_DEVICE): tags and channel taps come from the frames' own sync words and "
                          "headers instead of the caller"}
        del out2

    # ---- continuous IQ: the synchroniser in front of everything.  The batch's frames of a stream lie back to
    # back in memory, so [S, F * frame_samples] IS a continuous stream; every call feeds the next F frames' worth of samples
    # (the same samples again: the stream goes on) and gets the frames that were completed by them.
    from_stream = None
    if not a.no_from_iq:
        Fc = F + 2
        out3 = new_out(rx, S, Fc, dev_alloc, np)
        out3["meta"] = rx._new(dtl_b200.MEM_DEVICE, (S * Fc,), dtl_b200.frame_meta_dtype)
        out3["hres"] = rx._new(dtl_b200.MEM_DEVICE, (S * Fc,), dtl_b200.header_result_dtype)
        out3["n_frames"] = rx._new(dtl_b200.MEM_DEVICE, (S,), np.int32)
        out3["frame_start"] = rx._new(dtl_b200.MEM_DEVICE, (S * Fc,), np.int64)
        rx.reset(S)

        def step_stream():
            rx.rx_stream(iq, S, Fc, out=out3)

        for _ in range(3):
            step_stream()
        rx.sync()
        nfr = out3["n_frames"].cpu().numpy()
        hdr3 = out3["hres"].cpu().numpy().view(dtl_b200.header_result_dtype)["crc_ok"].reshape(S, Fc)
        ms_st = ctx.timed(rx, step_stream, min(a.steps, 20))
        rx.stage_times(reset=True)
        rx.profile(True)
        for _ in range(3):
            step_stream()
        rx.sync()
        st3 = rx.stage_times(reset=True)
        sync_ms = st3["sync"][0] / 3
        rx.profile(False)
        from_stream = {"value": samples_all / (ms_st * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_st,
                       "sync_front_end_ms_per_step": sync_ms, "frames_found_per_call": int(nfr.sum()), "frames_sent_per_call": S * F,
                       "headers_parsed_per_call": int(hdr3.sum()),
                       "stage_ms_per_step": {k: v[0] / 3 for k, v in st3.items() if v[0] > 0},
                       "api": "dtl_b200_rx_stream(mem=DTL_MEM_DEVICE): continuous IQ; Schmidl & Cox timing + fine CFO, trigger "
                              "regulariser, frame cut, then the path from IQ alone"}
        del out3
        rx.reset(S)

    # ---- TX FEC framer: user bytes -> chunks + tags for the same S x F frames, device-resident
    tx_framer = None
    if not a.no_from_iq:
        pay = 8 * F * ((cap * bps // 8) + 64)
        gen = torch.Generator(device=dev)
        gen.manual_seed(99 + rank)
        t_data = torch.randint(0, 256, (S, pay), generator=gen, device=dev, dtype=torch.uint8)
        t_c = torch.full((S * F,), int(a.cnst), device=dev, dtype=torch.int32)
        t_f = torch.full((S * F,), 1, device=dev, dtype=torch.int32)
        torch.cuda.synchronize(dev)
        fo = None

        def step_tx():
            nonlocal fo
            rx.reset(S)
            fo = rx.fec_frame(t_data, t_c, t_f, S, F)

        for _ in range(3):
            step_tx()
        rx.sync()
        used = float(fo["consumed"].sum().item())
        ms_tx = ctx.timed(rx, step_tx, min(a.steps, 20))
        tx_framer = {"ms_per_step": ms_tx, "frames": S * F, "payload_mbit_per_s": ctx.reduce(used * 8.0, "SUM") / (ms_tx * 1e-3) / 1e6,
                     "api": "dtl_b200_fec_frame(mem=DTL_MEM_DEVICE): CRC-32, TB split, LDPC encode, chunk packing, tags"}
        del t_data, fo

    # ---- the decoder north_star names: layered normalised min-sum, same input, TB-level disagreement with reference mode
    layered = latency = None
    configs = {}
    if not a.no_configs:
        other = dtl_b200.LDPC_REFERENCE if mode == dtl_b200.LDPC_LAYERED else dtl_b200.LDPC_LAYERED
        rx2 = dtl_b200.ofdm_receiver(product_config(a.cnst), device=local, ldpc_mode=other)
        out_l = new_out(rx2, S, F, lambda shape, dt: rx2._new(dtl_b200.MEM_DEVICE, shape, np.dtype(dt)), np)

        def step_l():
            rx2.reset(S)
            rx2.rx(iq, taps, meta, S, F, out=out_l)

        for _ in range(3):
            step_l()
        rx2.sync()
        bits_l = check_payload(out_l, batch, S, True, np)
        ok_l, its_l, _ = tb_table(out_l, S, np)
        ms_l = ctx.timed(rx2, step_l, min(a.steps, 50))
        tab_l, _ = stage_table(ctx, rx2, step_l, 10, alg, S * F)
        layered = {"ldpc_mode": "layered" if other == dtl_b200.LDPC_LAYERED else "reference",
                   "value": samples_all / (ms_l * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms_l,
                   "decoded_mbit_per_s": ctx.reduce(bits_l, "SUM") / (ms_l * 1e-3) / 1e6,
                   "ldpc_stage_ms": tab_l.get("ldpc", {}).get("ms_per_step"),
                   "speedup_of_ldpc_stage_vs_headline_mode": stage_tab["ldpc"]["ms_per_step"] / tab_l["ldpc"]["ms_per_step"],
                   "tbs": int(ok_l.size), "crc_ok_tbs": int(ok_l.sum()), "crc_ok_tbs_headline_mode": int(ok_ref.sum()),
                   "tbs_only_headline_mode_decodes": int((ok_ref & ~ok_l).sum()),
                   "tbs_only_this_mode_decodes": int((ok_l & ~ok_ref).sum()),
                   "note": "every CRC-passing TB of either mode carries the transmitted bytes (checked)"}
        rx2.close()
        del out_l
        # ---- what one work()-sized call costs: S = 1, F = 1, host buffers (a drop-in shim's default granularity)
        rx3 = dtl_b200.ofdm_receiver(product_config(a.cnst), device=local, ldpc_mode=mode)
        l_iq = rx3.host_alloc((1, fs), np.complex64)
        l_taps = rx3.host_alloc((1, N), np.complex64)
        l_meta = rx3.host_alloc((1,), dtl_b200.frame_meta_dtype)
        hq = iq[:F].cpu().numpy()
        ht = taps[:F].cpu().numpy()
        l_out = new_out(rx3, 1, 1, rx3.host_alloc, np)
        rx3.reset(1)
        lat = []
        for rep in range(8):
            for j in range(F):
                l_iq[0], l_taps[0], l_meta[0] = hq[j], ht[j], batch["meta_np"][j]
                t0 = time.perf_counter()
                rx3.rx(l_iq, l_taps, l_meta, 1, 1, out=l_out)
                lat.append((time.perf_counter() - t0) * 1e6)
        lat = sorted(lat[F:])
        latency = {"median_us": lat[len(lat) // 2], "p10_us": lat[len(lat) // 10], "p90_us": lat[9 * len(lat) // 10],
                   "calls": len(lat), "api": "dtl_b200_rx(S=1, F=1, mem=DTL_MEM_HOST): one tagged frame per call, consecutive frames of one stream"}
        rx3.close()
        # ---- BASELINE.json's other configurations
        k = max(5, min(a.steps, 20))
        configs["config1_no_fec_16qam"] = leg_config(ctx, dtl_b200, np, "config 1", 4, False, S, F, k)
        for c in (1, 2, 3, 4, 5, 6):
            configs["config3_%s" % CNST_NAMES[c].lower()] = leg_config(ctx, dtl_b200, np, "config 3", c, True, S, 16, k)
        configs["config5_fft1024_256qam"] = leg_config(ctx, dtl_b200, np, "config 5", 6, True, 256, 32, max(5, k // 2),
                                                       fft_len=1024, frame_len=5)

    if rank != 0:
        rx.close()
        if ctx.dist is not None:
            ctx.dist.destroy_process_group()
        return
    # ---- roofline of the dominant HBM-bound kernel + the LDPC stage -----------------------------------------------------
    peak, peak_src = ctx.peak, ctx.peak_src
    frames = S * F
    facts = ncu_facts()
    eq = stage_tab["eq_demap"]
    eq_ms = eq["ms_per_step"]
    achieved = alg["eq_demap"] * frames / (eq_ms * 1e-3) / 1e9
    tr = facts.get("eq_fast_kernel", {}).get("dram_bytes_per_frame")
    roofline = {"kernel": "eq_fast_kernel (frame equaliser + serialiser + soft demapper + sign words; CUDA events around the kernel on the library stream)",
                "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None if tr is None else tr * frames,
                "alg_bytes_per_launch": alg["eq_demap"] * frames, "ms_per_launch": eq_ms, "peak_source": peak_src}
    ld_ms = stage_tab["ldpc"]["ms_per_step"]
    cws = n_tbs * 13 if a.cnst == 4 else None
    bp = facts.get("ldpc_bp_ref_kernel" if a.ldpc_mode == "reference" else "ldpc_bp_lay_kernel", {})
    roofline_ldpc = {"stage": "ldpc_screen_kernel + bp_gather_kernel + ldpc_bp_%s_kernel" % ("ref" if a.ldpc_mode == "reference" else "lay"),
                     "bound": "FP64 pipe / warp issue (not HBM)" if a.ldpc_mode == "reference" else "shared-memory latency / warp issue",
                     "ms_per_launch": ld_ms, "share_of_step": stage_tab["ldpc"]["share"],
                     "decoded_gbit_per_s": decoded_bits / (ld_ms * 1e-3) / 1e9,
                     "codewords_per_s": None if cws is None else cws / (ld_ms * 1e-3),
                     "bp_iterations_per_s_approx": None if cws is None else its_ref * 13 / (ld_ms * 1e-3),
                     "mean_iterations_per_codeword_approx": None if not n_tbs else its_ref / n_tbs,
                     "fp64_pipe_pct": bp.get("fp64_pipe_pct"), "issue_active_pct": bp.get("issue_active_pct"),
                     "smem_gbs": bp.get("smem_gbs"), "smem_bank_conflict_share": bp.get("bank_conflict_share"),
                     "ncu_source": bp.get("source"),
                     "note": "iterations from the per-TB avg_it records (integer mean over the TB's 13 codewords): approximate"}
    cpu = None
    if not a.no_cpu_baseline:
        try:
            os.sched_setaffinity(0, ctx.cpus_all)        # the CPU leg uses every host core, not only the GPU's NUMA node
        except OSError:
            pass
        cpu = cpu_run(a, 8, 1, max_seconds=20.0)
        one = cpu_run(a, 3, 1, max_seconds=8.0, threads=1, streams=max(1, a.ref_streams // 16))
        cpu["single_thread"] = {kk: one[kk] for kk in ("value", "unit", "cores", "sample", "ms_per_step", "decoded_mbit_per_s")}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": ctx.world, "steps": a.steps, "warmup": max(a.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (FFT/equaliser/demapper) + %s" % ("f64 (LDPC sum-product, reference arithmetic)"
                                                         if a.ldpc_mode == "reference" else "f32 (LDPC layered min-sum)"),
        "data": "synthetic", "decoded_mbit_per_s": bits_all / (ms_step * 1e-3) / 1e6,
        "config": config_dict(a, S),
        "verified": {"tbs_per_step_per_gpu": n_tbs, "crc_ok_tbs": int(ok_ref.sum()), "check": "decoded bytes == transmitted payload"},
        "e2e": e2e, "e2e_from_iq": e2e_iq, "sustained": sustained, "rx_from_iq": from_iq, "rx_from_stream": from_stream, "tx_fec_framer": tx_framer,
        "h2d_ceiling": ceiling, "numa": ctx.numa,
        "gpu_launches": int(launches), "ms_per_step_serial": ms_serial, "clocks": clocks,
        "roofline": roofline, "roofline_ldpc": roofline_ldpc, "stages": stage_tab,
        "layered": layered, "latency": latency, "configs": configs, "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    rx.close()
    if ctx.dist is not None:
        ctx.dist.destroy_process_group()


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
