"""Stream sharding across the GPUs of one box: streams are independent, so rank r of W owns
a contiguous block of streams and runs the whole receive path on its own GPU.  Nothing crosses GPUs on the data
path; only per-GPU statistics are combined, with one tiny all-reduce per reporting interval (torch.distributed:
NCCL on GPUs, gloo in the CPU tests)."""
import numpy as np


def shard_streams(n_streams, world_size, rank):
    """(first stream, number of streams) of `rank`: contiguous blocks, sizes differ by at most one."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError("bad rank/world_size")
    base, extra = divmod(int(n_streams), int(world_size))
    start = rank * base + min(rank, extra)
    return start, base + (1 if rank < extra else 0)


def shard_frames(arr, n_streams, frames_per_stream, world_size, rank):
    """The rows of a stream-major [n_streams * frames_per_stream, ...] array that belong to `rank`."""
    s0, ns = shard_streams(n_streams, world_size, rank)
    return arr[s0 * frames_per_stream:(s0 + ns) * frames_per_stream]


class RxStats:
    """Per-GPU counters gathered on the host side (north_star: "only per-GPU stats are gathered")."""
    FIELDS = ("samples", "frames", "tbs", "tbs_crc_ok", "payload_bytes", "ldpc_iterations")

    def __init__(self, **kw):
        self.v = np.zeros(len(self.FIELDS), dtype=np.float64)
        for k, x in kw.items():
            self.v[self.FIELDS.index(k)] = x

    def __getitem__(self, k):
        return float(self.v[self.FIELDS.index(k)])

    def add_rx_result(self, out, n_streams, frames_per_stream, frame_samples):
        """accumulate the outputs of ofdm_receiver.rx (NumPy arrays)"""
        from ._capi import tb_result_dtype
        ntb = np.asarray(out["ntbs"])
        tbs = np.asarray(out["tbs"]).view(tb_result_dtype).reshape(n_streams, -1)
        self.v[0] += n_streams * frames_per_stream * frame_samples
        self.v[1] += n_streams * frames_per_stream
        self.v[2] += int(ntb.sum())
        for s in range(n_streams):
            t = tbs[s, :ntb[s]]
            self.v[3] += int(t["crc_ok"].sum())
            self.v[5] += int(t["avg_it"].sum())
        self.v[4] += int(np.asarray(out["nbytes"]).sum())
        return self

    def all_reduce(self, device=None):
        """sum over ranks (no-op without an initialised process group)"""
        import torch
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()):
            return self
        t = torch.from_numpy(self.v.copy())
        if device is not None:
            t = t.to(device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        self.v = t.cpu().numpy()
        return self


def max_over_ranks(x, device=None):
    """max of a scalar over ranks (timings are reported as the slowest rank's)"""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64)
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
