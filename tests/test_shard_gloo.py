"""N > 1 host logic on CPU: world_size-2 gloo process group, stream sharding and the per-GPU statistics gather."""
import os
import socket

import numpy as np
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "gr-dtl_b200"))
    import torch.distributed as dist
    from dtl_b200 import shard
    from dtl_b200._capi import tb_result_dtype
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    S, F, fs = 11, 3, 1920
    s0, ns = shard.shard_streams(S, world, rank)
    frames = np.arange(S * F * 2).reshape(S * F, 2)
    mine = shard.shard_frames(frames, S, F, world, rank)
    # a fake rx() result for this shard: stream s released s % 3 TBs, all but the first CRC-ok, 10 bytes each
    tb_cap = 4
    tbs = np.zeros((ns, tb_cap), dtype=tb_result_dtype)
    ntbs = np.array([(s0 + i) % 3 for i in range(ns)], dtype=np.int32)
    for i in range(ns):
        tbs[i, :ntbs[i]]["crc_ok"] = 1
        if ntbs[i]:
            tbs[i, 0]["crc_ok"] = 0
        tbs[i, :ntbs[i]]["avg_it"] = 2
    nbytes = np.array([10 * max(0, n - 1) for n in ntbs], dtype=np.int32)
    st = shard.RxStats().add_rx_result(dict(ntbs=ntbs, tbs=tbs, nbytes=nbytes), ns, F, fs).all_reduce()
    slowest = shard.max_over_ranks(1.0 + rank)
    q.put((rank, s0, ns, mine[0, 0].item(), mine.shape[0], st.v.tolist(), slowest))
    dist.barrier()
    dist.destroy_process_group()


def test_stream_sharding_and_stats_gather_world2():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    S, F = 11, 3
    (r0, a0, n0, first0, rows0, v0, m0), (r1, a1, n1, first1, rows1, v1, m1) = res
    assert (a0, n0, a1, n1) == (0, 6, 6, 5)                      # contiguous, sizes differ by at most one
    assert rows0 == n0 * F and rows1 == n1 * F and first0 == 0 and first1 == a1 * F * 2
    assert v0 == v1                                               # every rank holds the same totals
    ntb = [s % 3 for s in range(S)]
    want = [S * F * 1920, S * F, sum(ntb), sum(max(0, n - 1) for n in ntb), sum(10 * max(0, n - 1) for n in ntb), 2 * sum(ntb)]
    assert v0 == [float(x) for x in want]
    assert m0 == m1 == 2.0


def test_shard_streams_partitions_exactly():
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "gr-dtl_b200"))
    from dtl_b200 import shard
    for S in (0, 1, 7, 8, 1024, 8192, 8191):
        for W in (1, 2, 4, 8):
            parts = [shard.shard_streams(S, W, r) for r in range(W)]
            assert parts[0][0] == 0 and sum(n for _, n in parts) == S
            for (a, n), (b, _) in zip(parts, parts[1:]):
                assert a + n == b
            assert max(n for _, n in parts) - min(n for _, n in parts) <= 1
