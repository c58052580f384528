"""CPU, gloo, world_size 2: the sharding and metric-reduction logic of the N>1 path."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dlwp_benchmark_b200.ensemble import shard_range
from dlwp_benchmark_b200 import metrics as pmetrics
from oracle import metrics as ometrics


def test_shard_range_partitions():
    for total in (256, 7, 1, 0):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(5)
    o, t = torch.randn(7, 3, 2, 8, 16), torch.randn(7, 3, 2, 8, 16)
    a, b = shard_range(7, rank, world)                       # uneven split: 4 + 3 members
    sums, n = ometrics.partial_sums(o[a:b], t[a:b])          # CPU stand-in for the device kernel
    # (1) the caller knows the ensemble size: the sums alone are reduced, in place
    s1 = sums.clone()
    assert pmetrics.reduce_sums(s1) is s1
    rmse, acc = pmetrics.finalize(s1, 7 * 8 * 16)
    # (2) it does not: the per-rank counts ride in the same all-reduce as one extra element (rollout_metrics)
    buf = torch.cat([sums.reshape(-1), torch.tensor([float(n)], dtype=torch.float64)])
    pmetrics.reduce_sums(buf)
    rmse2, acc2 = pmetrics.finalize(buf[:-1].view_as(sums), buf[-1])
    assert torch.equal(rmse, rmse2) and torch.equal(acc, acc2)
    if rank == 0:
        torch.save({"rmse": rmse, "acc": acc, "n": float(buf[-1])}, out_path)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_metric_reduction_equals_single_process(tmp_path):
    out = str(tmp_path / "r.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    torch.manual_seed(5)
    o, t = torch.randn(7, 3, 2, 8, 16), torch.randn(7, 3, 2, 8, 16)
    rmse, acc = ometrics.rmse_acc(o, t)
    assert got["n"] == 7 * 8 * 16
    assert torch.allclose(got["rmse"], rmse, rtol=1e-12) and torch.allclose(got["acc"], acc, rtol=1e-12)
