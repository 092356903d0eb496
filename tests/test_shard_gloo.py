"""world_size-2 (and 3) gloo test of the level-sharding plan and the per-level exchange: every rank ends
every level with the full wire block, equal to the unsharded evaluation.  The per-gate work is a plaintext stand-in
(the engine has no CPU path); what is under test is the slicing, padding and exchange that bench.py runs over NCCL."""
import os
import socket
import sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, width, depth, stride, q):
    from helm_b200 import shard
    import bench
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pw = shard.padded_width(width, world)
    wires = torch.zeros(((depth + 1) * pw, stride), dtype=torch.int32)
    bits0 = (bench.splitmix64(2, width) & np.uint64(1)).astype(np.int32)
    wires[:width, 0] = torch.from_numpy(bits0)                    # word 0 carries the plaintext bit, the rest a row tag
    wires[:width, 1] = torch.arange(width, dtype=torch.int32)
    for lvl in range(1, depth + 1):
        ops, in0, in1 = bench.synth_level(lvl, width)
        g0, cnt = shard.level_slice(width, world, rank)
        prev = wires[(lvl - 1) * pw:(lvl - 1) * pw + width, 0].numpy()
        out = bench.plain_gate(ops[g0:g0 + cnt], prev[in0[g0:g0 + cnt]], prev[in1[g0:g0 + cnt]])
        blk = wires[lvl * pw:(lvl + 1) * pw]
        blk[g0:g0 + cnt, 0] = torch.from_numpy(out.astype(np.int32))
        blk[g0:g0 + cnt, 1] = torch.arange(g0, g0 + cnt, dtype=torch.int32) + 1000 * lvl
        shard.exchange_level(dist, blk, world, rank)
    q.put((rank, wires[:, :2].numpy().copy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,width", [(2, 64), (2, 37), (3, 50)])
def test_sharded_levels_equal_unsharded(world, width):
    import bench
    depth, stride = 4, 8
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, width, depth, stride, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    from helm_b200 import shard
    pw = shard.padded_width(width, world)
    cur = (bench.splitmix64(2, width) & np.uint64(1)).astype(bool)
    for lvl in range(1, depth + 1):
        ops, in0, in1 = bench.synth_level(lvl, width)
        cur = bench.plain_gate(ops, cur[in0], cur[in1])
        for r in range(world):
            got = res[r][lvl * pw:lvl * pw + width]
            assert np.array_equal(got[:, 0].astype(bool), cur), (lvl, r)
            assert np.array_equal(got[:, 1], np.arange(width) + 1000 * lvl), (lvl, r)   # every row came from its owner
    for r in range(1, world):
        assert np.array_equal(res[0], res[r])


def test_level_slice_plan():
    from helm_b200 import shard
    assert shard.level_slice(65536, 8, 3) == (3 * 8192, 8192)
    assert shard.level_slice(10, 4, 3) == (9, 1) and shard.level_slice(10, 4, 0) == (0, 3)
    assert shard.level_slice(167, 8, 5, min_width=1184) == (0, 0) and shard.level_slice(167, 8, 0, min_width=1184) == (0, 167)
    assert sum(shard.level_slice(1037, 8, r)[1] for r in range(8)) == 1037
    assert shard.padded_width(10, 4) == 12
