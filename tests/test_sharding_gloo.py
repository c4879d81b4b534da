"""N>1 path on the CPU: two ranks over gloo, each owning its own shard of publishers / recording
files, no data-path collective; the only exchanges are the ones bench.py uses for timing
(barrier, MAX over ranks) and, here, a gather to check coverage."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as tmp

from blendtorch import btt
from blendtorch.btt import synthetic as S


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, prefix, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        addrs = ["tcp://127.0.0.1:%d" % (20000 + i) for i in range(5)]
        mine = btt.shard(addrs, rank, world)
        assert mine == addrs[rank::world]
        ds = btt.FileDataset(prefix, rank=rank, world_size=world)
        keys = [(int(it["btid"]), int(it["frameid"])) for it in (ds[i] for i in range(len(ds)))]
        dist.barrier()
        # timing reduction as bench.py does it: MAX over ranks
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == float(world)
        n = torch.tensor([len(keys)], dtype=torch.int64)
        dist.all_reduce(n, op=dist.ReduceOp.SUM)
        np.save(os.path.join(out_dir, "keys_%d.npy" % rank), np.array(keys, dtype=np.int64))
        np.save(os.path.join(out_dir, "total_%d.npy" % rank), n.numpy())
    finally:
        dist.destroy_process_group()


def test_two_ranks_own_disjoint_shards(tmp_path):
    prefix = str(tmp_path / "rec")
    S.write_recordings(prefix, config=1, n_files=5, frames_per_file=3, h=8, w=8)
    world = 2
    tmp.spawn(_rank_main, args=(world, _free_port(), prefix, str(tmp_path)), nprocs=world, join=True)
    k0 = {tuple(x) for x in np.load(tmp_path / "keys_0.npy")}
    k1 = {tuple(x) for x in np.load(tmp_path / "keys_1.npy")}
    assert not (k0 & k1)                                     # disjoint
    assert k0 | k1 == {(f, i) for f in range(5) for i in range(3)}   # complete
    assert {f for f, _ in k0} == {0, 2, 4} and {f for f, _ in k1} == {1, 3}
    assert int(np.load(tmp_path / "total_0.npy")[0]) == 15
