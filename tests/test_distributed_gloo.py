"""CPU-side, world_size 2 over gloo: world sharding + the statistics/timing reductions bench.py uses at N > 1."""
import os
import socket
import numpy as np
import torch.multiprocessing as mp
from box2d_b200 import batch


def test_shard_ranges_cover_and_balance():
    for n in (1, 7, 8, 8192, 8191):
        for ws in (1, 2, 4, 8):
            r = [batch.shard_range(n, k, ws) for k in range(ws)]
            assert r[0][0] == 0 and r[-1][1] == n and all(r[i][1] == r[i + 1][0] for i in range(ws - 1))
            sizes = [e - b for b, e in r]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, port, out):
    import sys
    import torch.distributed as dist
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.dirname(here)); sys.path.insert(0, here)
    from box2d_b200 import api, scenes
    from oracle_backend import OracleBackend
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=2)
    begin, end = batch.shard_range(5, rank, 2)                 # 5 independent pyramid worlds over 2 ranks: 3 + 2
    w = api.World(OracleBackend(), scenes.world_def())         # gloo test of the host logic: the CPU oracle steps the worlds
    for k in range(begin, end):
        scenes.add_pyramid(w, 4, origin=(k * 15.0, 0.0), jitter_seed=k, jitter=0.001)
    for _ in range(5):
        w.Step(1 / 60, 4)
    b = w.ReadBodies()
    dyn = b[b["flags"] != 0]
    stats = batch.reduce_stats(batch.batch_stats(dyn), dist)
    ms, units = batch.reduce_timing([10.0 + rank, 3.0 - rank], len(dyn) * 4 * 5, dist)
    if rank == 0:
        np.save(out, np.concatenate([stats, ms, [units]]))
    w.Destroy()
    dist.destroy_process_group()


def test_two_rank_reduction(tmp_path):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / "r.npy")
    mp.spawn(_worker, args=(port, out), nprocs=2, join=True)
    r = np.load(out)
    assert r[0] == 5 * 10                                       # 5 worlds x 10 boxes
    assert r[4] == 11.0 and r[5] == 3.0                         # max over ranks
    assert r[6] == 5 * 10 * 4 * 5                               # units summed over ranks
