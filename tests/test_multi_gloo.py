"""world_size-2 CPU (gloo) test of the N>1 path: contiguous env sharding with no data-path collective and
the single all-gather of episode scores (magical-1_b200/shard.py)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import magical_b200  # noqa: F401
from magical_b200 import shard


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, num_envs, out_dir):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    lo, hi = shard.shard_range(num_envs, rank, world)
    local = torch.arange(lo, hi, dtype=torch.float32) * 0.5 + 1.0      # stand-in for this rank's eval scores
    full = shard.all_gather_scores(local, num_envs, rank, world)
    ids = ['B-Demo-v0', 'A-Demo-v0'] * (num_envs // 2) + ['C-Demo-v0'] * (num_envs % 2)
    gidx, local_ids = shard.shard_env_ids(ids, rank, world)
    perm = torch.as_tensor(gidx, dtype=torch.float32) * 0.5 + 1.0
    full2 = shard.all_gather_scores(perm, num_envs, rank, world, global_index=gidx)
    np.save(os.path.join(out_dir, f'r{rank}.npy'), np.stack([full.numpy(), full2.numpy()]))
    assert local_ids == sorted(local_ids)
    dist.destroy_process_group()


def test_shard_ranges_cover_batch():
    for n, w in ((4096, 8), (10, 3), (7, 8), (32768, 8)):
        ranges = [shard.shard_range(n, r, w) for r in range(w)]
        assert ranges[0][0] == 0 and ranges[-1][1] == n
        assert all(ranges[i][1] == ranges[i + 1][0] for i in range(w - 1))
        sizes = [b - a for a, b in ranges]
        assert max(sizes) - min(sizes) <= 1


def test_score_all_gather_world2(tmp_path):
    num_envs, world = 11, 2
    mp.spawn(_worker, args=(world, _free_port(), num_envs, str(tmp_path)), nprocs=world, join=True)
    want = np.arange(num_envs, dtype=np.float32) * 0.5 + 1.0
    for r in range(world):
        got = np.load(tmp_path / f'r{r}.npy')
        np.testing.assert_array_equal(got[0], want)
        np.testing.assert_array_equal(got[1], want)
