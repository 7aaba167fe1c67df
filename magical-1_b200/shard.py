"""Multi-GPU layout of the step path: envs are independent (each reference env owns its own pm.Space,
Viewer and RandomState: magical/base_env.py:112-123, 206), so a batch is split into contiguous blocks of
env indices, one block per rank, with NO inter-GPU traffic on the step path.  The only exchange is an
all-gather of per-env episode scores (the reference's info['eval_score'], base_env.py:289-302) when
episodes end; over NVLink/NVSwitch that is latency-bound, so it is one NCCL call on a flat f32 buffer."""
import numpy as np


def shard_range(num_envs, rank, world):
    """[lo, hi) of the global env indices owned by `rank` (contiguous blocks; sizes differ by <= 1)."""
    base, rem = divmod(int(num_envs), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_env_ids(env_ids, rank, world):
    """Env ids of a mixed batch owned by `rank`, sorted by task name inside the rank so that neighbouring
    lanes of the physics kernel see the same task layout.  Returns (global indices, ids)."""
    lo, hi = shard_range(len(env_ids), rank, world)
    idx = np.arange(lo, hi)
    order = np.argsort(np.asarray([env_ids[i] for i in idx], dtype=object), kind='stable')
    idx = idx[order]
    return idx, [env_ids[i] for i in idx]


def all_gather_scores(local_scores, num_envs, rank, world, global_index=None):
    """All-gather of the per-env eval scores of every rank into one [num_envs] tensor in global env order.
    `local_scores`: f32 tensor of this rank's block (CUDA tensor -> NCCL, CPU tensor -> gloo).
    `global_index`: global env index of each local score when the block was re-ordered (shard_env_ids)."""
    import torch
    import torch.distributed as dist
    sizes = [shard_range(num_envs, r, world)[1] - shard_range(num_envs, r, world)[0] for r in range(world)]
    n_max = max(sizes)
    dev = local_scores.device
    # one int32 buffer per rank: row 0 = the scores' f32 bit patterns, row 1 = global env indices (exact for any
    # batch size; -1 = padding), so that ONE collective moves both
    send = torch.full((2, n_max), -1, dtype=torch.int32, device=dev)
    n = local_scores.numel()
    send[0, :n] = local_scores.to(torch.float32).contiguous().view(torch.int32)
    lo = shard_range(num_envs, rank, world)[0]
    gidx = torch.arange(lo, lo + n, device=dev) if global_index is None else torch.as_tensor(global_index, device=dev)
    send[1, :n] = gidx.to(torch.int32)
    if world > 1 and dist.is_initialized():
        # flat buffers: the same call works for NCCL (CUDA tensors) and gloo (CPU tensors)
        recv = torch.empty(world * 2 * n_max, dtype=torch.int32, device=dev)
        dist.all_gather_into_tensor(recv, send.reshape(-1))
        recv = recv.reshape(world, 2, n_max)
    else:
        recv = send[None]
    # one scatter for all ranks: padding entries (index -1) are masked out
    idx = recv[:, 1, :].reshape(-1).long()
    val = recv[:, 0, :].reshape(-1).contiguous().view(torch.float32)
    keep = idx >= 0
    out = torch.zeros(num_envs, dtype=torch.float32, device=dev)
    out[idx[keep]] = val[keep]
    return out
