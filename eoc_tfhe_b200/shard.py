"""Gate batches over several GPUs of one node: one process per GPU, `torch.distributed` for the plumbing.

The path shards by independent units: gates of a level share only the read-only
cloud key, so
  * the key is built once on rank 0 (upload + conversion to the evaluation domain on its GPU) and
    replicated with one broadcast of the two resident blobs (NCCL over NVLink),
  * a level's gate list is cut into contiguous slices by rank (a MUX stays on one rank),
  * results are gathered once per level with one all-gather; circuits whose instances are independent
    (one adder per operand pair) shard by instance and need no per-level traffic at all.
There is no data-path collective inside a level.  Everything here is host logic around an `evaluate`
callable, so the same code runs under gloo on CPU tensors in the tests (with the CPU oracle as the
evaluator) and under NCCL on CUDA tensors in bench.py.
"""
import numpy as np


def slice_for_rank(count, world, rank):
    """Contiguous [start, end) of `count` units for `rank`: sizes differ by at most one, low ranks first."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    base, rem = divmod(int(count), world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def padded_slice_len(count, world):
    return -(-int(count) // world)


def broadcast_key(cloud_key, src=0, group=None):
    """Replicate the device-resident key: one broadcast per blob (FFT-domain BK, key-switch table).
    Returns the CUDA-event time of the broadcasts in ms."""
    import torch
    import torch.distributed as dist
    blobs = cloud_key.device_blob_tensors()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for t in blobs:
        dist.broadcast(t, src=src, group=group)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


def replicated_key(params, make_on_root, device, ks_variance=None, group=None):
    """CloudKey on every rank: `make_on_root()` runs on rank 0 only, the other ranks allocate an empty key
    and receive the blobs."""
    import torch.distributed as dist
    from .engine import CloudKey
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    ck = make_on_root() if rank == 0 else CloudKey.empty(params, ks_variance=ks_variance, device=device)
    ms = broadcast_key(ck, 0, group) if dist.is_initialized() and dist.get_world_size(group) > 1 else 0.0
    return ck, ms


def all_gather_rows(local, count, group=None):
    """All-gather of per-rank row slices (torch tensors [len_r][row]) cut by slice_for_rank into the full
    [count][row] tensor.  Slices are padded to a common length so that one all_gather_into_tensor does it."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    pad = padded_slice_len(count, world)
    row = local.shape[1]
    send = local
    if local.shape[0] != pad:
        send = torch.zeros((pad, row), dtype=local.dtype, device=local.device)
        send[:local.shape[0]] = local
    recv = torch.empty((world * pad, row), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(recv, send.contiguous(), group=group)
    if count == world * pad:
        return recv
    parts = []
    for r in range(world):
        s, e = slice_for_rank(count, world, r)
        parts.append(recv[r * pad:r * pad + (e - s)])
    return torch.cat(parts, dim=0)


def sharded_gate(evaluate, op, a, b=None, c=None, group=None):
    """One level of `count` independent gates over all ranks.

    `a`, `b`, `c` are the FULL input batches (torch tensors, identical on every rank, e.g. the previous
    level's gathered outputs); `evaluate(op, a_slice, b_slice, c_slice)` returns this rank's output rows
    (CloudKey.gate on the GPU).  Returns the full gathered output batch on every rank."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    count = a.shape[0]
    s, e = slice_for_rank(count, world, rank)
    cut = lambda t: None if t is None else t[s:e].contiguous()
    local = evaluate(op, cut(a), cut(b), cut(c))
    return all_gather_rows(local, count, group)


def sharded_instances(run_local, inputs, count, group=None):
    """Independent circuit instances (one adder per operand pair) sharded by instance: every rank runs
    `run_local(inputs_slice, n_local)` -> dict name -> tensor [n_local][...] on its contiguous slice, and
    the outputs are gathered once at the end (no per-level traffic)."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    s, e = slice_for_rank(count, world, rank)
    local = run_local({k: v[s:e] for k, v in inputs.items()}, e - s)
    out = {}
    for name, t in local.items():
        flat = t.reshape(e - s, -1)
        out[name] = all_gather_rows(flat, count, group).reshape((count,) + tuple(t.shape[1:]))
    return out
