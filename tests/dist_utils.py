"""Helpers to run a function on N gloo ranks (CPU) from pytest."""
import os
import socket
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _to_plain(obj):
    """Tensors -> numpy so results cross the process boundary by value (no shared-memory handles)."""
    if isinstance(obj, torch.Tensor):
        return obj.detach().cpu().float().numpy() if obj.dtype == torch.bfloat16 else obj.detach().cpu().numpy()
    if isinstance(obj, dict):
        return {k: _to_plain(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_plain(v) for v in obj)
    return obj


def _from_plain(obj):
    import numpy as np
    if isinstance(obj, np.ndarray):
        return torch.from_numpy(obj)
    if isinstance(obj, dict):
        return {k: _from_plain(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_from_plain(v) for v in obj)
    return obj


def _entry(rank: int, world: int, port: int, fn, args, queue) -> None:
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR='127.0.0.1',
                      MASTER_PORT=str(port))
    torch.set_num_threads(2)
    try:
        dist.init_process_group('gloo', rank=rank, world_size=world)
        out = fn(rank, world, *args)
        queue.put((rank, 'ok', _to_plain(out)))
    except Exception:  # noqa: BLE001
        queue.put((rank, 'error', traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def run_distributed(fn, world: int = 2, args: tuple = (), timeout: float = 300.0) -> list:
    ctx = mp.get_context('spawn')
    queue = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, queue)) for r in range(world)]
    for p in procs:
        p.start()
    results: dict[int, object] = {}
    try:
        for _ in range(world):
            rank, status, payload = queue.get(timeout=timeout)
            if status != 'ok':
                raise AssertionError(f'rank {rank} failed:\n{payload}')
            results[rank] = _from_plain(payload)
    finally:
        for p in procs:
            p.join(timeout=30)
            if p.is_alive():
                p.kill()
    return [results[r] for r in range(world)]
