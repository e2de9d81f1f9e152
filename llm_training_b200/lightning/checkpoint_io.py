"""Checkpoint save / load with the reference's on-disk layout.

* sharded (``save_distributed_checkpoint=True``, FSDP2 default): ``epoch=E-step=S.ckpt`` is a DIRECTORY holding
  torch-DCP files (``__<rank>_0.distcp`` + ``.metadata``) with keys ``state_dict.*`` (full-shape model tensors,
  LightningModule-relative names) and ``optimizer_0.*``, plus ``meta.pt`` with the rest of the checkpoint dict
  (reference: fsdp2_strategy.py:376-389).  The reference's ``convert_to_hf.py`` logic (read DCP metadata,
  allocate, ``dcp.load``, strip ``state_dict.``) reads it unchanged.
* full: one ``torch.save`` file written by rank 0 (reference :390-393).

Writing is balanced by hand: model tensor ``i`` is written by rank ``i % world`` (TP-sharded tensors are
re-assembled one at a time on their writer), so no rank ever stages more than ~1/world of the model on the
host.  Optimizer state (fp32 master + Adam moments of this rank's flat slices) is written by its owner under
rank-qualified keys; the fp32 masters ARE saved (the reference drops them).
"""
import logging
import os
from typing import Any

import torch
import torch.distributed as dist

logger = logging.getLogger(__name__)

_STATE = 'state_dict'


def _dcp():
    import torch.distributed.checkpoint as dcp
    return dcp


def _full_tensor(strategy, name: str, tensor: torch.Tensor, writer: int | None) -> torch.Tensor | None:
    """Full-shape CPU copy of a module tensor on its writer rank (None elsewhere when ``writer`` is set)."""
    info = getattr(tensor, '_tp_info', None)
    if info is not None and strategy.tp_size > 1:
        full = info.gather(tensor.detach())           # collective over the tp axis
    else:
        full = tensor.detach()
    if writer is not None and strategy.global_rank != writer:
        return None
    return full.to('cpu', copy=True)


class _Unsharded:
    """Every engine's parameters whole for the duration (ZeRO-3 engines gather, resident engines just sync)."""

    def __init__(self, strategy, writeback: bool = False) -> None:
        engines = list(getattr(strategy, 'engines', [])) + list(getattr(strategy, 'frozen_engines', []))
        self.ctxs = [e.unsharded(writeback=writeback) for e in engines]

    def __enter__(self):
        for c in self.ctxs:
            c.__enter__()
        return self

    def __exit__(self, *exc):
        for c in reversed(self.ctxs):
            c.__exit__(*exc)
        return False


def module_state_dict(strategy, balanced: bool) -> dict[str, torch.Tensor]:
    module = strategy.lightning_module
    world = strategy.world_size
    out: dict[str, torch.Tensor] = {}
    with _Unsharded(strategy):
        for i, (name, tensor) in enumerate(module.state_dict(keep_vars=True).items()):
            writer = (i % world) if balanced else None
            full = _full_tensor(strategy, name, tensor, writer)
            if full is not None:
                out[name] = full
    return out


def _optimizer_entries(strategy) -> dict[str, torch.Tensor]:
    entries: dict[str, torch.Tensor] = {}
    for i, opt in enumerate(strategy.trainer.optimizers):
        sd = opt.state_dict()
        for unit, states in sd.get('units', {}).items():
            for key, t in states.items():
                entries[f'optimizer_{i}.units.{unit}.{key}.rank{strategy.global_rank}'] = t.detach().to('cpu', copy=True)
    return entries


def _optimizer_meta(strategy) -> list[dict[str, Any]]:
    metas = []
    for opt in strategy.trainer.optimizers:
        sd = opt.state_dict()
        metas.append({k: v for k, v in sd.items() if k != 'units'})
    return metas


def save_checkpoint(strategy, checkpoint: dict[str, Any], filepath: str) -> None:
    filepath = strategy.broadcast(filepath)
    sharded = bool(strategy.save_distributed_checkpoint)
    checkpoint = dict(checkpoint)
    checkpoint['optimizer_meta'] = _optimizer_meta(strategy)
    checkpoint['world_size'] = strategy.world_size
    if sharded:
        if strategy.is_global_zero:
            os.makedirs(filepath, exist_ok=True)
        strategy.barrier()
        payload = {f'{_STATE}.{k}': v for k, v in module_state_dict(strategy, balanced=True).items()}
        payload.update(_optimizer_entries(strategy))
        no_dist = not (dist.is_initialized() and dist.get_world_size() > 1)
        _dcp().save(payload, checkpoint_id=filepath, no_dist=no_dist)
        if strategy.is_global_zero:
            torch.save(checkpoint, os.path.join(filepath, 'meta.pt'))
        return
    # single file: rank 0 holds everything
    state = module_state_dict(strategy, balanced=False)
    opt_entries = _optimizer_entries(strategy)
    gathered: list[dict] | None = None
    if dist.is_initialized() and dist.get_world_size() > 1:
        gathered = [None] * dist.get_world_size() if strategy.is_global_zero else None   # type: ignore[list-item]
        dist.gather_object(opt_entries, gathered, dst=0)
    else:
        gathered = [opt_entries]
    if strategy.is_global_zero:
        os.makedirs(os.path.dirname(os.path.abspath(filepath)), exist_ok=True)
        checkpoint[_STATE] = state
        merged: dict[str, torch.Tensor] = {}
        for part in gathered or []:
            merged.update(part)
        checkpoint['optimizer_states'] = merged
        torch.save(checkpoint, filepath)


def _load_module_state(strategy, state: dict[str, torch.Tensor]) -> None:
    module = strategy.lightning_module
    own = module.state_dict(keep_vars=True)
    missing = [k for k in own if k not in state]
    with torch.no_grad(), _Unsharded(strategy, writeback=True):
        for name, target in own.items():
            if name not in state:
                continue
            src = state[name]
            info = getattr(target, '_tp_info', None)
            if info is not None and strategy.tp_size > 1:
                src = info.scatter(src)
            target.copy_(src.to(target.device, dtype=target.dtype))
    if missing:
        logger.warning(f'checkpoint has no values for {len(missing)} tensors (kept as initialised): {missing[:5]}...')


def _load_optimizer_states(strategy, entries: dict[str, torch.Tensor], metas: list[dict]) -> None:
    for i, opt in enumerate(strategy.trainer.optimizers):
        suffix = f'.rank{strategy.global_rank}'
        prefix = f'optimizer_{i}.units.'
        units: dict[str, dict[str, torch.Tensor]] = {}
        for key, t in entries.items():
            if key.startswith(prefix) and key.endswith(suffix):
                unit, state_key = key[len(prefix):-len(suffix)].rsplit('.', 1)
                units.setdefault(unit, {})[state_key] = t
        meta = dict(metas[i]) if i < len(metas) else {}
        meta['units'] = units
        opt.load_state_dict(meta)
        if not units and hasattr(opt, 'sync_master_from_params'):
            opt.sync_master_from_params()


def load_checkpoint(strategy, filepath: str) -> dict[str, Any]:
    filepath = strategy.broadcast(filepath)
    if os.path.isdir(filepath) and (os.path.exists(os.path.join(filepath, 'latest'))
                                    or os.path.isdir(os.path.join(filepath, 'checkpoint'))):
        # a checkpoint the reference wrote under DeepSpeed (stage <= 2): weights and loop state are restored, the
        # optimizer state (DeepSpeed's own partition files) is rebuilt from the weights
        state, rest = _read_deepspeed_checkpoint(filepath)
        _load_module_state(strategy, state)
        logger.warning('resuming from a DeepSpeed checkpoint: optimizer state is rebuilt (fp32 masters from the '
                       'loaded weights, zero moments)')
        _load_optimizer_states(strategy, {}, [])
        return rest
    if os.path.isdir(filepath):
        dcp = _dcp()
        meta = torch.load(os.path.join(filepath, 'meta.pt'), map_location='cpu', weights_only=False)
        reader = dcp.FileSystemReader(filepath)
        md = reader.read_metadata()
        own = strategy.lightning_module.state_dict(keep_vars=True)
        wanted: dict[str, torch.Tensor] = {}
        for key, tmd in md.state_dict_metadata.items():
            if key.startswith(_STATE + '.'):
                if key[len(_STATE) + 1:] in own:
                    wanted[key] = torch.empty(tmd.size, dtype=tmd.properties.dtype)
            elif key.endswith(f'.rank{strategy.global_rank}') and meta.get('world_size') == strategy.world_size:
                wanted[key] = torch.empty(tmd.size, dtype=tmd.properties.dtype)
        no_dist = not (dist.is_initialized() and dist.get_world_size() > 1)
        dcp.load(wanted, checkpoint_id=filepath, no_dist=no_dist)
        _load_module_state(strategy, {k[len(_STATE) + 1:]: v for k, v in wanted.items() if k.startswith(_STATE + '.')})
        opt_entries = {k: v for k, v in wanted.items() if not k.startswith(_STATE + '.')}
        if meta.get('world_size') != strategy.world_size:
            logger.warning('checkpoint was written with a different world size: optimizer state is rebuilt '
                           '(fp32 masters from the loaded weights, zero moments)')
        _load_optimizer_states(strategy, opt_entries, meta.get('optimizer_meta', []))
        return meta
    ckpt = torch.load(filepath, map_location='cpu', weights_only=False)
    _load_module_state(strategy, ckpt.get(_STATE, {}))
    _load_optimizer_states(strategy, ckpt.get('optimizer_states', {}) if ckpt.get('world_size') == strategy.world_size
                           else {}, ckpt.get('optimizer_meta', []))
    return ckpt


def _read_deepspeed_checkpoint(path: str) -> tuple[dict[str, torch.Tensor], dict[str, Any]]:
    """Checkpoints written by the REFERENCE under its DeepSpeed strategy (six of its eight example configs, stage 2):
    a directory with ``latest`` (tag) and ``<tag>/mp_rank_00_model_states.pt`` whose ``module`` entry is the full
    bf16 state dict under the Lightning wrapper prefix ``_forward_module.``; the Lightning client state (loops,
    callbacks, the embedded ``config``) sits next to it in the same file.  DeepSpeed itself is not needed to read
    it.  ZeRO-3 checkpoints keep only placeholders in ``module`` (the weights are the fp32 partitions of the optimizer
    files): convert those with DeepSpeed's ``zero_to_fp32.py`` first."""
    tag = 'checkpoint'
    latest = os.path.join(path, 'latest')
    if os.path.exists(latest):
        tag = open(latest).read().strip() or tag
    states = os.path.join(path, tag, 'mp_rank_00_model_states.pt')
    if not os.path.exists(states):
        raise FileNotFoundError(f'{path}: DeepSpeed checkpoint without {tag}/mp_rank_00_model_states.pt')
    blob = torch.load(states, map_location='cpu', weights_only=False)
    module = blob.pop('module', None)
    if not module:
        raise ValueError(f'{states} has no `module` state dict')
    placeholders = [k for k, v in module.items() if torch.is_tensor(v) and v.numel() == 0]
    if len(placeholders) > len(module) // 2:
        raise ValueError(f'{path} is a ZeRO-3 checkpoint (parameters are partitioned into the optimizer files); '
                         "run DeepSpeed's zero_to_fp32.py on it first")
    prefix = '_forward_module.'
    state = {(k[len(prefix):] if k.startswith(prefix) else k): v for k, v in module.items()}
    return state, blob


def read_checkpoint_state_dict(path: str) -> tuple[dict[str, torch.Tensor], dict[str, Any]]:
    """Stand-alone reader for tools (``convert_to_hf.py``): returns (model state dict, rest of checkpoint).
    Handles a DCP directory with ``meta.pt``, a single file, and the DeepSpeed engine directories the reference's
    ZeRO-1/2 configs write (``<ckpt>/latest`` + ``<tag>/mp_rank_00_model_states.pt``)."""
    if os.path.isdir(path):
        if os.path.exists(os.path.join(path, 'latest')) or os.path.isdir(os.path.join(path, 'checkpoint')):
            return _read_deepspeed_checkpoint(path)
        dcp = _dcp()
        reader = dcp.FileSystemReader(path)
        md = reader.read_metadata()
        wanted = {k: torch.empty(t.size, dtype=t.properties.dtype) for k, t in md.state_dict_metadata.items()
                  if k.startswith(_STATE + '.')}
        dcp.load(wanted, checkpoint_id=path, no_dist=True)
        meta = torch.load(os.path.join(path, 'meta.pt'), map_location='cpu', weights_only=False)
        return {k[len(_STATE) + 1:]: v for k, v in wanted.items()}, meta
    ckpt = torch.load(path, map_location='cpu', weights_only=False)
    return ckpt.pop(_STATE), ckpt
