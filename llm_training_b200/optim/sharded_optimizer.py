"""Optimizer over the flat shards of :class:`parallel.FlatParamEngine`.

One fused kernel launch per sharding unit does: gradient clip scaling (from the device-side global
norm), AdamW on the fp32 master slice, and the bf16 write-back — followed by one in-place all-gather per
unit.  This replaces the reference's per-parameter ``MasterWeightsOptimizer.step`` (bf16->fp32 grad copy,
AdamW, fp32->bf16 copy; src/llm_training/optim/master_weight_wrapper.py:63-80), ``clip_grad_norm_``
(fsdp2_precision.py:165-169) and DeepSpeed's FusedAdam + ZeRO partitioned step.

AdamW-family user optimizers (torch.optim.AdamW/Adam, FusedAdamW, the DeepSpeed FusedAdam alias) run on
the fused kernel; any other optimizer class is instantiated over the fp32 master slices and stepped by
PyTorch (generic path).  The user's optimizer object stays the owner of ``param_groups`` so LR schedulers
and LearningRateMonitor keep working unchanged.
"""
from typing import Any

import torch
from torch.optim import Optimizer

from ..ops.adamw_op import adamw_step
from ..parallel.comm import Comm
from ..parallel.fsdp import FlatParamEngine

import os

from ..ops import _lib

# occupancy of the update kernels while they share the GPU with the next forward pass
_BACKGROUND_CTAS_PER_SM = int(os.environ.get('LLMT_OPT_BG_CTAS', '2'))

_ADAMW_LIKE = ('AdamW', 'FusedAdamW', 'FusedAdam', 'Adam', 'DeepSpeedCPUAdam')


class ShardedOptimizer(Optimizer):
    def __init__(self, optimizer: Optimizer, engines: list[FlatParamEngine], max_grad_norm: float | None = None,
                 norm_extra_axes: list[Comm] | None = None, offload: bool = False,
                 master_weights: bool = True) -> None:   # noqa: no super()
        """``master_weights=False`` (the reference's ``FSDP2Strategy(use_master_weights=False)``,
        fsdp2_strategy.py:59,249): no fp32 copy of low-precision parameters — the user's optimizer class steps the bf16
        slices directly (PyTorch ops; the reference README warns this "can significantly degrade training").
        ``offload``: keep the fp32 master slices and Adam moments in (pinned) host memory and run the update on
        the CPU — the reference's ``CPUOffloadPolicy`` / DeepSpeed ``offload_optimizer`` (fsdp2_strategy.py:58,
        deepspeed_strategy.py:22-37).  It frees 12 bytes per parameter slice of HBM at the price of a PCIe round trip
        and a CPU AdamW per step; with 180 GB per GPU it is an escape hatch, not the default."""
        self._optimizer = optimizer
        self.offload = bool(offload)
        self.engines = engines
        self.max_grad_norm = max_grad_norm
        self.norm_extra_axes = norm_extra_axes or []
        self.master_weights = bool(master_weights) or self.offload
        self.fused = type(optimizer).__name__ in _ADAMW_LIKE
        if not self.master_weights and any(u.param_shard.dtype != torch.float32 for e in engines for u in e.units):
            self.fused = False     # the fused kernel updates fp32 slices; without masters torch steps the bf16 ones
        if type(optimizer).__name__ == 'Adam' and any(g.get('weight_decay', 0) != 0 for g in optimizer.param_groups):
            self.fused = False     # Adam's L2 penalty is not decoupled decay
        # parameter -> index of the user's param group (per-group lr / weight_decay, e.g. no-decay groups)
        self._group_of = {id(p): gi for gi, g in enumerate(optimizer.param_groups) for p in g['params']}
        self.step_count = 0
        self.unit_state: list[dict[str, torch.Tensor]] = []
        for e in engines:
            for u in e.units:
                shard = u.param_shard
                st: dict[str, torch.Tensor] = {}
                if self.offload:
                    pin = shard.is_cuda
                    st['master'] = shard.detach().to('cpu', torch.float32).clone()
                    st['exp_avg'] = torch.zeros(u.shard_numel, dtype=torch.float32)
                    st['exp_avg_sq'] = torch.zeros(u.shard_numel, dtype=torch.float32)
                    if pin:
                        st = {k: v.pin_memory() for k, v in st.items()}
                    self.unit_state.append(st)
                    continue
                if shard.dtype != torch.float32 and self.master_weights:
                    st['master'] = shard.detach().float().clone()
                st['exp_avg'] = torch.zeros(u.shard_numel, dtype=torch.float32, device=shard.device)
                st['exp_avg_sq'] = torch.zeros(u.shard_numel, dtype=torch.float32, device=shard.device)
                self.unit_state.append(st)
        self._generic: Optimizer | None = None
        if not self.fused:
            self._build_generic()
        self.last_grad_norm: torch.Tensor | None = None
        # fp16 loss scaling (set by the precision plugin before every step): gradients are multiplied by
        # `grad_scale` (= 1 / loss scale) INSIDE the fused update; a non-finite gradient norm skips the step
        self.grad_scale: float = 1.0
        self.skip_nonfinite: bool = False
        self.last_step_skipped: bool = False

    # -- the user's optimizer stays the source of truth for hyper-parameters ---------------------------
    @property
    def param_groups(self):
        return self._optimizer.param_groups

    @param_groups.setter
    def param_groups(self, v):
        self._optimizer.param_groups = v

    @property
    def state(self):
        return self._optimizer.state

    @property
    def defaults(self):
        return self._optimizer.defaults

    def __getattr__(self, name: str) -> Any:
        if name.startswith('_'):
            raise AttributeError(name)
        return getattr(self._optimizer, name)

    def __repr__(self) -> str:
        return f'ShardedOptimizer({type(self._optimizer).__name__}, fused={self.fused})'

    @property
    def inner_class_name(self) -> str:
        return type(self._optimizer).__name__

    def _units(self):
        i = 0
        for e in self.engines:
            for u in e.units:
                yield e, u, self.unit_state[i]
                i += 1

    def _segments(self, u) -> list[tuple[int, int, int]]:
        """``[(param_group index, start in shard, length)]`` covering this rank's slice of unit ``u``: runs of
        consecutive parameters that belong to the same group of the user's optimizer (alignment gaps and the
        tail padding join the run before them).  One group -> one segment, the whole slice."""
        cached = getattr(u, '_opt_segments', None)
        if cached is not None:
            return cached
        runs: list[list[int]] = []
        if len(self._optimizer.param_groups) > 1:
            for _, p, _, start, length in u.shard_slices():
                gi = self._group_of.get(id(p), 0)
                if runs and runs[-1][0] == gi:
                    runs[-1][2] = start + length - runs[-1][1]
                else:
                    if runs:
                        runs[-1][2] = start - runs[-1][1]
                    runs.append([gi, start, length])
        if not runs:
            runs = [[0, 0, u.shard_numel]]
        runs[0][2] += runs[0][1]
        runs[0][1] = 0
        runs[-1][2] = u.shard_numel - runs[-1][1]
        u._opt_segments = [tuple(r) for r in runs]
        return u._opt_segments

    def _build_generic(self) -> None:
        """Any non-AdamW optimizer class: one fp32 leaf per (unit, group run), grouped like the user's param groups."""
        per_group: list[list[torch.nn.Parameter]] = [[] for _ in self._optimizer.param_groups]
        for _, u, st in self._units():
            m = st.get('master')
            if m is None:
                m = u.param_shard      # fp32 training: step the slice itself
            st['master_params'] = []
            for gi, start, length in self._segments(u):
                leaf = torch.nn.Parameter(m[start:start + length], requires_grad=True)
                st['master_params'].append((leaf, start, length))
                per_group[gi].append(leaf)
        groups = []
        for leaves, g in zip(per_group, self._optimizer.param_groups):
            groups.append({'params': leaves, **{k: v for k, v in g.items()
                                                 if k != 'params' and k in self._optimizer.defaults}})
        self._generic_groups = [i for i, g in enumerate(groups) if g['params']]
        self._generic = type(self._optimizer)([groups[i] for i in self._generic_groups])

    # -- step ------------------------------------------------------------------------------------------
    def sync_master_from_params(self) -> None:
        """Call after weights were loaded into the flat buffers (pre-trained weights / checkpoint)."""
        self.synchronize()
        for _, u, st in self._units():
            if 'master' in st:
                st['master'].copy_(u.param_shard)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for e in self.engines:
            e.finish_backward()
        norm_sq = None
        for e in self.engines:
            ns = e.grad_norm_sq(self.norm_extra_axes)
            norm_sq = ns if norm_sq is None else norm_sq.add_(ns)
        self.last_grad_norm = norm_sq.sqrt() * abs(self.grad_scale) if norm_sq is not None else None
        self.last_step_skipped = False
        if self.skip_nonfinite and norm_sq is not None and not bool(torch.isfinite(norm_sq).all()):
            self.last_step_skipped = True          # overflow under fp16 loss scaling: leave weights and state alone
            return loss
        self.step_count += 1
        for e in self.engines:
            e._slices_dirty = True        # ZeRO-3: the next parameter gather needs one cross-rank barrier first
        max_norm = float(self.max_grad_norm) if self.max_grad_norm else 0.0
        pushed_all: dict[int, bool] = {}
        if self.fused:
            hyper = [dict(lr=g['lr'], beta1=g.get('betas', (0.9, 0.999))[0], beta2=g.get('betas', (0.9, 0.999))[1],
                          eps=g.get('eps', 1e-8), weight_decay=g.get('weight_decay', 0.0), step=self.step_count,
                          max_norm=max_norm, norm_sq=norm_sq, grad_scale=self.grad_scale)
                     for g in self._optimizer.param_groups]
            if self._overlap_enabled():
                self._step_overlapped(hyper)
                return loss
            for e, u, st in self._units():
                pushed = self._update_unit(e, u, st, hyper)
                pushed_all[id(e)] = pushed_all.get(id(e), True) and pushed
        else:
            clip = self.grad_scale
            if max_norm > 0 and norm_sq is not None:
                clip = self.grad_scale * torch.clamp(max_norm / (norm_sq.sqrt() * abs(self.grad_scale) + 1e-6), max=1.0)
            for group, gi in zip(self._generic.param_groups, self._generic_groups):
                for k, v in self._optimizer.param_groups[gi].items():
                    if k != 'params':
                        group[k] = v
            for _, u, st in self._units():
                grad = u.local_grad().float() * clip
                for leaf, start, length in st['master_params']:
                    leaf.grad = grad[start:start + length].to(device=leaf.device, dtype=leaf.dtype)
            self._generic.step()
            for _, u, st in self._units():
                if 'master' in st:
                    u.param_shard.copy_(st['master'])
        for e in self.engines:
            if e.reshard:
                e._slices_dirty = True    # ZeRO-3: the slices are gathered again right before their next use; the
                continue                  # first gather after this step carries the one cross-rank barrier
            if e.nvlink and pushed_all.get(id(e), False):
                e.comm.barrier()          # every unit's new slice was pushed by its update kernel and has landed
            else:
                e.gather_params()         # at least one unit was updated locally only (fp32 / fp16 parameters,
                                          # several param groups, offload): exchange the slices
        return loss

    def _update_unit(self, e: FlatParamEngine, u, st: dict, hyper: list[dict]) -> bool:
        """Fused AdamW on one unit's slice (``hyper[i]`` = hyper-parameters of the user's param group ``i``).
        Returns True when the new parameters were also pushed to the peers."""
        shard = u.param_shard
        low = 'master' in st
        segments = self._segments(u)
        if self.offload:
            # gradient slice -> host, AdamW on the host copies (fp32 restatement), new weights -> device slice
            grad = u.local_grad().to('cpu', torch.float32)
            for gi, a, n in segments:
                h = dict(hyper[gi])
                if h.get('norm_sq') is not None:
                    h['norm_sq'] = h['norm_sq'].detach().to('cpu')
                adamw_step(st['master'][a:a + n], st['exp_avg'][a:a + n], st['exp_avg_sq'][a:a + n], grad[a:a + n],
                           None, **h)
            shard.copy_(st['master'], non_blocking=True)
            return False
        if e.nvlink and not e.reshard and low and shard.dtype == torch.bfloat16 and len(segments) == 1:
            # fused: AdamW on the fp32 slice + bf16 result stored into every peer's parameter arena
            e.comm.adamw_push(st['master'], st['exp_avg'], st['exp_avg_sq'], u.local_grad(), e.param_arena,
                              u.arena_offset * 2, u.shard_numel, **hyper[segments[0][0]])
            return True
        target = st['master'] if low else shard
        grad = u.local_grad()
        bf16_out = low and shard.dtype == torch.bfloat16
        for gi, a, n in segments:
            adamw_step(target[a:a + n], st['exp_avg'][a:a + n], st['exp_avg_sq'][a:a + n], grad[a:a + n],
                       shard[a:a + n] if bf16_out else None, **hyper[gi])
        if low and not bf16_out:
            shard.copy_(st['master'])
        return False

    def _overlap_enabled(self) -> bool:
        return not self.offload and all(e.overlap_optimizer and e._opt_stream is not None for e in self.engines)

    def _step_overlapped(self, hyper: list[dict]) -> None:
        """Per-unit update -> (push | gather) -> barrier -> grad clear on the engine's side stream, in the order the
        next forward pass touches the units; each unit's first use waits on its own event (engine forward
        pre-hooks), so the memory-bound AdamW and the parameter exchange run under the next step's GEMMs."""
        i = 0
        for e in self.engines:
            states = self.unit_state[i:i + len(e.units)]
            i += len(e.units)
            stream = e._opt_stream
            stream.wait_stream(torch.cuda.current_stream(e.device))
            order = sorted(range(len(e.units)), key=lambda k: (e.units[k].hook_seen, k))   # un-hooked units first
            background = _lib.kernels_available() and e.device.type == 'cuda'
            if background:
                _lib.lib().llmt_set_update_ctas_per_sm(_BACKGROUND_CTAS_PER_SM)
            with torch.cuda.stream(stream):
                for k in order:
                    u, st = e.units[k], states[k]
                    pushed = self._update_unit(e, u, st, hyper)
                    if pushed:
                        e.comm.barrier()              # this unit's slices have landed on every rank
                    else:
                        e.gather_unit(u)
                    if u.grad_live:
                        u.flat_grad.zero_()
                        u.grad_zeroed = True
                    if u.ready_event is None:
                        u.ready_event = torch.cuda.Event()
                    u.ready_event.record(stream)
                    u.ready_pending = True
            if background:
                _lib.lib().llmt_set_update_ctas_per_sm(8)

    def synchronize(self) -> None:
        """Make the current stream wait for in-flight unit updates (before reading parameters / state directly)."""
        for e in self.engines:
            e.synchronize_optimizer()

    def zero_grad(self, set_to_none: bool = True) -> None:
        for e in self.engines:
            e.zero_grad()

    # -- checkpointing: per-unit shards, keyed by unit name ---------------------------------------------
    def state_dict(self) -> dict:
        self.synchronize()
        units = {}
        for _, u, st in self._units():
            units[u.name] = {k: v for k, v in st.items() if k in ('master', 'exp_avg', 'exp_avg_sq')}
        sd = {'step': self.step_count, 'units': units,
              'param_groups': [{k: v for k, v in g.items() if k != 'params'} for g in self.param_groups]}
        if self._generic is not None:
            sd['generic'] = self._generic.state_dict()
        return sd

    def load_state_dict(self, state_dict: dict) -> None:
        self.synchronize()
        for e in self.engines:
            e._slices_dirty = True
        self.step_count = int(state_dict.get('step', 0))
        for _, u, st in self._units():
            saved = state_dict['units'].get(u.name, {})
            for k, v in saved.items():
                if k in st:
                    st[k].copy_(v)
            if 'master' in st and 'master' not in saved:
                st['master'].copy_(u.param_shard)
        for g, saved in zip(self.param_groups, state_dict.get('param_groups', [])):
            g.update(saved)
        if self._generic is not None and 'generic' in state_dict:
            self._generic.load_state_dict(state_dict['generic'])
