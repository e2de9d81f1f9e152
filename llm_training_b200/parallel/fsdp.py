"""Flat-parameter sharding engine (FSDP / ZeRO), B200-first.

What the reference gets from ``torch.distributed.fsdp.fully_shard`` per decoder layer
(src/llm_training/models/llama/llama_model.py:246-268) and from DeepSpeed ZeRO-1/2/3
(src/llm_training/lightning/strategy/deepspeed/deepspeed_strategy.py:16-102) is provided here by one
engine built around *flat per-unit buffers*:

* every sharding unit (embedding, each decoder layer, final norm, LM head) owns ONE contiguous parameter
  buffer and ONE contiguous gradient buffer; module parameters are views into them.  q|k|v and gate|up
  weights are therefore adjacent rows of one matrix, which is what lets ``ops.multi_linear`` run them as a
  single tcgen05 GEMM with no copy;
* rank ``r`` of the data-parallel axis owns slice ``r`` of each flat buffer: fp32 master weights and Adam
  moments exist only for that slice (ZeRO-1), gradients are reduce-scattered into that slice as soon as the
  unit's backward is complete (ZeRO-2, overlapped with the rest of backward on a side stream), and the
  updated bf16 slice is all-gathered back in place after the optimizer step;
* with 180 GB of HBM3e per GPU the bf16 parameters of an 8B model (16 GB) stay resident, so the gather is
  ONE in-place collective per unit per step instead of FSDP2's gather-before-forward + gather-before-backward
  (``reshard_after_forward``); ``reshard=True`` additionally drops a unit's full buffer between uses for
  models that need it (70B at dp=8).

Communication goes through :mod:`.comm` (gloo for CPU tests, NCCL or the NVLink kernels on GPUs).
"""
import logging
from typing import Any, Iterable

import torch
from torch import nn

from .comm import Comm, get_comm
from .mesh import MeshAxis

logger = logging.getLogger(__name__)

_ALIGN = 8   # elements: keeps every parameter 16-byte aligned for TMA / vector loads


class ShardedUnit:
    @staticmethod
    def layout(params: list[tuple[str, nn.Parameter]], world: int) -> tuple[list[int], int, int]:
        offsets, off = [], 0
        for _, p in params:
            off = (off + _ALIGN - 1) // _ALIGN * _ALIGN
            offsets.append(off)
            off += p.numel()
        chunk = _ALIGN * world
        return offsets, off, max(chunk, (off + chunk - 1) // chunk * chunk)

    def __init__(self, name: str, params: list[tuple[str, nn.Parameter]], world: int, rank: int,
                 device: torch.device, dtype: torch.dtype, grad_dtype: torch.dtype,
                 flat_param: torch.Tensor | None = None, flat_grad: torch.Tensor | None = None,
                 arena_offset: int = 0, reshard: bool = False, shard_store: torch.Tensor | None = None,
                 with_grad: bool = True) -> None:
        """``with_grad=False``: a frozen unit (e.g. DPO's reference model) — parameters are sharded / gathered like
        any other unit, but no gradient buffer, no reduce-scatter target and no optimizer state ever exist."""
        self.name = name
        self.with_grad = with_grad
        self.params = params
        self.world, self.rank = world, rank
        self.offsets, self.numel, self.padded_numel = self.layout(params, world)
        self.shard_numel = self.padded_numel // world
        self.arena_offset = arena_offset          # element offset inside the engine's symmetric arenas
        self.flat_param = flat_param if flat_param is not None else torch.zeros(self.padded_numel, dtype=dtype,
                                                                                 device=device)
        if not with_grad:
            self.flat_grad = torch.zeros(0, dtype=grad_dtype, device=device)
        else:
            self.flat_grad = flat_grad if flat_grad is not None else torch.zeros(self.padded_numel, dtype=grad_dtype,
                                                                                  device=device)
        for (_, p), o in zip(params, self.offsets):
            view = self.flat_param[o:o + p.numel()].view(p.shape)
            if not p.is_meta:
                view.copy_(p.data)
            p.data = view
            if with_grad:
                p.grad = self.flat_grad[o:o + p.numel()].view(p.shape)
        # reduce-scatter target: separate buffer so the in-flight collective never aliases live grads
        self.grad_shard = (self.flat_grad if (world == 1 or not with_grad)
                           else torch.zeros(self.shard_numel, dtype=grad_dtype, device=device))
        self.pending_grads = 0
        self.reduced = False
        self.work: Any = None
        # ZeRO-3 mode: the owned slice lives in its own persistent tensor; the full parameter / gradient buffers
        # only have storage between `unshard` and `reshard` (their tensor objects, and every view of them held
        # by parameters or saved by autograd, stay valid across the storage resize)
        self.reshard = reshard and world > 1
        self._shard: torch.Tensor | None = None
        if self.reshard:
            self._shard = shard_store if shard_store is not None else torch.zeros(
                self.shard_numel, dtype=dtype, device=device)
            self._shard.copy_(self.flat_param[rank * self.shard_numel:(rank + 1) * self.shard_numel])
        self.is_sharded = False       # flat_param currently has no storage
        self.grad_live = with_grad    # flat_grad currently has storage (and holds zeros or partial sums)
        self.in_backward = False
        self.gather_event: Any = None
        self.gather_pending = False
        # ZeRO-3 on NVLink: row-stacked weight groups whose consuming GEMM gathers them itself in forward (q|k|v and
        # gate|up of a decoder layer): key -> {'off': element offset, 'n': rows, 'k': cols, 'pending': bool}.  The
        # side-stream gather covers the rest of the unit.
        self.lazy_heads: dict[int, dict] = {}

    @property
    def head_pending(self) -> bool:
        return any(h['pending'] for h in self.lazy_heads.values())

    @property
    def param_shard(self) -> torch.Tensor:
        if self._shard is not None:
            return self._shard
        s = self.shard_numel
        return self.flat_param[self.rank * s:(self.rank + 1) * s]

    def _set_storage(self, t: torch.Tensor, present: bool) -> None:
        t.untyped_storage().resize_(self.padded_numel * t.element_size() if present else 0)

    def local_grad(self) -> torch.Tensor:
        return self.grad_shard[:self.shard_numel] if self.world == 1 else self.grad_shard

    def shard_slices(self) -> Iterable[tuple[str, nn.Parameter, int, int, int]]:
        """(name, param, start_in_param, start_in_shard, length) for the part of each parameter that
        falls into this rank's shard."""
        lo, hi = self.rank * self.shard_numel, (self.rank + 1) * self.shard_numel
        for (n, p), o in zip(self.params, self.offsets):
            a, b = max(o, lo), min(o + p.numel(), hi)
            if a < b:
                yield n, p, a - o, a - lo, b - a


class FlatParamEngine:
    def __init__(self, units: list[tuple[str, list[tuple[str, nn.Parameter]]]], dp_axis: MeshAxis | None,
                 device: torch.device, comm: Comm | None = None, grad_dtype: torch.dtype | None = None,
                 tp_axis: MeshAxis | None = None, reshard: bool = False, frozen: bool = False) -> None:
        """``frozen=True``: every unit is parameter-only (``ShardedUnit(with_grad=False)``); the engine then only
        gathers before use and drops after use (ZeRO-3 for a model that is never trained)."""
        self.frozen = bool(frozen)
        self.axis = dp_axis if dp_axis is not None else MeshAxis('data_parallel', [0], 0, None)
        # tensor-parallel peers hold different shards of most parameters but REPLICAS of a few (norm weights);
        # under sequence parallelism the replicas' gradients are partial sums over the tp axis
        self.tp_axis = tp_axis if (tp_axis is not None and tp_axis.size() > 1) else None
        self.comm = comm if comm is not None else get_comm(self.axis)
        self.device = device
        self.world_is_cuda = device.type == 'cuda'
        self.world, self.rank = self.axis.size(), self.axis.rank
        self.units: list[ShardedUnit] = []
        units = [(n, ps) for n, ps in units if ps]
        self.nvlink = bool(getattr(self.comm, 'is_nvlink', False)) and self.world > 1
        self.param_arena = self.grad_arena = self.shard_arena = None
        self.reshard = bool(reshard) and self.world > 1
        dtypes = {ps[0][1].dtype for _, ps in units}
        if self.reshard:
            # ZeRO-3: only the owned slices are persistent.  On NVLink they sit in ONE symmetric arena so that any
            # rank can pull any slice (csrc/kernels/comm.cu gather_shards / the weight-gathering GEMM); gradients
            # are reduced by the generic reduce-scatter of the backend.
            shard_sizes = [ShardedUnit.layout(ps, self.world)[2] // self.world for _, ps in units]
            if self.nvlink and len(dtypes) == 1:
                self.shard_arena = self.comm.alloc(sum(shard_sizes), next(iter(dtypes)))
                self.shard_arena.tensor.zero_()
            off = 0
            for (name, params), ssz in zip(units, shard_sizes):
                dtype = params[0][1].dtype
                assert all(p.dtype == dtype for _, p in params), f'unit {name}: mixed parameter dtypes'
                store = self.shard_arena.tensor[off:off + ssz] if self.shard_arena is not None else None
                self.units.append(ShardedUnit(name, params, self.world, self.rank, device, dtype, grad_dtype or dtype,
                                              arena_offset=off, reshard=True, shard_store=store,
                                              with_grad=not self.frozen))
                off += ssz
            # ZeRO-3 gradients on NVLink: a unit's full gradient buffer is ordinary (resizable) memory, so right
            # before its reduce-scatter it is staged into one of TWO symmetric slots that every peer can read (or the
            # switch can reduce).  Two are enough: the barrier that opens reduce-scatter j+1 proves every rank has
            # finished pulling slot (j mod 2), and the copy for unit j+2 is queued behind it on the same stream.
            self.grad_ring = None
            import os as _os
            if self.shard_arena is not None and not self.frozen and _os.environ.get('LLMT_ZERO3_NVLINK_GRADS', '1') != '0':
                biggest = max(ShardedUnit.layout(ps, self.world)[2] for _, ps in units)
                self.grad_ring = [self.comm.alloc(biggest, next(iter(dtypes))) for _ in range(2)]
                self._ring_next = 0
        elif self.nvlink and len(dtypes) == 1 and not self.frozen:
            # one symmetric allocation for all parameters and one for all gradients: peers address any unit as
            # arena base + offset, and a single rendezvous exchanges the handles
            dtype = next(iter(dtypes))
            sizes = [ShardedUnit.layout(ps, self.world)[2] for _, ps in units]
            total = sum(sizes)
            self.param_arena = self.comm.alloc(total, dtype)
            self.grad_arena = self.comm.alloc(total, dtype)
            self.param_arena.tensor.zero_()
            self.grad_arena.tensor.zero_()
            off = 0
            for (name, params), size in zip(units, sizes):
                self.units.append(ShardedUnit(name, params, self.world, self.rank, device, dtype, dtype,
                                              flat_param=self.param_arena.tensor[off:off + size],
                                              flat_grad=self.grad_arena.tensor[off:off + size], arena_offset=off))
                off += size
        else:
            self.nvlink = False
            for name, params in units:
                dtype = params[0][1].dtype
                assert all(p.dtype == dtype for _, p in params), f'unit {name}: mixed parameter dtypes'
                self.units.append(ShardedUnit(name, params, self.world, self.rank, device, dtype, grad_dtype or dtype,
                                              with_grad=not self.frozen))
        # `nvlink` = parameters travel through our kernels; `nvlink_grads` = gradients do too (resident mode: straight
        # from the symmetric gradient arena; ZeRO-3: through the staging ring)
        self.nvlink_grads = self.nvlink and (not self.reshard or getattr(self, 'grad_ring', None) is not None)
        import os
        self.fused_first_gemm = os.environ.get('LLMT_FUSED_FIRST_GEMM', '1') != '0'
        # ZeRO-3 on NVLink: the persistent slices change only in the optimizer step (and when weights are loaded), so
        # ONE cross-rank barrier after such a write covers every gather until the next one — a barrier per unit would
        # couple the ranks' forward and backward passes layer by layer
        self._slices_dirty = True
        self._forward_order: list[ShardedUnit] = []      # hooked units in the order the last forward used them
        self._norm_accumulating = False
        self.require_sync = True
        self._comm_stream = torch.cuda.Stream(device) if device.type == 'cuda' and self.world > 1 else None
        self._norm_sq = torch.zeros(1, dtype=torch.float32, device=device)
        # optimizer overlap (see ShardedOptimizer.step): per-unit updates run on a side stream while the next
        # forward pass starts; a unit's first use waits on the event recorded after its update
        self.overlap_optimizer = False
        self._opt_stream = self._comm_stream if self._comm_stream is not None else (
            torch.cuda.Stream(device) if device.type == 'cuda' else None)
        for u in self.units:
            u.ready_event = None          # recorded on the optimizer stream after the unit's update (+ gather)
            u.ready_pending = False
            u.hook_seen = False           # the unit's module runs a forward pre-hook (else: waited at model entry)
            u.grad_zeroed = False         # flat_grad was cleared by the optimizer stream after the update
        self._param_to_unit: dict[int, ShardedUnit] = {}
        for u in self.units:
            for _, p in u.params:
                self._param_to_unit[id(p)] = u
                if p.requires_grad:
                    p.register_post_accumulate_grad_hook(self._on_grad_ready)
        self._reset_pending()

    # ------------------------------------------------------------------ gradients
    def _reset_pending(self) -> None:
        for u in self.units:
            u.pending_grads = sum(1 for _, p in u.params if p.requires_grad)
            u.ready_ids = set()
            u.reduced = False
            u.work = None
        # Gradient reductions are issued in ONE order on every rank — last unit first, the order backward completes
        # them in — whatever order the gradients really arrive in: the NVLink collectives pair ranks up by a running
        # epoch, so a rank whose unit k got no gradient (unused parameters, a branch not taken on its batch) must not
        # let unit k-1 overtake it.  A unit that is not ready holds back the ones before it until `finish_backward`.
        self._next_reduce = len(self.units) - 1

    def _on_grad_ready(self, param: nn.Parameter) -> None:
        """One notification per parameter and backward pass.  A parameter can report twice — the wgrad GEMM that
        accumulated straight into the flat buffer reports it, and autograd's AccumulateGrad node still runs its
        post-accumulate hook for the ``None`` gradient the op returned — so readiness is tracked per parameter,
        not by counting calls."""
        u = self._param_to_unit[id(param)]
        u.grad_zeroed = False
        if id(param) in u.ready_ids:
            return
        u.ready_ids.add(id(param))
        u.pending_grads -= 1
        if u.pending_grads == 0 and self.require_sync:
            self._launch_ready_reductions()

    def _launch_ready_reductions(self, flush: bool = False) -> None:
        """Issue the reductions of the longest run of complete units, in the fixed order (see `_reset_pending`);
        ``flush``: the backward pass is over, issue everything that is left."""
        while self._next_reduce >= 0:
            u = self.units[self._next_reduce]
            if not flush and u.pending_grads > 0:
                return
            self._next_reduce -= 1
            if not u.reduced:
                self._launch_reduce(u)

    def _settle_fresh(self, u: ShardedUnit) -> None:
        """Weights whose zeroing was deferred to their first wgrad GEMM (`zero_grad`) but that received NO gradient
        this step still hold last step's values: clear them before anything reads the buffer."""
        for _, p in u.params:
            if getattr(p, '_llmt_grad_fresh', False):
                p._llmt_grad_fresh = False
                if p.grad is not None:
                    p.grad.zero_()

    def _launch_reduce(self, u: ShardedUnit) -> None:
        if u.reduced:
            return
        u.reduced = True
        self._settle_fresh(u)
        if self.world == 1:
            if self.tp_axis is not None:
                self._reduce(u)
            return
        if self._comm_stream is not None:
            self._comm_stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._comm_stream):
                self._reduce(u)
        else:
            self._reduce(u)
        if self.reshard:
            self._after_unit_backward(u)

    def _reduce(self, u: ShardedUnit) -> None:
        if self.tp_axis is not None and self.world > 1:
            import torch.distributed as dist
            for _, p in u.params:
                if getattr(p, '_tp_replicated', False) and p.grad is not None:
                    dist.all_reduce(p.grad, group=self.tp_axis.group)
        if self.world == 1:
            return                      # pure tensor parallel: the replicas are summed once, in finish_backward
        if self.nvlink_grads and self.reshard:
            slot = self.grad_ring[self._ring_next]
            self._ring_next ^= 1
            slot.tensor[:u.padded_numel].copy_(u.flat_grad)      # on the communication stream, behind the unit's backward
            self.comm.reduce_scatter_sumsq(slot, 0, u.shard_numel, u.grad_shard, self._norm_sq,
                                           accumulate_norm=self._norm_accumulating)
            self._norm_accumulating = True
        elif self.nvlink_grads:
            # barrier + pull of this rank's slice from every peer + squared-norm partial, one pass over NVLink
            esize = u.flat_grad.element_size()
            self.comm.reduce_scatter_sumsq(self.grad_arena, u.arena_offset * esize, u.shard_numel, u.grad_shard,
                                           self._norm_sq, accumulate_norm=self._norm_accumulating)
            self._norm_accumulating = True
        else:
            self.comm.reduce_scatter_into(u.grad_shard, u.flat_grad, average=True)

    def enable_direct_grad_accumulation(self, flag: bool = True) -> None:
        """Let the wgrad GEMMs accumulate straight into the flat gradient buffer (``ops.linear_op.wgrad``).  Only
        valid when every weight is used ONCE per backward pass (the engine counts one notification per
        parameter to know when a unit's gradients are complete) — CLM turns it on, DPO / ORPO do not."""
        def settle_before_accumulate(grad, p):
            # a gradient is about to arrive through autograd's AccumulateGrad (an op that does not use the direct
            # wgrad path: fused linear + loss, embedding): a deferred zeroing has to happen before the `+=`
            if getattr(p, '_llmt_grad_fresh', False):
                p._llmt_grad_fresh = False
                if p.grad is not None:
                    p.grad.zero_()
            return grad

        for u in self.units:
            for _, p in u.params:
                if p.requires_grad and p.dim() in (1, 2):          # GEMM weights and norm weights
                    p._llmt_direct_grad = flag
                    p._llmt_grad_ready = self._on_grad_ready
                    if flag and not getattr(p, '_llmt_settle_hook', False):
                        p.register_hook(lambda g, p=p: settle_before_accumulate(g, p))
                        p._llmt_settle_hook = True

    def set_requires_gradient_sync(self, flag: bool) -> None:
        """False on non-boundary micro-batches of gradient accumulation (reference:
        ``block_backward_sync``, fsdp2_strategy.py:430-442): grads keep accumulating locally."""
        self.require_sync = flag

    def finish_backward(self) -> None:
        """Make every unit's reduced gradient shard available on the current stream."""
        if self.frozen:
            return
        if not self.require_sync:
            self._rearm()
            return
        if self.reshard:
            for u in self.units:       # un-hooked units / units that saw no backward still need gradient storage
                if not u.reduced:
                    self._alloc_grad(u)
        self._launch_ready_reductions(flush=True)      # params that received no grad this step (unused / frozen siblings)
        if self._comm_stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self._comm_stream)
        self._replica_flat = None
        if self.tp_axis is not None and self.world == 1:
            self._allreduce_tp_replicas()
        if self.reshard:
            for u in self.units:
                u.in_backward = False
                self.reshard_unit(u)
                self._free_grad(u)

    def _allreduce_tp_replicas(self) -> None:
        """Pure tensor parallel (no data-parallel sharding): the gradients of the parameters replicated over the tp
        axis (norm weights, biases of row-parallel layers) are partial sums.  ONE all-reduce over their concatenation
        instead of one per parameter (65 for a 32-layer model); the summed copy also serves the norm correction."""
        import torch.distributed as dist
        views = [p.grad.view(-1) for u in self.units for _, p in u.params
                 if getattr(p, '_tp_replicated', False) and p.grad is not None]
        if not views:
            return
        flat = torch.cat(views)
        dist.all_reduce(flat, group=self.tp_axis.group)
        torch._foreach_copy_(views, list(flat.split([v.numel() for v in views])))
        self._replica_flat = flat

    def _rearm(self) -> None:
        for u in self.units:
            u.pending_grads = sum(1 for _, p in u.params if p.requires_grad)
            u.ready_ids = set()
            if self.reshard:              # non-boundary micro-batch: keep the partial gradients, drop the parameters
                u.in_backward = False
                self.reshard_unit(u)

    # ------------------------------------------------------------------ optimizer overlap
    def attach_module_hooks(self, model: nn.Module) -> None:
        """Forward pre-hooks that make a unit's first use wait for its (possibly still running) optimizer update."""
        by_name: dict[str, list[ShardedUnit]] = {}
        for u in self.units:
            by_name.setdefault(u.name.split('#')[0], []).append(u)
        unit_modules = dict(model.sharding_units()) if hasattr(model, 'sharding_units') else {}

        def make_hook(units: list[ShardedUnit]):
            def hook(module, args):
                for u in units:
                    u.hook_seen = True
                    self.wait_unit(u)
                    if self.reshard:
                        self._pre_forward(u)
            return hook

        def make_post_hook(units: list[ShardedUnit]):
            def hook(module, args, output):
                if not self.reshard:
                    return
                self._post_forward(units, output)
            return hook

        for name, module in unit_modules.items():
            if name in by_name and module is not model:
                module.register_forward_pre_hook(make_hook(by_name[name]))
                module.register_forward_hook(make_post_hook(by_name[name]))

        def root_hook(module, args):
            if self.reshard:
                if self.frozen and self._forward_order:     # no backward ever records the order for a frozen model
                    self._last_forward_order = list(self._forward_order)
                self._forward_order = []
            for u in self.units:
                if not u.hook_seen:
                    self.wait_unit(u)
                    if self.reshard:          # units used without a module call stay whole for the entire step
                        self.unshard(u)
                        self.wait_unshard(u)
                        if torch.is_grad_enabled():
                            self._alloc_grad(u)

        model.register_forward_pre_hook(root_hook)

    # ------------------------------------------------------------------ ZeRO-3: gather before use, drop after use
    def enable_fused_first_gemm(self, model: nn.Module) -> None:
        """ZeRO-3 + NVLink: let the column GEMMs that open each half of a decoder layer — the q|k|v projection and
        the gate|up projection (285 of a LLaMA-3-8B layer's 436 MB) — pull their own weights from the owners' slices
        while they run (``NVLinkComm.gemm_weight_gather``) instead of waiting for a gather."""
        if self.shard_arena is None or not hasattr(model, 'sharding_units'):
            return
        n_max = 0
        for _, module in model.sharding_units():
            groups = []
            attn, mlp = getattr(module, 'self_attn', None), getattr(module, 'mlp', None)
            if attn is not None and all(hasattr(attn, n) for n in ('q_proj', 'k_proj', 'v_proj')):
                groups.append([attn.q_proj.weight, attn.k_proj.weight, attn.v_proj.weight])
            elif attn is not None and hasattr(attn, 'qkv_proj'):
                groups.append([attn.qkv_proj.weight])
            if mlp is not None and all(hasattr(mlp, n) for n in ('gate_proj', 'up_proj')):
                groups.append([mlp.gate_proj.weight, mlp.up_proj.weight])
            elif mlp is not None and hasattr(mlp, 'gate_up_proj'):
                groups.append([mlp.gate_up_proj.weight])
            for ws in groups:
                u = self._param_to_unit.get(id(ws[0]))
                if u is None or any(self._param_to_unit.get(id(w)) is not u for w in ws):
                    continue
                offs = {id(p): o for (_, p), o in zip(u.params, u.offsets)}
                k = ws[0].shape[1]
                ok, expect = True, offs[id(ws[0])]
                for w in ws:
                    ok &= w.dim() == 2 and w.shape[1] == k and w.dtype == torch.bfloat16 and offs[id(w)] == expect
                    expect += w.numel()
                if not ok or k % 8:
                    continue
                key = id(ws[0])
                u.lazy_heads[key] = {'off': offs[id(ws[0])], 'n': sum(w.shape[0] for w in ws), 'k': k,
                                     'pending': False}
                n_max = max(n_max, u.lazy_heads[key]['n'])
                for w in ws:
                    w._llmt_lazy_unit = (self, u, key)
        if n_max:
            self._head_flags = torch.zeros(n_max // 256 + 2, dtype=torch.int32, device=self.device)

    def first_gemm(self, x: torch.Tensor, u: ShardedUnit, key: int) -> torch.Tensor:
        """``x @ W^T`` for the lazy weight group ``key`` of ``u`` while the GEMM itself gathers W."""
        h = u.lazy_heads[key]
        esize = u.flat_param.element_size()
        y = self.comm.gemm_weight_gather(x, self.shard_arena, u.arena_offset * esize, u.flat_param, h['off'],
                                         u.shard_numel, h['n'], self._head_flags)
        h['pending'] = False
        return y

    def complete_head(self, u: ShardedUnit, key: int | None = None) -> None:
        """Some other consumer needs lazy weights first: gather them now on the current stream."""
        esize = u.flat_param.element_size()
        for hk, h in u.lazy_heads.items():
            if h['pending'] and (key is None or hk == key):
                lo, hi = h['off'] * esize, (h['off'] + h['n'] * h['k']) * esize
                self.comm.gather_range(self.shard_arena, u.arena_offset * esize, u.shard_numel * esize, u.flat_param,
                                       lo, hi, barrier=False)
                h['pending'] = False

    def unshard(self, u: ShardedUnit, lazy_head: bool = False) -> None:
        """Give ``flat_param`` storage and start gathering every rank's slice into it on the side stream.  With
        ``lazy_head`` (forward pass) the q|k|v rows are left to the GEMM that consumes them."""
        if not u.is_sharded:
            return
        u._set_storage(u.flat_param, True)
        u.is_sharded = False
        lazy = bool(lazy_head and u.lazy_heads and self._comm_stream is not None and self.fused_first_gemm)
        for h in u.lazy_heads.values():
            h['pending'] = lazy
        main = torch.cuda.current_stream(self.device) if self._comm_stream is not None else None
        if self._comm_stream is not None:
            self._comm_stream.wait_stream(main)        # the slice may have just been written by the optimizer
            with torch.cuda.stream(self._comm_stream):
                self._gather_into_full(u)
                if u.gather_event is None:
                    u.gather_event = torch.cuda.Event()
                u.gather_event.record(self._comm_stream)
            u.gather_pending = True
        else:
            self._gather_into_full(u)

    def _gather_into_full(self, u: ShardedUnit) -> None:
        if self.shard_arena is not None:
            esize = u.flat_param.element_size()
            need_barrier, self._slices_dirty = self._slices_dirty, False
            if u.head_pending:
                # everything except the lazy groups: the ranges between them, one launch each, one barrier in front
                pos, first = 0, True
                spans = sorted((h['off'] * esize, (h['off'] + h['n'] * h['k']) * esize)
                               for h in u.lazy_heads.values() if h['pending'])
                for lo, hi in spans + [(u.padded_numel * esize, u.padded_numel * esize)]:
                    if lo > pos or (first and need_barrier):
                        self.comm.gather_range(self.shard_arena, u.arena_offset * esize, u.shard_numel * esize,
                                               u.flat_param, pos, max(lo, pos), barrier=first and need_barrier)
                        first = False
                    pos = max(pos, hi)
                return
            self.comm.gather_shards(self.shard_arena, u.arena_offset * esize, u.shard_numel * esize, u.flat_param,
                                    barrier=need_barrier)
        else:
            self.comm.all_gather_into(u.flat_param, u.param_shard)

    def wait_unshard(self, u: ShardedUnit, keep_lazy_head: bool = False) -> None:
        if u.gather_pending:
            torch.cuda.current_stream(self.device).wait_event(u.gather_event)
            u.gather_pending = False
        if u.head_pending and not keep_lazy_head:
            self.complete_head(u)

    def reshard_unit(self, u: ShardedUnit) -> None:
        if u.is_sharded or not self.reshard:
            return
        self.wait_unshard(u, keep_lazy_head=True)     # never free under an in-flight gather
        for h in u.lazy_heads.values():
            h['pending'] = False
        u._set_storage(u.flat_param, False)
        u.is_sharded = True

    def _pre_forward(self, u: ShardedUnit) -> None:
        fwd = not u.in_backward
        self.unshard(u, lazy_head=fwd)
        self.wait_unshard(u, keep_lazy_head=fwd)      # the q|k|v GEMM finishes the gather itself
        if fwd:
            if u not in self._forward_order:
                self._forward_order.append(u)
            nxt = self._next_in_order(u, +1)
            if nxt is not None:
                self.unshard(nxt, lazy_head=True)     # prefetch: its gather runs under this unit's compute

    def _next_in_order(self, u: ShardedUnit, step: int) -> ShardedUnit | None:
        order = self._last_forward_order if getattr(self, '_last_forward_order', None) else None
        if not order or u not in order:
            return None
        k = order.index(u) + step
        return order[k] if 0 <= k < len(order) else None

    def _post_forward(self, units: list[ShardedUnit], output: Any) -> None:
        if any(u.in_backward for u in units):
            return                        # activation recompute inside backward: parameters are needed right after
        if torch.is_grad_enabled():
            outs = output if isinstance(output, (tuple, list)) else (output,)
            for t in outs:
                if torch.is_tensor(t) and t.requires_grad:
                    t.register_hook(lambda g, units=units: self._pre_backward(units))
        for u in units:
            self.reshard_unit(u)

    def _pre_backward(self, units: list[ShardedUnit]) -> None:
        self._last_forward_order = list(self._forward_order) or getattr(self, '_last_forward_order', [])
        for u in units:
            if u.in_backward:
                continue
            u.in_backward = True
            self.unshard(u)
            self.wait_unshard(u)
            self._alloc_grad(u)
            prev = self._next_in_order(u, -1)
            if prev is not None and not prev.in_backward:
                self.unshard(prev)        # prefetch for the unit whose backward runs next
        return None

    def _alloc_grad(self, u: ShardedUnit) -> None:
        if not u.grad_live and u.with_grad:
            u._set_storage(u.flat_grad, True)
            direct = [getattr(p, '_llmt_direct_grad', False) and p.requires_grad for _, p in u.params]
            if self.require_sync and self.world_is_cuda and any(direct):
                # fresh storage: the weights whose first wgrad GEMM overwrites need no clearing (see zero_grad); the
                # rest of the buffer — other parameters, alignment gaps, tail padding — does, because the
                # reduce-scatter and the optimizer read every element of the slice
                pos = 0
                for (_, p), o, d in zip(u.params, u.offsets, direct):
                    end = o + p.numel()
                    if d:
                        p._llmt_grad_fresh = True
                        if o > pos:
                            u.flat_grad[pos:o].zero_()
                    else:
                        u.flat_grad[pos:end].zero_()
                    pos = end
                if pos < u.padded_numel:
                    u.flat_grad[pos:].zero_()
            else:
                u.flat_grad.zero_()
            u.grad_live = True
            u.grad_zeroed = True

    def _after_unit_backward(self, u: ShardedUnit) -> None:
        """The unit's gradients were handed to the reduce-scatter: drop its full parameter and gradient buffers."""
        if not u.hook_seen:
            return                        # un-hooked units are dropped in finish_backward
        self.reshard_unit(u)
        self._free_grad(u)

    def _free_grad(self, u: ShardedUnit) -> None:
        if u.grad_live and self.require_sync and self.world > 1:
            if self._comm_stream is not None:
                u.flat_grad.record_stream(self._comm_stream)     # the reduce-scatter may still be reading it
            u._set_storage(u.flat_grad, False)
            u.grad_live = False

    def wait_unit(self, u: ShardedUnit) -> None:
        if u.ready_pending:
            torch.cuda.current_stream(self.device).wait_event(u.ready_event)
            u.ready_pending = False

    def synchronize_optimizer(self) -> None:
        """Current stream waits for every in-flight unit update (checkpointing, evaluation, timing boundaries)."""
        for u in self.units:
            self.wait_unit(u)

    def zero_grad(self) -> None:
        for u in self.units:
            if not u.grad_live or not u.with_grad:           # ZeRO-3: no storage until the unit's next backward allocates + clears it
                continue
            for (_, p), o in zip(u.params, u.offsets):   # something may have replaced .grad (set_to_none)
                if p.grad is None or p.grad.data_ptr() != u.flat_grad.data_ptr() + o * u.flat_grad.element_size():
                    p.grad = u.flat_grad[o:o + p.numel()].view(p.shape)
            if not u.grad_zeroed:
                self.wait_unit(u)
                direct = [getattr(p, '_llmt_direct_grad', False) and p.requires_grad for _, p in u.params]
                if self.require_sync and self.world_is_cuda and any(direct):
                    # weights that accumulate straight from their wgrad GEMM are not cleared here: their first GEMM
                    # of the step overwrites (ops.linear_op.wgrad).  Everything else (norm weights, embeddings,
                    # alignment gaps, which are never written) is cleared as before.
                    for (_, p), o, d in zip(u.params, u.offsets, direct):
                        if d:
                            p._llmt_grad_fresh = True
                        else:
                            u.flat_grad[o:o + p.numel()].zero_()
                else:
                    u.flat_grad.zero_()
        self._reset_pending()

    def grad_norm_sq(self, extra_axes: list[Comm] | None = None) -> torch.Tensor:
        """Device scalar: squared L2 norm of the (already reduced) gradient, summed over the dp axis
        (and any extra axes the parameters are sharded over).  No host sync."""
        from ..ops.adamw_op import sumsq
        if self.nvlink_grads:
            # per-unit partial norms were accumulated by the reduce-scatter kernels; one tiny all-reduce is left
            self._norm_accumulating = False
            self._discount_tp_replicas()
            self.comm.all_reduce(self._norm_sq)
            for c in extra_axes or []:
                c.all_reduce(self._norm_sq)
            return self._norm_sq
        first = True
        for u in self.units:
            sumsq(u.local_grad(), self._norm_sq, accumulate=not first)
            first = False
        if first:
            self._norm_sq.zero_()
        self._discount_tp_replicas()
        self.comm.all_reduce(self._norm_sq)
        for c in extra_axes or []:
            c.all_reduce(self._norm_sq)
        return self._norm_sq

    def _discount_tp_replicas(self) -> None:
        """The norm is summed over the tp axis too; parameters replicated across it must count once."""
        if self.tp_axis is None or self.tp_axis.rank == 0:
            return
        flat = getattr(self, '_replica_flat', None)
        if flat is None:                 # sharded over dp as well: collect this rank's slices of the replicas
            parts = []
            for u in self.units:
                g = u.local_grad()
                parts += [g[start:start + length] for _, p, _, start, length in u.shard_slices()
                          if getattr(p, '_tp_replicated', False)]
            if not parts:
                return
            flat = torch.cat(parts)
        self._norm_sq -= flat.float().pow(2).sum()       # one pass instead of four launches per parameter

    # ------------------------------------------------------------------ parameters
    def gather_unit(self, u: ShardedUnit) -> None:
        """Rebuild ONE unit's full buffer from the updated slices (used by the overlapped optimizer step)."""
        if self.world == 1 or self.reshard:       # ZeRO-3: the next use gathers
            return
        if self.nvlink:
            esize = self.param_arena.tensor.element_size()
            self.comm.allgather_pull(self.param_arena, u.arena_offset * esize, u.shard_numel * esize)
            self.comm.barrier()
            return
        self.comm.all_gather_into(u.flat_param, u.param_shard)

    def gather_params(self) -> None:
        """After the optimizer updated every rank's slice: rebuild the full bf16 buffers in place."""
        if self.world == 1 or self.reshard:
            return
        if self.nvlink:
            esize = self.param_arena.tensor.element_size()
            for u in self.units:
                self.comm.allgather_pull(self.param_arena, u.arena_offset * esize, u.shard_numel * esize)
            self.comm.barrier()
            return
        for u in self.units:
            self.comm.all_gather_into(u.flat_param, u.param_shard)

    def broadcast_params(self, src_index: int = 0) -> None:
        for u in self.units:
            self.comm.broadcast(u.flat_param, src_index)
        self.sync_shards_from_full()

    # ------------------------------------------------------------------ ZeRO-3: whole-model access
    def sync_shards_from_full(self) -> None:
        """After something wrote the FULL parameters directly (weight loading, checkpoint restore, init
        broadcast): refresh the persistent owned slices.  No-op in resident mode."""
        if not self.reshard:
            return
        self._slices_dirty = True
        for u in self.units:
            if not u.is_sharded:
                s = u.shard_numel
                u._shard.copy_(u.flat_param[self.rank * s:(self.rank + 1) * s])

    def unshard_all(self) -> None:
        for u in self.units:
            self.wait_unit(u)
            if self.reshard:
                self.unshard(u)
                self.wait_unshard(u)

    def reshard_all(self) -> None:
        if self.reshard:
            for u in self.units:
                self.reshard_unit(u)

    def unsharded(self, writeback: bool = False):
        """Context manager: every parameter is whole inside (state_dict, export, evaluation helpers); with
        ``writeback`` the owned slices are refreshed from whatever was written."""
        import contextlib

        @contextlib.contextmanager
        def ctx():
            self.unshard_all()
            try:
                yield self
            finally:
                if writeback:
                    self.sync_shards_from_full()
                self.reshard_all()
        return ctx()

    def num_parameters(self) -> int:
        return sum(u.numel for u in self.units)


def _collect_units(model: nn.Module, only_trainable: bool = True
                   ) -> list[tuple[str, list[tuple[str, nn.Parameter]]]]:
    """Group parameters by the model's ``sharding_units()`` (fallback: one unit for everything)."""
    claimed: set[int] = set()
    groups: list[tuple[str, list[tuple[str, nn.Parameter]]]] = []
    unit_list = model.sharding_units() if hasattr(model, 'sharding_units') else [('', model)]
    for name, module in unit_list:
        params = []
        for pn, p in module.named_parameters():
            if id(p) in claimed or (only_trainable and not p.requires_grad):
                continue
            claimed.add(id(p))
            params.append((f'{name}.{pn}' if name else pn, p))
        # split by dtype so every flat buffer is homogeneous
        by_dtype: dict[torch.dtype, list] = {}
        for item in params:
            by_dtype.setdefault(item[1].dtype, []).append(item)
        for i, (dt, items) in enumerate(by_dtype.items()):
            groups.append((name if i == 0 else f'{name}#{dt}', items))
    rest = [(pn, p) for pn, p in model.named_parameters()
            if id(p) not in claimed and (p.requires_grad or not only_trainable)]
    if rest:
        groups.append(('_rest', rest))
    return groups


def _resolve_reshard(mode: Any, units: list, dp_mesh: MeshAxis | None, device: torch.device,
                     frozen: bool = False) -> bool:
    """``True`` / an int -> ZeRO-3 (full buffers only around a unit's use), ``False`` -> resident parameters
    (ZeRO-2), ``'auto'`` (our default) -> resident unless bf16 parameters + gradients would take more than 40% of
    the device.  ``LLMT_RESHARD=0|1`` overrides.  The reference's default is ``True``
    (fsdp2_strategy.py:56); at 180 GB per GPU resident is faster whenever it fits."""
    import os
    if dp_mesh is None or dp_mesh.size() <= 1:
        return False
    env = os.environ.get('LLMT_RESHARD')
    if env in ('0', '1'):
        return env == '1'
    if isinstance(mode, str) or mode is None:
        if device.type != 'cuda':
            return False
        need = (1 if frozen else 2) * sum(p.numel() * p.element_size() for _, ps in units for _, p in ps)
        return need > (0.2 if frozen else 0.4) * torch.cuda.get_device_properties(device).total_memory
    if not isinstance(mode, bool) and isinstance(mode, int):
        # FSDP2: an int N means "after forward keep the parameters sharded over N ranks" (a middle ground between
        # True and False).  One NVSwitch box has no cheaper sub-group than the full group, so N > 1 is the same
        # schedule as True here; N == 1 ("keep whole parameters") is False.
        logger.info(f'reshard_after_forward={mode}: treated as {mode > 1} (no intermediate shard group on one box)')
        return mode > 1
    return bool(mode)


def apply_fully_sharded(model: nn.Module, dp_mesh: MeshAxis | None, reshard_after_forward: bool | int | str = 'auto',
                        mp_policy: Any = None, offload_policy: Any = None, device: torch.device | None = None,
                        comm: Comm | None = None, **_: Any) -> FlatParamEngine:
    """Build the flat-parameter engine for ``model`` over the data-parallel axis and attach it as
    ``model._flat_engine``.  ``reshard_after_forward`` selects ZeRO-3 vs resident parameters (see
    ``_resolve_reshard``); ``offload_policy`` is accepted for API parity."""
    if getattr(model, '_flat_engine', None) is not None:
        return model._flat_engine
    if device is None:
        first = next((p for p in model.parameters() if not p.is_meta), None)
        device = first.device if first is not None else torch.device('cpu')
    # gradients are reduced in the parameter dtype (bf16 for `bf16-true`, like the reference's default
    # MixedPrecisionPolicy, fsdp2_precision.py:72-76); `.grad` must match the parameter dtype anyway
    tp_axis = getattr(getattr(model, 'tp_context', None), 'axis', None)
    # A model without a single trainable parameter (DPO's reference policy; the reference shards it with the same
    # mesh as the policy, lms/dpo/dpo.py:69-71) gets a parameter-only engine: it is worth sharding only in ZeRO-3
    # mode — resident parameters of a frozen model are just the model.
    frozen = not any(p.requires_grad for p in model.parameters())
    units = _collect_units(model, only_trainable=not frozen)
    reshard = _resolve_reshard(reshard_after_forward, units, dp_mesh, device, frozen=frozen)
    if frozen and not reshard:
        model._flat_engine = None
        return None
    engine = FlatParamEngine(units, dp_mesh, device, comm=comm, tp_axis=tp_axis, reshard=reshard, frozen=frozen)
    model._flat_engine = engine
    engine.attach_module_hooks(model)
    if engine.reshard:
        engine.enable_fused_first_gemm(model)
    if offload_policy is not None and getattr(offload_policy, 'pin_memory', None) is not None \
            and type(offload_policy).__name__ == 'CPUOffloadPolicy':
        logger.info('CPUOffloadPolicy: optimizer state (fp32 masters + Adam moments) is kept in host memory; bf16 '
                    'parameters and gradients stay on the device')
    return engine
