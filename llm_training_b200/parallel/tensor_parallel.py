"""Megatron-style tensor parallelism with sequence parallelism for the decoder models.

What the reference builds from DTensor ``parallelize_module`` plans (src/llm_training/models/llama/
llama_model.py:197-244, phi3_model.py:212-256) is done here on plain tensors:

  * q/k/v and gate/up are column-parallel (rows of the weight split by heads / features), o_proj and down_proj
    are row-parallel (input features split); local head counts are ``H / tp``
  * the residual stream and both RMSNorms live on ``[T/tp, d]`` token shards (sequence parallel):
    ``gather_sequence`` all-gathers before attention / MLP, the row-parallel projections reduce-scatter back
  * embedding and LM head are vocabulary-parallel; the loss never materialises full logits
    (``vocab_parallel_cross_entropy`` / ``vocab_parallel_logprob``)
  * fused weights (Phi-3 ``qkv_proj`` / ``gate_up_proj``) are sharded HEAD-ALIGNED: rank r holds
    ``[q_r; k_r; v_r]`` / ``[gate_r; up_r]`` — the reference's plain column split is wrong for pre-trained
    weights

Collectives go through torch.distributed (gloo on CPU, NCCL on GPU) inside autograd Functions.
"""
from dataclasses import dataclass
from typing import Any, Callable

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn

from ..ops import _lib, embedding, linear
from ..ops.gemm_op import gemm
from .mesh import MeshAxis


# ---------------------------------------------------------------------------------------------------
# sharding descriptors (used by checkpoint I/O and weight loading)
# ---------------------------------------------------------------------------------------------------
@dataclass
class TPInfo:
    full_shape: torch.Size
    scatter: Callable[[torch.Tensor], torch.Tensor]        # full tensor -> this rank's shard
    gather: Callable[[torch.Tensor], torch.Tensor]         # this rank's shard -> full tensor (collective)


def _all_gather_cat(local: torch.Tensor, axis: MeshAxis, dim: int) -> torch.Tensor:
    parts = [torch.empty_like(local) for _ in range(axis.size())]
    dist.all_gather(parts, local.contiguous(), group=axis.group)
    return torch.cat(parts, dim=dim)


def _plain_shard(axis: MeshAxis, dim: int) -> tuple[Callable, Callable]:
    def scatter(full: torch.Tensor) -> torch.Tensor:
        return full.chunk(axis.size(), dim=dim)[axis.rank].contiguous()

    def gather(local: torch.Tensor) -> torch.Tensor:
        return _all_gather_cat(local, axis, dim)
    return scatter, gather


def _blocked_shard(axis: MeshAxis, block_rows: list[int]) -> tuple[Callable, Callable]:
    """Row-blocked weight (e.g. [q; k; v]): every block is split evenly, rank r keeps the r-th piece of each."""
    tp = axis.size()

    def scatter(full: torch.Tensor) -> torch.Tensor:
        pieces = [b.chunk(tp, dim=0)[axis.rank] for b in full.split(block_rows, dim=0)]
        return torch.cat(pieces, dim=0).contiguous()

    def gather(local: torch.Tensor) -> torch.Tensor:
        parts = [torch.empty_like(local) for _ in range(tp)]
        dist.all_gather(parts, local.contiguous(), group=axis.group)
        local_rows = [r // tp for r in block_rows]
        blocks = []
        for b in range(len(block_rows)):
            blocks.append(torch.cat([p.split(local_rows, dim=0)[b] for p in parts], dim=0))
        return torch.cat(blocks, dim=0)
    return scatter, gather


def _shard_param(module: nn.Module, name: str, scatter: Callable, gather: Callable) -> None:
    p = getattr(module, name)
    if p is None:
        return
    full_shape = p.shape
    if p.is_meta:
        local = torch.empty(scatter(torch.empty(full_shape, device='meta')).shape, dtype=p.dtype, device='meta')
    else:
        local = scatter(p.data)
    new = nn.Parameter(local, requires_grad=p.requires_grad)
    new._tp_info = TPInfo(full_shape, scatter, gather)
    setattr(module, name, new)


# ---------------------------------------------------------------------------------------------------
# autograd collectives
# ---------------------------------------------------------------------------------------------------
class _GatherSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, axis: MeshAxis) -> torch.Tensor:
        ctx.axis = axis
        out = torch.empty((x.shape[0] * axis.size(), *x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, x.contiguous(), group=axis.group)
        return out

    @staticmethod
    def backward(ctx, g: torch.Tensor):
        return _reduce_scatter_rows(g.contiguous(), ctx.axis), None


class _ReduceScatterSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, axis: MeshAxis) -> torch.Tensor:
        ctx.axis = axis
        return _reduce_scatter_rows(x.contiguous(), axis)

    @staticmethod
    def backward(ctx, g: torch.Tensor):
        out = torch.empty((g.shape[0] * ctx.axis.size(), *g.shape[1:]), dtype=g.dtype, device=g.device)
        dist.all_gather_into_tensor(out, g.contiguous(), group=ctx.axis.group)
        return out, None


class _ScatterSeq(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: torch.Tensor, axis: MeshAxis) -> torch.Tensor:
        ctx.axis = axis
        return x.chunk(axis.size(), dim=0)[axis.rank].contiguous()

    @staticmethod
    def backward(ctx, g: torch.Tensor):
        out = torch.empty((g.shape[0] * ctx.axis.size(), *g.shape[1:]), dtype=g.dtype, device=g.device)
        dist.all_gather_into_tensor(out, g.contiguous(), group=ctx.axis.group)
        return out, None


def _reduce_scatter_rows(x: torch.Tensor, axis: MeshAxis) -> torch.Tensor:
    tp = axis.size()
    assert x.shape[0] % tp == 0, f'token count {x.shape[0]} is not divisible by the tensor-parallel size {tp}'
    if dist.get_backend(axis.group) == 'gloo':
        full = x.clone()
        dist.all_reduce(full, group=axis.group)
        return full.chunk(tp, dim=0)[axis.rank].contiguous()
    out = torch.empty((x.shape[0] // tp, *x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.reduce_scatter_tensor(out, x, group=axis.group)
    return out


class _VocabParallelLoss(torch.autograd.Function):
    """CE / token log-prob over vocabulary shards: each rank holds ``W[v0:v1]``; max / sum-exp / target logit are
    combined across ranks (reference: ``loss_parallel()``, lms/clm/clm.py:114-120 — 3 all-reduces of [T];
    explicit all-reduce of per-token log-probs, lms/dpo/dpo.py:89-108).  Returns per-token ``lse - tgt``
    (0 at ignored positions).  The hidden-state gradient is a partial sum (the caller's sequence gather turns it
    into a reduce-scatter)."""

    @staticmethod
    def forward(ctx, hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor, ignore_index: int,
                axis: MeshAxis, tp: 'TensorParallelContext | None' = None) -> torch.Tensor:
        """``tp`` given: ``hidden`` is the LOCAL sequence shard and the logits GEMM gathers it itself (land-once
        all-gather -> GEMM); backward then reduce-scatters dH inside the dgrad GEMM."""
        v_local = weight.shape[0]
        v0 = axis.rank * v_local
        local_labels = labels - v0
        in_shard = (local_labels >= 0) & (local_labels < v_local)
        # out-of-shard labels must not look like `ignore_index`
        local_labels = torch.where(in_shard, local_labels, torch.full_like(labels, -1))
        local_labels = torch.where(labels == ignore_index, torch.full_like(labels, ignore_index), local_labels)
        use_k = _lib.use_kernels(hidden) and weight.dtype == torch.bfloat16 and v_local % 8 == 0
        assert tp is None or use_k, 'fused vocab-parallel loss needs the CUDA kernels'
        if use_k:
            if tp is not None:
                logits, hidden = _gather_gemm(tp, hidden.contiguous(), weight, b_mn_major=False)
            else:
                logits = gemm(hidden.contiguous(), weight)
            rows = logits.shape[0]
            on_nvlink = tp is not None and tp.ring is not None and rows % 4 == 0
            if on_nvlink:
                # the loss kernel writes this shard's statistics straight into symmetric memory; ONE kernel then reads
                # every rank's block over NVLink and combines them (no all_gather / logsumexp / all_reduce launches)
                stats_buf, stats = tp.ring.next_f32(2 * rows)
                lse_local, tgt_local = stats[:rows], stats[rows:]
            else:
                lse_local = torch.empty(rows, dtype=torch.float32, device=hidden.device)
                tgt_local = torch.empty_like(lse_local)
            _lib.check(_lib.lib().llmt_ce_fwd(_lib.ptr(logits), _lib.ptr(local_labels), _lib.ptr(lse_local),
                                              _lib.ptr(tgt_local), _lib.c_int(rows), _lib.c_int(v_local),
                                              _lib.c_i64(logits.stride(0)), _lib.c_i64(ignore_index),
                                              _lib.c_f32(0.0), _lib.stream()), 'ce_fwd')
            if on_nvlink:
                tp.comm.barrier()                   # every rank's statistics are in place
                lse = torch.empty(rows, dtype=torch.float32, device=hidden.device)
                nll = torch.empty_like(lse)
                _lib.check(_lib.lib().llmt_comm_vocab_loss_combine(
                    stats_buf.peer_ptrs_at(0), _lib.c_int(tp.size), _lib.c_int(rows), _lib.ptr(labels),
                    _lib.c_i64(ignore_index), _lib.ptr(lse), _lib.ptr(nll), _lib.stream()), 'comm_vocab_loss_combine')
                ctx.save_for_backward(hidden, weight, logits, local_labels, lse)
                ctx.meta = (ignore_index, use_k, v_local)
                ctx.tp = tp
                return nll
        else:
            logits = F.linear(hidden, weight).float()
            lse_local = torch.logsumexp(logits, dim=-1)
            tgt_local = logits.gather(1, local_labels.clamp(0, v_local - 1).unsqueeze(1)).squeeze(1) * in_shard
        # combine shards: global lse = logsumexp over ranks of the local lse; target logit lives on one rank
        gathered = torch.empty(axis.size() * lse_local.numel(), dtype=torch.float32, device=hidden.device)
        dist.all_gather_into_tensor(gathered, lse_local.float().contiguous(), group=axis.group)
        lse = torch.logsumexp(gathered.view(axis.size(), -1), dim=0)
        tgt = tgt_local.float() * (labels != ignore_index)
        dist.all_reduce(tgt, group=axis.group)
        ctx.save_for_backward(hidden, weight, logits, local_labels, lse)
        ctx.meta = (ignore_index, use_k, v_local)
        ctx.tp = tp
        return (lse - tgt) * (labels != ignore_index)

    @staticmethod
    def backward(ctx, grad: torch.Tensor):
        hidden, weight, logits, local_labels, lse = ctx.saved_tensors
        ignore_index, use_k, v_local = ctx.meta
        scale = grad.float().contiguous()
        if use_k:
            rows = logits.shape[0]
            _lib.check(_lib.lib().llmt_ce_bwd_inplace(_lib.ptr(logits), _lib.ptr(local_labels), _lib.ptr(lse),
                                                      _lib.ptr(scale), _lib.c_int(rows), _lib.c_int(v_local),
                                                      _lib.c_i64(logits.stride(0)), _lib.c_i64(ignore_index),
                                                      _lib.c_f32(0.0), _lib.stream()), 'ce_bwd')
            if ctx.tp is not None:      # dH partial sums -> sequence shards, reduce-scattered by the dgrad GEMM
                dh = _fused_linear_reduce_scatter(ctx.tp, logits, weight, None, b_mn_major=True)
            else:
                dh = gemm(logits, weight, b_mn_major=True)
            dw = gemm(logits, hidden.contiguous(), a_mn_major=True, b_mn_major=True)
        else:
            probs = torch.exp(logits - lse.unsqueeze(1))
            onehot = torch.zeros_like(probs)
            ok = (local_labels >= 0) & (local_labels < v_local)
            onehot[ok, local_labels[ok]] = 1.0
            dlogits = (probs - onehot) * (scale * (local_labels != ignore_index)).unsqueeze(1)
            dh = (dlogits @ weight.float()).to(hidden.dtype)
            dw = (dlogits.t() @ hidden.float()).to(weight.dtype)
        return dh, dw, None, None, None, None


# ---------------------------------------------------------------------------------------------------
# fused collective GEMMs over NVLink symmetric memory (CUDA/bf16 path)
# ---------------------------------------------------------------------------------------------------
class _SymmetricRing:
    """Two alternating symmetric buffers per shape: buffer k is rewritten only after the barrier of use k+1, i.e.
    after every peer has finished reading it (see the protocol note in ``parallel/nvlink.py``)."""

    MAX_SHAPES = 16     # distinct (rows, cols) kept alive; variable-length batches would otherwise grow the heap forever

    def __init__(self, comm: Any) -> None:
        self.comm = comm
        self.buffers: dict[tuple, list] = {}     # insertion order = least recently used first
        self.turn: dict[tuple, int] = {}

    def next(self, rows: int, cols: int):
        key = (rows, cols)
        if key not in self.buffers:
            while len(self.buffers) >= self.MAX_SHAPES:
                # every rank sees the same sequence of shapes, so every rank evicts the same entry; a barrier makes
                # sure no peer is still reading the buffers that are about to be unmapped
                oldest = next(iter(self.buffers))
                self.comm.barrier()
                torch.cuda.current_stream().synchronize()
                del self.buffers[oldest], self.turn[oldest]
            self.buffers[key] = [self.comm.alloc(rows * cols, torch.bfloat16) for _ in range(2)]
            self.turn[key] = 0
        else:
            self.buffers[key] = self.buffers.pop(key)      # mark as most recently used
        self.turn[key] ^= 1
        buf = self.buffers[key][self.turn[key]]
        return buf, buf.tensor.view(rows, cols)

    def next_f32(self, numel: int):
        """A symmetric fp32 vector of ``numel`` elements (its own ring entry: two bf16 elements per float)."""
        buf, view = self.next(2 * numel, 1)
        return buf, view.view(-1).view(torch.float32)


def _gemm_dist(mode: int, a: torch.Tensor | None, a_ptrs, b: torch.Tensor, c: torch.Tensor | None, c_ptrs, world: int,
               rows_per_rank: int, m: int, n: int, k: int, lda: int, ldc: int, b_mn_major: bool) -> None:
    import ctypes
    null = (ctypes.c_void_p * 1)(0)
    _lib.check(_lib.lib().llmt_gemm_bf16_dist(
        _lib.c_int(mode), _lib.ptr(a), a_ptrs if a_ptrs is not None else null, _lib.ptr(b), _lib.ptr(c),
        c_ptrs if c_ptrs is not None else null, _lib.c_int(world), _lib.c_int(rows_per_rank), _lib.c_int(m),
        _lib.c_int(n), _lib.c_int(k), _lib.c_i64(lda), _lib.c_i64(b.stride(0)), _lib.c_i64(ldc),
        _lib.c_int(b_mn_major), _lib.c_int(2), _lib.c_int(0), _lib.stream()), 'gemm_bf16_dist')


_AG_FLAGS: dict[torch.device, torch.Tensor] = {}


def _gather_gemm(tp: 'TensorParallelContext', x_local: torch.Tensor, weight: torch.Tensor, b_mn_major: bool
                 ) -> tuple[torch.Tensor, torch.Tensor]:
    """``(all_gather(x_local) @ op(W), all_gather(x_local))`` in ONE kernel (csrc/kernels/gemm.cu, kDistA == 2):
    pull warps land every remote block in ``x_full`` once while the tensor cores work on the blocks already there."""
    rows, k = x_local.shape
    n = weight.shape[1] if b_mn_major else weight.shape[0]
    m = rows * tp.size
    cta_group = 2 if rows % 256 == 0 else 1
    buf, view = tp.ring.next(rows, k)
    view.copy_(x_local)
    tp.comm.barrier()                   # every rank's block is in place
    dev = x_local.device
    flags = _AG_FLAGS.get(dev)
    if flags is None or flags.numel() < m // 128:
        flags = _AG_FLAGS[dev] = torch.zeros(max(4096, m // 128), dtype=torch.int32, device=dev)
    x_full = torch.empty((m, k), dtype=x_local.dtype, device=dev)
    out = torch.empty((m, n), dtype=x_local.dtype, device=dev)
    _lib.check(_lib.lib().llmt_gemm_bf16_ag(
        buf.peer_ptrs_at(0), _lib.ptr(x_full), _lib.ptr(weight), _lib.ptr(out), _lib.ptr(flags),
        _lib.c_int(tp.rank), _lib.c_int(tp.size), _lib.c_int(rows), _lib.c_int(m), _lib.c_int(n), _lib.c_int(k),
        _lib.c_i64(k), _lib.c_i64(k), _lib.c_i64(weight.stride(0)), _lib.c_i64(n), _lib.c_int(b_mn_major),
        _lib.c_int(cta_group), _lib.c_int(0), _lib.stream()), 'gemm_bf16_ag')
    return out, x_full


class _FusedGatherLinear(torch.autograd.Function):
    """``all_gather(x_local) @ [W_1; W_2; ...]^T`` as ONE kernel (land-once gather inside the GEMM, see
    ``_gather_gemm``).  The gathered input the kernel leaves behind is what dW needs in backward; dx is ``dy @ W``
    fused with the reduce-scatter."""

    @staticmethod
    def forward(ctx, x_local: torch.Tensor, tp: 'TensorParallelContext', *weights: torch.Tensor) -> torch.Tensor:
        from ..ops.linear_op import _adjacent_rows_view
        weight = weights[0] if len(weights) == 1 else _adjacent_rows_view(weights)
        if weight is None:
            weight = torch.cat([w.detach() for w in weights], dim=0)
        x_local = x_local.contiguous()
        out, x_full = _gather_gemm(tp, x_local, weight, b_mn_major=False)
        # memory-tight runs (activation checkpointing on) keep the [T/tp, d] shard and gather it again for dW in
        # backward, like Megatron's sequence parallelism; otherwise the gathered input the kernel left behind is kept
        ctx.regather = tp.regather_in_backward
        ctx.save_for_backward(x_local if ctx.regather else x_full, weight)
        ctx.tp = tp
        ctx.splits = [w.shape[0] for w in weights]
        ctx.weights = weights
        return out

    @staticmethod
    def backward(ctx, dy: torch.Tensor):
        x_full, weight = ctx.saved_tensors
        tp = ctx.tp
        if ctx.regather:
            shard = x_full
            x_full = torch.empty((shard.shape[0] * tp.size, shard.shape[1]), dtype=shard.dtype, device=shard.device)
            dist.all_gather_into_tensor(x_full, shard, group=tp.axis.group)
        dy = dy.contiguous()
        from ..ops.linear_op import wgrad
        dws = wgrad(dy, x_full, ctx.weights)       # straight into the flat gradient buffer when the engine allows it
        dx = _fused_linear_reduce_scatter(tp, dy, weight, None, b_mn_major=True)   # dy @ W, reduce-scattered
        return (dx, None, *dws)


def _fused_linear_reduce_scatter(tp: 'TensorParallelContext', x: torch.Tensor, weight: torch.Tensor,
                                 residual: torch.Tensor | None, b_mn_major: bool) -> torch.Tensor:
    m, k = x.shape
    n = weight.shape[1] if b_mn_major else weight.shape[0]
    rows = m // tp.size
    buf, view = tp.ring.next(rows, n)
    if residual is not None:
        view.copy_(residual)            # the residual add rides in the reduction target
    else:
        view.zero_()
    tp.comm.barrier()                   # every target is initialised
    _gemm_dist(2, x, None, weight, None, buf.peer_ptrs_at(0), tp.size, rows, m, n, k, k, n, b_mn_major)
    tp.comm.barrier()                   # every contribution has landed
    return view.clone()


class _FusedLinearReduceScatter(torch.autograd.Function):
    """``reduce_scatter(x @ W_local^T) (+ residual)`` as ONE kernel: each output tile is TMA-reduce-added into its
    owner rank's buffer.  Backward: dx = all_gather(dy) @ W is the land-once gather GEMM, whose gathered dy then
    feeds dW."""

    @staticmethod
    def forward(ctx, x: torch.Tensor, weight: torch.Tensor, residual: torch.Tensor | None,
                tp: 'TensorParallelContext') -> torch.Tensor:
        ctx.save_for_backward(x, weight)
        ctx.tp = tp
        ctx.has_residual = residual is not None
        ctx.weights = (weight,)
        return _fused_linear_reduce_scatter(tp, x.contiguous(), weight, residual, b_mn_major=False)

    @staticmethod
    def backward(ctx, dy_local: torch.Tensor):
        x, weight = ctx.saved_tensors
        tp = ctx.tp
        dx, dy_full = _gather_gemm(tp, dy_local.contiguous(), weight, b_mn_major=True)   # dy @ W, gathered dy kept
        from ..ops.linear_op import wgrad
        dw = wgrad(dy_full, x.contiguous(), ctx.weights)[0]
        return dx, dw, (dy_local if ctx.has_residual else None), None


class _VocabParallelEmbedding(torch.autograd.Function):
    """Vocabulary-sharded embedding -> sequence-sharded activations without a reduction: the rank that owns a token's
    id stores its row straight into the buffer of the rank that owns the token's sequence shard
    (csrc/kernels/comm.cu, vocab_embedding_scatter_kernel).  Backward gathers the sequence shards of the gradient over
    NVLink (pull) and sums each local id's rows in sorted order (no atomics)."""

    @staticmethod
    def forward(ctx, flat_ids: torch.Tensor, weight: torch.Tensor, tp: 'TensorParallelContext', pad_local: int
                ) -> torch.Tensor:
        tokens, (v_local, dim) = flat_ids.numel(), weight.shape
        rows = tokens // tp.size
        buf, view = tp.ring.next(rows, dim)
        tp.comm.barrier()                   # nobody still reads this ring slot's previous contents
        _lib.check(_lib.lib().llmt_comm_vocab_embedding_scatter(
            _lib.ptr(flat_ids), _lib.ptr(weight), buf.peer_ptrs_at(0), _lib.c_int(tp.size), _lib.c_int(tokens),
            _lib.c_int(dim), _lib.c_i64(tp.rank * v_local), _lib.c_int(v_local), _lib.c_int(rows),
            _lib.c_i64(pad_local), _lib.stream()), 'comm_vocab_embedding_scatter')
        tp.comm.barrier()                   # every owner's rows have landed
        ctx.save_for_backward(flat_ids)
        ctx.tp, ctx.meta = tp, (v_local, dim, pad_local, weight.dtype)
        return view.clone()

    @staticmethod
    def backward(ctx, g: torch.Tensor):
        (flat_ids,) = ctx.saved_tensors
        tp = ctx.tp
        v_local, dim, pad_local, wdtype = ctx.meta
        tokens = flat_ids.numel()
        rows = tokens // tp.size
        buf, full = tp.ring.next(tokens, dim)
        full[tp.rank * rows:(tp.rank + 1) * rows].copy_(g.reshape(rows, dim))
        tp.comm.allgather_pull(buf, 0, rows * dim * 2)          # barrier + pull of every peer's shard
        local = flat_ids - tp.rank * v_local                     # ids of other shards fall outside [0, v_local)
        sorted_ids, perm = torch.sort(local, stable=True)
        dw = torch.empty((v_local, dim), dtype=wdtype, device=g.device)
        _lib.check(_lib.lib().llmt_embedding_bwd_sorted(_lib.ptr(sorted_ids), _lib.ptr(perm), _lib.ptr(full),
                                                        _lib.ptr(dw), _lib.c_int(tokens), _lib.c_int(dim),
                                                        _lib.c_int(v_local), _lib.c_i64(pad_local), _lib.stream()),
                   'embedding_bwd')
        return None, dw, None, None


# ---------------------------------------------------------------------------------------------------
# the context object handed to the models
# ---------------------------------------------------------------------------------------------------
class TensorParallelContext:
    def __init__(self, axis: MeshAxis) -> None:
        self.axis = axis
        self.size = axis.size()
        self.rank = axis.rank
        self.comm = None
        self.ring = None
        self.regather_in_backward = False     # set by apply_decoder_tensor_parallel (LLMT_TP_REGATHER / checkpointing)
        if torch.cuda.is_available() and dist.is_initialized() and dist.get_backend(axis.group) == 'nccl':
            from .comm import get_comm
            comm = get_comm(axis)
            if getattr(comm, 'is_nvlink', False):
                self.comm = comm
                self.ring = _SymmetricRing(comm)

    def _fused_ok(self, x: torch.Tensor, weight: torch.Tensor, rows_per_rank: int) -> bool:
        return (self.comm is not None and _lib.use_kernels(x) and weight.dtype == torch.bfloat16
                and rows_per_rank % 128 == 0 and self.size <= 8)

    def column_linear(self, x_local: torch.Tensor, weights: list[torch.Tensor],
                      biases: list[torch.Tensor | None] | None = None) -> torch.Tensor:
        """Sequence-sharded input ``[T/tp, d]`` -> ``[T, sum(out_i)/tp]``: all-gather + column-parallel projection(s),
        one fused kernel on the NVLink path (reference: PrepareModuleInput + ColwiseParallel,
        llama_model.py:223-229)."""
        from ..ops.linear_op import multi_linear
        no_bias = biases is None or all(b is None for b in biases)
        if no_bias and x_local.dim() == 2 and all(w.dtype == torch.bfloat16 for w in weights) \
                and self._fused_ok(x_local, weights[0], x_local.shape[0]):
            return _FusedGatherLinear.apply(x_local, self, *weights)
        return multi_linear(self.gather_sequence(x_local), weights, biases)

    def gather_sequence(self, x: torch.Tensor) -> torch.Tensor:
        return _GatherSeq.apply(x, self.axis)

    def scatter_sequence(self, x: torch.Tensor) -> torch.Tensor:
        return _ScatterSeq.apply(x, self.axis)

    def row_linear(self, x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor | None,
                   residual: torch.Tensor | None) -> torch.Tensor:
        """``x [T, in/tp] @ W_local^T`` -> partial sums -> reduce-scatter over tokens -> ``[T/tp, out]``."""
        if bias is None and x.dim() == 2 and x.shape[0] % self.size == 0 \
                and self._fused_ok(x, weight, x.shape[0] // self.size):
            return _FusedLinearReduceScatter.apply(x, weight, residual, self)
        partial = linear(x, weight)
        out = _ReduceScatterSeq.apply(partial, self.axis)
        if bias is not None:
            out = out + bias
        return out if residual is None else out + residual

    def vocab_parallel_embedding(self, input_ids: torch.Tensor, weight: torch.Tensor, padding_idx: int | None,
                                 noise_scale: torch.Tensor | None, seed: int) -> torch.Tensor:
        """Masked local lookup + reduce-scatter to token shards (reference: RowwiseParallel embedding,
        llama_model.py:205-208).  NEFTune noise is added AFTER the reduction, per token shard, with the scale
        computed from the global sequence length (fixes SURVEY Appendix B5)."""
        v_local = weight.shape[0]
        flat = input_ids.reshape(-1)
        rows_per_rank = flat.numel() // self.size
        if (self.ring is not None and _lib.use_kernels(weight) and weight.dtype == torch.bfloat16
                and weight.shape[1] % 8 == 0 and flat.numel() % self.size == 0 and weight.is_contiguous()):
            pad_local = -1 if padding_idx is None else padding_idx - self.rank * v_local
            out = _VocabParallelEmbedding.apply(flat.contiguous(), weight, self, pad_local)
            return self._add_neftune_noise(out, noise_scale, seed)
        local = flat - self.rank * v_local
        ok = (local >= 0) & (local < v_local)
        local_pad = None
        if padding_idx is not None and 0 <= padding_idx - self.rank * v_local < v_local:
            local_pad = padding_idx - self.rank * v_local      # the padding row lives in this shard
        rows = embedding(local.clamp(0, v_local - 1), weight, local_pad)
        rows = rows * ok.unsqueeze(1).to(rows.dtype)
        out = _ReduceScatterSeq.apply(rows, self.axis)
        return self._add_neftune_noise(out, noise_scale, seed)

    def _add_neftune_noise(self, out: torch.Tensor, noise_scale: torch.Tensor | None, seed: int) -> torch.Tensor:
        if noise_scale is not None:
            shard = noise_scale.chunk(self.size)[self.rank]
            gen = torch.Generator(device=out.device).manual_seed(seed * 1009 + self.rank)
            noise = torch.rand(out.shape, generator=gen, device=out.device, dtype=torch.float32) * 2 - 1
            out = out + (noise * shard.unsqueeze(1)).to(out.dtype)
        return out

    def lm_head(self, normed_local: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
        """Full logits ``[T, V]`` (API parity path; the training losses use the vocab-parallel functions)."""
        h = self.gather_sequence(normed_local)
        local = linear(h, weight)
        parts = [torch.empty_like(local) for _ in range(self.size)]
        dist.all_gather(parts, local.contiguous(), group=self.axis.group)
        parts[self.rank] = local
        return torch.cat(parts, dim=-1)

    def _token_nll(self, normed_local: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor, ignore_index: int
                   ) -> torch.Tensor:
        if normed_local.dim() == 2 and weight.shape[0] % 8 == 0 and weight.dtype == torch.bfloat16 \
                and self._fused_ok(normed_local, weight, normed_local.shape[0]):
            return _VocabParallelLoss.apply(normed_local, weight, labels.contiguous(), ignore_index, self.axis, self)
        h = self.gather_sequence(normed_local)
        return _VocabParallelLoss.apply(h, weight, labels.contiguous(), ignore_index, self.axis)

    def vocab_parallel_cross_entropy(self, normed_local: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                                     ignore_index: int = -100) -> torch.Tensor:
        nll = self._token_nll(normed_local, weight, labels, ignore_index)
        return nll.sum() / labels.ne(ignore_index).sum().clamp_min(1)

    def vocab_parallel_logprob(self, normed_local: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                               ignore_index: int = -100) -> torch.Tensor:
        return -self._token_nll(normed_local, weight, labels, ignore_index)


# ---------------------------------------------------------------------------------------------------
# applying the plan
# ---------------------------------------------------------------------------------------------------
def apply_decoder_tensor_parallel(model: nn.Module, tp_axis: MeshAxis, fused_qkv: bool, fused_gate_up: bool) -> None:
    tp = tp_axis.size()
    cfg = model.config
    n_heads = cfg.num_attention_heads
    n_kv = getattr(cfg, 'num_key_value_heads', None) or n_heads
    if n_heads % tp or n_kv % tp:
        raise ValueError(f'tensor_parallel_size={tp} must divide num_attention_heads={n_heads} and '
                         f'num_key_value_heads={n_kv} (the reference silently computes 0 heads, SURVEY §5.7)')
    if cfg.vocab_size % tp or cfg.intermediate_size % tp:
        raise ValueError(f'tensor_parallel_size={tp} must divide vocab_size and intermediate_size')
    head_dim = getattr(cfg, 'head_dim', None) or cfg.hidden_size // n_heads
    col = _plain_shard(tp_axis, 0)
    row = _plain_shard(tp_axis, 1)
    for layer in model.layers:
        attn, mlp = layer.self_attn, layer.mlp
        if fused_qkv:
            _shard_param(attn.qkv_proj, 'weight', *_blocked_shard(tp_axis, [n_heads * head_dim, n_kv * head_dim,
                                                                            n_kv * head_dim]))
        else:
            for proj in (attn.q_proj, attn.k_proj, attn.v_proj):
                _shard_param(proj, 'weight', *col)
                if proj.bias is not None:
                    _shard_param(proj, 'bias', *col)
        _shard_param(attn.o_proj, 'weight', *row)
        if fused_gate_up:
            _shard_param(mlp.gate_up_proj, 'weight',
                         *_blocked_shard(tp_axis, [cfg.intermediate_size, cfg.intermediate_size]))
        else:
            for proj in (mlp.gate_proj, mlp.up_proj):
                _shard_param(proj, 'weight', *col)
                if proj.bias is not None:
                    _shard_param(proj, 'bias', *col)
        _shard_param(mlp.down_proj, 'weight', *row)
        for norm in (layer.input_layernorm, layer.post_attention_layernorm):
            norm.weight._tp_replicated = True
        for proj in (attn.o_proj, mlp.down_proj):
            if getattr(proj, 'bias', None) is not None:
                proj.bias._tp_replicated = True
    tied = model.lm_head.weight is model.embed_tokens.weight
    _shard_param(model.embed_tokens, 'weight', *col)
    if tied:
        model.lm_head.weight = model.embed_tokens.weight
    else:
        _shard_param(model.lm_head, 'weight', *col)
    model.norm.weight._tp_replicated = True
    model.tp_context = TensorParallelContext(tp_axis)
    import os
    regather = os.environ.get('LLMT_TP_REGATHER', 'auto')
    model.tp_context.regather_in_backward = (bool(getattr(cfg, 'enable_gradient_checkpointing', False))
                                             if regather == 'auto' else regather == '1')
