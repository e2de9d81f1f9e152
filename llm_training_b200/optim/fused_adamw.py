"""AdamW backed by the fused sm_100a kernel (``csrc/kernels/optimizer.cu``).

Stands in for ``deepspeed.ops.adam.FusedAdam`` (reference config: config/examples/llama-3.1/
llama-3.1-8b_pt_example.yaml:44 — the class path is aliased in ``utils.imports.CLASS_ALIASES``) and for
``torch.optim.AdamW`` + ``MasterWeightsOptimizer``.  Used stand-alone it keeps an fp32 master copy per
low-precision parameter; under a strategy the trainer swaps it for :class:`ShardedOptimizer`, which runs the
same kernel over flat per-unit shards.
"""
import torch
from torch.optim import Optimizer

from ..ops.adamw_op import adamw_step


class FusedAdamW(Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas: tuple[float, float] = (0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, bias_correction: bool = True, adam_w_mode: bool = True,
                 amsgrad: bool = False, set_grad_none: bool = True, **_) -> None:
        if amsgrad:
            raise ValueError('FusedAdamW does not support amsgrad')
        if not adam_w_mode:
            raise ValueError('only decoupled weight decay (AdamW) is implemented')
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            beta1, beta2 = group['betas']
            for p in group['params']:
                if p.grad is None:
                    continue
                state = self.state[p]
                if not state:
                    state['step'] = 0
                    state['exp_avg'] = torch.zeros_like(p, dtype=torch.float32, memory_format=torch.contiguous_format)
                    state['exp_avg_sq'] = torch.zeros_like(p, dtype=torch.float32, memory_format=torch.contiguous_format)
                    if p.dtype != torch.float32:
                        state['master'] = p.detach().float().contiguous()
                state['step'] += 1
                low_precision = p.dtype != torch.float32
                master = state['master'] if low_precision else p.data
                grad = p.grad if p.grad.dtype in (torch.float32, torch.bfloat16) else p.grad.float()
                out = p.data if (low_precision and p.dtype == torch.bfloat16 and p.is_contiguous()) else None
                adamw_step(master.view(-1), state['exp_avg'].view(-1), state['exp_avg_sq'].view(-1),
                           grad.contiguous().view(-1), None if out is None else out.view(-1), lr=group['lr'],
                           beta1=beta1, beta2=beta2, eps=group['eps'], weight_decay=group['weight_decay'],
                           step=state['step'])
                if low_precision and out is None:
                    p.data.copy_(master)
        return loss
