"""2-rank ZeRO-3 debug run: resharding engine + direct wgrad accumulation + overlapped optimizer (torchrun)."""
import os
import sys
import traceback

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank = int(os.environ['RANK']); world = int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', device_id=torch.device('cuda', rank))
    from llm_training_b200.models import Llama, LlamaConfig
    from llm_training_b200.ops import fused_linear_cross_entropy, shift_labels
    from llm_training_b200.optim import ShardedOptimizer
    from llm_training_b200.parallel import apply_fully_sharded, init_mesh
    mesh = init_mesh(world, 1)
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=3,
                      num_attention_heads=4, num_key_value_heads=2)
    with torch.device('cuda'):
        model = Llama(cfg).to(torch.bfloat16)
    engine = apply_fully_sharded(model, mesh.dp, device=torch.device('cuda', rank), reshard_after_forward=True)
    engine.enable_direct_grad_accumulation(os.environ.get('DIRECT', '1') == '1')
    engine.overlap_optimizer = os.environ.get('OVERLAP', '1') == '1'
    opt = ShardedOptimizer(torch.optim.AdamW(model.parameters(), lr=1e-3), [engine], max_grad_norm=1.0)
    for step in range(4):
        ids = torch.randint(0, 1024, (2, 256), generator=torch.Generator().manual_seed(step * world + rank)).cuda()
        out = model(input_ids=ids, skip_logits=True)
        loss = fused_linear_cross_entropy(out.normed_hidden_states, model.lm_head.weight, shift_labels(ids).flatten())
        try:
            loss.backward()
        except Exception:
            traceback.print_exc()
            print('STATE', [(u.name, u.hook_seen, u.in_backward, u.is_sharded, u.grad_live, u.pending_grads, u.reduced)
                            for u in engine.units], flush=True)
            raise
        opt.step()
        opt.zero_grad()
        if rank == 0:
            print('step', step, float(loss), flush=True)
    opt.synchronize()
    torch.cuda.synchronize()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
