"""One LLaMA-3-8B-shaped training step (4 layers, full vocabulary, 8192 tokens) for ncu.

    python bench/profile_step.py                 # plain run, prints per-step time
    ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
        --log-file gpurun_out/launches.csv python bench/profile_step.py
Only the step between cudaProfilerStart/Stop is captured (after 2 warm-up steps)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_training_b200.lightning import FSDP2Strategy, Trainer  # noqa: E402
from llm_training_b200.lms import CLM, CLMConfig  # noqa: E402

LAYERS = int(os.environ.get('LAYERS', '4'))
cfg = dict(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=LAYERS,
           num_attention_heads=32, num_key_value_heads=8, max_position_embeddings=8192, rope_theta=500000.0)
lm = CLM(CLMConfig.model_validate(dict(init_weights=True, load_weights=False,
                                       model=dict(model_class='llm_training.models.Llama', model_config=cfg),
                                       optim=dict(optimizer_class='torch.optim.AdamW',
                                                  optimizer_kwargs=dict(lr=1e-5)))))
trainer = Trainer(strategy=FSDP2Strategy(data_parallel_size=1), precision='bf16-true', logger=False,
                  enable_progress_bar=False, enable_checkpointing=False, gradient_clip_val=1.0)
lm.trainer = trainer
trainer.lightning_module = lm
trainer.strategy.setup_environment(trainer)
with trainer.strategy.model_init_context():
    lm.configure_model()
trainer.strategy.setup(trainer)
dev = trainer.strategy.root_device
ids = torch.randint(0, 128256, (1, 8192), device=dev)
batch = dict(input_ids=ids, attention_mask=torch.ones_like(ids), position_ids=torch.arange(8192, device=dev)[None],
             labels=ids)
opt = trainer.optimizers[0]


def step():
    loss = trainer.strategy.training_step(batch, 0)
    trainer.strategy.backward(loss)
    trainer.strategy.optimizer_step(opt)
    opt.zero_grad()


for _ in range(2):
    step()
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.cudart().cudaProfilerStart()
s.record()
step()
trainer.strategy.synchronize_optimizer()
e.record()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print(f'step_ms {s.elapsed_time(e):.3f} layers {LAYERS}')
