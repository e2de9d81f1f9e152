"""``DeepSpeedStrategy``: the reference's ZeRO configuration surface on the native engine.

The reference wraps Lightning's DeepSpeed strategy (src/llm_training/lightning/strategy/deepspeed/
deepspeed_strategy.py:16-156); six of its eight example configs use ``stage: 2``.  DeepSpeed is not a
dependency here — the constructor accepts the same arguments and maps them onto the flat-parameter engine:

  stage 0/1/2   -> optimizer state and fp32 master weights sharded over all ranks (ZeRO-1), gradients
                   reduce-scattered per unit during backward (ZeRO-2), bf16 parameters resident and refreshed in
                   place by the optimizer kernel after the step
  stage 3       -> the engine's resharding mode (``reshard_after_forward=True``): parameters and gradients exist
                   whole only around a unit's forward / backward (parallel/fsdp.py)
  offload_optimizer -> fp32 masters + Adam moments in pinned host memory, AdamW on the CPU (optim/sharded_optimizer.py)
  offload_parameters / remote_device (CPU / NVMe parameter offload), quantised weights / gradients, cpu_checkpointing
                -> NOT implemented: a non-default value raises (LLMT_IGNORE_UNSUPPORTED=1 downgrades to a warning)
  bucket sizes / hpZ / partition_activations / aio knobs -> scheduling hints with no counterpart here (logged)
  loss scaling knobs -> fp16 GradScaler configuration
"""
import logging
from typing import Any

from ..base import Strategy

logger = logging.getLogger(__name__)


class DeepSpeedStrategy(Strategy):
    def __init__(self, accelerator: Any | None = None, zero_optimization: bool = True, stage: int = 2,
                 remote_device: str | None = None, offload_optimizer: bool = False, offload_parameters: bool = False,
                 offload_params_device: str = 'cpu', nvme_path: str = '/local_nvme', params_buffer_count: int = 5,
                 params_buffer_size: int = 100_000_000, max_in_cpu: int = 1_000_000_000,
                 offload_optimizer_device: str = 'cpu', optimizer_buffer_count: int = 4, block_size: int = 1048576,
                 queue_depth: int = 8, single_submit: bool = False, overlap_events: bool = True,
                 thread_count: int = 1, pin_memory: bool = False, sub_group_size: int = 1_000_000_000_000,
                 contiguous_gradients: bool = True, overlap_comm: bool = True, allgather_partitions: bool = True,
                 reduce_scatter: bool = True, allgather_bucket_size: int = 200_000_000,
                 reduce_bucket_size: int = 200_000_000, zero_allow_untested_optimizer: bool = True,
                 logging_batch_size_per_gpu: str | int = 'auto', config: Any | None = None,
                 logging_level: int | str = logging.WARN, parallel_devices: Any | None = None,
                 cluster_environment: Any | None = None, loss_scale: float = 0, initial_scale_power: int = 16,
                 loss_scale_window: int = 1000, hysteresis: int = 2, min_loss_scale: int = 1,
                 partition_activations: bool = False, cpu_checkpointing: bool = False,
                 contiguous_memory_optimization: bool = False, synchronize_checkpoint_boundary: bool = False,
                 load_full_weights: bool = False, precision_plugin: Any | None = None,
                 process_group_backend: str | None = None, exclude_frozen_parameters: bool = True,
                 raise_error_at_min_scale: bool | None = None, zero3_leaf_modules: list[type] | None = None,
                 stage3_max_live_parameters: int | float = 1e9, stage3_max_reuse_distance: int | float = 1e9,
                 stage3_prefetch_bucket_size: int | float = 5e8,
                 stage3_param_persistence_threshold: int | float = 1e6, zero_hpz_partition_size: int = 1,
                 zero_quantized_weights: bool = False, zero_quantized_gradients: bool = False) -> None:
        super().__init__(process_group_backend)
        if stage not in (0, 1, 2, 3):
            raise ValueError(f'ZeRO stage must be 0..3, got {stage}')
        self.stage = stage
        self.offload_optimizer = bool(offload_optimizer)
        self.reshard_after_forward = True if stage == 3 else False
        self.zero_optimization = zero_optimization
        self.exclude_frozen_parameters = exclude_frozen_parameters
        self.raise_error_at_min_scale = raise_error_at_min_scale
        self.zero3_leaf_modules = zero3_leaf_modules
        self.loss_scale_config = dict(loss_scale=loss_scale, initial_scale_power=initial_scale_power,
                                      loss_scale_window=loss_scale_window, hysteresis=hysteresis,
                                      min_loss_scale=min_loss_scale)
        self.save_distributed_checkpoint = True
        self.config = config
        if precision_plugin is not None:
            self.precision_plugin = precision_plugin
        # options that would change WHERE state lives or WHAT goes over the wire: refusing beats pretending
        unsupported = {
            'offload_parameters': offload_parameters, 'remote_device': remote_device,
            'zero_quantized_weights': zero_quantized_weights, 'zero_quantized_gradients': zero_quantized_gradients,
            'cpu_checkpointing': cpu_checkpointing,
        }
        active = [k for k, v in unsupported.items() if v]
        if active:
            import os
            msg = (f'DeepSpeedStrategy: {active} not implemented by the native engine (parameters, gradients and '
                   'optimizer state live in HBM, sharded per ZeRO stage; only `offload_optimizer` moves state to the '
                   'host)')
            if os.environ.get('LLMT_IGNORE_UNSUPPORTED', '0') != '1':
                raise NotImplementedError(msg + '; set LLMT_IGNORE_UNSUPPORTED=1 to train without them')
            logger.warning(msg + ' — ignored')
        hints = [k for k, v in {'partition_activations': partition_activations,
                                'zero_hpz_partition_size': zero_hpz_partition_size != 1}.items() if v]
        if hints:
            logger.warning('DeepSpeedStrategy: %s are scheduling hints with no counterpart here (ignored)', hints)

    @property
    def zero_stage_3(self) -> bool:
        return self.stage == 3

    @property
    def is_fp16(self) -> bool:
        """fp16 with dynamic loss scaling (reference: deepspeed_strategy.py:104-108)."""
        return self.precision_plugin.precision in ('16-mixed', '16-true')

    @property
    def skipped_steps(self) -> int:
        """Optimizer steps skipped because the fp16 gradients overflowed (shown in the progress bar upstream)."""
        return int(getattr(self.precision_plugin, 'skipped_steps', 0))

    def init_deepspeed(self) -> None:
        """DeepSpeed builds its engine here; the native engine is built in ``setup`` — kept for API parity."""

    def _mesh_sizes(self) -> tuple[int, int]:
        return self.world_size, 1

    def setup(self, trainer) -> None:
        if self.precision_plugin.precision == '16-mixed' and self.precision_plugin.scaler is not None:
            cfg = self.loss_scale_config
            if cfg['loss_scale'] == 0:
                import torch
                self.precision_plugin.scaler = torch.amp.GradScaler(
                    self.root_device.type, init_scale=2.0 ** cfg['initial_scale_power'],
                    growth_interval=cfg['loss_scale_window'])
            self.precision_plugin.growth_interval = int(cfg['loss_scale_window'])
            self.precision_plugin.min_scale = float(cfg['min_loss_scale'])
            self.precision_plugin.raise_error_at_min_scale = bool(self.raise_error_at_min_scale)
        super().setup(trainer)
