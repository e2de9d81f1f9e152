"""Rotary-embedding frequency tables: default, linear, dynamic-NTK, YaRN, LongRoPE, llama3.

Same public names as the reference (``RoPEConfig``, ``ROPE_INIT_FUNCTIONS``;
src/llm_training/ops/rope_utils.py:12-31, :289-296) and the same formulas (:33-283); organised
as a decorator registry with shared helpers.  Every initialiser returns
``(inv_freq fp32 [dim/2], attention_factor)``.
"""
import logging
import math
from typing import Any, Callable

import torch
from pydantic import BaseModel, model_validator

logger = logging.getLogger(__name__)

RopeInit = Callable[["RoPEConfig", torch.device | None, int | None], tuple[torch.Tensor, float]]

ROPE_INIT_FUNCTIONS: dict[str, RopeInit] = {}
ROPE_REQUIRED_KEYS: dict[str, tuple[set[str], set[str]]] = {}


def register_rope(name: str, required: set[str] = frozenset(), optional: set[str] = frozenset()):
    def deco(fn: RopeInit) -> RopeInit:
        ROPE_INIT_FUNCTIONS[name] = fn
        ROPE_REQUIRED_KEYS[name] = (set(required), set(optional))
        return fn
    return deco


class RoPEConfig(BaseModel):
    type: str
    base: float
    dim: int
    max_position_embeddings: int
    scaling_config: dict[str, Any] | None = None

    @model_validator(mode='after')
    def _check_keys(self) -> "RoPEConfig":
        if self.type not in ROPE_REQUIRED_KEYS:
            logger.warning(f"no key validation registered for rope type '{self.type}'")
            return self
        required, optional = ROPE_REQUIRED_KEYS[self.type]
        received = set((self.scaling_config or {}).keys()) - {'type', 'rope_type'}
        missing = required - received
        if missing:
            raise KeyError(f"rope_scaling for '{self.type}' is missing keys: {sorted(missing)}")
        unused = received - required - optional
        if unused:
            logger.warning(f"rope_scaling for '{self.type}' has unrecognised keys: {sorted(unused)}")
        sc = self.scaling_config or {}
        if 'factor' in required:
            f = sc['factor']
            if not isinstance(f, (int, float)) or f < 1.0:
                logger.warning(f"rope_scaling `factor` should be a number >= 1, got {f}")
        if self.type == 'longrope':
            for key in ('short_factor', 'long_factor'):
                val = sc[key]
                if not (isinstance(val, (list, tuple)) and all(isinstance(x, (int, float)) for x in val)):
                    logger.warning(f"rope_scaling `{key}` must be a list of numbers")
                elif len(val) != self.dim // 2:
                    logger.warning(f"rope_scaling `{key}` must have length {self.dim // 2}, got {len(val)}")
        if self.type == 'llama3':
            lo, hi = sc['low_freq_factor'], sc['high_freq_factor']
            if hi <= lo:
                logger.warning('rope_scaling `high_freq_factor` must exceed `low_freq_factor`')
            if sc['original_max_position_embeddings'] >= self.max_position_embeddings:
                logger.warning('`original_max_position_embeddings` should be smaller than max_position_embeddings')
        return self


def _exponents(dim: int, device) -> torch.Tensor:
    return torch.arange(0, dim, 2, dtype=torch.int64, device=device).float() / dim


def _base_inv_freq(base: float, dim: int, device) -> torch.Tensor:
    return 1.0 / (base ** _exponents(dim, device))


@register_rope('default')
def _default(config: RoPEConfig, device=None, seq_len=None):
    return _base_inv_freq(config.base, config.dim, device), 1.0


@register_rope('linear', required={'factor'})
def _linear(config: RoPEConfig, device=None, seq_len=None):
    # scaling positions by 1/factor == scaling frequencies by 1/factor
    return _base_inv_freq(config.base, config.dim, device) / config.scaling_config['factor'], 1.0


@register_rope('dynamic', required={'factor'}, optional={'original_max_position_embeddings'})
def _dynamic_ntk(config: RoPEConfig, device=None, seq_len=None):
    dim, mpe = config.dim, config.max_position_embeddings
    factor = config.scaling_config['factor']
    seq_len = seq_len if (seq_len is not None and seq_len > mpe) else mpe
    base = config.base * ((factor * seq_len / mpe) - (factor - 1)) ** (dim / (dim - 2))
    return _base_inv_freq(base, dim, device), 1.0


@register_rope('yarn', required={'factor'},
               optional={'attention_factor', 'beta_fast', 'beta_slow', 'original_max_position_embeddings'})
def _yarn(config: RoPEConfig, device=None, seq_len=None):
    sc = config.scaling_config
    base, dim, mpe, factor = config.base, config.dim, config.max_position_embeddings, sc['factor']
    attention_factor = sc.get('attention_factor')
    if attention_factor is None:
        attention_factor = 0.1 * math.log(factor) + 1.0
    beta_fast = sc.get('beta_fast') or 32
    beta_slow = sc.get('beta_slow') or 1

    def dim_for_rotations(n_rot: float) -> float:
        return dim * math.log(mpe / (n_rot * 2 * math.pi)) / (2 * math.log(base))

    low = max(math.floor(dim_for_rotations(beta_fast)), 0)
    high = min(math.ceil(dim_for_rotations(beta_slow)), dim - 1)
    if low == high:
        high += 0.001
    ramp = ((torch.arange(dim // 2, dtype=torch.float32, device=device) - low) / (high - low)).clamp(0, 1)
    extrapolation_weight = 1.0 - ramp
    extrapolated = _base_inv_freq(base, dim, device)
    interpolated = extrapolated / factor
    inv_freq = interpolated * (1 - extrapolation_weight) + extrapolated * extrapolation_weight
    return inv_freq, float(attention_factor)


@register_rope('longrope', required={'short_factor', 'long_factor', 'factor'},
               optional={'attention_factor', 'original_max_position_embeddings'})
def _longrope(config: RoPEConfig, device=None, seq_len=None):
    sc = config.scaling_config
    mpe, factor = config.max_position_embeddings, sc['factor']
    seq_len = seq_len or mpe * factor
    attention_factor = sc.get('attention_factor')
    if attention_factor is None:
        attention_factor = 1.0 if factor <= 1.0 else math.sqrt(1 + math.log(factor) / math.log(mpe))
    chosen = sc['long_factor'] if seq_len > mpe else sc['short_factor']
    ext = torch.tensor(chosen, dtype=torch.float32, device=device)
    inv_freq = 1.0 / (ext * config.base ** _exponents(config.dim, device))
    return inv_freq, float(attention_factor)


@register_rope('llama3', required={'factor', 'original_max_position_embeddings', 'low_freq_factor',
                                   'high_freq_factor'})
def _llama3(config: RoPEConfig, device=None, seq_len=None):
    sc = config.scaling_config
    factor, lo, hi = sc['factor'], sc['low_freq_factor'], sc['high_freq_factor']
    old_len = sc['original_max_position_embeddings']
    inv_freq = _base_inv_freq(config.base, config.dim, device)
    wavelen = 2 * math.pi / inv_freq
    scaled = torch.where(wavelen > old_len / lo, inv_freq / factor, inv_freq)
    smooth = (old_len / wavelen - lo) / (hi - lo)
    blended = (1 - smooth) * scaled / factor + smooth * scaled
    medium = (wavelen >= old_len / hi) & (wavelen <= old_len / lo)
    return torch.where(medium, blended, scaled), 1.0
