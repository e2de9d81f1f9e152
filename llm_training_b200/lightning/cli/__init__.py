from .cli import LightningCLI
from .config import instantiate, load_config, parse_overrides
from .utils import HFTokenizer

__all__ = ['LightningCLI', 'HFTokenizer', 'instantiate', 'load_config', 'parse_overrides']
