from .clm import CLM
from .clm_config import CLMConfig
