from .phi3_config import Phi3Config
from .phi3_model import Phi3
