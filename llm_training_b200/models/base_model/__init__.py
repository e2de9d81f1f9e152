from .base_model import BaseModel
from .base_model_config import BaseModelConfig
