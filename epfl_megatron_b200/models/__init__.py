from .fused_layer_norm import MixedFusedLayerNorm as LayerNorm, RMSNorm  # noqa: F401
from .distributed import DistributedDataParallel  # noqa: F401
from .gpt_model import GPTModel  # noqa: F401
from .llama_model import LlamaModel  # noqa: F401
from .falcon_model import FalconModel  # noqa: F401
from .language_model import get_language_model  # noqa: F401
from .module import Float16Module  # noqa: F401
from .enums import ModelType  # noqa: F401
