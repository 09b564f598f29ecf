from .fused_layer_norm import MixedFusedLayerNorm as LayerNorm, RMSNorm  # noqa: F401
from .distributed import DistributedDataParallel  # noqa: F401
from .gpt_model import GPTModel  # noqa: F401
from .llama_model import LlamaModel  # noqa: F401
from .falcon_model import FalconModel  # noqa: F401
from .language_model import get_language_model  # noqa: F401
from .module import Float16Module  # noqa: F401
from .enums import ModelType  # noqa: F401
from .bert_model import BertModel  # noqa: F401,E402
from .t5_model import T5Model  # noqa: F401,E402
from .classification import Classification  # noqa: F401,E402
from .multiple_choice import MultipleChoice  # noqa: F401,E402
