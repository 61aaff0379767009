"""Model zoo: the reference's VGG and DeepNN, the toy regressor, and the BASELINE.json MLP / GPT-2 configs."""
from .vgg import VGG  # noqa: F401
from .deepnn import DeepNN  # noqa: F401
from .toy import ToyLinear  # noqa: F401
from .mlp import MLP  # noqa: F401
from .gpt2 import GPT2, GPT2Config  # noqa: F401
