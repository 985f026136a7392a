from .ppo import PPObj, layer_init  # noqa: F401
from .common import Predictor  # noqa: F401
