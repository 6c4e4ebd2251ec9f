from .net import Net  # noqa: F401
from .mlp import WideMLP  # noqa: F401
