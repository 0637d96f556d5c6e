"""convnets_b200 — B200-native (sm_100a) convolutional training hot path behind the ConvNets
architecture zoo.  See DESIGN.md."""
__version__ = "0.1.0"

from .convert import convert  # noqa: E402,F401
from .graph import EvalGraph, StepGraph, enable_eval_graph, enable_step_graph  # noqa: E402,F401
