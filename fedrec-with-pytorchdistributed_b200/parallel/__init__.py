from .comm import init_distributed, shutdown_distributed, DistInfo  # noqa: F401
from .aggregate import GradAggregator, ParamAggregator  # noqa: F401
