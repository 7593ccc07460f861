from .export import export_servable, latest_export, signature  # noqa: F401
from .predictor import Predictor  # noqa: F401
from .server import BatchingScorer, ModelServer  # noqa: F401
