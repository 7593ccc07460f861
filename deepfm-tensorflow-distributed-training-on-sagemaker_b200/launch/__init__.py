from .local import LocalEstimator, hyperparameters_to_argv  # noqa: F401
