"""Importable alias for the framework package.

The source tree lives in ``deepfm-tensorflow-distributed-training-on-sagemaker_b200/``
(the directory name the build spec asks for).  A hyphenated directory cannot be
imported, so this shim points ``deepfm_b200.__path__`` at it and runs its
``__init__``.  ``import deepfm_b200.models`` etc. then resolve inside that tree,
and ``python -m deepfm_b200 --task_type train ...`` runs its ``__main__``.
"""
import os as _os

_ROOT = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
_SRC = _os.path.join(_ROOT, "deepfm-tensorflow-distributed-training-on-sagemaker_b200")
__path__ = [_SRC]
REPO_ROOT = _ROOT
SRC_DIR = _SRC

with open(_os.path.join(_SRC, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_SRC, "__init__.py"), "exec"))
