from .tf1 import TF1Optimizer, OPT_IDS, canonical_name, SLOT_SUFFIXES, SLOT_INIT  # noqa: F401
