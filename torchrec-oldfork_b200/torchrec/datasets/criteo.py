"""Criteo column names used to shape synthetic data (reference torchrec/datasets/criteo.py:29-43)."""
INT_FEATURE_COUNT = 13
CAT_FEATURE_COUNT = 26
DEFAULT_LABEL_NAME = "label"
DEFAULT_INT_NAMES = [f"int_{i}" for i in range(INT_FEATURE_COUNT)]
DEFAULT_CAT_NAMES = [f"cat_{i}" for i in range(CAT_FEATURE_COUNT)]
