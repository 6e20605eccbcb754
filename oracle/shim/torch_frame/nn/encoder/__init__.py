from .stypewise_encoder import StypeWiseFeatureEncoder  # noqa: F401
