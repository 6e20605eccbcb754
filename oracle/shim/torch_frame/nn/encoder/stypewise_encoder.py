class StypeWiseFeatureEncoder:  # annotation-only use at models/fused.py:22,50
    pass
