"""vlb-b200: B200-native (sm_100a) implementation of the phantom_vlb training-step hot path."""
__version__ = "0.1.0"
