"""cnn-itmo_b200 — B200-native (sm_100a) implementation of the CNN-ITMO U-Net hot path."""
__all__ = []
