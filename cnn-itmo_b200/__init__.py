"""cnn-itmo_b200 — B200-native (sm_100a) implementation of the one hot path of LynxHack/CNN-ITMO:
the Keras U-Net forward pass predict.py runs (SDR -> Reinhard-tone-mapped HDR estimate).

    from cnn_itmo_b200 import U_net, load_model      # drop-in for model.py:204 / predict.py:24

Importing the package does not need a GPU; the first call that does real work loads libitmo_b200.so
and fails loudly if it is missing — there is no CPU fallback.
"""
from .graph import LAYERS, count_params, forward_flops_per_image, summary_text  # noqa: F401
from .model import Model, U_net, initial_weights, load_model  # noqa: F401
from . import frames  # noqa: F401
from . import callbacks  # noqa: F401
from . import datagen  # noqa: F401
from . import train  # noqa: F401
from .checkpoint import read_checkpoint, write_checkpoint  # noqa: F401

__all__ = ['U_net', 'load_model', 'Model', 'initial_weights', 'read_checkpoint', 'write_checkpoint', 'LAYERS',
           'count_params', 'forward_flops_per_image', 'summary_text']
