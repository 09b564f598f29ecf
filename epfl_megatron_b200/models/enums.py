"""Enumerations shared by models, arguments and checkpoints.

The reference pickles ``PositionEmbeddingType`` members inside the ``args`` Namespace of
every checkpoint (megatron/arguments.py:245-246, megatron/model/enums.py:6-28), so the
``megatron.model.enums`` facade re-exports exactly these classes.
"""
import enum


class ModelType(enum.Enum):
    encoder_or_decoder = 1
    encoder_and_decoder = 2


class LayerType(enum.Enum):
    encoder = 1
    decoder = 2


class AttnType(enum.Enum):
    self_attn = 1
    cross_attn = 2


class AttnMaskType(enum.Enum):
    padding = 1
    causal = 2


class PositionEmbeddingType(enum.Enum):
    rotary = 1
    absolute = 2

    def __str__(self):  # argparse prints choices with str()
        return self.name
