"""Model zoo: the tutorial Transformer LM in eager and fused (sm_100a) bodies."""
from .transformer import (CONFIGS, Decoder, Encoder, PositionalEncoding,  # noqa: F401
                          TransformerConfig, build_layer, build_layers, build_stage, build_stages,
                          count_parameters)
