"""lipvq -- B200-native LipVQ-VAE action tokenizer (hand-written sm_100a kernels behind the reference's module API).

Importing this package loads liblipvq.so; it raises ImportError when the extension has not been built.
"""
from . import _cabi
from ._cabi import (ACT_GELU, ACT_GIVEN, ACT_NONE, ACT_RELU, ACT_SIGMOID, ALGO_AUTO, ALGO_SIMT_F32, ALGO_TCGEN05, EMA_LAPLACE,
                    EMA_NORMALIZE, EMA_SIMPLE, METRIC_DOT, METRIC_L2, METRIC_L2SQ, launch_count)
from . import ops
from .modules import LFQQuantizer, LFQQuantizerEMA, LipschitzMLP, LLFQVAE_V4
from .classic import VQVAE
from .hvq import (ActionDecoder, ActionEncoder, DilatedTemporalConvLayer, HierarchicalVQVAE, MSTCNStage)
from .hvq_trainer import HvqTrainer
from .hq import HierarchicalLFQHVQVAE, LFQQuantizerEMA_KMeans
from .bin_action import AdaptiveBinActionEmbedding
from .dataset import ActionSequenceDataset
from . import icrt, icl_algo

__all__ = ["ops", "LLFQVAE_V4", "LipschitzMLP", "LFQQuantizer", "LFQQuantizerEMA", "LFQQuantizerEMA_KMeans",
           "HierarchicalLFQHVQVAE", "VQVAE", "HierarchicalVQVAE", "HvqTrainer", "ActionEncoder", "ActionDecoder", "MSTCNStage",
           "DilatedTemporalConvLayer", "AdaptiveBinActionEmbedding", "ActionSequenceDataset", "icrt", "icl_algo", "launch_count"]
