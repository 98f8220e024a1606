"""B200-native encode/decode hot path of the learned point-cloud codec.

Import name: ``pcc_b200`` (the repo-root ``pcc_b200.py`` loader registers this directory, whose
on-disk name is not a Python identifier, under that name).
"""
from . import _lib, ops, me, compressai_compat, codec, synthetic, frames, pipeline, slab, scene, bitstream, qmap, autograd, loss  # noqa: F401
from .codec import ColorModel, OURS_CONFIG, ABLATION_CONFIG, init_synthetic, export_params, load_config  # noqa: F401
from ._lib import PccError  # noqa: F401
