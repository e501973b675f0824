"""norm_b200 -- B200-native CUDA backend for norm's image-effect path.

Product = norm_b200/libnorm_cuda.so (C ABI in include/norm_cuda.h); this package is the thin
host-side mirror of the reference's effect API over it.  Importing never builds or falls
back: a missing extension raises at first use.
"""
from ._lib import (LIB_PATH, NORM_CS_RAW, NORM_CS_REFERENCE, NORM_EXPAND_COMPAT, NORM_EXPAND_SPEC,
                   NORM_MAX_BLUR_RADIUS, NORM_OUT_SRGB, NormFeatureUniform, NormPngDesc, PROTOTYPES, load)
from .api import (Band, Circle, Context, DeviceImage, DrawUniform, EffectPipeline, EffectPipelineBuilder, FeatureUniform, NormError, Png,
                  Texture, blur_taps, front_lut, plan_band, plan_frames, reference_pipeline, ShapeUniform, srgb_tables,
                  ssim_from_moments)

from . import sharding  # noqa: E402,F401

__all__ = [n for n in dir() if not n.startswith("_")]
