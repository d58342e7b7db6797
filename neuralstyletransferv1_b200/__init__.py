"""B200-native (sm_100a) per-frame stylisation hot path of TrentMahaffey/NeuralStyleTransferV1.

Public surface (mirrors the reference's):
  TransformerNet, TransformerNetNST, ReCoNet, ReCoNetModel   -- drop-in nn.Modules
  Engine, composite_regions, io_preset_affine                -- the fast uint8 / compositing API
  sharding.shard_range                                        -- frame-batch partitioning over GPUs
"""
from .engine import (Engine, composite_from_crops, composite_regions, composite_regions_batch, flow_ema,  # noqa: F401
                     io_preset_affine)
from .frame_pipeline import FramePipeline, MultiGpuFramePipeline  # noqa: F401
from .region_masks import (MorphAnimation, RegionBlender, assign_models_to_regions, feather_masks, generate_region_masks,  # noqa: F401
                           rotate_all_masks, voronoi_masks, warp_all_masks_organic)
from .temporal import LabEma, motion_blend  # noqa: F401
from .post_chain import PostChain  # noqa: F401
from .region_anim import BlendAnimation, ScaleAnimation, animated_scale, animated_weights, harmonic_value  # noqa: F401
from .region_crops import RegionConfig, RegionCrop, RegionCropStyler, mask_bboxes, prepare_region_crops  # noqa: F401
from .modules import ReCoNet, TransformerNet, TransformerNetNST  # noqa: F401
from .reconet_model import ReCoNetModel  # noqa: F401
from .sharding import shard_range, shard_sizes  # noqa: F401

__all__ = ["Engine", "composite_regions", "composite_regions_batch", "composite_from_crops", "flow_ema", "FramePipeline",
           "MultiGpuFramePipeline", "voronoi_masks", "generate_region_masks", "rotate_all_masks", "warp_all_masks_organic",
           "MorphAnimation", "RegionBlender", "assign_models_to_regions", "feather_masks", "LabEma", "motion_blend", "PostChain", "BlendAnimation", "ScaleAnimation", "animated_scale", "animated_weights", "harmonic_value", "RegionConfig", "RegionCrop", "RegionCropStyler", "mask_bboxes", "prepare_region_crops", "io_preset_affine", "TransformerNet", "TransformerNetNST",
           "ReCoNet", "ReCoNetModel", "shard_range", "shard_sizes"]
