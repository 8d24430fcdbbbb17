"""ghost_b200 - B200-native (sm_100a) face-swap inference hot path behind GHOST's operator API.

The numerical work lives in ghost_b200/libghost_b200.so (hand-written CUDA behind the C ABI of
include/ghost_b200.h); this package is the host-side mirror of the reference's Python interface.
"""
from . import _lib
from .api import (AEI_Net, IResNet100, iresnet100, faceshifter_batch, transform_target_to_torch,
                  normalize_and_torch, normalize_and_torch_batch, get_final_image, get_final_video, paste_back, paste_boxes, SwapPipeline,
                  LIPSPADEGenerator, Pix2PixModel, face_enhancement, crop_faces, match_identities, embed_and_swap)
from . import masks
from .shard import shard_range, FrameSharder, PeerGather

__all__ = ["AEI_Net", "IResNet100", "iresnet100", "faceshifter_batch", "transform_target_to_torch",
           "normalize_and_torch", "normalize_and_torch_batch", "get_final_image", "get_final_video", "paste_back",
           "SwapPipeline", "paste_boxes", "shard_range", "FrameSharder", "PeerGather", "LIPSPADEGenerator", "Pix2PixModel", "face_enhancement", "crop_faces", "match_identities", "embed_and_swap", "masks"]
__version__ = "0.1.0"
