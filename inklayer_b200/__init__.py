"""inklayer_b200: B200-native drop-in for InkLayer's SAM segmentation hot path.

    from inklayer_b200 import build_sam, SamPredictor      # same surface as `from segment_anything import ...`
"""
from .config import CONFIGS, SamConfig  # noqa: F401
from .predictor import SamPredictor  # noqa: F401
from .resize import ResizeLongestSide  # noqa: F401
from .sam import (Sam, build_sam, build_sam_from_state_dict, build_sam_vit_b, build_sam_vit_h, build_sam_vit_l,  # noqa: F401
                  sam_model_registry)
