"""inklayer_b200: B200-native drop-in for InkLayer's SAM segmentation hot path."""
from .config import CONFIGS, SamConfig  # noqa: F401
