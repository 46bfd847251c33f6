"""Raw image I/O, synthetic images, timing and clock sampling."""
from .rawio import output_name, read_block, read_raw, write_block, write_raw
from .synth import random_image, waterfall_image

__all__ = ["read_raw", "write_raw", "read_block", "write_block", "output_name", "random_image", "waterfall_image"]
