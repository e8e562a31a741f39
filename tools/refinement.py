#!/usr/bin/env python
"""Drop-in for the reference's tools/refinement.py (same argv: --config <yaml> --cls <query>, same
config keys, same <final_output_dir>/<cls>/<scene>.pth outputs that evaluation/eval/eval_scannet200.py
reads), with the mask algebra on the B200 through libbff.so. CLIP ("ViT-L/14", the reference's model)
is loaded with the `clip` package exactly like the reference does. No CPU fallback."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "beyond-fixed-forms_b200"))

from bff_b200.refine import main  # noqa: E402
# the reference's importable functions, same names and signatures, served by the GPU library
from bff_b200.reference_api import (  # noqa: E402,F401
    rle_decode, calculate_iou_between_stages)

if __name__ == "__main__":
    sys.exit(main())
