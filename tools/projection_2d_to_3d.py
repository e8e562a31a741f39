#!/usr/bin/env python
"""Drop-in for the reference's tools/projection_2d_to_3d.py (same argv: --config <yaml> --cls <query>,
same config keys, same <mask_3d_dir>/<cls>/<scene>.pth outputs and checkpoint yaml), running the
lifting / aggregation / filtering on the B200 through libbff.so. Under torchrun the frames of each
scene are sharded over the ranks (NCCL). No CPU fallback."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "beyond-fixed-forms_b200"))

from bff_b200.stage2 import main  # noqa: E402
# the reference's importable functions, same names and signatures, served by the GPU library
from bff_b200.reference_api import (  # noqa: E402,F401
    compute_projected_pts_tensor, compute_visibility_mask_tensor, compute_visible_masked_pts_tensor, aggregate, calculate_iou, find_unconnected_subgraphs_tensor, solve_overlapping, cfg)

if __name__ == "__main__":
    sys.exit(main())
