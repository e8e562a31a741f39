#!/usr/bin/env python
"""Refinement (tools/refinement.py of the reference) for SEVERAL classes in one process:
    python tools/refinement_multi.py --config configs/config.yaml --cls "office chair" table ...
Writes what one reference run per class writes (final_output_dir/<cls>/<scene>.pth, checkpoints/refinement_checkpoint_<cls>.yaml);
the stage-1 masks of a scene are read and decoded once for all classes and CLIP is loaded once. Body: bff_b200/refine.py."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "beyond-fixed-forms_b200"))

from bff_b200.refine import main_multi  # noqa: E402

if __name__ == "__main__":
    sys.exit(main_multi())
