#!/usr/bin/env python
"""Stages 2 and 3 of evaluation/eval/run_evl.py (lifting, refinement) for many classes in ONE process:
    python tools/run_evl_b200.py --config configs/config.yaml --cls "office chair" table ... [--batch 20]
Same stage files, same process_checkpoint.txt; a scene's inputs are loaded and uploaded once per batch of classes.
Body: bff_b200/run_classes.py."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "beyond-fixed-forms_b200"))

from bff_b200.run_classes import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
