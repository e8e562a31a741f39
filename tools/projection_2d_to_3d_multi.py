#!/usr/bin/env python
"""Addition to the reference's surface: lifting for SEVERAL classes in one process,
`--config <yaml> --cls <q1> <q2> ...`. Writes exactly the files one reference run per class would write; the
projection / visibility / viewed-count work of a scene is done once and shared by all classes."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "beyond-fixed-forms_b200"))

from bff_b200.stage2 import main_multi  # noqa: E402

if __name__ == "__main__":
    sys.exit(main_multi())
