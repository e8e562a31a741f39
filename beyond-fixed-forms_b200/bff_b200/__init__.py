"""bff_b200 -- B200-native (sm_100a) 2D->3D lifting, aggregation/filtering and refinement for
Beyond-Fixed-Forms. Host code is Python/PyTorch; all device work is in libbff.so (C ABI,
include/bff.h). There is no CPU fallback."""
from . import _cabi  # noqa: F401

__all__ = ["_cabi"]
__version__ = "0.1.0"
