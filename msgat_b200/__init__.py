"""msgat_b200 -- B200-native (sm_100a) implementation of the MS-GAT attention hot path.

``msgat_b200.models`` mirrors the reference's ``src/models`` package; ``msgat_b200.ops``
holds the autograd Functions over the C ABI of ``csrc/libmsgat_sm100.so``."""
from .models import msgat48, msgat72, msgat96  # noqa: F401
