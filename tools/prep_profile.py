"""One launch of each input-path kernel at the config-2 shapes, for `ncu --set full` (profiles/r1_input_path_ncu.md)."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from roma_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
planar = torch.randint(0, 256, (8, 3, 864, 864), dtype=torch.uint8, device=dev)
hwc = torch.randint(0, 256, (8, 768, 1024, 3), dtype=torch.uint8, device=dev)
for _ in range(2):
    ops.prepare_images(planar, (560, 560))                            # ATen-antialias semantics, uint8 tensor batches
    ops.prepare_images(planar, (864, 864))                            # normalise only
    ops.prepare_images_pil(hwc, (560, 560), channels_last=True)       # Pillow semantics, interleaved as decoded
    ops.prepare_images_pil(hwc, (864, 864), channels_last=True)
torch.cuda.synchronize()
