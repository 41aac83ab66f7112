"""The memory-bound pre-processing kernels once each (after one warm call) for an ncu capture:
   ncu --set full --clock-control none --import-source on -k regex:'gaussian|noise|impulse|copy_u8|resample|normalize' \
       --launch-skip <warm launches> -o gpurun_out/membound_full python tools/prof_membound.py"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cid_b200 import noiser as nz  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(5)
big = torch.randint(0, 256, (256, 512, 512, 3), dtype=torch.uint8, device=dev, generator=g)  # 201 MB > L2
small = big[:4].contiguous()
calls = [
    lambda t: nz.add_gaussian_noise(t, 0.0, 25.0, seed=1),
    lambda t: nz.add_uniform_noise(t, 0.0, 25.0, seed=2),
    lambda t: nz.add_speckle_noise(t, 0.0, 0.1, seed=3),
    lambda t: nz.add_salt_pepper_noise(t, 0.02, 0.02, seed=4),
    lambda t: nz.resize_bicubic(t, (256, 256)),
    lambda t: nz.normalize_to_tensor(t),
]
if hasattr(nz, "add_gaussian_unit"):
    calls.append(lambda t: nz.add_gaussian_unit(t, seed=6))
    calls.append(lambda t: nz.add_salt_pepper_unit(t, seed=7))
for f in calls:       # warm (module load, coefficient tables): small inputs
    f(small)
torch.cuda.synchronize()
print("WARM_DONE")
for f in calls:
    f(big)
torch.cuda.synchronize()
