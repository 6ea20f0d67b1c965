"""One launch of every RAD augmentation kernel (uint8 and float32, 1024 x 9 x 100 x 100) for an ncu capture:
ncu --set full --clock-control none -k regex:'augment|grayscale' -o gpurun_out/r01_rad_prof python tools/rad_profile.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import s2p_b200.data_augs as R  # noqa: E402

np.random.seed(0)
B = 1024
for x in (torch.randint(0, 256, (B, 9, 100, 100), device="cuda", dtype=torch.uint8), torch.rand(B, 9, 100, 100, device="cuda")):
    R.random_crop(x, out=84)
    R.random_translate(x, 108)
    R.random_cutout(x)
    R.random_flip(x, p=.5)
    R.random_rotation(x, p=.9)
R.random_grayscale(x, p=.5)
torch.cuda.synchronize()
print("rad_profile: 11 launches")
