import sys, time, cProfile, pstats
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from s2p_b200 import data_augs as rad
x = torch.randint(0, 256, (1024, 9, 100, 100), dtype=torch.uint8, device='cuda')
for _ in range(5): rad.random_crop(x, out=84)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200): rad.random_crop(x, out=84)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("host us/call", (t1 - t0) / 200 * 1e6, "incl drain", (t2 - t0) / 200 * 1e6)
pr = cProfile.Profile(); pr.enable()
for _ in range(200): rad.random_crop(x, out=84)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats('tottime').print_stats(14)
