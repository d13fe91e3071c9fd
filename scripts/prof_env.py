"""A few DCML env steps at one shape (for ncu): python scripts/prof_env.py B W"""
import sys, torch
sys.path.insert(0, '.')
from mat_dcml_b200 import DcmlVecEnv
B, W = int(sys.argv[1]), int(sys.argv[2]); L = W + 1
env = DcmlVecEnv(B, W, seed=1)
obs, sob, ava = env.reset_device()
g = torch.Generator(device="cuda").manual_seed(0)
for _ in range(4):
    act = torch.cat([(torch.rand(B, W, device="cuda", generator=g) < 0.5).float() * ava[:, :W, 1], torch.rand(B, 1, device="cuda", generator=g)], 1)
    obs, sob, rew, done, delay, pay, ava = env.step_device(act)
torch.cuda.synchronize()
print("ok", float(rew.mean()))
