import sys, torch
sys.path.insert(0, '.')
from mat_dcml_b200 import DcmlVecEnv
for B, W in ((16384, 50), (8192, 20), (4096, 100)):
    env = DcmlVecEnv(B, W, seed=1); L = W + 1
    obs, sob, ava = env.reset_device()
    g = torch.Generator(device="cuda").manual_seed(0)
    act = torch.cat([(torch.rand(B, W, device="cuda", generator=g) < 0.5).float() * ava[:, :W, 1], torch.rand(B, 1, device="cuda", generator=g)], 1)
    for _ in range(3): env.step_device(act)
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): env.step_device(act)
    e1.record(); torch.cuda.synchronize()
    print("B=%d W=%d env.step %.3f ms" % (B, W, e0.elapsed_time(e1) / 20))
