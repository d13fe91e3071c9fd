"""Wall-clock breakdown of one end-to-end rollout step through the host-array API (bench.py's e2e_step) at C3."""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import bench
from mat_dcml_b200 import TransformerPolicy, DcmlVecEnv, SharedReplayBuffer, ValueNorm
B, W, T = 16384, 50, 50; L = W + 1
dev = torch.device("cuda")
env = DcmlVecEnv(B, W, seed=1, device=dev, host_dtype=np.float32, replicate_share_obs=False)
pol = TransformerPolicy(bench.model_args(), [7], [W + 2], env.action_space[0], L, device=dev)
buf = SharedReplayBuffer(bench.model_args(episode_length=T, n_rollout_threads=B), L, [7], [W + 2], env.action_space[0], "DCML", device=dev)
obs, sob, ava = env.reset()
acc = {}
def tick(name, t0):
    torch.cuda.synchronize(); t1 = time.perf_counter(); acc[name] = acc.get(name, 0.0) + (t1 - t0); return t1
for i in range(13):
    if i == 3: acc.clear()
    torch.cuda.synchronize(); t = time.perf_counter()
    v, a, lp, _, _ = pol.get_actions(None, obs, None, None, None, ava)
    t = tick("get_actions (H2D obs/ava + kernels + D2H a/lp)", t)
    v = v.cpu()
    t = tick("values.cpu()", t)
    obs, sob, rew, dones, infos, ava = env.step(a.numpy().reshape(B, L, 1))
    t = tick("env.step (H2D act + kernel + D2H outputs)", t)
    buf.insert(sob[:, 0, :], obs, None, None, a, lp, v, rew[:, 0, 0], 1.0 - dones[:, 0].astype(np.float32), None, None, ava)
    t = tick("buf.insert (async H2D, then waited)", t)
for k, x in acc.items():
    print("%-52s %7.3f ms" % (k, x / 10 * 1e3))
print("total %.3f ms (synchronised after every call; the bench does not wait for the insert)" % (sum(acc.values()) / 10 * 1e3))
