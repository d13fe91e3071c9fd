"""Wall-clock split of one end-to-end rollout step (host arrays in and out): python scripts/e2e_breakdown.py [B W]"""
import sys, time, numpy as np, torch
sys.path.insert(0, '.')
import bench
from mat_dcml_b200 import TransformerPolicy, DcmlVecEnv, SharedReplayBuffer
B, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (16384, 50)
L, T = W + 1, 50
dev = torch.device("cuda")
env = DcmlVecEnv(B, W, seed=1, device=dev, host_dtype=np.float32, replicate_share_obs=False)
pol = TransformerPolicy(bench.model_args(), [7], [W + 2], env.action_space[0], L, device=dev)
buf = SharedReplayBuffer(bench.model_args(episode_length=T, n_rollout_threads=B), L, [7], [W + 2], env.action_space[0], "DCML", device=dev)
obs, sob, ava = env.reset()
acc = {}
def tick(name, t0):
    torch.cuda.synchronize(); t1 = time.perf_counter(); acc[name] = acc.get(name, 0.0) + (t1 - t0); return t1
for it in range(23):
    if it == 3: acc.clear()
    torch.cuda.synchronize(); t = time.perf_counter(); t00 = t
    v, a, lp, _, _ = pol.get_actions(None, obs, None, None, None, ava); t = tick("get_actions", t)
    v = pol.transformer.to_host(v); t = tick("v.cpu", t)
    obs, sob, rew, dones, infos, ava = env.step(a.numpy().reshape(B, L, 1)); t = tick("env.step", t)
    buf.insert(sob[:, 0, :], obs, None, None, a, lp, v, rew[:, 0, 0], 1.0 - dones[:, 0].astype(np.float32), None, None, ava); t = tick("insert", t)
    acc["total"] = acc.get("total", 0.0) + (t - t00)
print({k: round(v / 20 * 1e3, 3) for k, v in acc.items()})
