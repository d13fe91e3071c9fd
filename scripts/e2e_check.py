import sys, time, numpy as np, torch
sys.path.insert(0, '.')
from mat_dcml_b200.ma_transformer import MultiAgentTransformer
from mat_dcml_b200 import DcmlVecEnv
from oracle import mat_oracle as mo
B, W = 16384, 50; L = W + 1
m = MultiAgentTransformer(L + 1, 7, 2, L, 2, 64, 2, device=torch.device('cuda'), action_type='Semi_Discrete')
m.load_state_dict(mo.full_state_dict(mo.make_weights(1, L), L))
for dt, rep in ((np.float32, False), (np.float64, True)):
    env = DcmlVecEnv(B, W, seed=1, host_dtype=dt, replicate_share_obs=rep)
    env2 = DcmlVecEnv(B, W, seed=1, host_dtype=dt, replicate_share_obs=rep); env2.HOST_SLICE_MIN = 1 << 40
    obs, sob, ava = env.reset(); o2, s2, a2 = env2.reset()
    assert np.array_equal(obs, o2)
    # sliced vs one-piece paths give identical results
    m.rollout_step = 5
    a, lp, v = m.get_actions(None, obs, ava)
    m.rollout_step = 5; m.HOST_SLICE_MIN = 1 << 40
    b, lq, w = m.get_actions(None, obs, ava)
    m.HOST_SLICE_MIN = 2048
    assert torch.equal(a, b) and torch.equal(lp, lq) and torch.equal(v, w), "sliced get_actions differs"
    r1 = env.step(a.numpy().reshape(B, L, 1)); r2 = env2.step(a.numpy().reshape(B, L, 1))
    for x, y in zip(r1, r2):
        if isinstance(x, np.ndarray): assert np.array_equal(x, y)
    assert [r1[4][i][0]["delay"] for i in (0, 5000, B - 1)] == [r2[4][i][0]["delay"] for i in (0, 5000, B - 1)]
    obs, ava = r1[0], r1[5]
    for name, e_ in (("sliced", env), ("one-piece", env2)):
        m.HOST_SLICE_MIN = 2048 if name == "sliced" else 1 << 40
        for it in range(13):
            if it == 3: torch.cuda.synchronize(); t0 = time.perf_counter()
            a, lp, v = m.get_actions(None, obs, ava); v = v.cpu()
            obs, sob, rew, dones, infos, ava = e_.step(a.numpy().reshape(B, L, 1))
        torch.cuda.synchronize()
        print(dt.__name__, name, "%.2f ms/step" % ((time.perf_counter() - t0) / 10 * 1e3), flush=True)
