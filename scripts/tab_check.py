"""Table-mode decode against the per-env-cache decode at bench sizes (separate processes: the switch is read once).
    python scripts/tab_check.py            # runs both modes for every case and compares
    python scripts/tab_check.py dump OUT   # one mode (MATAS_DECODE_TAB from the environment) -> OUT.npz
"""
import os, subprocess, sys
import numpy as np
CASES = [(16384, 50, False, 1), (4096, 200, False, 1), (1024, 20, True, 10), (2048, 50, True, 2), (512, 100, True, 7), (300, 20, True, 1)]
if len(sys.argv) > 2 and sys.argv[1] == "dump":
    import torch
    sys.path.insert(0, '.')
    from mat_dcml_b200.ma_transformer import MultiAgentTransformer
    from mat_dcml_b200 import DcmlVecEnv
    from oracle import mat_oracle as mo
    out = {}
    for i, (B, W, det, stride) in enumerate(CASES):
        L = W + 1
        m = MultiAgentTransformer(L + 1, 7, 2, L, 2, 64, 2, device=torch.device('cuda'), action_type='Semi_Discrete')
        m.load_state_dict(mo.full_state_dict(mo.make_weights(3 + i, L), L)); m.output_device = torch.device('cuda')
        m.rollout_seed = 1234
        env = DcmlVecEnv(B, W, seed=2 + i)
        obs, sob, ava = env.reset_device()
        a, lp, v = m.get_actions(None, obs, ava, deterministic=det, stride=stride)
        out["a%d" % i], out["lp%d" % i] = a.cpu().numpy(), lp.cpu().numpy()
    np.savez(sys.argv[2], **out)
    sys.exit(0)
for tab in ("0", "1"):
    subprocess.check_call([sys.executable, __file__, "dump", "/tmp/tab%s" % tab], env=dict(os.environ, MATAS_DECODE_TAB=tab))
x, y = np.load("/tmp/tab0.npz"), np.load("/tmp/tab1.npz")
ok = True
for i, (B, W, det, stride) in enumerate(CASES):
    L = W + 1
    a0, a1 = x["a%d" % i].reshape(B, L), y["a%d" % i].reshape(B, L)
    l0, l1 = x["lp%d" % i].reshape(B, L), y["lp%d" % i].reshape(B, L)
    same = (a0[:, :-1] == a1[:, :-1]).all(1)
    dl = np.abs(l0[same] - l1[same]).max() if same.any() else float("nan")
    da = np.abs(a0[same, -1] - a1[same, -1]).max() if same.any() else float("nan")
    print("B=%d W=%d det=%d stride=%d: identical discrete rows %.4f, max |dlogp| %.2e, max |d a_cont| %.2e" % (B, W, det, stride, same.mean(), dl, da))
    ok = ok and same.mean() > 0.99 and dl < 1e-4 and da < 1e-4
print("OK" if ok else "MISMATCH")
sys.exit(0 if ok else 1)
