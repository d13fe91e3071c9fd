import sys, numpy as np, torch
sys.path.insert(0, '.')
from mat_dcml_b200.ma_transformer import MultiAgentTransformer
from mat_dcml_b200 import DcmlVecEnv
from oracle import mat_oracle as mo
dev = torch.device("cuda:0")
W, L, B = 20, 21, int(sys.argv[1]) if len(sys.argv) > 1 else 48
sd = mo.make_weights(1, L)
m = MultiAgentTransformer(L + 1, 7, 2, L, 2, 64, 2, device=dev, action_type="Semi_Discrete")
m.load_state_dict(mo.full_state_dict(sd, L))
env = DcmlVecEnv(B, W, seed=1, device=dev)
obs, sob, ava = env.reset()
a, lp, v = m.get_actions(None, obs, ava, deterministic=True, stride=10)
with torch.no_grad():
    lp2, v2, ent = m(None, obs, a, ava)
lpo, vo, eo_ = mo.evaluate_actions(sd, torch.as_tensor(obs), a.clone(), torch.as_tensor(ava), L)
d = (lp2.cpu() - lp).abs().reshape(B, L)
print("max |lp2-lp| %.3e at" % d.max(), np.unravel_index(int(d.argmax()), d.shape))
d2 = (lp2.cpu().reshape(-1) - lpo.reshape(-1)).abs().reshape(B, L)
print("max |lp2-oracle_eval| %.3e at" % d2.max(), np.unravel_index(int(d2.argmax()), d2.shape), "values err %.3e" % (v2.cpu().reshape(-1) - vo.reshape(-1)).abs().max())
print("rows with err>1e-4:", sorted(set(np.nonzero(d2.numpy() > 1e-4)[0].tolist()))[:20])
