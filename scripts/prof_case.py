import sys, torch
sys.path.insert(0, '.')
from mat_dcml_b200.ma_transformer import MultiAgentTransformer
from mat_dcml_b200 import DcmlVecEnv
from oracle import mat_oracle as mo
B, W = int(sys.argv[1]), int(sys.argv[2]); L = W + 1
m = MultiAgentTransformer(L + 1, 7, 2, L, 2, 64, 2, device=torch.device('cuda'), action_type='Semi_Discrete')
m.load_state_dict(mo.full_state_dict(mo.make_weights(1, L), L)); m.output_device = torch.device('cuda')
env = DcmlVecEnv(B, W, seed=1)
obs, sob, ava = env.reset_device()
for _ in range(2):
    a, lp, v = m.get_actions(None, obs, ava, deterministic=False)
torch.cuda.synchronize()
print("ok", float(lp.sum()))
