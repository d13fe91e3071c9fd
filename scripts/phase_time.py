import sys, ctypes, torch
sys.path.insert(0, '.')
from mat_dcml_b200.ma_transformer import MultiAgentTransformer
from mat_dcml_b200 import DcmlVecEnv, _cabi
from oracle import mat_oracle as mo
B, W = int(sys.argv[1]), int(sys.argv[2]); L = W + 1
m = MultiAgentTransformer(L + 1, 7, 2, L, 2, 64, 2, device=torch.device('cuda'), action_type='Semi_Discrete')
m.load_state_dict(mo.full_state_dict(mo.make_weights(1, L), L)); m.output_device = torch.device('cuda')
env = DcmlVecEnv(B, W, seed=1)
obs, sob, ava = env.reset_device()
m.get_actions(None, obs, ava)
_cabi.check(_cabi.lib().matas_policy_decode_phases(m._h(), 1, None))
m.get_actions(None, obs, ava); torch.cuda.synchronize()
out = (ctypes.c_ulonglong * 16)()
_cabi.check(_cabi.lib().matas_policy_decode_phases(m._h(), 0, out))
names = ["inputs", "query GEMMs x2", "self attention", "out GEMMs x2+LN1", "cross inputs", "cross attention", "out GEMMs x2+LN2", "MLP.0", "MLP.2+LN3", "head GEMM", "pick"]
tot = sum(out[:11])
for n, c in zip(names, out):
    print("%-20s %12d cyc  %5.1f%%" % (n, c, 100.0 * c / tot))
print("total %d cycles = %.2f ms at 1.965 GHz" % (tot, tot / 1.965e6))
