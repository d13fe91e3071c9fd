"""Time get_actions (encoder + decode) at one shape for the current MATAS_* environment: python scripts/decode_sweep.py B W [det stride]"""
import sys, torch
sys.path.insert(0, '.')
from mat_dcml_b200.ma_transformer import MultiAgentTransformer
from mat_dcml_b200 import DcmlVecEnv
from oracle import mat_oracle as mo
B, W = int(sys.argv[1]), int(sys.argv[2]); L = W + 1
det = len(sys.argv) > 3 and sys.argv[3] == "1"
stride = int(sys.argv[4]) if len(sys.argv) > 4 else 1
m = MultiAgentTransformer(L + 1, 7, 2, L, 2, 64, 2, device=torch.device('cuda'), action_type='Semi_Discrete')
m.load_state_dict(mo.full_state_dict(mo.make_weights(1, L), L)); m.output_device = torch.device('cuda')
env = DcmlVecEnv(B, W, seed=1)
obs, sob, ava = env.reset_device()
def timeit(fn, n=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
t_enc = timeit(lambda: m.get_values(None, obs))
t_act = timeit(lambda: m.get_actions(None, obs, ava, deterministic=det, stride=stride))
print("B=%d W=%d det=%d: encoder %.3f ms, get_actions %.3f ms, decode ~%.3f ms" % (B, W, det, t_enc, t_act, t_act - t_enc), flush=True)
