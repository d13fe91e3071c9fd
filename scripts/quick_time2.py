import sys, os, torch
sys.path.insert(0, '.')
from mat_dcml_b200.ma_transformer import MultiAgentTransformer
from mat_dcml_b200 import DcmlVecEnv
from oracle import mat_oracle as mo
def timeit(fn, n=3, warm=1):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
B, W = int(sys.argv[1]), int(sys.argv[2]); L = W + 1
m = MultiAgentTransformer(L + 1, 7, 2, L, 2, 64, 2, device=torch.device('cuda'), action_type='Semi_Discrete')
m.load_state_dict(mo.full_state_dict(mo.make_weights(1, L), L)); m.output_device = torch.device('cuda')
env = DcmlVecEnv(B, W, seed=1)
obs, sob, ava = env.reset_device()
t_val = timeit(lambda: m.get_values(None, obs))
t_sto = timeit(lambda: m.get_actions(None, obs, ava, deterministic=False))
a, lp, v = m.get_actions(None, obs, ava, deterministic=False)
with torch.no_grad():
    t_ev = timeit(lambda: m(None, obs, a, ava))
print("TOKEN=%s DECODE=%s B=%d W=%d: encoder %.2f ms, get_actions %.2f ms (decode %.2f), evaluate %.2f ms" % (
    os.environ.get("MATAS_RPT_TOKEN"), os.environ.get("MATAS_RPT_DECODE"), B, W, t_val, t_sto, t_sto - t_val, t_ev), flush=True)
