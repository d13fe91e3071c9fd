import sys, torch, numpy as np
sys.path.insert(0, '.')
from mat_dcml_b200.ma_transformer import MultiAgentTransformer
M, L = int(sys.argv[1]), int(sys.argv[2])
torch.manual_seed(0)
m = MultiAgentTransformer(L + 1, 7, 2, L, 2, 64, 2, device=torch.device('cuda'), action_type='Semi_Discrete')
obs = torch.rand(M, L, 7, device='cuda'); ava = torch.ones(M, L, 2, device='cuda')
act = (torch.rand(M, L, 1, device='cuda') < 0.5).float()
def ev(tag):
    e = torch.cuda.Event(enable_timing=True); e.record(); return e
for it in range(3):
    m.flat_grad(zero=True)
    e0 = ev(0)
    with torch.no_grad():
        m(None, obs, act, ava)
    e1 = ev(1)
    lp, v, ent = m(None, obs, act, ava)
    e2 = ev(2)
    (lp.sum() + v.sum() + ent.sum()).backward()
    e3 = ev(3)
    torch.cuda.synchronize()
print("M=%d L=%d: evaluate (inference) %.2f ms, evaluate (train, saves) %.2f ms, backward %.2f ms -> %.0f rows/s fwd+bwd" % (
    M, L, e0.elapsed_time(e1), e1.elapsed_time(e2), e2.elapsed_time(e3), M / ((e1.elapsed_time(e3)) * 1e-3)))
