"""torchrun --nproc-per-node 2 scripts/dist_check.py : the PPO update sharded over 2 GPUs (NCCL all-reduce of
the flat gradient + global denominators) must reproduce the reference's single-process ppo_update fixture."""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from test_train_gpu import _targs
from oracle import mat_oracle as mo
from mat_dcml_b200 import TransformerPolicy, Action_Space, MATTrainer

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "ppo_L11.npz"))
L, M = int(g["L"]), int(g["M"])
assert M % world == 0
args = _targs()
pol = TransformerPolicy(args, [7], [L + 1], Action_Space(2, semi_index=-1, extra=True), L, device=dev)
pol.transformer.load_state_dict(mo.full_state_dict(mo.make_weights(int(g["wseed"]), L), L))
tr = MATTrainer(args, pol, L, device=dev)
tr.value_normalizer.stats.copy_(torch.from_numpy(g["vn_stats_before"]))
rows = slice(rank * (M // world) * L, (rank + 1) * (M // world) * L)
f = lambda k, dt=np.float32: g[k][rows].astype(dt)
sample = (None, f("obs"), None, None, f("actions"), f("value_preds"), f("returns"), None, f("active"), f("old_logp"), f("adv"), f("ava"))
vl, gn, pl, ent, _, ratio = tr.ppo_update(sample)
tot = torch.stack([vl.detach(), pl.detach(), ent.detach()])          # every rank holds the global losses (they ride on the gradient all-reduce)
ok = abs(tot[0].item() - float(g["value_loss"])) < 1e-4 * max(1, abs(float(g["value_loss"]))) and \
     abs(tot[1].item() - float(g["policy_loss"])) < 1e-4 and abs(tot[2].item() - float(g["entropy"])) < 1e-4 and \
     abs(float(gn) - float(g["grad_norm"])) < 2e-3 * float(g["grad_norm"])
params = dict(pol.transformer.named_parameters())
for name, st in zip(g["grad_names"], g["grad_stats"]):
    gr = params[str(name)].grad.detach().cpu().numpy()
    scale = max(st[1], 1e-4 * float(g["grad_norm"]))
    ok = ok and abs(np.sqrt((gr.astype(np.float64) ** 2).sum()) - st[1]) < 3e-3 * scale
# the all-reduced flat gradient equals the single-GPU gradient of the concatenated batch.
# The same update on the whole fixture batch, on this GPU alone (the trainer's collectives switched off):
import mat_dcml_b200.mat_trainer as mt
dist_grad = pol.transformer.flat_grad().detach().clone()
pol1 = TransformerPolicy(args, [7], [L + 1], Action_Space(2, semi_index=-1, extra=True), L, device=dev)
pol1.transformer.load_state_dict(mo.full_state_dict(mo.make_weights(int(g["wseed"]), L), L))
tr1 = MATTrainer(args, pol1, L, device=dev)
tr1.value_normalizer.stats.copy_(torch.from_numpy(g["vn_stats_before"]))
fa = lambda k, dt=np.float32: g[k].astype(dt)
full = (None, fa("obs"), None, None, fa("actions"), fa("value_preds"), fa("returns"), None, fa("active"), fa("old_logp"), fa("adv"), fa("ava"))
world_fn, mt._world = mt._world, (lambda: 1)
try:
    vl1, gn1, pl1, ent1, _, _ = tr1.ppo_update(full)
finally:
    mt._world = world_fn
one_grad = pol1.transformer.flat_grad().detach()
gerr = float((dist_grad - one_grad).abs().max()) / float(one_grad.abs().max())
ok = ok and gerr < 2e-4 and abs(float(gn) - float(gn1)) < 1e-4 * float(gn1)
ok = ok and float((pol.transformer.flat_parameters() - pol1.transformer.flat_parameters()).abs().max()) < 1e-6
print("rank %d: max |all-reduced grad - single-GPU grad| / max |grad| = %.2e" % (rank, gerr))
print("entropy %.6f (ref %.6f)" % (tot[2].item(), float(g["entropy"])))
print("rank %d: value_loss %.6f (ref %.6f) policy_loss %.6f (ref %.6f) grad_norm %.5f (ref %.5f) -> %s"
      % (rank, tot[0].item(), float(g["value_loss"]), tot[1].item(), float(g["policy_loss"]), float(gn), float(g["grad_norm"]),
         "OK" if ok else "MISMATCH"), flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
