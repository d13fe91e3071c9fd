"""Where one PPO minibatch goes: torch profiler table of MATTrainer.ppo_update at the bench's PPO shape."""
import argparse, sys, torch
sys.path.insert(0, '.')
import bench
from mat_dcml_b200 import TransformerPolicy, MATTrainer, DcmlVecEnv, SharedReplayBuffer
dev = torch.device("cuda")
N, W, T = 1024, 20, 50
L = W + 1
args = bench.model_args(episode_length=T, n_rollout_threads=N, clip_param=0.2, ppo_epoch=1, num_mini_batch=4,
                        value_loss_coef=1.0, entropy_coef=0.01, max_grad_norm=10.0, huber_delta=10.0, use_max_grad_norm=True,
                        use_clipped_value_loss=True, use_huber_loss=True, use_value_active_masks=True)
env = DcmlVecEnv(N, W, seed=1)
pol = TransformerPolicy(args, [7], [W + 2], env.action_space[0], L, device=dev)
pol.transformer.output_device = dev
tr = MATTrainer(args, pol, L, device=dev)
buf = SharedReplayBuffer(args, L, [7], [W + 2], env.action_space[0], "DCML")
obs, sob, ava = env.reset_device()
buf.obs[0].copy_(obs); buf._share_obs[0].copy_(env.sob); buf.available_actions[0].copy_(ava)
for t in range(T):
    v, a, lp, _, _ = pol.get_actions(None, obs.reshape(-1, 7), None, None, None, ava.reshape(-1, 2))
    obs, sob, rew, done, delay, pay, ava = env.step_device(a.reshape(N, L))
    buf.insert(env.sob, obs, None, None, a.reshape(N, L, 1), lp.reshape(N, L, 1), v.reshape(N, L, 1), rew, 1.0 - done.float(), None, None, ava)
tr.train(buf)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); tr.train(buf); e1.record(); torch.cuda.synchronize()
print("one train() = %d epochs x %d minibatches: %.2f ms" % (tr.ppo_epoch, tr.num_mini_batch, e0.elapsed_time(e1)))
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=60))
