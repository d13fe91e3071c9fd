"""Device-resident runner for the MAT-AS / DCML training loop.

Mirrors the reference's ``DCMLRunner.run / warmup / collect / insert`` (``dcml_runner.py:22-288``) and
``Runner.compute / train`` (``mat_src/mat/runner/shared/base_runner.py:286-325,429-435``) for
``algorithm_name == "mat"``: the same call order, the same arrays in the rollout buffer, the same masks, the same
per-episode statistics (average reward / delay / payment of finished episodes).  What differs is where the data
lives: observations, actions, rewards and masks stay on the GPU between the policy, the simulator and the buffer
(no ``_t2n`` / ``np.concatenate`` / ``np.split`` per step), ``share_obs`` is kept as one row per env, and the
statistics are accumulated on the device and read back once per logging interval.

    runner = DcmlRunner(dict(all_args=args, envs=DcmlVecEnv(...), num_agents=L, device=torch.device("cuda")))
    logs = runner.run()          # list of dicts, one per logged episode

``eval`` is the reference's deterministic evaluation loop (``dcml_runner.py:319-448``) on the same kernels.
TensorBoard / W&B writers and model-directory handling are outside the hot path (SURVEY.md section 8 "not
rebuilt"); ``log_train`` returns the numbers instead of writing them.
"""
import time

import numpy as np
import torch

from . import _cabi
from .mat_trainer import MATTrainer
from .shared_buffer import SharedReplayBuffer
from .transformer_policy import TransformerPolicy


class DcmlRunner:
    def __init__(self, config):
        self.all_args = a = config["all_args"]
        self.envs = config["envs"]
        self._eval_envs = config.get("eval_envs")
        self.device = torch.device(config.get("device", "cuda"))
        if self.device.type != "cuda":
            raise _cabi.MatasError("DcmlRunner needs a CUDA device (there is no CPU fallback)")
        self.num_agents = config["num_agents"]
        if a.algorithm_name != "mat":
            raise NotImplementedError("algorithm_name %r: only 'mat' is wired for DCML (base_runner.py:198)" % a.algorithm_name)
        # base_runner.py:20-45
        self.algorithm_name, self.experiment_name = a.algorithm_name, getattr(a, "experiment_name", "")
        self.num_env_steps, self.episode_length = a.num_env_steps, a.episode_length
        self.n_rollout_threads = a.n_rollout_threads
        self.use_linear_lr_decay = getattr(a, "use_linear_lr_decay", False)
        self.use_centralized_V = getattr(a, "use_centralized_V", True)
        if not self.use_centralized_V:
            # base_runner.py:190-193 would hand the 7-wide local observation to the critic input; the MAT encoder ignores
            # that input (encode_state=False) and the buffer here is sized for the (W+2)-wide share_obs row
            raise NotImplementedError("use_centralized_V=False is not wired (the MAT policy never reads share_obs)")
        self.log_interval = getattr(a, "log_interval", 1)
        self.save_interval = getattr(a, "save_interval", 0)
        self.save_dir = config.get("save_dir")
        env = self.envs
        if env.N != self.n_rollout_threads or env.L != self.num_agents:
            raise ValueError("env batch (%d x %d) does not match n_rollout_threads x num_agents (%d x %d)"
                             % (env.N, env.L, self.n_rollout_threads, self.num_agents))
        # base_runner.py:198-230
        self.policy = TransformerPolicy(a, env.observation_space[0], env.share_observation_space[0], env.action_space[0],
                                        self.num_agents, device=self.device)
        self.policy.transformer.output_device = self.device          # keep actions / log-probs on the GPU
        self.policy.transformer.env_offset = env.env_offset
        self.trainer = MATTrainer(a, self.policy, self.num_agents, device=self.device)
        self.buffer = SharedReplayBuffer(a, self.num_agents, env.observation_space[0], env.share_observation_space[0],
                                         env.action_space[0], a.env_name, device=self.device)

    @property
    def eval_envs(self):
        """Evaluation envs (config["eval_envs"], else a second batch like the training one on another seed, built on first use)."""
        if self._eval_envs is None:
            e = self.envs
            self._eval_envs = type(e)(e.N, e.W, e.base_workload, seed=e.seed + 1, device=self.device, double_reset=e.double_reset,
                                      env_offset=e.env_offset, shannon=e.shannon, fixed=e.fixed, standalone=e.standalone)
        return self._eval_envs

    # ---- dcml_runner.py:126-142 ------------------------------------------------------------------------
    def warmup(self):
        obs, share_obs, ava = self.envs.reset_device()
        self.buffer.obs[0].copy_(obs)
        self.buffer._share_obs[0].copy_(self.envs.sob if self.use_centralized_V else obs[:, 0])
        self.buffer.available_actions[0].copy_(ava)

    # ---- dcml_runner.py:144-245 ("mat" branch) -----------------------------------------------------------
    @torch.no_grad()
    def collect(self, step):
        self.trainer.prep_rollout()
        b, N, L = self.buffer, self.n_rollout_threads, self.num_agents
        value, action, logp, _, _ = self.policy.get_actions(None, b.obs[step].reshape(N * L, -1), None, None, None,
                                                            b.available_actions[step].reshape(N * L, -1))
        return value.reshape(N, L, -1), action.reshape(N, L, -1), logp.reshape(N, L, -1), None, None

    # ---- dcml_runner.py:247-288 ----------------------------------------------------------------------------
    def insert(self, data):
        obs, share_row, rewards, dones_env, available_actions, values, actions, action_log_probs = data
        # every agent of an env finishes together (DCML_BID_FIRST_MA_ENV_SingleProcess.py:262-267), so
        # masks = 1 - done per env and active_masks stay 1 (dcml_runner.py:264-268)
        masks = 1.0 - dones_env.float()
        self.buffer.insert(share_row, obs, None, None, actions, action_log_probs, values, rewards, masks, None, None,
                           available_actions)

    # ---- base_runner.py:286-325 ------------------------------------------------------------------------------
    @torch.no_grad()
    def compute(self):
        self.trainer.prep_rollout()
        b = self.buffer
        next_values = self.policy.get_values(None, b.obs[-1].reshape(-1, b.obs.shape[-1]), None, None)
        b.compute_returns(next_values.reshape(b.value_preds[-1].shape), self.trainer.value_normalizer)

    # ---- base_runner.py:429-435 -------------------------------------------------------------------------------
    def train(self):
        self.trainer.prep_training()
        train_infos = self.trainer.train(self.buffer)
        self.buffer.after_update()
        return train_infos

    def save(self, episode):
        if self.save_dir is not None:
            self.policy.save(self.save_dir, episode)          # base_runner.py:437-449

    def log_train(self, train_infos, total_num_steps):
        train_infos = dict(train_infos)
        train_infos["average_step_rewards"] = float(self.buffer.rewards.mean())     # dcml_runner.py:300-303
        train_infos["total_num_steps"] = total_num_steps
        return train_infos

    # ---- dcml_runner.py:319-448 ("mat" branch) ----------------------------------------------------------------
    @torch.no_grad()
    def eval(self, total_num_steps, stride=2, eval_envs=None):
        """Deterministic evaluation: ``policy.act(deterministic=True, stride=stride)`` for ``total_num_steps`` steps of
        the evaluation envs (preset contexts via ``DcmlVecEnv.set_context`` reproduce the reference's benchmark
        mode).  Returns the reference's 4-tuple: mean reward / delay / payment over the episodes that finished,
        and the mean time of one batched decision in seconds (measured on the device)."""
        envs = self.eval_envs if eval_envs is None else eval_envs
        if envs is self.envs:
            # the reference evaluates on separate eval_envs (dcml_runner.py:319-338); resetting the training envs here would
            # leave buffer.obs[0] describing a context the envs no longer hold
            raise ValueError("eval() needs evaluation envs of its own, not the training envs")
        self.trainer.prep_rollout()
        N, L = envs.N, envs.L
        f64 = dict(dtype=torch.float64, device=self.device)
        one_r, one_d, one_p = torch.zeros(N, **f64), torch.zeros(N, **f64), torch.zeros(N, **f64)
        fin = torch.zeros(4, **f64)                  # finished episodes: count, reward, delay, payment sums
        obs, _, ava = envs.reset_device()
        t_ms, e0, e1 = 0.0, torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(total_num_steps):
            e0.record()
            actions, _ = self.policy.act(None, obs.reshape(N * L, -1), None, None, ava.reshape(N * L, -1), deterministic=True,
                                         stride=stride)
            e1.record()
            obs, _, rewards, dones, delay, payment, ava = envs.step_device(actions.reshape(N, L, 1))
            one_r += rewards.double(); one_d += delay.double(); one_p += payment.double()
            d = dones.bool()
            dm = d.double()
            fin += torch.stack([dm.sum(), (one_r * dm).sum(), (one_d * dm).sum(), (one_p * dm).sum()])
            one_r.masked_fill_(d, 0.0); one_d.masked_fill_(d, 0.0); one_p.masked_fill_(d, 0.0)
            e1.synchronize()
            t_ms += e0.elapsed_time(e1)
        cnt, r, dl, pay = fin.tolist()
        nan = float("nan")
        return ((r / cnt, dl / cnt, pay / cnt) if cnt > 0 else (nan, nan, nan)) + (t_ms * 1e-3 / max(total_num_steps, 1),)

    # ---- dcml_runner.py:22-124 -----------------------------------------------------------------------------
    def run(self, episodes=None, verbose=False):
        self.warmup()
        start = time.time()
        if episodes is None:
            episodes = int(self.num_env_steps) // self.episode_length // self.n_rollout_threads
        N = self.n_rollout_threads
        f64 = dict(dtype=torch.float64, device=self.device)
        ep_reward, ep_delay, ep_payment = torch.zeros(N, **f64), torch.zeros(N, **f64), torch.zeros(N, **f64)
        done_sum = torch.zeros(4, **f64)           # finished episodes: count, reward, delay, payment sums
        logs = []
        for episode in range(episodes):
            if self.use_linear_lr_decay:
                self.policy.lr_decay(episode, episodes)
            for step in range(self.episode_length):
                values, actions, action_log_probs, _, _ = self.collect(step)
                obs, _, rewards, dones, delay, payment, ava = self.envs.step_device(actions)
                d = dones.bool()
                ep_reward += rewards.double(); ep_delay += delay.double(); ep_payment += payment.double()
                dm = d.double()
                done_sum += torch.stack([dm.sum(), (ep_reward * dm).sum(), (ep_delay * dm).sum(), (ep_payment * dm).sum()])
                ep_reward.masked_fill_(d, 0.0); ep_delay.masked_fill_(d, 0.0); ep_payment.masked_fill_(d, 0.0)
                self.insert((obs, self.envs.sob if self.use_centralized_V else obs[:, 0], rewards, dones, ava, values, actions,
                             action_log_probs))
            self.compute()
            train_infos = self.train()
            total_num_steps = (episode + 1) * self.episode_length * N
            if self.save_interval and (episode % self.save_interval == 0 or episode == episodes - 1):
                self.save(episode)
            if episode % self.log_interval == 0:
                info = self.log_train(train_infos, total_num_steps)
                info["fps"] = int(total_num_steps / max(time.time() - start, 1e-9))
                cnt, r, dl, pay = done_sum.tolist()
                if cnt > 0:
                    info.update(aver_episode_rewards=r / cnt, aver_episode_delays=dl / cnt, aver_episode_payments=pay / cnt,
                                done_episodes=int(cnt))
                    done_sum.zero_()
                if verbose:
                    print("updates %d/%d episodes, total num timesteps %d, FPS %d, %s"
                          % (episode, episodes, total_num_steps, info["fps"], {k: (round(v, 5) if isinstance(v, float) else v)
                                                                                   for k, v in info.items()}))
                logs.append(info)
        return logs
