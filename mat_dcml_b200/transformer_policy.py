"""Drop-in for the reference's ``TransformerPolicy``
(mat_src/mat/algorithms/mat/algorithm/transformer_policy.py:9-255): same constructor, methods,
argument meaning and return shapes; the model behind it is the CUDA-backed
``mat_dcml_b200.ma_transformer.MultiAgentTransformer``.
"""
import os

import numpy as np
import torch

from .ma_transformer import MultiAgentTransformer


def _space_shape(space):
    # mat/utils/util.py:28-36
    name = space.__class__.__name__
    if name == "Box":
        return space.shape
    if name == "list":
        return space
    raise NotImplementedError


def _check(x):
    return torch.from_numpy(x) if isinstance(x, np.ndarray) else x


class TransformerPolicy:
    def __init__(self, args, obs_space, cent_obs_space, act_space, num_agents, device=torch.device("cpu")):
        self.device = torch.device(device)
        self.algorithm_name = args.algorithm_name
        self.lr, self.opti_eps, self.weight_decay = args.lr, args.opti_eps, args.weight_decay
        self._use_policy_active_masks = args.use_policy_active_masks
        semi_index = None
        cls = act_space.__class__.__name__
        if cls == "Action_Space":                                    # transformer_policy.py:29-35
            if act_space.semi_index != 0:
                self.action_type, semi_index = "Semi_Discrete", act_space.semi_index
            else:
                self.action_type = "Discrete"
        elif cls in ("Box", "Available_Continous_Space"):
            raise NotImplementedError("continuous act paths are outside the MAT-AS hot path (SURVEY.md 2.1 row 3)")
        else:
            self.action_type = "Discrete"
        if self.algorithm_name not in ("mat", "mat_dec"):
            raise NotImplementedError("algorithm_name %r: only 'mat' is wired for DCML "
                                      "(mat_src/mat/runner/shared/base_runner.py:198)" % self.algorithm_name)
        self.obs_dim = _space_shape(obs_space)[0]
        self.share_obs_dim = _space_shape(cent_obs_space)[0]
        self.act_dim = act_space.n
        self.act_discrete_dim = self.act_dim
        self.act_num = self.act_output_num = self.act_prob_dim = 1
        self.num_agents = num_agents
        self.tpdv = dict(dtype=torch.float32, device=self.device)
        self.transformer = MultiAgentTransformer(self.share_obs_dim, self.obs_dim, self.act_dim, num_agents,
                                                 n_block=args.n_block, n_embd=args.n_embd, n_head=args.n_head,
                                                 encode_state=args.encode_state, device=self.device,
                                                 action_type=self.action_type, dec_actor=args.dec_actor,
                                                 share_actor=args.share_actor,
                                                 semi_index=-1 if semi_index is None else semi_index)
        if getattr(args, "env_name", None) == "hands":
            self.transformer.zero_std()
        if self.device.type == "cuda":
            # same update as the reference's torch.optim.Adam (transformer_policy.py:103-105), fused with the
            # gradient clipping over the flat parameter buffer (flat_adam.py, csrc/adam.cu)
            from .flat_adam import FlatAdam
            self.optimizer = FlatAdam(self.transformer, lr=self.lr, eps=self.opti_eps, weight_decay=self.weight_decay)
        else:
            # a CPU-resident policy only holds parameters (checkpoint conversion): nothing here can produce gradients
            self.optimizer = torch.optim.Adam(self.transformer.parameters(), lr=self.lr, eps=self.opti_eps,
                                              weight_decay=self.weight_decay)

    def lr_decay(self, episode, episodes):
        lr = self.lr - (self.lr * (episode / float(episodes)))       # mat/utils/util.py:17-21
        for g in self.optimizer.param_groups:
            g["lr"] = lr

    def get_actions(self, cent_obs, obs, rnn_states_actor, rnn_states_critic, masks, available_actions=None,
                    deterministic=False, stride=1, n_objective=1):
        obs = _check(obs).reshape(-1, self.num_agents, self.obs_dim)
        if available_actions is not None:
            available_actions = _check(available_actions).reshape(-1, self.num_agents, self.act_dim)
        actions, action_log_probs, values = self.transformer.get_actions(None, obs, available_actions, deterministic,
                                                                        stride=stride)
        actions = actions.view(-1, self.act_output_num)
        action_log_probs = action_log_probs.view(-1, self.act_prob_dim)
        values = values.view(-1, n_objective)
        if rnn_states_actor is not None:
            rnn_states_actor = _check(rnn_states_actor).to(**self.tpdv)
        if rnn_states_critic is not None:
            rnn_states_critic = _check(rnn_states_critic).to(**self.tpdv)
        return values, actions, action_log_probs, rnn_states_actor, rnn_states_critic

    def get_values(self, cent_obs, obs, rnn_states_critic, masks, available_actions=None, n_objective=1):
        obs = _check(obs).reshape(-1, self.num_agents, self.obs_dim)
        return self.transformer.get_values(None, obs, available_actions).view(-1, n_objective)

    def evaluate_actions(self, cent_obs, obs, rnn_states_actor, rnn_states_critic, actions, masks,
                         available_actions=None, active_masks=None, n_objective=1, active_denominator=None):
        """transformer_policy.py:181-217.  ``active_denominator`` (optional, not in the reference) replaces
        ``active_masks.sum()`` by the global count when the minibatch is sharded over ranks."""
        obs = _check(obs).reshape(-1, self.num_agents, self.obs_dim)
        actions = _check(actions).reshape(-1, self.num_agents, self.act_output_num)
        if available_actions is not None:
            available_actions = _check(available_actions).reshape(-1, self.num_agents, self.act_dim)
        action_log_probs, values, entropy = self.transformer(None, obs, actions, available_actions)
        action_log_probs = action_log_probs.view(-1, self.act_prob_dim)
        entropy = entropy.view(-1, self.act_prob_dim)
        values = values.view(-1, n_objective)
        if self._use_policy_active_masks and active_masks is not None:
            am = _check(active_masks).to(**self.tpdv)
            entropy = (entropy * am).sum() / (am.sum() if active_denominator is None else active_denominator)
        else:
            entropy = entropy.mean()
        return values, action_log_probs, entropy

    def act(self, cent_obs, obs, rnn_states_actor, masks, available_actions=None, deterministic=True, stride=2):
        rnn_states_critic = None if rnn_states_actor is None else np.zeros_like(rnn_states_actor)
        _, actions, _, rnn_states_actor, _ = self.get_actions(cent_obs, obs, rnn_states_actor, rnn_states_critic, masks,
                                                              available_actions, deterministic, stride=stride)
        return actions, rnn_states_actor

    def save(self, save_dir, episode):
        torch.save(self.transformer.state_dict(), str(save_dir) + "/transformer_" + str(episode) + ".pt")

    def restore(self, model_dir):
        self.transformer.load_state_dict(torch.load(model_dir, map_location=self.device))

    def train(self):
        self.transformer.train()

    def eval(self):
        self.transformer.eval()
