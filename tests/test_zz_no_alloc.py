"""The allocation-free act entry point (matas_policy_act_autoregressive_static).

It exists for hosts that may not allocate device memory once they are running (a decision server
whose GPU memory is owned by another framework's allocator, or a latency-critical loop that cannot
take the device-wide sync of cudaMalloc/cudaFree): its scratch is a fixed array inside the library.
This file is collected last on purpose, and the test also passes when run alone in a fresh process.
It calls the entry point through the C ABI twice in a row -- the second call finds the first call's
leftovers in the shared array -- and checks each result against the regular, allocating entry point and
against the CPU oracle (same named Philox draws for the stochastic call).
"""
import ctypes

import numpy as np
import pytest
import torch

from oracle import mat_oracle as mo
from oracle import philox as px
from test_policy_gpu import assert_actions

pytestmark = pytest.mark.gpu
TOL = 1e-4      # the log-prob tolerance of the other policy tests


def _act(fn, m, rep, ava, B, det, stride, step):
    from mat_dcml_b200 import _cabi
    L = m.n_agent
    actions = torch.full((B, L), -7.0, device="cuda")
    logp = torch.full((B, L), -7.0, device="cuda")
    _cabi.check(fn(m._h(), _cabi.ptr(rep), _cabi.ptr(ava), B, int(det), stride, 4242, step, 3, None, None,
                   _cabi.ptr(actions), _cabi.ptr(logp), _cabi.stream_ptr()))
    torch.cuda.synchronize()
    return actions.cpu().numpy(), logp.cpu().numpy()


def test_static_scratch_matches_allocating_entry_point(cuda_device):
    from mat_dcml_b200 import _cabi, MatasError
    from mat_dcml_b200.ma_transformer import MultiAgentTransformer
    lib = _cabi.lib()
    L = 21
    m = MultiAgentTransformer(L + 1, 7, 2, L, 2, 64, 2, device=torch.device("cuda"), action_type="Semi_Discrete")
    sd = mo.make_weights(41, L)
    m.load_state_dict(mo.full_state_dict(sd, L))
    m.sync_weights()
    rng = np.random.RandomState(5)
    # two calls in a row: different batch sizes, inputs and modes, so stale scratch would show
    for call, (B, det, stride) in enumerate(((90, False, 1), (37, True, 10))):
        obs = torch.from_numpy(rng.uniform(0, 1.1, size=(B, L, 7)).astype(np.float32)).cuda()
        ava = torch.ones(B, L, 2, device="cuda")
        ava[:, :-1, 1] = torch.from_numpy((rng.uniform(size=(B, L - 1)) > 0.3).astype(np.float32)).cuda()
        _, rep = m.encode(obs)
        rep = rep.contiguous()
        a_new, lp_new = _act(lib.matas_policy_act_autoregressive_static, m, rep, ava, B, det, stride, 10 + call)
        a_old, lp_old = _act(lib.matas_policy_act_autoregressive, m, rep, ava, B, det, stride, 10 + call)
        assert np.isfinite(a_new).all() and np.isfinite(lp_new).all() and not (a_new == -7.0).any()
        np.testing.assert_array_equal(a_new[:, :-1], a_old[:, :-1])
        np.testing.assert_allclose(a_new[:, -1], a_old[:, -1], atol=TOL)
        np.testing.assert_allclose(lp_new, lp_old, atol=TOL)
        # ... and against the oracle: the stochastic call with the kernel's own named draws (seed 4242, env offset 3)
        u = None if det else px.policy_uniforms(4242, 10 + call, B, L, env0=3)
        z = None if det else px.policy_normals(4242, 10 + call, B, L - 1, env0=3)
        _, ao, lpo = mo.get_actions(sd, obs.cpu().numpy().reshape(-1, 7), ava.cpu().numpy(), L, deterministic=det, stride=stride, u=u, z=z)
        assert_actions(a_new, ao.numpy(), lp_new, lpo.numpy(), L, u=u)
    # beyond the stated limit (2048 env x agent pairs) the call is refused, not redirected
    B = 2048 // L + 1
    rep = torch.zeros(B, L, 64, device="cuda")
    ava = torch.ones(B, L, 2, device="cuda")
    with pytest.raises(MatasError, match="static scratch limit"):
        _act(lib.matas_policy_act_autoregressive_static, m, rep, ava, B, True, 1, 0)
