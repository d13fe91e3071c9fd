"""CPU checks: the C-ABI library builds, loads and exports every symbol include/matas_b200.h declares;
host-side mirror of the reference interface (module tree, state_dict, shapes, error behaviour)."""
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT


@pytest.fixture(scope="module")
def lib():
    from mat_dcml_b200 import build, _cabi
    build.build()
    return _cabi.lib()


def _declared():
    src = open(os.path.join(ROOT, "include", "matas_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b((?:matas|dcml)_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported(lib):
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), "libmatas_b200.so does not export %s" % n
    from mat_dcml_b200 import _cabi
    assert set(_cabi.exported_symbols()) == set(names), "ctypes table and header disagree"


def test_header_cites_reference_for_every_entry_point():
    src = open(os.path.join(ROOT, "include", "matas_b200.h")).read()
    for fn in _declared():
        if fn in ("matas_last_error", "matas_version", "matas_policy_destroy", "dcml_env_destroy",
                  "matas_policy_param_count"):
            continue
        i = src.index(fn + "(")
        block = src[src.rfind("/*", 0, i):i]
        assert re.search(r"\.py:\d+", block) or "no counterpart in the reference" in block, fn


def test_no_device_calls_fail_loudly(lib):
    """Without a GPU the product path must raise, never fall back."""
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mat_dcml_b200 import MatasError
    from mat_dcml_b200.ma_transformer import MultiAgentTransformer
    from mat_dcml_b200.dcml_env import DcmlVecEnv
    m = MultiAgentTransformer(22, 7, 2, 21, 2, 64, 2, action_type="Semi_Discrete")
    with pytest.raises(MatasError):
        m.get_values(None, np.zeros((21, 7), np.float32))
    with pytest.raises(MatasError):
        DcmlVecEnv(4, 20, device="cpu")


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "mat_dcml_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+\.*oracle", txt, flags=re.M), f


def test_module_matches_reference_state_dict_layout():
    from mat_dcml_b200.ma_transformer import MultiAgentTransformer
    from oracle import mat_oracle as mo
    for L, semi in ((101, True), (21, True), (13, False)):
        m = MultiAgentTransformer(L + 1, 7, 2, L, 2, 64, 2, action_type="Semi_Discrete" if semi else "Discrete")
        names = [n for n, _ in m.named_parameters()]
        want = mo.param_shapes(L, 7, L + 1, semi=semi)
        assert names == [n for n, _ in want]
        for (n, p), (_, s) in zip(m.named_parameters(), want):
            assert tuple(p.shape) == tuple(s), n
        sd = m.state_dict()
        assert len(sd) == len(want) + 6 and all(k in sd for k in mo.mask_names())
        w = mo.make_weights(3, L, semi=semi)
        m.load_state_dict(mo.full_state_dict(w, L), strict=True)
        flat = torch.cat([w[n].reshape(-1) for n, _ in want])
        assert torch.equal(m.flat_parameters(), flat)      # parameters are views into the flat buffer
        g = m.flat_grad()
        g.fill_(2.0)
        assert all(float(p.grad.min()) == 2.0 for p in m.parameters())
    assert sum(p.numel() for p in MultiAgentTransformer(102, 7, 2, 101, 2, 64, 2, action_type="Semi_Discrete").parameters()) == 151469


def test_policy_constructor_contract():
    import argparse
    from mat_dcml_b200 import TransformerPolicy, Action_Space
    args = argparse.Namespace(algorithm_name="mat", lr=5e-4, opti_eps=1e-5, weight_decay=0, use_policy_active_masks=True,
                              n_block=2, n_embd=64, n_head=2, encode_state=False, dec_actor=False, share_actor=False,
                              env_name="DCML")
    pol = TransformerPolicy(args, [7], [22], Action_Space(2, semi_index=-1, extra=True), 21)
    assert pol.action_type == "Semi_Discrete" and pol.act_dim == 2 and pol.obs_dim == 7 and pol.share_obs_dim == 22
    assert isinstance(pol.optimizer, torch.optim.Adam) and hasattr(pol.transformer, "parameters")
    pol.lr_decay(5, 10)
    assert abs(pol.optimizer.param_groups[0]["lr"] - 2.5e-4) < 1e-12
    assert TransformerPolicy(args, [7], [22], Action_Space(2, semi_index=0), 21).action_type == "Discrete"
    args.algorithm_name = "mat_gru"
    with pytest.raises(NotImplementedError):
        TransformerPolicy(args, [7], [22], Action_Space(2, semi_index=-1), 21)
