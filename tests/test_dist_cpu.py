"""N>1 host logic on CPU: two gloo ranks, each with half of the batch, must produce the statistics
the reference computes over the whole batch, and the reference arm of bench.py only
speaks on rank 0."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import argparse
        from mat_dcml_b200.mat_trainer import MATTrainer
        from oracle import ppo_oracle as po
        args = argparse.Namespace(n_rollout_threads=4, clip_param=0.2, ppo_epoch=1, num_mini_batch=1, value_loss_coef=1.0,
                                  entropy_coef=0.01, max_grad_norm=10.0, huber_delta=10.0, use_max_grad_norm=True,
                                  use_clipped_value_loss=True, use_huber_loss=True, use_valuenorm=True,
                                  use_value_active_masks=True, use_policy_active_masks=True)
        tr = MATTrainer(args, None, 5, device=torch.device("cpu"))
        rng = np.random.RandomState(0)
        ret = torch.from_numpy((rng.standard_normal((40, 1)) * 30 - 100).astype(np.float32))
        adv = torch.from_numpy(rng.standard_normal((8, 4, 5, 1)).astype(np.float32))
        act = torch.from_numpy((rng.uniform(size=(8, 4, 5, 1)) > 0.2).astype(np.float32))
        val = torch.from_numpy(rng.standard_normal((40, 1)).astype(np.float32))
        vpred = torch.from_numpy(rng.standard_normal((40, 1)).astype(np.float32))
        am = torch.from_numpy((rng.uniform(size=(40, 1)) > 0.2).astype(np.float32))
        half = slice(rank * 20, rank * 20 + 20)
        # value loss incl. ValueNorm.update over the global batch: the two shard losses add up to the full loss
        vl = tr.cal_value_loss(val[half], vpred[half], ret[half], am[half])
        tot = vl.clone()
        dist.all_reduce(tot)
        ref_vn = po.ValueNormState()
        ref_vn.update(ret)
        e1 = ref_vn.normalize(ret) - val
        e2 = ref_vn.normalize(ret) - (vpred + (val - vpred).clamp(-0.2, 0.2))
        want = (torch.max(po.huber(e1, 10.0), po.huber(e2, 10.0)) * am).sum() / am.sum()
        # the same with the ValueNorm moments of the epoch's minibatches reduced up front (what train() does): one collective
        tr2 = MATTrainer(args, None, 5, device=torch.device("cpu"))
        mom = tr2.minibatch_moments([(None,) * 6 + (ret[half],) + (None,) * 5])
        vl2 = tr2.cal_value_loss(val[half], vpred[half], ret[half], am[half], vn_moments=mom[0], numerator_only=True)
        cnt = am[half].sum()
        both = torch.stack([vl2, cnt])
        dist.all_reduce(both)
        assert abs(float(both[0] / both[1]) - float(want)) < 1e-5 * max(1.0, abs(float(want)))
        np.testing.assert_allclose(tr2.value_normalizer.stats.numpy(), ref_vn.stats(), rtol=1e-5)
        nadv = tr.normalize_advantages(adv[:, rank * 2:rank * 2 + 2], act[:, rank * 2:rank * 2 + 2])
        want_adv = po.normalize_advantages(adv.numpy(), torch.cat([act, act[:1]]).numpy())[:, rank * 2:rank * 2 + 2]
        q.put((rank, float(tot), float(want), tr.value_normalizer.stats.numpy().copy(), ref_vn.stats(),
               float(np.abs(nadv.numpy() - want_adv).max())))
    finally:
        dist.destroy_process_group()


def test_two_rank_statistics_match_global_batch():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, tot, want, stats, ref_stats, adv_err in out:
        assert abs(tot - want) < 1e-5 * max(1.0, abs(want)), (rank, tot, want)
        np.testing.assert_allclose(stats, ref_stats, rtol=1e-5)
        assert adv_err < 1e-5


def test_reference_arm_speaks_only_on_rank0():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""
