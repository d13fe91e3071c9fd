#!/usr/bin/env python
"""Rollout throughput of the MAT-AS decision-and-rollout hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c3] [--no-extras]

One "step" is one rollout step over the whole batch of synthetic envs: policy decision for all
L agents (autoregressive decode), DCML env step incl. reset, buffer insert, and -- every
episode_length steps -- bootstrap values + GAE; the episode boundary is placed inside the timed region.
Prints ONE JSON line (see DESIGN.md "Measurement"): the headline workload (BASELINE configs[2], "c3") with
`roofline`, `cpu_baseline`, `e2e`; the other BASELINE shapes ride along under `workloads` (c2, c4 rollout, the c4
full PPO training step, the c5 agent-count sweep) and the PPO update under `ppo_update`, each with its own roofline.
Under torchrun every rank owns its own shard of envs (weak scaling, no data-path collective).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (envs per GPU, workers, episode_length, deterministic, stride, description)
    "c2": (1024, 20, 50, True, 10, "BASELINE configs[1]: deterministic batched decision-making, 1024 envs x 20 workers, stride 10"),
    "c3": (16384, 50, 50, False, 1, "BASELINE configs[2]: rollout, 16384 envs x 50 workers per GPU, stochastic decode"),
    "c4": (8192, 20, 50, False, 1, "BASELINE configs[3] rollout part: 8192 envs x 20 workers"),
    "c5_10": (4096, 10, 50, False, 1, "BASELINE configs[4]: 4096 envs x 10 workers"),
    "c5_20": (4096, 20, 50, False, 1, "BASELINE configs[4]: 4096 envs x 20 workers"),
    "c5_50": (4096, 50, 50, False, 1, "BASELINE configs[4]: 4096 envs x 50 workers"),
    "c5_100": (4096, 100, 50, False, 1, "BASELINE configs[4]: 4096 envs x 100 workers"),
    "c5_150": (4096, 150, 50, False, 1, "BASELINE configs[4]: 4096 envs x 150 workers"),
    "c5_200": (4096, 200, 50, False, 1, "BASELINE configs[4]: 4096 envs x 200 workers"),
}
EXTRAS = ("c2", "c4", "c5_10", "c5_20", "c5_50", "c5_100", "c5_150", "c5_200")
METRIC, UNIT = "rollout_env_steps_per_sec", "env-steps/s"


def decode_schedule_positions(L, det, stride):
    """Agent positions the decode kernel visits in one launch (mirror of DecSched in csrc/policy_fwd.h,
    Semi_Discrete mode): stochastic = every position once; deterministic = per stride block a speculative
    pass and, unless it was the last block, a commit pass (transformer_act.py:164-207)."""
    if not det:
        return list(range(L))
    nd, out = L - 1, []
    s, e, go = 0, 1, True
    while e <= L and go:
        if e == L:
            go = False
        out += list(range(s, e))                 # speculative pass
        if not go:
            break
        out += list(range(s + 1, e))             # commit pass
        if e < nd:
            s, e = e, min(e + stride, nd)
        else:
            s, e = e, min(e + 1, L)
    return out


def model_args(**over):
    ns = argparse.Namespace(algorithm_name="mat", env_name="DCML", lr=5e-5, opti_eps=1e-5, weight_decay=0,
                            use_policy_active_masks=True, n_block=2, n_embd=64, n_head=2, encode_state=False,
                            dec_actor=False, share_actor=False, hidden_size=64, recurrent_N=1, gamma=0.99,
                            gae_lambda=0.95, use_gae=True, use_popart=True, use_valuenorm=True,
                            use_proper_time_limits=False)
    for k, v in over.items():
        setattr(ns, k, v)
    return ns


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        return {}


def load_traffic():
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
    except (OSError, ValueError):
        return {}


# ----------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the oracle port of the reference's CPU path (torch-CPU policy that
# re-runs the decoder per agent like transformer_act.py:77-98 + single-process Python env loop)
# ----------------------------------------------------------------------------------------------
def cpu_rollout_sample(W, n_envs, steps, warmup, deterministic, stride, threads=None):
    import torch
    from oracle import env_oracle as eo, mat_oracle as mo, philox as px
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    L = W + 1
    sd = mo.make_weights(1, L, profile="init")
    env = eo.DcmlEnvOracle(n_envs, W, seed=1)
    obs, sob, ava = env.reset()
    times = []
    for t in range(warmup + steps):
        t0 = time.perf_counter()
        u = px.policy_uniforms(1, t, n_envs, L)
        z = px.policy_normals(1, t, n_envs, L - 1)
        v, a, lp = mo.get_actions(sd, obs.reshape(-1, 7), ava.reshape(-1, 2), L, deterministic=deterministic,
                                  stride=stride, u=u, z=z)
        obs, sob, rew, dones, infos, ava = env.step(a.numpy().reshape(n_envs, L, 1))
        if t >= warmup:
            times.append(time.perf_counter() - t0)
    sec = float(np.sum(times))
    return n_envs * steps / sec, sec / steps, cores


def run_reference(ns):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B, W, T, det, stride, desc = WORKLOADS[ns.workload]
    n_envs = ns.cpu_envs
    value, spp, cores = cpu_rollout_sample(W, n_envs, ns.steps, ns.warmup, det, stride)
    # the reference's own default is ONE intra-op thread (mat_src/mat/config.py:168 n_training_threads = 1)
    v1, spp1, _ = cpu_rollout_sample(W, n_envs, max(1, min(ns.steps, 2)), 1, det, stride, threads=1)
    sample = "%d envs x %d workers, %d timed rollout steps (policy decision + env step) of the same workload" % (n_envs, W, ns.steps)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": ns.gpus, "steps": ns.steps,
            "warmup": ns.warmup, "ms_per_step": spp * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": ns.workload + ": " + desc, "sample": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                             "one_thread": {"value": v1, "unit": UNIT, "cores": 1,
                                            "note": "the reference's default n_training_threads = 1 (mat_src/mat/config.py:168)"}},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ----------------------------------------------------------------------------------------------
# clocks sampler
# ----------------------------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append([x.strip() for x in ln.split(",")])

    def mark(self):
        return len(self.rows)

    def stop(self, start=0, end=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        rows = [r for r in self.rows[start:end] if len(r) >= 7] or [r for r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        f = lambda s: float(s) if s.replace(".", "", 1).isdigit() else float("nan")
        sm = [f(r[0]) for r in rows]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": f(rows[0][1]), "power_w_max": max(f(r[2]) for r in rows),
                "samples": len(rows), "reasons": reasons}


# ----------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------
class Ctx:
    """torch / distributed handles shared by the measurements of one process."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback "
                             "(use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.peaks, self.traffic = load_peaks(), load_traffic()
        self.fp32_peak = None

    def timed(self, step_fn, start_fn, K, Wm):
        """Wm untimed + K timed calls of step_fn; CUDA events on the current stream, barrier + synchronize on both sides,
        MAX over ranks.  Returns milliseconds for the K calls."""
        torch, dist = self.torch, self.dist
        start_fn()
        for i in range(Wm):
            step_fn(i)
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for i in range(K):
            step_fn(Wm + i)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=self.dev)
        if self.world > 1:
            dist.barrier()
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def ffma_peak(self):
        if self.fp32_peak is None:
            from mat_dcml_b200 import _cabi
            v = ctypes.c_double()
            _cabi.check(_cabi.lib().matas_fp32_peak(20000, ctypes.byref(v)))
            self.fp32_peak = v.value
        return self.fp32_peak


def decode_roofline(ctx, name, B, L, det, stride, dec_ms, n_timed, step_ms):
    """Roofline of the autoregressive decode kernel (the dominant kernel of a rollout step).

    HBM bound on the kernel's own byte model: at every position j of every pass the decode re-reads the j+1 cached
    256-byte layer-input rows of its two cached attentions (block-1 self and cross; block-0 self-attention runs on
    token-code counts and reads nothing, block-0 cross-attention gathers its keys from an L2-resident table), writes one
    new row per cached attention, reads 4 folded query rows + the obs_rep row (256 B each) and writes the outputs.
    SURVEY 8(d)'s own terms are printed beside it: its byte budget for a cache that never leaves the chip
    (52 L B per launch) and its KV-cached MAC count against the tensor peak / 3 (three fp16 products per GEMM)."""
    peaks = ctx.peaks
    visits = decode_schedule_positions(L, det, stride)
    n_cached = 2
    alg_bytes = B * (sum(n_cached * 256 * (j + 1) + n_cached * 256 + 5 * 256 + 16 for j in visits))
    mac = L * (86336 + 256 * (L + 1)) * B                      # KV-cached decoder MACs per launch
    hbm_peak = peaks.get("hbm_gbs", 6500.0)
    hbm_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6.5 TB/s (B200_PROFILING.md)"
    tensor_peak = peaks.get("bf16_tflops_sustained", 1400.0)
    tensor_src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained)" if peaks else "fallback 1.4 PFLOP/s sustained (B200_PROFILING.md)"
    sec = dec_ms * 1e-3
    achieved_gbs = alg_bytes / sec / 1e9
    tflops = 2.0 * mac / sec / 1e12
    traffic = ctx.traffic.get(name)
    bytes_8d = 52 * L * B
    fp32 = ctx.ffma_peak()
    return {"kernel": "decode_tc_kernel", "bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
            "frac": achieved_gbs / hbm_peak, "traffic": traffic, "peak_source": hbm_src, "launch_ms": dec_ms,
            "launches_timed": n_timed, "share_of_step": dec_ms / step_ms, "algorithmic_bytes_per_launch": alg_bytes,
            "cached_attentions": n_cached,
            "survey_8d": {"algorithmic_bytes_8d": bytes_8d,
                          "traffic_ratio": (traffic / bytes_8d) if traffic else alg_bytes / bytes_8d,
                          "traffic_ratio_of": "ncu dram bytes" if traffic else "the kernel's byte model (no ncu capture for this shape)",
                          "algorithmic_flops_per_launch": 2.0 * mac, "achieved_tflops": tflops,
                          "tensor_peak_third_tflops": tensor_peak / 3.0, "frac_tensor_third": tflops / (tensor_peak / 3.0),
                          "tensor_peak_source": tensor_src, "fp32_ffma_peak_tflops": fp32, "frac_fp32": tflops / fp32 if fp32 else None,
                          "note": "the dense stages run on tcgen05 as three fp16 products (hi/lo split, 1e-4 parity bar); the "
                                  "attention scans run on the CUDA cores at ~2 flop per cache byte, which is why the per-env "
                                  "cache lives in HBM and the kernel sits under the HBM roof, not the tensor roof"}}


def measure_rollout(ctx, name, steps, warmup, e2e_steps, ref_contract_steps=0):
    """Device-resident and end-to-end rollout throughput of one BASELINE shape on this rank's shard of envs."""
    torch = ctx.torch
    from mat_dcml_b200 import TransformerPolicy, DcmlVecEnv, SharedReplayBuffer, ValueNorm, _cabi
    dev, world, rank = ctx.dev, ctx.world, ctx.rank
    B, W, T, det, stride, desc = WORKLOADS[name]
    L = W + 1
    torch.manual_seed(1)
    env = DcmlVecEnv(B, W, seed=1, device=dev, env_offset=rank * B, host_dtype=np.float32, replicate_share_obs=False)
    pol = TransformerPolicy(model_args(), [7], [W + 2], env.action_space[0], L, device=dev)
    mdl = pol.transformer
    mdl.env_offset = rank * B
    buf = SharedReplayBuffer(model_args(episode_length=T, n_rollout_threads=B), L, [7], [W + 2], env.action_space[0],
                             "DCML", device=dev)
    vn = ValueNorm(1, device=dev)
    vn.update(torch.randn(256, 1, device=dev) * 30 - 100)
    launches = {"n": 0, "gae": 0}
    n_enc = 5                                # embed+QKV, then per block tensor-core attention + block kernel
    state, host = {}, {}

    def warm_start():
        obs, sob, ava = env.reset_device()
        state["obs"], state["ava"] = obs, ava
        buf.obs[0].copy_(obs); buf._share_obs[0].copy_(env.sob); buf.available_actions[0].copy_(ava)
        buf.step = 0

    def device_step(i):
        mdl.output_device = dev
        a, lp, v = mdl.get_actions(None, state["obs"], state["ava"], deterministic=det, stride=stride)
        obs, sob, rew, done, delay, pay, ava = env.step_device(a)
        buf.insert(env.sob, obs, None, None, a, lp, v, rew, 1.0 - done.float(), None, None, ava)
        launches["n"] += n_enc + 1 + 1          # encoder, decode, env step (the weight repack runs only when weights change)
        if buf.step == 0:                        # episode boundary: bootstrap + GAE (base_runner.py:286-325)
            nv = mdl.get_values(None, obs)
            buf.compute_returns(nv, vn)
            buf.after_update()
            launches["n"] += n_enc + 1          # encoder, GAE
            launches["gae"] += 1
        state["obs"], state["ava"] = obs, ava

    def e2e_start(e=env):
        obs, sob, ava = e.reset()
        host["obs"], host["ava"] = obs, ava
        buf.step = 0

    def make_e2e_step(e):
        def e2e_step(i):
            mdl.output_device = torch.device("cpu")
            v, a, lp, _, _ = pol.get_actions(None, host["obs"], None, None, None, host["ava"], deterministic=det, stride=stride)
            v = pol.transformer.to_host(v)
            obs, sob, rew, dones, infos, ava = e.step(a.numpy().reshape(B, L, 1))
            buf.insert(sob[:, 0, :], obs, None, None, a, lp, v, rew[:, 0, 0], 1.0 - dones[:, 0].astype(np.float32), None, None, ava)
            if buf.step == 0:
                nv = pol.get_values(None, obs, None, None)
                buf.compute_returns(nv, vn)
                buf.after_update()
            host["obs"], host["ava"] = obs, ava
        return e2e_step

    def boundary_inside(K):
        # put the episode boundary (bootstrap + GAE + after_update) in the middle of the K timed steps
        buf.step = (T - max(1, K // 2)) % T

    # device-resident arm, with per-launch events on the decode kernel
    _cabi.check(_cabi.lib().matas_policy_timing(mdl._h(), 1))
    ctx.timed(device_step, warm_start, 0, max(warmup, 3))             # untimed warm-up (allocations, L2 / clock ramp)
    _cabi.check(_cabi.lib().matas_policy_timing(mdl._h(), 1))          # reset counters
    launches["n"] = launches["gae"] = 0
    boundary_inside(steps)
    mark = ctx.clocks.mark() if ctx.clocks else 0
    ms = ctx.timed(device_step, lambda: None, steps, 0)
    mark_end = ctx.clocks.mark() if ctx.clocks else 0
    tms, tn = ctypes.c_double(), ctypes.c_long()
    _cabi.check(_cabi.lib().matas_policy_decode_time(mdl._h(), ctypes.byref(tms), ctypes.byref(tn)))
    _cabi.check(_cabi.lib().matas_policy_timing(mdl._h(), 0))
    out = {"workload": name + ": " + desc, "envs_per_gpu": B, "workers": W, "agents": L, "deterministic": det, "stride": stride,
           "value": world * B * steps / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / steps, "steps": steps,
           "gpu_launches": launches["n"], "episode_boundaries_in_timed_region": launches["gae"],
           "clock_marks": (mark, mark_end)}
    dec_ms = tms.value / max(tn.value, 1)
    out["roofline"] = decode_roofline(ctx, name, B, L, det, stride, dec_ms, tn.value, ms / steps)
    if e2e_steps > 0:
        K = max(1, min(steps, e2e_steps))
        step_fn = make_e2e_step(env)
        ctx.timed(step_fn, e2e_start, 0, 3)
        boundary_inside(K)
        e2e_ms = ctx.timed(step_fn, lambda: None, K, 0)
        # host -> device per step: policy inputs (obs, ava), env actions, and the buffer insert of host arrays
        # (share_obs row, obs, ava, actions, log-probs, values, reward, mask); device -> host: values, actions,
        # log-probs, and the env outputs (obs, ava, share_obs row, reward, done, delay, payment)
        h2d = B * L * (7 + 2) * 4 + B * L * 4 + (B * (W + 2) + B * L * (7 + 2) + 3 * B * L + 2 * B) * 4
        d2h = 3 * B * L * 4 + B * L * (7 + 2) * 4 + B * (W + 2) * 4 + B * 13
        out["e2e"] = {"value": world * B * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                      "steps": K, "ms_per_step": e2e_ms / K, "host_dtype": "float32", "replicate_share_obs": False,
                      "note": "host numpy arrays in and out of TransformerPolicy.get_actions / DcmlVecEnv.step / SharedReplayBuffer.insert; "
                              "fp32 observations and a broadcast view of the per-env share_obs row (the reference's own float64 + "
                              "replicated contract is timed under e2e_reference_contract)"}
    if ref_contract_steps > 0:
        # the reference's own host contract: float64 observations / rewards, int64 flags, L materialised copies of share_obs
        env64 = DcmlVecEnv(B, W, seed=1, device=dev, env_offset=rank * B, host_dtype=np.float64, replicate_share_obs=True)
        K = ref_contract_steps
        step_fn = make_e2e_step(env64)
        ctx.timed(step_fn, lambda: e2e_start(env64), 0, 2)
        e2e_ms = ctx.timed(step_fn, lambda: None, K, 0)
        h2d = B * L * 7 * 8 + B * L * 2 * 8 + B * L * 4 + (B * (W + 2) * 8 + B * L * 7 * 8 + B * L * 2 * 8 + 3 * B * L * 4 + 2 * B * 8)
        d2h = 3 * B * L * 4 + B * L * 7 * 8 + B * L * 2 * 8 + B * L * (W + 2) * 8 + B * L * 8 + B * L + B * 8
        out["e2e_reference_contract"] = {"value": world * B * K / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / K, "steps": K,
                                         "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "host_dtype": "float64",
                                         "replicate_share_obs": True}
        env64.close()
    env.close()
    del buf, pol, mdl
    torch.cuda.empty_cache()
    return out


def ppo_block(ctx, full_step=True):
    """PPO update samples/sec and the full PPO training step of BASELINE configs[3] at its per-GPU share
    (8192 envs x 20 workers x T=50 over 8 GPUs = 1024 envs per GPU; weak scaling in the number of GPUs).

    ppo_update: ONE epoch (get_values + GAE + advantage norm + 4 minibatches of forward, backward, gradient all-reduce,
    clip, Adam); a row is one (env, t) sample carrying all 21 agents (shared_buffer.py:248).
    full_step : rollout of T = 50 steps + Runner.compute + MATTrainer.train with ppo_epoch = 15 x 4 minibatches
    (dcml_runner.py:43-85, base_runner.py:286-325,429-435, mat_trainer.py:158-217) as one number."""
    torch, dist, dev, world, rank = ctx.torch, ctx.dist, ctx.dev, ctx.world, ctx.rank
    from mat_dcml_b200 import TransformerPolicy, DcmlVecEnv, SharedReplayBuffer, MATTrainer
    N, W, T = 1024, 20, 50
    L = W + 1
    a = model_args(episode_length=T, n_rollout_threads=N, clip_param=0.2, ppo_epoch=1, num_mini_batch=4,
                   value_loss_coef=1.0, entropy_coef=0.01, max_grad_norm=10.0, huber_delta=10.0, use_max_grad_norm=True,
                   use_clipped_value_loss=True, use_huber_loss=True, use_value_active_masks=True)
    env = DcmlVecEnv(N, W, seed=2, device=dev, env_offset=rank * N)
    pol = TransformerPolicy(a, [7], [W + 2], env.action_space[0], L, device=dev)
    pol.transformer.output_device = dev
    pol.transformer.env_offset = rank * N
    tr = MATTrainer(a, pol, L, device=dev)
    buf = SharedReplayBuffer(a, L, [7], [W + 2], env.action_space[0], "DCML", device=dev)
    st = {}

    def rollout():
        obs, ava = st["obs"], st["ava"]
        for t in range(T):
            act, lp, v = pol.transformer.get_actions(None, obs, ava)
            obs, sob, rew, done, delay, pay, ava = env.step_device(act)
            buf.insert(env.sob, obs, None, None, act, lp, v, rew, 1.0 - done.float(), None, None, ava)
        st["obs"], st["ava"] = obs, ava

    obs, sob, ava = env.reset_device()
    st["obs"], st["ava"] = obs, ava
    buf.obs[0].copy_(obs); buf._share_obs[0].copy_(env.sob); buf.available_actions[0].copy_(ava)
    rollout()
    tr.train(buf)                                   # warm-up epoch (allocations)
    info = {}
    ms = ctx.timed(lambda i: info.update(tr.train(buf)), lambda: None, 1, 0)
    rows = world * N * T
    mac_row = L * (140096 + 768 * L)                # evaluate forward MACs per row; backward = 2 x forward
    # forward + backward of ONE minibatch alone (no collective, no optimizer): the kernels the update is made of
    M = N * T // 4
    sample = next(iter(buf.feed_forward_generator_transformer(buf.advantages, 4)))
    mdl = pol.transformer

    def fwd_bwd(i):
        mdl.flat_grad(zero=True)
        lp, v, ent = mdl(None, sample[1], sample[4], sample[11])
        (lp.mean() + v.mean() + ent.mean()).backward()
    fb_ms = ctx.timed(fwd_bwd, lambda: None, 5, 2) / 5
    tflops = 3 * 2.0 * mac_row * M / (fb_ms * 1e-3) / 1e12
    tensor_peak = ctx.peaks.get("bf16_tflops_sustained", 1400.0)
    fp32 = ctx.ffma_peak()
    out = {"rows_per_s": rows / (ms * 1e-3), "agent_samples_per_s": rows * L / (ms * 1e-3), "ms_per_epoch": ms,
           "rows_per_epoch": rows, "minibatches": 4, "envs_per_gpu": N, "workers": W, "T": T,
           "includes": "get_values + GAE + advantage norm + 4 x (forward, backward, grad all-reduce, clip, Adam)",
           "value_loss": info["value_loss"], "grad_norm": info["actor_grad_norm"],
           "roofline": {"kernel": "evaluate forward + backward of one minibatch (%d rows x %d agents): the tcgen05 token-parallel "
                                  "kernels of policy_tc.cu / attn_tc.cu and the backward of policy_bwd.cu, bwd_*_tc.cu, attn_bwd_tile.cu"
                                  % (M, L),
                        "bound": "tensor", "achieved": tflops, "peak": tensor_peak / 3.0, "unit": "TFLOP/s",
                        "frac": tflops / (tensor_peak / 3.0), "traffic": ctx.traffic.get("ppo_fwd_bwd"),
                        "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained / 3 (every GEMM is three fp16 products)" if ctx.peaks
                                       else "fallback 1.4 PFLOP/s / 3 (B200_PROFILING.md)",
                        "launch_ms": fb_ms, "algorithmic_flops": 3 * 2.0 * mac_row * M, "fp32_ffma_peak_tflops": fp32,
                        "frac_fp32": tflops / fp32 if fp32 else None,
                        "note": "SURVEY 8(d): L (140,096 + 768 L) MAC per row forward, backward = 2 x forward; measured with CUDA "
                                "events over 5 repetitions on the launching stream"}}
    if full_step:
        tr.ppo_epoch = 15

        def full(i):
            rollout()
            nv = pol.get_values(None, buf.obs[-1].reshape(-1, 7), None, None)
            buf.compute_returns(nv.reshape(buf.value_preds[-1].shape), tr.value_normalizer)
            tr.train(buf)
            buf.after_update()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        full(0)                                     # warm-up iteration
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0.record(); rollout(); e1.record()
        nv = pol.get_values(None, buf.obs[-1].reshape(-1, 7), None, None)
        buf.compute_returns(nv.reshape(buf.value_preds[-1].shape), tr.value_normalizer)
        tr.train(buf); buf.after_update()
        e2.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e2), e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tot, roll = t.tolist()
        out["full_step"] = {"workload": "BASELINE configs[3]: full PPO training step, %d envs x %d workers per GPU (8192 over 8 GPUs), T = %d, "
                                        "15 epochs x 4 minibatches" % (N, W, T),
                            "value": world * N * T / (tot * 1e-3), "unit": UNIT, "ms_per_training_step": tot, "rollout_ms": roll,
                            "compute_and_update_ms": tot - roll, "update_rows_per_s": 15 * rows / ((tot - roll) * 1e-3)}
    env.close()
    return out


def run_ours(ns):
    ctx = Ctx()
    torch, dist, world, rank = ctx.torch, ctx.dist, ctx.world, ctx.rank
    ctx.clocks = Clocks(ctx.local) if rank == 0 else None
    B, W, T, det, stride, desc = WORKLOADS[ns.workload]
    head = measure_rollout(ctx, ns.workload, ns.steps, ns.warmup, ns.e2e_steps,
                           ref_contract_steps=3 if (ns.extras and world == 1) else 0)
    clk = ctx.clocks.stop(*head.pop("clock_marks")) if ctx.clocks else None
    # ---- secondary metric: PPO update rows/s + the full PPO training step (BASELINE configs[3] shape per GPU) ----
    ppo = ppo_block(ctx, full_step=ns.extras) if ns.ppo else None
    # ---- the other BASELINE shapes (short runs; every rank takes part: max over ranks like the headline) ----------
    others = {}
    if ns.extras:
        for name in EXTRAS:
            if name == ns.workload:
                continue
            r = measure_rollout(ctx, name, ns.extra_steps, 3, min(3, ns.e2e_steps))
            r.pop("clock_marks")
            others[name] = r
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    cpu = None
    if world == 1:
        cv, cspp, cores = cpu_rollout_sample(W, ns.cpu_envs, 3, 2, det, stride)
        cpu = {"value": cv, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "%d envs x %d workers, 3 timed rollout steps (after 2 warm-up) of the same workload on the host (oracle port of "
                         "the reference: torch-CPU policy without KV cache + single-process env loop)" % (ns.cpu_envs, W)}
        if ns.extras:
            c1, _, _ = cpu_rollout_sample(W, ns.cpu_envs, 1, 1, det, stride, threads=1)
            cpu["one_thread"] = {"value": c1, "unit": UNIT, "cores": 1,
                                 "note": "the reference's default n_training_threads = 1 (mat_src/mat/config.py:168); 1 timed step"}
    line = {"metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": ns.steps, "warmup": max(ns.warmup, 3),
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": head["workload"], "envs_per_gpu": B, "workers": W, "agents": W + 1,
                       "episode_length": T, "deterministic": det, "stride": stride,
                       "episode_boundaries_in_timed_region": head["episode_boundaries_in_timed_region"],
                       "e2e_host_dtype": "float32", "e2e_replicate_share_obs": False,
                       "l2": "inputs larger than L2: K/V cache scratch + token buffers + rollout buffer are GBs per step"},
            "clocks": clk, "gpu_launches": head["gpu_launches"], "e2e": head.get("e2e"),
            "e2e_reference_contract": head.get("e2e_reference_contract"),
            "roofline": head["roofline"], "cpu_baseline": cpu, "ppo_update": ppo, "workloads": others}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


_STDOUT_FD = None


def emit(line):
    """The ONE JSON line of the contract, written to the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _STDOUT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_STDOUT_FD, data)


def main():
    # Libraries print to stdout behind our back (NCCL writes its version banner there when the communicator is
    # created): point file descriptor 1 at stderr for the whole run and keep the original for the JSON line.
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-envs", type=int, default=64, help="envs in the bounded CPU sample")
    ap.add_argument("--e2e-steps", type=int, default=25)
    ap.add_argument("--extra-steps", type=int, default=10, help="timed steps of each of the other BASELINE shapes")
    ap.add_argument("--no-ppo", dest="ppo", action="store_false", help="skip the PPO-update block")
    ap.add_argument("--no-extras", dest="extras", action="store_false",
                    help="headline workload only (no other shapes, no full training step, no reference-contract e2e)")
    ns = ap.parse_args()
    if ns.impl == "reference":
        run_reference(ns)
    else:
        run_ours(ns)


if __name__ == "__main__":
    main()
