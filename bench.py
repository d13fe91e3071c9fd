#!/usr/bin/env python
"""Rollout throughput of the MAT-AS decision-and-rollout hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c3]

One "step" is one rollout step over the whole batch of synthetic envs: policy decision for all
L agents (stochastic autoregressive decode), DCML env step incl. reset, buffer insert, and -- every
episode_length steps -- bootstrap values + GAE.  Prints ONE JSON line (see DESIGN.md "Measurement").
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
    "c5_100": (4096, 100, 50, False, 1, "BASELINE configs[4]: 4096 envs x 100 workers"),
    "c5_200": (4096, 200, 50, False, 1, "BASELINE configs[4]: 4096 envs x 200 workers"),
}
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


# ----------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the oracle port of the reference's CPU path (torch-CPU policy that
# re-runs the decoder per agent like transformer_act.py:77-98 + single-process Python env loop)
# ----------------------------------------------------------------------------------------------
def cpu_rollout_sample(W, n_envs, steps, warmup, deterministic, stride):
    import torch
    from oracle import env_oracle as eo, mat_oracle as mo, philox as px
    cores = os.cpu_count() or 1
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
    sample = "%d envs x %d workers, %d timed rollout steps (policy decision + env step) of the same workload" % (n_envs, W, ns.steps)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": ns.gpus, "steps": ns.steps,
            "warmup": ns.warmup, "ms_per_step": spp * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": ns.workload + ": " + desc, "sample": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
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

    def stop(self, start=0):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        rows = [r for r in self.rows[start:] if len(r) >= 7] or [r for r in self.rows if len(r) >= 7]
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
def run_ours(ns):
    import torch
    import torch.distributed as dist
    from mat_dcml_b200 import TransformerPolicy, DcmlVecEnv, SharedReplayBuffer, ValueNorm, _cabi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, W, T, det, stride, desc = WORKLOADS[ns.workload]
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
    launches = {"n": 0}

    # ---- device-resident rollout step (inputs already in HBM) --------------------------------
    state = {}

    def warm_start():
        obs, sob, ava = env.reset_device()
        state["obs"], state["ava"] = obs, ava
        buf.obs[0].copy_(obs); buf._share_obs[0].copy_(env.sob); buf.available_actions[0].copy_(ava)
        buf.step = 0

    n_enc = 5 if L <= 64 else 3             # embed+QKV, then per block [tiled attention (n_agent <= 64)] + block kernel

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
        state["obs"], state["ava"] = obs, ava

    # ---- end-to-end step through the reference-facing API with HOST arrays -------------------
    host = {}

    def e2e_start():
        obs, sob, ava = env.reset()
        host["obs"], host["ava"] = obs, ava
        buf.step = 0

    def e2e_step(i):
        mdl.output_device = torch.device("cpu")
        v, a, lp, _, _ = pol.get_actions(None, host["obs"], None, None, None, host["ava"], deterministic=det, stride=stride)
        v = v.cpu()
        obs, sob, rew, dones, infos, ava = env.step(a.numpy().reshape(B, L, 1))
        buf.insert(sob[:, 0, :], obs, None, None, a, lp, v, rew[:, 0, 0], 1.0 - dones[:, 0].astype(np.float32), None, None, ava)
        if buf.step == 0:
            nv = pol.get_values(None, obs, None, None)
            buf.compute_returns(nv, vn)
            buf.after_update()
        host["obs"], host["ava"] = obs, ava

    def timed(step_fn, start_fn, K, Wm):
        start_fn()
        for i in range(Wm):
            step_fn(i)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for i in range(K):
            step_fn(Wm + i)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.barrier()
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    clocks = Clocks(local) if rank == 0 else None
    # device-resident arm, with per-launch events on the decode kernel
    _cabi.check(_cabi.lib().matas_policy_timing(mdl._h(), 1))
    launches["n"] = 0
    timed(device_step, warm_start, 0, max(ns.warmup, 3))          # untimed warm-up (allocations, L2/clock ramp)
    _cabi.check(_cabi.lib().matas_policy_timing(mdl._h(), 1))      # reset counters
    launches["n"] = 0
    mark = clocks.mark() if clocks else 0
    ms = timed(device_step, lambda: None, ns.steps, 0)
    tms, tn = ctypes.c_double(), ctypes.c_long()
    _cabi.check(_cabi.lib().matas_policy_decode_time(mdl._h(), ctypes.byref(tms), ctypes.byref(tn)))
    _cabi.check(_cabi.lib().matas_policy_timing(mdl._h(), 0))
    clk = clocks.stop(mark) if clocks else None
    n_launch = launches["n"]
    value = world * B * ns.steps / (ms * 1e-3)
    # end-to-end arm
    e2e_K = max(1, min(ns.steps, ns.e2e_steps))
    e2e_ms = timed(e2e_step, e2e_start, e2e_K, 3)
    e2e_value = world * B * e2e_K / (e2e_ms * 1e-3)
    # host -> device per step: policy inputs (obs, ava), env actions, and the buffer insert of host arrays
    # (share_obs row, obs, ava, actions, log-probs, values, reward, mask); device -> host: values, actions,
    # log-probs, and the env outputs (obs, ava, share_obs row, reward, done, delay, payment)
    h2d = B * L * (7 + 2) * 4 + B * L * 4 + (B * (W + 2) + B * L * (7 + 2) + 3 * B * L + 2 * B) * 4
    d2h = 3 * B * L * 4 + B * L * (7 + 2) * 4 + B * (W + 2) * 4 + B * 13

    # ---- secondary metric: PPO update rows/s (BASELINE configs[3] shape per GPU) ----------------
    ppo = None
    if ns.ppo:
        ppo = ppo_block(ns, torch, dist, dev, world, rank)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (the autoregressive decode) -------------------------
    dec_ms = tms.value / max(tn.value, 1)
    mac = L * (86336 + 256 * (L + 1)) * B                      # KV-cached decoder MACs per launch
    n_steps_sched = None
    achieved = 2.0 * mac / (dec_ms * 1e-3) / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    tensor_peak = peaks.get("bf16_tflops_sustained", 1400.0)
    peak_src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained)" if peaks else "fallback 1.4 PFLOP/s sustained (B200_PROFILING.md)"
    fp32 = ctypes.c_double()
    _cabi.check(_cabi.lib().matas_fp32_peak(20000, ctypes.byref(fp32)))
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json"))).get(ns.workload)
    except (OSError, ValueError):
        pass
    # HBM bound: at every position j of every pass the decode re-reads the j+1 cached 256-byte layer-input rows
    # of its cached attentions, writes one new row per cached attention, reads 4 folded query rows + the obs_rep row
    # (256 B each) and writes the outputs; arithmetic intensity is ~7 flop/byte, below the fp32 ridge, so HBM is the
    # roof.  Cached attentions: block-1 self and cross (block-0 self-attention runs on token-code counts and reads
    # nothing; block-0 cross-attention gathers its keys from shared L2-resident tables, decode_tc.cu: x1_table_kernel /
    # x1_spec_table_kernel, computed once per weight update); with MATAS_DECODE_TAB=0 block-0 cross is cached per env too.
    visits = decode_schedule_positions(L, det, stride)
    tab = os.environ.get("MATAS_DECODE_TAB", "1") != "0" and os.environ.get("MATAS_TC", "1") != "0"
    n_cached = 2 if tab else 3
    alg_bytes = B * (sum(n_cached * 256 * (j + 1) + n_cached * 256 + 5 * 256 + 16 for j in visits))
    hbm_peak = peaks.get("hbm_gbs", 6500.0)
    hbm_src = ("measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6.5 TB/s (B200_PROFILING.md)")
    achieved_gbs = alg_bytes / (dec_ms * 1e-3) / 1e9
    roofline = {"kernel": "decode_tc_kernel", "bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved_gbs / hbm_peak, "traffic": traffic, "peak_source": hbm_src,
                "launch_ms": dec_ms, "launches_timed": tn.value, "share_of_step": dec_ms / (ms / ns.steps),
                "algorithmic_bytes_per_launch": alg_bytes, "cached_attentions": n_cached,
                "compute": {"algorithmic_flops_per_launch": 2.0 * mac, "achieved_tflops": achieved,
                            "fp32_ffma_peak_tflops": fp32.value, "frac_fp32": achieved / fp32.value if fp32.value else None,
                            "bf16_tensor_peak_tflops": tensor_peak, "tensor_peak_source": peak_src,
                            "note": "projections run as fp32 FFMA (1e-4 parity bar); ~7 flop/byte puts the kernel under "
                                    "the HBM roof for both the fp32 and the tensor-core ridge"}}
    cpu = None
    if world == 1:
        cv, cspp, cores = cpu_rollout_sample(W, ns.cpu_envs, 3, 2, det, stride)
        cpu = {"value": cv, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "%d envs x %d workers, 3 timed rollout steps (after 2 warm-up) of the same workload on the host (oracle port of "
                         "the reference: torch-CPU policy without KV cache + single-process env loop)" % (ns.cpu_envs, W)}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": ns.steps, "warmup": max(ns.warmup, 3),
            "ms_per_step": ms / ns.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": ns.workload + ": " + desc, "envs_per_gpu": B, "workers": W, "agents": L,
                       "episode_length": T, "deterministic": det, "stride": stride,
                       "l2": "inputs larger than L2: K/V cache scratch + token buffers + rollout buffer are GBs per step"},
            "clocks": clk, "gpu_launches": n_launch,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_K, "ms_per_step": e2e_ms / e2e_K},
            "roofline": roofline, "cpu_baseline": cpu, "ppo_update": ppo}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def ppo_block(ns, torch, dist, dev, world, rank):
    """PPO update samples/sec: one epoch (4 minibatches) of MATTrainer.train at BASELINE configs[3]'s
    per-GPU share (8192 envs x 20 workers x T=50 over 8 GPUs = 1024 envs per GPU), forward + backward +
    gradient all-reduce + Adam.  A row is one (env, t) sample carrying all 21 agents (shared_buffer.py:248)."""
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
    obs, sob, ava = env.reset_device()
    buf.obs[0].copy_(obs); buf._share_obs[0].copy_(env.sob); buf.available_actions[0].copy_(ava)
    for t in range(T):
        act, lp, v = pol.transformer.get_actions(None, obs, ava)
        obs, sob, rew, done, delay, pay, ava = env.step_device(act)
        buf.insert(env.sob, obs, None, None, act, lp, v, rew, 1.0 - done.float(), None, None, ava)
    tr.train(buf)                                   # warm-up epoch (allocations)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    info = tr.train(buf)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    rows = world * N * T
    return {"rows_per_s": rows / (ms.item() * 1e-3), "agent_samples_per_s": rows * L / (ms.item() * 1e-3), "ms_per_epoch": ms.item(),
            "rows_per_epoch": rows, "minibatches": 4, "envs_per_gpu": N, "workers": W, "T": T,
            "includes": "get_values + GAE + advantage norm + 4 x (forward, backward, grad all-reduce, clip, Adam)",
            "value_loss": info["value_loss"], "grad_norm": info["actor_grad_norm"]}


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
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-ppo", dest="ppo", action="store_false", help="skip the PPO-update block")
    ns = ap.parse_args()
    if ns.impl == "reference":
        run_reference(ns)
    else:
        run_ours(ns)


if __name__ == "__main__":
    main()
