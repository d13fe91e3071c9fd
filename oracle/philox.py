"""Counter-based random streams shared by the CUDA kernels, the oracle and the
reference-injection harness (test infrastructure -- see oracle/__init__.py).

The reference draws from Python's ``random`` and ``numpy.random`` global
streams in a data-dependent order, which no parallel
kernel can replay.  Parity is therefore defined on *named* draws: every random
quantity of the DCML simulator and of the stochastic policy is a pure function
``philox4x32_10(key=seed, counter=(env, step, tag, block))`` where
``tag = stream << 16 | index``.  The same definition is implemented in
``mat_dcml_b200/csrc/philox.cuh``; ``tests/golden/make_golden.py`` feeds these
values to the unmodified reference through shims of ``random`` / ``np.random``.

Streams (``index`` is the worker or agent number):

==========  ===  ==========================================================
name        id   use
==========  ===  ==========================================================
RETRY        1   n-th geometric-retry uniform of worker ``index`` during
                 ``Worker.process`` (block n//4, lane n%4)
JITTER       2   20 workload jitter uniforms of worker ``index`` (blocks 0-4)
RESET        3   block 16*r+0: lanes = disable count, arrive slot, R, C;
                 block 16*r+1: lane 0 = master Pr (unused), lane 1 = done coin
WORKER       4   block r: lane 0 = worker Pr uniform, lane 1 = disable key
POLICY       8   block 0: lane 0 = categorical uniform of agent ``index``;
                 lanes 0-3 = Box-Muller inputs of the continuous agent
==========  ===  ==========================================================

``r`` is the reset epoch within a step (1 only for the second reset that the
reference's ShareDummyVecEnv performs when an env reports done).
"""
import numpy as np

S_RETRY, S_JITTER, S_RESET, S_WORKER, S_POLICY = 1, 2, 3, 4, 8

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)
_SH = np.uint64(32)


def philox4x32(c0, c1, c2, c3, seed):
    """Philox-4x32-10.  Counter words broadcast as uint32 arrays; returns 4 uint32 arrays."""
    c0, c1, c2, c3 = np.broadcast_arrays(*[np.asarray(c, dtype=np.uint64) & _MASK for c in (c0, c1, c2, c3)])
    k0 = int(seed) & 0xFFFFFFFF
    k1 = (int(seed) >> 32) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> _SH, p0 & _MASK
        hi1, lo1 = p1 >> _SH, p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)) & _MASK, lo1, (hi0 ^ c3 ^ np.uint64(k1)) & _MASK, lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def tag(stream, index=0):
    return (int(stream) << 16) | int(index)


def u01(x):
    """24-bit uniform in [0,1): exactly representable in fp32 and fp64."""
    return (np.asarray(x, dtype=np.uint32) >> np.uint32(8)).astype(np.float64) * (1.0 / 16777216.0)


def bounded(x, n):
    """Integer in [0, n) by multiply-shift of a uint32."""
    return ((np.asarray(x, dtype=np.uint64) * np.uint64(n)) >> _SH).astype(np.int64)


def block(seed, env, step, stream, index, blk):
    """One Philox block as 4 python ints."""
    r = philox4x32(env, step, tag(stream, index), blk, seed)
    return tuple(int(np.asarray(v).reshape(-1)[0]) for v in r)


def seq_uniform(seed, env, step, stream, index, n):
    """n-th uniform of a sequential stream (block n//4, lane n%4)."""
    return float(u01(block(seed, env, step, stream, index, n // 4)[n % 4]))


def policy_uniforms(seed, step, batch, n_agent, env0=0):
    """(batch, n_agent) categorical uniforms used by the stochastic decode."""
    env = (np.arange(batch, dtype=np.uint64) + np.uint64(env0))[:, None]
    agent = np.arange(n_agent, dtype=np.uint64)[None, :]
    x0, _, _, _ = philox4x32(env, step, (np.uint64(S_POLICY) << np.uint64(16)) | agent, 0, seed)
    return u01(x0).astype(np.float32)


def policy_normals(seed, step, batch, agent, env0=0):
    """(batch, 2) standard normals for the continuous agent (Box-Muller on lanes 0-3)."""
    env = np.arange(batch, dtype=np.uint64) + np.uint64(env0)
    x = philox4x32(env, step, tag(S_POLICY, agent), 0, seed)
    u = [(np.asarray(v, dtype=np.uint32) >> np.uint32(8)).astype(np.float64) for v in x]
    # (u+1)/2^24 in (0,1] keeps log finite
    r0 = np.sqrt(-2.0 * np.log((u[0] + 1.0) / 16777216.0))
    r1 = np.sqrt(-2.0 * np.log((u[2] + 1.0) / 16777216.0))
    z0 = r0 * np.cos(2.0 * np.pi * (u[1] / 16777216.0))
    z1 = r1 * np.cos(2.0 * np.pi * (u[3] / 16777216.0))
    return np.stack([z0, z1], axis=1).astype(np.float32)
