"""mat_dcml_b200: B200-native MAT-AS decision-and-rollout hot path.

Public surface mirrors the reference's for this path:
  TransformerPolicy, MultiAgentTransformer, the four transformer_act entry points,
  DcmlVecEnv (vec-env contract), SharedReplayBuffer.compute_returns, ValueNorm, MATTrainer,
  DcmlRunner (device-resident DCMLRunner.run / collect / insert / compute / train).
Everything computes in libmatas_b200.so (hand-written sm_100a CUDA, C ABI in include/matas_b200.h);
there is no CPU or eager fallback.
"""
from ._cabi import MatasError  # noqa: F401


def __getattr__(name):
    import importlib
    table = {"TransformerPolicy": "transformer_policy", "MultiAgentTransformer": "ma_transformer",
             "DcmlVecEnv": "dcml_env", "Action_Space": "dcml_env", "SharedReplayBuffer": "shared_buffer",
             "ValueNorm": "valuenorm", "MATTrainer": "mat_trainer", "DcmlRunner": "dcml_runner"}
    if name in table:
        return getattr(importlib.import_module("." + table[name], __name__), name)
    raise AttributeError(name)
