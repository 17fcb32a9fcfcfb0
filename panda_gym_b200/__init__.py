"""panda_gym_b200 -- B200-native batched Panda simulator behind panda-gym's gymnasium / VecEnv API.

Hot path only: env step / reset / compute_reward as hand-written sm_100a CUDA kernels
behind a C ABI (include/panda_b200.h); Python here is the host-side mirror of the reference interface.
"""
from .config import TrainConfig  # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):
    # heavy imports (torch, the CUDA library) only when the env classes are actually used
    if name in ("PandaVecEnv", "make_vec_env"):
        from . import vec_env
        return getattr(vec_env, name)
    if name in ("PandaReachAOEnv", "PandaReachEnv", "make", "register_envs", "register_reach_ao"):
        from . import envs
        return getattr(envs, name)
    if name in ("PandaBatch",):
        from . import sim
        return getattr(sim, name)
    raise AttributeError(name)
