"""hmp2g_b200 -- B200-native DCA (Decentralized Collective Assault) batched env step + observation + GAE.

Drop-in for the hot path of binary-husky/hmp2g: MISSION/dca behind UTIL/shm_env.SuperpoolEnv.  The compute lives in
libdca_b200.so (hand-written sm_100a CUDA, C-ABI in include/dca_b200.h); this package is the thin host mirror of the
reference's Python interfaces.  No CPU fallback exists.
"""
from .config import ChainVar, GlobalConfig, ScenarioConfig, inject, load_jsonc  # noqa: F401

__all__ = ['ChainVar', 'GlobalConfig', 'ScenarioConfig', 'inject', 'load_jsonc', 'DcaVecEnv', 'DcaSuperpoolEnv',
           'DcaEnvGroups', 'make_parallel_envs', 'gae_return', 'DeviceRollout', 'DeviceRunner', 'MtScenarioConfig', 'DcaMultiTeamVecEnv',
           'DcaMultiTeamSuperpoolEnv']


def __getattr__(name):  # torch is imported lazily so that config-only users do not pay for it
    if name in ('DcaVecEnv', 'DcaSuperpoolEnv', 'DcaEnvGroups', 'make_parallel_envs', 'pod_config'):
        from . import vec_env
        return getattr(vec_env, name)
    if name in ('gae_return', 'normalized_advantage'):
        from . import gae
        return getattr(gae, name)
    if name == 'DeviceRollout':
        from . import rollout
        return rollout.DeviceRollout
    if name in ('MtScenarioConfig', 'DcaMultiTeamVecEnv', 'DcaMultiTeamSuperpoolEnv'):
        from . import multiteam
        return getattr(multiteam, name)
    if name == 'DeviceRunner':
        from . import runner
        return runner.DeviceRunner
    raise AttributeError(name)
