"""ORACLE / TEST INFRASTRUCTURE -- runs only where /root/reference exists.

Imports the *unmodified* reference DCA environment read-only from
/root/reference (two import shims: gym, func_timeout; Cython built out of tree by
pyximport) and drives it under the keyed RNG stream of oracle/dca_rng.py.
Only ``np.random.*`` draw functions are swapped while the reference's
``reset_world`` / ``observation`` run; every line of reference arithmetic executes
as shipped.  Used by oracle/make_golden.py to produce tests/golden/*.npz and by
tests that pin the C oracle when the reference tree is present.

Reference entry points driven here:
  MISSION/dca/collective_assault_parallel_run.py:103-117 (make_collective_assault_env)
  MISSION/dca/collective_assault_parallel_run.py:192-270 (step / reset)
  UTIL/shm_env.py:18-122 (EnvAutoReset)
"""
import contextlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF_ROOT = os.environ.get('HMP_REFERENCE_ROOT', '/root/reference')

sys.path.insert(0, os.path.dirname(HERE))  # oracle/
import dca_rng  # noqa: E402

_READY = False


def reference_available():
    return os.path.isdir(os.path.join(REF_ROOT, 'MISSION', 'dca'))


def setup_reference():
    """Make the reference importable (idempotent)."""
    global _READY
    if _READY:
        return
    assert reference_available(), 'reference tree not present'
    work = os.path.join(REPO, 'build', 'ref_work')
    os.makedirs(work, exist_ok=True)
    os.chdir(work)  # reference code writes ./TEMP relative to cwd
    for p in (REF_ROOT, os.path.join(HERE, 'shim')):
        if p not in sys.path:
            sys.path.insert(0, p)
    import pyximport
    pyximport.install(build_dir=os.path.join(REPO, 'build', 'pyx'), inplace=False, language_level=3,
                      setup_args={'include_dirs': np.get_include()})
    _READY = True


# --------------------------------------------------------------------------------------
# keyed replacements for the np.random draw functions
# --------------------------------------------------------------------------------------
@contextlib.contextmanager
def _patched_np_random(**funcs):
    saved = {k: getattr(np.random, k) for k in funcs}
    try:
        for k, f in funcs.items():
            setattr(np.random, k, f)
        yield
    finally:
        for k, f in saved.items():
            setattr(np.random, k, f)


class _ResetDraws(object):
    """Sequential reader over the keyed reset draws, in reset_world's call order
    (collective_assault_env.py:64, :70-73, :85, :97/:120)."""

    def __init__(self, seed, env_gid, episode, n_guards, n_attackers):
        self.d = dca_rng.reset_draws(seed, env_gid, episode, n_guards + n_attackers)
        self.G, self.A = n_guards, n_attackers
        self.n_rand = 0
        self.n_randn = 0

    def rand(self, *shape):
        assert shape == ()
        i = self.n_rand
        self.n_rand += 1
        return self.d['theta_u'] if i == 0 else self.d['agent_u'][i - 1]

    def randn(self, *shape):
        assert shape == (2,)
        i = self.n_randn
        self.n_randn += 1
        return self.d['agent_n'][i].copy()

    def randint(self, low, high):
        return low + ((int(self.d['fluct_word']) * (high - low)) >> 32)

    def choice(self, arr, size, replace):
        assert not replace and len(arr) == self.A
        keys = self.d['agent_key'][self.G:]
        order = sorted(range(self.A), key=lambda a: (int(keys[a]), a))
        return np.array(order[:size], dtype=np.int64)


class _ObsDraws(object):
    """np.random.shuffle / np.random.rand replacement for one (guard, list) pair;
    item_random_mv (collective_assault_env.py:217-226) calls shuffle then rand."""

    def __init__(self, words):
        self.words = words

    def shuffle(self, arr):
        dca_rng.keyed_shuffle(arr, self.words)

    def rand(self, n):
        return dca_rng.keyed_jam_uniforms(n, self.words)


DEFAULT_CFG = dict(num_guards=50, num_attackers=50, size=5.0, random_jam_prob=0.05, introduce_terrain=True,
                   terrain_parameters=[0.05, 0.2], MaxEpisodeStep=180, init_distance=4, half_death_reward=True,
                   enable_attacker_fluctuation=False, attacker_fluctuation=20, REWARD_DEBUG=False,
                   REWARD_DEBUG_Value=1, DISALBE_RED_FUNCTION=False, agent_acc=3.0, agent_max_speed=1.0,
                   agent_max_rot=0.17, agent_shoot_base_radius=0.4, agent_shoot_win=float(np.pi / 4))


class RefDcaEnv(object):
    """One reference env (wrapped in the reference's EnvAutoReset when autoreset=True)."""

    def __init__(self, cfg, seed=0, env_gid=0, autoreset=True, np_seed=12345):
        setup_reference()
        from config import GlobalConfig
        from MISSION.dca.collective_assault_parallel_run import ScenarioConfig
        full = dict(DEFAULT_CFG)
        full.update(cfg)
        self.cfg = full
        for k, v in full.items():
            assert hasattr(ScenarioConfig, k), k
            setattr(ScenarioConfig, k, v)
        G, A = full['num_guards'], full['num_attackers']
        ScenarioConfig.N_AGENT_EACH_TEAM = [G]
        ScenarioConfig.AGENT_ID_EACH_TEAM = [range(0, G)]
        ScenarioConfig.num_entity = A
        ScenarioConfig.num_object = A + G
        GlobalConfig.env_name = 'dca'
        self.G, self.A, self.N = G, A, G + A
        self.seed, self.env_gid = seed, env_gid
        self.episode = -1
        self.obs_counter = -1
        self._list_idx = 0
        # class-level cache of the flat table size (collective_assault_env.py:181-204) is per-process state:
        # clear it so envs of different N can live in one harness process
        from MISSION.dca.collective_assault_env import collective_assaultEnvV1
        collective_assaultEnvV1.raw_obs_size = -1
        np.random.seed(np_seed)  # planner's kmeans2 seeds still come from the global stream
        if autoreset:
            from UTIL.shm_env import EnvAutoReset
            self.wrapper = EnvAutoReset({'env_name': 'dca', 'proc_index': env_gid})
            self.env = self.wrapper._env
        else:
            from MISSION.dca.collective_assault_parallel_run import make_collective_assault_env
            self.wrapper = None
            self.env = make_collective_assault_env('dca', env_gid)
        self.world = self.env.world
        self._install_hooks()

    # -- hooks ------------------------------------------------------------------------
    def _install_hooks(self):
        env = self.env
        orig_reset_world = env.reset_callback
        orig_observation = env.observation_callback

        def keyed_reset_world():
            self.episode += 1
            d = _ResetDraws(self.seed, self.env_gid, self.episode, self.G, self.A)
            with _patched_np_random(rand=d.rand, randn=d.randn, randint=d.randint, choice=d.choice):
                orig_reset_world()

        def keyed_observation(agent, world, *a, **kw):
            if agent.iden == 0:
                self.obs_counter += 1
            state = {'lst': 0}

            def shuffle(arr):
                w = dca_rng.obs_words(self.seed, self.env_gid, self.obs_counter, agent.iden, state['lst'])
                state['draw'] = _ObsDraws(w)
                state['draw'].shuffle(arr)

            def rand(n):
                d = state['draw']
                state['lst'] += 1
                return d.rand(n)

            with _patched_np_random(shuffle=shuffle, rand=rand):
                return orig_observation(agent, world, *a, **kw)

        env.reset_callback = keyed_reset_world
        env.observation_callback = keyed_observation

    # -- state ------------------------------------------------------------------------
    def export_state(self):
        ag = self.world.agents
        f = lambda name, dt: np.array([getattr(a, name) for a in ag], dtype=dt)
        return dict(
            pos=np.array([a.pos for a in ag], dtype=np.float64),
            vel=np.array([a.vel for a in ag], dtype=np.float64),
            atk_rad=f('atk_rad', np.float64), terrain=f('terrain', np.float64),
            alive=f('alive', np.uint8), just_died=f('justDied', np.uint8),
            hit=f('hit', np.uint8), was_hit=f('wasHit', np.uint8),
            num_hit=f('numHit', np.int32), num_was_hit=f('numWasHit', np.int32),
            init_theta=np.float64(self.world.init_theta), time_step=np.int32(self.world.time_step),
            n_alive_guards=np.int32(self.world.numAliveGuards),
            n_alive_attackers=np.int32(self.world.numAliveAttackers),
            obs_counter=np.int32(self.obs_counter), episode=np.int32(self.episode))

    def attacker_force(self):
        """Planner output of the last step: attacker.act[0:2] (cheat_script_ai.py:183-184)."""
        return np.array([a.act[:2] for a in self.world.agents[self.G:]], dtype=np.float64)

    # -- env surface ------------------------------------------------------------------
    def reset(self):
        ob, info = (self.wrapper.reset() if self.wrapper is not None else self.env.reset())
        return np.asarray(ob), info

    def step(self, act):
        target = self.wrapper if self.wrapper is not None else self.env
        ob, reward, done, info = target.step(np.asarray(act, dtype=np.float64))
        return np.asarray(ob), np.asarray(reward, dtype=np.float64), bool(done), info
