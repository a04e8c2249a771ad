"""ORACLE / TEST INFRASTRUCTURE -- ctypes front-end of oracle/dca_oracle.c.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package (hmp2g_b200) never does.
Parity status: pinned against the reference executed here (tests/golden, oracle/make_golden.py).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, 'libdca_oracle.so')
OBS_ROWS, OBS_ROW = 12, 16


def build(force=False):
    srcs = [os.path.join(HERE, 'dca_oracle.c'), os.path.join(HERE, 'dca_planner.c'), os.path.join(HERE, 'mt_oracle.c')]
    hdr = os.path.join(HERE, 'dca_oracle.h')
    if (not force and os.path.exists(LIB_PATH)
            and os.path.getmtime(LIB_PATH) >= max(os.path.getmtime(f) for f in srcs + [hdr, os.path.join(HERE, 'mt_oracle.h')])):
        return LIB_PATH
    cmd = ['gcc', '-O2', '-std=c11', '-ffp-contract=off', '-fno-fast-math', '-fopenmp', '-fPIC', '-shared',
           '-o', LIB_PATH] + srcs + ['-lm']
    subprocess.check_call(cmd)
    return LIB_PATH


class OrcConfig(C.Structure):
    _fields_ = [('num_guards', C.c_int32), ('num_attackers', C.c_int32),
                ('agent_acc', C.c_double), ('agent_max_speed', C.c_double), ('agent_max_rot', C.c_double),
                ('agent_shoot_base_radius', C.c_double), ('agent_shoot_win', C.c_double),
                ('random_jam_prob', C.c_double), ('introduce_terrain', C.c_int32),
                ('terrain_A', C.c_double), ('terrain_B', C.c_double), ('max_episode_step', C.c_int32),
                ('init_distance', C.c_double), ('enable_attacker_fluctuation', C.c_int32),
                ('attacker_fluctuation', C.c_int32), ('size', C.c_double), ('half_death_reward', C.c_int32),
                ('reward_debug', C.c_int32), ('reward_debug_value', C.c_double),
                ('disable_red_function', C.c_int32), ('seed', C.c_uint64)]


_P = C.c_void_p


class OrcState(C.Structure):
    _fields_ = [(n, _P) for n in ('pos', 'vel', 'atk_rad', 'terrain', 'alive', 'just_died', 'hit', 'was_hit',
                                  'num_hit', 'num_was_hit', 'init_theta', 'time_step', 'n_alive_guards',
                                  'n_alive_attackers', 'obs_counter', 'episode')]


class OrcOut(C.Structure):
    _fields_ = [(n, _P) for n in ('obs', 'tobs', 'reward', 'done', 'win', 'slot_id', 'nearest')]


DEFAULT_CFG = dict(num_guards=50, num_attackers=50, agent_acc=3.0, agent_max_speed=1.0, agent_max_rot=0.17,
                   agent_shoot_base_radius=0.4, agent_shoot_win=float(np.pi / 4), random_jam_prob=0.0,
                   introduce_terrain=False, terrain_parameters=[0, 0], MaxEpisodeStep=180, init_distance=4,
                   enable_attacker_fluctuation=False, attacker_fluctuation=20, size=5.0, half_death_reward=True,
                   REWARD_DEBUG=False, REWARD_DEBUG_Value=1, DISALBE_RED_FUNCTION=False)


def make_config(cfg, seed=0):
    """cfg uses ScenarioConfig attribute names (collective_assault_parallel_run.py:30-101)."""
    f = dict(DEFAULT_CFG)
    unknown = set(cfg) - set(f)
    assert not unknown, unknown
    f.update(cfg)
    return OrcConfig(f['num_guards'], f['num_attackers'], f['agent_acc'], f['agent_max_speed'], f['agent_max_rot'],
                     f['agent_shoot_base_radius'], f['agent_shoot_win'], f['random_jam_prob'],
                     int(f['introduce_terrain']), float(f['terrain_parameters'][0]), float(f['terrain_parameters'][1]),
                     f['MaxEpisodeStep'], float(f['init_distance']), int(f['enable_attacker_fluctuation']),
                     f['attacker_fluctuation'], f['size'], int(f['half_death_reward']), int(f['REWARD_DEBUG']),
                     float(f['REWARD_DEBUG_Value']), int(f['DISALBE_RED_FUNCTION']), seed)


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB_PATH)
        _lib.orc_reset.restype = C.c_int
        _lib.orc_step.restype = C.c_int
        _lib.orc_gae.restype = C.c_int
        _lib.orc_plan_centroid.restype = C.c_int
        _lib.orc_plan_core.restype = C.c_int
        _lib.orc_kmeans.restype = C.c_int
        _lib.orc_plan.restype = C.c_int
    return _lib


STATE_SPEC = [('pos', np.float64, 2), ('vel', np.float64, 2), ('atk_rad', np.float64, 0), ('terrain', np.float64, 0),
              ('alive', np.uint8, 0), ('just_died', np.uint8, 0), ('hit', np.uint8, 0), ('was_hit', np.uint8, 0),
              ('num_hit', np.int32, 0), ('num_was_hit', np.int32, 0)]
ENV_SPEC = [('init_theta', np.float64), ('time_step', np.int32), ('n_alive_guards', np.int32),
            ('n_alive_attackers', np.int32), ('obs_counter', np.uint32), ('episode', np.uint32)]


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleVecEnv(object):
    """E envs on the CPU oracle.  State/outputs are plain numpy arrays (attributes)."""

    def __init__(self, cfg, n_env, seed=0, env_gid0=0, nthreads=1):
        self.cfg = make_config(cfg, seed)
        self.E, self.G, self.A = n_env, self.cfg.num_guards, self.cfg.num_attackers
        self.N = self.G + self.A
        self.env_gid0, self.nthreads = env_gid0, nthreads
        E, N, G, A = self.E, self.N, self.G, self.A
        self.state = {}
        for name, dt, k in STATE_SPEC:
            self.state[name] = np.zeros((E, N, k) if k else (E, N), dtype=dt)
        for name, dt in ENV_SPEC:
            self.state[name] = np.zeros((E,), dtype=dt)
        # counters start at -1 so the first reset is episode 0 / obs pass 0
        self.state['obs_counter'][:] = np.uint32(0xFFFFFFFF)
        self.state['episode'][:] = np.uint32(0xFFFFFFFF)
        self.obs = np.zeros((E, G, OBS_ROWS * OBS_ROW), np.float32)
        self.tobs = np.zeros((E, G, OBS_ROWS * OBS_ROW), np.float32)
        self.reward = np.zeros((E, G), np.float64)
        self.done = np.zeros((E,), np.uint8)
        self.win = np.zeros((E,), np.uint8)
        self.slot_id = np.full((E, G, OBS_ROWS), -1, np.int16)
        self.nearest = np.full((E, A), -1, np.int32)

    def _structs(self):
        st = OrcState(*[_ptr(self.state[n]) for n, _ in OrcState._fields_])
        out = OrcOut(_ptr(self.obs), _ptr(self.tobs), _ptr(self.reward), _ptr(self.done), _ptr(self.win),
                     _ptr(self.slot_id), _ptr(self.nearest))
        return st, out

    def reset(self):
        st, out = self._structs()
        rc = lib().orc_reset(C.byref(self.cfg), C.byref(st), C.byref(out), self.E, C.c_int64(self.env_gid0),
                             self.nthreads)
        assert rc == 0, rc
        return self.obs

    def step(self, actions, att_force, rng_words=None, autoreset=True):
        actions = np.ascontiguousarray(actions, dtype=np.float64).reshape(self.E, self.G)
        att_force = np.ascontiguousarray(att_force, dtype=np.float64).reshape(self.E, self.A, 2)
        if rng_words is not None:
            rng_words = np.ascontiguousarray(rng_words, dtype=np.uint32).reshape(self.E, self.G, 2, 12)
        st, out = self._structs()
        rc = lib().orc_step(C.byref(self.cfg), C.byref(st), C.byref(out), self.E, C.c_int64(self.env_gid0),
                            _ptr(actions), _ptr(att_force), _ptr(rng_words), int(autoreset), self.nthreads)
        assert rc == 0, rc
        return self.obs, self.reward, self.done, self.win

    def plan_centroid(self):
        f = np.zeros((self.E, self.A, 2), np.float64)
        st, _ = self._structs()
        lib().orc_plan_centroid(C.byref(self.cfg), C.byref(st), self.E, _ptr(f), self.nthreads)
        return f

    def import_state(self, e, d):
        """d: dict from RefDcaEnv.export_state() / golden file (single env)."""
        for name, dt, _ in STATE_SPEC:
            self.state[name][e] = np.asarray(d[name]).astype(dt)
        for name, dt in ENV_SPEC:
            self.state[name][e] = np.asarray(d[name]).astype(dt)


def gae(reward, value, dead, traj_len=None, gamma=0.99, tau=0.95, threat_limit=11, gamma_fwd=1.0):
    """reward f64 [T,E,G], value f32 [T,E,G], dead bool [T,E,G] -> (ret f32, threat f32, pushed reward f64)."""
    reward = np.ascontiguousarray(reward, np.float64)
    value = np.ascontiguousarray(value, np.float32)
    dead = np.ascontiguousarray(dead, np.uint8)
    T, E, G = reward.shape
    ret = np.zeros((T, E, G), np.float32)
    threat = np.zeros((T, E, G), np.float32)
    pushed = np.zeros((T, E, G), np.float64)
    tl = None if traj_len is None else np.ascontiguousarray(traj_len, np.int32)
    rc = lib().orc_gae(_ptr(reward), _ptr(value), _ptr(dead), _ptr(tl), _ptr(ret), _ptr(threat), _ptr(pushed),
                       T, E, G, C.c_double(gamma), C.c_double(tau), C.c_double(gamma_fwd), int(threat_limit))
    assert rc == 0
    return ret, threat, pushed


# ---------------------------------------------------------------------------------------------- planner (dca_planner.c)
def plan_core(gpos, galive, apos, aalive, label, cent, disable_red=False):
    """Deterministic part of attackers_policy_1 (cheat_script_ai.py:49-184) for ONE env, given the team labels [A]
    and k-means centroids [K,2].  Returns attacker.act[0:2] as f64 [A,2]."""
    gpos = np.ascontiguousarray(gpos, np.float64); apos = np.ascontiguousarray(apos, np.float64)
    galive = np.ascontiguousarray(galive, np.uint8); aalive = np.ascontiguousarray(aalive, np.uint8)
    label = np.ascontiguousarray(label, np.int32); cent = np.ascontiguousarray(cent, np.float64)
    G, A, K = gpos.shape[0], apos.shape[0], cent.shape[0]
    force = np.zeros((A, 2), np.float64)
    rc = lib().orc_plan_core(G, A, _ptr(gpos), _ptr(galive), _ptr(apos), _ptr(aalive), _ptr(label), K, _ptr(cent),
                             int(bool(disable_red)), _ptr(force))
    assert rc == 0
    return force


def kmeans(pts, k, seed, gid, counter):
    """get_groups' kmeans2(pts, k, iter=20, minit='++') under the keyed stream.  Returns (label[n], cent[k,2], raised)."""
    pts = np.ascontiguousarray(pts, np.float64)
    n = pts.shape[0]
    label = np.zeros((n,), np.int32)
    cent = np.zeros((k, 2), np.float64)
    rc = lib().orc_kmeans(_ptr(pts), n, int(k), C.c_uint64(int(seed)), C.c_int64(int(gid)), C.c_uint32(int(counter)),
                          _ptr(label), _ptr(cent))
    assert rc >= 0
    return label, cent, rc


class OraclePlanner(object):
    """attackers_policy_1 with its persistent team state (start_flag survives resets) for E envs."""

    def __init__(self, env):
        self.env = env
        E, A = env.E, env.A
        self.K = max(A // 5, 1)
        self.start_flag = np.ones((E,), np.uint8)
        self.small = np.zeros((E,), np.uint8)
        self.counter = np.zeros((E,), np.uint32)
        self.label1 = np.zeros((E, A), np.int32)
        self.cent1 = np.zeros((E, self.K, 2), np.float64)

    def plan(self):
        env = self.env
        force = np.zeros((env.E, env.A, 2), np.float64)
        st, _ = env._structs()
        rc = lib().orc_plan(C.byref(env.cfg), C.byref(st), env.E, C.c_int64(env.env_gid0), _ptr(self.start_flag),
                            _ptr(self.small), _ptr(self.counter), _ptr(self.label1), _ptr(self.cent1), _ptr(force),
                            env.nthreads)
        assert rc == 0
        return force
