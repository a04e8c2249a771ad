"""ORACLE / TEST INFRASTRUCTURE -- ctypes wrapper of oracle/mt_oracle.c (the dca_multiteam restatement).
Only tests/ may import this."""
import ctypes as C

import numpy as np

import dca_oracle

MT_ROWS, MT_ROW, MT_MAX_TEAMS = 12, 17, 8
DEFAULT_CFG = dict(N_AGENT_EACH_TEAM=[25, 25], size=5.0, random_jam_prob=0.0, introduce_terrain=False,
                   terrain_parameters=[0, 0], MaxEpisodeStep=150, StepsToNarrowCircle=100, init_distance=2.5,
                   half_death_reward=True)


class MtConfig(C.Structure):
    _fields_ = [('n_team', C.c_int32), ('team_size', C.c_int32 * MT_MAX_TEAMS), ('size', C.c_double),
                ('random_jam_prob', C.c_double), ('introduce_terrain', C.c_int32), ('terrain_A', C.c_double),
                ('terrain_B', C.c_double), ('max_episode_step', C.c_int32), ('steps_to_narrow', C.c_int32),
                ('init_distance', C.c_double), ('half_death_reward', C.c_int32), ('seed', C.c_uint64)]


STATE_FIELDS = (('pos', np.float64, 2), ('vel', np.float64, 2), ('atk_rad', np.float64, 0), ('terrain', np.float64, 0),
                ('alive', np.uint8, 0), ('just_died', np.uint8, 0), ('hit', np.uint8, 0), ('was_hit', np.uint8, 0),
                ('bsa', np.uint8, 0))
ENV_FIELDS = (('init_theta', np.float64), ('time_step', np.int32), ('obs_counter', np.uint32), ('episode', np.uint32))


class MtState(C.Structure):
    _fields_ = [(n, C.c_void_p) for n, _, _ in STATE_FIELDS] + [(n, C.c_void_p) for n, _ in ENV_FIELDS]


class MtOut(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ('obs', 'tobs', 'reward', 'done', 'ranking', 'slot_id')]


def make_config(cfg, seed=0):
    full = dict(DEFAULT_CFG)
    full.update(cfg)
    sizes = list(full['N_AGENT_EACH_TEAM'])
    c = MtConfig()
    c.n_team = len(sizes)
    for i, n in enumerate(sizes):
        c.team_size[i] = n
    c.size, c.random_jam_prob = full['size'], full['random_jam_prob']
    c.introduce_terrain = int(full['introduce_terrain'])
    c.terrain_A, c.terrain_B = full['terrain_parameters']
    c.max_episode_step, c.steps_to_narrow = full['MaxEpisodeStep'], full['StepsToNarrowCircle']
    c.init_distance, c.half_death_reward, c.seed = full['init_distance'], int(full['half_death_reward']), seed
    return c, full


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleMtEnv(object):
    def __init__(self, cfg, n_env, seed=0, env_gid0=0):
        dca_oracle.build()
        self.lib = C.CDLL(dca_oracle.LIB_PATH)
        self.lib.mt_reset.argtypes = [C.POINTER(MtConfig), C.POINTER(MtState), C.POINTER(MtOut), C.c_int, C.c_int64]
        self.lib.mt_step.argtypes = [C.POINTER(MtConfig), C.POINTER(MtState), C.POINTER(MtOut), C.c_int, C.c_int64,
                                     C.c_void_p, C.c_int]
        self.cfg, self.full = make_config(cfg, seed)
        self.E, self.T = n_env, self.cfg.n_team
        self.N = sum(self.full['N_AGENT_EACH_TEAM'])
        self.gid0 = env_gid0
        E, N = self.E, self.N
        self.state = {n: np.zeros((E, N) + ((d,) if d else ()), dt) for n, dt, d in STATE_FIELDS}
        self.state.update({n: np.zeros((E,), dt) for n, dt in ENV_FIELDS})
        self.state['episode'][:] = np.uint32(0xFFFFFFFF)      # first reset -> episode 0
        self.state['obs_counter'][:] = np.uint32(0xFFFFFFFF)  # first observation pass -> counter 0
        self.obs = np.zeros((E, N, MT_ROWS * MT_ROW), np.float32)
        self.tobs = np.zeros_like(self.obs)
        self.reward = np.zeros((E, N), np.float64)
        self.done = np.zeros((E,), np.uint8)
        self.ranking = np.zeros((E, self.T), np.int32)
        self.slot_id = np.zeros((E, N, MT_ROWS), np.int16)

    def _structs(self):
        st = MtState(*[_p(self.state[n]) for n, _, _ in STATE_FIELDS], *[_p(self.state[n]) for n, _ in ENV_FIELDS])
        out = MtOut(_p(self.obs), _p(self.tobs), _p(self.reward), _p(self.done), _p(self.ranking), _p(self.slot_id))
        return st, out

    def reset(self):
        st, out = self._structs()
        assert self.lib.mt_reset(C.byref(self.cfg), C.byref(st), C.byref(out), self.E, self.gid0) == 0
        return self.obs

    def step(self, actions, autoreset=True):
        a = np.ascontiguousarray(actions, dtype=np.float64).reshape(self.E, self.N)
        st, out = self._structs()
        assert self.lib.mt_step(C.byref(self.cfg), C.byref(st), C.byref(out), self.E, self.gid0, _p(a),
                                int(autoreset)) == 0
        return self.obs, self.reward, self.done, self.ranking

    def import_state(self, e, d):
        for n, _, _ in STATE_FIELDS:
            self.state[n][e] = d[n]
        for n, _ in ENV_FIELDS:
            self.state[n][e] = d[n]
