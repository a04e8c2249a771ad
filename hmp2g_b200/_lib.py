"""ctypes binding of include/dca_b200.h.  There is no CPU fallback: a missing library is an error."""
import ctypes as C
import os

from . import build as _build

_P = C.c_void_p


class DcaConfig(C.Structure):
    _fields_ = [('num_guards', C.c_int32), ('num_attackers', C.c_int32),
                ('agent_acc', C.c_double), ('agent_max_speed', C.c_double), ('agent_max_rot', C.c_double),
                ('agent_shoot_base_radius', C.c_double), ('agent_shoot_win', C.c_double),
                ('random_jam_prob', C.c_double), ('introduce_terrain', C.c_int32),
                ('terrain_A', C.c_double), ('terrain_B', C.c_double), ('max_episode_step', C.c_int32),
                ('init_distance', C.c_double), ('enable_attacker_fluctuation', C.c_int32),
                ('attacker_fluctuation', C.c_int32), ('size', C.c_double), ('half_death_reward', C.c_int32),
                ('reward_debug', C.c_int32), ('reward_debug_value', C.c_double),
                ('disable_red_function', C.c_int32), ('seed', C.c_uint64), ('env_gid0', C.c_int64)]


class DcaBuffers(C.Structure):
    _fields_ = [(n, _P) for n in ('obs', 'tobs', 'reward', 'done', 'win', 'alive', 'slot_id', 'nearest')]


STATE_FIELDS = ('pos', 'vel', 'atk_rad', 'terrain', 'alive', 'just_died', 'hit', 'was_hit', 'num_hit', 'num_was_hit',
                'init_theta', 'time_step', 'n_alive_guards', 'n_alive_attackers', 'obs_counter', 'episode')


class DcaState(C.Structure):
    _fields_ = [(n, _P) for n in STATE_FIELDS]


class DcaRunnerState(C.Structure):
    _fields_ = [(n, _P) for n in ('pause', 'suffered', 'cur_step', 'ep_cnt', 'reward_sum', 'counters', 'fin_slot',
                                  'log_meta', 'log_reward')] + [('log_capacity', C.c_int32)]


class DcaMtConfig(C.Structure):
    _fields_ = [('n_team', C.c_int32), ('team_size', C.c_int32 * 8), ('size', C.c_double),
                ('random_jam_prob', C.c_double), ('introduce_terrain', C.c_int32), ('terrain_A', C.c_double),
                ('terrain_B', C.c_double), ('max_episode_step', C.c_int32), ('steps_to_narrow', C.c_int32),
                ('init_distance', C.c_double), ('half_death_reward', C.c_int32), ('seed', C.c_uint64),
                ('env_gid0', C.c_int64)]


class DcaMtBuffers(C.Structure):
    _fields_ = [(n, _P) for n in ('obs', 'tobs', 'reward', 'done', 'ranking', 'alive', 'slot_id')]


MT_STATE_FIELDS = ('pos', 'vel', 'atk_rad', 'terrain', 'alive', 'just_died', 'hit', 'was_hit', 'bsa', 'init_theta',
                   'time_step', 'obs_counter', 'episode')


class DcaMtState(C.Structure):
    _fields_ = [(n, _P) for n in MT_STATE_FIELDS]


EXPORTS = ('dca_abi_version', 'dca_last_error', 'dca_create', 'dca_destroy', 'dca_bind_buffers', 'dca_reset',
           'dca_step', 'dca_observe', 'dca_plan', 'dca_plan_centroid', 'dca_export_state', 'dca_import_state', 'dca_gae',
           'dca_kernel_launches', 'dca_debug_phase_clocks', 'dca_set_option', 'dca_runner_init', 'dca_runner_update',
           'dca_runner_report', 'dca_histpool_roll', 'dca_adv_moments', 'dca_adv_normalize', 'dca_adv_moments_static', 'dca_norm_moments', 'dca_norm_apply', 'dca_conc_attend', 'dca_conc_attend2', 'dca_policy_encode', 'dca_policy_linear',
           'dca_mt_create', 'dca_mt_destroy',
           'dca_mt_bind_buffers', 'dca_mt_reset', 'dca_mt_step', 'dca_mt_export_state', 'dca_mt_import_state', 'dca_mt_debug_phase_clocks')

_lib = None


class DcaError(RuntimeError):
    pass


def load(build_if_missing=False):
    """Load libdca_b200.so (in-tree).  Raises DcaError if it is absent -- no fallback path exists."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if not os.path.exists(path):
        if build_if_missing:
            _build.build()
        else:
            raise DcaError('%s is missing: run `python -m hmp2g_b200.build` (or __graft_entry__.build()). '
                           'hmp2g_b200 has no CPU fallback.' % path)
    lib = C.CDLL(path)
    lib.dca_last_error.restype = C.c_char_p
    lib.dca_last_error.argtypes = [_P]
    lib.dca_create.argtypes = [C.POINTER(DcaConfig), C.c_int, C.c_int, C.POINTER(_P)]
    lib.dca_destroy.argtypes = [_P]
    lib.dca_bind_buffers.argtypes = [_P, C.POINTER(DcaBuffers)]
    lib.dca_reset.argtypes = [_P, _P, _P]
    lib.dca_step.argtypes = [_P, _P, _P, _P, _P, C.c_int, _P]
    lib.dca_observe.argtypes = [_P, _P, _P]
    lib.dca_plan_centroid.argtypes = [_P, _P, _P]
    lib.dca_plan.argtypes = [_P, _P, _P, _P, _P]
    lib.dca_set_option.argtypes = [_P, C.c_int, C.c_int]
    lib.dca_export_state.argtypes = [_P, C.POINTER(DcaState), _P]
    lib.dca_import_state.argtypes = [_P, C.POINTER(DcaState), _P]
    lib.dca_gae.argtypes = [_P, _P, _P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                            C.c_int, _P]
    lib.dca_runner_init.argtypes = [C.POINTER(DcaRunnerState), C.c_int, C.c_int, _P]
    lib.dca_runner_update.argtypes = [C.POINTER(DcaRunnerState), C.c_int, C.c_int, C.c_int, _P, _P, _P, _P]
    lib.dca_runner_report.argtypes = [C.POINTER(DcaRunnerState), C.c_int, C.c_int64, C.c_int, _P, C.c_int, _P, _P]
    lib.dca_histpool_roll.argtypes = [_P, _P, _P, _P, _P, C.c_int, C.c_int, _P]
    lib.dca_adv_moments.argtypes = [_P, _P, _P, C.c_int, C.c_uint64, _P, _P, _P]
    lib.dca_adv_normalize.argtypes = [_P, _P, _P, C.c_int, C.c_uint64, _P, _P, _P]
    lib.dca_adv_moments_static.argtypes = [_P, _P, _P, C.c_int, C.c_uint64, _P, _P]
    lib.dca_norm_moments.argtypes = [_P, C.c_uint64, C.c_int, _P, _P, _P]
    lib.dca_norm_apply.argtypes = [_P, _P, _P, C.c_uint64, C.c_int, _P, _P, _P]
    lib.dca_policy_encode.argtypes = [_P, _P, _P, _P, _P, _P, _P, C.c_uint64, _P, _P, _P, _P]
    lib.dca_conc_attend.argtypes = [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_uint64, _P, _P, _P]
    lib.dca_conc_attend2.argtypes = [_P, _P, _P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, _P, _P, _P, _P, _P]
    lib.dca_policy_linear.argtypes = [_P, C.c_int, C.c_int, _P, C.c_int, C.c_int, _P, C.c_int, C.c_int, _P, C.c_int, C.c_int,
                                      _P, C.c_int, C.c_int, C.c_uint64, _P, C.c_int, _P]
    lib.dca_mt_create.argtypes = [C.POINTER(DcaMtConfig), C.c_int, C.c_int, C.POINTER(_P)]
    lib.dca_mt_destroy.argtypes = [_P]
    lib.dca_mt_bind_buffers.argtypes = [_P, C.POINTER(DcaMtBuffers)]
    lib.dca_mt_reset.argtypes = [_P, _P, _P]
    lib.dca_mt_step.argtypes = [_P, _P, _P, C.c_int, _P]
    lib.dca_mt_export_state.argtypes = [_P, C.POINTER(DcaMtState), _P]
    lib.dca_mt_import_state.argtypes = [_P, C.POINTER(DcaMtState), _P]
    lib.dca_mt_debug_phase_clocks.argtypes = [_P, _P]
    lib.dca_kernel_launches.argtypes = [_P, C.POINTER(C.c_int64)]
    lib.dca_debug_phase_clocks.argtypes = [_P, _P]
    for name in EXPORTS:
        getattr(lib, name)  # AttributeError if the symbol is not exported
        if name != 'dca_last_error':
            getattr(lib, name).restype = C.c_int
    _lib = lib
    return lib


def check(rc, handle=None):
    if rc != 0:
        msg = load().dca_last_error(handle)
        raise DcaError('dca_b200 error %d: %s' % (rc, msg.decode() if msg else '?'))
