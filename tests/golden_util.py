"""Helpers shared by the oracle-vs-golden (CPU) and CUDA-vs-oracle/golden (GPU) tests."""
import glob
import json
import os

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN_DIR = os.path.join(REPO, 'tests', 'golden')
ORACLE_KEYS = ('num_guards', 'num_attackers', 'agent_acc', 'agent_max_speed', 'agent_max_rot',
               'agent_shoot_base_radius', 'agent_shoot_win', 'random_jam_prob', 'introduce_terrain',
               'terrain_parameters', 'MaxEpisodeStep', 'init_distance', 'enable_attacker_fluctuation',
               'attacker_fluctuation', 'size', 'half_death_reward', 'REWARD_DEBUG', 'REWARD_DEBUG_Value',
               'DISALBE_RED_FUNCTION')
STATE_KEYS = ('pos', 'vel', 'atk_rad', 'terrain', 'alive', 'just_died', 'hit', 'was_hit', 'num_hit', 'num_was_hit',
              'init_theta', 'time_step', 'n_alive_guards', 'n_alive_attackers', 'obs_counter', 'episode')
INT_KEYS = ('alive', 'just_died', 'hit', 'was_hit', 'num_hit', 'num_was_hit', 'time_step', 'n_alive_guards',
            'n_alive_attackers', 'obs_counter', 'episode')
FLOAT_KEYS = ('pos', 'vel', 'atk_rad', 'terrain', 'init_theta')


def golden_cases():
    return sorted(os.path.basename(p)[4:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, 'dca_*.npz')))


def load_case(name):
    z = np.load(os.path.join(GOLDEN_DIR, 'dca_%s.npz' % name), allow_pickle=False)
    d = {k: z[k] for k in z.files}
    cfg = json.loads(str(d['cfg_json']))
    d['cfg'] = {k: cfg[k] for k in ORACLE_KEYS}
    d['seed'] = int(d['seed'])
    d['gid'] = int(d['gid'])
    return d


def state0(case):
    return {k: case['s0_' + k] for k in STATE_KEYS}


def slot_ids_from_obs(obs):
    """[..., G, 192] obs -> [..., G, 12] entity id per row, -1 for all-zero rows (column 0 = alive = 1 on
    every gathered row, column 6 = iden; collective_assault_env.py:237-252)."""
    o = obs.reshape(obs.shape[:-1] + (12, 16))
    valid = o[..., 0] != 0
    return np.where(valid, o[..., 6].astype(np.int32), -1)


def assert_close_rel(a, b, rtol, name, atol=0.0):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    both_nan = np.isnan(a) & np.isnan(b)
    err = np.abs(a - b)
    lim = rtol * np.maximum(np.abs(a), np.abs(b)) + atol
    bad = ~both_nan & ~(err <= lim)
    assert not bad.any(), '%s: %d mismatches, max abs err %g' % (name, bad.sum(), np.nanmax(np.where(bad, err, 0)))
