"""conc_4hist history pool: oracle restatement and CUDA kernel against the trajectories recorded
from the reference's ShellEnvWrapper / roll_hisory (tests/golden/histpool_cases.npz).  Pure data movement: bit-exact."""
import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, 'oracle'))
GOLDEN = os.path.join(REPO, 'tests', 'golden', 'histpool_cases.npz')
CASES = ['mixed', 'nopause']


def load(name):
    z = np.load(GOLDEN)
    return {k.split('__', 1)[1]: z[k] for k in z.files if k.startswith(name + '__')}


def same_bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


def same_values(a, b):
    """equal, NaN == NaN (the reference's NaN payload is numpy's default quiet NaN, as is ours)"""
    return np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize('name', CASES)
def test_histpool_oracle_matches_reference(name):
    from histpool_oracle import histpool_roll
    c = load(name)
    T, E, G = c['obs'].shape[:3]
    his = np.zeros((E, G, 12, 16), np.float32)
    for t in range(T):
        feed, his = histpool_roll(c['obs'][t], his, c['pause'][t], c['suffered'][t])
        assert same_values(feed, c['feed'][t]), t
        assert same_bits(his, c['his_pool'][t]), t


@pytest.mark.gpu
@pytest.mark.parametrize('name', CASES)
def test_cuda_histpool_matches_reference(name):
    import torch
    assert torch.cuda.is_available()
    from hmp2g_b200.rollout import HistoryPool
    c = load(name)
    T, E, G = c['obs'].shape[:3]
    hp = HistoryPool(E, G)
    dev = hp.device
    for t in range(T):
        feed = hp.roll(torch.from_numpy(c['obs'][t]).to(dev).view(E, G, 12, 16),
                       pause=torch.from_numpy(c['pause'][t].astype(np.uint8)).to(dev),
                       suffered_reset=torch.from_numpy(c['suffered'][t].astype(np.uint8)).to(dev))
        assert same_values(feed.cpu().numpy(), c['feed'][t]), t
        assert same_bits(hp.his_pool.cpu().numpy(), c['his_pool'][t]), t


@pytest.mark.gpu
def test_cuda_histpool_full_size_on_env_observations():
    """4096 envs x 50 guards fed by the env kernel's own observations (zero rows and the NaN-rows option), runner
    pause / reset masks from DeviceRunner, against the oracle."""
    import torch
    from hmp2g_b200.rollout import HistoryPool
    from hmp2g_b200.runner import DeviceRunner
    from hmp2g_b200.vec_env import DcaVecEnv
    from hmp2g_b200.config import ScenarioConfig
    from histpool_oracle import histpool_roll
    E = 4096
    scfg = type('S', (ScenarioConfig,), dict(MaxEpisodeStep=12))
    for nan_rows in (False, True):
        env = DcaVecEnv(E, scfg=scfg, seed=3, empty_rows_nan=nan_rows)
        G = env.G
        run = DeviceRunner(E, G, align_episode=True, report_interval=64)
        hp = HistoryPool(E, G)
        env.reset()
        his = np.zeros((E, G, 12, 16), np.float32)
        g = torch.Generator(device=env.device)
        g.manual_seed(0)
        n_paused = 0
        for t in range(30):
            pause, suffered = run.pause, run.info['Env-Suffered-Reset']
            n_paused += int(pause.sum())
            feed = hp.roll(env.obs, pause=pause, suffered_reset=suffered)
            if t % 6 == 0 or t >= 26:
                want_feed, his = histpool_roll(env.obs.cpu().numpy().reshape(E, G, 192), his,
                                               pause.cpu().numpy().astype(bool), suffered.cpu().numpy().astype(bool))
                assert same_values(feed.cpu().numpy(), want_feed), (nan_rows, t)
                assert same_bits(hp.his_pool.cpu().numpy(), his), (nan_rows, t)
            else:
                his = hp.his_pool.cpu().numpy()
            act = torch.randint(0, 7, (E, G), generator=g, device=env.device, dtype=torch.int32)
            # paused envs keep their previous (obs, reward, done), shm_env.py:37-41.  Early on (before any done) a
            # random half is held back as well, so that the time-limit endings spread over several steps and the
            # runner really pauses the early finishers
            hold = torch.rand((E,), generator=g, device=env.device) < (0.5 if t < 6 else 0.0)
            env.step(act, skip=(pause | hold.to(torch.uint8)))
            run.update(env.done, env.win, env.reward)
        assert int(run.counters[1]) >= E and int(run.counters[2]) >= 1 and n_paused > E
