"""Runner bookkeeping: the oracle restatement and the CUDA kernels against the trajectory recorded
from the reference's own Runner.update_runner / _checkout_interested_agents (tests/golden/runner_cases.npz)."""
import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, 'oracle'))
GOLDEN = os.path.join(REPO, 'tests', 'golden', 'runner_cases.npz')
CASES = ['align', 'free', 'burst', 'one_env']


def load(name):
    z = np.load(GOLDEN)
    return {k.split('__', 1)[1]: z[k] for k in z.files if k.startswith(name + '__')}


@pytest.mark.parametrize('name', CASES)
def test_runner_oracle_matches_reference(name):
    from runner_oracle import RunnerOracle
    c = load(name)
    E, G, T, align, ri = [int(x) for x in c['spec']]
    orc = RunnerOracle(E, G, align, ri)
    for t in range(T):
        assert np.array_equal(orc.pause, c['pause_in'][t]), t
        orc.update(c['done'][t], c['win'][t], c['reward'][t])
        assert np.array_equal(orc.pause, c['pause'][t]), t
        assert np.array_equal(orc.suffered, c['suffered'][t]), t
        assert np.array_equal(orc.cur_step, c['cur_step'][t]), t
        assert np.array_equal(orc.ep_cnt, c['ep_cnt'][t]), t
        assert np.array_equal(orc.reward_sum, c['reward_sum'][t]), t
        assert (orc.n_frame, orc.n_episode) == (int(c['n_frame'][t]), int(c['n_episode'][t])), t
    assert len(orc.reports) == len(c['reports']) > 0
    assert np.array_equal(np.array(orc.reports, np.float64), c['reports'])


def _drive(run, c, T, check):
    import torch
    dev = run.device
    seen = []
    for t in range(T):
        assert np.array_equal(run.pause.cpu().numpy().astype(bool), c['pause_in'][t]), t
        run.update(torch.from_numpy(c['done'][t].astype(np.uint8)).to(dev),
                   torch.from_numpy(c['win'][t].astype(np.uint8)).to(dev),
                   torch.from_numpy(c['reward'][t].astype(np.float32)).to(dev))
        p = run.poll()
        seen += [(r['episode'], r['win_rate'], r['mean_reward'][0], r['top_reward'][0], r['acc_win_ratio'])
                 for r in p['reports']]
        check(t, p)
    return seen


@pytest.mark.gpu
@pytest.mark.parametrize('name', CASES)
def test_cuda_runner_matches_reference(name):
    import torch
    assert torch.cuda.is_available()
    from hmp2g_b200.runner import DeviceRunner
    c = load(name)
    E, G, T, align, ri = [int(x) for x in c['spec']]
    run = DeviceRunner(E, G, align_episode=bool(align), report_interval=ri)

    def check(t, p):
        i = run.info
        assert np.array_equal(i['ENV-PAUSE'].cpu().numpy().astype(bool), c['pause'][t]), t
        assert np.array_equal(i['Env-Suffered-Reset'].cpu().numpy().astype(bool), c['suffered'][t]), t
        assert np.array_equal(i['Current-Obs-Step'].cpu().numpy(), c['cur_step'][t]), t
        assert np.array_equal(i['Thread-Episode-Cnt'].cpu().numpy(), c['ep_cnt'][t]), t
        assert np.array_equal(i['Latest-Reward-Sum'].cpu().numpy(), c['reward_sum'][t]), t
        assert (p['current_n_frame'], p['current_n_episode']) == (int(c['n_frame'][t]), int(c['n_episode'][t])), t
    seen = _drive(run, c, T, check)
    assert np.array_equal(np.array(seen, np.float64), c['reports'])


@pytest.mark.gpu
@pytest.mark.parametrize('align', [False, True])
def test_cuda_runner_4096_envs_vs_oracle(align):
    """Full size: 4096 envs x 50 guards, time-limit bursts (thousands of episodes per step), against the oracle."""
    import torch
    from hmp2g_b200.runner import DeviceRunner
    from runner_oracle import RunnerOracle
    E, G, ri = 4096, 50, 64
    rng = np.random.RandomState(11)
    run = DeviceRunner(E, G, align_episode=align, report_interval=ri)
    orc = RunnerOracle(E, G, align, ri)
    dev = run.device
    left = np.where(rng.rand(E) < 0.7, 30, rng.randint(5, 30, size=E))
    seen = []
    last = None
    for t in range(95):
        P = orc.pause.copy()
        assert np.array_equal(run.pause.cpu().numpy().astype(bool), P), t
        reward = rng.choice([0.0, 1.0, -0.5, 0.5, 3.5], size=(E, G)).astype(np.float32)
        left = np.where(P, left, left - 1)
        done = (left == 0) & ~P
        win = done & (rng.rand(E) < 0.3)
        left = np.where(done, np.where(rng.rand(E) < 0.7, 30, rng.randint(5, 30, size=E)), left)
        if last is not None:  # a paused env returns its cached previous result (shm_env.py:37-41)
            reward[P], done[P], win[P] = last[0][P], last[1][P], last[2][P]
        last = (reward.copy(), done.copy(), win.copy())
        orc.update(done, win, reward.astype(np.float64))
        run.update(torch.from_numpy(done.astype(np.uint8)).to(dev), torch.from_numpy(win.astype(np.uint8)).to(dev),
                   torch.from_numpy(reward).to(dev))
        p = run.poll()
        seen += [(r['episode'], r['win_rate'], r['mean_reward'][0], r['top_reward'][0], r['acc_win_ratio'])
                 for r in p['reports']]
        assert (p['current_n_frame'], p['current_n_episode']) == (orc.n_frame, orc.n_episode), t
    i = run.info
    assert np.array_equal(i['Latest-Reward-Sum'].cpu().numpy(), orc.reward_sum)
    assert np.array_equal(i['Current-Obs-Step'].cpu().numpy(), orc.cur_step)
    assert np.array_equal(i['Thread-Episode-Cnt'].cpu().numpy(), orc.ep_cnt)
    assert orc.n_episode > 2 * E and len(orc.reports) == orc.n_episode // ri
    assert np.array_equal(np.array(seen, np.float64), np.array(orc.reports, np.float64))


@pytest.mark.gpu
def test_cuda_device_resident_loop_runner_histpool_rollout_gae():
    """The loop of INTEGRATION.md §3: env kernel + planner + DeviceRunner (align_episode) + HistoryPool + DeviceRollout
    + GAE, all on device tensors.  Checks the pieces against each other: frames counted by the runner == steps
    finalized by the rollout buffer, paused envs record nothing, returns are finite where valid."""
    import torch
    import hmp2g_b200
    from hmp2g_b200.config import ScenarioConfig
    from hmp2g_b200.rollout import HistoryPool
    E = 256
    scfg = type('S', (ScenarioConfig,), dict(num_guards=20, num_attackers=25, MaxEpisodeStep=14, init_distance=1.5))
    env = hmp2g_b200.DcaVecEnv(E, scfg=scfg, seed=9, empty_rows_nan=True)
    run = hmp2g_b200.DeviceRunner(E, env.G, align_episode=True, report_interval=32, test_interval=100)
    hist = HistoryPool(E, env.G)
    buf = hmp2g_b200.DeviceRollout(env)
    env.reset()
    g = torch.Generator(device=env.device)
    g.manual_seed(1)
    steps_finalized = episodes = tests_due = n_reports = paused_steps = 0
    for t in range(70):
        feed = hist.roll(env.obs, pause=run.pause, suffered_reset=run.info['Env-Suffered-Reset'])
        running = run.pause == 0
        paused_steps += int((~running).sum())
        # a stand-in policy: random actions, value = number of visible rows (NaN-aware, like the reference's masks)
        act = torch.randint(0, 7, (E, env.G), generator=g, device=env.device, dtype=torch.int32)
        value = (~torch.isnan(feed[..., 0])).sum(-1).float()
        alive = buf.alive_guards().clone()
        # hold a random half back during the first steps so that the envs do not finish in lockstep
        hold = (torch.rand((E,), generator=g, device=env.device) < (0.5 if t < 5 else 0.0)).to(torch.uint8)
        env.step(act, skip=run.pause | hold)
        stepped = running & (hold == 0)
        finished = buf.add(alive, value, env.reward, env.done, active=stepped)
        # the runner only knows its own pauses: give it done = 0 for the envs held back by the test
        done_seen = env.done * (hold == 0).to(torch.uint8)
        run.update(done_seen, env.win, env.reward)
        st = run.poll()
        tests_due += st['tests_due']
        n_reports += len(st['reports'])
        if bool(finished.any()):
            lens = buf.length.clone()
            ret, threat, adv, valid = buf.finalize(finished)
            assert torch.isfinite(ret[valid]).all() and torch.isfinite(adv[valid]).all()
            steps_finalized += int(lens[finished].sum())
            episodes += int(finished.sum())
    assert episodes == run.current_n_episode >= 2 * E
    # the runner also counts the steps an env was held back by the test (it believes they ran)
    assert run.current_n_frame >= steps_finalized and paused_steps > 0
    assert tests_due == run.current_n_episode // 100 and n_reports == run.current_n_episode // 32
