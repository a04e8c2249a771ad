"""ORACLE / TEST INFRASTRUCTURE -- generates tests/golden/runner_cases.npz by EXECUTING THE REFERENCE's
task_runner.Runner.update_runner / _checkout_interested_agents (task_runner.py:95-138, :231-287) on synthetic
(done, reward, info) streams.  Run in the build container only (needs /root/reference):

    python oracle/make_golden_runner.py

The Runner methods are called unbound on a plain namespace object carrying the attributes Runner.__init__ /
init_runner would have set (no envs, no algorithms, no logger); the env side is a stand-in that follows
EnvAutoReset's contract (UTIL/shm_env.py:33-74): a paused env (NaN action row) returns its cached previous result.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(HERE, 'ref_harness'))
sys.path.insert(0, HERE)
OUT = os.path.join(REPO, 'tests', 'golden', 'runner_cases.npz')

CASES = {
    'align': dict(E=37, G=6, T=160, align=True, report_interval=16, len_lo=4, len_hi=18, seed=1),
    'free': dict(E=37, G=6, T=160, align=False, report_interval=10, len_lo=4, len_hi=18, seed=2),
    # every env ends on the same step (time-limit episodes): many episodes per step, several windows per step
    'burst': dict(E=70, G=3, T=60, align=False, report_interval=16, len_lo=9, len_hi=10, seed=3),
    'one_env': dict(E=1, G=2, T=40, align=True, report_interval=3, len_lo=2, len_hi=6, seed=4),
}


class _Mcv(object):
    def rec(self, *a, **k):
        pass

    def rec_show(self, *a, **k):
        pass


def run_case(spec):
    from ref_env import setup_reference
    setup_reference()
    import task_runner
    from config import GlobalConfig as cfg
    Runner = task_runner.Runner
    E, G, T = spec['E'], spec['G'], spec['T']
    rng = np.random.RandomState(spec['seed'])

    class SC(object):
        N_AGENT_EACH_TEAM = [G]
        AGENT_ID_EACH_TEAM = [list(range(G))]
    cfg.ScenarioConfig = SC
    reports = []
    me = types.SimpleNamespace()
    me.n_thread, me.n_agent, me.n_team = E, G, 1
    me.RewardAsUnity = False
    me.align_episode = spec['align']
    me.report_interval = spec['report_interval']
    me.train_time_testing, me.test_only, me.test_interval = False, False, 10 ** 9
    me.max_n_episode = 10 ** 9
    me.current_n_frame = me.current_n_episode = 0
    me._exit_early_ = False
    me.top_rewards = me.test_top_rewards = None
    me.mcv = _Mcv()
    me.note = 'golden'
    me.test_envs = None
    me.info_runner = {}

    def checkout(info_runner, testing=False):
        wr, mr = Runner._checkout_interested_agents(me, info_runner, testing)
        reports.append((me.current_n_episode, float(wr[0]), float(mr[0]), float(me.top_rewards[0]),
                        float(np.mean(me.info_runner['t0_win_cnt_avg']))))
        return wr, mr
    me._checkout_interested_agents = checkout
    # init_runner (task_runner.py:77-93) with a stand-in for self.envs.reset()
    me.envs = types.SimpleNamespace(reset=lambda: (np.zeros((E, G, 4), np.float32),
                                                   np.array([{} for _ in range(E)], dtype=object)))
    Runner.init_runner(me)

    left = rng.randint(spec['len_lo'], spec['len_hi'] + 1, size=E)
    cached = None
    rec = {k: [] for k in ('done', 'win', 'reward', 'pause_in', 'pause', 'suffered', 'cur_step', 'ep_cnt', 'reward_sum',
                           'n_frame', 'n_episode')}
    for t in range(T):
        P = me.info_runner['ENV-PAUSE'].copy()  # rows MMPlatform.act would fill with NaN (multi_team.py:62-66)
        obs = np.zeros((E, G, 4), np.float32)
        reward = rng.choice([0.0, 1.0, -0.5, 0.5, 3.0], size=(E, G))
        done = np.zeros((E,), bool)
        info = np.empty((E,), dtype=object)
        win = np.zeros((E,), bool)
        for i in range(E):
            if P[i]:
                assert cached is not None
                reward[i], done[i], info[i], win[i] = cached[0][i], cached[1][i], cached[2][i], cached[3][i]
                continue
            left[i] -= 1
            if left[i] == 0:
                done[i] = True
                win[i] = rng.rand() < 0.4
                info[i] = {'win': bool(win[i]), 'obs-echo': None, 'others': None}
                left[i] = rng.randint(spec['len_lo'], spec['len_hi'] + 1)
            else:
                info[i] = {}
        cached = (reward.copy(), done.copy(), info.copy(), win.copy())
        Runner.update_runner(me, done, obs, reward, info)
        rec['pause_in'].append(P)
        rec['done'].append(done)
        rec['win'].append(win)
        rec['reward'].append(reward)
        ir = me.info_runner
        rec['pause'].append(ir['ENV-PAUSE'].copy())
        rec['suffered'].append(ir['Env-Suffered-Reset'].copy())
        rec['cur_step'].append(ir['Current-Obs-Step'].copy())
        rec['ep_cnt'].append(ir['Thread-Episode-Cnt'].copy())
        rec['reward_sum'].append(ir['Latest-Reward-Sum'].copy())
        rec['n_frame'].append(me.current_n_frame)
        rec['n_episode'].append(me.current_n_episode)
    out = {k: np.array(v) for k, v in rec.items()}
    out['reports'] = np.array(reports, np.float64).reshape(len(reports), 5)
    out['spec'] = np.array([E, G, T, int(spec['align']), spec['report_interval']], np.int64)
    return out


if __name__ == '__main__':
    allrec = {}
    for name, spec in CASES.items():
        r = run_case(spec)
        for k, v in r.items():
            allrec['%s__%s' % (name, k)] = v
        print(name, 'episodes', int(r['n_episode'][-1]), 'frames', int(r['n_frame'][-1]), 'reports', len(r['reports']),
              'paused-steps', int(r['pause_in'].sum()))
    np.savez_compressed(OUT, **allrec)
    print(OUT, '%.1f KB' % (os.path.getsize(OUT) / 1024))
