"""ORACLE / TEST INFRASTRUCTURE -- CPU restatement of task_runner.Runner.update_runner (task_runner.py:95-138) and of
the report arithmetic of _checkout_interested_agents (:231-275), vectorised over envs.  Pinned to the reference by
tests/golden/runner_cases.npz (made by oracle/make_golden_runner.py, which executes the reference's own methods).
Only tests/ may import this."""
import numpy as np


class RunnerOracle(object):
    def __init__(self, E, G, align_episode, report_interval):
        self.E, self.G, self.align, self.ri = E, G, bool(align_episode), int(report_interval)
        self.pause = np.zeros(E, bool)                 # ENV-PAUSE                 task_runner.py:85
        self.suffered = np.ones(E, bool)               # Env-Suffered-Reset        :84
        self.cur_step = np.zeros(E, np.int64)          # Current-Obs-Step          :86
        self.ep_cnt = np.zeros(E, np.int64)            # Thread-Episode-Cnt        :89
        self.reward_sum = np.zeros((E, G), np.float64)  # Latest-Reward-Sum        :88
        self.n_frame = self.n_episode = 0
        self.recent_sum, self.recent_win = [], []
        self.top = None
        self.win_avg = []
        self.reports = []

    def update(self, done, win, reward):
        P = self.pause.copy()
        R = ~P
        self.reward_sum[R] += reward[R]                # :103
        self.cur_step[R] += 1                          # :104
        self.suffered[:] = done                        # :106
        for i in np.nonzero(R & done)[0]:              # env-index order, :105-133
            self.n_frame += int(self.cur_step[i])
            self.n_episode += 1
            self.recent_sum.append(self.reward_sum[i].copy())
            self.recent_win.append(1 if win[i] else 0)
            self.reward_sum[i] = 0
            self.cur_step[i] = 0
            self.ep_cnt[i] += 1
            if self.align:
                self.pause[i] = True
            if self.n_episode % self.ri == 0:          # :124-128 -> :231-275
                mean_reward = float(np.stack(self.recent_sum).mean())
                win_rate = float(np.array(self.recent_win).mean())
                self.top = mean_reward if self.top is None else max(self.top, mean_reward)
                self.win_avg.append(win_rate)
                self.reports.append((self.n_episode, win_rate, mean_reward, self.top, float(np.mean(self.win_avg))))
                self.recent_sum, self.recent_win = [], []
        if self.align and self.pause.all():            # :135
            self.pause[:] = False
