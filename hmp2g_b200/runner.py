"""Runner bookkeeping on the device: task_runner.Runner.init_runner / update_runner / _checkout_interested_agents
(task_runner.py:77-138, :231-287; test-run twin :155-217) without the per-env Python loop.

The state lives in torch CUDA tensors (named after the reference's info_runner keys) and is advanced by
dca_runner_update (hmp2g_b200/csrc/dca_runner.cuh).  `pause` is handed to DcaVecEnv.step(skip=...) -- the ENV-PAUSE
rows that MMPlatform.act turns into NaN actions (multi_team.py:62-66).
"""
import ctypes as C

import numpy as np
import torch

from . import _lib


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class DeviceRunner(object):
    def __init__(self, n_env, n_agent, agent_id_each_team=None, align_episode=True, report_interval=64,
                 test_interval=None, max_n_episode=None, device=None, log_capacity=None):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise _lib.DcaError('hmp2g_b200 needs a CUDA device; there is no CPU fallback')
        self.device = torch.device(device if device is not None else 'cuda')
        assert self.device.type == 'cuda'
        if self.device.index is None:  # 'cuda' means the current device, exactly as DcaVecEnv resolves it
            self.device = torch.device('cuda', torch.cuda.current_device())
        self.E, self.G = int(n_env), int(n_agent)
        self.align_episode = bool(align_episode)
        self.report_interval = int(report_interval)
        self.test_interval = None if test_interval is None else int(test_interval)
        self.max_n_episode = max_n_episode
        teams = agent_id_each_team if agent_id_each_team is not None else [list(range(self.G))]
        self.n_team = len(teams)
        toa = np.full((self.G,), -1, np.int32)  # agents outside every team count for none (AGENT_ID_EACH_TEAM)
        for t, ids in enumerate(teams):
            toa[np.asarray(list(ids), dtype=np.int64)] = t
        dev = self.device
        self._team_of_agent = torch.from_numpy(toa).to(dev)
        # a report window is read back after the step that completes it: up to E more episodes may be logged by then
        self.log_capacity = int(log_capacity or (self.E + max(self.report_interval, 1)))
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)
        self.info = {
            'ENV-PAUSE': z((self.E,), torch.uint8), 'Env-Suffered-Reset': z((self.E,), torch.uint8),
            'Current-Obs-Step': z((self.E,), torch.int32), 'Thread-Episode-Cnt': z((self.E,), torch.int32),
            'Latest-Reward-Sum': z((self.E, self.G), torch.float64),
        }
        self.counters = z((8,), torch.int64)
        self._fin_slot = z((self.E,), torch.int32)
        self.log_meta = z((self.log_capacity, 4), torch.int64)
        self.log_reward = z((self.log_capacity, self.G), torch.float64)
        self._report_out = z((self.n_team + 1,), torch.float64)
        self._h_counters = torch.zeros((8,), dtype=torch.int64, pin_memory=True)
        self._state = _lib.DcaRunnerState(
            _ptr(self.info['ENV-PAUSE']), _ptr(self.info['Env-Suffered-Reset']), _ptr(self.info['Current-Obs-Step']),
            _ptr(self.info['Thread-Episode-Cnt']), _ptr(self.info['Latest-Reward-Sum']), _ptr(self.counters),
            _ptr(self._fin_slot), _ptr(self.log_meta), _ptr(self.log_reward), self.log_capacity)
        self.init_runner()

    @property
    def pause(self):
        """uint8 [E]: the skip mask of the next env step."""
        return self.info['ENV-PAUSE']

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def init_runner(self):
        _lib.check(self.lib.dca_runner_init(C.byref(self._state), self.E, self.G, self._stream()))
        self.current_n_frame = 0
        self.current_n_episode = 0
        self._reported = 0  # episodes covered by reports so far
        self._tests_done = 0
        self.top_rewards = None
        self.win_cnt_avg = [[] for _ in range(self.n_team)]
        self._exit_early_ = False

    def update(self, done, win, reward):
        """done uint8 [E], win uint8 [E], reward float32 [E,G] of the step just taken (device tensors)."""
        for t, dt, shape in ((done, torch.uint8, (self.E,)), (win, torch.uint8, (self.E,)),
                             (reward, torch.float32, (self.E, self.G))):
            if t.dtype != dt or tuple(t.shape) != shape or not t.is_cuda or not t.is_contiguous():
                raise _lib.DcaError('update: expected contiguous cuda %s %s, got %s %s'
                                    % (dt, shape, t.dtype, tuple(t.shape)))
        _lib.check(self.lib.dca_runner_update(C.byref(self._state), self.E, self.G, int(self.align_episode),
                                              _ptr(done), _ptr(win), _ptr(reward), self._stream()))

    def report_window(self, first_episode, n_episodes):
        """(mean reward per team, win rate) over episodes first_episode .. first_episode+n_episodes-1 (1-based)."""
        _lib.check(self.lib.dca_runner_report(C.byref(self._state), self.G, int(first_episode), int(n_episodes),
                                              _ptr(self._team_of_agent), self.n_team, _ptr(self._report_out),
                                              self._stream()))
        out = self._report_out.cpu().numpy()
        return [float(x) for x in out[:self.n_team]], float(out[self.n_team])

    def poll(self):
        """Read the two global counters back (64 bytes) and emit what the reference's loop would have emitted since
        the last poll: one report per completed report_interval window (in order), the number of test runs that fell
        due, and the exit flag.  Call after every update() to report exactly when the reference does; calling it less
        often only delays the reports (the windows stay the same) as long as fewer than log_capacity - report_interval
        episodes finish in between."""
        self._h_counters.copy_(self.counters, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        self.current_n_frame = int(self._h_counters[0])
        prev, self.current_n_episode = self.current_n_episode, int(self._h_counters[1])
        reports = []
        while self.report_interval > 0 and self.current_n_episode - self._reported >= self.report_interval:
            mean_reward, win_rate = self.report_window(self._reported + 1, self.report_interval)
            self._reported += self.report_interval
            if self.top_rewards is None:
                self.top_rewards = list(mean_reward)
            self.top_rewards = [max(a, b) for a, b in zip(self.top_rewards, mean_reward)]
            # single-team DCA: the win rate belongs to team 0 (task_runner.py:263-267)
            self.win_cnt_avg[0].append(win_rate)
            reports.append({'episode': self._reported, 'mean_reward': mean_reward, 'win_rate': win_rate,
                            'top_reward': list(self.top_rewards),
                            'acc_win_ratio': float(np.mean(self.win_cnt_avg[0]))})
        tests_due = 0
        if self.test_interval:
            tests_due = self.current_n_episode // self.test_interval - prev // self.test_interval
        if self.max_n_episode is not None and self.current_n_episode >= self.max_n_episode:
            self._exit_early_ = True
        return {'current_n_frame': self.current_n_frame, 'current_n_episode': self.current_n_episode,
                'reports': reports, 'tests_due': tests_due, 'exit_early': self._exit_early_}
