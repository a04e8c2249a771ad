"""ORACLE / TEST INFRASTRUCTURE -- fixtures for the conc_4hist policy rollout (BASELINE.json
configs[3] = RESULT/50RL-55opp).  Runs only where the reference is staged (oracle/stage_ref.py; build container).

Everything recorded here is produced by the UNMODIFIED reference code:
  * ALGORITHM/conc_4hist/net.py Net (h_dim 48, n_focus_on 3, dual_conc) with the shipped RESULT/50RL-55opp/model.pt,
  * ALGORITHM/common/norm.py DynamicNorm (running moments, exclude_nan / exclude_one_hot),
  * ALGORITHM/conc_4hist/shell_env.py ShellEnvWrapper.solve_duplicate (+ cython roll_hisory),
  * the DCA env behind SmartPool + SuperpoolEnv + EnvAutoReset (MISSION/dca, 50 guards vs 55 scripted attackers,
    REWARD_DEBUG 3, terrain [0.15, 0.2]: RESULT/50RL-55opp/test-50RL-55opp.jsonc).

  python oracle/make_golden_policy.py net              # deterministic: Net.evaluate path + DynamicNorm updates
  python oracle/make_golden_policy.py rollout 512      # statistical: win rate / episode length over >= 512 episodes

writes tests/golden/policy_net_4hist.npz and tests/golden/policy_rollout_50v55.json.
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
OUT = os.path.join(REPO, 'tests', 'golden')
sys.path.insert(0, HERE)
import stage_ref  # noqa: E402
from ref_pool_bench import SCENARIOS, inject  # noqa: E402

ALG = dict(n_focus_on=3, gamma_in_reward_forwarding=True, gamma_in_reward_forwarding_value=0.95, gamma=0.99,
           ignore_test_mode=True)  # the AlgorithmConfig group of test-50RL-55opp.jsonc that the rollout reads


def setup(n_env, fold, seed):
    stage_ref.activate()
    import numpy as np  # noqa: F401
    from config import GlobalConfig as cfg
    cfg.env_name = 'dca'
    cfg.num_threads = n_env
    cfg.fold = fold
    cfg.seed = seed
    cfg.device = 'cpu'
    cfg.use_float64 = False
    cfg.cfg_ready = True   # what main.py's prepare_args sets once the JSONC groups are injected (UTIL/config_args.py)
    from MISSION.env_router import load_ScenarioConfig
    load_ScenarioConfig()
    inject(cfg.ScenarioConfig, SCENARIOS['50v55'])
    return cfg


def load_net(cfg):
    import torch
    from ALGORITHM.conc_4hist.foundation import AlgorithmConfig
    for k, v in ALG.items():
        setattr(AlgorithmConfig, k, v)
    from ALGORITHM.conc_4hist.net import Net
    net = Net(rawob_dim=cfg.ScenarioConfig.obs_vec_length, n_action=cfg.ScenarioConfig.n_actions,
              use_normalization=AlgorithmConfig.use_normalization, n_focus_on=AlgorithmConfig.n_focus_on,
              actor_attn_mod=AlgorithmConfig.actor_attn_mod, dual_conc=AlgorithmConfig.dual_conc)
    sd = torch.load(os.path.join(stage_ref.TREE, 'RESULT', '50RL-55opp', 'model.pt'), map_location='cpu')
    net.load_state_dict(sd)
    return net


class Shell(object):
    """The history-pool half of ShellEnvWrapper.interact_with_env (shell_env.py:29-45), solve_duplicate unmodified."""

    def __init__(self, cfg):
        from ALGORITHM.conc_4hist.shell_env import ShellEnvWrapper
        self.w = ShellEnvWrapper.__new__(ShellEnvWrapper)
        self.w.core_dim = cfg.ScenarioConfig.obs_vec_length
        self.his = None

    def feed(self, obs, rst):
        import numpy as np
        from UTIL.tensor_ops import my_view
        # the env hands back float64 once anyone is dead (zeros(192) rows are float64); the values are
        # float32 numbers and roll_hisory is typed float32 (conc_4hist/cython_func.pyx:6)
        obs = obs.astype(np.float32)
        if self.his is None:
            self.his = my_view(np.zeros_like(obs), [0, 0, -1, self.w.core_dim])
        self.his[rst] = 0
        obs_feed, nxt = self.w.solve_duplicate(obs.copy(), self.his.copy())
        self.his = nxt
        return obs_feed


def gen_net():
    """Deterministic fixtures: (1) Net._act in eval mode (argmax actions, value, log-prob) and the evaluate path (probs,
    threat) on real observations of a reference rollout; (2) DynamicNorm running-moment updates in training mode."""
    import numpy as np
    import torch
    cfg = setup(n_env=2, fold=1, seed=4242)
    from MISSION.env_router import make_parallel_envs
    from UTIL.shm_pool import SmartPool
    net = load_net(cfg)
    pool = SmartPool(fold=1, proc_num=2, base_seed=4242)
    rec = {}
    try:
        envs = make_parallel_envs(pool)
        obs, _ = envs.reset()
        shell = Shell(cfg)
        rst = np.ones(2, bool)
        feeds = []
        torch.manual_seed(0)
        for t in range(60):
            feed = shell.feed(obs, rst)
            if t % 6 == 0:
                feeds.append(feed.copy())
            with torch.no_grad():
                act, value, logp = net.act(feed, test_mode=False)
            obs, rew, done, info = envs.step(act.astype(np.float64))
            rst = done.copy()
    finally:
        pool.party_over()
    x = np.concatenate(feeds, 0).astype(np.float32)          # [20, 50, 24, 16] with NaN rows / dead agents
    rec['obs_feed'] = x
    sd0 = {k: v.clone() for k, v in net.state_dict().items()}
    net.eval()                                               # DynamicNorm frozen: a pure function of the checkpoint
    with torch.no_grad():
        xt = torch.from_numpy(x)
        act, value, logp = net._act(xt.clone(), test_mode=True)
        v2, logp2, ent, probs, others = net._act(xt.clone(), test_mode=True, eval_mode=True, eval_actions=act)
    rec['eval_norm_mean'] = net._batch_norm.mean.numpy().copy()
    rec['eval_norm_var'] = net._batch_norm.var.numpy().copy()
    rec['eval_act'] = act.numpy().astype(np.int32)
    rec['eval_value'] = value.numpy()[..., 0]
    rec['eval_logp'] = logp.numpy()[..., 0]
    rec['eval_probs'] = probs.numpy()
    rec['eval_threat'] = others['threat'].numpy()[..., 0]
    # DynamicNorm in training mode, from the checkpoint's moments, three batches
    net.load_state_dict(sd0)
    net.train()
    bn = net._batch_norm
    for i in range(3):
        xb = torch.from_numpy(x[i * 5:(i + 1) * 5].copy())
        with torch.no_grad():
            y = bn(xb)
        rec['norm%d_mean' % i] = bn.mean.numpy().copy()
        rec['norm%d_var' % i] = bn.var.numpy().copy()
        rec['norm%d_n' % i] = bn.n_sample.numpy().copy()
        rec['norm%d_filter' % i] = bn.one_hot_filter.numpy().copy()
        rec['norm%d_out' % i] = y.numpy()
    rec['norm_n0'] = sd0['_batch_norm.n_sample'].numpy().copy()
    for k, v in sd0.items():   # the checkpoint itself, as arrays (the GPU box has no reference tree)
        rec['sd/' + k] = v.numpy().copy()
    path = os.path.join(OUT, 'policy_net_4hist.npz')
    np.savez_compressed(path, **rec)
    print('policy_net_4hist %.1f KB; obs_feed %s, NaN rows %.1f%%' % (
        os.path.getsize(path) / 1024, x.shape, 100 * np.isnan(x).any(-1).mean()))


def gen_rollout(n_episodes, procs, fold):
    """Statistical fixture: the shipped policy (sampled actions: ignore_test_mode) against the scripted attackers on the
    reference env; win rate, episode length and per-agent reward sums over the first n_episodes finished episodes."""
    import numpy as np
    import torch
    E = procs * fold
    cfg = setup(n_env=E, fold=fold, seed=22333)
    from MISSION.env_router import make_parallel_envs
    from UTIL.shm_pool import SmartPool
    net = load_net(cfg)   # training flag on, as in the reference run (the policy is never .eval()'d): norm keeps updating
    pool = SmartPool(fold=fold, proc_num=procs, base_seed=22333)
    wins, lens, rsum = [], [], []
    t0 = time.time()
    try:
        envs = make_parallel_envs(pool)
        obs, _ = envs.reset()
        shell = Shell(cfg)
        rst = np.ones(E, bool)
        cur_len = np.zeros(E, int)
        cur_r = np.zeros(E)
        torch.manual_seed(1)
        while len(wins) < n_episodes:
            feed = shell.feed(obs, rst)
            with torch.no_grad():
                act, value, logp = net.act(feed, test_mode=False)
            obs, rew, done, info = envs.step(act.astype(np.float64))
            cur_len += 1
            cur_r += rew.mean(-1)
            for e in np.nonzero(done)[0]:
                wins.append(bool(info[e]['win'])); lens.append(int(cur_len[e])); rsum.append(float(cur_r[e]))
                cur_len[e] = 0; cur_r[e] = 0
            rst = done.copy()
            if len(wins) and len(wins) % 64 < done.sum():
                print('%d episodes, win rate %.3f, mean len %.1f, %.0f s' % (
                    len(wins), np.mean(wins), np.mean(lens), time.time() - t0), flush=True)
    finally:
        pool.party_over()
    wins, lens, rsum = wins[:n_episodes], lens[:n_episodes], rsum[:n_episodes]
    out = {'what': 'RESULT/50RL-55opp/model.pt (conc_4hist, sampled actions) vs scripted attackers on the unmodified '
                   'reference env (SmartPool %d procs x fold %d); first %d finished episodes' % (procs, fold, n_episodes),
           'episodes': len(wins), 'win_rate': float(np.mean(wins)), 'mean_len': float(np.mean(lens)),
           'std_len': float(np.std(lens)), 'mean_reward_sum': float(np.mean(rsum)), 'std_reward_sum': float(np.std(rsum)),
           'seconds': time.time() - t0}
    with open(os.path.join(OUT, 'policy_rollout_50v55.json'), 'w') as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == '__main__':
    what = sys.argv[1] if len(sys.argv) > 1 else 'net'
    if what == 'net':
        gen_net()
    else:
        n = int(sys.argv[2]) if len(sys.argv) > 2 else 512
        procs = int(sys.argv[3]) if len(sys.argv) > 3 else (os.cpu_count() or 1)
        gen_rollout(n, procs, fold=int(sys.argv[4]) if len(sys.argv) > 4 else 4)
