"""CPU tests of the host side: config injection, the C-ABI surface, product/oracle separation, sharding (gloo x2)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

JSONC = r'''
{
    // shipped RESULT/test-50+50 style config (stale module + key names)
    "config.py->GlobalConfig": { "note": "x", "env_name": "collective_assault", "num_threads": "16", "fold": "1", },
    "MISSION.collective_assault.collective_assault_parallel_run.py->ScenarioConfig": {
        "size": "5", /* block
        comment */ "random_jam_prob": 0.05, "introduce_terrain": "True",
        "terrain_parameters": [0.05,0.2], "num_steps": "180", "half_death_reward": "True",
        "num_guards": 60, "render": "False",
        "TEAM_NAMES": [ "ALGORITHM.conc.foundation->ReinforceAlgorithmFoundation" ],
    },
    "ALGORITHM.conc.foundation.py->AlgorithmConfig": { "lr": 0.0005, "url": "http://a//b" }
}
'''


@pytest.fixture()
def fresh_cfg():
    from hmp2g_b200 import config
    snap_s = config.reset_defaults(config.ScenarioConfig)
    snap_g = config.reset_defaults(config.GlobalConfig)
    yield config
    config.reset_defaults(config.ScenarioConfig, snap_s)
    config.reset_defaults(config.GlobalConfig, snap_g)


def test_jsonc_injection_coercion_and_chainvars(fresh_cfg):
    import json
    cfg = fresh_cfg
    data = json.loads(cfg.strip_jsonc(JSONC))
    assert data['ALGORITHM.conc.foundation.py->AlgorithmConfig']['url'] == 'http://a//b'
    rep = cfg.inject(data)
    S = cfg.ScenarioConfig
    assert S.size == 5.0 and isinstance(S.size, float)
    assert S.introduce_terrain is True and S.half_death_reward is True and S.render is False
    assert S.MaxEpisodeStep == 180 and ('num_steps', 'MaxEpisodeStep') in rep.renamed
    assert S.num_guards == 60
    # chained twins follow num_guards (UTIL/config_args.py:121-144)
    assert S.N_AGENT_EACH_TEAM == [60] and S.AGENT_ID_EACH_TEAM == [range(0, 60)]
    assert S.num_object == 110 and S.uid_dictionary['attackers_uid'] == range(60, 110)
    assert S.num_entity == 50  # not chained to num_guards
    assert set(rep.altered_chained) == {'N_AGENT_EACH_TEAM', 'AGENT_ID_EACH_TEAM', 'num_object', 'uid_dictionary'}
    assert cfg.GlobalConfig.env_name == 'dca' and cfg.GlobalConfig.num_threads == 16
    assert rep.ignored_groups == ['ALGORITHM.conc.foundation.py->AlgorithmConfig']


def test_injection_errors_like_reference(fresh_cfg):
    cfg = fresh_cfg
    key = 'MISSION.dca.collective_assault_parallel_run.py->ScenarioConfig'
    with pytest.raises(AssertionError):
        cfg.inject({key: {'no_such_item': 1}})           # unknown key (config_args.py:285)
    with pytest.raises(AssertionError):
        cfg.inject({key: {'num_guards': 50.5}})          # int default, float override (:299)
    with pytest.raises(AssertionError):
        cfg.inject({key: {'introduce_terrain': 'yes'}})  # bool must be True/False (:296)
    with pytest.raises(AssertionError):
        cfg.inject({key: {'terrain_parameters': 3}})     # list default needs a list (:305)
    cfg.inject({key: {'num_guards': '40', 'N_AGENT_EACH_TEAM': [7], 'REWARD_DEBUG_Value': 3}})
    assert cfg.ScenarioConfig.num_guards == 40
    assert cfg.ScenarioConfig.N_AGENT_EACH_TEAM == [7]   # explicitly overridden twin is not recomputed (:127)


def _declared_functions():
    names = set()
    for fn in os.listdir(os.path.join(REPO, 'include')):
        if fn.endswith('.h'):
            text = open(os.path.join(REPO, 'include', fn)).read()
            text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
            names |= set(re.findall(r'\b(dca_[a-z_0-9]+)\s*\(', text))
    return names


def test_c_abi_library_exports_every_declared_symbol():
    from hmp2g_b200 import _lib, build
    build.build()
    lib = ctypes.CDLL(build.LIB_PATH)
    declared = _declared_functions()
    assert len(declared) >= 12 and declared == set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert _lib.load().dca_abi_version() == 2


def test_abi_struct_layout_matches_header():
    """ctypes mirrors vs the sizes the C compiler gives include/dca_b200.h."""
    from hmp2g_b200 import _lib
    src = '#include <stdio.h>\n#include "dca_b200.h"\nint main(){printf("%zu %zu %zu %zu\\n", sizeof(DcaConfig), sizeof(DcaBuffers), sizeof(DcaState), sizeof(DcaRunnerState));return 0;}'
    exe = os.path.join(REPO, 'build', 'abi_sizes')
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    subprocess.run(['gcc', '-x', 'c', '-', '-I', os.path.join(REPO, 'include'), '-o', exe], input=src.encode(), check=True)
    sizes = [int(x) for x in subprocess.check_output([exe]).split()]
    assert sizes == [ctypes.sizeof(_lib.DcaConfig), ctypes.sizeof(_lib.DcaBuffers), ctypes.sizeof(_lib.DcaState),
                     ctypes.sizeof(_lib.DcaRunnerState)]


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    from hmp2g_b200 import _lib
    from hmp2g_b200.vec_env import DcaVecEnv
    with pytest.raises(_lib.DcaError):
        DcaVecEnv(4)
    lib = _lib.load()
    h = ctypes.c_void_p()
    cfg = _lib.DcaConfig()
    cfg.num_guards = cfg.num_attackers = 50
    assert lib.dca_create(ctypes.byref(cfg), 4, 0, ctypes.byref(h)) == -4  # DCA_E_NODEVICE
    assert b'no CPU fallback' in lib.dca_last_error(None)


def test_policy_entry_points_reject_bad_arguments_before_any_launch():
    """Argument validation of the policy kernels' entry points happens on the host, before a device is touched: null
    buffers, empty batches, widths the kernels do not cover, misaligned inputs -> DCA_E_INVALID with a message."""
    from hmp2g_b200 import _lib
    lib = _lib.load()
    P = ctypes.c_void_p
    buf = (ctypes.c_float * 64)()
    a = P(ctypes.addressof(buf))
    inv = -1  # DCA_E_INVALID
    lin = lambda **kw: lib.dca_policy_linear(kw.get('a', a), kw.get('lda', 48), kw.get('ka', 48), kw.get('wa', a), 48, 1,
                                             kw.get('b', None), kw.get('ldb', 0), kw.get('kb', 0), kw.get('wb', None), 0, 0,
                                             None, kw.get('n', 48), 1, kw.get('rows', 1), kw.get('out', a), kw.get('ldo', 48), None)
    assert lin(a=None) == inv and lin(out=None) == inv and lin(rows=0) == inv
    assert lin(n=49) == inv and lin(ka=46) == inv and lin(lda=50) == inv and lin(ldo=40) == inv
    assert lin(kb=48) == inv                       # a second input without its buffers
    assert lin(a=P(ctypes.addressof(buf) + 4)) == inv   # inputs are read 16 bytes at a time
    assert b'dca_policy_linear' in lib.dca_last_error(None)
    assert lib.dca_policy_encode(None, a, a, a, a, a, a, 1, a, a, a, None) == inv
    assert lib.dca_conc_attend(a, a, a, a, 0, 11, 2, 1, a, a, None) == inv   # entity 0 is the agent itself
    assert lib.dca_norm_moments(None, 1, 16, a, a, None) == inv


def test_product_never_touches_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use oracle/."""
    pat = re.compile(r'(import|from)\s+\S*oracle|dca_oracle|oracle/|libdca_oracle|\borc_[a-z]+')
    for root, _, files in os.walk(os.path.join(REPO, 'hmp2g_b200')):
        for fn in files:
            if fn.endswith(('.py', '.cu', '.cuh', '.h')):
                m = pat.search(open(os.path.join(root, fn)).read())
                assert m is None, (fn, m.group(0))


def test_shard_envs_partitions():
    from hmp2g_b200.sharding import shard_envs
    for total, world in [(32768, 8), (4096, 1), (10, 3), (7, 7)]:
        spans = [shard_envs(total, r, world) for r in range(world)]
        assert spans[0][0] == 0 and sum(n for _, n in spans) == total
        for (g0, n0), (g1, _) in zip(spans, spans[1:]):
            assert g0 + n0 == g1


_WORKER = r'''
import os, sys, torch
sys.path.insert(0, %r)
import torch.distributed as dist
from hmp2g_b200.sharding import init_distributed, shard_envs, RolloutStats
from hmp2g_b200.gae import allreduce_moments, moments_mean_std
rank, world, _ = init_distributed('gloo')
gid0, n = shard_envs(10, rank, world)
g = torch.Generator().manual_seed(0)
reward = torch.randn(10, 5, generator=g); done = (torch.rand(10, generator=g) < 0.5).to(torch.uint8)
win = (torch.rand(10, generator=g) < 0.5).to(torch.uint8)
ret = torch.randn(10, 5, generator=g); val = torch.randn(10, 5, generator=g)
st = RolloutStats('cpu'); sl = slice(gid0, gid0 + n)
st.update(reward[sl], done[sl], win[sl])
tot = st.reduce()
ref = RolloutStats('cpu'); ref.update(reward, done, win)
for k, v in zip(ref.FIELDS, ref.buf.tolist()):
    assert abs(v - tot[k]) < 1e-9, (k, v, tot[k])
a = (ret[sl] - val[sl]).double()     # what dca_adv_moments accumulates on each rank's shard
stats = allreduce_moments(torch.stack([a.sum(), (a * a).sum(), torch.tensor(float(a.numel()), dtype=torch.float64)]))
mean, std = moments_mean_std(stats)
full = (ret - val).double()          # ppo.py:53-55 over the pooled batch (numpy std: ddof 0)
assert abs(float(mean) - float(full.mean())) < 1e-12 and abs(float(std) - float(full.std(unbiased=False))) < 1e-12
dist.barrier(); dist.destroy_process_group()
sys.stdout.write('rank' + str(rank) + 'ok\n'); sys.stdout.flush()
'''


def test_two_rank_gloo_stats_and_advantage_moments(tmp_path):
    script = tmp_path / 'worker.py'
    script.write_text(_WORKER % REPO)
    env = dict(os.environ, MASTER_ADDR='127.0.0.1', OMP_NUM_THREADS='1')
    port = 29600 + os.getpid() % 300
    out = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2',
                          '--master-addr', '127.0.0.1', '--master-port', str(port), str(script)],
                         env=env, capture_output=True, text=True, timeout=280)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.count('ok') == 2, out.stdout


def test_multiteam_jsonc_injection_and_chainvars():
    """The dca_multiteam ScenarioConfig group (MISSION/dca_multiteam/collective_assault_parallel_run.py:27-77) is
    injected with the reference's rules: coercion to the default's type, chained AGENT_ID_EACH_TEAM recomputed."""
    from hmp2g_b200 import config
    from hmp2g_b200.multiteam import MtScenarioConfig, pod_mt_config
    snap = config.reset_defaults(MtScenarioConfig)
    try:
        rep = config.inject({
            'MISSION.dca_multiteam.collective_assault_parallel_run.py->ScenarioConfig': {
                'N_AGENT_EACH_TEAM': [7, 9, 8], 'random_jam_prob': 0.2, 'introduce_terrain': True,
                'terrain_parameters': [0.05, 0.2], 'MaxEpisodeStep': 45, 'StepsToNarrowCircle': 15, 'size': 3},
            'ALGORITHM.conc_4hist.foundation.py->AlgorithmConfig': {'train_traj_needed': 64},
        })
        assert rep.applied and rep.ignored_groups == ['ALGORITHM.conc_4hist.foundation.py->AlgorithmConfig']
        assert MtScenarioConfig.N_AGENT_EACH_TEAM == [7, 9, 8]
        assert [list(r) for r in MtScenarioConfig.AGENT_ID_EACH_TEAM] == [list(range(0, 7)), list(range(7, 16)),
                                                                         list(range(16, 24))]
        assert isinstance(MtScenarioConfig.size, float) and MtScenarioConfig.size == 3.0
        pod = pod_mt_config(MtScenarioConfig, seed=5, env_gid0=11)
        assert (pod.n_team, list(pod.team_size)[:3], pod.max_episode_step, pod.steps_to_narrow) == (3, [7, 9, 8], 45, 15)
        assert (pod.terrain_A, pod.terrain_B, pod.seed, pod.env_gid0) == (0.05, 0.2, 5, 11)
        with pytest.raises(Exception):
            config.inject({'hmp2g_b200.multiteam.py->MtScenarioConfig': {'no_such_key': 1}})
    finally:
        config.reset_defaults(MtScenarioConfig, snap)


def test_dynamic_norm_merge_matches_reference_moments_on_cpu():
    """The host half of DynamicNormDevice -- merging batch statistics (sum, sum of squares, non-unit counts, rows) into
    the running moments, ALGORITHM/common/norm.py:47-63 -- against the reference's DynamicNorm over three batches
    (tests/golden/policy_net_4hist.npz); the statistics themselves are formed with numpy here (on the device:
    dca_norm_moments), split in two halves and summed like an all-reduce over two ranks would."""
    import torch
    from hmp2g_b200.policy import DynamicNormDevice
    z = np.load(os.path.join(REPO, 'tests', 'golden', 'policy_net_4hist.npz'))
    bn = DynamicNormDevice(16)
    bn.load_state_dict({k[len('sd/_batch_norm.'):]: torch.from_numpy(z[k]) for k in z.files if k.startswith('sd/_batch_norm.')})
    x = z['obs_feed']

    def stats_of(rows):
        rows = rows.reshape(-1, 16)
        r = rows[~np.isnan(rows).any(-1)].astype(np.float64)
        return np.concatenate([r.sum(0), (r * r).sum(0), ((r != 0) & (r != 1) & (r != -1)).sum(0), [r.shape[0]]])
    for i in range(3):
        xb = x[i * 5:(i + 1) * 5]
        stats = stats_of(xb[:2]) + stats_of(xb[2:])   # two "ranks"
        bn.merge(torch.from_numpy(stats))
        np.testing.assert_allclose(bn.mean.numpy(), z['norm%d_mean' % i], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(bn.var.numpy(), z['norm%d_var' % i], rtol=1e-5, atol=1e-6)
        assert int(bn.n_sample.item()) == int(z['norm%d_n' % i][0])
        assert np.array_equal(bn.one_hot_filter.numpy(), z['norm%d_filter' % i])
