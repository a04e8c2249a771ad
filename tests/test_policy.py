"""conc_4hist policy rollout on device tensors (BASELINE.json configs[3]): the normalisation
kernels, the network and the rollout loop against fixtures produced by the unmodified reference
(oracle/make_golden_policy.py: ALGORITHM/conc_4hist/net.py Net + ALGORITHM/common/norm.py DynamicNorm with the shipped
RESULT/50RL-55opp/model.pt, on observations of the reference env)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import golden_util as gu

pytestmark = pytest.mark.gpu


def _fixture():
    return np.load(os.path.join(gu.GOLDEN_DIR, 'policy_net_4hist.npz'))


def _net(z, torch):
    from hmp2g_b200.policy import ConcNet4Hist
    net = ConcNet4Hist(rawob_dim=16, n_action=7, n_focus_on=3)
    sd = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith('sd/')}
    net.load_state_dict(sd)   # strict: same names, shapes and dtypes as the reference's state_dict
    return net.cuda()


def test_norm_moments_and_apply_match_numpy():
    import torch
    from hmp2g_b200.policy import DynamicNormDevice
    z = _fixture()
    x = z['obs_feed']
    bn = DynamicNormDevice(16).cuda()
    xt = torch.from_numpy(x).cuda()
    stats = bn.batch_stats(xt).cpu().numpy()
    rows = x.reshape(-1, 16)
    ok = ~np.isnan(rows).any(-1)
    r64 = rows[ok].astype(np.float64)
    np.testing.assert_allclose(stats[:16], r64.sum(0), rtol=1e-12, atol=1e-9)
    np.testing.assert_allclose(stats[16:32], (r64 * r64).sum(0), rtol=1e-12, atol=1e-9)
    assert np.array_equal(stats[32:48], ((r64 != 0) & (r64 != 1) & (r64 != -1)).sum(0))
    assert stats[48] == ok.sum()
    again = bn.batch_stats(xt).cpu().numpy()
    assert np.array_equal(stats.view(np.uint64), again.view(np.uint64))  # fixed summation order
    # apply with the checkpoint's moments: float32 arithmetic like torch's
    mean, var = z['eval_norm_mean'], z['eval_norm_var']
    bn.mean.data.copy_(torch.from_numpy(mean)); bn.var.data.copy_(torch.from_numpy(var))
    y, mask = bn.apply_norm(xt)
    want = np.clip((rows - mean) / np.sqrt(var + np.float32(1e-8)), -10, 10)
    want_mask = np.isnan(rows).any(-1)
    assert np.array_equal(mask.cpu().numpy().reshape(-1).astype(bool), want_mask)
    got = y.cpu().numpy().reshape(-1, 16)
    assert np.array_equal(got[want_mask], np.zeros_like(got[want_mask]))
    assert np.array_equal(got[~want_mask].view(np.uint32), want[~want_mask].astype(np.float32).view(np.uint32))


def test_dynamic_norm_updates_match_reference():
    """DynamicNorm in training mode from the checkpoint's moments over three batches (norm.py:47-63): running mean /
    variance / sample count / one-hot filter after each batch, and the normalised batch."""
    import torch
    z = _fixture()
    net = _net(z, torch)
    bn = net._batch_norm
    bn.train()
    x = z['obs_feed']
    assert int(bn.n_sample.item()) == int(z['norm_n0'][0])
    for i in range(3):
        xb = torch.from_numpy(x[i * 5:(i + 1) * 5].copy()).cuda()
        y, mask = bn(xb)
        np.testing.assert_allclose(bn.mean.cpu().numpy(), z['norm%d_mean' % i], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(bn.var.cpu().numpy(), z['norm%d_var' % i], rtol=1e-5, atol=1e-6)
        assert int(bn.n_sample.item()) == int(z['norm%d_n' % i][0])
        assert np.array_equal(bn.one_hot_filter.cpu().numpy(), z['norm%d_filter' % i])
        want = z['norm%d_out' % i]
        got = y.cpu().numpy()
        nan = np.isnan(want)
        assert np.array_equal(mask.cpu().numpy().astype(bool), nan.any(-1))
        np.testing.assert_allclose(got[~nan], want[~nan], rtol=1e-5, atol=1e-5)
        assert (got[nan] == 0).all()


def test_net_matches_reference_net_on_reference_observations():
    """ConcNet4Hist with the shipped checkpoint vs the reference Net (eval mode: frozen normalisation) on 1000 agents'
    observations from a reference rollout: action probabilities to 1e-4, value / threat to 2e-3, arg-max actions equal wherever
    the reference's top two probabilities are more than 1e-4 apart."""
    import torch
    z = _fixture()
    net = _net(z, torch).eval()
    x = torch.from_numpy(z['obs_feed']).cuda()
    logits, value, threat = net.heads(x)
    probs = torch.softmax(logits, -1).cpu().numpy()
    # value / threat ride on the attention weights, a float32 softmax over logits in the hundreds (the checkpoint's
    # attention is nearly one-hot): a 1e-3 absolute difference in a logit -- what the reference's own CPU and CUDA runs
    # differ by -- moves a weight by 1e-3 relative.  The action head only depends on WHICH entities are picked.
    np.testing.assert_allclose(value.cpu().numpy(), z['eval_value'], rtol=2e-3, atol=2e-3)
    np.testing.assert_allclose(probs, z['eval_probs'], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(threat.cpu().numpy(), z['eval_threat'], rtol=2e-3, atol=2e-3)
    act, v2, logp = net.act(x, test_mode=True)
    top2 = np.sort(z['eval_probs'], -1)[..., -2:]
    clear = (top2[..., 1] - top2[..., 0]) > 1e-4
    assert clear.mean() > 0.9
    assert np.array_equal(act.cpu().numpy()[clear], z['eval_act'][clear])
    np.testing.assert_allclose(logp.cpu().numpy()[clear], z['eval_logp'][clear], rtol=1e-4, atol=1e-4)
    # sampling: frequencies follow the probabilities (one agent's distribution, many draws through the batch axis)
    one = x[:1, :1].expand(400, 50, 24, 16).contiguous()
    g = torch.Generator(device='cuda'); g.manual_seed(5)
    a, _, _ = net.act(one, test_mode=False, generator=g)
    freq = np.bincount(a.cpu().numpy().reshape(-1), minlength=7) / a.numel()
    assert np.abs(freq - probs[0, 0]).max() < 0.02


def test_encoder_normalisation_is_bit_identical_to_the_division_kernel():
    """The normalised rows and row masks that dca_policy_encode writes next to the encodings must agree bit for bit with
    dca_norm_apply (the stand-alone DynamicNorm apply kernel, pinned against the reference's fixtures): on the reference
    observations and on random rows / random moments (tiny and huge spreads, NaN rows, a constant feature)."""
    import torch
    from hmp2g_b200.policy import ConcNet4Hist
    z = _fixture()
    net = _net(z, torch).eval()
    g = torch.Generator(device='cuda'); g.manual_seed(11)
    feeds = [torch.from_numpy(z['obs_feed']).cuda()]
    for scale in (1e-3, 1.0, 37.0):
        xb = torch.randn((64, 50, 24, 16), device='cuda', generator=g) * scale
        xb[torch.rand((64, 50, 24), device='cuda', generator=g) < 0.3] = float('nan')
        xb[..., 5] = 0.0   # a constant feature: (0 - mean) / sd and exact zeros
        feeds.append(xb)
    bn = net._batch_norm
    for i, x in enumerate(feeds):
        if i:
            bn.mean.copy_(torch.randn(16, device='cuda', generator=g) * (10.0 ** (i - 2)))
            bn.var.copy_(torch.rand(16, device='cuda', generator=g) * (10.0 ** (2 * i - 4)) + 1e-12)
        xn, dead, _ = net.encode_fused(x.contiguous())
        want, mask = bn.apply_norm(x.contiguous())
        assert torch.equal(xn.reshape(-1).view(torch.int32), want.reshape(-1).view(torch.int32)), i
        assert torch.equal(dead.reshape(-1), mask.reshape(-1)), i


def test_policy_linear_kernel_matches_float64_reference():
    """dca_policy_linear: act(a Wa^T + b Wb^T + bias) for the layer shapes of the policy -- ragged row counts, inputs
    that are column slices of wider tensors, weight halves and transposed weights, N = 48 / 24 / 7 / 1."""
    import torch
    from hmp2g_b200.policy import _plin
    g = torch.Generator(device='cuda')
    g.manual_seed(7)
    rnd = lambda *s: torch.randn(*s, generator=g, device='cuda', dtype=torch.float32)
    for B in (1, 127, 1000, 4133):
        wide = rnd(B, 256)
        va = rnd(B, 48)
        cases = [
            (wide[:, :48], rnd(48, 48).t(), None, None, None, False),                       # x @ M
            (wide[:, 64:112], rnd(48, 96)[:, :48], va, rnd(48, 96)[:, 48:], rnd(48), True),  # cat((a, b)) -> Linear, ReLU
            (wide, rnd(48, 256), None, None, rnd(48), True),                                # forward layer, K = 256
            (wide[:, :192], rnd(48, 192), None, None, rnd(48), True),                       # K = 192
            (va, rnd(24, 48), None, None, rnd(24), True),                                   # 48 -> 24
            (wide[:, 8:32], rnd(7, 24), None, None, rnd(7), False),                         # 24 -> 7
            (va, rnd(1, 48), None, None, rnd(1), False),                                    # 48 -> 1
        ]
        for i, (a, wa, b, wb, bias, relu) in enumerate(cases):
          for ffma in (False, True):   # 3xTF32 tensor-core path (default) and the float32 FFMA path
              got = _plin(a, wa, b, wb, bias, relu, ffma=ffma)
              want = a.double() @ wa.double().t()
              if b is not None:
                  want = want + b.double() @ wb.double().t()
              if bias is not None:
                  want = want + bias.double()
              if relu:
                  want = want.clamp_min(0)
              scale = float(want.abs().max()) + 1.0
              assert got.shape == want.shape
              err = float((got.double() - want).abs().max())
              assert err < 4e-6 * scale, (B, i, ffma, err, scale)   # float32-class: a single tf32 pass would be ~1e-3


def test_fused_attention_kernel_matches_the_torch_path():
    """dca_conc_attend (scores, masked softmax, attended encoding, stable top-k gather in one kernel) against the same
    steps written with torch ops (bmm / softmax / stable argsort / gather / cat): identical picks, values to float32
    rounding -- on the reference observations and on a large synthetic batch with every masking pattern (no live entity,
    fewer live entities than n_focus, dead agents)."""
    import torch
    from hmp2g_b200.policy import ConcNet4Hist
    z = _fixture()
    net = _net(z, torch).eval()
    g = torch.Generator(device='cuda'); g.manual_seed(7)
    big = torch.randn((512, 50, 24, 16), device='cuda', generator=g) * 2
    rows = torch.rand((512, 50, 24), device='cuda', generator=g)
    big[rows < 0.45] = float('nan')
    big[(torch.rand((512, 50), device='cuda', generator=g) < 0.1)] = float('nan')      # dead agents
    few = torch.rand((512, 50), device='cuda', generator=g) < 0.15                      # at most one live entity per list
    keep = torch.zeros((24,), dtype=torch.bool, device='cuda'); keep[0] = keep[3] = keep[14] = True
    sel = few[:, :, None] & ~keep[None, None, :]
    big[sel] = float('nan')
    for x in (torch.from_numpy(z['obs_feed']).cuda(), big):
        ConcNet4Hist.fused = True
        try:
            lf, vf, tf = net.heads(x)
        finally:
            ConcNet4Hist.fused = False
        try:
            lt, vt, tt = net.heads(x)
        finally:
            ConcNet4Hist.fused = True
        ConcNet4Hist.attend_both = False   # the one-list attention kernel, twice: bit-identical to the two-list kernel
        try:
            l1, v1, t1 = net.heads(x)
        finally:
            ConcNet4Hist.attend_both = True
        assert torch.equal(l1, lf) and torch.equal(v1, vf) and torch.equal(t1, tf)
        # picks can only differ where the two softmax implementations disagree on whether a score underflows to exactly 0
        bad = ((lf - lt).abs() > 1e-4 + 1e-4 * lt.abs()).any(-1).float().mean().item()
        print('fused vs torch: %.5f%% of agents with a different pick' % (100 * bad))
        assert bad < 1e-3
        assert torch.allclose(vf, vt, rtol=1e-3, atol=1e-3), float((vf - vt).abs().max())
        assert torch.allclose(tf, tt, rtol=1e-3, atol=1e-3), float((tf - tt).abs().max())


def test_policy_rollout_50v55_statistics_match_reference_run():
    """BASELINE configs[3] end to end on the device: the shipped conc_4hist policy (sampled actions) against the scripted
    attackers, 256 envs x their first two episodes = 512 episodes, vs the same policy on the unmodified reference env
    (tests/golden/policy_rollout_50v55.json, 512 episodes): win rate and mean episode length within sampling error.
    The GAE pass runs on every finished trajectory with the configuration's gamma_in_reward_forwarding (0.95)."""
    import torch
    from hmp2g_b200.config import ScenarioConfig
    from hmp2g_b200.policy import PolicyRollout
    from hmp2g_b200.vec_env import DcaVecEnv
    ref = json.load(open(os.path.join(gu.GOLDEN_DIR, 'policy_rollout_50v55.json')))
    z = _fixture()
    net = _net(z, torch).train()   # the reference never calls .eval(): the normalisation keeps updating (norm.py:47)
    scfg = type('S', (ScenarioConfig,), dict(num_guards=50, num_attackers=55, size=5.0, random_jam_prob=0.05,
                                             introduce_terrain=True, terrain_parameters=[0.15, 0.2], MaxEpisodeStep=180,
                                             half_death_reward=True, REWARD_DEBUG=True, REWARD_DEBUG_Value=3))
    E, per_env = 256, 2
    env = DcaVecEnv(E, scfg=scfg, seed=22333, empty_rows_nan=True)
    roll = PolicyRollout(env, net, gamma=0.99, tau=0.95, gamma_fwd=0.95, seed=1)
    ep = np.zeros(E, int); cur = np.zeros(E, int); cur_r = np.zeros(E)
    lens, wins, rsum = [], [], []
    n_fin_traj = 0
    for t in range(per_env * 181 + 2):
        fin = roll.step()
        f = fin.cpu().numpy(); w = env.win.cpu().numpy()
        cur += 1
        cur_r += env.reward.mean(1).cpu().numpy()   # per-agent reward of the step, averaged over the guards
        for e in np.nonzero(f)[0]:
            if ep[e] < per_env:
                lens.append(cur[e]); wins.append(bool(w[e])); rsum.append(cur_r[e])
            ep[e] += 1; cur[e] = 0; cur_r[e] = 0
        if f.any():
            ret, threat, adv, valid = roll.buf.finalize(fin)
            assert torch.isfinite(ret[valid]).all() and torch.isfinite(adv[valid]).all()
            n_fin_traj += int(f.sum())
        if (ep >= per_env).all():
            break
    assert len(lens) == E * per_env and n_fin_traj >= E * per_env
    assert int(roll.episodes.item()) == n_fin_traj
    win_rate, mean_len = float(np.mean(wins)), float(np.mean(lens))
    print('policy rollout: %d episodes, win rate %.3f (reference %.3f), mean length %.1f +- %.1f (reference %.1f +- %.1f)'
          % (len(lens), win_rate, ref['win_rate'], mean_len, np.std(lens), ref['mean_len'], ref['std_len']))
    n_ref = ref['episodes']
    p = 0.5 * (win_rate + ref['win_rate'])
    se_p = np.sqrt(max(p * (1 - p), 1e-4) * (1.0 / len(lens) + 1.0 / n_ref))
    assert abs(win_rate - ref['win_rate']) <= 4 * se_p + 0.01
    se_len = np.sqrt(np.var(lens) / len(lens) + ref['std_len'] ** 2 / n_ref)
    assert abs(mean_len - ref['mean_len']) <= 4 * se_len + 2.0
    se_r = np.sqrt(np.var(rsum) / len(rsum) + ref['std_reward_sum'] ** 2 / n_ref)
    print('mean reward sum %.3f +- %.3f (reference %.3f +- %.3f)' % (np.mean(rsum), np.std(rsum), ref['mean_reward_sum'],
                                                                    ref['std_reward_sum']))
    assert abs(np.mean(rsum) - ref['mean_reward_sum']) <= 4 * se_r + 0.05


def test_policy_rollout_step_as_cuda_graph_equals_eager_steps():
    """PolicyRollout.capture(): one env step of the whole loop (history roll, normalisation update, network, draw,
    planner, step, rollout buffer) replayed as a CUDA graph gives the same trajectory as eager steps -- same seeds, test
    mode (arg-max actions) so both arms take identical actions.  The eager arm plans inside env.step on the one stream;
    the captured arm runs the planner on its side stream under the policy forward (plan_ahead), so the test also pins
    that fork / join."""
    import torch
    from hmp2g_b200.config import ScenarioConfig
    from hmp2g_b200.policy import PolicyRollout
    from hmp2g_b200.vec_env import DcaVecEnv
    z = _fixture()
    scfg = type('S', (ScenarioConfig,), dict(num_guards=50, num_attackers=55, size=5.0, random_jam_prob=0.05,
                                             introduce_terrain=True, terrain_parameters=[0.15, 0.2], MaxEpisodeStep=40,
                                             half_death_reward=True, REWARD_DEBUG=True, REWARD_DEBUG_Value=3))
    arms = []
    for graph in (False, True):
        net = _net(z, torch).train()
        # (the captured arm with DCA_OPT_PIPELINE_STEPS on: launches on a capturing stream fall back to ordinary ones)
        env = DcaVecEnv(64, scfg=scfg, seed=5, empty_rows_nan=True, pipeline_steps=graph)
        roll = PolicyRollout(env, net, gamma_fwd=0.95, seed=1, test_mode=True, plan_ahead=graph)
        step = roll.capture(warm=3) if graph else None
        if not graph:
            for _ in range(3):
                roll.step()
        n_fin = 0
        for t in range(50):
            fin = step() if graph else roll.step()
            n_fin += int(fin.sum().item())
        torch.cuda.synchronize()
        arms.append((env.obs.clone(), env.reward.clone(), env.alive.clone(), roll.buf.reward.clone(), roll.buf.value.clone(),
                     roll.buf.length.clone(), int(roll.episodes.item()), net._batch_norm.mean.clone(),
                     int(net._batch_norm.n_sample.item()), n_fin))
    a, b = arms
    assert a[6] == b[6] and a[6] >= 64 and a[8] == b[8] and a[9] == b[9]
    assert torch.equal(a[0].view(torch.int32), b[0].view(torch.int32))  # observations carry NaN rows: compare bits
    for i in (1, 2, 3, 5):
        assert torch.equal(a[i], b[i]), i
    assert torch.allclose(a[4], b[4], rtol=1e-5, atol=1e-5) and torch.allclose(a[7], b[7], rtol=1e-6, atol=1e-7)
