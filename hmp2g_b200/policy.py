"""The conc_4hist policy rollout on device tensors (BASELINE.json configs[3]).

Host-side mirror of what the reference runs between two env steps when a conc_4hist team drives DCA
(paths relative to /root/reference):

  ShellEnvWrapper.interact_with_env   ALGORITHM/conc_4hist/shell_env.py:25-99   -> hmp2g_b200.rollout.HistoryPool (k_histpool)
  Net._act / Net.act                  ALGORITHM/conc_4hist/net.py:163-233       -> ConcNet4Hist.act (below)
  DynamicNorm.forward                 ALGORITHM/common/norm.py:31-69            -> DynamicNormDevice (dca_norm_moments /
                                                                                    dca_norm_apply + the merge)
  Concentration.forward               ALGORITHM/common/conc.py:28-63            -> ConcCore
  Categorical(logits).sample()        ALGORITHM/conc_4hist/net.py:141-153       -> on the device, no numpy hop
  trajectory / GAE                    ALGORITHM/conc_4hist/trajectory.py         -> hmp2g_b200.rollout.DeviceRollout (k_gae)

The network is the reference's torch network: same parameter names, shapes and dtypes, so a shipped checkpoint
(`RESULT/50RL-55opp/model.pt`, `torch.save(policy.state_dict())` foundation.py:171) loads with `load_state_dict`.  The
module definitions are this repo's own: the attention scores and the attended value are formed without projecting every
entity (q' = (vs Wq) Wk^T, compat = ve . q'; Va = (score ve) Wv -- the same numbers up to float32 rounding), which cuts
two of the three [rows, 48] x [48, 48] products over the 24 entities of every agent.  PyTorch is the policy's home in the
reference too (cuBLAS GEMMs); the env / history / normalisation / GAE kernels around it are this library's.
"""
import ctypes as C
import math

import torch
import torch.nn as nn

from . import _lib


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class DynamicNormDevice(nn.Module):
    """DynamicNorm(input_size, only_for_last_dim=True, exclude_one_hot=True, exclude_nan=True), norm.py:14-69.
    Parameters as in the reference (mean, var float32 [D]; n_sample int64 [1]; one_hot_filter bool [D])."""

    def __init__(self, input_size=16, group=None):
        super().__init__()
        self.input_size = input_size
        self.mean = nn.Parameter(torch.zeros(input_size), requires_grad=False)
        self.var = nn.Parameter(torch.ones(input_size), requires_grad=False)
        self.n_sample = nn.Parameter(torch.zeros(1, dtype=torch.long), requires_grad=False)
        self.one_hot_filter = nn.Parameter(torch.ones(input_size, dtype=torch.bool), requires_grad=False)
        self.group = group  # torch.distributed group for the moments all-reduce (None = default group if initialised)
        self._scratch = None

    def batch_stats(self, x):
        """(sum x[D], sum x^2[D], values outside {-1,0,1}[D], rows) over the rows of x without NaN: f64 [3D+1] on the device."""
        lib = _lib.load()
        x2 = x.reshape(-1, self.input_size)
        assert x2.is_cuda and x2.dtype == torch.float32 and x2.is_contiguous()
        if self._scratch is None or self._scratch.device != x2.device:
            self._scratch = torch.empty(592 * 49, dtype=torch.float64, device=x2.device)
        stats = torch.empty(49, dtype=torch.float64, device=x2.device)
        stream = C.c_void_p(torch.cuda.current_stream(x2.device).cuda_stream)
        _lib.check(lib.dca_norm_moments(_ptr(x2), x2.shape[0], self.input_size, _ptr(self._scratch), _ptr(stats), stream))
        return stats

    @torch.no_grad()
    def update(self, x):
        """The training-mode half of DynamicNorm.forward (norm.py:47-63) from the batch statistics; with
        torch.distributed initialised the statistics of all ranks are summed first (one all-reduce of 49 doubles)."""
        stats = self.batch_stats(x)
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            torch.distributed.all_reduce(stats, group=self.group)
        return self.merge(stats)

    @torch.no_grad()
    def merge(self, stats):
        D = self.input_size
        s, q, nonunit, n = stats[:D], stats[D:2 * D], stats[2 * D:3 * D], stats[3 * D]
        nb = torch.clamp(n, min=1.0)
        b_mean = s / nb
        b_var = torch.clamp(q / nb - b_mean * b_mean, min=0.0)     # torch.var(unbiased=False)
        n0 = self.n_sample.to(torch.float64)
        tot = n0 + n
        mean0, var0 = self.mean.to(torch.float64), self.var.to(torch.float64)
        delta = b_mean - mean0
        new_mean = mean0 + delta * n / torch.clamp(tot, min=1.0)
        m2 = var0 * n0 + b_var * n + delta * delta * n0 * n / torch.clamp(tot, min=1.0)
        new_var = torch.clamp(m2 / torch.clamp(tot, min=1.0), min=0.01, max=1000.0)
        some = n > 0                                                # an empty batch changes nothing (norm.py:41-44)
        filt = self.one_hot_filter & ~((nonunit > 0) & some)
        keep = filt | ~some
        # in place: the parameters keep their addresses (a captured CUDA graph replays these updates)
        self.one_hot_filter.data.copy_(filt)
        self.mean.data.copy_(torch.where(keep, self.mean, new_mean.to(torch.float32)))
        self.var.data.copy_(torch.where(keep, self.var, new_var.to(torch.float32)))
        self.n_sample.data.copy_((n0 + n).round().to(torch.long).reshape(1))
        return self

    def apply_norm(self, x, out=None):
        """-> (normalised x with NaN -> 0, row mask uint8 = isnan(x).any(-1)); out may be x itself."""
        lib = _lib.load()
        shape = x.shape
        x2 = x.reshape(-1, self.input_size)
        assert x2.is_cuda and x2.dtype == torch.float32 and x2.is_contiguous()
        y = torch.empty_like(x2) if out is None else out.reshape(-1, self.input_size)
        mask = torch.empty((x2.shape[0],), dtype=torch.uint8, device=x2.device)
        stream = C.c_void_p(torch.cuda.current_stream(x2.device).cuda_stream)
        _lib.check(lib.dca_norm_apply(_ptr(x2), _ptr(self.mean.data), _ptr(self.var.data), x2.shape[0], self.input_size,
                                      _ptr(y), _ptr(mask), stream))
        return y.reshape(shape), mask.reshape(shape[:-1])

    def forward(self, x):
        if self.training:
            self.update(x)
        return self.apply_norm(x)


def _linear_cat(lin, a, b):
    """lin(cat((a, b), -1)) without materialising the concatenation: a W[:, :ka]^T + b W[:, ka:]^T + bias."""
    ka = a.shape[-1]
    y = torch.addmm(lin.bias, a, lin.weight[:, :ka].t())
    return y.addmm_(b, lin.weight[:, ka:].t())


def _plin(a, wa, b=None, wb=None, bias=None, relu=False, ffma=False):
    """act(a wa^T + b wb^T + bias) on the device in one kernel (dca_policy_linear): a [B,ka] / b [B,kb] float32 with unit
    inner stride (column slices of wider tensors are fine), wa [n,ka] / wb [n,kb] any 2-D views (x @ M: pass M.t()).
    Arithmetic: three tf32 tensor-core products per multiply on head + tail splits of the operands (float32-class
    accuracy); ffma=True: plain float32 FFMA."""
    lib = _lib.load()
    B, n = a.shape[0], wa.shape[0]
    assert a.dtype == torch.float32 and a.stride(1) == 1 and wa.shape[1] == a.shape[1]
    out = torch.empty((B, n), dtype=torch.float32, device=a.device)
    if b is not None:
        assert b.dtype == torch.float32 and b.stride(1) == 1 and wb.shape == (n, b.shape[1]) and b.shape[0] == B
    stream = C.c_void_p(torch.cuda.current_stream(a.device).cuda_stream)
    _lib.check(lib.dca_policy_linear(
        _ptr(a), a.stride(0), a.shape[1], _ptr(wa), wa.stride(0), wa.stride(1),
        None if b is None else _ptr(b), 0 if b is None else b.stride(0), 0 if b is None else b.shape[1],
        None if b is None else _ptr(wb), 0 if b is None else wb.stride(0), 0 if b is None else wb.stride(1),
        None if bias is None else _ptr(bias), n, int(bool(relu)) | (2 if ffma else 0), B, _ptr(out), out.stride(0), stream))
    return out


class ConcCore(nn.Module):
    """Concentration(n_focus_on, h_dim, skip_connect=True, skip_connect_dim), ALGORITHM/common/conc.py:7-63."""

    def __init__(self, n_focus_on, h_dim, skip_connect_dim):
        super().__init__()
        self.n_focus_on = n_focus_on
        self.h_dim = h_dim
        self.CT_W_query = nn.Parameter(torch.empty(h_dim, h_dim))
        self.CT_W_key = nn.Parameter(torch.empty(h_dim, h_dim))
        self.CT_W_val = nn.Parameter(torch.empty(h_dim, h_dim))
        self.CT_motivate_mlp = nn.Sequential(nn.Linear(h_dim * 2, h_dim), nn.ReLU(inplace=True))
        self.AT_forward_mlp = nn.Sequential(nn.Linear((n_focus_on + 1) * (h_dim + skip_connect_dim), h_dim),
                                            nn.ReLU(inplace=True))
        for p in self.parameters():  # conc.py:23-26
            stdv = 1. / math.sqrt(p.size(-1))
            p.data.uniform_(-stdv, stdv)

    def forward(self, vs, ve, ve_dead, ze, zs):
        """vs [B,h], ve [B,n,h], ve_dead bool [B,n], ze [B,n,d], zs [B,d] -> (v_C [B,h], v_M [B,h])."""
        qk = (vs @ self.CT_W_query) @ self.CT_W_key.t()                       # [B,h]
        compat = torch.einsum('bnh,bh->bn', ve, qk) * (1.0 / math.sqrt(self.h_dim))
        compat = compat.masked_fill(ve_dead, -math.inf)
        score = torch.nan_to_num(torch.softmax(compat, dim=-1), 0.0)          # no live entity: all zero (conc.py:39-41)
        va = torch.einsum('bn,bnh->bh', score, ve) @ self.CT_W_val
        v_m = self.CT_motivate_mlp(torch.cat((vs, va), -1))
        # The attention is often so peaked that float32 softmax gives several live entities a score of exactly 0, tied
        # with the masked ones; torch.argsort's order among equal scores is unspecified (and differs between the CPU
        # and the CUDA sort), so the reference's pick is implementation-defined there.  Stable order = lowest entity
        # index first, which is what the reference computes on the CPU.
        idx = torch.argsort(score, dim=-1, descending=True, stable=True)[..., :self.n_focus_on]
        pv = torch.gather(ve, 1, idx.unsqueeze(-1).expand(-1, -1, ve.shape[-1]))   # gather first, concatenate the few
        pz = torch.gather(ze, 1, idx.unsqueeze(-1).expand(-1, -1, ze.shape[-1]))
        stack = torch.cat((torch.cat((vs, zs), -1).unsqueeze(1), torch.cat((pv, pz), -1)), 1)  # [B, n_focus_on + 1, h + d]
        v_c = self.AT_forward_mlp(stack.reshape(stack.shape[0], -1))
        return v_c, v_m

    def qk_fused(self, vs):
        """(vs W_query) W_key^T = vs (W_query W_key^T): the 48 x 48 product once per call instead of a second [B,48] pass."""
        return _plin(vs, (self.CT_W_query @ self.CT_W_key.t()).t())

    def post_fused(self, vs, va, stack):
        """The layers behind the attention kernel: -> (v_C, v_M)."""
        lin, fwd = self.CT_motivate_mlp[0], self.AT_forward_mlp[0]
        fold = self.CT_W_val @ lin.weight[:, self.h_dim:].t()   # (va W_val) Wm2^T = va (W_val Wm2^T)
        v_m = _plin(vs, lin.weight[:, :self.h_dim], va, fold.t(), lin.bias, relu=True)
        v_c = _plin(stack, fwd.weight, bias=fwd.bias, relu=True)
        return v_c, v_m

    def forward_fused(self, v, x, dead_u8, first, n):
        """The same on the whole [B,24,*] tensors with the part between the matrix products in one kernel
        (dca_conc_attend: scores, masked softmax, attended encoding, stable top-n_focus gather): v f32 [B,24,h] entity
        encodings, x f32 [B,24,d] normalised observations, dead_u8 [B,24]; the list is entities first .. first+n-1."""
        lib = _lib.load()
        B = v.shape[0]
        vs = v[:, 0]
        qk = self.qk_fused(vs)
        va = torch.empty((B, self.h_dim), dtype=torch.float32, device=v.device)
        stack = torch.empty((B, (self.n_focus_on + 1) * (self.h_dim + x.shape[-1])), dtype=torch.float32, device=v.device)
        stream = C.c_void_p(torch.cuda.current_stream(v.device).cuda_stream)
        _lib.check(lib.dca_conc_attend(_ptr(v), _ptr(x), _ptr(dead_u8), _ptr(qk), first, n, self.n_focus_on, B, _ptr(va),
                                       _ptr(stack), stream))
        return self.post_fused(vs, va, stack)


class ConcNet4Hist(nn.Module):
    """Net(rawob_dim=16, n_action=7, n_focus_on, dual_conc=True) of ALGORITHM/conc_4hist/net.py:87-233 for the rollout
    (`act`): 24 entities per agent = self, 11 current, 12 from the history pool (div_entity :170-181)."""

    H = 48

    def __init__(self, rawob_dim=16, n_action=7, n_focus_on=3, group=None):
        super().__init__()
        h = self.H
        self.rawob_dim, self.n_action, self.n_focus_on = rawob_dim, n_action, n_focus_on
        self._batch_norm = DynamicNormDevice(rawob_dim, group=group)
        self.AT_obs_encoder = nn.Sequential(nn.Linear(rawob_dim, h), nn.ReLU(inplace=True), nn.Linear(h, h))
        self.MIX_conc_core_f = ConcCore(n_focus_on - 1, h, rawob_dim)
        self.MIX_conc_core_h = ConcCore(n_focus_on, h, rawob_dim)
        self.CT_get_value = nn.Sequential(nn.Linear(2 * h, h), nn.ReLU(inplace=True), nn.Linear(h, 1))
        self.CT_get_threat = nn.Sequential(nn.Linear(2 * h, h), nn.ReLU(inplace=True), nn.Linear(h, 1))
        self.AT_get_logit_db = nn.Sequential(nn.Linear(2 * h, h), nn.ReLU(inplace=True),
                                             nn.Linear(h, h // 2), nn.ReLU(inplace=True), nn.Linear(h // 2, n_action))

    @classmethod
    def from_checkpoint(cls, path, device, n_focus_on=3, group=None):
        sd = torch.load(path, map_location='cpu')
        net = cls(rawob_dim=sd['AT_obs_encoder.0.weight'].shape[1], n_action=sd['AT_get_logit_db.4.weight'].shape[0],
                  n_focus_on=n_focus_on, group=group)
        net.load_state_dict(sd)
        return net.to(device)

    fused = True  # dca_policy_encode / dca_conc_attend instead of torch elementwise + GEMM + bmm / sort / gather passes
    attend_both = True  # fused path: dca_conc_attend2 (both entity lists in one kernel) instead of two dca_conc_attend

    def encode_fused(self, obs_feed):
        """DynamicNorm apply + mask + AT_obs_encoder in one kernel: -> (x f32 [B,24,D], dead u8 [B,24], v f32 [B,24,H])."""
        lib = _lib.load()
        x2 = obs_feed.reshape(-1, self.rawob_dim)
        assert x2.is_cuda and x2.dtype == torch.float32 and x2.is_contiguous()
        R = x2.shape[0]
        xn = torch.empty_like(x2)
        dead = torch.empty((R,), dtype=torch.uint8, device=x2.device)
        v = torch.empty((R, self.H), dtype=torch.float32, device=x2.device)
        bn, enc = self._batch_norm, self.AT_obs_encoder
        stream = C.c_void_p(torch.cuda.current_stream(x2.device).cuda_stream)
        _lib.check(lib.dca_policy_encode(_ptr(x2), _ptr(bn.mean.data), _ptr(bn.var.data), _ptr(enc[0].weight.data),
                                         _ptr(enc[0].bias.data), _ptr(enc[2].weight.data), _ptr(enc[2].bias.data), R,
                                         _ptr(xn), _ptr(dead), _ptr(v), stream))
        return xn.reshape(-1, 24, self.rawob_dim), dead.reshape(-1, 24), v.reshape(-1, 24, self.H)

    def _tail_fused(self, vf_c, vf_m, vh_c, vh_m, need_threat=True):
        """_tail with every layer as one dca_policy_linear launch (the concatenations become second inputs)."""
        h = self.H
        db, cv, ct = self.AT_get_logit_db, self.CT_get_value, self.CT_get_threat
        a1 = _plin(vf_c, db[0].weight[:, :h], vh_c, db[0].weight[:, h:], db[0].bias, relu=True)
        logits = _plin(_plin(a1, db[2].weight, bias=db[2].bias, relu=True), db[4].weight, bias=db[4].bias)
        c1 = _plin(vf_m, cv[0].weight[:, :h], vh_m, cv[0].weight[:, h:], cv[0].bias, relu=True)
        value = _plin(c1, cv[2].weight, bias=cv[2].bias)[:, 0]
        if not need_threat:
            return logits, value, None
        t1 = _plin(vf_m, ct[0].weight[:, :h], vh_m, ct[0].weight[:, h:], ct[0].bias, relu=True)
        t = _plin(t1, ct[2].weight, bias=ct[2].bias)[:, 0]
        r = 0.5 * 11
        return logits, value, (torch.tanh(t / r) + 1.0) * r

    def _tail(self, vf_c, vf_m, vh_c, vh_m, need_threat=True):
        db = self.AT_get_logit_db
        logits = db[4](torch.relu_(db[2](torch.relu_(_linear_cat(db[0], vf_c, vh_c)))))
        value = self.CT_get_value[2](torch.relu_(_linear_cat(self.CT_get_value[0], vf_m, vh_m)))[:, 0]
        if not need_threat:  # Net.act returns (act, value, log-prob) only (net.py:231): the rollout never reads the threat
            return logits, value, None
        t = self.CT_get_threat[2](torch.relu_(_linear_cat(self.CT_get_threat[0], vf_m, vh_m)))[:, 0]
        r = 0.5 * 11
        return logits, value, (torch.tanh(t / r) + 1.0) * r                   # re_scale, net.py:226-229

    @torch.no_grad()
    def heads(self, obs_feed, chunk=1 << 18, need_threat=True):
        """obs_feed f32 [..., 24, D] (empty rows NaN, a dead agent all NaN) -> (logits [..., n_action], value [...],
        threat [...]).  On the device with the fused kernels everything runs in one pass; the torch path processes
        `chunk` agents at a time."""
        lead = obs_feed.shape[:-2]
        if self.fused and obs_feed.is_cuda and self.H == 48 and self.rawob_dim == 16:
            if self._batch_norm.training:
                self._batch_norm.update(obs_feed)
            x, dead_u8, v = self.encode_fused(obs_feed)
            cf, ch = self.MIX_conc_core_f, self.MIX_conc_core_h
            if self.attend_both:   # both entity lists in one attention kernel (dca_conc_attend2)
                vs, B, h, d = v[:, 0], v.shape[0], self.H, self.rawob_dim
                qk_f, qk_h = cf.qk_fused(vs), ch.qk_fused(vs)
                new = lambda n: torch.empty((B, n), dtype=torch.float32, device=v.device)
                va_f, va_h = new(h), new(h)
                st_f, st_h = new((cf.n_focus_on + 1) * (h + d)), new((ch.n_focus_on + 1) * (h + d))
                stream = C.c_void_p(torch.cuda.current_stream(v.device).cuda_stream)
                _lib.check(_lib.load().dca_conc_attend2(_ptr(v), _ptr(x), _ptr(dead_u8), _ptr(qk_f), _ptr(qk_h), 1, 11,
                                                        cf.n_focus_on, 12, ch.n_focus_on, B, _ptr(va_f), _ptr(st_f),
                                                        _ptr(va_h), _ptr(st_h), stream))
                vf_c, vf_m = cf.post_fused(vs, va_f, st_f)
                vh_c, vh_m = ch.post_fused(vs, va_h, st_h)
            else:
                vf_c, vf_m = cf.forward_fused(v, x, dead_u8, 1, 11)
                vh_c, vh_m = ch.forward_fused(v, x, dead_u8, 12, 12)
            logits, value, threat = self._tail_fused(vf_c, vf_m, vh_c, vh_m, need_threat)
            return (logits.reshape(*lead, self.n_action), value.reshape(lead),
                    threat.reshape(lead) if need_threat else None)
        x, dead = self._batch_norm(obs_feed)
        x = x.reshape(-1, 24, self.rawob_dim)
        dead = dead.reshape(-1, 24).bool()
        B = x.shape[0]
        logits = torch.empty((B, self.n_action), dtype=torch.float32, device=x.device)
        value = torch.empty((B,), dtype=torch.float32, device=x.device)
        threat = torch.empty((B,), dtype=torch.float32, device=x.device)
        for b0 in range(0, B, chunk):
            xs, ds = x[b0:b0 + chunk], dead[b0:b0 + chunk]
            v = self.AT_obs_encoder(xs)
            vf_c, vf_m = self.MIX_conc_core_f(v[:, 0], v[:, 1:12], ds[:, 1:12], xs[:, 1:12], xs[:, 0])
            vh_c, vh_m = self.MIX_conc_core_h(v[:, 0], v[:, 12:24], ds[:, 12:24], xs[:, 12:24], xs[:, 0])
            logits[b0:b0 + chunk], value[b0:b0 + chunk], threat[b0:b0 + chunk] = self._tail(vf_c, vf_m, vh_c, vh_m)
        return logits.reshape(*lead, self.n_action), value.reshape(lead), threat.reshape(lead)

    @torch.no_grad()
    def act(self, obs_feed, test_mode=False, generator=None):
        """Net.act: (action int32 [...], value f32 [...], log-prob f32 [...]).  test_mode: arg-max (net.py:147), else one
        Categorical draw per agent (Gumbel arg-max of the logits, from `generator` on the device)."""
        logits, value, _ = self.heads(obs_feed, need_threat=False)
        logp_all = torch.log_softmax(logits, -1)
        if test_mode:
            act = torch.argmax(logp_all, -1)
        else:
            u = torch.rand(logits.shape, dtype=torch.float32, device=logits.device, generator=generator)
            act = torch.argmax(logp_all - torch.log(-torch.log(u.clamp_min(1e-20))), -1)
        logp = torch.gather(logp_all, -1, act.unsqueeze(-1)).squeeze(-1)
        return act.to(torch.int32), value, logp


class PolicyRollout(object):
    """The device-resident loop of INTEGRATION.md section 3 for BASELINE configs[3]: scripted attackers (k_plan) + env
    step (k_step) + history pool (k_histpool) + conc_4hist policy + rollout buffer + GAE (k_gae), no host hop.
    env: DcaVecEnv(empty_rows_nan=True) whose guards the policy drives."""

    def __init__(self, env, net, gamma=0.99, tau=0.95, gamma_fwd=None, seed=0, test_mode=False, plan_ahead=True):
        from .rollout import DeviceRollout, HistoryPool
        self.env, self.net, self.test_mode = env, net, test_mode
        # plan_ahead: the scripted attackers' planner (k_plan) needs the env state, not the guards' actions -- it runs on
        # a side stream under the policy forward (one warp per env chasing dependent chains: it leaves the SMs' issue
        # slots to the policy's GEMMs) and the step joins both.  Same launches, same results; also inside a CUDA graph
        # (fork / join by events = parallel branches of the captured graph).
        self._side = torch.cuda.Stream(device=env.device) if plan_ahead else None
        self.hist = HistoryPool(env.E, env.G, device=env.device)
        self.buf = DeviceRollout(env, gamma=gamma, tau=tau, gamma_fwd=gamma_fwd)
        self.gen = torch.Generator(device=env.device)
        self.gen.manual_seed(seed)
        self.suffered = torch.ones((env.E,), dtype=torch.uint8, device=env.device)  # Env-Suffered-Reset: all, at start
        self.episodes = torch.zeros((), dtype=torch.int64, device=env.device)
        self.wins = torch.zeros((), dtype=torch.int64, device=env.device)
        self.steps_in_finished = torch.zeros((), dtype=torch.int64, device=env.device)
        self.obs = env.reset()

    def step(self, collect=True):
        """One env step of every env: history roll -> policy -> plan -> step -> rollout buffer.  Returns the mask of
        finished envs (device bool [E])."""
        env = self.env
        force = None
        if self._side is not None:
            cur = torch.cuda.current_stream(env.device)
            fork, join = torch.cuda.Event(), torch.cuda.Event()
            fork.record(cur)
            self._side.wait_event(fork)
            with torch.cuda.stream(self._side):
                force = env.plan()
                join.record(self._side)
        feed = self.hist.roll(self.obs, suffered_reset=self.suffered)
        act, value, _ = self.net.act(feed, test_mode=self.test_mode, generator=self.gen)
        alive_before = self.buf.alive_guards().clone() if collect else None
        lens_before = self.buf.length.clone() if collect else None
        if force is not None:
            cur.wait_event(join)
        self.obs, reward, done, win = env.step(act, force)   # force None: plan() inside, the scripted attackers of World.step
        fin = done != 0
        if collect:
            self.buf.add(alive_before, value, reward, done)
            self.steps_in_finished += torch.where(fin, lens_before + 1, torch.zeros_like(lens_before)).sum()
        self.suffered.copy_(done)   # in place: a captured graph reads the same buffer next step
        self.episodes += fin.sum()
        self.wins += (fin & (win != 0)).sum()
        return fin

    def capture(self, warm=3, collect=True):
        """Capture one step() as a CUDA graph (every op of the loop is an asynchronous launch on the current stream: the
        library's kernels through the C-ABI, torch's GEMMs, the Categorical draw) and return a callable that replays it
        -- one cudaGraphLaunch per env step instead of ~80 launches, which is what bounds a 4 ms step on the host.
        The replayed step writes the same tensors every time (`finished` below, env.obs / reward / done, the rollout
        buffer).  (DCA_OPT_PIPELINE_STEPS keeps host-side epoch counters: the library launches normally while capturing.)"""
        for _ in range(warm):   # allocations, lazy module loads and per-function attributes happen outside the capture
            self.step(collect)
        torch.cuda.synchronize(self.env.device)
        graph = torch.cuda.CUDAGraph()
        if hasattr(graph, 'register_generator_state'):
            graph.register_generator_state(self.gen)
        with torch.cuda.graph(graph):
            fin = self.step(collect)
        self._graph, self.finished = graph, fin

        def replay():
            graph.replay()
            return fin
        return replay
