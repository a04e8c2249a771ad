"""ScenarioConfig for DCA + JSONC config injection, host-side mirror of the reference interface.

Mirrors (same attribute names, value coercion and error behaviour; own implementation):
  * ScenarioConfig            /root/reference/MISSION/dca/collective_assault_parallel_run.py:30-101
  * ChainVar / `_cv` twins    /root/reference/UTIL/config_args.py:12-15, 121-144
  * group injection           /root/reference/UTIL/config_args.py:88-119 ("<module>.py-><Class>" keys)
  * type coercion             /root/reference/UTIL/config_args.py:284-313
`commentjson` is absent from the image, so JSONC (// and /* */ comments, trailing commas) is stripped here.

Stale-name handling: the shipped RESULT/test-50+50 and test-100+100 configs name the module
`MISSION.collective_assault...` (renamed to MISSION/dca) and the key `num_steps` (now MaxEpisodeStep); both are
mapped explicitly and reported in `InjectionReport.renamed`.
"""
import json
import math
import re


class ChainVar(object):
    """A value derived from other config items; recomputed when one of `chained_with` is overridden and the
    derived item itself is not (UTIL/config_args.py:12-15, 121-144)."""

    def __init__(self, chain_func, chained_with):
        self.chain_func = chain_func
        self.chained_with = chained_with


class ScenarioConfig(object):
    """DCA scenario parameters.  Read ONCE when an env object is constructed (POD snapshot handed to the
    kernels), exactly like the reference reads them in __init__ (core.py:72-73, collective_assault_env.py:10-13)."""
    num_guards = 50
    num_attackers = 50
    agent_acc = 3.0
    agent_max_speed = 1.0
    agent_max_rot = 0.17
    agent_shoot_base_radius = 0.4
    agent_shoot_win = math.pi / 4
    random_jam_prob = 0.0
    introduce_terrain = False
    terrain_parameters = [0, 0]
    MaxEpisodeStep = 180
    init_distance = 4
    enable_attacker_fluctuation = False
    attacker_fluctuation = 20

    size = 5.0
    render = False
    render_with_unity = False
    render_ip_with_unity = '127.0.0.1:11088'

    map_ = 'default'
    benchmark = False
    N_TEAM = 1

    N_AGENT_EACH_TEAM = [50, ]
    N_AGENT_EACH_TEAM_cv = ChainVar(lambda num_guards: [num_guards, ], chained_with=['num_guards'])

    AGENT_ID_EACH_TEAM = [range(0, 50), ]
    AGENT_ID_EACH_TEAM_cv = ChainVar(lambda num_guards: [range(0, num_guards), ], chained_with=['num_guards'])

    TEAM_NAMES = ['ALGORITHM.Starcraft.star_foundation->StarFoundation', ]
    obs_vec_length = 16

    num_entity = num_attackers
    num_entity_cv = ChainVar(lambda num_attackers: num_attackers, chained_with=['num_attackers'])

    num_object = num_attackers + num_guards
    num_object_cv = ChainVar(lambda num_attackers, num_guards: num_attackers + num_guards,
                             chained_with=['num_attackers', 'num_guards'])

    uid_dictionary = {'agent_uid': range(0, 50), 'entity_uid': range(50, 100),
                      'guards_uid': range(0, 50), 'attackers_uid': range(50, 100)}
    uid_dictionary_cv = ChainVar(lambda num_guards, num_attackers: {
        'agent_uid': range(0, num_guards),
        'entity_uid': range(num_guards, num_attackers + num_guards),
        'guards_uid': range(0, num_guards),
        'attackers_uid': range(num_guards, num_attackers + num_guards)},
        chained_with=['num_guards', 'num_attackers'])

    dec_dictionary = {'alive': 0, 'pos': range(1, 3), 'ang': 3, 'vel': range(4, 6), 'id': 6}
    StateProvided = False
    AvailActProvided = False
    RewardAsUnity = False
    ObsAsUnity = False
    EntityOriented = True

    MCOM_DEBUG = False
    Terrain_DEBUG = False
    DISALBE_RED_FUNCTION = False
    half_death_reward = True
    REWARD_DEBUG = False
    REWARD_DEBUG_Value = 1

    n_actions = 7


class GlobalConfig(object):
    """The few GlobalConfig items the env layer reads (config.py of the reference; env_router.py:68-117)."""
    env_name = 'dca'
    num_threads = 64
    fold = 1
    seed = 0
    device = 'cuda'
    ScenarioConfig = ScenarioConfig


# group keys that resolve to the classes above; the first two are stale names still found in shipped configs
_SCENARIO_GROUPS = (
    'MISSION.collective_assault.collective_assault_parallel_run.py->ScenarioConfig',
    'MISSION.collective_assault.collective_assault_parallel_run->ScenarioConfig',
    'MISSION.dca.collective_assault_parallel_run.py->ScenarioConfig',
    'MISSION.dca.collective_assault_parallel_run->ScenarioConfig',
    'hmp2g_b200.config.py->ScenarioConfig',
)
_MT_SCENARIO_GROUPS = (
    'MISSION.dca_multiteam.collective_assault_parallel_run.py->ScenarioConfig',
    'MISSION.dca_multiteam.collective_assault_parallel_run->ScenarioConfig',
    'hmp2g_b200.multiteam.py->MtScenarioConfig',
)
_GLOBAL_GROUPS = ('config.py->GlobalConfig', 'hmp2g_b200.config.py->GlobalConfig')
_STALE_KEYS = {'num_steps': 'MaxEpisodeStep'}
_STALE_ENV_NAMES = {'collective_assault': 'dca'}


def strip_jsonc(text):
    """Remove // line comments, /* */ block comments and trailing commas, leaving string literals intact."""
    out, i, n, in_str = [], 0, len(text), False
    while i < n:
        ch = text[i]
        if in_str:
            out.append(ch)
            if ch == '\\' and i + 1 < n:
                out.append(text[i + 1])
                i += 1
            elif ch == '"':
                in_str = False
        elif ch == '"':
            in_str = True
            out.append(ch)
        elif text.startswith('//', i):
            while i < n and text[i] != '\n':
                i += 1
            continue
        elif text.startswith('/*', i):
            j = text.find('*/', i + 2)
            i = n if j < 0 else j + 2
            continue
        else:
            out.append(ch)
        i += 1
    return re.sub(r',(\s*[}\]])', r'\1', ''.join(out))


def load_jsonc(path):
    with open(path, encoding='utf-8') as f:
        return json.loads(strip_jsonc(f.read()))


def coerce_like(original, new_value, key):
    """Coerce an override to the type of the default (UTIL/config_args.py:284-313): float, bool ('True'/'False'
    strings allowed), int (must be integral), str, list, dict; anything else is an error."""
    if isinstance(original, float):
        return float(new_value)
    if isinstance(original, bool):
        if new_value == 'True':
            return True
        if new_value == 'False':
            return False
        if isinstance(new_value, bool):
            return new_value
        raise AssertionError(('enter True or False, but have:', new_value))
    if isinstance(original, int):
        assert int(new_value) == float(new_value), \
            'warning, this var **%s** has an int default, but given a float override!' % key
        return int(new_value)
    if isinstance(original, str):
        return new_value
    if isinstance(original, list):
        assert isinstance(new_value, list)
        return new_value
    if isinstance(original, dict):
        assert isinstance(new_value, dict)
        return new_value
    raise AssertionError('not support this type')


def _is_chained_key(key):
    for suf in ('_cv', '_CV'):
        if key.endswith(suf):
            return True, key[:-len(suf)]
    return False, key


def override_config(conf_class, new_cfg):
    """Apply one JSONC group to a config class; returns the list of chained items that were recomputed."""
    for key, value in new_cfg.items():
        if value is None:
            continue
        assert hasattr(conf_class, key), (conf_class, 'has no such config item: **%s**' % key)
        setattr(conf_class, key, coerce_like(getattr(conf_class, key), value, key))
    altered = []
    for key in list(vars(conf_class)):
        is_chain, o_key = _is_chained_key(key)
        if not is_chain or o_key in new_cfg:
            continue
        assert hasattr(conf_class, o_key), 'twin var does not have original'
        cv = getattr(conf_class, key)
        if not any(dep in new_cfg for dep in cv.chained_with):
            continue
        setattr(conf_class, o_key, cv.chain_func(*[getattr(conf_class, v) for v in cv.chained_with]))
        altered.append(o_key)
    return altered


class InjectionReport(object):
    def __init__(self):
        self.applied, self.ignored_groups, self.renamed, self.altered_chained = [], [], [], []


def inject(json_data, scenario_cls=ScenarioConfig, global_cls=GlobalConfig):
    """Inject a parsed JSONC dict.  Groups addressed to the DCA ScenarioConfig / GlobalConfig are applied (with the
    reference's coercion rules); groups of modules outside this package (ALGORITHM.*) are listed in
    report.ignored_groups.  Unknown keys inside an applied group raise, as in the reference."""
    rep = InjectionReport()
    for group, body in json_data.items():
        assert '->' in group, group
        if group in _SCENARIO_GROUPS:
            body = dict(body)
            if 'collective_assault.collective' in group:
                rep.renamed.append((group, _SCENARIO_GROUPS[2]))
            for old, new in _STALE_KEYS.items():
                if old in body and not hasattr(scenario_cls, old):
                    body[new] = body.pop(old)
                    rep.renamed.append((old, new))
            rep.altered_chained += override_config(scenario_cls, body)
            rep.applied.append(group)
        elif group in _MT_SCENARIO_GROUPS:  # the sibling env MISSION/dca_multiteam
            from .multiteam import MtScenarioConfig
            rep.altered_chained += override_config(MtScenarioConfig, dict(body))
            rep.applied.append(group)
        elif group in _GLOBAL_GROUPS:
            known = {k: v for k, v in body.items() if hasattr(global_cls, k)}
            if known.get('env_name') in _STALE_ENV_NAMES:
                rep.renamed.append((known['env_name'], _STALE_ENV_NAMES[known['env_name']]))
                known['env_name'] = _STALE_ENV_NAMES[known['env_name']]
            override_config(global_cls, known)
            rep.applied.append(group)
        else:
            rep.ignored_groups.append(group)
    return rep


def reset_defaults(conf_class, snapshot=None):
    """Snapshot / restore class attributes (tests construct many configs in one process)."""
    if snapshot is None:
        return {k: v for k, v in vars(conf_class).items() if not k.startswith('__')}
    for k, v in snapshot.items():
        setattr(conf_class, k, v)
    return snapshot
