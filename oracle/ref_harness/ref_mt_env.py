"""ORACLE / TEST INFRASTRUCTURE -- runs only where /root/reference exists.

Drives the *unmodified* reference `dca_multiteam` env (MISSION/dca_multiteam, wrapped in UTIL/shm_env.EnvAutoReset)
under the keyed RNG stream: only np.random draw functions are swapped while reset_world / observation / the laser loop
run.  The laser hit draw (core.py:188) is keyed by the (shooter, target) pair the calling frame is looking at.
"""
import sys

import numpy as np

from ref_env import setup_reference, _patched_np_random, _ObsDraws
import dca_rng

SITE_LASER = 4


class RefMtEnv(object):
    def __init__(self, cfg, seed=0, env_gid=0):
        setup_reference()
        from config import GlobalConfig
        from UTIL.tensor_ops import arrange_id
        from MISSION.dca_multiteam.collective_assault_parallel_run import ScenarioConfig
        import mt_oracle
        full = dict(mt_oracle.DEFAULT_CFG)
        full.update(cfg)
        self.cfg = full
        for k, v in full.items():
            assert hasattr(ScenarioConfig, k), k
            setattr(ScenarioConfig, k, v)
        ScenarioConfig.N_TEAM = len(full['N_AGENT_EACH_TEAM'])
        ScenarioConfig.AGENT_ID_EACH_TEAM = arrange_id(full['N_AGENT_EACH_TEAM'])
        GlobalConfig.env_name = 'dca_multiteam'
        self.N = sum(full['N_AGENT_EACH_TEAM'])
        self.seed, self.env_gid = seed, env_gid
        self.episode = -1
        self.obs_counter = -1
        from MISSION.dca_multiteam.envs.collective_assault_env import collective_assaultEnvV1
        collective_assaultEnvV1.raw_obs_size = -1
        # the constructor itself calls reset_world() once with the global stream (envs/collective_assault_env.py:39);
        # nothing of it survives the first keyed reset
        from UTIL.shm_env import EnvAutoReset
        self.wrapper = EnvAutoReset({'env_name': 'dca_multiteam', 'proc_index': env_gid})
        self.env = self.wrapper._env
        self.world = self.env.world
        self._install_hooks()

    def _install_hooks(self):
        env, world = self.env, self.world
        orig_reset_world = env.reset_callback
        orig_observation = env.observation_callback
        orig_laser = world.apply_laser_effect
        N = self.N

        def keyed_reset_world():
            self.episode += 1
            ww = dca_rng.reset_world_words(self.seed, self.env_gid, self.episode)
            seq = [ww[0] * 2.0 ** -32, ww[1] * 2.0 ** -32]
            seq += [dca_rng.reset_agent_words(self.seed, self.env_gid, self.episode, i)[0] * 2.0 ** -32 for i in range(N)]
            it = iter(seq)
            with _patched_np_random(rand=lambda *s: next(it)):
                orig_reset_world()

        def keyed_observation(agent, world_, *a, **kw):
            if agent.iden == 0:
                self.obs_counter += 1
            state = {'lst': 0}

            def shuffle(arr):
                w = dca_rng.obs_words(self.seed, self.env_gid, self.obs_counter, agent.iden, state['lst'])
                state['draw'] = _ObsDraws(w)
                state['draw'].shuffle(arr)

            def rand(n):
                d = state['draw']
                state['lst'] += 1
                return d.rand(n)

            with _patched_np_random(shuffle=shuffle, rand=rand):
                return orig_observation(agent, world_, *a, **kw)

        def keyed_laser():
            def rand():
                f = sys._getframe(1).f_locals  # apply_laser_effect's frame: entity = shooter, entity_b = target
                i, j = f['entity'].uid, f['entity_b'].uid
                w = dca_rng._words(self.seed, self.env_gid, self.obs_counter, i * N + j, SITE_LASER, 1)
                return w[0] * 2.0 ** -32
            with _patched_np_random(rand=rand):
                orig_laser()

        env.reset_callback = keyed_reset_world
        env.observation_callback = keyed_observation
        world.apply_laser_effect = keyed_laser

    def export_state(self):
        ag = self.world.agents
        f = lambda name, dt: np.array([getattr(a, name) for a in ag], dtype=dt)
        bsa = np.zeros(len(ag), np.uint8)
        for a in self.world.before_step_alive_agents:
            bsa[a.uid] = 1
        return dict(pos=np.array([a.pos for a in ag], dtype=np.float64), vel=np.array([a.vel for a in ag], dtype=np.float64),
                    atk_rad=f('atk_rad', np.float64), terrain=f('terrain', np.float64), alive=f('alive', np.uint8),
                    just_died=f('justDied', np.uint8), hit=f('hit', np.uint8), was_hit=f('wasHit', np.uint8), bsa=bsa,
                    init_theta=np.float64(self.world.init_theta), time_step=np.int32(self.world.time_step),
                    obs_counter=np.uint32(self.obs_counter & 0xFFFFFFFF), episode=np.uint32(self.episode & 0xFFFFFFFF))

    def reset(self):
        ob, info = self.wrapper.reset()
        return np.asarray(ob), info

    def step(self, act):
        ob, reward, done, info = self.wrapper.step(np.asarray(act, dtype=np.float64))
        return np.asarray(ob), np.asarray(reward, dtype=np.float64), bool(done), info
