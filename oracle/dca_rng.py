"""ORACLE / TEST INFRASTRUCTURE -- not part of the product path.

Keyed random stream shared by the reference harness, the C oracle and the CUDA
kernels.  The reference draws from the process-global ``np.random`` stream at
three sites:

* reset:  /root/reference/MISSION/dca/collective_assault_env.py:64 (theta),
          :70-73 (attacker fluctuation), :85 (1 uniform / agent), :97,:120
          (2 normals / agent);
* observation: ``item_random_mv`` :217-226 -- ``np.random.shuffle`` then
          ``np.random.rand(len)`` per (alive guard, hostile list / friend list).

BASELINE.json's north_star asks for "the reference's RNG draws replaced by a
shared pre-generated stream".  The stream here is *keyed* (counter based,
Philox4x32-10) rather than sequential so that a GPU thread can produce the
draws of (env, observation-pass, guard, list) without knowing how many draws
other guards consumed (dead guards draw nothing in the reference).

Counter layout (4 x u32):   c0 = global env id
                            c1 = per-env pass counter (obs passes / episodes)
                            c2 = item  (obs: guard*2 + list ; reset: 0 = world, 1+i = agent i)
                            c3 = site | (block << 8)      site: 1 = obs, 2 = reset
Key (2 x u32): (seed & 0xffffffff, seed >> 32).

Observation words (12 per list, 3 Philox blocks; word w = block w//4, lane w%4):
    jam uniform k (k = 0..5)      -> word JAM_W[k],  u = word * 2**-32
    shuffle draw for position i   -> word SHUF_W[i], j = (word * (i+1)) >> 32
The interleave makes short visible lists need only block 0.
Fisher-Yates runs i = n-1 .. 1 (same direction as numpy's legacy shuffle).

Reset words: world item (c2 = 0): word0 -> theta uniform, word1 -> fluctuation
count; agent item (c2 = 1+i): word0 -> atk_rad uniform, word1/word2 -> Box-Muller
pair (u1 = (w1+1)*2**-32, u2 = w2*2**-32), word3 -> exclusion sort key.
"""
import numpy as np

SITE_OBS = 1
SITE_RESET = 2
JAM_W = (0, 1, 3, 5, 7, 9)
SHUF_W = (None, 2, 4, 6, 8, 10)  # index = position i (1..5)
WORDS_PER_LIST = 12

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10.  All inputs broadcastable integer arrays.
    Returns a tuple of four uint32 arrays."""
    c0, c1, c2, c3 = np.broadcast_arrays(*[np.asarray(x, dtype=np.uint64) & _MASK
                                           for x in (c0, c1, c2, c3)])
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        n0 = (hi1 ^ c1 ^ np.uint64(k0)) & _MASK
        n1 = lo1
        n2 = (hi0 ^ c3 ^ np.uint64(k1)) & _MASK
        n3 = lo0
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return tuple(x.astype(np.uint32) for x in (c0, c1, c2, c3))


def _words(seed, env_gid, counter, item, site, nblocks):
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    out = []
    for b in range(nblocks):
        out.extend(int(x) for x in philox4x32(env_gid, counter, item, site | (b << 8), k0, k1))
    return out


def obs_words(seed, env_gid, obs_counter, guard, lst):
    """12 u32 words for (env, observation pass, guard, list); list 0 = hostile, 1 = friend."""
    return _words(seed, env_gid, obs_counter, guard * 2 + lst, SITE_OBS, 3)


def obs_words_all(seed, env_gids, obs_counters, n_guards):
    """uint32 [E, G, 2, 12] stream for a whole batch (what `rng_stream` in the C-ABI takes)."""
    env_gids = np.asarray(env_gids, dtype=np.uint64).reshape(-1, 1, 1, 1)
    obs_counters = np.asarray(obs_counters, dtype=np.uint64).reshape(-1, 1, 1, 1)
    item = (np.arange(n_guards, dtype=np.uint64).reshape(1, -1, 1, 1) * np.uint64(2)
            + np.arange(2, dtype=np.uint64).reshape(1, 1, 2, 1))
    blk = np.arange(3, dtype=np.uint64).reshape(1, 1, 1, 3)
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    w = philox4x32(env_gids, obs_counters, item, np.uint64(SITE_OBS) | (blk << np.uint64(8)), k0, k1)
    # w[l] has shape [E,G,2,3(block)] -> words [E,G,2,3,4] -> [E,G,2,12]
    return np.stack(w, axis=-1).reshape(w[0].shape[0], n_guards, 2, 12)


def keyed_shuffle(arr, words):
    """In-place Fisher-Yates of `arr` (len <= 6) with the list's words."""
    n = len(arr)
    for i in range(n - 1, 0, -1):
        j = (int(words[SHUF_W[i]]) * (i + 1)) >> 32
        arr[i], arr[j] = arr[j], arr[i]
    return arr


def keyed_jam_uniforms(n, words):
    return np.array([int(words[JAM_W[k]]) * 2.0 ** -32 for k in range(n)], dtype=np.float64)


def reset_world_words(seed, env_gid, episode):
    return _words(seed, env_gid, episode, 0, SITE_RESET, 1)


def reset_agent_words(seed, env_gid, episode, agent):
    return _words(seed, env_gid, episode, 1 + agent, SITE_RESET, 1)


def reset_draws(seed, env_gid, episode, n_agents):
    """The draws reset_world consumes, in the reference's call order.

    Returns dict(theta_u, fluct_word, agent_u[N], agent_n[N,2], agent_key[N])."""
    w = reset_world_words(seed, env_gid, episode)
    out = dict(theta_u=w[0] * 2.0 ** -32, fluct_word=w[1],
               agent_u=np.zeros(n_agents), agent_n=np.zeros((n_agents, 2)),
               agent_key=np.zeros(n_agents, dtype=np.uint32))
    for i in range(n_agents):
        a = reset_agent_words(seed, env_gid, episode, i)
        out['agent_u'][i] = a[0] * 2.0 ** -32
        u1 = (a[1] + 1) * 2.0 ** -32
        u2 = a[2] * 2.0 ** -32
        r = np.sqrt(-2.0 * np.log(u1))
        out['agent_n'][i, 0] = r * np.cos(2.0 * np.pi * u2)
        out['agent_n'][i, 1] = r * np.sin(2.0 * np.pi * u2)
        out['agent_key'][i] = a[3]
    return out
