"""ORACLE / TEST INFRASTRUCTURE -- numpy restatement of the conc_4hist history pool: ShellEnvWrapper.interact_with_env
(ALGORITHM/conc_4hist/shell_env.py:25-66), solve_duplicate (:68-99), roll_hisory (cython_func.pyx:32-53), in the
uncompacted layout (paused envs keep their slot; their feed is NaN).  Pinned to the reference by
tests/golden/histpool_cases.npz (oracle/make_golden_histpool.py).  Only tests/ may import this."""
import numpy as np


def histpool_roll(obs, his_pool, pause, suffered):
    """obs [E,G,192] f32, his_pool [E,G,12,16] f32 -> (feed [E,G,24,16], next his_pool [E,G,12,16])."""
    E, G = obs.shape[:2]
    new = obs.reshape(E, G, 12, 16).copy()
    prev = his_pool.copy()
    prev[suffered] = 0                                            # shell_env.py:37
    feed = np.concatenate((new, prev), axis=-2)                   # :74
    feed[(feed == 0).all(-1)] = np.nan                            # :77
    new[(new == 0).all(-1)] = np.nan                              # :78
    valid = ~np.isnan(new).any(-1)                                # :79
    valid[:, :, 0] = False                                        # :82
    nxt = np.zeros_like(prev)
    for e in range(E):                                            # cython_func.pyx:44-52
        for a in range(G):
            rows = new[e, a][valid[e, a]]
            n = len(rows)
            nxt[e, a, :n] = rows
            nxt[e, a, n:] = prev[e, a, :12 - n]
    alive = ~np.isnan(new[:, :, 0]).any(-1)                       # :94
    feed[~alive] = np.nan                                         # :96
    feed[pause] = np.nan       # not fed at all in the reference (obs[~P], :40)
    nxt[pause] = 0                                                # :44
    return feed, nxt
