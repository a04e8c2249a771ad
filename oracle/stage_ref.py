"""ORACLE / TEST INFRASTRUCTURE -- stages the UNMODIFIED reference path into the git-ignored oracle/_ref/.

The reference's own CPU implementation of the hot path is Python + Cython (MISSION/dca, UTIL/shm_env.py,
UTIL/shm_pool.pyx ...).  /root/reference does not exist on the GPU box, so `__graft_entry__.build()` (run in the build
container, where the tree is present) calls `stage()`: it copies exactly the files listed in MANIFEST below from where
they lie under /root/reference into oracle/_ref/tree/ (git-ignored, NOT gpurun-ignored: it travels to the GPU box
like our own built .so files), compiles the three .pyx files the path needs with cython + gcc next to their sources
(so no pyximport build is needed at run time), copies the shipped checkpoints the policy-rollout tests load, and writes
oracle/_ref/MANIFEST.json (sha256 per file).  Nothing staged is ever committed, modified or imported by the product
package; users are bench.py's `--impl reference` / `cpu_baseline` legs and tests/ (SmartPool + SuperpoolEnv timing,
Runner + MMPlatform + ALGORITHM driving our DcaSuperpoolEnv).

    python oracle/stage_ref.py            # stage (idempotent)
    python oracle/stage_ref.py --check    # is a usable staged tree present?
"""
import hashlib
import json
import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get('HMP_REFERENCE_ROOT', '/root/reference')
OUT = os.path.join(HERE, '_ref')
TREE = os.path.join(OUT, 'tree')
SHIM = os.path.join(HERE, 'ref_harness', 'shim')  # gym / func_timeout import shims (ours, committed)

# files (or whole directories) of the reference that the path and its callers import -- found by running the path
# under an import trace (SmartPool + make_parallel_envs + SuperpoolEnv.step; task_runner.Runner + multi_team.MMPlatform
# + ALGORITHM.random / conc / conc_4hist)
MANIFEST = [
    'config.py', 'task_runner.py', 'multi_team.py',
    'MISSION/env_router.py', 'MISSION/common', 'MISSION/dca', 'MISSION/dca_multiteam',
    'UTIL/__init__.py', 'UTIL/colorful.py', 'UTIL/config_args.py', 'UTIL/data_struct.py', 'UTIL/exp_helper.py',
    'UTIL/file_lock.py', 'UTIL/gpu_share.py', 'UTIL/hmp_daemon.py', 'UTIL/network.py', 'UTIL/shm_env.py',
    'UTIL/shm_pool.pyx', 'UTIL/tensor_ops.py', 'UTIL/hidden_print.py',
    'ALGORITHM/common', 'ALGORITHM/conc', 'ALGORITHM/conc_4hist', 'ALGORITHM/random',
    'RESULT/50RL-55opp/model.pt', 'RESULT/50RL-55opp/test-50RL-55opp.jsonc',
    'RESULT/test-50+50/model.pt', 'RESULT/test-50+50/test-50+50.jsonc',
    'RESULT/test-100+100/model.pt', 'RESULT/test-100+100/test-100+100.jsonc',
]
PYX = ['UTIL/shm_pool.pyx', 'MISSION/dca/cython_func.pyx', 'MISSION/dca_multiteam/cython_func.pyx',
       'ALGORITHM/conc_4hist/cython_func.pyx', 'ALGORITHM/common/cython_func.pyx']


def reference_present():
    return os.path.isdir(os.path.join(REF_ROOT, 'MISSION', 'dca'))


def staged():
    """True if a staged tree with its compiled extensions is present (what the GPU box sees)."""
    if not os.path.exists(os.path.join(OUT, 'MANIFEST.json')):
        return False
    suffix = sysconfig.get_config_var('EXT_SUFFIX')
    return all(os.path.exists(os.path.join(TREE, p[:-4] + suffix)) for p in PYX[:2])


def _sha(path):
    h = hashlib.sha256()
    with open(path, 'rb') as f:
        h.update(f.read())
    return h.hexdigest()


def _build_pyx(rel):
    """cython + gcc -> <module><EXT_SUFFIX> next to the staged .pyx (what pyximport would build, main.py:65)."""
    import numpy
    src = os.path.join(TREE, rel)
    if not os.path.exists(src):
        return None
    work = os.path.join(OUT, 'build')
    os.makedirs(work, exist_ok=True)
    cfile = os.path.join(work, rel.replace('/', '_')[:-4] + '.c')
    subprocess.check_call([sys.executable, '-m', 'cython', '-3', src, '-o', cfile], cwd=TREE)
    out = src[:-4] + sysconfig.get_config_var('EXT_SUFFIX')
    subprocess.check_call(['gcc', '-O2', '-fPIC', '-shared', '-fno-strict-aliasing', '-w',
                           '-I', sysconfig.get_paths()['include'], '-I', numpy.get_include(), cfile, '-o', out])
    return out


def stage(force=False):
    """Copy + compile.  Returns the staged tree path, or None when /root/reference is absent and nothing is staged."""
    if not reference_present():
        return TREE if staged() else None
    if staged() and not force:
        return TREE
    if os.path.isdir(TREE):
        shutil.rmtree(TREE)
    os.makedirs(TREE)
    files = {}
    for rel in MANIFEST:
        src = os.path.join(REF_ROOT, rel)
        if os.path.isdir(src):
            for root, _, names in os.walk(src):
                for n in names:
                    if n.endswith(('.pyc', '.so', '.c')):
                        continue
                    s = os.path.join(root, n)
                    r = os.path.relpath(s, REF_ROOT)
                    d = os.path.join(TREE, r)
                    os.makedirs(os.path.dirname(d), exist_ok=True)
                    shutil.copyfile(s, d)
                    files[r] = _sha(d)
        elif os.path.exists(src):
            d = os.path.join(TREE, rel)
            os.makedirs(os.path.dirname(d), exist_ok=True)
            shutil.copyfile(src, d)
            files[rel] = _sha(d)
    for root, dirs, names in os.walk(TREE):  # staged copies are writable (the reference tree is read-only)
        for n in names:
            os.chmod(os.path.join(root, n), 0o644)
    built = [os.path.relpath(p, TREE) for p in (_build_pyx(rel) for rel in PYX) if p]
    with open(os.path.join(OUT, 'MANIFEST.json'), 'w') as f:
        json.dump({'reference_root': REF_ROOT, 'files': files, 'built': built,
                   'note': 'unmodified copies; git-ignored; never imported by hmp2g_b200'}, f, indent=1, sort_keys=True)
    return TREE


def activate(workdir=None):
    """Make the staged reference importable in this process (sys.path: shims + staged tree; cwd: a writable scratch
    directory, the reference writes ./TEMP and ./RESULT relative to it).  Returns the tree path."""
    assert staged(), 'oracle/_ref is not staged: run __graft_entry__.build() where /root/reference exists'
    workdir = workdir or os.path.join(OUT, 'work')
    os.makedirs(workdir, exist_ok=True)
    os.chdir(workdir)
    for p in (SHIM, TREE):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    for k in ('OMP_NUM_THREADS', 'MKL_NUM_THREADS', 'OPENBLAS_NUM_THREADS', 'NUM_THREADS'):  # main.py:24-29
        os.environ[k] = '1'
    return TREE


if __name__ == '__main__':
    if '--check' in sys.argv:
        print('staged' if staged() else 'not staged')
        sys.exit(0 if staged() else 1)
    print(stage(force='--force' in sys.argv))
