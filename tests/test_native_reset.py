"""The native reset engine (csrc/mg_reset.cpp) against its Python specification
(magical_b200/tasks.py, which consumes a real numpy RandomState): byte-identical scenes and identical
RandomState streams afterwards, for every registered base env and several seeds.  CPU only."""
import ctypes as C
import time

import numpy as np
import pytest

import magical_b200 as m
from magical_b200 import _lib as L, tasks

SKIP_KW = ('res_hw', 'fps', 'phys_steps', 'phys_iter', 'max_episode_steps')


def _kwargs(env_id):
    sp = m.spec(env_id)
    kw = {k: v for k, v in sp['kwargs'].items() if k not in SKIP_KW}
    rand_dynamics = kw.pop('rand_dynamics', False)
    return sp['entry_point'], sp['max_episode_steps'], rand_dynamics, kw


def _base_ids():
    m.register_envs()
    return [i for i in m.ALL_REGISTERED_ENVS if m.EnvName(i).preproc is None] + ['MoveToCorner-Demo-DebugReward-v0']


def test_mt19937_stream_matches_numpy():
    """The engine advances RandomState.get_state() exactly like numpy does for the draws the tasks make."""
    ep, horizon, rd, kw = _kwargs('ClusterColour-TestAll-v0')   # uniform, randint, choice, shuffle, rejection
    for seed in (0, 1, 12345, 2 ** 32 - 1):
        a, b = np.random.RandomState(seed), np.random.RandomState(seed)
        try:
            tasks.sample_scene(ep, a, horizon, rand_dynamics=rd, **kw)
            tasks.sample_scene_native(ep, b, horizon, rand_dynamics=rd, **kw)
        except tasks.PlacementError:
            pass
        sa, sb = a.get_state(), b.get_state()
        assert sa[2] == sb[2] and (sa[1] == sb[1]).all()
        assert a.randint(0, 2 ** 31 - 1) == b.randint(0, 2 ** 31 - 1)


@pytest.mark.parametrize('env_id', _base_ids())
def test_native_scene_equals_python_scene(env_id):
    ep, horizon, rd, kw = _kwargs(env_id)
    n_ok = 0
    for seed in range(12):
        a, b = np.random.RandomState(seed), np.random.RandomState(seed)
        for episode in range(2):       # consecutive resets of one env continue one stream
            err_py = err_nat = None
            try:
                want = tasks.sample_scene(ep, a, horizon, rand_dynamics=rd, **kw)
            except tasks.PlacementError as e:
                err_py = e
            try:
                got = tasks.sample_scene_native(ep, b, horizon, rand_dynamics=rd, **kw)
            except tasks.PlacementError as e:
                err_nat = e
            assert (err_py is None) == (err_nat is None), (env_id, seed, episode)
            sa, sb = a.get_state(), b.get_state()
            assert sa[2] == sb[2] and (sa[1] == sb[1]).all(), (env_id, seed, episode)
            if err_py is None:
                assert bytes(want) == bytes(got), (env_id, seed, episode)
                n_ok += 1
    assert n_ok > 0 or 'TestAll' in env_id or 'TestJitter' in env_id


def test_batch_engine_matches_sequential_and_is_faster():
    env_id = 'ClusterColour-TestAll-v0'
    ep, horizon, rd, kw = _kwargs(env_id)
    n = 192
    spec = tasks.reset_spec(ep, horizon, rd, **kw)
    rngs_a = [np.random.RandomState(100 + i) for i in range(n)]
    rngs_b = [np.random.RandomState(100 + i) for i in range(n)]
    t0 = time.perf_counter()
    want = []
    for r in rngs_a:
        try:
            want.append(tasks.sample_scene(ep, r, horizon, rand_dynamics=rd, **kw))
        except tasks.PlacementError:
            want.append(None)
    t_py = time.perf_counter() - t0
    ok = [i for i, w in enumerate(want) if w is not None]
    t0 = time.perf_counter()
    got = tasks.sample_scenes_native([spec] * len(ok), [rngs_b[i] for i in ok], n_threads=4)
    t_nat = time.perf_counter() - t0
    for k, i in enumerate(ok):
        assert bytes(want[i]) == bytes(got[k]), i
        assert (rngs_a[i].get_state()[1] == rngs_b[i].get_state()[1]).all()
    print(f'{len(ok)} ClusterColour-TestAll resets: python {t_py * 1e3:.0f} ms, native x4 threads {t_nat * 1e3:.0f} ms')
    assert t_nat < t_py
