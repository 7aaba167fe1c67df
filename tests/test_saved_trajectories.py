"""Demo-trajectory format and re-preprocessing (magical/saved_trajectories.py:14-149).  The GPU tests push stored
full-resolution frames through mg_frames_downsample / mg_frames_stack and compare with cv2.resize + numpy
restatements of the wrapper stack (benchmarks/__init__.py:51-216) and with the live env under the same
preprocessor."""
import collections
import gzip
import pickle
import sys
import types

import numpy as np
import pytest


def _ref_wrapper_stack(allo, ego, preproc):
    """numpy/cv2 restatement of the reference wrapper stacks over a stored trajectory (test oracle)."""
    import cv2

    def rs(x):
        return cv2.resize(x, (96, 96), interpolation=cv2.INTER_AREA)
    n = allo.shape[0]
    out = []
    if preproc == 'LoResStack':
        da, de = collections.deque(maxlen=4), collections.deque(maxlen=4)
        for t in range(n):
            fa, fe = rs(allo[t]), rs(ego[t])
            for _ in range(4 if t == 0 else 1):
                da.append(fa); de.append(fe)
            out.append(collections.OrderedDict(allo=np.concatenate(da, -1), ego=np.concatenate(de, -1)))
        return out
    n_allo, n_ego = {'LoRes3EA': (1, 3), 'LoRes4E': (0, 4), 'LoRes4A': (4, 0), 'LoResCHW4E': (0, 4), 'LoResCHW4A': (0, 4)}[preproc]
    da, de = collections.deque(maxlen=n_allo), collections.deque(maxlen=n_ego)
    for t in range(n):
        for _ in range(max(n_allo, n_ego) if t == 0 else 1):
            da.append(allo[t]); de.append(ego[t])
        past = np.concatenate([*da, *de], axis=-1)
        o = collections.OrderedDict(allo=rs(allo[t]), ego=rs(ego[t]), past_obs=rs(past))
        if 'CHW' in preproc:
            o = collections.OrderedDict((k, np.moveaxis(v, -1, 0)) for k, v in o.items())
        out.append(o)
    return out


def test_load_demos_rewrites_trajectory_classes(tmp_path):
    from magical_b200.saved_trajectories import MAGICALTrajectory, load_demos
    # a pickle that refers to the reference's own class path, as its demo files do
    mod = types.ModuleType('magical.saved_trajectories')
    Fake = collections.namedtuple('MAGICALTrajectory', ['acts', 'obs', 'rews', 'infos'])
    Fake.__module__ = 'magical.saved_trajectories'
    mod.MAGICALTrajectory = Fake
    sys.modules['magical'] = types.ModuleType('magical')
    sys.modules['magical.saved_trajectories'] = mod
    try:
        traj = Fake(acts=np.arange(3), obs={'allo': np.zeros((4, 2, 2, 3), np.uint8)}, rews=np.zeros(3), infos=[{}] * 3)
        path = tmp_path / 'demo-0.pkl.gz'
        with gzip.GzipFile(path, 'wb') as fp:
            pickle.dump({'env_name': 'MoveToCorner-Demo-v0', 'trajectory': traj, 'score': 0.5}, fp)
    finally:
        del sys.modules['magical.saved_trajectories'], sys.modules['magical']
    demos = list(load_demos([str(path)]))
    assert len(demos) == 1 and demos[0]['env_name'] == 'MoveToCorner-Demo-v0' and demos[0]['score'] == 0.5
    t = demos[0]['trajectory']
    assert isinstance(t, MAGICALTrajectory) and (t.acts == np.arange(3)).all() and t.obs['allo'].shape == (4, 2, 2, 3)


def test_splice_and_unknown_preproc():
    from magical_b200.saved_trajectories import splice_in_preproc_name
    assert splice_in_preproc_name('MoveToCorner-Demo-v0', 'LoResStack') == 'MoveToCorner-Demo-LoResStack-v0'
    with pytest.raises(AssertionError):
        splice_in_preproc_name('MoveToCorner-Demo-v0', 'NoSuchPreproc')


def test_wrapper_arg_is_refused():
    from magical_b200.saved_trajectories import preprocess_demos_with_wrapper
    with pytest.raises(NotImplementedError):
        preprocess_demos_with_wrapper([], 'MoveToCorner-Demo-v0', wrapper=lambda e: e)


@pytest.mark.gpu
@pytest.mark.parametrize('preproc', ['LoRes3EA', 'LoRes4E', 'LoRes4A', 'LoResStack', 'LoResCHW4E', 'LoResCHW4A'])
def test_preprocess_frames_matches_cv2_and_numpy(preproc):
    """Random frames (so every rounding tie and channel is exercised): bit-exact against cv2 + numpy."""
    from magical_b200.saved_trajectories import preprocess_frames
    rng = np.random.RandomState(3)
    n = 7
    allo = rng.randint(0, 256, size=(n, 384, 384, 3)).astype(np.uint8)
    ego = rng.randint(0, 256, size=(n, 384, 384, 3)).astype(np.uint8)
    allo[1, :8] = 8 * (rng.randint(0, 32, size=(8, 384, 3)))      # multiples of 8: sums hit the .5 ties
    got = preprocess_frames(allo, ego, preproc)
    want = _ref_wrapper_stack(allo, ego, preproc)
    assert list(got) == list(want[0])
    for key in got:
        assert got[key].dtype == np.uint8
        for t in range(n):
            assert got[key][t].shape == want[t][key].shape, (key, got[key][t].shape, want[t][key].shape)
            assert (got[key][t] == want[t][key]).all(), (preproc, key, t)


@pytest.mark.gpu
def test_demo_replay_equals_live_preprocessed_env():
    """Frames recorded from the un-preprocessed env, re-preprocessed as a demo, equal what the LoRes4E env
    returned live for the same seed and actions (the property preprocess_demos_with_wrapper exists for)."""
    import magical_b200 as m
    from magical_b200.saved_trajectories import MAGICALTrajectory, preprocess_demos_with_wrapper
    T = 12
    acts = np.random.RandomState(1).randint(0, 18, size=T)
    raw = m.make('MatchRegions-Demo-v0', seed=7)
    frames = [raw.reset()]
    rews, infos = [], []
    for a in acts:
        o, r, d, info = raw.step(a)
        frames.append(o); rews.append(r); infos.append(info)
    raw.close()
    traj = MAGICALTrajectory(acts=acts, obs=np.stack(frames, axis=0) if False else frames, rews=np.asarray(rews), infos=infos)
    new, = preprocess_demos_with_wrapper([traj], 'MatchRegions-Demo-v0', preproc_name='LoRes4E')
    assert isinstance(new, MAGICALTrajectory) and len(new.obs) == T + 1 and len(new.acts) == T
    live = m.make('MatchRegions-Demo-LoRes4E-v0', seed=7)
    o = live.reset()
    for t in range(T + 1):
        assert list(new.obs[t]) == ['allo', 'ego', 'past_obs']
        for key in o:
            assert (new.obs[t][key] == o[key]).all(), (t, key)
        if t < T:
            o, r, d, info = live.step(acts[t])
            assert r == new.rews[t] and info['eval_score'] == new.infos[t]['eval_score']
    live.close()


@pytest.mark.gpu
def test_frames_entry_points_reject_bad_arguments():
    import ctypes as C
    import torch
    import magical_b200._lib as L
    lib = L.lib()
    assert lib.mg_frames_downsample(0, None, 0, None, None) == 0              # empty trajectory: nothing to do
    assert lib.mg_frames_downsample(0, None, 3, None, None) == -1
    x = torch.zeros(16 + 384 * 384 * 3, dtype=torch.uint8, device='cuda')
    out = torch.zeros(96 * 96 * 3, dtype=torch.uint8, device='cuda')
    assert lib.mg_frames_downsample(0, x.data_ptr() + 1, 1, out.data_ptr(), None) == -1   # misaligned
    assert b'aligned' in lib.mg_last_error(None)
    v = (C.c_int32 * 1)(5)
    a = (C.c_int32 * 1)(0)
    assert lib.mg_frames_stack(0, out.data_ptr(), out.data_ptr(), 1, 1, C.addressof(v), C.addressof(a), 0, out.data_ptr(), None) == -1
