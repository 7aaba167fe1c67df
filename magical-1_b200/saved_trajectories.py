"""Demo-trajectory format and re-preprocessing, the data format either side of the step path.

Mirrors magical/saved_trajectories.py:
  MAGICALTrajectory (:14-22)            same NamedTuple (acts, obs, rews, infos)
  load_demos (:37-49)                   gzip + pickle of {'env_name', 'trajectory', 'score'} dictionaries, with
                                        the class references of older pickles rewritten (:25-34)
  splice_in_preproc_name (:52-60)       re-exported from the package
  preprocess_demos_with_wrapper (:87-149)
The reference replays every stored 384x384 frame through a mock env wrapped in the preprocessor's gym wrappers
(one cv2.resize + one np.concatenate per frame and key).  Here the stored frames of a trajectory go to the GPU
once and the two kernels of csrc/mg_frames.cu (mg_frames_downsample, mg_frames_stack) produce every
observation of the trajectory; the result has the reference's structure (an object array of per-step
OrderedDicts in key order allo, ego[, past_obs]).  There is no CPU path: without the CUDA library this raises.
"""
import collections
import ctypes as C
import gzip
from pickle import Unpickler
from typing import List, NamedTuple, Optional

import numpy as np

from . import _lib as L
from .env import PREPROC_SPECS


class MAGICALTrajectory(NamedTuple):
    """Trajectory representation compatible with imitation's trajectory data class
    (saved_trajectories.py:14-22): len(obs) == len(acts) + 1."""
    acts: np.ndarray
    obs: dict
    rews: np.ndarray
    infos: Optional[List[dict]]


# classes that demo pickles written by older code refer to (saved_trajectories.py:25-34), plus the reference's
# own module path, so that its demo files load without the reference installed
_TRAJECTORY_CLASSES = {
    ('imitation.util.rollout', 'Trajectory'),
    ('milbench.baselines.saved_trajectories', 'MILBenchTrajectory'),
    ('magical.saved_trajectories', 'MAGICALTrajectory'),
}


class _TrajRewriteUnpickler(Unpickler):
    def find_class(self, module, name):
        if (module, name) in _TRAJECTORY_CLASSES:
            return MAGICALTrajectory
        return super().find_class(module, name)


def load_demos(demo_paths, rewrite_traj_cls=True, verbose=False):
    """Generator over the demo dictionaries stored at `demo_paths` (saved_trajectories.py:37-49)."""
    total = len(demo_paths)
    for k, path in enumerate(demo_paths, start=1):
        if verbose:
            print(f"Loading '{path}' ({k}/{total})")
        with gzip.GzipFile(path, 'rb') as fp:
            yield (_TrajRewriteUnpickler(fp) if rewrite_traj_cls else Unpickler(fp)).load()


def splice_in_preproc_name(base_env_name, preproc_name):
    from . import splice_in_preproc_name as _splice
    return _splice(base_env_name, preproc_name)


def _obs_column(obs, key):
    """Frames of one key of a stored trajectory as one u8 [T+1, H, W, 3] array.  Stored observations are either
    a dict of stacked arrays or a sequence of per-step dicts (what np.stack of dicts leaves behind)."""
    if isinstance(obs, dict):
        arr = obs[key]
    else:
        arr = np.stack([o[key] for o in obs], axis=0)
    arr = np.ascontiguousarray(arr, dtype=np.uint8)
    if arr.ndim != 4 or arr.shape[1:] != (L.RES, L.RES, 3):
        raise ValueError(f"demo frames for '{key}' must be u8 [T+1, {L.RES}, {L.RES}, 3], got {arr.shape}")
    return arr


def preprocess_frames(allo, ego, preproc_name, device=0):
    """The observations `preproc_name`'s wrapper stack yields over a stored trajectory.
    allo, ego: u8 [N, 384, 384, 3] (frame 0 = the reset observation).  Returns an OrderedDict of u8 numpy arrays
    with leading dimension N, keys and layouts as the env returns them for that preprocessor."""
    import torch
    if preproc_name not in PREPROC_SPECS:
        raise KeyError(f"no preprocessor named '{preproc_name}', options are {', '.join(PREPROC_SPECS)}")
    if not torch.cuda.is_available():
        raise RuntimeError('magical_b200 needs a CUDA device: there is no CPU path')
    spec = PREPROC_SPECS[preproc_name]
    lib = L.lib()
    dev = torch.device('cuda', int(device))
    n = allo.shape[0]
    assert ego.shape[0] == n
    st = torch.cuda.current_stream(dev).cuda_stream

    def check(rc):
        if rc != 0:
            raise L.MgError(f'libmagical_b200 error {rc}: {lib.mg_last_error(None).decode()}')

    lo = {}
    for view, frames in ((L.VIEW_ALLO, allo), (L.VIEW_EGO, ego)):
        full = torch.from_numpy(frames).to(dev)
        out = torch.empty((n, L.LORES, L.LORES, 3), dtype=torch.uint8, device=dev)
        check(lib.mg_frames_downsample(dev.index, L.ptr(full), n, L.ptr(out), st))
        lo[view] = out
        del full
    groups = spec['groups']
    views = (C.c_int32 * len(groups))(*[g[0] for g in groups])
    ages = (C.c_int32 * len(groups))(*[g[1] for g in groups])
    ch = 3 * len(groups)
    shape = (n, ch, L.LORES, L.LORES) if spec['chw'] else (n, L.LORES, L.LORES, ch)
    stacked = torch.empty(shape, dtype=torch.uint8, device=dev)
    check(lib.mg_frames_stack(dev.index, L.ptr(lo[L.VIEW_ALLO]), L.ptr(lo[L.VIEW_EGO]), n, len(groups),
                              C.addressof(views), C.addressof(ages), int(spec['chw']), L.ptr(stacked), st))
    out = collections.OrderedDict()
    if spec['kind'] == 'stack':
        # EagerDictFrameStack: each key is its own 4-frame stack (benchmarks/__init__.py:51-82)
        out['allo'] = stacked[..., :12].cpu().numpy()
        out['ego'] = stacked[..., 12:].cpu().numpy()
        return out
    for key, view in (('allo', L.VIEW_ALLO), ('ego', L.VIEW_EGO)):
        img = lo[view].permute(0, 3, 1, 2).contiguous() if spec['chw'] else lo[view]
        out[key] = img.cpu().numpy()
    out['past_obs'] = stacked.cpu().numpy()
    return out


def preprocess_demos_with_wrapper(trajectories, orig_env_name, preproc_name=None, wrapper=None, device=0):
    """Preprocess stored trajectories with one of the built-in pipelines (saved_trajectories.py:87-149).
    `wrapper` (an arbitrary gym wrapper constructor) is not supported: the pipelines are kernels here, not
    gym wrappers.  Returns trajectories of the input's type whose `obs` is an object array of per-step
    OrderedDicts, exactly what the reference's np.stack over per-step dicts yields."""
    from . import spec as env_spec
    if wrapper is not None:
        raise NotImplementedError('only the registered preprocessors (preproc_name=...) are available')
    assert preproc_name is not None
    env_spec(orig_env_name)          # same failure as gym.make(orig_env_name) for unknown ids
    out = []
    for traj in trajectories:
        n_acts = len(traj.acts)
        allo, ego = _obs_column(traj.obs, 'allo'), _obs_column(traj.obs, 'ego')
        if allo.shape[0] != n_acts + 1:
            raise ValueError(f'trajectory has {n_acts} actions but {allo.shape[0]} observations')
        cols = preprocess_frames(allo, ego, preproc_name, device=device)
        obs = np.empty(n_acts + 1, dtype=object)
        for t in range(n_acts + 1):
            obs[t] = collections.OrderedDict((k, v[t]) for k, v in cols.items())
        infos = [dict(i or {}) for i in traj.infos] if traj.infos is not None else [{} for _ in range(n_acts)]
        out.append(type(traj)(obs=obs, acts=np.stack(list(traj.acts), axis=0), rews=np.stack(list(traj.rews), axis=0),
                              infos=infos))
    return out
