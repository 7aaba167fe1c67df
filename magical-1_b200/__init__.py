"""magical_b200 -- B200-native batched MAGICAL simulator behind the reference's env-id API.

Drop-in surface (reference: magical/__init__.py:1-4, magical/benchmarks/__init__.py:308-432,
1017-1102): ``register_envs()``, ``ALL_REGISTERED_ENVS``, ``DEMO_ENVS_TO_TEST_ENVS_MAP``,
``AVAILABLE_PREPROCESSORS``, ``EnvName``, ``update_magical_env_name``, ``splice_in_preproc_name``
and ``make(id)`` standing in for ``gym.make(id)`` (gym itself is not a dependency).
``make(id, num_envs=B, device=...)`` returns the vectorised env.
"""
import collections
import re

import numpy as np

from . import _lib, tasks  # noqa: F401
from ._registry_table import BASE_ENVS
from .env import PREPROC_SPECS, BatchedMagicalEnv

__all__ = ['register_envs', 'make', 'EnvName', 'update_magical_env_name', 'splice_in_preproc_name',
           'ALL_REGISTERED_ENVS', 'DEMO_ENVS_TO_TEST_ENVS_MAP', 'AVAILABLE_PREPROCESSORS',
           'BatchedMagicalEnv', 'MagicalEnv']

DEFAULT_RES = (384, 384)                                      # benchmarks/__init__.py:28
COMMON_KWARGS = dict(res_hw=DEFAULT_RES, fps=8, phys_steps=10, phys_iter=10)   # :434-437
AVAILABLE_PREPROCESSORS = list(PREPROC_SPECS)
_ENV_NAME_RE = re.compile(
    r'^(?P<name_prefix>[^-]+)(?P<demo_test_spec>-(Demo|Test[^-]*))'
    r'(?P<env_name_suffix>(-[^-]+)*)(?P<version_suffix>-v\d+)$')
_REGISTERED = False
_REGISTRY = collections.OrderedDict()   # id -> dict(entry_point, kwargs, max_episode_steps, preproc)
DEMO_ENVS_TO_TEST_ENVS_MAP = collections.OrderedDict()
ALL_REGISTERED_ENVS = []


class EnvName:
    """Parser for '<Task>-<Demo|TestX>[-<Preproc>]-v<N>' (benchmarks/__init__.py:350-424)."""

    def __init__(self, env_name):
        match = _ENV_NAME_RE.match(env_name)
        if match is None:
            raise ValueError(f"env name '{env_name}' does not match _ENV_NAME_RE spec")
        groups = match.groupdict()
        self.name_prefix = groups['name_prefix']
        self.demo_test_spec = groups['demo_test_spec']
        self.env_name_suffix = groups['env_name_suffix']
        self.version_suffix = groups['version_suffix']
        assert env_name == self.env_name
        if not self.is_test:
            assert self.demo_env_name == self.env_name, (self.demo_env_name, self.env_name)

    @property
    def env_name(self):
        return self.name_prefix + self.demo_test_spec + self.env_name_suffix + self.version_suffix

    @property
    def is_test(self):
        return self.demo_test_spec.startswith('-Test')

    @property
    def demo_env_name(self):
        return self.name_prefix + '-Demo' + self.env_name_suffix + self.version_suffix

    @property
    def task(self):
        return self.name_prefix

    @property
    def variant(self):
        return self.demo_test_spec.strip('-')

    @property
    def preproc(self):
        return self.env_name_suffix.strip('-') if self.env_name_suffix else None

    @property
    def version(self):
        return self.version_suffix.strip('-')


def update_magical_env_name(env_name, *, task=None, variant=None, preproc=None, version=None):
    """benchmarks/__init__.py:318-347."""
    ename = EnvName(env_name)
    parts = [task if task is not None else ename.task, variant if variant is not None else ename.variant]
    if preproc is None:
        preproc = ename.preproc
    if preproc is not None:
        parts.append(preproc)
    parts.append(version if version is not None else ename.version)
    return '-'.join(parts)


def splice_in_preproc_name(base_env_name, preproc_name):
    """saved_trajectories.py:52-60 (the sanity check on the preprocessor name included)."""
    assert preproc_name in PREPROC_SPECS, \
        f"no preprocessor named '{preproc_name}', options are {', '.join(PREPROC_SPECS)}"
    return update_magical_env_name(base_env_name, preproc=preproc_name)


def register_envs():
    """Register every env id of the reference (benchmarks/__init__.py:427-1102); idempotent."""
    global _REGISTERED
    if _REGISTERED:
        return False
    _REGISTERED = True
    for entry_point, env_name, ep_len, env_kwargs in BASE_ENVS:
        kwargs = {'max_episode_steps': ep_len, **COMMON_KWARGS, **env_kwargs}
        ALL_REGISTERED_ENVS.append(env_name)
        _REGISTRY[env_name] = dict(entry_point=entry_point, kwargs=kwargs, max_episode_steps=ep_len, preproc=None)
        for preproc in AVAILABLE_PREPROCESSORS:
            new_name = update_magical_env_name(env_name, preproc=preproc)
            ALL_REGISTERED_ENVS.append(new_name)
            _REGISTRY[new_name] = dict(entry_point=entry_point, kwargs=kwargs, max_episode_steps=ep_len,
                                       preproc=preproc)
    train_to_test = {}
    demo_envs = set()
    for name in ALL_REGISTERED_ENVS:
        parsed = EnvName(name)
        if parsed.is_test:
            train_to_test.setdefault(parsed.demo_env_name, []).append(parsed.env_name)
        else:
            demo_envs.add(parsed.env_name)
    assert demo_envs == train_to_test.keys()
    DEMO_ENVS_TO_TEST_ENVS_MAP.update(sorted((k, tuple(v)) for k, v in train_to_test.items()))
    # MoveToCorner-Demo-DebugReward[-<Preproc>]-v0: the reference registers the un-wrapped entry point for
    # all of them, so the preprocessor is not applied (benchmarks/__init__.py:1075-1100)
    mtc_len = 80
    debug_kwargs = {'debug_reward': True, 'max_episode_steps': mtc_len, 'rand_shape_colour': False,
                    'rand_shape_type': False, 'rand_poses': False, **COMMON_KWARGS}
    ep = 'magical.benchmarks.move_to_corner:MoveToCornerEnv'
    for name in ['MoveToCorner-Demo-DebugReward-v0'] + [f'MoveToCorner-Demo-DebugReward-{p}-v0'
                                                         for p in AVAILABLE_PREPROCESSORS]:
        ALL_REGISTERED_ENVS.append(name)
        _REGISTRY[name] = dict(entry_point=ep, kwargs=debug_kwargs, max_episode_steps=mtc_len, preproc=None)
    return True


def spec(env_id):
    register_envs()
    if env_id not in _REGISTRY:
        EnvName(env_id)   # ValueError for malformed ids, like the reference
        raise KeyError(f'No registered env with id: {env_id}')
    return _REGISTRY[env_id]


class Discrete:
    """Minimal gym.spaces.Discrete(18) (base_env.py:109)."""

    def __init__(self, n):
        self.n = n
        self.np_random = np.random.RandomState()
        self.dtype = np.int64
        self.shape = ()

    def seed(self, seed=None):
        self.np_random = np.random.RandomState(seed)
        return [seed]

    def sample(self):
        return int(self.np_random.randint(self.n))

    def contains(self, x):
        return 0 <= int(x) < self.n


class Box:
    def __init__(self, low, high, shape, dtype=np.uint8):
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape), dtype


class MagicalEnv:
    """Single env with the reference's gym protocol (reset / step / render / seed / close), returning
    numpy arrays, float reward, bool done and info={'eval_score': float} like BaseEnv.step
    (base_env.py:267-307) under the registered wrapper stack and gym's TimeLimit."""

    metadata = {'render.modes': ['rgb_array']}

    def __init__(self, env_id, device=0, seed=None, **overrides):
        sp = spec(env_id)
        kwargs = {**sp['kwargs'], **overrides}
        self.spec_id = env_id
        self.max_episode_steps = kwargs['max_episode_steps']
        self.fps = kwargs['fps']
        self._task = sp['entry_point']
        self._batched = BatchedMagicalEnv(sp['entry_point'], kwargs, self.max_episode_steps, preproc=sp['preproc'],
                                          num_envs=1, device=device, seed=seed, auto_reset=False)
        self.action_space = Discrete(_lib.NUM_ACTIONS)
        self.observation_space = collections.OrderedDict()
        self._elapsed = None
        self._pap = self._task.endswith('PickAndPlaceEnv')

    def seed(self, seed=None):
        return self._batched.seed(seed)

    @property
    def episode_steps(self):
        return int(self._batched.episode_steps[0])

    def _to_numpy(self, obs):
        out = collections.OrderedDict((k, v[0].cpu().numpy().copy()) for k, v in obs.items())
        if self._pap:
            # PickAndPlaceEnv.reset/step extras (pick_and_place.py:81-86, 104-112)
            sc = self._batched.scenes[0]
            blk = sc.blocks[0]
            out['target_type'] = int(blk.type)
            out['target_colour'] = int(blk.colour)
            out['target_position'] = np.asarray([sc.target_pos[0], sc.target_pos[1]])
        for k, v in out.items():
            if isinstance(v, np.ndarray) and v.ndim == 3:
                self.observation_space[k] = Box(0, 255, v.shape, np.uint8)
        return out

    def reset(self):
        self._elapsed = 0
        return self._to_numpy(self._batched.reset())

    def step(self, action):
        a = int(action)
        if not 0 <= a < _lib.NUM_ACTIONS:
            raise KeyError(a)    # ACTION_ID_TO_FLAGS[int(action)], base_env.py:125-127
        obs, rew, done, info = self._batched.step([a])
        self._elapsed += 1
        done = bool(done[0].item())
        info = {'eval_score': float(info['eval_score'][0].item())}
        if self._elapsed >= self.max_episode_steps:
            info['TimeLimit.truncated'] = not done
            done = True
        return self._to_numpy(obs), float(rew[0].item()), done, info

    def render(self, mode='rgb_array'):
        obs = self._batched.render(mode)
        return collections.OrderedDict((k, v[0].cpu().numpy().copy()) for k, v in obs.items())

    def get_bodies(self):
        return self._batched.get_bodies()[0]

    def close(self):
        self._batched.close()


def make(env_id, num_envs=None, device=0, seed=None, obs_keys=None, **overrides):
    """gym.make(env_id) stand-in.  num_envs=None -> MagicalEnv (reference protocol, numpy);
    num_envs=B -> BatchedMagicalEnv (torch tensors on `device`, auto-reset)."""
    sp = spec(env_id)
    if num_envs is None:
        return MagicalEnv(env_id, device=device, seed=seed, **overrides)
    kwargs = {**sp['kwargs'], **overrides}
    return BatchedMagicalEnv(sp['entry_point'], kwargs, kwargs['max_episode_steps'], preproc=sp['preproc'],
                             num_envs=num_envs, device=device, seed=seed, obs_keys=obs_keys)


def make_mixed(env_ids, device=0, seed=None, preproc='LoRes4E', obs_keys=('past_obs',)):
    """One batch holding a different registered base env id per env (BASELINE config 5)."""
    specs = []
    for env_id in env_ids:
        sp = spec(env_id)
        specs.append((sp['entry_point'], dict(sp['kwargs']), sp['max_episode_steps']))
    return BatchedMagicalEnv(None, {}, max(s[2] for s in specs), preproc=preproc, num_envs=len(specs), device=device,
                             seed=seed, obs_keys=obs_keys, env_ids=specs)
