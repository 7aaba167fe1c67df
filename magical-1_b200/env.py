"""Batched MAGICAL env on one B200 and its single-env gym-style view.

Mirrors the reference's env protocol for the hot path:
  BaseEnv.reset / step / render / seed / close      magical/base_env.py:134-141, 190-356
  preprocessor wrapper stacks (LoRes*)               magical/benchmarks/__init__.py:51-306
All simulation, rendering, resizing and frame stacking happens in libmagical_b200.so; this
module samples resets on the host (tasks.py), owns the torch tensors that receive the results
and keeps the per-env episode counters.  torch is used for device memory and streams only.
"""
import collections
import os
import queue
import threading
import ctypes as C

import numpy as np

from . import _lib as L
from . import tasks

# preprocessor -> list of (view, age) channel groups of the stacked tensor, oldest first
# (FlattenFrameStack: allo block before ego block, benchmarks/__init__.py:122-129, 269-306)
_E, _A = L.VIEW_EGO, L.VIEW_ALLO
PREPROC_SPECS = collections.OrderedDict([
    ('LoRes3EA', dict(groups=[(_A, 0), (_E, 2), (_E, 1), (_E, 0)], kind='flat', chw=False)),
    ('LoRes4E', dict(groups=[(_E, 3), (_E, 2), (_E, 1), (_E, 0)], kind='flat', chw=False)),
    ('LoRes4A', dict(groups=[(_A, 3), (_A, 2), (_A, 1), (_A, 0)], kind='flat', chw=False)),
    ('LoResStack', dict(groups=[(_A, 3), (_A, 2), (_A, 1), (_A, 0), (_E, 3), (_E, 2), (_E, 1), (_E, 0)],
                        kind='stack', chw=False)),
    ('LoResCHW4E', dict(groups=[(_E, 3), (_E, 2), (_E, 1), (_E, 0)], kind='flat', chw=True)),
    # registered with allo_frames=0, ego_frames=4 in the reference too (benchmarks/__init__.py:301-306)
    ('LoResCHW4A', dict(groups=[(_E, 3), (_E, 2), (_E, 1), (_E, 0)], kind='flat', chw=True)),
])


def _torch():
    import torch
    return torch


class _SceneView:
    """List-like view of the scene every env was last reset to.  The scenes live as raw mg_scene bytes
    ([B, sizeof(mg_scene)] uint8) so that thousands of resets per step cost one vectorised copy; an item is
    materialised as a ctypes Scene copy on access."""

    def __init__(self, raw):
        self._raw = raw

    def __len__(self):
        return self._raw.shape[0]

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[k] for k in range(*i.indices(len(self)))]
        return L.Scene.from_buffer_copy(self._raw[int(i)].tobytes())

    def __iter__(self):
        return (self[i] for i in range(len(self)))


class _ResetFeeder:
    """Samples, on a host thread, the initial states of the episodes that will start on FUTURE steps while the GPU
    runs the current ones.  Episode ends are known in advance (an env finishes when its step counter reaches its
    horizon), so the thread walks the schedule ahead of the stepping loop: for every coming step with finishing envs
    it draws those envs' next scenes from their own RandomState streams (native reset engine, reference draw order)
    and queues them.  The streams therefore run ahead of the episodes; `cancel()` rolls every unconsumed draw back, so
    whoever looks at the streams (reset(), seed(), rngs, set_state ...) sees them where a synchronous env would."""

    CHUNK_ENVS = 512      # envs sampled per call of the native engine (it starts its threads per call)

    def __init__(self, env, depth=16):
        self.env = env
        self.q = queue.Queue(maxsize=depth)
        self.lock = threading.Lock()
        self.items = []           # produced, not yet consumed: (step, ids, scenes_bytes, error, keys_before, pos_before)
        self.stop = False
        h = env._horizons
        self.next_finish = env._abs_step + np.maximum(h - env._episode_steps, 1)
        self.max_span = max(1, int(h.min()) - 1)     # an env appears at most once per chunk
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def _run(self):
        env, H = self.env, self.env._horizons
        nscene = C.sizeof(L.Scene)
        while not self.stop:
            # the next few steps that have finishing envs, up to CHUNK_ENVS envs, in one call of the engine
            s0 = int(self.next_finish.min())
            steps, ids_per_step, total = [], [], 0
            s = s0
            while s - s0 < self.max_span and total < self.CHUNK_ENVS and len(steps) < 8:
                ids = np.nonzero(self.next_finish == s)[0]
                if len(ids):
                    steps.append(s); ids_per_step.append(ids); total += len(ids)
                s += 1
            ids_all = np.concatenate(ids_per_step)
            keys_before, pos_before = env._mt_keys[ids_all].copy(), env._mt_pos[ids_all].copy()
            err, raw = None, None
            try:
                scenes = env._sample_many(ids_all, packed=True)
                raw = np.frombuffer(scenes, np.uint8).reshape(len(ids_all), nscene)
            except Exception as exc:       # noqa: BLE001  (PlacementError etc.: re-raised by the consumer)
                err = exc
            self.next_finish[ids_all] += H[ids_all]
            off = 0
            for s, ids in zip(steps, ids_per_step):
                n = len(ids)
                item = (s, ids, None if raw is None else raw[off:off + n], err, keys_before[off:off + n], pos_before[off:off + n])
                off += n
                with self.lock:
                    self.items.append(item)
                while not self.stop:
                    try:
                        self.q.put(item, timeout=0.05)
                        break
                    except queue.Full:
                        continue

    def get(self, step, finished):
        """Scenes ([n, sizeof(mg_scene)] uint8) of the envs finishing on `step`; None if the schedule does not match
        (the caller then cancels the feeder and samples synchronously)."""
        item = self.q.get()
        with self.lock:
            self.items = [it for it in self.items if it is not item]
        s, ids, raw, err, kb, pb = item
        if s != step or len(ids) != len(finished) or (ids != finished).any():
            with self.lock:
                self.items.insert(0, item)       # not consumed: roll it back with the rest
            return None
        if err is not None:
            raise err
        return raw

    def cancel(self):
        """Stop the thread and put every stream back to where the consumed episodes left it."""
        self.stop = True
        while self.thread.is_alive():
            try:
                self.q.get_nowait()
            except queue.Empty:
                pass
            self.thread.join(timeout=0.01)
        env = self.env
        for s, ids, raw, err, kb, pb in reversed(self.items):
            env._mt_keys[ids], env._mt_pos[ids] = kb, pb
        self.items = []


class BatchedMagicalEnv:
    """`num_envs` independent copies of one registered env id, stepped together on one GPU.

    reset() -> obs dict of uint8 torch tensors on the device
    step(actions[int, B]) -> (obs, reward f32[B], done bool[B], {'eval_score': f32[B]})
    Envs that reach their horizon are reset on the host (their own RandomState, reference draw
    order) before the step's render, so the returned observation of a finished env is the first
    observation of its next episode; reward/done/eval_score still belong to the finished one.
    """

    def __init__(self, entry_point, env_kwargs, max_episode_steps, preproc=None, num_envs=1, device=0,
                 seed=None, obs_keys=None, auto_reset=True, env_ids=None):
        torch = _torch()
        if not torch.cuda.is_available():
            raise RuntimeError('magical_b200 needs a CUDA device: there is no CPU path')
        if isinstance(device, str):
            device = torch.device(device).index or 0
        self.device_index = int(device)
        self.device = torch.device('cuda', self.device_index)
        self.num_envs = int(num_envs)
        self.preproc = preproc
        self.max_episode_steps = int(max_episode_steps)
        self.auto_reset = auto_reset
        # resets that raise PlacementError in the reference (geom.py:256-259) raise here too; > 0 re-samples such
        # envs that many times on their own stream first (what calling reset() again does)
        self.placement_retries = 0
        # one (entry point, kwargs, horizon) per env allows mixed batches (BASELINE config 5)
        if env_ids is None:
            self._specs = [(entry_point, dict(env_kwargs), self.max_episode_steps)] * self.num_envs
        else:
            self._specs = list(env_ids)
            assert len(self._specs) == self.num_envs
        self._horizons = np.asarray([s[2] for s in self._specs], np.int64)
        self._reset_spec_cache = {}
        self._py_reset = os.environ.get('MG_PY_RESET', '0') == '1'
        self._reset_threads = max(1, len(os.sched_getaffinity(0)) if hasattr(os, 'sched_getaffinity') else (os.cpu_count() or 1))
        spec = PREPROC_SPECS[preproc] if preproc is not None else dict(groups=[], kind='none', chw=False)
        self._pspec = spec
        groups = spec['groups']
        self.n_channels = 3 * len(groups)
        if obs_keys is None:
            obs_keys = self.default_obs_keys(preproc)
        self.obs_keys = tuple(obs_keys)

        cfg = L.Config()
        cfg.device, cfg.num_envs, cfg.n_groups, cfg.chw = self.device_index, self.num_envs, len(groups), int(spec['chw'])
        for g, (view, age) in enumerate(groups):
            cfg.group_view[g], cfg.group_age[g] = view, age
        self._lib = L.lib()
        h = C.c_void_p()
        rc = self._lib.mg_create(C.byref(cfg), C.byref(h))
        if rc != 0:
            raise L.MgError(f'mg_create failed ({rc}): {self._lib.mg_last_error(None).decode()}')
        self._h = h
        B = self.num_envs
        with torch.cuda.device(self.device):
            if groups:
                shape = (B, self.n_channels, L.LORES, L.LORES) if spec['chw'] else (B, L.LORES, L.LORES, self.n_channels)
                self._obs = torch.empty(shape, dtype=torch.uint8, device=self.device)
            else:
                self._obs = None
            self._reward = torch.zeros(B, dtype=torch.float32, device=self.device)
            self._score = torch.zeros(B, dtype=torch.float32, device=self.device)
            self._done = torch.zeros(B, dtype=torch.uint8, device=self.device)
            self._actions = torch.zeros(B, dtype=torch.int32, device=self.device)
        self._extra = {}
        self._res = [None, None]
        self._feeder = None
        # auto-reset samples the coming episodes on a host thread while the GPU steps (False: synchronously, on the
        # step that needs them; same scenes either way)
        self.async_reset = not self._py_reset
        # optional override of where auto-reset takes its scenes from: callable(env_indices) -> [n, sizeof(mg_scene)]
        # uint8 (bench.py's fixed-count scene pool); None: the envs' own RandomState streams
        self.reset_source = None
        self._abs_step = 0
        self._spec_bytes = None
        self.seed(seed)
        self._episode_steps = np.zeros(B, np.int64)
        self._scene_raw = np.zeros((B, C.sizeof(L.Scene)), np.uint8)
        self._needs_reset = True

    @staticmethod
    def default_obs_keys(preproc):
        if preproc is None:
            return ('allo', 'ego')
        if PREPROC_SPECS[preproc]['kind'] == 'stack':
            return ('allo', 'ego')
        return ('allo', 'ego', 'past_obs')

    # ------------------------------------------------------------------ plumbing
    def _stream(self):
        return _torch().cuda.current_stream(self.device).cuda_stream

    def _check(self, rc):
        L.check(self._h, rc)

    @property
    def scenes(self):
        """The scene (initial state) every env was last reset to: list-like of ctypes Scene copies."""
        return _SceneView(self._scene_raw)

    def _cancel_feeder(self):
        if self._feeder is not None:
            self._feeder.cancel()
            self._feeder = None

    def close(self):
        self._cancel_feeder()
        if getattr(self, '_h', None) is not None and self._h:
            self._lib.mg_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ reset
    def seed(self, seed=None):
        """BaseEnv.seed (base_env.py:134-141); env i gets RandomState(seed + i)."""
        self._cancel_feeder()
        if seed is None:
            seed = np.random.randint(0, (1 << 31) - 1)
        self._seed = int(seed)
        # the streams live as raw MT19937 states (RandomState.get_state()[1:3]) so that the native reset engine
        # can advance thousands of them without a Python object per env; `rngs` rebuilds RandomStates on demand
        B = self.num_envs
        self._mt_keys = np.empty((B, 624), np.uint32)
        self._mt_pos = np.empty(B, np.int32)
        for i in range(B):
            st = np.random.RandomState((self._seed + i) % (1 << 32)).get_state()
            self._mt_keys[i], self._mt_pos[i] = st[1], st[2]
        return [self._seed]

    @property
    def rngs(self):
        """np.random.RandomState per env, positioned where the env's stream currently is (copies)."""
        self._cancel_feeder()      # rolls back the draws of episodes that have not started yet
        out = []
        for i in range(self.num_envs):
            r = np.random.RandomState(0)
            r.set_state(('MT19937', self._mt_keys[i], int(self._mt_pos[i]), 0, 0.0))
            out.append(r)
        return out

    def _task_kwargs(self, i):
        ep, kw, horizon = self._specs[i]
        kw = dict(kw)
        rand_dynamics = kw.pop('rand_dynamics', False)
        for k in ('res_hw', 'fps', 'phys_steps', 'phys_iter', 'max_episode_steps'):
            kw.pop(k, None)
        return ep, horizon, rand_dynamics, kw

    def _reset_spec(self, i):
        key = id(self._specs[i])
        if key not in self._reset_spec_cache:
            ep, horizon, rand_dynamics, kw = self._task_kwargs(i)
            self._reset_spec_cache[key] = tasks.reset_spec(ep, horizon, rand_dynamics, **kw)
        return self._reset_spec_cache[key]

    def _sample(self, i):
        """One env through the Python specification (tasks.py), on the env's own stream."""
        ep, horizon, rand_dynamics, kw = self._task_kwargs(i)
        r = np.random.RandomState(0)
        r.set_state(('MT19937', self._mt_keys[i], int(self._mt_pos[i]), 0, 0.0))
        try:
            return tasks.sample_scene(ep, r, horizon, rand_dynamics=rand_dynamics, **kw)
        finally:
            st = r.get_state()
            self._mt_keys[i], self._mt_pos[i] = st[1], st[2]

    def _sample_many(self, env_ids, packed=False):
        """BaseEnv.reset()'s sampling for the given envs: the native engine (csrc/mg_reset.cpp) on all host
        cores, each env on its own RandomState stream; MG_PY_RESET=1 runs the Python specification instead.
        Unlike a sequential loop, a PlacementError is raised after every env's stream has advanced."""
        env_ids = np.asarray(env_ids, np.int64)
        if self._py_reset:
            out = [self._sample(int(i)) for i in env_ids]
            return L.scenes_array(out) if packed else out
        n = len(env_ids)
        if self._spec_bytes is None:
            sb = np.zeros((self.num_envs, C.sizeof(L.ResetSpec)), np.uint8)
            for i in range(self.num_envs):
                sb[i] = np.frombuffer(self._reset_spec(i), np.uint8)
            self._spec_bytes = sb
        specs_np = np.ascontiguousarray(self._spec_bytes[env_ids])
        specs = (L.ResetSpec * n).from_buffer(specs_np)
        keys = np.ascontiguousarray(self._mt_keys[env_ids])
        pos = np.ascontiguousarray(self._mt_pos[env_ids])
        scenes = (L.Scene * n)()
        status = np.zeros(n, np.int32)
        rc = self._lib.mg_host_sample_scenes(C.addressof(specs), n, keys.ctypes.data, pos.ctypes.data,
                                             C.addressof(scenes), status.ctypes.data, max(1, min(self._reset_threads, n // 8)))
        if rc != 0:
            raise L.MgError(f'mg_host_sample_scenes failed ({rc})')
        for _ in range(self.placement_retries):
            # what a caller of the reference does after a PlacementError: reset() again, the env's stream going on
            bad = np.nonzero(status == L.ERR_PLACEMENT)[0]
            if not len(bad):
                break
            for k in bad:
                one = np.zeros(1, np.int32)
                self._lib.mg_host_sample_scenes(C.addressof(specs) + int(k) * C.sizeof(L.ResetSpec), 1, keys[k:k + 1].ctypes.data,
                                                pos[k:k + 1].ctypes.data, C.addressof(scenes) + int(k) * C.sizeof(L.Scene),
                                                one.ctypes.data, 1)
                status[k] = one[0]
        self._mt_keys[env_ids], self._mt_pos[env_ids] = keys, pos
        bad = np.nonzero(status)[0]
        if len(bad):
            if (status[bad] == L.ERR_PLACEMENT).all():
                raise tasks.PlacementError(f'Could not place an entity in env(s) {env_ids[bad].tolist()[:8]}')
            raise L.MgError(f'native reset failed for envs {env_ids[bad].tolist()[:8]}: {status[bad].tolist()[:8]}')
        return scenes if packed else [scenes[k] for k in range(n)]

    def set_scenes(self, env_ids, scenes):
        """Install explicit initial states (used by reset and by the parity tests)."""
        self.set_scenes_packed(env_ids, L.scenes_array(list(scenes)), len(scenes))

    def set_scenes_packed(self, env_ids, packed, n):
        """Same as set_scenes for a pre-packed ctypes array of Scene (reset pools)."""
        self._cancel_feeder()
        env_ids = np.ascontiguousarray(env_ids, np.int32)
        self._check(self._lib.mg_set_scenes(self._h, env_ids.ctypes.data, int(n), C.addressof(packed), self._stream()))
        self._scene_raw[env_ids] = np.frombuffer(packed, np.uint8)[:int(n) * C.sizeof(L.Scene)].reshape(int(n), -1)
        self._episode_steps[env_ids] = 0

    def reset(self, scenes=None):
        self._cancel_feeder()
        ids = np.arange(self.num_envs)
        if scenes is None:
            self.set_scenes_packed(ids, self._sample_many(ids, packed=True), self.num_envs)
        else:
            self.set_scenes(ids, scenes)
        self._check(self._lib.mg_reset_obs(self._h, L.ptr(self._obs), self._stream()))
        self._needs_reset = False
        return self._make_obs()

    # ------------------------------------------------------------------ step
    def _next_scenes(self, finished):
        """[n, sizeof(mg_scene)] uint8: the next initial states of the envs finishing on the coming step, from the
        feeder thread (sampled while earlier steps ran) or, failing that, sampled now."""
        if self.reset_source is not None:
            return np.asarray(self.reset_source(finished), np.uint8).reshape(len(finished), -1)
        if self.async_reset:
            if self._feeder is None:
                self._feeder = _ResetFeeder(self)
            raw = self._feeder.get(self._abs_step + 1, finished)
            if raw is not None:
                return raw
            self._cancel_feeder()          # the schedule was changed under the feeder: start over
        packed = self._sample_many(finished, packed=True)
        return np.frombuffer(packed, np.uint8).reshape(len(finished), -1)

    def _queue_resets(self, extra=None):
        """Auto-reset: hand the next episodes of the envs that finish on the coming step (plus `extra`, envs the caller
        wants restarted on that step) to mg_queue_resets: they are installed between that step's physics and its
        render.  Returns the env indices."""
        finished = np.nonzero(self._episode_steps + 1 >= self._horizons)[0]
        if extra is not None and len(extra):
            finished = np.union1d(finished, np.asarray(extra, np.int64))
        if len(finished):
            raw = np.ascontiguousarray(self._next_scenes(finished))
            ids32 = finished.astype(np.int32)
            self._check(self._lib.mg_queue_resets(self._h, ids32.ctypes.data, len(finished), raw.ctypes.data, self._stream()))
            self._scene_raw[finished] = raw
        return finished

    def step_into(self, actions, reward, done, score, restart=None):
        """step() without the conveniences: `actions` is an int32 CUDA tensor, results go into the caller's tensors
        (f32 / u8 / f32, CUDA) and the observation into the env's own buffer (returned by `observation()`); nothing
        is cloned or converted.  Used by roll-out loops that own their buffers (bench.py).  `restart`: env indices to
        restart on this step although they have not reached their horizon (auto_reset only)."""
        if self._needs_reset:
            raise RuntimeError('call reset() before step()')
        st = self._stream()
        if self._obs is not None:
            finished = self._queue_resets(restart) if self.auto_reset else ()
            if not self.auto_reset:
                self._cancel_feeder()
            self._check(self._lib.mg_step(self._h, L.ptr(actions), L.ptr(self._obs), L.ptr(reward), L.ptr(done), L.ptr(score), st))
            self._episode_steps += 1
            self._abs_step += 1
            if len(finished):
                self._episode_steps[finished] = 0
            return
        # no preprocessor (full-resolution observations are rendered on demand): physics, resets, ring-less render
        self._check(self._lib.mg_step_physics(self._h, L.ptr(actions), L.ptr(reward), L.ptr(done), L.ptr(score), st))
        self._episode_steps += 1
        self._abs_step += 1
        finished = np.nonzero(self._episode_steps >= self._horizons)[0]
        if len(finished) and self.auto_reset:
            self.set_scenes_packed(finished, self._sample_many(finished, packed=True), len(finished))
        self._check(self._lib.mg_step_render(self._h, L.ptr(self._obs), st))

    def observation(self):
        return self._make_obs()

    def step(self, actions):
        torch = _torch()
        if self._needs_reset:
            raise RuntimeError('call reset() before step()')
        if not (isinstance(actions, torch.Tensor) and actions.is_cuda and actions.dtype == torch.int32):
            a = np.asarray(actions).astype(np.int64).reshape(-1)
            if a.shape[0] != self.num_envs or ((a < 0) | (a >= L.NUM_ACTIONS)).any():
                raise KeyError(f'actions must be {self.num_envs} ints in [0, {L.NUM_ACTIONS})')
            self._actions.copy_(torch.from_numpy(a.astype(np.int32)), non_blocking=False)
            actions = self._actions
        # results go into one of two alternating sets of buffers, so what a call returns stays valid until the call
        # after the next one (no per-step clones; `done` is the u8 buffer viewed as bool)
        k = self._abs_step & 1
        if self._res[k] is None:
            self._res[k] = (torch.zeros_like(self._reward), torch.zeros_like(self._done), torch.zeros_like(self._score))
        reward, done, score = self._res[k]
        self.step_into(actions, reward, done, score)
        return self._make_obs(), reward, done.view(torch.bool), {'eval_score': score}

    def step_host(self, actions_host, obs_host, reward_host, done_host, score_host):
        """mg_step_host: one env-step with HOST buffers (numpy / pinned torch CPU tensors); the H2D copy of
        the actions and the D2H copies of obs / reward / done / eval_score are inside the call.  With auto_reset the
        envs that finish on this step start their next episode inside it, as in step()."""
        finished = ()
        if self.auto_reset and self._obs is not None:
            finished = self._queue_resets()
        else:
            self._cancel_feeder()
        self._check(self._lib.mg_step_host(self._h, L.ptr(actions_host), L.ptr(obs_host), L.ptr(reward_host),
                                           L.ptr(done_host), L.ptr(score_host)))
        self._episode_steps += 1
        self._abs_step += 1
        if len(finished):
            self._episode_steps[finished] = 0

    def set_host_assembly(self, mode='auto', chunks=0, threads=0):
        """mg_set_host_assembly: let step_host send only the NEW frame of part of the batch over PCIe and build those
        envs' stacked observations with host threads (same bytes; single-view HWC stacks such as LoRes4E only).
        mode: 'off', 'auto' (the handle balances link and host) or 'fixed' (`chunks` of the 16 parts)."""
        m = {'off': 0, 'auto': 1, 'fixed': 2}[mode]
        self._check(self._lib.mg_set_host_assembly(self._h, m, int(chunks), int(threads)))

    def host_assembly(self):
        """(host-assembled parts, parts, host threads, last step's copy ms, last step's host ms)"""
        a, b, t = C.c_int(0), C.c_int(0), C.c_int(0)
        cm, hm = C.c_double(0), C.c_double(0)
        self._lib.mg_get_host_assembly(self._h, C.byref(a), C.byref(b), C.byref(t), C.byref(cm), C.byref(hm))
        return a.value, b.value, t.value, cm.value, hm.value

    def set_timing(self, enable):
        self._check(self._lib.mg_set_timing(self._h, int(enable)))

    def kernel_ms(self, which, reset=True):
        """(accumulated CUDA-event ms, launches) of the physics (0) / render (1) kernels."""
        n = C.c_int64(0)
        ms = self._lib.mg_kernel_ms(self._h, int(which), int(reset), C.byref(n))
        return float(ms), int(n.value)

    # ------------------------------------------------------------------ observations
    def render_view(self, view, full_res):
        torch = _torch()
        n = L.RES if full_res else L.LORES
        key = (view, bool(full_res))
        if key not in self._extra:
            self._extra[key] = torch.empty((self.num_envs, n, n, 3), dtype=torch.uint8, device=self.device)
        out = self._extra[key]
        self._check(self._lib.mg_render(self._h, int(view), int(bool(full_res)), L.ptr(out), self._stream()))
        return out

    def _make_obs(self):
        """Observation dict in the reference's key order (allo, ego[, past_obs])."""
        spec = self._pspec
        obs = collections.OrderedDict()
        if spec['kind'] == 'none':
            for key, view in (('allo', _A), ('ego', _E)):
                if key in self.obs_keys:
                    obs[key] = self.render_view(view, True)
            return obs
        if spec['kind'] == 'stack':
            # EagerDictFrameStack: every key becomes its own 4-frame stack (benchmarks/__init__.py:51-82)
            if 'allo' in self.obs_keys:
                obs['allo'] = self._obs[..., :12]
            if 'ego' in self.obs_keys:
                obs['ego'] = self._obs[..., 12:]
            return obs
        for key, view in (('allo', _A), ('ego', _E)):
            if key in self.obs_keys:
                img = self.render_view(view, False)
                obs[key] = img.permute(0, 3, 1, 2) if spec['chw'] else img
        if 'past_obs' in self.obs_keys:
            obs['past_obs'] = self._obs
        return obs

    def render(self, mode='rgb_array'):
        if mode != 'rgb_array':
            raise NotImplementedError("only mode='rgb_array' is available (no window system on this box)")
        return collections.OrderedDict(allo=self.render_view(_A, True), ego=self.render_view(_E, True))

    # ------------------------------------------------------------------ state access
    def get_bodies(self):
        """f64 [B, 16, 6]: x, y, angle, vx, vy, w of the six robot bodies then the blocks."""
        out = np.zeros((self.num_envs, L.MAX_BODIES, 6))
        self._check(self._lib.mg_get_bodies(self._h, out.ctypes.data))
        return out

    def score(self):
        torch = _torch()
        out = torch.empty(self.num_envs, dtype=torch.float32, device=self.device)
        self._check(self._lib.mg_score(self._h, L.ptr(out), self._stream()))
        return out

    def debug_reward(self):
        torch = _torch()
        out = torch.empty(self.num_envs, dtype=torch.float32, device=self.device)
        self._check(self._lib.mg_debug_reward(self._h, L.ptr(out), self._stream()))
        return out

    def status(self, clear=False):
        """int32 [B] of sticky health flags (_lib.STATUS_ARB_OVERFLOW: a contact was dropped because the per-env
        arbiter cache was full; _lib.STATUS_NONFINITE: a body state became NaN / Inf).  Synchronous."""
        out = np.zeros(self.num_envs, np.int32)
        self._check(self._lib.mg_get_status(self._h, out.ctypes.data, int(bool(clear))))
        return out

    def get_state(self, env_ids=None):
        ids = np.arange(self.num_envs, dtype=np.int32) if env_ids is None else np.ascontiguousarray(env_ids, np.int32)
        _torch().cuda.synchronize(self.device)
        sz = self._lib.mg_state_size(self._h)
        buf = np.zeros((len(ids), sz), np.uint8)
        self._check(self._lib.mg_get_state(self._h, ids.ctypes.data, len(ids), buf.ctypes.data))
        return dict(blobs=buf, episode_steps=self._episode_steps[ids].copy())

    def set_state(self, state, env_ids=None):
        ids = np.arange(self.num_envs, dtype=np.int32) if env_ids is None else np.ascontiguousarray(env_ids, np.int32)
        self._cancel_feeder()
        buf = np.ascontiguousarray(state['blobs'], np.uint8)
        self._check(self._lib.mg_set_state(self._h, ids.ctypes.data, len(ids), buf.ctypes.data))
        self._episode_steps[ids] = state['episode_steps']
        self._needs_reset = False

    @property
    def episode_steps(self):
        return self._episode_steps.copy()

    def launch_count(self):
        return int(self._lib.mg_launch_count(self._h))
