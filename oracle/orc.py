"""ctypes front-end of the CPU ORACLE (oracle/*.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, ``__graft_entry__.smoke()`` and bench.py's ``cpu_baseline`` /
``--impl reference`` legs may import this module; the product package
(magical-1_b200/) never does.  PARITY UNPINNED -- see oracle/orc.h.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, '_build', 'liborc.so')

MAX_BLOCKS = 10
MAX_GOALS = 3
RES = 384
LORES = 96

SQUARE, PENTAGON, STAR, CIRCLE = range(4)
RED, GREEN, BLUE, YELLOW = range(4)
(TASK_MOVE_TO_CORNER, TASK_MOVE_TO_REGION, TASK_MATCH_REGIONS, TASK_MAKE_LINE,
 TASK_FIND_DUPE, TASK_FIX_COLOUR, TASK_CLUSTER_COLOUR, TASK_CLUSTER_SHAPE,
 TASK_PICK_AND_PLACE) = range(9)
ROLE_NONE, ROLE_TARGET, ROLE_DISTRACTOR = range(3)


class Block(C.Structure):
    _fields_ = [('type', C.c_int32), ('colour', C.c_int32), ('role', C.c_int32), ('pad_', C.c_int32),
                ('x', C.c_double), ('y', C.c_double), ('angle', C.c_double)]


class Goal(C.Structure):
    _fields_ = [('colour', C.c_int32), ('want_block', C.c_int32),
                ('cx', C.c_double), ('cy', C.c_double), ('h', C.c_double), ('w', C.c_double)]


class Scene(C.Structure):
    """Mirror of mg_scene (include/mg_scene.h)."""
    _fields_ = [('task', C.c_int32), ('horizon', C.c_int32), ('n_blocks', C.c_int32), ('n_goals', C.c_int32),
                ('robot_first', C.c_int32), ('debug_reward', C.c_int32), ('target_block', C.c_int32),
                ('n_valid_targets', C.c_int32),
                ('phys', C.c_double * 5), ('target_pos', C.c_double * 2),
                ('robot', (C.c_double * 3) * 6),
                ('blocks', Block * MAX_BLOCKS), ('goals', Goal * MAX_GOALS)]


def build(force=False):
    """Compile oracle/_build/liborc.so with the committed Makefile."""
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith(('.c', '.h'))]
    srcs.append(os.path.join(_HERE, '..', 'include', 'mg_scene.h'))
    if (not force and os.path.exists(_LIB_PATH)
            and all(os.path.getmtime(_LIB_PATH) >= os.path.getmtime(s) for s in srcs)):
        return _LIB_PATH
    subprocess.run(['make', '-C', _HERE, '-B', '_build/liborc.so'], check=True,
                   stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        try:
            build()
        except Exception:
            if not os.path.exists(_LIB_PATH):
                raise
        L = C.CDLL(_LIB_PATH)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.POINTER(Scene)]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_env_step.argtypes = [C.c_void_p, C.c_int]
        L.orc_score.restype = C.c_double
        L.orc_score.argtypes = [C.c_void_p]
        L.orc_debug_reward.restype = C.c_double
        L.orc_debug_reward.argtypes = [C.c_void_p]
        L.orc_num_bodies.restype = C.c_int
        L.orc_num_bodies.argtypes = [C.c_void_p]
        L.orc_get_bodies.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_render.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.orc_downsample4.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.orc_robot_default_poses.argtypes = [C.c_double, C.c_double, C.c_double, C.c_void_p]
        L.orc_bench_rollout.restype = C.c_double
        L.orc_bench_rollout.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        _lib = L
    return _lib


DEFAULT_PHYS = (3.0, 1.0, 4.0, 1.5, 0.1)  # base_env.py:53-57 defaults


def robot_default_poses(x, y, angle):
    out = np.zeros((6, 3))
    lib().orc_robot_default_poses(float(x), float(y), float(angle), out.ctypes.data)
    return out


def make_scene(task, horizon, robot_pose=None, robot_bodies=None, blocks=(), goals=(), robot_first=False,
               phys=DEFAULT_PHYS, debug_reward=False, target_block=0, n_valid_targets=0,
               target_pos=(0.0, 0.0)):
    """blocks: (type, colour, x, y, angle[, role]); goals: (colour, cx, cy, h, w[, want_block])."""
    s = Scene()
    s.task, s.horizon = task, horizon
    s.n_blocks, s.n_goals = len(blocks), len(goals)
    s.robot_first = int(robot_first)
    s.debug_reward = int(debug_reward)
    s.target_block, s.n_valid_targets = target_block, n_valid_targets
    for i in range(5):
        s.phys[i] = phys[i]
    s.target_pos[0], s.target_pos[1] = target_pos
    if robot_bodies is None:
        (x, y), a = robot_pose
        robot_bodies = robot_default_poses(x, y, a)
    for i in range(6):
        for k in range(3):
            s.robot[i][k] = float(robot_bodies[i][k])
    for i, b in enumerate(blocks):
        s.blocks[i].type, s.blocks[i].colour = int(b[0]), int(b[1])
        s.blocks[i].x, s.blocks[i].y, s.blocks[i].angle = float(b[2]), float(b[3]), float(b[4])
        s.blocks[i].role = int(b[5]) if len(b) > 5 else ROLE_NONE
    for i, g in enumerate(goals):
        s.goals[i].colour = int(g[0])
        s.goals[i].cx, s.goals[i].cy, s.goals[i].h, s.goals[i].w = map(float, g[1:5])
        s.goals[i].want_block = int(g[5]) if len(g) > 5 else -1
    return s


class Env:
    """One oracle env: BaseEnv.step / render / score_on_end_of_traj for a fixed scene."""

    def __init__(self, scene):
        self._L = lib()
        self.scene = scene
        self._h = self._L.orc_create(C.byref(scene))
        self.n_bodies = self._L.orc_num_bodies(self._h)
        self.steps = 0

    def close(self):
        if self._h:
            self._L.orc_destroy(self._h)
            self._h = None

    __del__ = close

    def step(self, action):
        a = int(action)
        if not 0 <= a < 18:
            raise KeyError(a)
        self._L.orc_env_step(self._h, a)
        self.steps += 1

    def bodies(self):
        out = np.zeros((self.n_bodies, 6))
        self._L.orc_get_bodies(self._h, out.ctypes.data)
        return out

    def score(self):
        return self._L.orc_score(self._h)

    def debug_reward(self):
        return self._L.orc_debug_reward(self._h)

    def render(self, view):
        out = np.empty((RES, RES, 3), np.uint8)
        self._L.orc_render(self._h, int(view), out.ctypes.data)
        return out


def downsample4(img):
    img = np.ascontiguousarray(img, np.uint8)
    h, w, c = img.shape
    out = np.empty((h // 4, w // 4, c), np.uint8)
    lib().orc_downsample4(img.ctypes.data, h, w, c, out.ctypes.data)
    return out
