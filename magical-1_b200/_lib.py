"""ctypes binding of libmagical_b200.so (C ABI: include/mg_b200.h).

The library is built in-tree by ``csrc/build.sh`` (``__graft_entry__.build()``).  There is
no Python or CPU implementation of the step path behind this module: if the shared object is
missing, or no CUDA device is usable, the calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('MG_LIB_PATH') or os.path.join(_HERE, 'libmagical_b200.so')   # MG_LIB_PATH: dev builds (tools/time_variants.py)

MAX_BLOCKS = 10
MAX_GOALS = 3
MAX_GROUPS = 8
MAX_BODIES = 16
RES = 384
LORES = 96
NUM_ACTIONS = 18
VIEW_ALLO, VIEW_EGO = 0, 1
ENT_ROBOT, ENT_BLOCK, ENT_GOAL = 0, 1, 2

# include/mg_scene.h enums
SQUARE, PENTAGON, STAR, CIRCLE = range(4)          # entities.py:564-570 (SHAPE_TYPES order)
RED, GREEN, BLUE, YELLOW = range(4)                # entities.py:571-577 (SHAPE_COLOURS order)
(TASK_MOVE_TO_CORNER, TASK_MOVE_TO_REGION, TASK_MATCH_REGIONS, TASK_MAKE_LINE, TASK_FIND_DUPE,
 TASK_FIX_COLOUR, TASK_CLUSTER_COLOUR, TASK_CLUSTER_SHAPE, TASK_PICK_AND_PLACE) = range(9)
ROLE_NONE, ROLE_TARGET, ROLE_DISTRACTOR = range(3)
RB_ROBOT, RB_CONTROL, RB_PUPIL_L, RB_PUPIL_R, RB_FINGER_L, RB_FINGER_R = range(6)


class Block(C.Structure):
    _fields_ = [('type', C.c_int32), ('colour', C.c_int32), ('role', C.c_int32), ('pad_', C.c_int32),
                ('x', C.c_double), ('y', C.c_double), ('angle', C.c_double)]


class Goal(C.Structure):
    _fields_ = [('colour', C.c_int32), ('want_block', C.c_int32),
                ('cx', C.c_double), ('cy', C.c_double), ('h', C.c_double), ('w', C.c_double)]


class Scene(C.Structure):
    """Mirror of ``mg_scene`` (include/mg_scene.h): what BaseEnv.reset() leaves behind."""
    _fields_ = [('task', C.c_int32), ('horizon', C.c_int32), ('n_blocks', C.c_int32), ('n_goals', C.c_int32),
                ('robot_first', C.c_int32), ('debug_reward', C.c_int32), ('target_block', C.c_int32),
                ('n_valid_targets', C.c_int32),
                ('phys', C.c_double * 5), ('target_pos', C.c_double * 2),
                ('robot', (C.c_double * 3) * 6),
                ('blocks', Block * MAX_BLOCKS), ('goals', Goal * MAX_GOALS)]


class Config(C.Structure):
    _fields_ = [('device', C.c_int32), ('num_envs', C.c_int32), ('n_groups', C.c_int32),
                ('group_view', C.c_int32 * MAX_GROUPS), ('group_age', C.c_int32 * MAX_GROUPS),
                ('chw', C.c_int32), ('reserved', C.c_int32 * 5)]


class ResetSpec(C.Structure):
    """mg_reset_spec: task, horizon, rand_dynamics and the task's boolean kwargs in header order."""
    _fields_ = [('task', C.c_int32), ('horizon', C.c_int32), ('rand_dynamics', C.c_int32),
                ('kw', C.c_int32 * 8), ('reserved', C.c_int32)]


ERR_PLACEMENT = -4
NOALLOC_MAX_ENVS = 32                              # mg_get_state_noalloc: envs per call
STATUS_ARB_OVERFLOW, STATUS_NONFINITE = 1, 2      # mg_get_status bits

# every symbol include/mg_b200.h declares: name -> (restype, argtypes)
_VP = C.c_void_p
SYMBOLS = {
    'mg_create': (C.c_int, [C.POINTER(Config), C.POINTER(_VP)]),
    'mg_destroy': (C.c_int, [_VP]),
    'mg_last_error': (C.c_char_p, [_VP]),
    'mg_set_scenes': (C.c_int, [_VP, _VP, C.c_int, _VP, _VP]),
    'mg_queue_resets': (C.c_int, [_VP, _VP, C.c_int, _VP, _VP]),
    'mg_reset_obs': (C.c_int, [_VP, _VP, _VP]),
    'mg_step': (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    'mg_step_physics': (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP]),
    'mg_step_render': (C.c_int, [_VP, _VP, _VP]),
    'mg_step_host': (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP]),
    'mg_set_host_assembly': (C.c_int, [_VP, C.c_int, C.c_int, C.c_int]),
    'mg_get_host_assembly': (C.c_int, [_VP, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                       C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    'mg_host_stack_frames': (C.c_int, [_VP, _VP, C.c_int, C.c_int, _VP, C.c_int, _VP, C.c_int]),
    'mg_render': (C.c_int, [_VP, C.c_int, C.c_int, _VP, _VP]),
    'mg_score': (C.c_int, [_VP, _VP, _VP]),
    'mg_debug_reward': (C.c_int, [_VP, _VP, _VP]),
    'mg_get_bodies': (C.c_int, [_VP, _VP]),
    'mg_get_status': (C.c_int, [_VP, _VP, C.c_int]),
    'mg_state_size': (C.c_size_t, [_VP]),
    'mg_get_state': (C.c_int, [_VP, _VP, C.c_int, _VP]),
    'mg_set_state': (C.c_int, [_VP, _VP, C.c_int, _VP]),
    'mg_get_state_noalloc': (C.c_int, [_VP, _VP, C.c_int, _VP]),
    'mg_launch_count': (C.c_int64, [_VP]),
    'mg_set_timing': (C.c_int, [_VP, C.c_int]),
    'mg_kernel_ms': (C.c_double, [_VP, C.c_int, C.c_int, C.POINTER(C.c_int64)]),
    'mg_frames_downsample': (C.c_int, [C.c_int, _VP, C.c_int64, _VP, _VP]),
    'mg_frames_stack': (C.c_int, [C.c_int, _VP, _VP, C.c_int64, C.c_int, _VP, _VP, C.c_int, _VP, _VP]),
    'mg_host_palette': (C.c_int, [_VP]),
    'mg_host_entity_overlaps': (C.c_int, [C.POINTER(Scene), C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint32]),
    'mg_host_sample_scene': (C.c_int, [C.POINTER(ResetSpec), _VP, _VP, C.POINTER(Scene)]),
    'mg_host_sample_scenes': (C.c_int, [_VP, C.c_int, _VP, _VP, _VP, _VP, C.c_int]),
}

_lib = None


def lib():
    """Load the CUDA library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f'{LIB_PATH} is missing: build it with magical-1_b200/csrc/build.sh '
                '(__graft_entry__.build()); this package has no CPU fallback')
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class MgError(RuntimeError):
    pass


def check(handle, rc):
    if rc != 0:
        msg = lib().mg_last_error(handle)
        raise MgError(f'libmagical_b200 error {rc}: {msg.decode() if msg else "?"}')


def scenes_array(scenes):
    """Pack a list of Scene into one contiguous ctypes array."""
    arr = (Scene * len(scenes))()
    for i, s in enumerate(scenes):
        C.memmove(C.byref(arr, i * C.sizeof(Scene)), C.byref(s), C.sizeof(Scene))
    return arr


def copy_scene(s):
    out = Scene()
    C.memmove(C.byref(out), C.byref(s), C.sizeof(Scene))
    return out


def ptr(x):
    """Device/host pointer of a torch tensor, numpy array or ctypes object (None -> NULL)."""
    if x is None:
        return None
    if hasattr(x, 'data_ptr'):
        return x.data_ptr()
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    return C.addressof(x)
