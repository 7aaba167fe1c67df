"""Host-side reset: the nine task ``on_reset`` samplers of the reference, restated so that a
``np.random.RandomState`` is consumed draw for draw like the reference consumes ``self.rng``
and the result is one plain ``Scene`` record (include/mg_scene.h) for the CUDA side.

Reference: magical/base_env.py:190-246 (reset, phys-var sampling), magical/geom.py:116-384
(pm_randomise_pose / pm_randomise_all_poses / randomise_hw / pm_shift_bodies) and the
``on_reset`` of magical/benchmarks/{move_to_corner,move_to_region,match_regions,make_line,
find_dupe,fix_colour,cluster,pick_and_place}.py.  The overlap test of the rejection sampler
(``space.shape_query``, geom.py:226-233) is ``mg_host_entity_overlaps`` of the C ABI, which runs
the same narrow phase as the device kernels.
"""
import math
import warnings

import numpy as np

from . import _lib as L
from ._lib import (BLUE, CIRCLE, GREEN, PENTAGON, RED, SQUARE, STAR, YELLOW, Scene)

ROBOT_RAD = 0.2                        # base_env.py:62
SHAPE_RAD = ROBOT_RAD * 0.6            # base_env.py:64
ARENA_LRBT = (-1.0, 1.0, -1.0, 1.0)    # base_env.py:65
RAND_GOAL_MIN_SIZE, RAND_GOAL_MAX_SIZE = 0.5, 0.8
JITTER_PCT = 0.05
JITTER_POS_BOUND = max(ARENA_LRBT) * JITTER_PCT / 2.0
JITTER_ROT_BOUND = JITTER_PCT * np.pi
JITTER_TARGET_BOUND = JITTER_PCT * (RAND_GOAL_MAX_SIZE - RAND_GOAL_MIN_SIZE) / 2

# phys_vars.py / base_env.py:49-57: (default, (lo, hi)) in declaration order
PHYS_VARS = ((3, (2.2, 3.5)), (1, (0.7, 1.5)), (4, (2.5, 4.5)), (1.5, (1.0, 1.8)), (0.1, (0.07, 0.15)))

ALL_COLOURS = np.arange(4)   # stands in for en.SHAPE_COLOURS (red, green, blue, yellow): same rng.choice draws
ALL_TYPES = np.arange(4)     # en.SHAPE_TYPES (square, pentagon, star, circle)

ROBOT = ('robot', 0)


class PlacementError(Exception):
    """geom.PlacementError (geom.py:111-113)."""


def _rotated(x, y, ang):
    """pymunk Vec2d.rotated."""
    c, s = math.cos(ang), math.sin(ang)
    return x * c - y * s, x * s + y * c


def robot_body_poses(pos, angle):
    """Poses of the six robot bodies right after Robot.setup (entities.py:243-254, 267-277, 313-331)."""
    x, y = float(pos[0]), float(pos[1])
    angle = float(angle)
    out = [[x, y, angle], [x, y, angle], [0.0, 0.0, angle], [0.0, 0.0, angle]]
    for side in (-1, 1):
        lower, upper = (-0.0, math.pi / 8) if side < 0 else (-(math.pi / 8), 0.0)
        delta = upper if side < 0 else lower
        rx, ry = _rotated(side * ROBOT_RAD * 0.45, ROBOT_RAD * 0.1, angle)
        out.append([x + rx, y + ry, angle + delta])
    return out


class SceneBuilder:
    """The mutable stand-in for (space, entity list) during on_reset."""

    def __init__(self, task, horizon, rng, rand_dynamics):
        self.rng = rng
        s = self.scene = Scene()
        s.task, s.horizon = task, horizon
        # BaseEnv.reset: phys vars are drawn before on_reset (base_env.py:210-215)
        for i, (default, (lo, hi)) in enumerate(PHYS_VARS):
            s.phys[i] = rng.uniform(lo, hi) if rand_dynamics else default
        self.robot_added = False

    # ---- entity creation (add order is fixed by Scene: goals, [robot], blocks, [robot]) ----
    def add_robot(self, pos, angle, first=False):
        s = self.scene
        for b, pose in enumerate(robot_body_poses(pos, angle)):
            for k in range(3):
                s.robot[b][k] = pose[k]
        if first:
            assert s.n_blocks == 0
            s.robot_first = 1
        self.robot_added = True

    def add_block(self, shape_type, colour, pos, angle, role=L.ROLE_NONE):
        s = self.scene
        i = s.n_blocks
        b = s.blocks[i]
        b.type, b.colour, b.role = int(shape_type), int(colour), int(role)
        b.x, b.y, b.angle = float(pos[0]), float(pos[1]), float(angle)
        s.n_blocks = i + 1
        return ('block', i)

    def add_goal(self, x, y, h, w, colour, want_block=-1):
        """GoalRegion(x, y, h, w, colour): (x, y) is the top-left corner (entities.py:783-790)."""
        s = self.scene
        i = s.n_goals
        g = s.goals[i]
        g.colour, g.want_block = int(colour), int(want_block)
        g.h, g.w = float(h), float(w)
        g.cx, g.cy = x + w / 2, y - h / 2
        s.n_goals = i + 1
        return ('goal', i)

    # ---- pose access: main body of an entity ----
    def get_pose(self, ent):
        s = self.scene
        kind, i = ent
        if kind == 'robot':
            return s.robot[0][0], s.robot[0][1], s.robot[0][2]
        if kind == 'block':
            return s.blocks[i].x, s.blocks[i].y, s.blocks[i].angle
        return s.goals[i].cx, s.goals[i].cy, 0.0

    def shift(self, ent, position=None, angle=None):
        """geom.pm_shift_bodies (geom.py:362-384)."""
        s = self.scene
        kind, i = ent
        px, py, pa = self.get_pose(ent)
        if angle is None:
            angle = pa
        if position is None:
            position = (px, py)
        nx, ny, angle = float(position[0]), float(position[1]), float(angle)
        if kind == 'robot':
            root_x, root_y, root_a = px, py, pa
            new = []
            for b in range(6):
                bx, by, ba = s.robot[b][0], s.robot[b][1], s.robot[b][2]
                dx, dy = _rotated(bx - root_x, by - root_y, angle - root_a)
                new.append((nx + dx, ny + dy, angle + (ba - root_a)))
            for b in range(6):
                s.robot[b][0], s.robot[b][1], s.robot[b][2] = new[b]
        elif kind == 'block':
            dx, dy = _rotated(0.0, 0.0, angle - pa)
            s.blocks[i].x, s.blocks[i].y, s.blocks[i].angle = nx + dx, ny + dy, angle + (pa - pa)
        else:
            s.goals[i].cx, s.goals[i].cy = nx, ny

    def _save(self, ent):
        s = self.scene
        kind, i = ent
        if kind == 'robot':
            return [[s.robot[b][k] for k in range(3)] for b in range(6)]
        return self.get_pose(ent)

    def _restore(self, ent, saved):
        s = self.scene
        kind, i = ent
        if kind == 'robot':
            for b in range(6):
                for k in range(3):
                    s.robot[b][k] = saved[b][k]
        elif kind == 'block':
            s.blocks[i].x, s.blocks[i].y, s.blocks[i].angle = saved
        else:
            s.goals[i].cx, s.goals[i].cy = saved[0], saved[1]

    def all_entities(self):
        s = self.scene
        ents = [('goal', g) for g in range(s.n_goals)] + [('block', b) for b in range(s.n_blocks)]
        if self.robot_added:
            ents.append(ROBOT)
        return ents

    def _overlaps(self, ent, active):
        kind, i = ent
        robot_active = int(ROBOT in active and ent != ROBOT)
        bmask = gmask = 0
        for k, j in active:
            if (k, j) == ent:
                continue
            if k == 'block':
                bmask |= 1 << j
            elif k == 'goal':
                gmask |= 1 << j
        code = {'robot': L.ENT_ROBOT, 'block': L.ENT_BLOCK, 'goal': L.ENT_GOAL}[kind]
        return bool(L.lib().mg_host_entity_overlaps(self.scene, code, i, robot_active, bmask, gmask))

    def randomise_pose(self, ent, active, rand_pos=True, rand_rot=True, rel_pos_linf_limit=None,
                       rel_rot_limit=None):
        """geom.pm_randomise_pose (geom.py:116-264); `active` = entities whose shapes can be hit."""
        assert rand_pos or rand_rot
        rng = self.rng
        saved = self._save(ent)
        orig_x, orig_y, orig_angle = self.get_pose(ent)
        arena_l, arena_r, arena_b, arena_t = ARENA_LRBT
        if rel_pos_linf_limit is not None:
            assert 0 <= rel_pos_linf_limit
            pos_x_minmax = (max(arena_l, orig_x - rel_pos_linf_limit), min(arena_r, orig_x + rel_pos_linf_limit))
            pos_y_minmax = (max(arena_b, orig_y - rel_pos_linf_limit), min(arena_t, orig_y + rel_pos_linf_limit))
        else:
            pos_x_minmax = (arena_l, arena_r)
            pos_y_minmax = (arena_b, arena_t)
        if rel_rot_limit is not None:
            assert 0 <= rel_rot_limit
            rot_min, rot_max = orig_angle - rel_rot_limit, orig_angle + rel_rot_limit
        else:
            rot_min, rot_max = -np.pi, np.pi
        max_tries = 10000
        n_tries = 0
        while n_tries < max_tries:
            if rand_pos:
                new_pos = (rng.uniform(*pos_x_minmax), rng.uniform(*pos_y_minmax))
            else:
                new_pos = (orig_x, orig_y)
            new_angle = rng.uniform(rot_min, rot_max) if rand_rot else orig_angle
            self.shift(ent, position=new_pos, angle=new_angle)
            if not self._overlaps(ent, active):
                break
            n_tries += 1
        else:
            self._restore(ent, saved)
            raise PlacementError(f'Could not place {ent} after {n_tries} attempts')
        if n_tries > max_tries // 10:
            warnings.warn(f'Took {n_tries}>{max_tries // 10} samples to place shape.')
        return n_tries

    def randomise_all_poses(self, entities, rand_pos=True, rand_rot=True, rel_pos_linf_limits=None,
                            rel_rot_limits=None, ignore=(), max_retries=10):
        """geom.pm_randomise_all_poses (geom.py:281-341)."""
        n = len(entities)

        def listify(v):
            return list(v) if isinstance(v, (list, tuple)) else [v] * n
        pos_limits, rot_limits = listify(rel_pos_linf_limits), listify(rel_rot_limits)
        rand_pos, rand_rot = listify(rand_pos), listify(rand_rot)
        others = [e for e in self.all_entities() if e not in entities and e not in ignore]
        for retry in range(max_retries):
            active = list(others)
            for ent, pl, rl, rp, rr in zip(entities, pos_limits, rot_limits, rand_pos, rand_rot):
                active.append(ent)
                try:
                    self.randomise_pose(ent, active, rand_pos=rp, rand_rot=rr, rel_pos_linf_limit=pl,
                                        rel_rot_limit=rl)
                except PlacementError:
                    if retry == max_retries - 1:
                        raise
                    break
            else:
                break


def randomise_hw(min_side, max_side, rng, current_hw=None, linf_bound=None):
    """geom.randomise_hw (geom.py:344-359)."""
    minima = np.asarray((min_side, min_side))
    maxima = np.asarray((max_side, max_side))
    if linf_bound is not None:
        current_hw = np.asarray(current_hw)
        minima = np.maximum(minima, current_hw - linf_bound)
        maxima = np.minimum(maxima, current_hw + linf_bound)
    h, w = rng.uniform(minima, maxima)
    return h, w


# --------------------------------------------------------------------------- the nine tasks

def _choice(rng, arr, size=None):
    """rng.choice over en.SHAPE_COLOURS / en.SHAPE_TYPES (object arrays of 4): same draws on arange(4)."""
    r = rng.choice(arr, size=size)
    return int(r) if size is None else [int(v) for v in r.tolist()]


def reset_move_to_corner(b, rand_shape_colour=False, rand_shape_type=False, rand_poses=False, debug_reward=False):
    """MoveToCornerEnv.on_reset (move_to_corner.py:31-65)."""
    rng = b.rng
    robot_pos = rng.rand(2)
    b.add_robot(robot_pos, 0.55 * math.pi, first=True)
    colour, stype = RED, SQUARE
    if rand_shape_colour:
        colour = _choice(rng, ALL_COLOURS)
    if rand_shape_type:
        stype = _choice(rng, ALL_TYPES)
    blk = b.add_block(stype, colour, (0.1, -0.65), 0.13 * math.pi)
    b.scene.debug_reward = int(debug_reward)
    if rand_poses:
        b.randomise_all_poses([ROBOT, blk], rand_pos=True, rand_rot=True,
                              rel_pos_linf_limits=JITTER_POS_BOUND, rel_rot_limits=JITTER_ROT_BOUND)


def reset_move_to_region(b, rand_poses_minor=False, rand_poses_full=False, rand_goal_colour=False):
    """MoveToRegionEnv.on_reset (move_to_region.py:30-83)."""
    assert not (rand_poses_minor and rand_poses_full)
    rng = b.rng
    x, y, h, w = (-0.62, -0.17, 0.76, 0.75)
    if rand_poses_minor or rand_poses_full:
        hw_bound = JITTER_TARGET_BOUND if rand_poses_minor else None
        h, w = randomise_hw(RAND_GOAL_MIN_SIZE, RAND_GOAL_MAX_SIZE, rng, current_hw=(h, w), linf_bound=hw_bound)
    colour = _choice(rng, ALL_COLOURS) if rand_goal_colour else BLUE
    goal = b.add_goal(x, y, h, w, colour)
    b.add_robot((0.058, 0.53), -2.13)
    if rand_poses_minor or rand_poses_full:
        if rand_poses_minor:
            pos_limits, rot_limits = JITTER_POS_BOUND, [None, JITTER_ROT_BOUND]
        else:
            pos_limits = rot_limits = None
        b.randomise_all_poses([goal, ROBOT], rand_pos=True, rand_rot=(False, True),
                              rel_pos_linf_limits=pos_limits, rel_rot_limits=rot_limits)


def reset_match_regions(b, rand_target_colour=False, rand_shape_type=False, rand_shape_count=False,
                        rand_layout_minor=False, rand_layout_full=False):
    """MatchRegionsEnv.on_reset (match_regions.py:44-191)."""
    if rand_shape_count:
        assert rand_layout_full and rand_shape_type and rand_target_colour
    rng = b.rng
    robot_pose = ((-0.5, 0.1), -math.pi * 1.2)
    target_colour = _choice(rng, ALL_COLOURS) if rand_target_colour else GREEN
    distractor_colours = [c for c in (RED, GREEN, BLUE, YELLOW) if c != target_colour]
    target_h, target_w, target_x, target_y = 0.7, 0.6, 0.1, 0.7
    if rand_layout_minor or rand_layout_full:
        hw_bound = JITTER_TARGET_BOUND if rand_layout_minor else None
        target_h, target_w = randomise_hw(RAND_GOAL_MIN_SIZE, RAND_GOAL_MAX_SIZE, rng,
                                          current_hw=(target_h, target_w), linf_bound=hw_bound)
    sensor = b.add_goal(target_x, target_y, target_h, target_w, target_colour)
    default_target_types = [STAR, SQUARE]
    default_distractor_types = [[], [PENTAGON], [CIRCLE, PENTAGON]]
    default_target_poses = [(0.8, -0.7, 2.37), (-0.68, 0.72, 1.28)]
    default_distractor_poses = [[], [(-0.05, -0.2, -1.09)], [(-0.75, -0.55, 2.78), (0.3, -0.82, -1.15)]]
    if rand_shape_count:
        target_count = rng.randint(1, 2 + 1)
        distractor_counts = [rng.randint(0, 2 + 1) for _ in distractor_colours]
    else:
        target_count = len(default_target_types)
        distractor_counts = [len(lst) for lst in default_distractor_types]
    if rand_shape_type:
        target_types = [_choice(rng, ALL_TYPES) for _ in range(target_count)]
        distractor_types = [[_choice(rng, ALL_TYPES) for _ in range(dc)] for dc in distractor_counts]
    else:
        target_types, distractor_types = default_target_types, default_distractor_types
    if rand_layout_full:
        target_poses = [(0, 0, 0)] * target_count
        distractor_poses = [[(0, 0, 0)] * dc for dc in distractor_counts]
    else:
        target_poses, distractor_poses = default_target_poses, default_distractor_poses
    shape_ents = []
    for st, (x, y, a) in zip(target_types, target_poses):
        shape_ents.append(b.add_block(st, target_colour, (x, y), a, role=L.ROLE_TARGET))
    for dcol, dtypes, dposes in zip(distractor_colours, distractor_types, distractor_poses):
        for st, (x, y, a) in zip(dtypes, dposes):
            shape_ents.append(b.add_block(st, dcol, (x, y), a, role=L.ROLE_DISTRACTOR))
    b.add_robot(*robot_pose)
    if rand_layout_minor or rand_layout_full:
        all_ents = [sensor, ROBOT, *shape_ents]
        if rand_layout_minor:
            pos_limits, rot_limits = JITTER_POS_BOUND, JITTER_ROT_BOUND
        else:
            pos_limits = rot_limits = None
        rand_rot = [False] + [True] * (len(all_ents) - 1)
        b.randomise_all_poses(all_ents, rand_pos=True, rand_rot=rand_rot,
                              rel_pos_linf_limits=pos_limits, rel_rot_limits=rot_limits)


def reset_make_line(b, rand_colours=False, rand_shapes=False, rand_count=False, rand_layout_minor=False,
                    rand_layout_full=False):
    """MakeLineEnv.on_reset (make_line.py:93-140)."""
    if rand_count:
        assert rand_layout_full and rand_shapes and rand_colours
    rng = b.rng
    block_shapes = [STAR, CIRCLE, STAR, PENTAGON]
    block_colours = [BLUE, YELLOW, RED, GREEN]
    block_poses = [((0.790, -0.820), -0.721), ((-0.177, 0.383), -1.733),
                   ((-0.051, -0.128), 2.696), ((-0.292, -0.745), -0.159)]
    if rand_count:
        n_blocks = rng.randint(3, 4 + 1)
        block_poses = block_poses[:1] * n_blocks
    else:
        n_blocks = len(block_shapes)
    if rand_colours:
        block_colours = _choice(rng, ALL_COLOURS, size=n_blocks)
    if rand_shapes:
        block_shapes = _choice(rng, ALL_TYPES, size=n_blocks)
    blocks = [b.add_block(st, col, pos, ang) for st, col, (pos, ang) in zip(block_shapes, block_colours, block_poses)]
    b.add_robot((0.702, -0.255), 0.347)
    if rand_layout_minor or rand_layout_full:
        if rand_layout_minor:
            pos_limits, rot_limit = JITTER_POS_BOUND, JITTER_ROT_BOUND
        else:
            pos_limits = rot_limit = None
        b.randomise_all_poses([ROBOT, *blocks], rand_pos=True, rand_rot=True,
                              rel_pos_linf_limits=pos_limits, rel_rot_limits=rot_limit)


def reset_find_dupe(b, rand_colours=False, rand_shapes=False, rand_count=False, rand_layout_minor=False,
                    rand_layout_full=False):
    """FindDupeEnv.on_reset (find_dupe.py:72-201)."""
    if rand_count:
        assert rand_layout_full and rand_shapes and rand_colours
    rng = b.rng
    query_colour, query_shape = YELLOW, PENTAGON
    out_colours = [GREEN, RED, RED, YELLOW, BLUE, query_colour]
    out_shapes = [PENTAGON, CIRCLE, CIRCLE, SQUARE, STAR, query_shape]
    out_poses = [((-0.066751, 0.7552), -2.9266), ((-0.05195, 0.31468), 1.5418), ((0.57528, -0.46865), -2.2141),
                 ((0.40594, -0.74977), 0.24582), ((0.45254, 0.3681), -1.0834), ((0.76849, -0.10652), 0.10028)]
    n_out = len(out_colours)
    if rand_count:
        n_out = rng.randint(1, 5 + 1) + 1
    n_distractors = n_out - 1
    if rand_colours:
        query_colour = _choice(rng, ALL_COLOURS)
        out_colours = _choice(rng, ALL_COLOURS, size=n_distractors)
        out_colours.append(query_colour)
    if rand_shapes:
        query_shape = _choice(rng, ALL_TYPES)
        out_shapes = _choice(rng, ALL_TYPES, size=n_distractors)
        out_shapes.append(query_shape)
    xyhw = (-0.72, -0.22, 0.67, 0.72)
    if rand_layout_minor or rand_layout_full:
        hw_bound = JITTER_TARGET_BOUND if rand_layout_minor else None
        hw = randomise_hw(RAND_GOAL_MIN_SIZE, RAND_GOAL_MAX_SIZE, rng, current_hw=xyhw[2:], linf_bound=hw_bound)
        xyhw = (*xyhw[:2], *hw)
    sensor = b.add_goal(*xyhw, query_colour)
    if rand_count:
        out_poses = [((0, 0), 0)] * n_out
    outside = []
    for st, col, (pos, ang) in zip(out_shapes, out_colours, out_poses):
        is_target = col == query_colour and st == query_shape
        outside.append(b.add_block(st, col, pos, ang, role=L.ROLE_TARGET if is_target else L.ROLE_DISTRACTOR))
    query = b.add_block(query_shape, query_colour, (-0.33, -0.49), -0.51, role=L.ROLE_TARGET)
    b.add_robot((-0.57, 0.25), 3.83)
    if rand_layout_minor or rand_layout_full:
        if rand_layout_minor:
            pos_limits, rot_limit = JITTER_POS_BOUND, JITTER_ROT_BOUND
        else:
            pos_limits = rot_limit = None
        all_ents = [sensor, ROBOT, *outside]
        rand_rot = [False] + [True] * (len(all_ents) - 1)
        b.randomise_all_poses(all_ents, rand_pos=True, rand_rot=rand_rot, rel_pos_linf_limits=pos_limits,
                              rel_rot_limits=rot_limit, ignore=[query])
        query_pos_limit = max(0, min(xyhw[2:]) / 2 - SHAPE_RAD / 2)
        if rand_layout_minor:
            query_pos_limit = min(JITTER_POS_BOUND, query_pos_limit)
        b.shift(query, position=b.get_pose(sensor)[:2])
        active = [e for e in b.all_entities() if e != sensor]
        b.randomise_pose(query, active, rand_pos=True, rand_rot=True, rel_pos_linf_limit=query_pos_limit,
                         rel_rot_limit=rot_limit)


def reset_fix_colour(b, rand_colours=False, rand_shapes=False, rand_count=False, rand_layout_minor=False,
                     rand_layout_full=False):
    """FixColourEnv.on_reset (fix_colour.py:67-191)."""
    if rand_count:
        assert rand_layout_full and rand_shapes and rand_colours
    rng = b.rng
    block_colours = [GREEN, GREEN, BLUE]
    region_colours = [GREEN, GREEN, RED]
    block_shapes = [PENTAGON, SQUARE, PENTAGON]
    block_poses = [((0.289, 0.030), 0.307), ((0.133, -0.561), 1.699), ((-0.336, 0.000), -1.529)]
    region_xyhws = [(-0.032, 0.348, 0.427, 0.468), (0.019, -0.391, 0.460, 0.458), (-0.681, 0.196, 0.498, 0.418)]
    n_regions = len(block_colours)
    if rand_count:
        n_regions = rng.randint(2, 3 + 1)
        block_poses = block_poses[:1] * n_regions
        region_xyhws = region_xyhws[:1] * n_regions
    if rand_colours:
        region_colours = _choice(rng, ALL_COLOURS, size=n_regions)
        block_colours = list(region_colours)
        odd_idx = rng.randint(len(block_colours))
        new_col_idx = rng.randint(4 - 1)
        if new_col_idx == block_colours[odd_idx]:
            new_col_idx += 1
        block_colours[odd_idx] = int(new_col_idx)
    if rand_shapes:
        block_shapes = _choice(rng, ALL_TYPES, size=n_regions)
    if rand_layout_minor or rand_layout_full:
        hw_bound = JITTER_TARGET_BOUND if rand_layout_minor else None
        region_xyhws = [(x, y, *randomise_hw(0.4, 0.5, rng, current_hw=hw, linf_bound=hw_bound))
                        for x, y, *hw in region_xyhws]
    sensors = []
    for i, (colour, xyhw, bcol) in enumerate(zip(region_colours, region_xyhws, block_colours)):
        # fix_colour.py:127-141: the region keeps its block only when the colours agree
        sensors.append(b.add_goal(*xyhw, colour, want_block=i if bcol == colour else -1))
    blocks = [b.add_block(st, col, pos, ang) for st, col, (pos, ang) in zip(block_shapes, block_colours, block_poses)]
    b.add_robot((0.368, 0.586), 0.718)
    if rand_layout_minor or rand_layout_full:
        if rand_layout_minor:
            pos_limits, rot_limit = JITTER_POS_BOUND, JITTER_ROT_BOUND
        else:
            pos_limits = rot_limit = None
        rand_rot = [False] * n_regions + [True]
        b.randomise_all_poses([*sensors, ROBOT], rand_pos=True, rand_rot=rand_rot, rel_pos_linf_limits=pos_limits,
                              rel_rot_limits=rot_limit, ignore=blocks)
        for blk, sensor in zip(blocks, sensors):
            b.shift(blk, position=b.get_pose(sensor)[:2])
        for blk, sensor, (_, _, *sensor_hw) in zip(blocks, sensors, region_xyhws):
            block_pos_limit = max(0, min(sensor_hw) / 2 - SHAPE_RAD)
            if rand_layout_minor:
                block_pos_limit = min(JITTER_POS_BOUND, block_pos_limit)
            active = [e for e in b.all_entities() if e != sensor]
            b.randomise_pose(blk, active, rand_pos=True, rand_rot=True, rel_pos_linf_limit=block_pos_limit,
                             rel_rot_limit=rot_limit)


CLUSTER_DEFAULTS = {
    'colour': dict(  # ClusterColourEnv, cluster.py:219-256
        robot=((0.71692, -0.34374), 0.83693),
        colours=[BLUE, BLUE, BLUE, GREEN, GREEN, RED, YELLOW, YELLOW],
        shapes=[CIRCLE, STAR, SQUARE, PENTAGON, PENTAGON, SQUARE, STAR, PENTAGON],
        poses=[((-0.5147, 0.14149), -0.38871), ((-0.1347, -0.71414), 1.0533), ((-0.74247, -0.097592), 1.1571),
               ((-0.077363, -0.42964), -0.64379), ((0.51978, 0.1853), -1.1762), ((-0.5278, -0.21642), 2.9356),
               ((-0.54039, 0.48292), 0.072818), ((-0.16761, 0.64303), -2.3255)]),
    'type': dict(    # ClusterShapeEnv, cluster.py:259-297
        robot=((0.286, -0.202), -1.878),
        colours=[YELLOW, BLUE, RED, RED, GREEN, YELLOW, BLUE, GREEN],
        shapes=[SQUARE, PENTAGON, PENTAGON, PENTAGON, CIRCLE, STAR, STAR, CIRCLE],
        poses=[((-0.414, 0.297), -1.731), ((0.068, 0.705), 2.184), ((0.821, 0.220), 0.650),
               ((-0.461, -0.749), -2.673), ((0.867, -0.149), -2.215), ((-0.785, -0.140), -0.405),
               ((-0.305, -0.226), 1.341), ((0.758, -0.708), -2.140)]),
}


def reset_cluster(b, cluster_by, rand_shape_colour=False, rand_shape_type=False, rand_layout_minor=False,
                  rand_layout_full=False, rand_shape_count=False):
    """BaseClusterEnv.on_reset (cluster.py:67-164)."""
    assert not (rand_layout_minor and rand_layout_full)
    if rand_shape_count:
        assert rand_layout_full and rand_shape_type and rand_shape_colour
    rng = b.rng
    d = CLUSTER_DEFAULTS[cluster_by]
    if rand_shape_count:
        n_shapes = rng.randint(7, 10 + 1)
        poses = [((0, 0), 0)] * n_shapes
    else:
        n_shapes = len(d['colours'])
        poses = d['poses']
    if rand_shape_colour:
        colours = [RED, GREEN, BLUE, YELLOW]
        colours.extend([_choice(rng, ALL_COLOURS) for _ in range(n_shapes - len(colours))])
        rng.shuffle(colours)
    else:
        colours = d['colours']
    if rand_shape_type:
        shape_types = [SQUARE, PENTAGON, STAR, CIRCLE]
        shape_types.extend([_choice(rng, ALL_TYPES) for _ in range(n_shapes - len(shape_types))])
        rng.shuffle(shape_types)
    else:
        shape_types = d['shapes']
    blocks = [b.add_block(st, col, pos, ang) for (pos, ang), col, st in zip(poses, colours, shape_types)]
    b.add_robot(*d['robot'])
    if rand_layout_full or rand_layout_minor:
        if rand_layout_full:
            pos_limit = rot_limit = None
        else:
            pos_limit, rot_limit = JITTER_POS_BOUND, JITTER_ROT_BOUND
        b.randomise_all_poses([ROBOT, *blocks], rand_pos=True, rand_rot=True, rel_pos_linf_limits=pos_limit,
                              rel_rot_limits=rot_limit)


def reset_pick_and_place(b, rand_shape_colour=False, rand_shape_type=False, rand_poses=False, debug_reward=False):
    """PickAndPlaceEnv.on_reset (pick_and_place.py:31-79)."""
    rng = b.rng
    b.add_robot((0.0, 0.0), 0.55 * math.pi, first=True)
    blocks, specs = [], []
    for _ in range(3):
        colour, stype = RED, SQUARE
        if rand_shape_colour:
            colour = _choice(rng, ALL_COLOURS)
        if rand_shape_type:
            stype = _choice(rng, ALL_TYPES)
        blocks.append(b.add_block(stype, colour, (0.1, -0.65), 0.13 * math.pi))
        specs.append((stype, colour))
    target_position = rng.rand(2) * 2 - 1
    valid = [i for i, sp in enumerate(specs) if sp == specs[0]]
    target = int(rng.choice(np.arange(len(valid))))
    s = b.scene
    s.target_block, s.n_valid_targets = valid[target], len(valid)
    s.target_pos[0], s.target_pos[1] = float(target_position[0]), float(target_position[1])
    s.debug_reward = int(debug_reward)
    if rand_poses:
        b.randomise_all_poses([ROBOT, *blocks], rand_pos=True, rand_rot=True)


# entry point (as registered by the reference, benchmarks/__init__.py:440-1014) -> (task id, sampler, fixed kwargs)
ENTRY_POINTS = {
    'magical.benchmarks.move_to_corner:MoveToCornerEnv': (L.TASK_MOVE_TO_CORNER, reset_move_to_corner, {}),
    'magical.benchmarks.move_to_region:MoveToRegionEnv': (L.TASK_MOVE_TO_REGION, reset_move_to_region, {}),
    'magical.benchmarks.match_regions:MatchRegionsEnv': (L.TASK_MATCH_REGIONS, reset_match_regions, {}),
    'magical.benchmarks.make_line:MakeLineEnv': (L.TASK_MAKE_LINE, reset_make_line, {}),
    'magical.benchmarks.find_dupe:FindDupeEnv': (L.TASK_FIND_DUPE, reset_find_dupe, {}),
    'magical.benchmarks.fix_colour:FixColourEnv': (L.TASK_FIX_COLOUR, reset_fix_colour, {}),
    'magical.benchmarks.cluster:ClusterColourEnv': (L.TASK_CLUSTER_COLOUR, reset_cluster, {'cluster_by': 'colour'}),
    'magical.benchmarks.cluster:ClusterShapeEnv': (L.TASK_CLUSTER_SHAPE, reset_cluster, {'cluster_by': 'type'}),
    'magical.benchmarks.pick_and_place:PickAndPlaceEnv': (L.TASK_PICK_AND_PLACE, reset_pick_and_place, {}),
}


def sample_scene(entry_point, rng, max_episode_steps, rand_dynamics=False, **task_kwargs):
    """BaseEnv.reset() up to (not including) the first render: returns the Scene."""
    task, fn, fixed = ENTRY_POINTS[entry_point]
    b = SceneBuilder(task, max_episode_steps, rng, rand_dynamics)
    fn(b, **fixed, **task_kwargs)
    return b.scene


# --------------------------------------------------------------------------- native engine (csrc/mg_reset.cpp)

# order of each task's boolean kwargs in mg_reset_spec.kw (include/mg_b200.h)
_KW_ORDER = {
    L.TASK_MOVE_TO_CORNER: ('rand_shape_colour', 'rand_shape_type', 'rand_poses', 'debug_reward'),
    L.TASK_PICK_AND_PLACE: ('rand_shape_colour', 'rand_shape_type', 'rand_poses', 'debug_reward'),
    L.TASK_MOVE_TO_REGION: ('rand_poses_minor', 'rand_poses_full', 'rand_goal_colour'),
    L.TASK_MATCH_REGIONS: ('rand_target_colour', 'rand_shape_type', 'rand_shape_count', 'rand_layout_minor',
                           'rand_layout_full'),
    L.TASK_MAKE_LINE: ('rand_colours', 'rand_shapes', 'rand_count', 'rand_layout_minor', 'rand_layout_full'),
    L.TASK_FIND_DUPE: ('rand_colours', 'rand_shapes', 'rand_count', 'rand_layout_minor', 'rand_layout_full'),
    L.TASK_FIX_COLOUR: ('rand_colours', 'rand_shapes', 'rand_count', 'rand_layout_minor', 'rand_layout_full'),
    L.TASK_CLUSTER_COLOUR: ('rand_shape_colour', 'rand_shape_type', 'rand_layout_minor', 'rand_layout_full',
                            'rand_shape_count'),
    L.TASK_CLUSTER_SHAPE: ('rand_shape_colour', 'rand_shape_type', 'rand_layout_minor', 'rand_layout_full',
                           'rand_shape_count'),
}


def _check_flag_combinations(task, kw):
    """The assertions the reference's on_resets make on their flags (e.g. cluster.py:56-65)."""
    if task in (L.TASK_MAKE_LINE, L.TASK_FIND_DUPE, L.TASK_FIX_COLOUR) and kw.get('rand_count'):
        assert kw.get('rand_layout_full') and kw.get('rand_shapes') and kw.get('rand_colours')
    if task == L.TASK_MATCH_REGIONS and kw.get('rand_shape_count'):
        assert kw.get('rand_layout_full') and kw.get('rand_shape_type') and kw.get('rand_target_colour')
    if task in (L.TASK_CLUSTER_COLOUR, L.TASK_CLUSTER_SHAPE):
        assert not (kw.get('rand_layout_minor') and kw.get('rand_layout_full'))
        if kw.get('rand_shape_count'):
            assert kw.get('rand_layout_full') and kw.get('rand_shape_type') and kw.get('rand_shape_colour')
    if task == L.TASK_MOVE_TO_REGION:
        assert not (kw.get('rand_poses_minor') and kw.get('rand_poses_full'))


def reset_spec(entry_point, max_episode_steps, rand_dynamics=False, **task_kwargs):
    """mg_reset_spec for one registered env (entry point + its kwargs)."""
    task, _, _ = ENTRY_POINTS[entry_point]
    names = _KW_ORDER[task]
    unknown = set(task_kwargs) - set(names)
    if unknown:
        raise TypeError(f'unexpected keyword arguments for {entry_point}: {sorted(unknown)}')
    _check_flag_combinations(task, task_kwargs)
    spec = L.ResetSpec()
    spec.task, spec.horizon, spec.rand_dynamics = task, int(max_episode_steps), int(bool(rand_dynamics))
    for k, name in enumerate(names):
        spec.kw[k] = int(bool(task_kwargs.get(name, False)))
    return spec


def sample_scene_native(entry_point, rng, max_episode_steps, rand_dynamics=False, **task_kwargs):
    """sample_scene through the C++ engine: same Scene bytes, same RandomState stream afterwards."""
    spec = reset_spec(entry_point, max_episode_steps, rand_dynamics, **task_kwargs)
    name, key, pos, has_gauss, cached = rng.get_state()
    key = np.ascontiguousarray(key, np.uint32).copy()
    cpos = L.C.c_int32(int(pos))
    scene = Scene()
    rc = L.lib().mg_host_sample_scene(L.C.byref(spec), key.ctypes.data, L.C.addressof(cpos), L.C.byref(scene))
    rng.set_state((name, key, int(cpos.value), has_gauss, cached))
    if rc == L.ERR_PLACEMENT:
        raise PlacementError('Could not place an entity (native reset engine)')
    if rc != 0:
        raise RuntimeError(f'mg_host_sample_scene failed ({rc})')
    return scene


def sample_scenes_native(specs, rngs, n_threads=None):
    """Batch form: specs (ResetSpec per env) and RandomStates per env -> list of Scene (PlacementError for any
    failure, like a sequential loop over the envs would raise).  Runs on n_threads host threads."""
    import os
    n = len(specs)
    if n == 0:
        return []
    if n_threads is None:
        n_threads = min(n, len(os.sched_getaffinity(0)) if hasattr(os, 'sched_getaffinity') else (os.cpu_count() or 1))
    spec_arr = (L.ResetSpec * n)(*specs)
    keys = np.empty((n, 624), np.uint32)
    pos = np.empty(n, np.int32)
    states = [r.get_state() for r in rngs]
    for i, st in enumerate(states):
        keys[i] = st[1]
        pos[i] = st[2]
    scenes = (Scene * n)()
    status = np.zeros(n, np.int32)
    rc = L.lib().mg_host_sample_scenes(L.C.addressof(spec_arr), n, keys.ctypes.data, pos.ctypes.data,
                                       L.C.addressof(scenes), status.ctypes.data, int(n_threads))
    if rc != 0:
        raise RuntimeError(f'mg_host_sample_scenes failed ({rc})')
    for i, (r, st) in enumerate(zip(rngs, states)):
        r.set_state((st[0], keys[i], int(pos[i]), st[3], st[4]))
    bad = np.nonzero(status)[0]
    if len(bad):
        if (status[bad] == L.ERR_PLACEMENT).all():
            raise PlacementError(f'Could not place an entity in env(s) {bad.tolist()[:8]} (native reset engine)')
        raise RuntimeError(f'native reset failed for envs {bad.tolist()[:8]}: {status[bad].tolist()[:8]}')
    return [L.copy_scene(scenes[i]) for i in range(n)]
