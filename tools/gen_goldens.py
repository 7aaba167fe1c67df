"""Dev-time generator of tests/golden/*.json from the REFERENCE ITSELF (run in the build container only,
where /root/reference is mounted; the GPU box never reads it).

`import magical` fails here (pymunk / pygame / gym are not installed), so the pieces that are plain
numpy / stdlib are executed directly from the reference's source text:
  * magical/style.py, magical/phys_vars.py            -> constants.json (palette, line widths, phys vars)
  * make_line.longest_line + MakeLineEnv.score_on_end_of_traj          (make_line.py:31-71, 142-152)
  * BaseClusterEnv.score_on_end_of_traj                                 (cluster.py:166-216)
  * MoveToCornerEnv.score_on_end_of_traj / debug_shaped_reward          (move_to_corner.py:67-100)
  * PickAndPlaceEnv.score_on_end_of_traj / debug_shaped_reward          (pick_and_place.py:88-124)
Each method body is compiled into a stand-in class of the same name (so name mangling matches) whose
`self` carries just the attributes the method reads.
"""
import ast
import importlib.util
import itertools as it
import json
import os
import sys
import textwrap
import types

import numpy as np

REF = '/root/reference/magical'
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden')


def load_standalone(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, name + '.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def method_source(path, cls_name, fn_names):
    src = open(path).read()
    tree = ast.parse(src)
    out = []
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == cls_name:
            for item in node.body:
                if isinstance(item, ast.FunctionDef) and item.name in fn_names:
                    out.append(textwrap.indent(textwrap.dedent(ast.get_source_segment(src, item)), '    '))
    return f'class {cls_name}:\n' + '\n'.join(out) + '\n'


def function_source(path, fn_name):
    src = open(path).read()
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name == fn_name:
            return ast.get_source_segment(src, node)
    raise KeyError(fn_name)


def body(x, y):
    pos = types.SimpleNamespace(x=float(x), y=float(y))
    pos_arr = np.asarray([float(x), float(y)])

    class _P(types.SimpleNamespace):
        def __array__(self, dtype=None, copy=None):
            return pos_arr
    return types.SimpleNamespace(shape_body=types.SimpleNamespace(position=_P(x=float(x), y=float(y))),
                                 robot_body=types.SimpleNamespace(position=_P(x=float(x), y=float(y))))


def main():
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.RandomState(20240)
    style = load_standalone('style')
    phys = load_standalone('phys_vars')

    def u8(c):
        return [int(v) for v in np.round(np.asarray(c) * 255)]
    consts = {
        'colours': {k: u8(v) for k, v in style.COLOURS_RGB.items()},
        'dark': {k: u8(style.darken_rgb(v)) for k, v in style.COLOURS_RGB.items()},
        'light2': {k: u8(style.lighten_rgb(v, 2)) for k, v in style.COLOURS_RGB.items()},
        'light4': {k: u8(style.lighten_rgb(v, 4)) for k, v in style.COLOURS_RGB.items()},
        'background': u8(style.lighten_rgb(style.COLOURS_RGB['grey'], times=4)),
        'ARENA_ZOOM_OUT': style.ARENA_ZOOM_OUT, 'GOAL_LINE_THICKNESS': style.GOAL_LINE_THICKNESS,
        'SHAPE_LINE_THICKNESS': style.SHAPE_LINE_THICKNESS, 'ROBOT_LINE_THICKNESS': style.ROBOT_LINE_THICKNESS,
    }
    # base_env.py:49-57 PhysicsVariables (the class body is literal PhysVar(...) calls)
    src = open(os.path.join(REF, 'base_env.py')).read()
    pv = []
    for node in ast.parse(src).body:
        if isinstance(node, ast.ClassDef) and node.name == 'PhysicsVariables':
            for item in node.body:
                if isinstance(item, ast.Assign):
                    default, bounds = (ast.literal_eval(a) for a in item.value.args)
                    pv.append([item.targets[0].id, default, list(bounds)])
    consts['phys_vars'] = pv
    # PhysVar.sample draw order check: 5 uniforms
    r2 = np.random.RandomState(7)
    consts['phys_sample_seed7'] = [phys.PhysVar(d, b).sample(r2) for _, d, b in pv]
    json.dump(consts, open(os.path.join(OUT, 'constants.json'), 'w'), indent=1)

    cases = {'make_line': [], 'cluster': [], 'move_to_corner': [], 'pick_and_place': []}
    ns = {'np': np, 'it': it}
    exec(function_source(os.path.join(REF, 'benchmarks/make_line.py'), 'longest_line'), ns)
    exec(method_source(os.path.join(REF, 'benchmarks/make_line.py'), 'MakeLineEnv', ['score_on_end_of_traj']), ns)
    SHAPE_RAD = 0.2 * 0.6
    for k in range(60):
        n = 3 + k % 2
        if k % 3 == 0:   # near-collinear layouts
            t = np.sort(rng.uniform(-0.8, 0.8, n))
            d = rng.uniform(-1, 1, 2); d /= np.linalg.norm(d)
            pts = t[:, None] * d[None] * rng.uniform(0.3, 1.0) + rng.normal(0, 0.05, (n, 2))
        else:
            pts = rng.uniform(-0.9, 0.9, (n, 2))
        env = ns['MakeLineEnv']()
        env._blocks = [body(x, y) for x, y in pts]
        env.inlier_dist, env.max_sep = SHAPE_RAD * 1.5, SHAPE_RAD * 3.5
        cases['make_line'].append({'points': pts.tolist(), 'score': float(env.score_on_end_of_traj()),
                                   'longest_line': int(ns['longest_line'](pts, env.inlier_dist, env.max_sep))})

    exec(method_source(os.path.join(REF, 'benchmarks/cluster.py'), 'BaseClusterEnv', ['score_on_end_of_traj']), ns)
    colours = ['red', 'green', 'blue', 'yellow']
    types_ = ['square', 'pentagon', 'star', 'circle']
    for k in range(60):
        n = 7 + k % 4
        by_type = k % 2
        cols = list(rng.choice(4, n)); tys = list(rng.choice(4, n))
        labels = tys if by_type else cols
        names = [(types_ if by_type else colours)[v] for v in labels]
        if k % 3 == 0:   # clustered layouts
            centres = rng.uniform(-0.7, 0.7, (4, 2))
            pts = np.asarray([centres[v] for v in labels]) + rng.normal(0, 0.04 + 0.05 * (k % 5), (n, 2))
        else:
            pts = rng.uniform(-0.9, 0.9, (n, 2))
        env = ns['BaseClusterEnv']()
        vals = np.unique(np.asarray(names, dtype='object'))
        blocks = {}
        for (x, y), name in zip(pts, names):
            blocks.setdefault(name, []).append(body(x, y))
        setattr(env, '_BaseClusterEnv__characteristic_values', vals)
        setattr(env, '_BaseClusterEnv__blocks_by_characteristic', blocks)
        cases['cluster'].append({'points': pts.tolist(), 'colours': [int(c) for c in cols], 'types': [int(t) for t in tys],
                                 'by_type': int(by_type), 'score': float(env.score_on_end_of_traj())})

    exec(method_source(os.path.join(REF, 'benchmarks/move_to_corner.py'), 'MoveToCornerEnv',
                       ['score_on_end_of_traj', 'debug_shaped_reward']), ns)
    for k in range(40):
        rp, sp = rng.uniform(-1, 1, 2), rng.uniform(-1, 1, 2)
        env = ns['MoveToCornerEnv']()
        env.robot = env._robot = body(*rp)
        setattr(env, '_MoveToCornerEnv__shape_ref', body(*sp))
        cases['move_to_corner'].append({'robot': rp.tolist(), 'block': sp.tolist(),
                                        'score': float(env.score_on_end_of_traj()),
                                        'debug_reward': float(env.debug_shaped_reward())})

    exec(method_source(os.path.join(REF, 'benchmarks/pick_and_place.py'), 'PickAndPlaceEnv',
                       ['score_on_end_of_traj', 'debug_shaped_reward']), ns)
    for k in range(40):
        rp, sp, tp = rng.uniform(-1, 1, 2), rng.uniform(-1, 1, 2), rng.uniform(-1, 1, 2)
        env = ns['PickAndPlaceEnv']()
        env.SHAPE_RAD = SHAPE_RAD
        env.robot = body(*rp)
        env.target_shape = body(*sp)
        env.valid_target_shapes = [env.target_shape] * (1 + k % 3)
        env.target_position = tp
        cases['pick_and_place'].append({'robot': rp.tolist(), 'block': sp.tolist(), 'target': tp.tolist(),
                                        'n_valid': 1 + k % 3, 'score': float(env.score_on_end_of_traj()),
                                        'debug_reward': float(env.debug_shaped_reward())})
    json.dump(cases, open(os.path.join(OUT, 'reference_scores.json'), 'w'))
    print({k: len(v) for k, v in cases.items()})


if __name__ == '__main__':
    sys.exit(main())
