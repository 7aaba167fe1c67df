"""The C-ABI library loads without a GPU and exports every symbol include/mg_b200.h declares; the ctypes
binding covers them all.  No compute entry point is called here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'mg_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(mg_[a-z_0-9]+)\s*\(', text)))


def test_header_symbols_are_exported_and_bound():
    import magical_b200
    L = magical_b200._lib
    lib = L.lib()
    names = declared_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), f'{name} declared in include/mg_b200.h but not exported'
        assert name in L.SYMBOLS, f'{name} has no ctypes prototype'
    assert set(L.SYMBOLS) == set(names)


def test_scene_struct_matches_header():
    """sizeof(mg_scene) computed from the header by the C compiler == the ctypes mirror."""
    import subprocess
    import tempfile

    import magical_b200
    src = '#include <stdio.h>\n#include "mg_b200.h"\nint main(){printf("%zu %zu %zu %zu", sizeof(mg_scene), sizeof(mg_block), sizeof(mg_goal), sizeof(mg_config));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, 't.c')
        open(c, 'w').write(src)
        exe = os.path.join(d, 't')
        subprocess.run(['gcc', '-I', os.path.join(ROOT, 'include'), c, '-o', exe], check=True)
        sizes = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    L = magical_b200._lib
    assert sizes == [C.sizeof(L.Scene), C.sizeof(L.Block), C.sizeof(L.Goal), C.sizeof(L.Config)]


def test_create_fails_loudly_without_gpu():
    import torch

    import magical_b200
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    L = magical_b200._lib
    cfg = L.Config()
    cfg.num_envs = 4
    h = C.c_void_p()
    rc = L.lib().mg_create(C.byref(cfg), C.byref(h))
    assert rc == -3 and b'no CUDA device' in L.lib().mg_last_error(None)
    with pytest.raises(RuntimeError):
        magical_b200.make('MoveToCorner-Demo-v0')


def test_palette_matches_reference_style():
    """mg_host_palette against tests/golden/constants.json (generated from the reference's style.py)."""
    import json

    import magical_b200
    g = json.load(open(os.path.join(ROOT, 'tests', 'golden', 'constants.json')))
    out = np.zeros((32, 3), np.uint8)
    n = magical_b200._lib.lib().mg_host_palette(out.ctypes.data)
    pal = out[:n].tolist()
    order = ['red', 'green', 'blue', 'yellow']
    want = [g['background'], [255, 255, 255], g['colours']['grey'], g['dark']['grey'], g['light4']['grey'], [26, 26, 26]]
    want += [g['colours'][c] for c in order] + [g['dark'][c] for c in order] + [g['light2'][c] for c in order]
    assert pal == want
