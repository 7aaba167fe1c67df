"""CPU side of the host-assembled end-to-end step: mg_host_stack_frames (the worker code of mg_step_host with
mg_set_host_assembly on) against FlattenFrameStack's layout written out in numpy
(/root/reference/magical/benchmarks/__init__.py:122-147: frames concatenated along the channel axis, oldest first; after
a reset the first frame is repeated)."""
import numpy as np
import pytest

import magical_b200  # noqa: F401
from magical_b200 import _lib as L


def _expect(ring, newest, ages, since):
    n = ring.shape[0]
    out = np.empty((n, 96, 96, 3 * len(ages)), np.uint8)
    for e in range(n):
        for g, a in enumerate(ages):
            slot = (newest - min(a, int(since[e]))) & 3
            out[e, :, :, 3 * g:3 * g + 3] = ring[e, slot]
    return out


@pytest.mark.parametrize('ages,threads,misalign', [((3, 2, 1, 0), 1, 0), ((3, 2, 1, 0), 5, 0), ((3, 2, 1, 0), 3, 1),
                                                   ((2, 1, 0), 2, 0), ((0,), 1, 0), ((3, 3, 0, 1), 4, 0)])
def test_host_stack_frames_matches_numpy(ages, threads, misalign):
    lib = L.lib()
    rs = np.random.RandomState(len(ages) * 7 + threads)
    n = 37
    ring = rs.randint(0, 256, size=(n, 4, 96, 96, 3)).astype(np.uint8)
    since = rs.randint(0, 4, size=n).astype(np.uint8)
    ages_a = np.asarray(ages, np.int32)
    for newest in range(4):
        buf = np.zeros(n * 96 * 96 * 3 * len(ages) + 64, np.uint8)
        base = (-buf.ctypes.data) % 64 + misalign          # 64-byte aligned start, or one byte off it
        out = buf[base:base + n * 96 * 96 * 3 * len(ages)]
        rc = lib.mg_host_stack_frames(ring.ctypes.data, out.ctypes.data, n, newest, ages_a.ctypes.data, len(ages),
                                      since.ctypes.data, threads)
        assert rc == 0
        np.testing.assert_array_equal(out.reshape(n, 96, 96, 3 * len(ages)), _expect(ring, newest, ages, since))
        assert not buf[:base].any() and not buf[base + out.size:].any()      # nothing written outside


def test_host_stack_frames_rejects_bad_arguments():
    lib = L.lib()
    ring = np.zeros((1, 4, 96, 96, 3), np.uint8)
    out = np.zeros((1, 96, 96, 12), np.uint8)
    since = np.zeros(1, np.uint8)
    bad_age = np.asarray([4, 0, 0, 0], np.int32)
    assert lib.mg_host_stack_frames(ring.ctypes.data, out.ctypes.data, 1, 0, bad_age.ctypes.data, 4, since.ctypes.data, 1) != 0
    ok_age = np.asarray([3, 2, 1, 0], np.int32)
    assert lib.mg_host_stack_frames(None, out.ctypes.data, 1, 0, ok_age.ctypes.data, 4, since.ctypes.data, 1) != 0
    assert lib.mg_host_stack_frames(ring.ctypes.data, out.ctypes.data, 1, 0, ok_age.ctypes.data, 9, since.ctypes.data, 1) != 0
