"""--validate / ffs_self_check (csrc/api_selfcheck.inl): the reference's --validate re-runs its CPU
DIALS code per image and compares the strong pixels (spotfinder.cc:1010-1052); this build holds no
CPU implementation, so the second opinion is the device's exact per-pixel routine (direct window
sums, no pruning) and the two strong-pixel planes are compared.  The check itself is checked here:
it agrees with the oracle's count on good runs and it fires when the fused kernel is made to
drop its resolve stage (developer knob FFS_FUSED_DEBUG=2)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle
from test_gpu_cli import BIN, run, write_stream

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_self_check_through_the_abi(dtype):
    g = synth.small_geometry(520, 300)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 5, dtype, config=2, n_spots=40)
    with Spotfinder(g.width, g.height, dtype, mask, n_slots=3) as sf:
        assert sf._L.ffs_self_check(sf._ctx, 1) == 0
        for i, f in enumerate(frames):
            sf.submit_copy(i, f)
            res, _ = sf.wait()
            exact, diff = C.c_uint32(), C.c_uint32()
            assert sf._L.ffs_last_self_check(sf._ctx, C.byref(exact), C.byref(diff)) == 0
            want = int(oracle.dispersion(f, mask).sum())
            assert (res.num_strong_pixels, exact.value, diff.value) == (want, want, 0)
        # with a trusted range and a gain map the exact pass follows suit
        gain = np.random.default_rng(3).uniform(0.7, 1.4, mask.shape)
        sf.set_gain(gain)
        sf.submit_copy(9, frames[0])
        res, _ = sf.wait()
        exact, diff = C.c_uint32(), C.c_uint32()
        sf._L.ffs_last_self_check(sf._ctx, C.byref(exact), C.byref(diff))
        want = int(oracle.dispersion_gain(frames[0], mask, gain).sum())
        assert (res.num_strong_pixels, exact.value, diff.value) == (want, want, 0)
        assert sf._L.ffs_self_check(sf._ctx, 0) == 0
        assert sf._L.ffs_last_self_check(sf._ctx, C.byref(exact), C.byref(diff)) != 0   # off: nothing to report
    with Spotfinder(g.width, g.height, dtype, mask, n_slots=1, algorithm="dispersion_extended") as sf:
        assert sf._L.ffs_self_check(sf._ctx, 1) != 0                                    # standard dispersion only


def test_validate_flag_prints_the_reference_lines(tmp_path):
    g = synth.small_geometry(300, 200)
    mask = synth.make_mask(g)
    frames = synth.make_frames(g, mask, 4, np.uint16, config=8, n_spots=12)
    write_stream(tmp_path, frames, mask, 16)
    det = '{"pixel_size_x": 0.075, "pixel_size_y": 0.075, "beam_center_x": 11.0, "beam_center_y": 7.5, "distance": 150.0}'
    rc, lines, out, err = run("spotfinder", [str(tmp_path), "--threads", "2", "--validate", "--detector", det])
    assert rc == 0 and err == ""
    got = {int(m.group(1)): int(m.group(2)) for m in re.finditer(r"Image\s+(\d+): Compared: \x1b\[32mMatch (\d+) px", out)}
    assert got == {i: int(oracle.dispersion(f, mask).sum()) for i, f in enumerate(frames)}
    assert "Mismatch" not in out and len(lines) == 4
    # the fused kernel with its resolve stage switched off finds nothing; the exact pass notices
    env = dict(os.environ, FFS_DEV_KNOBS="1", FFS_FUSED_DEBUG="2")
    r = subprocess.run([str(BIN / "spotfinder"), str(tmp_path), "--threads", "2", "--validate", "--detector", det],
                       capture_output=True, text=True, env=env, timeout=120)
    assert r.returncode == 1 and r.stdout.count("Mismatch (0 px from kernel)") == 4
    # extended: not validated, runs as usual
    rc, lines, out, err = run("spotfinder", [str(tmp_path), "--threads", "2", "--validate", "--algorithm",
                                              "dispersion_extended", "--detector", det])
    assert rc == 0 and "not validating" in out and "finished image" in out
