"""The reference's own CUDA kernels (spotfinder/kernels/thresholding.cu, erosion.cu, wrappers
spotfinder/spotfinder.cu:148-189,213-347), compiled in place for sm_100 (oracle/Makefile target
`refgpu`), run beside the product's kernels on the same frames.

They are the "kernel to beat" (bench.py times them), NOT the parity oracle: they evaluate the
tests in fp32 ratio form, default to min_count 3 (N2/N8), skip masked neighbours in
the erosion (N6) and need n > 0 in the final pass (N7), while the product follows the CPU
algorithm (standalone.cc / baseline.cpp) bit for bit.  What this file pins is that the two
GPU implementations agree wherever those documented divergences cannot bite, so a gross error
in either would show."""
import numpy as np
import pytest

from ffs_b200 import synth
from ffs_b200.spotfinder import Spotfinder
from oracle import oracle

pytestmark = pytest.mark.gpu


def _frames(dtype, n=2):
    g = synth.small_geometry(1200, 600)
    mask = synth.make_mask(g)
    # no hot pixels: the reference evaluates pixel*pixel in the narrow type (N13, thresholding.cu:95-97)
    frames = synth.make_frames(g, mask, n, dtype, config=6, n_spots=60, n_hot=0)
    frames[:, mask == 0] = 0  # gap value irrelevant (masked) but keeps N13 out of the picture entirely
    return g, mask, frames


@pytest.mark.parametrize("dtype", [np.uint16, np.uint32])
def test_dispersion_agrees_with_reference_gpu_kernel(dtype):
    if not oracle.refgpu_available(np.dtype(dtype).itemsize):
        pytest.skip("oracle/_ref/libffs_refgpu*.so not built")
    g, mask, frames = _frames(dtype)
    ref, ms = oracle.refgpu_run(frames, mask, oracle.ALGO_DISPERSION, min_count=2)
    assert ms > 0
    with Spotfinder(g.width, g.height, dtype, mask, n_slots=len(frames)) as sf:
        sf.upload(frames)
        sf.threshold(0, len(frames))
        for i in range(len(frames)):
            got = sf.strong_mask(i)
            want = oracle.dispersion(frames[i], mask)
            assert np.array_equal(got, want)  # the product is bit-exact against the CPU algorithm
            n_strong = int(want.sum())
            n_diff = int((got != ref[i]).sum())
            # fp32 ratio form vs exact integer/fp64: only borderline pixels may flip
            assert n_strong > 100 and n_diff <= max(2, 0.01 * n_strong), (n_strong, n_diff)


def test_extended_overlaps_reference_gpu_kernel():
    if not oracle.refgpu_available(2):
        pytest.skip("oracle/_ref/libffs_refgpu.so not built")
    g, mask, frames = _frames(np.uint16)
    ref, _ = oracle.refgpu_run(frames, mask, oracle.ALGO_DISPERSION_EXTENDED, min_count=2)
    with Spotfinder(g.width, g.height, np.uint16, mask, n_slots=len(frames), algorithm="dispersion_extended") as sf:
        sf.upload(frames)
        sf.threshold(0, len(frames))
        for i in range(len(frames)):
            got = sf.strong_mask(i)
            assert np.array_equal(got, oracle.dispersion_extended(frames[i], mask))
            # away from masked pixels (N6) the two implementations find the same spots
            inter = int((got & ref[i]).sum())
            union = int((got | ref[i]).sum())
            assert union > 100 and inter / union > 0.9, (inter, union)


def test_reference_gpu_kernels_only_timer_runs():
    if not oracle.refgpu_available(2):
        pytest.skip("oracle/_ref/libffs_refgpu.so not built")
    g, mask, frames = _frames(np.uint16)
    for algo in (oracle.ALGO_DISPERSION, oracle.ALGO_DISPERSION_EXTENDED):
        _, ms = oracle.refgpu_run(frames, mask, algo, repeats=3, kernels_only=True)
        assert 0 < ms < 1000
