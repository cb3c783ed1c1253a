"""The C++ host mirror (pathtracing_b200/host): builds on CPU, fails loudly without a GPU, and on
the GPU box renders the reference frame to a PPM that equals the oracle's image."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "pathtracing_b200", "host", "pt_render")


def test_cli_is_built():
    from pathtracing_b200 import _lib
    _lib.build()
    assert os.path.exists(CLI)


def test_cli_fails_loudly_without_a_device(tmp_path):
    from pathtracing_b200 import _lib
    if _lib.load().ptb_device_count() > 0:
        pytest.skip("a CUDA device is present")
    r = subprocess.run([CLI, "--size", "64x36", "--out", str(tmp_path / "x.ppm")], capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr
    assert not (tmp_path / "x.ppm").exists()


def test_cli_rejects_bad_arguments():
    r = subprocess.run([CLI, "--bogus"], capture_output=True, text=True)
    assert r.returncode == 2 and "usage" in r.stderr


@pytest.mark.gpu
def test_cli_renders_the_reference_frame(tmp_path, oracle):
    from pathtracing_b200.image_io import read_ppm
    out = tmp_path / "frame.ppm"
    r = subprocess.run([CLI, "--size", "1920x1080", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    ref = oracle.csmain_frame(oracle.frame_params(1920, 1080))["rgba8"]
    assert np.array_equal(read_ppm(str(out)), ref[..., :3])


@pytest.mark.gpu
def test_cli_progressive_mode(tmp_path, oracle):
    from pathtracing_b200.image_io import read_ppm
    out = tmp_path / "acc.ppm"
    r = subprocess.run([CLI, "--size", "320x180", "--spp", "4", "--frames", "2", "--out", str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    p = oracle.frame_params(320, 180)
    acc = oracle.accumulate(p, 0, 4, 0, True)
    acc = oracle.accumulate(p, 4, 4, 0, True, accum=acc)
    assert np.array_equal(read_ppm(str(out)), oracle.resolve(acc, 8)[..., :3])


@pytest.mark.gpu
def test_cli_seed_is_honoured(tmp_path, oracle):
    from pathtracing_b200.image_io import read_ppm
    out = tmp_path / "seed.ppm"
    r = subprocess.run([CLI, "--size", "320x180", "--spp", "3", "--seed", "77", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    acc = oracle.accumulate(oracle.frame_params(320, 180), 0, 3, 77, True)
    assert np.array_equal(read_ppm(str(out)), oracle.resolve(acc, 3)[..., :3])
    assert not np.array_equal(read_ppm(str(out)), oracle.resolve(oracle.accumulate(oracle.frame_params(320, 180), 0, 3, 0, True), 3)[..., :3])
