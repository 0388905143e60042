"""CPU-side checks: the C-ABI library loads and exports every symbol include/visnav_b200.h
declares, the host mirror fails loudly without a GPU (no fallback), host-side helpers."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import visnav_b200
from visnav_b200 import keypoints as kp
from visnav_b200 import sharding
from visnav_b200 import workloads as wl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    txt = open(os.path.join(ROOT, "include", "visnav_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(vslam_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    syms = _header_symbols()
    assert len(syms) >= 30
    L = ctypes.CDLL(kp.lib_path())
    missing = [s for s in syms if not hasattr(L, s)]
    assert not missing, missing
    assert sorted(kp.ABI_SYMBOLS) == syms      # the python mirror binds exactly the header


def test_library_is_sm100a_and_uses_tma():
    out = subprocess.run(["cuobjdump", "-lelf", kp.lib_path()], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    sass = subprocess.run(["cuobjdump", "-sass", kp.lib_path()], capture_output=True, text=True).stdout
    assert "UTMALDG" in sass and "UBLKCP" in sass and "POPC" in sass


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(kp.VslamError):
        kp.Context(0)
    kd = kp.KeypointsData()
    with pytest.raises(kp.VslamError):
        visnav_b200.detectKeypointsAndDescriptors(np.zeros((64, 64), np.uint8), kd, 10, True)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "visual-slam_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "liboracle" not in txt and "visnav_oracle" not in txt, f
    hdr = open(os.path.join(ROOT, "include", "visnav", "keypoints.h")).read()
    assert "oracle" not in hdr


def test_image_marshalling_keeps_pitch():
    big = np.arange(100 * 120, dtype=np.uint8).reshape(100, 120)
    view = big[5:95, 8:108]
    a, w, h, pitch = kp._img(view)
    assert (w, h, pitch) == (100, 90, 120) and a.ctypes.data == view.ctypes.data    # no copy: pitch honoured
    a, w, h, pitch = kp._img(big[:, ::2])
    assert pitch == w == 60                                                        # strided columns are copied
    with pytest.raises(kp.VslamError):
        kp._img(np.zeros((4, 4), np.float32))


def test_workloads_are_deterministic_and_shaped():
    a, b = wl.synth_stereo_pair(3, 160, 120)
    a2, b2 = wl.synth_stereo_pair(3, 160, 120)
    assert a.shape == (120, 160) and a.dtype == np.uint8 and np.array_equal(a, a2) and np.array_equal(b, b2)
    batch = wl.synth_stereo_batch(3, 96, 64, first=3, workers=1)
    assert np.array_equal(batch[0, 0], wl.synth_stereo_pair(3, 96, 64)[0])
    d = wl.random_descriptors(50, 1)
    f = wl.flip_bits(d, 25, np.random.default_rng(0))
    x = d ^ f
    assert all(bin(int.from_bytes(r.tobytes(), "little")).count("1") == 25 for r in x)
    q, db = wl.reloc_query(20, 5000, 25)
    assert q.shape == (20, 8) and db.shape == (5000, 8)
    jobs = wl.keyframe_pair_jobs(500)
    assert len(jobs) == 124750 and (jobs[:, 0] < jobs[:, 1]).all()
    lv = wl.pyramid_levels(np.full((65, 33), 7, np.uint8), 3)
    assert [l.shape for l in lv] == [(65, 33), (32, 16), (16, 8)] and (lv[2] == 7).all()


def test_shard_range_partitions():
    for n in (0, 1, 7, 2000, 124750):
        for world in (1, 2, 3, 8):
            r = [sharding.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1


def test_bench_reference_arm_prints_contract_line():
    env = dict(os.environ, OMP_NUM_THREADS="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0"], capture_output=True, text=True, env=env, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    import json
    line = json.loads(p.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["value"] > 0 and line["unit"] == "stereo pairs/s"
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["cores"] >= 1
