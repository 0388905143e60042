"""The C++ drop-in header (include/visnav/keypoints.h) compiled against stub Pangolin / common_types
headers and driven like the reference drives it (slam.cpp:1125-1143)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "test_dropin")


def _build():
    cmd = ["g++", "-std=c++14", "-O2", "-I", os.path.join(ROOT, "tests", "cpp", "stubs"), "-I", os.path.join(ROOT, "include"),
           "-o", EXE, os.path.join(ROOT, "tests", "cpp", "test_dropin.cpp"), "-L", os.path.join(ROOT, "visual-slam_b200", "csrc"),
           "-lvisnav_b200", "-Wl,-rpath," + os.path.join(ROOT, "visual-slam_b200", "csrc")]
    subprocess.run(cmd, check=True)


def test_dropin_header_compiles_and_fails_loudly_without_gpu():
    import torch
    _build()
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([EXE, "/dev/null", "/dev/null", "64", "64"], capture_output=True, text=True)
    assert p.returncode in (2, 3)      # unreadable input or "no usable sm_100 GPU": never a silent CPU result


@pytest.mark.gpu
def test_dropin_header_matches_oracle(tmp_path, oracle_port):
    import visnav_b200  # noqa: F401
    from visnav_b200 import workloads as wl
    _build()
    left, right = wl.synth_stereo_pair(11)
    left.tofile(tmp_path / "l.raw")
    right.tofile(tmp_path / "r.raw")
    p = subprocess.run([EXE, str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), "752", "480"], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout[-500:] + p.stderr[-500:]
    lines = p.stdout.splitlines()
    head = lines[0].split()
    xl, al, dl = oracle_port.detect_describe(left, 1500)
    xr, ar, dr = oracle_port.detect_describe(right, 1500)
    want = oracle_port.match(dl, dr, 70, 1.2)
    assert int(head[1]) == len(xl) and int(head[3]) == len(xr) and int(head[5]) == len(want) and head[7] == "1"
    got_m = np.array([[int(v) for v in ln.split()[1:]] for ln in lines if ln.startswith("m ")], np.int32)
    assert np.array_equal(got_m, want)
    ks = [ln.split() for ln in lines if ln.startswith("k ")]
    assert np.array_equal(np.array([[float(k[1]), float(k[2])] for k in ks]), xl)
    assert np.abs(np.array([float(k[3]) for k in ks]) - al).max() <= 1e-5
    # bitset::to_string prints bit 255 first
    bits = np.array([[int(c) for c in k[4][::-1]] for k in ks], np.uint8)
    words = np.packbits(bits.reshape(len(ks), 8, 32), axis=2, bitorder="little").view(np.uint32).reshape(len(ks), 8)
    assert np.array_equal(words, dl)


EXE_LM = os.path.join(ROOT, "tests", "cpp", "test_dropin_landmarks")


def _build_lm():
    cmd = ["g++", "-std=c++14", "-O2", "-I", os.path.join(ROOT, "tests", "cpp", "stubs"), "-I", os.path.join(ROOT, "include"),
           "-o", EXE_LM, os.path.join(ROOT, "tests", "cpp", "test_dropin_landmarks.cpp"), "-L",
           os.path.join(ROOT, "visual-slam_b200", "csrc"), "-lvisnav_b200",
           "-Wl,-rpath," + os.path.join(ROOT, "visual-slam_b200", "csrc")]
    subprocess.run(cmd, check=True)


def test_landmark_dropin_compiles_and_fails_loudly_without_gpu():
    import torch
    _build_lm()
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([EXE_LM], capture_output=True, text=True)
    assert p.returncode == 3 and "no usable sm_100 GPU" in p.stdout


@pytest.mark.gpu
def test_landmark_dropin_header():
    """find_matches_landmarks_b200 on std containers shaped like the reference's: every second keypoint
    has one landmark 2.5 px away with 3 flipped bits -> 20 matches (feature 2k, track 500 + 2k)."""
    _build_lm()
    p = subprocess.run([EXE_LM], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout[-500:] + p.stderr[-500:]
    lines = p.stdout.splitlines()
    assert lines[0] == "matches 20"
    got = [tuple(int(v) for v in ln.split()[1:]) for ln in lines[1:]]
    assert got == [(j, 500 + j) for j in range(0, 40, 2)]


EXE_GEO = os.path.join(ROOT, "tests", "cpp", "test_dropin_geometry")


def _build_geo():
    cmd = ["g++", "-std=c++14", "-O2", "-I", os.path.join(ROOT, "tests", "cpp", "stubs"), "-I", os.path.join(ROOT, "include"),
           "-o", EXE_GEO, os.path.join(ROOT, "tests", "cpp", "test_dropin_geometry.cpp"), "-L",
           os.path.join(ROOT, "visual-slam_b200", "csrc"), "-lvisnav_b200",
           "-Wl,-rpath," + os.path.join(ROOT, "visual-slam_b200", "csrc")]
    subprocess.run(cmd, check=True)


def test_geometry_dropin_compiles_and_fails_loudly_without_gpu():
    import torch
    _build_geo()
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([EXE_GEO], capture_output=True, text=True)
    assert p.returncode == 3 and "no usable sm_100 GPU" in p.stdout


@pytest.mark.gpu
def test_geometry_dropin_header():
    """findInliersEssential_b200 / project_landmarks_b200 on a pinhole rig with a pure x baseline: partners in the
    same row are inliers, every third (10 rows off) is not; landmarks behind the camera or outside the image
    are dropped and the rest project to fx*x/z+cx, fy*y/z+cy."""
    _build_geo()
    p = subprocess.run([EXE_GEO], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout[-500:] + p.stderr[-500:]
    lines = p.stdout.splitlines()
    assert lines[0] == "inliers 20"
    assert [tuple(int(v) for v in ln.split()[1:]) for ln in lines if ln.startswith("i ")] == \
        [(i, i) for i in range(30) if i % 3 != 2]
    proj = [ln.split()[1:] for ln in lines if ln.startswith("p ")]
    want = [(700 + j, 400 * (-3.0 + 0.5 * j) / (2.0 + j) + 376, 400 * (0.1 * j) / (2.0 + j) + 240) for j in range(2, 12)]
    want = [w for w in want if 0 <= w[1] <= 752 and 0 <= w[2] <= 480]
    assert [int(r[0]) for r in proj] == [w[0] for w in want]
    assert np.abs(np.array([[float(r[1]), float(r[2])] for r in proj]) - np.array([w[1:] for w in want])).max() < 1e-8


EXE_STORE = os.path.join(ROOT, "tests", "cpp", "test_dropin_store")


def _build_store():
    cmd = ["g++", "-std=c++14", "-O2", "-I", os.path.join(ROOT, "tests", "cpp", "stubs"), "-I", os.path.join(ROOT, "include"),
           "-o", EXE_STORE, os.path.join(ROOT, "tests", "cpp", "test_dropin_store.cpp"), "-L",
           os.path.join(ROOT, "visual-slam_b200", "csrc"), "-lvisnav_b200",
           "-Wl,-rpath," + os.path.join(ROOT, "visual-slam_b200", "csrc")]
    subprocess.run(cmd, check=True)


def test_keyframe_store_header_compiles_and_fails_loudly_without_gpu():
    import torch
    _build_store()
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([EXE_STORE], capture_output=True, text=True)
    assert p.returncode == 3 and "no usable sm_100 GPU" in p.stdout


@pytest.mark.gpu
def test_keyframe_store_header():
    """KeyframeStore (include/visnav/keyframe_store_b200.h): keyframes appended one by one (ragged, one empty), the
    newest matched against all earlier ones in one call == matchDescriptors on the same pairs."""
    _build_store()
    p = subprocess.run([EXE_STORE], capture_output=True, text=True)
    assert p.returncode == 0 and p.stdout.startswith("ok 7 jobs"), p.stdout[-400:] + p.stderr[-400:]
    assert int(p.stdout.split()[3]) > 1000


EXE_BOW = os.path.join(ROOT, "tests", "cpp", "test_dropin_bow")


def _build_bow():
    cmd = ["g++", "-std=c++14", "-O2", "-I", os.path.join(ROOT, "tests", "cpp", "stubs"), "-I", os.path.join(ROOT, "include"),
           "-o", EXE_BOW, os.path.join(ROOT, "tests", "cpp", "test_dropin_bow.cpp"), "-L",
           os.path.join(ROOT, "visual-slam_b200", "csrc"), "-lvisnav_b200",
           "-Wl,-rpath," + os.path.join(ROOT, "visual-slam_b200", "csrc")]
    subprocess.run(cmd, check=True)


def test_bow_header_compiles_and_fails_loudly_without_gpu():
    import torch
    _build_bow()
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([EXE_BOW], capture_output=True, text=True)
    assert p.returncode == 3 and "no usable sm_100 GPU" in p.stdout


@pytest.mark.gpu
def test_bow_header(tmp_path):
    """include/visnav/bow_b200.h driven from C++ (Vocabulary::loadFromTextFile, compute_bow_vector, score,
    BowDatabase) == the CPU oracles: cv::ORB restatement -> DBoW2 restatement on the same images."""
    from oracle import oracle as oo
    from visnav_b200 import workloads as wl
    _build_bow()
    voc = wl.synth_vocabulary(8, 3, 13, stop=0.05)
    vpath = os.path.join(str(tmp_path), "voc.txt")
    wl.write_vocabulary_text(vpath, voc, trailing_newline=True)
    imgs = np.stack([wl.synth_image(700 + i, 320, 240) for i in range(3)] + [np.roll(wl.synth_image(700, 320, 240), -9, 1)])
    ipath = os.path.join(str(tmp_path), "imgs.bin")
    imgs.tofile(ipath)
    p = subprocess.run([EXE_BOW, vpath, "320", "240", str(len(imgs)), ipath], capture_output=True, text=True)
    assert p.returncode == 0, p.stdout[-400:] + p.stderr[-400:]
    lines = p.stdout.strip().split("\n")
    assert lines[0] == "voc 8 3 512"
    o = oo.BowVocab.from_text(vpath)
    want = []
    for i, img in enumerate(imgs):
        kp, desc = oo.orb_detect_compute(img, 1500)
        ids, vals, fw, fn, fwt = o.transform(desc, 4)
        want.append((ids, vals))
        tag, nw, total, nnodes, listed = lines[1 + i].split()
        assert tag == "v" and int(nw) == len(ids) and int(listed) == int((fwt > 0).sum())
        assert int(nnodes) == len(set(fn[fwt > 0]))        # levelsup 4 > L = 3: every feature sits under the root (node 0)
        s = 0.0
        for x in vals:
            s += x
        assert float(total) == s
    n = len(imgs)
    scores = np.array([[o.score(a, b) for b in want] for a in want])
    got = np.array([[float(x) for x in lines[1 + n + i].split()[1:]] for i in range(n)])
    assert np.array_equal(got.view(np.uint64), scores.view(np.uint64))
    for i in range(n):
        q = np.array([float(x) for x in lines[1 + 2 * n + 2 * i].split()[1:]])
        c = [int(x) for x in lines[2 + 2 * n + 2 * i].split()[1:]]
        assert np.array_equal(q.view(np.uint64), scores[i].view(np.uint64))
        assert c == [len(np.intersect1d(want[i][0], b[0])) for b in want]
    assert scores[0, 3] > scores[0, 1] and scores[0, 3] > scores[0, 2]     # the shifted view of image 0 scores highest
