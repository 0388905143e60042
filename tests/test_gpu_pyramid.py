"""Multi-scale extension (BASELINE config 5): device pyramid + per-level detect/describe against the
oracle applied to the numpy pyramid of the same image (workloads.pyramid_levels: integer 2x2 mean,
bit-exact).  There is no reference behaviour for the pyramid itself; per level
the bar is the usual one: corners ordered-equal, angles <= 1e-5 rad, descriptors bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _check(ctx, oracle_port, imgs, caps):
    from visnav_b200 import workloads as wl
    xy, ang, desc, nkp, off = ctx.pyramid_detect_describe_batch(imgs, caps)
    for i, img in enumerate(imgs):
        levels = wl.pyramid_levels(img, len(caps))
        for k, lv in enumerate(levels):
            n = int(nkp[i, k])
            if lv.shape[0] < 39 or lv.shape[1] < 39:
                assert n == 0
                continue
            wxy, wang, wdesc = oracle_port.detect_describe(np.ascontiguousarray(lv), int(caps[k]))
            o = int(off[k])
            assert n == len(wxy), (i, k, n, len(wxy))
            assert np.array_equal(xy[i, o:o + n], wxy), (i, k)
            assert n == 0 or np.abs(ang[i, o:o + n] - wang).max() <= 1e-5
            assert np.array_equal(desc[i, o:o + n], wdesc), (i, k)


def test_pyramid_small_odd_sizes(ctx, oracle_port):
    from visnav_b200 import workloads as wl
    imgs = np.stack([wl.synth_image(40 + s, 333, 201) for s in range(3)])
    _check(ctx, oracle_port, imgs, [400, 120, 40])


def test_pyramid_config5_shape(ctx, oracle_port):
    from visnav_b200 import workloads as wl
    img = wl.synth_image(77, 1920, 1080)
    _check(ctx, oracle_port, img[None], [3012, 753, 188, 47])


def test_pyramid_single_level_equals_batch(ctx):
    from visnav_b200 import workloads as wl
    imgs = np.stack([wl.synth_image(3 + s, 320, 240) for s in range(2)])
    xy, ang, desc, nkp, _ = ctx.pyramid_detect_describe_batch(imgs, [300])
    bxy, bang, bdesc, bn = ctx.detect_describe_batch(imgs, 300)
    assert np.array_equal(nkp[:, 0], bn)
    for i in range(2):
        n = int(bn[i])
        assert np.array_equal(xy[i, :n], bxy[i, :n]) and np.array_equal(desc[i, :n], bdesc[i, :n])
        assert np.array_equal(ang[i, :n], bang[i, :n])
