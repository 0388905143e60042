cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_match.py tests/test_gpu_match_mma.py tests/test_sharding_gloo.py -x -q -m gpu > gpurun_out/r2t_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2t_pytest.log
tail -4 gpurun_out/r2t_pytest.log
python - <<'PY'
import sys, json; sys.path.insert(0, '.')
import numpy as np, visnav_b200
from visnav_b200 import workloads as wl
from visnav_b200.keypoints import VSLAM_OPT_MATCH_MMA
ctx = visnav_b200.Context(0)
q, db = wl.reloc_query()
out = {}
for mode in (1, 0):
    ctx.set_option(VSLAM_OPT_MATCH_MMA, mode)
    best = 1e9
    for _ in range(5):
        m = ctx.match(q, db, 70, 1.2); best = min(best, ctx.last_match_ms())
    out["c4_mma%d_ms" % mode] = best
    out["c4_mma%d_dist_per_s" % mode] = 3e8 / (best * 1e-3)
print(json.dumps(out)); print(ctx.mma_timeouts())
PY
