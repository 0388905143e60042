# usage: bash tools/gpu_resp_iter.sh <tag>   -- response-kernel iteration: detection parity tests + bench line
cd $GRAFT_REPO_ROOT
T=$1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_detect_describe.py tests/test_gpu_golden.py -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -4 gpurun_out/${T}_pytest.log
timeout 300 python bench.py --no-cpu-baseline --no-extras > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; tail -3 gpurun_out/${T}_bench.err
python -c "
import json; b=json.load(open('gpurun_out/${T}_bench.json')); print(b['value'], b['e2e']['value'], b['stage_ms'], b['roofline']['frac'])"
