cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2o_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2o_pytest.log
tail -6 gpurun_out/r2o_pytest.log
timeout 200 python tools/probes/single_pair.py > gpurun_out/r2o_single_pair.json 2> gpurun_out/r2o_single_pair.err; cat gpurun_out/r2o_single_pair.json; tail -2 gpurun_out/r2o_single_pair.err
timeout 400 python bench.py > gpurun_out/r2o_bench.json 2> gpurun_out/r2o_bench.err; tail -3 gpurun_out/r2o_bench.err
python -c "
import json; b=json.load(open('gpurun_out/r2o_bench.json')); print(b['value'], b['e2e']['value'], b['stage_ms']); print(json.dumps(b['configs'], indent=1)); print(b['roofline_match']['frac'], b['roofline_describe']['frac'], b['cpu_baseline'])"
