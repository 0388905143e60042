cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log
tail -8 gpurun_out/r2c_pytest.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; tail -3 gpurun_out/r2c_bench.err
python -c "
import json; b=json.load(open('gpurun_out/r2c_bench.json')); print(b['value'], b['e2e']['value'], b['stage_ms'], b['roofline_match']['distances_per_s'])"
timeout 300 python bench.py --no-cpu-baseline --steps 2 --warmup 3 --e2e-steps 1 > gpurun_out/r2c_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r2c_launches.csv python bench.py --no-cpu-baseline --steps 2 --warmup 3 --e2e-steps 1 > gpurun_out/r2c_ncu.log 2>&1
tail -2 gpurun_out/r2c_ncu.log
