# usage: bash tools/gpu_final_n1.sh <tag>  -- end-of-round evidence at N = 1: full GPU suite, smoke, bench (both arms), single pair, parity report
cd $GRAFT_REPO_ROOT
T=$1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -3 gpurun_out/${T}_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
timeout 600 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/${T}_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_reference.json 2> gpurun_out/${T}_reference.err; echo "ref rc=$?"
timeout 200 python tools/probes/single_pair.py > gpurun_out/${T}_single_pair.json 2> gpurun_out/${T}_single_pair.err; cat gpurun_out/${T}_single_pair.json
timeout 600 python tools/parity_report.py > gpurun_out/${T}_parity.json 2> gpurun_out/${T}_parity.err; tail -c 600 gpurun_out/${T}_parity.json
python -c "
import json; b=json.load(open('gpurun_out/${T}_bench.json')); print(b['value'], b['e2e']['value'], b['stage_ms'], b['roofline']['frac'], b['roofline_match']['frac'], b['roofline_match']['frac_executed'], b.get('cpu_baseline',{}).get('value')); print({k:(v.get('distances_per_s') or v.get('pairs_per_s') or v.get('ms')) for k,v in b['configs'].items()})
r=json.load(open('gpurun_out/${T}_reference.json')); print(r.get('value'), r.get('cpu_baseline'))"
