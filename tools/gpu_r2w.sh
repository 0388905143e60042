cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
FUZZ_SECONDS=200 FUZZ_SEED=7 timeout 400 python tools/fuzz_parity.py > gpurun_out/r2w_fuzz.json 2> gpurun_out/r2w_fuzz.err; echo "fuzz rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/fuzz.json')); print({k:v for k,v in d.items()})"
timeout 600 python tools/parity_report.py > gpurun_out/r2w_parity.log 2>&1; echo "parity rc=$?"; tail -5 gpurun_out/r2w_parity.log; ls gpurun_out | grep parity
