cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 300 tools/probes/bin/copy_probe 4 > gpurun_out/r2p_copy_n8.jsonl 2> gpurun_out/r2p_copy_n8.err; cat gpurun_out/r2p_copy_n8.jsonl
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2p_bench_n8.json 2> gpurun_out/r2p_bench_n8.err; tail -2 gpurun_out/r2p_bench_n8.err
python -c "
import json; b=json.load(open('gpurun_out/r2p_bench_n8.json')); print(b['n_gpus'], b['value'], b['e2e']['value'], b['stage_ms']); print(b['configs']['c5_single_scale']['pairs_per_s'], b['configs']['c3_sample']['distances_per_s'])"
