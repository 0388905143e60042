# usage: bash tools/gpu_final_n8.sh <tag>   (gpurun --gpus 8)  -- the bench line at N = 8, torchrun as the driver launches it
cd $GRAFT_REPO_ROOT
T=$1
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/${T}_bench_n8.json 2> gpurun_out/${T}_bench_n8.err; echo "rc=$?"; tail -2 gpurun_out/${T}_bench_n8.err
python -c "
import json; b=json.load(open('gpurun_out/${T}_bench_n8.json')); print(b['n_gpus'], b['value'], b['e2e']['value'], b['stage_ms']); print(b['configs']['c5_single_scale']['pairs_per_s'], b['configs']['c3_sample']['distances_per_s'])"
