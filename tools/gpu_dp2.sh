#!/bin/bash
# 2-GPU check: NCCL equivalence test + the 2-rank bench line
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi -L | head -4
timeout 900 python -m pytest tests/test_gpu_dp2.py -m gpu -x -q > gpurun_out/r50_dp2.log 2>&1; echo "dp2 tests exit $?"; tail -3 gpurun_out/r50_dp2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/r50_bench_n2.json 2> gpurun_out/r50_bench_n2.err
echo "bench n2 exit $?"; tail -1 gpurun_out/r50_bench_n2.json | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > gpurun_out/r50_bench_ref_n2.json 2> gpurun_out/r50_bench_ref_n2.err
echo "ref n2 exit $?"; tail -1 gpurun_out/r50_bench_ref_n2.json | cut -c1-300
