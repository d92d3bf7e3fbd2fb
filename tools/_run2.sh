set -x
timeout 300 python -m pytest tests/test_gpu_shard.py tests/test_gpu_lamb.py -x -q 2>&1 | tail -8
timeout 300 python -m pytest tests/test_gpu_peer.py -x -q -k "sharded" 2>&1 | tail -8
B="--no-train --no-config5 --no-e2e --no-cpu-baseline"
for mode in sharded spec; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 10 --step-mode $mode $B > gpurun_out/s3_n2_$mode.log 2>&1; tail -c 2500 gpurun_out/s3_n2_$mode.log
done
for u in 1; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 50 --warmup 10 --step-mode sharded --unroll $u $B > gpurun_out/s3_n2_sharded_u$u.log 2>&1; tail -c 700 gpurun_out/s3_n2_sharded_u$u.log
done
