N=${1:-4}
B="--no-train --no-config5 --no-e2e --no-cpu-baseline --no-parity"
run() { tag=$1; shift; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 10 $B "$@" > gpurun_out/s3_n${N}_$tag.log 2>&1; python - gpurun_out/s3_n${N}_$tag.log <<'PY'
import sys,json
ok=False
for l in open(sys.argv[1]):
    if l.startswith('{"metric"'):
        d=json.loads(l); ok=True; print(sys.argv[1], "ms/step", round(d["ms_per_step"],4), "kernel", round(d["roofline"]["kernel_ms_avg"],4))
if not ok: print(sys.argv[1], "NO LINE"); print(open(sys.argv[1]).read()[-800:])
PY
}
B200_SHARD_DEBUG=1 run dbg_nopush --step-mode sharded
B200_SHARD_DEBUG=2 run dbg_nopull --step-mode sharded
B200_SHARD_DEBUG=3 run dbg_local --step-mode sharded
run cta1 --step-mode sharded --ctas-per-sm 1 --unroll 1
