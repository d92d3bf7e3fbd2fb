N=${1:-4}
timeout 400 python -m pytest tests/test_gpu_peer.py -x -q -k "sharded" 2>&1 | tail -4
B="--no-train --no-config5 --no-e2e --no-cpu-baseline"
run() { tag=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 50 --warmup 10 $B "$@" > gpurun_out/s3_n${N}_$tag.log 2>&1; python - gpurun_out/s3_n${N}_$tag.log <<'PY'
import sys,json
ok=False
for l in open(sys.argv[1]):
    if l.startswith('{"metric"'):
        d=json.loads(l); ok=True; print(sys.argv[1], "ms/step", round(d["ms_per_step"],4), "kernel", round(d["roofline"]["kernel_ms_avg"],4), "parity", d.get("parity",{}).get("all_ranks_ok"), d.get("error"))
if not ok: print(sys.argv[1], "NO LINE"); print(open(sys.argv[1]).read()[-1500:])
PY
}
run sharded --step-mode sharded
run sharded_u1 --step-mode sharded --unroll 1
run sharded_t512 --step-mode sharded --threads 512 --unroll 1
run spec --step-mode spec
