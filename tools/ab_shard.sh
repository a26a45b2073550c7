run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $1 --steps 5 --warmup 2 --no-cpu --no-e2e --no-queries $2 2>gpurun_out/err_$3.txt | tail -1 > gpurun_out/$3.json; python - <<PY
import json
d=json.loads(open("gpurun_out/$3.json").read())
print("$3", d["ms_per_step"], d.get("verify",{}) and d["verify"].get("match"), d["shard"]["stages_ms"] if d.get("shard") else None, d["roofline"]["stages_ms"])
PY
}
