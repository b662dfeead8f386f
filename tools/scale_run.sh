# bench.py at N = 1, 2, 4, 8 on one 8-GPU box, one JSON line per N into gpurun_out/scale.jsonl; then the 2-GPU tests
set -x
: > gpurun_out/scale.jsonl
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu --no-configs 2>gpurun_out/scale_n1.err | tail -1 >> gpurun_out/scale.jsonl
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 20 --warmup 3 2>gpurun_out/scale_n$n.err | tail -1 >> gpurun_out/scale.jsonl
done
timeout 600 python -m pytest tests/test_gpu_multi.py -v 2>&1 | tail -15 > gpurun_out/multi_gpu_tests.log
python - <<'PY'
import json
for l in open("gpurun_out/scale.jsonl"):
    j = json.loads(l)
    print(j["n_gpus"], j["ms_per_step"], j["value"], j["parity_check"]["max_rel"], j.get("per_rank"))
PY
