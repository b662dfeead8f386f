# Split of a PSO step on N GPUs (N = first argument): step time and, from three runs, the device time of each of its
# three launches (the engine's own events around one stage per run).  Output: one line per stage.
N=${1:-1}
for st in 1 2 3; do
  if [ "$N" = 1 ]; then
    python bench.py --steps 20 --warmup 3 --no-cpu --no-configs --draws ${2:-1000000} --stage $st 2>/dev/null | tail -1
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + st)) \
        bench.py --gpus $N --steps 20 --warmup 3 --stage $st 2>/dev/null | tail -1
  fi | python -c "
import sys, json
j = json.loads(sys.stdin.read())
pr = j.get('per_rank') or {}
print(j['n_gpus'], 'GPU(s)  step %.4f ms  %-26s %.4f ms' % (j['ms_per_step'], j['stage_timed'], j['roofline']['pso_step']['kernel_ms']), ' per rank:', [round(x, 4) for x in pr.get('kernel_ms', [])])"
done
