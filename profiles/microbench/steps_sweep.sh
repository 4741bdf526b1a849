# does the per-step time depend on the length of the timed region?  (N = 1 and N = 2 ranks)
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
p=29600
for k in 20 50 100 300; do for rep in 1 2; do
  python bench.py --steps $k --warmup 5 --no-cpu-baseline --e2e-steps 3 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('N=1 steps', d['steps'], round(d['ms_per_step']*1e3,2))"
  p=$((p+1)); $TR --master-port $p bench.py --gpus 2 --steps $k --warmup 5 --no-cpu-baseline --e2e-steps 3 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('N=2 steps', d['steps'], round(d['ms_per_step']*1e3,2))"
done; done
