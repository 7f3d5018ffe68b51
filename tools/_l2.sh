mkdir -p gpurun_out/l2
timeout 900 python tools/multi_gpu_checks.py --gpus 2 --quick --out gpurun_out/l2/multi2 > gpurun_out/l2/multi2.log 2>&1; tail -4 gpurun_out/l2/multi2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522"
timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/l2/bench_n2.json 2> gpurun_out/l2/e2.log; tail -c 300 gpurun_out/l2/e2.log
PMD_TAIL_LL=0 timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/l2/bench_n2_noll.json 2> gpurun_out/l2/e3.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/l2/bench*.json')):
    d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['ms_per_step'], d['e2e']['ms_per_step'], d['exposed_comm'], d['impl_detail']['grad_allreduce'], d['launches_per_step'])
s=json.load(open('gpurun_out/l2/multi2/summary.json')); print(s['ok'], {k:v['weights_bitwise_equal_across_ranks'] for k,v in s['algos'].items()}, {k:v['loss_matches_single_rank'] for k,v in s['algos'].items()})
PY
