mkdir -p gpurun_out/p4
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521"
timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/p4/bench_n2.json 2> gpurun_out/p4/e2.log; tail -c 300 gpurun_out/p4/e2.log
python - <<'PY'
import json
d=json.loads(open('gpurun_out/p4/bench_n2.json').read().strip().splitlines()[-1]); print('N2', d['ms_per_step'], d['e2e']['ms_per_step'], d['exposed_comm_ms'], d['loss_after_bench'])
PY
