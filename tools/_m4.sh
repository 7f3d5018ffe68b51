mkdir -p gpurun_out/m4
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -k "pool or head or addend or engine" 2>&1 | tail -6 > gpurun_out/m4/pytest.log; tail -6 gpurun_out/m4/pytest.log
timeout 300 python bench.py --gpus 1 --steps 40 --warmup 5 > gpurun_out/m4/bench_n1.json 2> gpurun_out/m4/e1.log; tail -c 200 gpurun_out/m4/e1.log
timeout 300 python bench.py --gpus 1 --steps 10 --warmup 3 --model res50 --image-size 224 --per-gpu-batch 32 > gpurun_out/m4/bench_res50_224_n1.json 2> gpurun_out/m4/e2.log; tail -c 200 gpurun_out/m4/e2.log
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 --model res50 --per-gpu-batch 256 > gpurun_out/m4/bench_res50_32_n1.json 2> gpurun_out/m4/e3.log; tail -c 200 gpurun_out/m4/e3.log
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 3 --model res34 --per-gpu-batch 256 > gpurun_out/m4/bench_res34_32_n1.json 2> gpurun_out/m4/e4.log; tail -c 200 gpurun_out/m4/e4.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/m4/bench*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['launches_per_step'], d['achieved_tflops_per_gpu'])
    except Exception as e: print(f,'ERR',e)
PY
