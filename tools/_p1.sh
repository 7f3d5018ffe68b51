mkdir -p gpurun_out/p1
A18="bench.py --no-graph --steps 1 --warmup 3 --no-exposed-comm"
A50="bench.py --no-graph --steps 1 --warmup 3 --no-exposed-comm --model res50 --image-size 224 --per-gpu-batch 32"
python $A18 > gpurun_out/p1/plain18.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 420 --csv --log-file gpurun_out/p1/launches18.csv python $A18 > gpurun_out/p1/ncu18.log 2>&1
python $A50 > gpurun_out/p1/plain50.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 560 -c 700 --csv --log-file gpurun_out/p1/launches50.csv python $A50 > gpurun_out/p1/ncu50.log 2>&1
python $A18 > gpurun_out/p1/plain18b.log 2>&1 && ncu --set full --clock-control none --import-source on -s 192 -c 64 -o gpurun_out/p1/step18 python $A18 > gpurun_out/p1/ncu18full.log 2>&1
ls -la gpurun_out/p1; tail -2 gpurun_out/p1/ncu50.log
ncu -i gpurun_out/p1/step18.ncu-rep --page raw --csv > gpurun_out/p1/step18_raw.csv 2>/dev/null; wc -c gpurun_out/p1/step18_raw.csv; rm -f gpurun_out/p1/step18.ncu-rep
