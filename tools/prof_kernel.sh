# ncu --set full capture of one kernel of the c2 training step: tools/prof_kernel.sh <kernel regex> <out name> [workload]
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$1 -s 3 -c 1 -o gpurun_out/$2 python bench.py --workload ${3:-c2} --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$2.log 2>&1; tail -1 gpurun_out/ncu_$2.log | cut -c1-100
