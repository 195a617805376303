#!/bin/bash
# Round-2 evidence on one GPU: bench line, ncu launch list of the same command, one `--set full` capture of the contraction
# kernel, the rerun stress.  Run on the GPU box: bash scripts/collect_r02.sh ; results land in gpurun_out/.
set -u
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference.json 2> gpurun_out/r02_bench_reference.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-extras --no-parity > gpurun_out/r02_ncu_launches.log 2>&1
python scripts/launch_summary.py gpurun_out/r02_launches.csv 3 > gpurun_out/r02_launches_summary.txt 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 220 -c 6 -f -o gpurun_out/r02_gemm_full \
  python bench.py --steps 2 --warmup 3 --no-extras --no-parity > gpurun_out/r02_ncu_full.log 2>&1
python scripts/ncu_summary.py gpurun_out/r02_gemm_full.ncu-rep 25 > gpurun_out/r02_gemm_full_summary.txt 2>&1
timeout 900 python scripts/stress_race.py > gpurun_out/r02_stress.txt 2>&1
cuobjdump -sass rten_fork_b200/librten_gpu.so 2>/dev/null | grep -E "UTCHMMA|UTCQMMA|UTMALDG|UTMASTG|UTCBAR|LDTM|STTM|UBLKCP" | awk '{print $2}' | sed 's/\..*//' | sort | uniq -c > gpurun_out/r02_sass_mnemonics.txt
rm -f gpurun_out/r02_gemm_full.ncu-rep
ls -la gpurun_out
