#!/bin/bash
# one GPU-box visit: parity tests, bench (both arms), kernel timing sets, ncu launch list + full captures
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 200 python tools/quick_bench.py basic > gpurun_out/quick_bench.log 2>&1; echo "quick rc=$?"
timeout 100 python tools/quick_bench.py layer >> gpurun_out/quick_bench.log 2>&1
timeout 200 python tools/quick_bench.py cols > gpurun_out/cols.log 2>&1; echo "cols rc=$?"
for t in q8_0 q4_0 q5_K; do timeout 100 python tools/quick_bench.py cols $t >> gpurun_out/cols.log 2>&1; done
bash tools/mmq_set.sh > gpurun_out/mmq_set.log 2>&1; echo "mmq set rc=$?"
timeout 200 python tools/pp512_bench.py q8_0 512 > gpurun_out/pp512.log 2>&1
timeout 200 python tools/pp512_bench.py q4_k_m 512 >> gpurun_out/pp512.log 2>&1
timeout 200 python tools/moe_bench.py >> gpurun_out/pp512.log 2>&1; echo "pp512/moe rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gemv -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:gemv_kq -s 130 -c 8 -f -o gpurun_out/gemv_full \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:mmq_kernel -s 2 -c 1 -f -o gpurun_out/mmq_full \
    python tools/mmq_one.py q8_0 14336 4096 512 3 > gpurun_out/ncu_mmq.log 2>&1; echo "ncu mmq rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemv_mma -s 4 -c 1 -f -o gpurun_out/mma_full \
    python tools/quick_bench.py one q4_K 128256 4096 8 > gpurun_out/ncu_mma.log 2>&1; echo "ncu mma rc=$?"
nvidia-smi > gpurun_out/nvidia_smi.txt; nproc >> gpurun_out/nvidia_smi.txt; lscpu | head -20 >> gpurun_out/nvidia_smi.txt
