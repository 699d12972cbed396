#!/bin/bash
# Round-2 evidence run (one gpurun call): GPU test-suite, smoke, both bench arms, ncu launch list + one full capture of
# k_encode_frame at 1080p.  Everything lands in gpurun_out/ and is copied to profiles/ by hand afterwards.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -6 > gpurun_out/r2_pytest_gpu.log
python __graft_entry__.py smoke 2>&1 | tail -2 >> gpurun_out/r2_pytest_gpu.log
( time python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err ) 2>> gpurun_out/r2_bench_n1.err
( time python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err ) 2>> gpurun_out/r2_bench_ref.err
# launch list of one encoder process of the bench's workload (4 frames of 1080p LDB medium), then the full capture of the
# third k_encode_frame launch (a P frame)
python profiles/r2/enc_cmd.py LDB_medium_complexity 1920 1080 4 32 > gpurun_out/enc_cmd.txt
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_enc1080.csv $(cat gpurun_out/enc_cmd.txt) > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_encode_frame -s 2 -c 1 -f -o gpurun_out/r2_enc1080 $(cat gpurun_out/enc_cmd.txt) > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -12
