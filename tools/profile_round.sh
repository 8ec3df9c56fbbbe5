set -x
python bench.py --workload viterbi --reads 65536 --steps 2 > gpurun_out/plain_v.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:viterbi_stream -s 3 -c 1 -o gpurun_out/r02_viterbi python bench.py --workload viterbi --reads 65536 --steps 2 > gpurun_out/ncu_v.log 2>&1
python tools/quick_duplex.py 9472 500 32 0 1 > gpurun_out/plain_d.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:duplex_packed -c 1 -o gpurun_out/r02_duplex python tools/quick_duplex.py 9472 500 32 0 1 > gpurun_out/ncu_d.log 2>&1
python tools/quick_beam.py 3552 2000 32 0.0 > gpurun_out/plain_b32.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:beam_search_kernel -c 1 -o gpurun_out/r02_beam32 python tools/quick_beam.py 3552 2000 32 0.0 > gpurun_out/ncu_b32.log 2>&1
python tools/quick_beam.py 18944 2000 5 0.1 > gpurun_out/plain_b5.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:beam_small_kernel -c 1 -o gpurun_out/r02_beam5 python tools/quick_beam.py 18944 2000 5 0.1 > gpurun_out/ncu_b5.log 2>&1
python bench.py --steps 2 --warmup 1 > gpurun_out/plain_all.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r02.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_all.log 2>&1
for f in gpurun_out/ncu_*.log; do tail -n 2 "$f"; done
