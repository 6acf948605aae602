set -x
timeout 600 python -m pytest tests -m gpu -v 2>&1 | grep -v "^$" > gpurun_out/r2_pytest_gpu.log; tail -2 gpurun_out/r2_pytest_gpu.log
K='regex:search_pair|rowprep|group_list|refine_cand|accumulate'
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k "$K" -o /tmp/cfg4 python scripts/ncu_capture.py 100000000 128 1000 euclidean > gpurun_out/ncu_cfg4.log 2>&1; tail -2 gpurun_out/ncu_cfg4.log
ncu -i /tmp/cfg4.ncu-rep --page raw --csv > gpurun_out/r2_ncu_full_cfg4_raw.csv
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k "$K" -o /tmp/cfg5 python scripts/ncu_capture.py 20000000 256 500 cosine > gpurun_out/ncu_cfg5.log 2>&1; tail -2 gpurun_out/ncu_cfg5.log
ncu -i /tmp/cfg5.ncu-rep --page raw --csv > gpurun_out/r2_ncu_full_cfg5_raw.csv
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k "$K" -o /tmp/cfg3 python scripts/ncu_capture.py 10000000 64 100 euclidean > gpurun_out/ncu_cfg3.log 2>&1; tail -2 gpurun_out/ncu_cfg3.log
ncu -i /tmp/cfg3.ncu-rep --page raw --csv > gpurun_out/r2_ncu_full_cfg3_raw.csv
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_ncu_launches_100M.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-configs --upload-gb 0 > gpurun_out/ncu_launches.log 2>&1; tail -c 200 gpurun_out/ncu_launches.log
ls -la gpurun_out/*.csv
