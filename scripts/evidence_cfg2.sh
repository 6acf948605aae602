set -x
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 600 python -m pytest tests -m gpu -v 2>&1 | grep -v "^$" > gpurun_out/r2_pytest_gpu.log; tail -1 gpurun_out/r2_pytest_gpu.log
K='regex:search_pair|rowprep|group_list|refine_cand|accumulate'
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k "$K" -o /tmp/cfg2 python scripts/ncu_capture.py 1000000 784 10 euclidean > gpurun_out/ncu_cfg2.log 2>&1; tail -2 gpurun_out/ncu_cfg2.log
ncu -i /tmp/cfg2.ncu-rep --page raw --csv > gpurun_out/r2_ncu_full_cfg2_raw.csv
ls -la gpurun_out/r2_ncu_full_cfg2_raw.csv
