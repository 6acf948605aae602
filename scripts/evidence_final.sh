# One call: tests, ncu --set full captures of the four shapes, traffic.json, default bench line (all into gpurun_out/).
timeout 300 python -m pytest tests -m gpu -v 2>&1 | grep -v "^$" > gpurun_out/r2_pytest_gpu.log
tail -1 gpurun_out/r2_pytest_gpu.log
grep -q " failed" gpurun_out/r2_pytest_gpu.log && { grep -n "FAILED\|Error" gpurun_out/r2_pytest_gpu.log | head -5; exit 1; }
K='regex:search_pair|rowprep|group_list|refine_cand|accumulate'
for cfg in "4 100000000 128 1000 euclidean" "5 20000000 256 500 cosine" "3 10000000 64 100 euclidean" "2 1000000 784 10 euclidean"; do
  set -- $cfg
  timeout 400 ncu --set full --clock-control none --import-source on --profile-from-start off -k "$K" -o /tmp/cfg$1 python scripts/ncu_capture.py $2 $3 $4 $5 > gpurun_out/ncu_cfg$1.log 2>&1
  ncu -i /tmp/cfg$1.ncu-rep --page raw --csv > gpurun_out/r2_ncu_full_cfg$1_raw.csv
  cp gpurun_out/r2_ncu_full_cfg$1_raw.csv profiles/
done
python scripts/ncu_traffic.py profiles/r2_ncu_full_cfg4_raw.csv:100000000:128:1000 profiles/r2_ncu_full_cfg5_raw.csv:20000000:256:500 profiles/r2_ncu_full_cfg3_raw.csv:10000000:64:100 profiles/r2_ncu_full_cfg2_raw.csv:1000000:784:10 > /dev/null
cp profiles/traffic.json gpurun_out/traffic.json
timeout 300 python bench.py > gpurun_out/r2_bench_100M_1gpu.json 2> gpurun_out/bench.err
tail -c 150 gpurun_out/bench.err; ls -la gpurun_out/r2_ncu_full_cfg*_raw.csv gpurun_out/traffic.json gpurun_out/r2_bench_100M_1gpu.json
