set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2g_smoke.log 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2g_pytest.log
python bench.py --steps 5 > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2g_bench_ref.json 2> gpurun_out/r2g_bench_ref.err
timeout 600 python tools/dbench.py 262144 32768 3 > gpurun_out/r2g_dbench.jsonl 2> gpurun_out/r2g_dbench.err
DBENCH_ONLY=3,32,0,1 ncu --set full --clock-control none --import-source on -k regex:lz4e_decompress_wave -c 6 -f -o gpurun_out/r2g_wave python tools/dbench.py 262144 32768 1 > gpurun_out/r2g_ncu.log 2>&1
tail -3 gpurun_out/r2g_pytest.log
cat gpurun_out/r2g_bench.json
cat gpurun_out/r2g_dbench.jsonl
