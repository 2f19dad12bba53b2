set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
python bench.py --steps 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2a_bench_ref.json 2> gpurun_out/r2a_bench_ref.err
LZ4E_B200_GIN_MIN=1024 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2a_ginmin1024.json 2>&1
LZ4E_B200_DGROUP=4 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2a_dgroup4.json 2>&1
LZ4E_B200_DGROUP=8 LZ4E_B200_DSTAGE=1 python bench.py --steps 5 --no-e2e --no-cpu --no-replay > gpurun_out/r2a_dgroup8_stage.json 2>&1
tail -3 gpurun_out/r2a_pytest.log
