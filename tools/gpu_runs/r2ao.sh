set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2ao_smoke.log 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2ao_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2ao_pytest.log
( time python bench.py > gpurun_out/r2ao_bench.json 2> gpurun_out/r2ao_bench.err ) 2> gpurun_out/r2ao_time.log
( time python bench.py --impl reference > gpurun_out/r2ao_bench_ref.json 2> gpurun_out/r2ao_bench_ref.err ) 2>> gpurun_out/r2ao_time.log
