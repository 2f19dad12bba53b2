set -x
DBENCH_ONLY=3,32,0,0 python tools/dbench.py 262144 32768 5 > gpurun_out/r2e_wave.jsonl 2>&1
DBENCH_ONLY=3,32,0,1 python tools/dbench.py 262144 32768 5 >> gpurun_out/r2e_wave.jsonl 2>&1
LZ4E_B200_DVARIANT=3 timeout 900 python -m pytest tests -m gpu -x -q -k "decompress or roundtrip or config or malformed or stress" > gpurun_out/r2e_pytest.log 2>&1
tail -3 gpurun_out/r2e_pytest.log
cat gpurun_out/r2e_wave.jsonl
