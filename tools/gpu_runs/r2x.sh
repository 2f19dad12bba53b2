set -x
for c in 2 3; do
CLASS_ONLY=$c ncu --set full --clock-control none --import-source on -k regex:lz4e_compress_kernel -c 1 -s 1 -f -o gpurun_out/r2x_class$c python tools/class_bench.py 32768 1 > gpurun_out/r2x_ncu$c.log 2>&1
done
