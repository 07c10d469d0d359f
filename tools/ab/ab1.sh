set -x
timeout 600 python -m pytest tests/test_gpu_full_size.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
for i in 1 2; do
RT_B200_LIB=tools/ab/noorder.so timeout 300 python tools/prim_bench.py
timeout 300 python tools/prim_bench.py
done
RT_B200_LIB=tools/ab/noorder.so timeout 300 python tools/shard_quick.py
timeout 300 python tools/shard_quick.py
