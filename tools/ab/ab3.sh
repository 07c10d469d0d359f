timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for i in 1 2; do
RT_B200_LIB=tools/ab/noasync.so timeout 300 python tools/prim_bench.py
timeout 300 python tools/prim_bench.py
done
RT_B200_LIB=tools/ab/noasync.so timeout 300 python tools/shard_quick.py
timeout 300 python tools/shard_quick.py
RT_B200_LIB=tools/ab/noasync.so timeout 300 python tools/deep_quick.py
timeout 300 python tools/deep_quick.py
