for b in probe_noorder probe_order probe_bottomup; do RT_B200_LIB=tools/ab/$b.so timeout 200 python tools/busy_probe.py; done
