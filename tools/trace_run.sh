make -C speechseparation_b200/csrc clean >/dev/null; make -C speechseparation_b200/csrc -j EXTRA=-DBSB_TRACE 2>&1 | tail -1
timeout -s KILL 120 python tools/trace_tc.py
