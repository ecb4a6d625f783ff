# pipeline timeline of the tensor-core Linear kernel: builds with -DBSB_TRACE, traces launch $BSB_TRACE_LAUNCH of
# bsb_bandsplit (0 = pre.0, 1 = pre.2, 2 = fc.0 ...), restores the normal build afterwards
make -C speechseparation_b200/csrc clean >/dev/null; make -C speechseparation_b200/csrc -j EXTRA=-DBSB_TRACE 2>&1 | tail -1
timeout -s KILL 120 python tools/trace_tc.py
