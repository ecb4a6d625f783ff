set -e
python tools/prof_hop.py ${ROWS:-2} 6 > gpurun_out/hop_plain.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum --clock-control none --cache-control none -k "regex:^(grouped_|lstm_|stream_|tc_)" -c 200 --csv --log-file gpurun_out/hop_launches.csv python tools/prof_hop.py ${ROWS:-2} 6 > gpurun_out/hop_ncu.log 2>&1
tail -2 gpurun_out/hop_ncu.log
