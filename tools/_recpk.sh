mkdir -p gpurun_out/r02pk
timeout 120 python tools/gram_probe.py 64 19 64 3 > /dev/null 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02pk/l64.csv python tools/gram_probe.py 64 19 64 3 > gpurun_out/r02pk/ncu64.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:rec_.*_pk_kernel -s 4 -c 2 -o gpurun_out/r02pk/prof_recpk python tools/gram_probe.py 64 19 64 3 > gpurun_out/r02pk/ncu_full.log 2>&1
ls gpurun_out/r02pk
