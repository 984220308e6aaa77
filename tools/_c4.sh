mkdir -p gpurun_out/r02c4
timeout 200 python tools/c4_profile.py > gpurun_out/r02c4/prof.txt 2>&1
C4_NOPROF=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 1500 --csv --log-file gpurun_out/r02c4/launches.csv python tools/c4_profile.py 30 > gpurun_out/r02c4/ncu.log 2>&1
head -80 gpurun_out/r02c4/prof.txt
