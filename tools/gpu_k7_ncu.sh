# ncu evidence for K7: launch list + one full-set capture of each kernel (after the same command exited 0 without ncu)
timeout 120 python tools/k7_once.py > gpurun_out/k7_once.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/k7_once.log; exit 1; }
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_pw' --csv --log-file gpurun_out/k7_launches.csv python tools/k7_once.py > gpurun_out/k7_ncu1.log 2>&1; echo "ncu1 rc=$?"
grep -c k_pw gpurun_out/k7_launches.csv; grep k_pw gpurun_out/k7_launches.csv | tail -8 | cut -c1-260
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_pwconv' -s 4 -c 4 -o gpurun_out/prof_k7 python tools/k7_once.py > gpurun_out/k7_ncu2.log 2>&1; echo "ncu2 rc=$?"
ls -la gpurun_out/prof_k7.ncu-rep
