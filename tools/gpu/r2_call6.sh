# Round 2 (1 GPU): ONE ncu --set full capture of every kernel family (launcher exited 0 without ncu in call 6a:
# gpurun_out/ncu_all_plain.log). The report stays on the box (> 64 MiB); its raw page comes back as CSV.
set -x
timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:b200:: -c 150 -f -o /tmp/ncu_all_r2 python tools/gpu/ncu_all.py > gpurun_out/ncu_all.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_all.log
ncu -i /tmp/ncu_all_r2.ncu-rep --page raw --csv > gpurun_out/ncu_all_r2_raw.csv 2>gpurun_out/ncu_export.err; ls -la gpurun_out/ncu_all_r2_raw.csv /tmp/ncu_all_r2.ncu-rep
