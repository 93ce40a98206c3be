cd /root/repo
python tools/refine_step.py > gpurun_out/refine_step_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:head_kernel -s 2 -c 2 -f -o gpurun_out/r2_head_full python tools/refine_step.py > gpurun_out/r2_ncu_full.log 2>&1
python tools/refine_step.py > gpurun_out/refine_step_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_final.csv python tools/refine_step.py > gpurun_out/refine_step_ncu.log 2>&1
echo done
