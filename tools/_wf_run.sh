cd /root/repo
for slots in 1048576 4194304; do
ZB_WAVEFRONT=1 ZB_WF_SLOTS=$slots timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/wf_launches_$slots.csv python tools/tune.py --scene weekend --spp 64 --reps 1 > gpurun_out/wf_ncu_$slots.log 2>&1
done
ZB_WAVEFRONT=1 ZB_WF_SLOTS=2097152 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/wf_launches_mesh.csv python tools/tune.py --scene mesh --spp 16 --reps 1 > gpurun_out/wf_ncu_mesh.log 2>&1
ls -la gpurun_out
