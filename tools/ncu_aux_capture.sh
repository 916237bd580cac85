set -x
cd $GRAFT_REPO_ROOT
N="ncu --set full --clock-control none --import-source on -f"
timeout 120 python tools/aux_step.py 20000 > /dev/null 2>&1 && timeout 120 python tools/loss_step.py 512x512 3 >/dev/null 2>&1 || echo PLAIN_FAILED
timeout 300 $N -k regex:'nearest_vertex_grid_kernel|vertex_grid_build_kernel' -c 5 -o /tmp/grid python tools/aux_step.py 20000 > gpurun_out/ncu_grid.log 2>&1
timeout 300 $N -k regex:'nearest_vertex_grid_kernel' -c 3 -o /tmp/grid100k python tools/aux_step.py 100000 >> gpurun_out/ncu_grid.log 2>&1
timeout 300 $N -k regex:image_losses -c 3 -o /tmp/loss512 python tools/loss_step.py 512x512 3 > gpurun_out/ncu_loss.log 2>&1
timeout 600 $N -k regex:'preprocess_forward_kernel|preprocess_backward_kernel' -c 8 -o /tmp/fused python tools/fused_bench.py 20000 > gpurun_out/ncu_fused.log 2>&1
for f in grid grid100k loss512 fused; do ncu -i /tmp/$f.ncu-rep --page raw --csv > gpurun_out/r02_q_ncu_full_$f.raw.csv 2>/dev/null; done
ls -la gpurun_out/
