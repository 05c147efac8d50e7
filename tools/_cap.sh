export VMASB200_JIT_TE=32
bash tools/profile_lines.sh r2t football@1048576 > /dev/null 2>&1
grep -E "stalled_(no_instruction|barrier|long_scoreboard|wait|short_scoreboard|branch_resolving|selected|not_selected|math|lg_thr|mio).*per_issue_active|issue_active.avg|registers_per_thread \[|warps_active|gpu__time_duration|inst_executed.sum \[|occupancy_limit|grid_size|block_size" gpurun_out/r2t/step_kernel_football_1048576.txt
