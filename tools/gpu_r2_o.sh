timeout 200 python tools/frame_time.py c4 reps=2 trace=2 > gpurun_out/trace2_c4_w1.log 2>&1
timeout 200 python tools/frame_time.py c4 reps=2 trace=2 world=8 > gpurun_out/trace2_c4_w8.log 2>&1
tail -n 45 gpurun_out/trace2_c4_w8.log
