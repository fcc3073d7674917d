set -x
python bench.py --steps 2 --warmup 3 --skip-extract > gpurun_out/plain_match.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01_v8_match_launches.csv python bench.py --steps 2 --warmup 3 --skip-extract > gpurun_out/ncu_match1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:match_fused -s 2 -c 1 -o gpurun_out/prof_match python bench.py --steps 2 --warmup 3 --skip-extract > gpurun_out/ncu_match2.log 2>&1
python tools/time_vit.py dinov2_vitb14 32 > gpurun_out/plain_vit.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 600 -c 220 --csv --log-file gpurun_out/r01_v8_vit_launches.csv python tools/time_vit.py dinov2_vitb14 32 > gpurun_out/ncu_vit.log 2>&1
python tools/time_post.py 256 768 37 37 --once > gpurun_out/plain_post.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:moments|response|select|sample_pack' -c 40 -o gpurun_out/prof_post python tools/time_post.py 256 768 37 37 --once > gpurun_out/ncu_post.log 2>&1
ls -la gpurun_out/*.ncu-rep
