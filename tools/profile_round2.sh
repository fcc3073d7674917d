# Round-2 evidence run (one GPU).  Every profiled command first exits 0 without ncu.
set -x
O=gpurun_out
python bench.py --steps 10 --warmup 3 > $O/r02_bench.json 2> $O/r02_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r02_bench_reference.json 2> $O/r02_bench_reference.err
python bench.py --steps 2 --warmup 3 --skip-extract --skip-variants > $O/plain_match.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:match|emit|merge|scan_counts|pack_matches" -c 120 --csv --log-file $O/r02_match_launches.csv python bench.py --steps 2 --warmup 3 --skip-extract --skip-variants > $O/ncu_match1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:match_fused2_kernel -s 3 -c 1 -o $O/r02_match_fused2 python bench.py --steps 2 --warmup 3 --skip-extract --skip-variants > $O/ncu_match2.log 2>&1
python tools/time_guided.py 8 8192 128 > $O/r02_time_guided.txt 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:match_top2_kernel -s 2 -c 1 -o $O/r02_match_guided python tools/time_guided.py 8 8192 128 > $O/ncu_guided.log 2>&1
python tools/time_match_dims.py 16 8192 > $O/r02_time_match_dims.txt 2>&1
VC_MATCH_NO_PRUNE=1 python tools/time_match_dims.py 16 8192 > $O/r02_time_match_dims_noprune.txt 2>&1
ls -la $O/*.ncu-rep
