# Development aid: the measurement batch behind profiles/r02_*.{json,jsonl,txt} of the end of round 2, as one gpurun
# command (`gpurun --timeout 3000 -- bash tools/final_batch.sh`). The effort sweep needs a library with the effort
# knobs compiled in; build it here, before the snapshot travels (tools/_variants/ is git-ignored, not gpurun-ignored):
#   nvcc <flags of devspace_b200/build.py> -DDSG_LZ_RUNTIME_EFFORT -o tools/_variants/libdsgpack_effort.so csrc/*.cu csrc/*.cc
set -x
cd $GRAFT_REPO_ROOT
O=gpurun_out
(time timeout 900 python -m pytest tests -m gpu -x -q) > $O/r02_gputests_final.log 2>&1; tail -3 $O/r02_gputests_final.log
timeout 500 python bench.py --impl reference --steps 3 --warmup 3 > $O/r02_bench_reference_arm.json 2> $O/ref_err.log
timeout 700 python bench.py --steps 3 --warmup 3 > $O/r02_bench_n1_c4shard.json 2> $O/bench_err.log; cut -c1-400 $O/r02_bench_n1_c4shard.json
(LZ_PROF=1 timeout 300 python tools/debug_deflate.py 100000000; LZ_PROF=1 LZ_KINDS=bin timeout 300 python tools/debug_deflate.py 1000000000; LZ_PROF=1 LZ_KINDS=text timeout 300 python tools/debug_deflate.py 1000000000) > $O/r02_lz_prof_final.txt 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:lz_deflate -c 1 -f -o $O/r02c_lz env LZ_KINDS=bin python tools/debug_deflate.py 20000000 > $O/ncu_log.txt 2>&1
timeout 200 python tools/small_latency.py > $O/r02_small_latency.jsonl 2>&1
timeout 400 python tools/sweep_geometry.py 512 > $O/r02_geometry_sweep.jsonl 2>&1
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/mm tools/micro_match.cu 2>/dev/null && timeout 60 /tmp/mm > $O/r02_micro_match.txt 2>&1
timeout 300 python tools/e2e_segments.py 16 0 256 512 1024 > $O/r02_e2e_segments.jsonl 2>&1
timeout 300 python tools/size_ratios.py > $O/r02_size_ratios.json 2>&1
timeout 300 python tools/crossover.py > $O/r02_crossover.jsonl 2>&1
cp devspace_b200/libdsgpack.so /tmp/default.so; cp tools/_variants/libdsgpack_effort.so devspace_b200/libdsgpack.so
timeout 600 python tools/sweep_effort.py 512 r2d96,r2d64,r2d32,r1d160,r1d96,r1d64,r1d32,r1d16,r1d8,r3,s1,s4 > $O/r02_effort_sweep.jsonl 2>&1
cp /tmp/default.so devspace_b200/libdsgpack.so
tail -2 $O/r02_effort_sweep.jsonl
