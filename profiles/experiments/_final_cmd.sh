TAG=r2_v13
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/${TAG}_pytest_gpu_tail.txt; cat gpurun_out/${TAG}_pytest_gpu_tail.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 400 python bench.py --steps 200 --warmup 20 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo bench rc=$?
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference_arm.json 2>/dev/null; echo ref rc=$?
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --profile-only --steps 3 --warmup 3 > gpurun_out/${TAG}_ncu1.log 2>&1; echo ncu1 rc=$?
KRE="pose_fwd_kernel|pose_bwd_kernel|preprocess_kernel|preprocess_bwd_kernel|onesweep_pass_kernel|bin_place_kernel|blend_fwd_kernel|blend_bwd_kernel"
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"${KRE}" -s 33 -c 11 -o gpurun_out/${TAG}_full -f python bench.py --profile-only --steps 3 --warmup 3 > gpurun_out/${TAG}_ncu2.log 2>&1; echo ncu2 rc=$?
ls -la gpurun_out/${TAG}_*
