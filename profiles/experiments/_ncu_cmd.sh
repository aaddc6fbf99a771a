TAG=r2_v13
KRE="pose_fwd_kernel|pose_bwd_kernel|preprocess_kernel|preprocess_bwd_kernel|onesweep_pass_kernel|bin_place_kernel|blend_fwd_kernel|blend_bwd_kernel"
timeout 500 ncu --set full --clock-control none --import-source on -k regex:"${KRE}" -s 11 -c 11 -o gpurun_out/${TAG}_full -f python bench.py --profile-only --steps 3 --warmup 3 > gpurun_out/${TAG}_ncu2.log 2>&1; echo ncu2 rc=$?
