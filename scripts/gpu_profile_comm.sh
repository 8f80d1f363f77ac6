mkdir -p gpurun_out
timeout 120 python scripts/ncu_target.py > gpurun_out/ncu_target_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/ncu_target_plain.log; exit 1; }
tail -1 gpurun_out/ncu_target_plain.log
timeout 240 ncu --set full --clock-control none --import-source on -k regex:"k_coll|k_encode_push|k_recv_apply|k_band|k_refine|k_sample_hist" \
  --launch-skip 10 --launch-count 6 -o gpurun_out/prof_comm_r20 python scripts/ncu_target.py > gpurun_out/ncu_r20.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_r20.log; ls -la gpurun_out/prof_comm_r20.ncu-rep
