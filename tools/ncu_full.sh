timeout 300 python tools/ncu_target.py || exit 1
timeout 1500 ncu --set full --clock-control none --import-source on \
  -k regex:'k_trace_stream_spec|k_path_persistent|k_ltc|k_primary|k_denoise_atrous|k_shadow_resolve|k_shadow_gen' -s 20 -c 14 \
  -o gpurun_out/r1o_full -f python tools/ncu_target.py > gpurun_out/r1o_full.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/r1o_full.ncu-rep
