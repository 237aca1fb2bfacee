O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/c3_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/c3_pytest.log
for m in 0 8 16 32 64 128 auto; do
  if [ $m = auto ]; then unset VPIC_B200_AP_M; else export VPIC_B200_AP_M=$m; fi
  tools/gpu_call.sh c3_m$m vpic_b200/libvpic_b200.so
done
