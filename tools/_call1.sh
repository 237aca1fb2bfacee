O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/c1_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/c1_pytest.log
for cv in 100 86 70; do
  VPIC_B200_AP_CARVEOUT=$cv tools/gpu_call.sh c1_cv$cv vpic_b200/libvpic_b200.so
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_advance_p --launch-skip 8 -c 2 -o $O/r1j_ap -f python bench.py --cells 128x128x128 --steps 4 --warmup 4 --no-cpu --no-e2e --no-yee > $O/c1_ncu.log 2>&1; echo "ncu rc=$?"
ls -la $O
