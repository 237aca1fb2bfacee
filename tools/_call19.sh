O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_c1_energy_history.py -m gpu -x -q > $O/c19_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/c19_pytest.log
tools/gpu_call.sh c19 vpic_b200/libvpic_b200.so
tools/gpu_call.sh c19b vpic_b200/libvpic_b200.so
