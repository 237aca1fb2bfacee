timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_c1_energy_history.py -m gpu -x -q 2>&1 | tail -n 2
tools/gpu_call.sh ab vpic_b200/libvpic_b200.so
tools/gpu_call.sh ab2 vpic_b200/libvpic_b200.so
