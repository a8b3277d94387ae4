python -m pytest tests/test_gpu_image.py -x -q > gpurun_out/s8_pytest.log 2>&1; echo pytest=$?; tail -3 gpurun_out/s8_pytest.log
python scripts/profile_trace.py wavefront 64 | tail -1
echo single pipe; FS2_TUNE=$((32+8)) python scripts/profile_trace.py wavefront 64 | tail -1
echo pool 4M; FS2_TUNE=$((4096*256+8)) python scripts/profile_trace.py wavefront 64 | tail -1
echo pool 16M; FS2_TUNE=$((16384*256+8)) python scripts/profile_trace.py wavefront 64 | tail -1
