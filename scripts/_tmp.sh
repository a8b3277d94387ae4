python scripts/profile_trace.py wavefront 64 | tail -1
for b in 4 5; do echo "blocks=$b"; FS2_CUDA_LIB=$PWD/final-stage-path-tracer-2-0_b200/_exp/lib_b$b.so python scripts/profile_trace.py wavefront 64 | tail -1; done
