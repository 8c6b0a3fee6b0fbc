mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -3
python tools/pso_host_profile.py > gpurun_out/r02_pso_host_profile_v2.log 2>&1; echo "profile exit $?"; grep -E "per swarm|synchronize|_integral|copy_|_counts_packed" gpurun_out/r02_pso_host_profile_v2.log | cut -c1-150
bash tools/r02_small_prof.sh 2>&1 | tee gpurun_out/r02_small_profile_v2.log | head -12
