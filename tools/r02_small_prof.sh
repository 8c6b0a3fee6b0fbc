mkdir -p gpurun_out
cd gaussian-process_b200/csrc
NVCC=/usr/local/cuda/bin/nvcc
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -DGPK_SMALL_PROFILE"
$NVCC $FLAGS -c gpk_loss.cu -o /tmp/gpk_loss_prof.o && $NVCC -shared -gencode arch=compute_100a,code=sm_100a -o ../../gpurun_out/libgpk_prof.so build/gpk_api.o build/gpk_cov.o build/gpk_diag.o build/gpk_gemm.o build/gpk_kmeans.o /tmp/gpk_loss_prof.o build/gpk_oz.o build/gpk_pso.o build/gpk_surface.o build/gpk_csv.o build/gpk_finish.o
cd ../..
GPK_LIB=$PWD/gpurun_out/libgpk_prof.so python tools/small_profile.py 100
GPK_LIB=$PWD/gpurun_out/libgpk_prof.so python tools/small_profile.py 21
GPK_LIB=$PWD/gpurun_out/libgpk_prof.so python tools/small_profile.py 64
GPK_LIB=$PWD/gpurun_out/libgpk_prof.so python tools/small_profile.py 128
rm -f gpurun_out/libgpk_prof.so
