# usage (on the GPU box): bash tools/ab_build.sh "-DBVC_LONG_GROUP=4"   -> rebuilds libbvc_gpu.so with extra nvcc flags
cd basevarc_b200/csrc && touch bvc_kernels.cu && make EXTRA="$1" > /dev/null 2>&1; cd ../..
