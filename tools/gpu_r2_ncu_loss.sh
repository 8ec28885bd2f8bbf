# one ncu --set full capture (with source) of the training-mode loss kernel, built with $VFD_NVCC_DEFINES
mkdir -p gpurun_out
python -m vfdepth_b200.build > gpurun_out/build_ncu.log 2>&1 || { tail -5 gpurun_out/build_ncu.log; exit 1; }
timeout 600 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph --no-extras > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${KERNELS:-warp_loss_bwd_kernel}" -c ${COUNT:-1} -o gpurun_out/${OUT:-prof_loss2} python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-graph --no-extras > gpurun_out/ncu_loss.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_loss.log
cp vfdepth_b200/lib/libvfdepth_b200.so gpurun_out/lib_${OUT:-prof_loss2}.so
