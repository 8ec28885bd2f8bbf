mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o /tmp/umma_mn_probe tools/umma_mn_probe.cu > gpurun_out/mn_probe.log 2>&1
timeout 60 /tmp/umma_mn_probe >> gpurun_out/mn_probe.log 2>&1; echo "mn_probe rc=$?"
cat gpurun_out/mn_probe.log
