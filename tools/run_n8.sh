set -x
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $T --master-port 29521 bench.py --gpus 8 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err; echo bench rc=$?
timeout 300 $T --master-port 29523 tools/fused_exchange_diag.py --gaussians 6000000 --width 1920 --height 1080 2> gpurun_out/r2_fused_n8.err | grep "^{" > gpurun_out/r2_fused_exchange_diag_6M_n8.json; echo diag rc=$?
timeout 300 $T --master-port 29525 tools/bands_diag.py 6000000 3840 2160 > gpurun_out/r2_bands_diag_6M_4k_n8.log 2> gpurun_out/r2_bands_n8.err; echo bands rc=$?
timeout 600 $T --master-port 29527 bench.py --impl reference --gpus 8 --steps 20 --warmup 3 > gpurun_out/r2_bench_ref_n8.json 2> gpurun_out/r2_bench_ref_n8.err; echo ref rc=$?
ls gpurun_out | head -40
grep -h "NVLS\|nranks\|Connected" gpurun_out/nccl_*.log 2>/dev/null | head -8
tail -2 gpurun_out/r2_bench_n8.err gpurun_out/r2_fused_n8.err gpurun_out/r2_bands_n8.err
