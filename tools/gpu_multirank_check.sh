mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 tools/check_multirank_host.py 2>&1 | grep -v "OMP_NUM\|\*\*\*\*\|^$" | tail -n 6
NG=2 TAG=r1o bash tools/gpu_round2.sh | tail -n 3 | cut -c1-200
