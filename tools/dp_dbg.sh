export CUDA_DEVICE_MAX_CONNECTIONS=32
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for st in 1 2; do for p in 0 1; do
echo "== steps $st p2p $p"
ACT_DP_STEPS=$st ACT_DP_DIAG=1 ACT_CONCURRENT=0 ACT_DP_P2P=$p timeout 200 $TR --master-port 295$st$p tools/dp_check.py 2>&1 | grep "dp_check" | grep -v "rank"
done; done
