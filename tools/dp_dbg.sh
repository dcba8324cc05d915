TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for d in 0 1 2; do
ACT_DP_TIMING_DIAG=$d ACT_CONCURRENT=0 ACT_DP_P2P=1 timeout 200 $TR --master-port 2952$d tools/timeline.py 8 2>&1 | tail -1
python tools/timeline_print.py | grep -E "dp_adamw" | sed "s/^/diag$d /"
done
