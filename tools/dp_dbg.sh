TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 120 python -m pytest tests/test_dp_p2p.py -q -m gpu --no-header -p no:cacheprovider 2>&1 | tail -2
ACT_DP_P2P=1 timeout 200 $TR --master-port 29512 tools/dp_check.py 2>&1 | grep -E "dp_check" | grep -v "rank" | grep -E "step 1|identical"
ACT_DP_P2P=1 timeout 200 $TR --master-port 29521 tools/timeline.py 8 2>&1 | tail -1
ACT_DP_P2P=1 timeout 200 $TR --master-port 29531 bench.py --gpus 2 --steps 20 --warmup 5 --no-infer 2>gpurun_out/dp_v.err | tail -1 | cut -c1-140
