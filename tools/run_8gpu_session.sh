#!/bin/bash
# Round-end evidence run on one 8-GPU box: multi-GPU correctness subset at 8 ranks, the headline and every variant of
# BASELINE.json at 8 GPUs, the library-only rung, the NVLink probe.  Every command has its own timeout.
O=gpurun_out/r2_8gpu; mkdir -p $O
DDL_TEST_NPROC=8 timeout 400 python -m pytest tests/test_multigpu.py -x -q -k "sync_colocated or sync_vs_nccl or sync_nvls_on or async_adam or async_dedicated" > $O/mgpu_tests_n8.log 2>&1; tail -2 $O/mgpu_tests_n8.log
b() { name=$1; shift; timeout 150 python bench.py --gpus 8 --steps 100 --warmup 10 "$@" > $O/$name.json 2> $O/$name.err; python - "$O/$name.json" <<'PY'
import sys, json
try:
    d = json.loads(open(sys.argv[1]).read())
    print(sys.argv[1].split('/')[-1], round(d["ms_per_step"], 4), int(d["value"]), d.get("params_bitwise_equal", d.get("async_all_pushes_applied")), d.get("exposed_push_pull_ms"), d.get("transport_stats", {}).get("nvls"))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
b headline_sync_sharding_greedy
b sync_unsharded --variant mnist_sync --no-exposed
b sync_sharding_block --variant mnist_sync_sharding --no-exposed
b sync_sharding_round_robin --variant mnist_sync_sharding --shard round_robin --no-exposed
b sync_sharding_flat --variant mnist_sync_sharding --shard flat --no-exposed
b async_dedicated_1ps_7w --variant mnist_async --layout dedicated --num-ps 1 --no-exposed
b async_unsharded_colocated --variant mnist_async --no-exposed
b async_sharding_block --variant mnist_async_sharding --no-exposed
b async_sharding_greedy --variant mnist_async_sharding_greedy --no-exposed
b widened_x6_sync_greedy --width-mult 6 --steps 30 --warmup 5 --no-exposed
b torch_graph --impl torch_graph
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29546 tools/p2p_probe.py --out $O/nvlink_peak_n8.json > $O/p2p.log 2>&1; tail -1 $O/p2p.log | cut -c1-100
