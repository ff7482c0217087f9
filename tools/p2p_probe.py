"""NVLink / NVSwitch probe: peer bandwidth and round-trip latency between the GPUs of one box, measured with this
library's own kernels over cudaIpc-mapped memory (the same mapping the fused transport uses).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/p2p_probe.py [--out profiles/nvlink_peak.json]

Reported (rank 0 <-> rank 1 unless noted; CUDA events, median of 5, max over the ranks taking part):
  copy_kernel_gbs_unidir   rank 0 copies 256 MiB out of rank 1's memory (float4 loads over NVLink, local stores)
  copy_kernel_gbs_bidir    rank 0 and rank 1 do so at the same time (per direction)
  store_kernel_gbs_unidir  rank 0 writes 256 MiB into rank 1's memory (remote stores)
  memcpy_gbs_unidir        cudaMemcpyAsync device-to-device through the mapped pointer (copy engine)
  all_ranks_gbs            every rank r reads from rank (r + 1) % N at the same time (per GPU, NVSwitch under full load)
  ld_rtt_us / ld_rtt_idle_us     one thread, dependent 4-byte remote loads back to back / after 50 us of link idle
  st_fence_rtt_us / ..._idle_us  one thread, remote store + fence.sys (the wait behind every "delivered" signal)
  local_ld_us              the same dependent-load loop on local memory, for scale
  *_peer_polling / *_self_polling   the round trips while 102 blocks of the peer / of rank 0 poll a flag (ld.acquire.sys)
"""
import argparse, ctypes as C, json, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from ddl_b200.ops import native
from ddl_b200.parallel.transport.fused import SymmetricMemory


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--mib", type=int, default=256)
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    L = native.lib()
    nbytes = a.mib << 20
    mem = SymmetricMemory(nbytes, rank, world, dev)
    local = torch.empty(nbytes // 4, dtype=torch.float32, device=dev)
    mem.tensor().normal_()
    out4 = torch.zeros(4, dtype=torch.int64, device=dev)
    s = native.stream_ptr
    peer = (rank + 1) % world

    def sync():
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()

    def timed(fn, active):
        """median ms over 5 runs of fn() on the ranks in `active` (others idle), max over ranks"""
        ts = []
        for _ in range(6):
            sync()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if rank in active:
                fn()
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) if rank in active else 0.0], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t.item()))
        return statistics.median(ts[1:])

    gbs = lambda ms: nbytes / (ms * 1e-3) / 1e9
    res = {"world": world, "bytes": nbytes, "gpu": torch.cuda.get_device_name(dev)}
    blocks = 148 * 8
    read_peer = lambda: native.check(L.ddl_p2p_copy(local.data_ptr(), mem.ptrs[peer], nbytes, blocks, s()), "p2p_copy")
    write_peer = lambda: native.check(L.ddl_p2p_copy(mem.ptrs[peer], local.data_ptr(), nbytes, blocks, s()), "p2p_copy")
    res["copy_kernel_gbs_unidir"] = gbs(timed(read_peer, {0}))
    res["copy_kernel_gbs_bidir"] = gbs(timed(read_peer, {0, 1}))
    res["store_kernel_gbs_unidir"] = gbs(timed(write_peer, {0}))
    res["memcpy_gbs_unidir"] = gbs(timed(lambda: native.check(L.ddl_memcpy_async(local.data_ptr(), mem.ptrs[peer], nbytes, s()), "memcpy"), {0}))
    res["all_ranks_gbs"] = gbs(timed(read_peer, set(range(world))))
    res["local_copy_kernel_gbs"] = gbs(timed(lambda: native.check(L.ddl_p2p_copy(local.data_ptr(), mem.local, nbytes, blocks, s()), "copy"), {0}))
    probe_latency(res, L, mem, out4, rank, world, peer, nbytes, dev, sync, s)
    sync()
    if rank == 0:
        print(json.dumps(res, indent=1))
        if a.out:
            json.dump(res, open(a.out, "w"), indent=1)
    mem.close()
    dist.destroy_process_group()


def probe_latency(res, L, mem, out4, rank, world, peer, nbytes, dev, sync, s):
    def lat(ptr, idle_ns, store, iters=200):
        sync()
        if rank == 0:
            native.check(L.ddl_p2p_latency(ptr, nbytes // 4, iters, idle_ns, store, out4.data_ptr(), s()), "p2p_latency")
            torch.cuda.synchronize()
            v = out4.cpu().tolist()
            return {"mean_us": v[0] / iters / 1e3, "first_us": v[1] / 1e3, "max_us": v[2] / 1e3}
        return None

    for name, ptr, idle, store in [("ld_rtt", mem.ptrs[peer], 0, 0), ("ld_rtt_idle", mem.ptrs[peer], 50000, 0),
                                   ("st_fence_rtt", mem.ptrs[peer], 0, 1), ("st_fence_rtt_idle", mem.ptrs[peer], 50000, 1),
                                   ("local_ld", mem.local, 0, 0), ("local_st_fence", mem.local, 0, 1)]:
        r = lat(ptr, idle, store)
        if r:
            res[name + "_us"] = r
    # the same round trips while the PEER's SMs poll a flag in its own memory with ld.acquire.sys (102 blocks x 2 threads,
    # what a rank does while it waits in a sync-kernel barrier), and while rank 0's own SMs do
    spin_out = torch.zeros(1, dtype=torch.int64, device=dev)
    side = torch.cuda.Stream(device=dev)
    for who, label in ((1, "peer_polling"), (0, "self_polling")):
        for name, store in (("ld_rtt", 0), ("st_fence_rtt", 1)):
            sync()
            if rank == who:
                with torch.cuda.stream(side):
                    native.check(L.ddl_p2p_spin(mem.local, 102, 3_000_000, spin_out.data_ptr(), native.stream_ptr()), "spin")
            dist.barrier()
            if rank == 0:
                native.check(L.ddl_p2p_latency(mem.ptrs[peer], nbytes // 4, 200, 0, store, out4.data_ptr(), s()), "p2p_latency")
                torch.cuda.synchronize()
                v = out4.cpu().tolist()
                res["%s_%s_us" % (name, label)] = {"mean_us": v[0] / 200 / 1e3, "first_us": v[1] / 1e3, "max_us": v[2] / 1e3}
            torch.cuda.synchronize()


if __name__ == "__main__":
    main()
