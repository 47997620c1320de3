"""Host<->device copy bandwidth with N ranks copying AT THE SAME TIME (one process per GPU, torchrun), the traffic pattern of
the end-to-end bench: every rank streams 78.6 MB of pinned frames to its GPU and 14.8 MB of results back.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/pcie_probe_ranks.py

Rank 0 prints one JSON line: per-rank and aggregate GB/s for H2D alone, D2H alone, and both directions together, each
timed with CUDA events between barriers (max over ranks), plus the NUMA binding used (bench.py's bind_to_gpu_numa_node)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    bench.bind_to_gpu_numa_node(local)
    H, D = 78643200, 14795776
    x = torch.empty(H, dtype=torch.uint8).pin_memory()
    dx = torch.empty(H, dtype=torch.uint8, device="cuda")
    y = torch.empty(D, dtype=torch.uint8).pin_memory()
    dy = torch.empty(D, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(h2d, d2h, reps=40):
        def once():
            if h2d:
                with torch.cuda.stream(s1):
                    dx.copy_(x, non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    y.copy_(dy, non_blocking=True)
        for _ in range(5):
            once()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s1.wait_event(e0); s2.wait_event(e0)
        for _ in range(reps):
            once()
        f1, f2 = torch.cuda.Event(), torch.cuda.Event()
        f1.record(s1); f2.record(s2)
        torch.cuda.current_stream().wait_event(f1); torch.cuda.current_stream().wait_event(f2)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        s = ms.item() * 1e-3 / reps
        return s

    out = {"n_gpus": world, "h2d_bytes": H, "d2h_bytes": D}
    s = timed(True, False); out["h2d_only_GBps_per_rank"] = H / s / 1e9; out["h2d_only_GBps_total"] = world * H / s / 1e9
    s = timed(False, True); out["d2h_only_GBps_per_rank"] = D / s / 1e9; out["d2h_only_GBps_total"] = world * D / s / 1e9
    s = timed(True, True)
    out["both_h2d_GBps_per_rank"] = H / s / 1e9
    out["both_total_GBps"] = world * (H + D) / s / 1e9
    out["both_frames_per_s_bound_640x480"] = world * 256 / s   # 256 frames in, their results out: the PCIe-only e2e ceiling
    try:
        out["cpu_affinity"] = len(os.sched_getaffinity(0))
    except Exception:
        pass
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
