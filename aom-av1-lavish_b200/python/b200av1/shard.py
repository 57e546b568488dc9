"""Frame sharding across the GPUs of one box: frames are independent,
so rank r of W simply owns frames r, r+W, r+2W, ... -- no data-path collective.  The only
communication is the measurement: a barrier, max-over-ranks of the device time and the sum
of the units processed (torch.distributed: NCCL on GPUs, gloo in the CPU tests)."""


def frames_of_rank(n_frames, rank, world):
    """indices of the frames rank `rank` processes (round-robin, config 5 of BASELINE.json)"""
    return list(range(rank, n_frames, world))


def aggregate_throughput(units_local, seconds_local, dist=None, device=None):
    """whole-job units/s = sum of units over ranks / max time over ranks"""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return units_local / seconds_local, units_local, seconds_local
    import torch
    t = torch.tensor([seconds_local], dtype=torch.float64, device=device)
    u = torch.tensor([float(units_local)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return float(u.item()) / float(t.item()), float(u.item()), float(t.item())
