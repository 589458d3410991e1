"""Gradient aggregation plumbing: contiguous buckets of the flat fp32 fusion buffer, all-reduced
(SUM) with torch.distributed — NCCL over NVLink/NVSwitch on the GPUs, gloo in the CPU tests.
The 1/R of SyncReplicasOptimizer's mean (models/resnet_model.py:104-116; R = replicas_to_aggregate
= number of ranks) is applied inside the fused optimizer kernel, not here."""


def plan_buckets(offsets, n_flat, ready_after, cap_elems, total_launches):
    """offsets[i] / ready_after[i]: flat offset of trainable variable i (creation order) and the
    number of backward launches after which its gradient is final.  Gradients complete in
    REVERSE creation order, so buckets are cut from the end of the buffer.  Returns
    [(lo, hi, ready)] with hi descending, covering [0, n_flat), ready monotone."""
    buckets = []
    hi = n_flat
    ready = 0
    for i in range(len(offsets) - 1, -1, -1):
        ready = max(ready, ready_after[i])
        lo = offsets[i]
        if hi - lo >= cap_elems:
            buckets.append((lo, hi, ready))
            hi = lo
            ready = 0
    if hi > 0:
        buckets.append((0, hi, total_launches))
    m = 0
    out = []
    for lo, hi_, r in buckets:
        m = max(m, r)
        out.append((lo, hi_, m))
    return out


def allreduce_buckets(flat, buckets, group):
    import torch.distributed as dist
    for lo, hi, _ in buckets:
        dist.all_reduce(flat[lo:hi], op=dist.ReduceOp.SUM, group=group)
