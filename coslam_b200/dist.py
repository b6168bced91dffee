"""Ray-sharded mapping bundle adjustment over several GPUs.

One process per GPU (torch.distributed, NCCL over NVLink/NVSwitch).  The hash table, the decoder and the poses are
replicated; the globally sampled ray list is cut into contiguous, equally sized shards.  Per step there are exactly two
exchange points, both sums:
  AR-1  the 3 int64 mask counts + 4 double loss sums, inside RenderLossFunction, so that every rank finishes the SAME
        global losses (the loss normalisers are batch-global, model/utils.py:111-115);
  AR-2  one flat fp32 bucket [table grad | decoder grads | extra (pose) grads] after the backward.
Tracking stays on one GPU.  The host-side logic here runs unchanged on the gloo backend (tests/test_dist_cpu.py).
"""
import torch
import torch.distributed as dist


def shard_bounds(n_items, world_size, rank):
    """Contiguous shard [lo, hi) of rank `rank`; the union over ranks is exactly range(n_items), in order, so the global
    ray list (and therefore the keyframe ray indices) is identical to the single-GPU run."""
    base, rem = divmod(int(n_items), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def equal_shard(n_items, world_size, rank):
    """Shard for the fused loss head, which assumes equal shard sizes (n_total = n_local * world): the list is truncated
    to a multiple of the world size like DistributedSampler(drop_last=True)."""
    per = int(n_items) // int(world_size)
    return rank * per, (rank + 1) * per


class GradBucket(object):
    """Flat fp32 bucket for AR-2.  pack() copies the gradients in, all_reduce() sums over the group, unpack() writes
    the (optionally averaged) result back into the .grad tensors."""

    def __init__(self, params, group=None):
        self.params = [p for p in params]
        self.group = group
        self.flat = None

    def _ensure(self):
        n = sum(p.numel() for p in self.params)
        ref = self.params[0]
        if self.flat is None or self.flat.numel() != n or self.flat.device != ref.device:
            self.flat = torch.zeros(n, dtype=torch.float32, device=ref.device)

    def pack(self):
        self._ensure()
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is None:
                self.flat[off:off + n].zero_()
            else:
                self.flat[off:off + n].copy_(p.grad.reshape(-1))
            off += n
        return self.flat

    def all_reduce(self):
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)

    def unpack(self, scale=1.0):
        off = 0
        for p in self.params:
            n = p.numel()
            g = self.flat[off:off + n].view_as(p)
            if scale != 1.0:
                g = g * scale
            if p.grad is None:
                p.grad = g.clone()
            else:
                p.grad.copy_(g)
            off += n


class ShardedMapper(object):
    """Drives one mapping step of `model` (a coslam_b200 JointEncoding) on this rank's shard of the ray batch."""

    def __init__(self, model, loss_weights, extra_params=(), group=None):
        self.model = model
        self.group = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        self.world = dist.get_world_size(self.group) if self.group is not None else 1
        self.model.process_group = self.group if self.world > 1 else None
        self.loss_weights = loss_weights       # (rgb, depth, sdf, fs) of coslam.py:140-160
        params = [p for p in model.parameters() if p.requires_grad and p.numel() > 0] + list(extra_params)
        self.bucket = GradBucket(params, self.group)

    def step(self, rays_o, rays_d, target_rgb, target_d):
        """forward + backward on the local shard, then AR-2.  Losses returned are the GLOBAL ones."""
        ret = self.model.forward(rays_o, rays_d, target_rgb, target_d)
        w = self.loss_weights
        loss = w[0] * ret['rgb_loss'] + w[1] * ret['depth_loss'] + w[2] * ret['sdf_loss'] + w[3] * ret['fs_loss']
        loss.backward()
        if self.world > 1:
            self.bucket.pack()
            self.bucket.all_reduce()
            self.bucket.unpack()
        return loss, ret
