"""CorrBlock / AltCorrBlock with the reference's interface (dbaf/modules/corr.py:23-139), backed
by the native volume+pyramid kernel and the single-launch pyramid lookup.

Differences from the reference are internal only: the volume and its pyramid come from one
native call instead of torch.matmul + 3x avg_pool2d, and the four per-level lookups are one
launch.  Shapes and values of everything a caller sees are the same on the inference path
(float16 feature maps under torch.no_grad, what DBA-Fusion runs: dbaf/depth_video.py:64, dbaf/dbaf.py:53).

Restriction: the native volume kernel is float16 and not an autograd op.  Float32 feature maps are cast to half
(the reference's float32 training path keeps float32 volumes), and no gradient reaches the feature maps through
CorrBlock; with gradients enabled on the inputs the constructors raise instead of silently cutting the graph.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

import droid_backends


class CorrSampler(torch.autograd.Function):
    """dbaf/modules/corr.py:6-20"""

    @staticmethod
    def forward(ctx, volume, coords, radius):
        ctx.save_for_backward(volume, coords)
        ctx.radius = radius
        corr, = droid_backends.corr_index_forward(volume, coords, radius)
        return corr

    @staticmethod
    def backward(ctx, grad_output):
        volume, coords = ctx.saved_tensors
        grad_output = grad_output.contiguous()
        grad_volume, = droid_backends.corr_index_backward(volume, coords, grad_output, ctx.radius)
        return grad_volume, None, None


class CorrBlock:
    """dbaf/modules/corr.py:23-71.  fmap1, fmap2: [batch, num, dim, ht, wd] (float16 on the
    inference path)."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=3):
        self.num_levels = num_levels
        self.radius = radius
        batch, num, dim, ht, wd = fmap1.shape
        if torch.is_grad_enabled() and (fmap1.requires_grad or fmap2.requires_grad):
            raise RuntimeError("CorrBlock (B200): inference only -- the native volume kernel has no backward; build the "
                               "volume with torch.matmul + avg_pool2d to train through it")
        f1 = fmap1.reshape(batch * num, dim, ht, wd).contiguous()
        f2 = fmap2.reshape(batch * num, dim, ht, wd).contiguous()
        if f1.dtype != torch.float16:
            # the training path of the reference runs this in float32; the native kernel is fp16
            f1, f2 = f1.half(), f2.half()
        self.corr_pyramid = droid_backends.corr_volume_pyramid(f1, f2, num_levels)

    @classmethod
    def from_frames(cls, fmaps1, fmaps2, ii, jj, num_levels=4, radius=3):
        """The construction of dbaf/covisible_graph.py:129-131, CorrBlock(video.fmaps[ii,0][None],
        video.fmaps[jj,c][None]), without materialising the gathered maps: fmaps1 [n1,dim,ht,wd] and fmaps2
        [n2,dim,ht,wd] are the per-frame stacks (e.g. video.fmaps[:,0] and video.fmaps[:,c]); same pyramid,
        bit for bit."""
        self = cls.__new__(cls)
        self.num_levels, self.radius = num_levels, radius
        f1 = fmaps1 if fmaps1.dtype == torch.float16 else fmaps1.half()
        f2 = f1 if fmaps2 is fmaps1 else (fmaps2 if fmaps2.dtype == torch.float16 else fmaps2.half())
        self.corr_pyramid = droid_backends.corr_volume_pyramid_indexed(f1.contiguous(), f2.contiguous(), ii, jj,
                                                                       num_levels)
        return self

    def __call__(self, coords):
        batch, num, ht, wd, _ = coords.shape
        coords = coords.permute(0, 1, 4, 2, 3).contiguous().view(batch * num, 2, ht, wd)
        if self.radius == 3 and not torch.is_grad_enabled():
            out = droid_backends.corr_lookup_pyramid(self.corr_pyramid, coords.float(), self.radius)
            return out.view(batch, num, -1, ht, wd)
        out_pyramid = []
        for i in range(self.num_levels):
            corr = CorrSampler.apply(self.corr_pyramid[i], coords / 2 ** i, self.radius)
            out_pyramid.append(corr.view(batch, num, -1, ht, wd))
        return torch.cat(out_pyramid, dim=2)

    def cat(self, other):
        for i in range(self.num_levels):
            self.corr_pyramid[i] = torch.cat([self.corr_pyramid[i], other.corr_pyramid[i]], 0)
        return self

    def __getitem__(self, index):
        for i in range(self.num_levels):
            self.corr_pyramid[i] = self.corr_pyramid[i][index]
        return self

    @staticmethod
    def corr(fmap1, fmap2):
        """All-pairs correlation only (level 0), dbaf/modules/corr.py:63-71."""
        batch, num, dim, ht, wd = fmap1.shape
        f1 = fmap1.reshape(batch * num, dim, ht, wd).contiguous().half()
        f2 = fmap2.reshape(batch * num, dim, ht, wd).contiguous().half()
        vol, = droid_backends.corr_volume_pyramid(f1, f2, 1)
        return vol.view(batch, num, ht, wd, ht, wd)


class CorrLayer(torch.autograd.Function):
    """dbaf/modules/corr.py:74-88"""

    @staticmethod
    def forward(ctx, fmap1, fmap2, coords, r):
        ctx.r = r
        ctx.save_for_backward(fmap1, fmap2, coords)
        corr, = droid_backends.altcorr_forward(fmap1, fmap2, coords, ctx.r)
        return corr

    @staticmethod
    def backward(ctx, grad_corr):
        fmap1, fmap2, coords = ctx.saved_tensors
        grad_corr = grad_corr.contiguous()
        g1, g2, gc = droid_backends.altcorr_backward(fmap1, fmap2, coords, grad_corr, ctx.r)
        return g1, g2, gc, None


class AltCorrBlock:
    """Low-memory correlation: nothing is materialised, each lookup dots features on the fly.
    Same constructor / call signature and output layout as dbaf/modules/corr.py:91-139."""

    def __init__(self, fmaps, num_levels=4, radius=3):
        self.num_levels, self.radius = num_levels, radius
        B, N, C, H, W = fmaps.shape
        level = fmaps.reshape(B * N, C, H, W) / 4.0          # same pre-scaling as CorrBlock.corr
        self.pyramid = []                                     # level l: [B, N, H>>l, W>>l, C], pixel-major
        for l in range(num_levels):
            self.pyramid.append(level.permute(0, 2, 3, 1).contiguous().view(B, N, H >> l, W >> l, C))
            if l + 1 < num_levels:
                level = F.avg_pool2d(level, 2, stride=2)

    def _lookup_level(self, l, coords, ii, jj):
        # coords: [B, E, S, H, W, 2] in level-0 pixels
        B, E, S, H, W, _ = coords.shape
        src = self.pyramid[0][:, ii].flatten(0, 1).float().contiguous()
        dst = self.pyramid[l][:, jj].flatten(0, 1).float().contiguous()
        xy = (coords * (0.5 ** l)).reshape(B * E, S, H, W, 2).contiguous()
        out = CorrLayer.apply(src, dst, xy, self.radius)      # [B*E, S, 49, H, W]
        return out.view(B, E, S, -1, H, W).permute(0, 1, 3, 4, 5, 2)

    def __call__(self, coords, ii, jj):
        single = coords.dim() == 5                            # [B, E, H, W, 2] -> one coordinate set
        if single:
            coords = coords.unsqueeze(-2)
        coords = coords.permute(0, 1, 4, 2, 3, 5)             # [B, E, S, H, W, 2]
        corr = torch.cat([self._lookup_level(l, coords, ii, jj) for l in range(self.num_levels)], dim=2)
        if single:
            corr = corr.squeeze(-1)
        return corr.contiguous()


class CorrPool:
    """CorrBlock whose pyramids live in a fixed pool of `capacity` slots, so that adding and removing edges
    never copies a volume.  The reference re-creates its pyramid tensors on every edge change
    (`CorrBlock.cat`, `CorrBlock.__getitem__` with a boolean mask, dbaf/modules/corr.py:52-60, called from
    dbaf/covisible_graph.py:131-132 and :166-167): adding one keyframe's edges to a 48-edge graph copies the whole
    2 GB pyramid.  Here `append` builds the new edges' volumes straight into free slots and `keep` only edits the
    edge -> slot table; the lookup reads through that table.  Edge order is the order of the calls, as in the
    reference (new edges at the end, removal keeps the order of the survivors), and the lookup result is
    bit-identical to a CorrBlock built from scratch on the surviving edge list."""

    def __init__(self, capacity, ht, wd, num_levels=4, radius=3, device="cuda", dtype=torch.float16):
        if radius != 3:
            raise RuntimeError("CorrPool supports radius 3 (the only one the pipeline uses)")
        self.capacity, self.ht, self.wd = int(capacity), int(ht), int(wd)
        self.num_levels, self.radius = int(num_levels), int(radius)
        self.pool = [torch.empty(self.capacity, ht, wd, ht >> l, wd >> l, dtype=dtype, device=device)
                     for l in range(self.num_levels)]
        self.slots = torch.empty(0, dtype=torch.int32, device=device)      # slot of every live edge, in edge order
        self._slots_host = []                                              # the same list on the host
        self._free = list(range(self.capacity - 1, -1, -1))               # host-side free list (pop from the end)

    def __len__(self):
        return int(self.slots.numel())

    def append(self, fmaps1, fmaps2, ii, jj):
        """Add the edges (ii[e], jj[e]): correlate fmaps1[ii[e]] with fmaps2[jj[e]] (per-frame stacks [n,C,ht,wd])."""
        n_new = int(ii.numel())
        if n_new == 0:
            return
        if n_new > len(self._free):
            raise RuntimeError(f"CorrPool is full: {n_new} new edges, {len(self._free)} free slots")
        new = [self._free.pop() for _ in range(n_new)]
        dev = self.slots.device
        new_t = torch.tensor(new, dtype=torch.int32, device=dev)
        f1 = fmaps1 if fmaps1.dtype == torch.float16 else fmaps1.half()
        f2 = f1 if fmaps2 is fmaps1 else (fmaps2 if fmaps2.dtype == torch.float16 else fmaps2.half())
        droid_backends.corr_volume_pyramid_pool(f1.contiguous(), f2.contiguous(), ii, jj, new_t, self.pool)
        self.slots = torch.cat([self.slots, new_t])
        self._slots_host.extend(new)

    def keep(self, mask):
        """Drop the edges where `mask` is False (the reference's `corr = corr[~remove_mask]`)."""
        keep = mask.detach().to("cpu", torch.bool).tolist()      # the one host read: the mask itself
        if len(keep) != len(self._slots_host):
            raise RuntimeError("mask must have one entry per live edge")
        gone = [s for s, k in zip(self._slots_host, keep) if not k]
        self._free.extend(reversed(gone))
        self._slots_host = [s for s, k in zip(self._slots_host, keep) if k]
        self.slots = torch.tensor(self._slots_host, dtype=torch.int32, device=self.slots.device)

    def __call__(self, coords):
        """coords [batch=1, num, ht, wd, 2] as for CorrBlock.__call__ -> [1, num, levels*49, ht, wd]."""
        batch, num, ht, wd, _ = coords.shape
        if batch != 1 or num != len(self):
            raise RuntimeError("coords must be [1, number of live edges, ht, wd, 2]")
        c = coords.permute(0, 1, 4, 2, 3).contiguous().view(num, 2, ht, wd)
        out = droid_backends.corr_lookup_pyramid_pool(self.pool, c.float(), self.slots, self.radius)
        return out.view(1, num, -1, ht, wd)
