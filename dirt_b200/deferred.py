"""Deferred-shading helpers around the rasteriser.

samples/deferred.py:62-118 in the reference builds a G-buffer with THREE rasterise calls over the
same geometry (world position, albedo, world normal -- each 3 channels, the op's limit,
csrc/hwc.h:27) and then dilates position / normal at the silhouette with framework ops
(:87-91).  Visibility does not depend on the attributes, so here:
  rasterise_gbuffer   one rasterise call (one visibility pass) carrying all attribute sets as
                      channels (up to 16), split back into the caller's sets;
  dilate_silhouette   the mask + 3x3 max-pool + blend of :87-91 as one fused kernel each way.
"""
import torch

from . import _lib, rasterise_ops


class _Dilate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        lib = _lib.load()
        x = x.contiguous()
        out = torch.empty_like(x)
        B, H, W, C = x.shape
        with torch.cuda.device(x.device):
            _lib.check(lib.dirt_dilate_silhouette(x.data_ptr(), out.data_ptr(), B, H, W, C,
                                                  torch.cuda.current_stream(x.device).cuda_stream))
        ctx.save_for_backward(x)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        (x,) = ctx.saved_tensors
        lib = _lib.load()
        grad_out = grad_out.contiguous().to(torch.float32)
        grad_x = torch.empty_like(x)
        B, H, W, C = x.shape
        with torch.cuda.device(x.device):
            _lib.check(lib.dirt_dilate_silhouette_grad(x.data_ptr(), grad_out.data_ptr(), grad_x.data_ptr(), B, H, W, C,
                                                       torch.cuda.current_stream(x.device).cuda_stream))
        return grad_x


def dilate_silhouette(gbuffer):
    """Replaces every -inf (background) value of `gbuffer` [H,W,C] or [B,H,W,C] by the maximum of
    its 3x3 neighbourhood, as samples/deferred.py:87-91 does with a max-pool; differentiable."""
    if not isinstance(gbuffer, torch.Tensor) or not gbuffer.is_cuda:
        raise RuntimeError("dirt_b200.dilate_silhouette needs a CUDA tensor; there is no CPU fallback")
    if gbuffer.dtype != torch.float32 or gbuffer.dim() not in (3, 4):
        raise _lib.InvalidArgumentError(-1, "dilate_silhouette expects a float32 [H,W,C] or [B,H,W,C] tensor")
    batched = gbuffer.dim() == 4
    out = _Dilate.apply(gbuffer if batched else gbuffer.unsqueeze(0))
    return out if batched else out[0]


def rasterise_gbuffer(backgrounds, vertices, attribute_sets, faces):
    """One visibility pass for several per-vertex attribute sets.

    backgrounds[i] [B|1,H,W,Ci] and attribute_sets[i] [B|1,V,Ci] pair up; sum(Ci) <= 16.
    vertices [B|1,V,4] clip space, faces [B|1,F,3].  Returns a list of [B,H,W,Ci] tensors, the same
    values three separate `rasterise_batch` calls would give (samples/deferred.py:63-83)."""
    if len(backgrounds) != len(attribute_sets) or not backgrounds:
        raise _lib.InvalidArgumentError(-1, "rasterise_gbuffer expects one background per attribute set")
    widths = [int(a.shape[-1]) for a in attribute_sets]
    if sum(widths) > _lib.MAX_CHANNELS:
        raise _lib.InvalidArgumentError(-1, "rasterise_gbuffer: %d channels exceed the limit of %d" % (sum(widths), _lib.MAX_CHANNELS))
    B = max([int(t.shape[0]) for t in backgrounds] + [int(t.shape[0]) for t in attribute_sets])
    expand = lambda t: t if t.shape[0] == B else t.expand(B, *t.shape[1:])
    bg = torch.cat([expand(b) for b in backgrounds], dim=-1)
    Ba = max(int(t.shape[0]) for t in attribute_sets)
    at = torch.cat([a if a.shape[0] == Ba else a.expand(Ba, *a.shape[1:]) for a in attribute_sets], dim=-1)
    pixels = rasterise_ops.rasterise_batch(bg, vertices, at, faces)
    return list(torch.split(pixels, widths, dim=-1))
