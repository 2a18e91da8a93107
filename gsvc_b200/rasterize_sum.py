"""rasterize_gaussians_sum with the reference's signature and autograd contract
(gsplat/gsplat/rasterize_sum.py:14-254 of the reference).

One C call launches the whole forward on torch's current stream and NOTHING waits for the device:
the reference blocks the host on `cum_tiles_hit[-1].item()` before it can launch anything past the
scan (utils.py:117); here the intersection count stays on the device (the kernels handle
the "no intersection" case themselves), a copy of it lands in a pinned mailbox that the host reads
whenever it next gets round to it (capacity hint for later calls, lazy overflow check), and the
forward + backward can be captured in a CUDA graph.

Default path (`gs2d_rasterize_sum_fwd_ranked`): per-tile counts + last-block scan, scatter, tile
kernel that orders each tile's batch by Gaussian id while staging it -- the same tile ranges and
the same order as the reference's cumsum -> map -> torch.sort -> gather -> bin edges chain.  The
step-by-step chain with the custom onesweep sort (`gs2d_rasterize_sum_fwd`) serves the public
binning utilities, callers with non-zero depths, the N-channel / untruncated options and
`options.sync_free = False`.

Capacity: the packed records of one call live in a buffer sized from the intersection counts seen
so far for that image size (4x headroom).  The first call of a size waits for its count once.  If a
later call still outgrows its buffer, that call's image is filled with NaN (never silently wrong),
and the next call into this module -- or that call's backward -- raises with the buffer regrown.
"""
from __future__ import annotations

import collections
from typing import Optional

import torch
from torch import Tensor
from torch.autograd import Function

from . import _lib
from . import cuda as _C
from .options import options

_RING = 64
GS2D_MAX_PER_TILE = 256   # forward.cu:566-614
_CAP_QUANTUM = 16384
_HEADROOM = 4             # capacity hint = this many times the largest intersection count seen


class CapacityOverflow(RuntimeError):
    """an earlier sync-free call had more intersections than its buffer held (its image was filled
    with NaN); the buffer has been regrown -- run the iteration again"""


class _DeviceCtx:
    """per-device mailbox ring, capacity hints and cached layouts"""

    def __init__(self, dev: torch.device):
        self.mail = torch.zeros((_RING, 8), dtype=torch.int32).pin_memory()     # sync-free path, slot = seq % ring
        self.mail_np = self.mail.numpy()
        self.mail_ptr = self.mail.data_ptr()
        self.mail_sync = torch.zeros((_RING, 8), dtype=torch.int32).pin_memory()   # step-by-step path (event waits)
        self.mail_sync_np = self.mail_sync.numpy()
        self.events = []
        with torch.cuda.device(dev):
            for _ in range(_RING):
                e = torch.cuda.Event()
                e.record()   # creates the underlying cudaEvent_t
                self.events.append(e)
        self.slot = 0
        self.last_num_intersects = 0
        self.cap = {}       # (H, W) -> capacity in intersections
        self.cap_n = {}     # (H, W) -> number of Gaussians of the calls the hint was learnt from
        self.layouts = {}   # (n, cap, tx, ty) -> StateLayout
        # sync-free path: sequence numbers, the calls whose mail has not been read yet, byte sizes
        self.seq = 0
        self.pending = collections.deque()   # (seq, slot, cap, (H, W))
        self.opstate = {}                    # (cap, H, W) -> bytes
        # calls made under stream capture re-write their mailbox at every replay: each gets a slot of
        # this second ring (allocated here: pinning memory is not allowed while a stream captures)
        self.graph_mail = torch.zeros((_RING, 8), dtype=torch.int32).pin_memory()
        self.graph_mail_np = self.graph_mail.numpy()
        self.graph_used = 0
        self.dev = dev

    def opstate_bytes(self, cap: int, H: int, W: int) -> int:
        key = (cap, H, W)
        b = self.opstate.get(key)
        if b is None:
            b = int(_lib.load().gs2d_opstate_bytes(cap, H, W))
            if b == 0:
                _lib.check(1, "opstate_bytes")
            if len(self.opstate) > 256:
                self.opstate.clear()
            self.opstate[key] = b
        return b

    def _read(self, seq: int, slot: int, cap: int, key) -> bool:
        """consume one landed mailbox record: capacity hint, overflow -> regrow (returns True)"""
        m = self.mail_np[slot]
        num, overflow = int(m[0]), int(m[1])
        self.last_num_intersects = num
        if overflow or num * _HEADROOM // 2 > cap or key not in self.cap:
            self.cap[key] = max(self.cap.get(key, 0), _round_cap(_HEADROOM * num))
        return bool(overflow)

    def poll(self, wait_for: int = 0):
        """read every mailbox record that has landed (and, wait_for != 0, wait for that call's);
        raises CapacityOverflow if one of them reports that its call outgrew its buffer"""
        overflowed = None
        while self.pending:
            seq, slot, cap, key = self.pending[0]
            if int(self.mail_np[slot][4]) != seq:
                if wait_for and seq <= wait_for:
                    self._wait(slot, seq)
                else:
                    break
            self.pending.popleft()
            if self._read(seq, slot, cap, key) and overflowed is None:
                overflowed = (seq, int(self.mail_np[slot][0]), cap)
        if overflowed is not None:
            raise CapacityOverflow(
                f"rasterize_gaussians_sum: call #{overflowed[0]} had {overflowed[1]} intersections but room for "
                f"{overflowed[2]}; its image was filled with NaN.  The buffer has been regrown: run the iteration again "
                "(gsvc_b200.options.sync_free = False makes every call wait for its own count instead)")

    def _wait(self, slot: int, seq: int):
        rec = self.mail_np[slot]
        spins = 0
        while int(rec[4]) != seq:
            spins += 1
            if spins > 200_000:
                torch.cuda.current_stream(self.dev).synchronize()
                if int(rec[4]) != seq:
                    raise RuntimeError("rasterize_gaussians_sum: the forward finished without reporting")

    def layout(self, n: int, cap: int, tx: int, ty: int) -> _lib.StateLayout:
        key = (n, cap, tx, ty)
        lay = self.layouts.get(key)
        if lay is None:
            lay = _lib.StateLayout()
            _lib.check(_lib.load().gs2d_state_layout_for(n, cap, tx, ty, lay), "state_layout_for")
            if len(self.layouts) > 256:
                self.layouts.clear()
            self.layouts[key] = lay
        return lay


_ctx_by_device: dict = {}


def _device_ctx(dev: torch.device) -> _DeviceCtx:
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    c = _ctx_by_device.get(idx)
    if c is None:
        c = _ctx_by_device[idx] = _DeviceCtx(torch.device("cuda", idx))
    return c


_ones_cache: dict = {}


def _ones(c: int, dev: torch.device) -> Tensor:
    """the default background (rasterize_sum.py:57-64 builds a new ones(C) per call); read-only here"""
    key = (c, dev.index if dev.index is not None else torch.cuda.current_device() if dev.type == "cuda" else -1)
    t = _ones_cache.get(key)
    if t is None:
        t = _ones_cache[key] = torch.ones(c, dtype=torch.float32, device=dev)
    return t


def _round_cap(v: int) -> int:
    return max(_CAP_QUANTUM, (int(v) + _CAP_QUANTUM - 1) // _CAP_QUANTUM * _CAP_QUANTUM)


def rasterize_gaussians_sum(
    xys: Tensor,
    depths: Tensor,
    radii: Tensor,
    conics: Tensor,
    num_tiles_hit: Tensor,
    colors: Tensor,
    opacity: Tensor,
    img_height: int,
    img_width: int,
    BLOCK_H: int = 16,
    BLOCK_W: int = 16,
    background: Optional[Tensor] = None,
    return_alpha: Optional[bool] = False,
) -> Tensor:
    """Accumulated-sum rasterisation of 2D Gaussians (no transmittance, no background term,
    first 256 Gaussians of every tile -- the reference's behaviour).

    Differentiable w.r.t. xys, conics, colors and opacity.  Returns out_img [H,W,3] or
    (out_img, out_alpha [H,W]) when return_alpha is set (out_alpha is all zeros, as in the
    reference where final_Ts is the constant 1)."""
    if colors.dtype == torch.uint8:
        colors = colors.float() / 255   # rasterize_sum.py:53-55

    if background is not None:
        assert background.shape[0] == colors.shape[-1], (
            f"incorrect shape of background color tensor, expected shape {colors.shape[-1]}")
        if background.dtype != torch.float32:
            background = background.float()
    else:
        background = _ones(colors.shape[-1], colors.device)

    if xys.ndimension() != 2 or xys.size(1) != 2:
        raise ValueError("xys must have dimensions (N, 2)")
    if colors.ndimension() != 2:
        raise ValueError("colors must have dimensions (N, D)")

    # set by this package's projections (depths identically zero); trusted only while nobody wrote
    # into the tensor since (torch bumps ._version on in-place ops)
    zero_depth = getattr(depths, "_gs2d_zero_depth", None) == depths._version
    return _RasterizeGaussiansSum.apply(
        xys.contiguous(), depths.contiguous(), radii.contiguous(), conics.contiguous(), num_tiles_hit.contiguous(),
        colors.contiguous(), opacity.contiguous(), img_height, img_width, BLOCK_H, BLOCK_W, background.contiguous(),
        return_alpha, zero_depth,
    )


class _RasterizeGaussiansSum(Function):
    @staticmethod
    def forward(ctx, xys, depths, radii, conics, num_tiles_hit, colors, opacity, img_height, img_width,
                BLOCK_H=16, BLOCK_W=16, background=None, return_alpha=False, zero_depth=False):
        if BLOCK_H != 16 or BLOCK_W != 16:
            # the reference's binning and shared-memory batch are compile-time 16x16
            raise ValueError("rasterize_gaussians_sum supports only BLOCK_H = BLOCK_W = 16")
        channels = int(colors.shape[-1])
        if channels != 3 and not options.nd_channels:
            # the reference dispatches other channel counts to an unbound function
            raise NotImplementedError("rasterize_gaussians_sum supports only 3 colour channels "
                                      "(gsvc_b200.options.nd_channels = True enables 1..32)")
        # SURVEY 8f.4 options; the defaults select the reference-exact 3-channel kernels.  The
        # reference's N-channel kernel has no per-tile limit (forward.cu:664), so neither has ours.
        max_tile = int(options.max_per_tile) if channels == 3 else 0
        nd = channels != 3 or max_tile != GS2D_MAX_PER_TILE
        dev = xys.device
        if not xys.is_cuda:
            raise RuntimeError("rasterize_gaussians_sum needs CUDA tensors (there is no CPU path)")
        for t, nm in ((xys, "xys"), (conics, "conics"), (colors, "colors"), (opacity, "opacity")):
            if t.dtype != torch.float32:
                raise RuntimeError(f"{nm} must be float32")
        H, W = int(img_height), int(img_width)
        n = xys.size(0)
        tx, ty = (W + 15) // 16, (H + 15) // 16
        radii = radii.to(torch.int32)
        num_tiles_hit = num_tiles_hit.to(torch.int32)
        need_grad = any(ctx.needs_input_grad[i] for i in (0, 3, 5, 6))

        dctx = _device_ctx(dev)
        lib = _lib.load()
        ctx.img_width, ctx.img_height = W, H
        ctx.BLOCK_H, ctx.BLOCK_W = BLOCK_H, BLOCK_W
        ctx.num_points = n
        ctx.nd, ctx.channels, ctx.max_tile = nd, channels, max_tile
        ctx.general = False
        ctx.ranked = False

        if not nd and zero_depth and n > 0 and options.binning == "ranked":
            # ---- default: count / scatter binning, order restored in the tile kernels; sync-free ----
            out_img, final_idx, state = _forward_ranked(ctx, dctx, lib, dev, n, H, W, xys, radii, conics, colors,
                                                        opacity, background, need_grad)
            if need_grad:
                ctx.save_for_backward(xys, conics, colors, opacity, state, final_idx)
            else:
                ctx.save_for_backward(xys, conics, colors, opacity)
            ctx.num_intersects = None      # stays on the device (dctx.last_num_intersects: last one the host saw)
            if return_alpha:
                return out_img, torch.zeros(H, W, dtype=torch.float32, device=dev)
            return out_img

        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("rasterize_gaussians_sum: only the default path (3 channels, depths from this package's "
                               "projections, options.binning = 'ranked') can be captured in a CUDA graph")
        general = None
        with _C._guard(dev):
            stream = torch.cuda.current_stream(dev).cuda_stream
            cap = dctx.cap.get((H, W), 0)
            if cap < 1:
                cap = _round_cap(max(4 * n, 65536))
            while True:
                lay = dctx.layout(n, cap, tx, ty)
                state = torch.empty((lay.total_bytes,), dtype=torch.uint8, device=dev)
                out_img = torch.empty((H, W, channels), dtype=torch.float32, device=dev)
                final_idx = torch.empty((H, W), dtype=torch.int32, device=dev) if need_grad else None
                slot = dctx.slot
                dctx.slot = (slot + 1) % _RING
                ev = dctx.events[slot]
                mail = dctx.mail_sync[slot]
                if nd:
                    _lib.check(lib.gs2d_rasterize_sum_nd_fwd(
                        n, channels, xys.data_ptr(), None if zero_depth else depths.data_ptr(), radii.data_ptr(),
                        conics.data_ptr(), num_tiles_hit.data_ptr(), colors.data_ptr(), opacity.data_ptr(), H, W,
                        cap, max_tile, state.data_ptr(), lay.total_bytes, out_img.data_ptr(),
                        None if final_idx is None else final_idx.data_ptr(), mail.data_ptr(), ev.cuda_event,
                        stream), "rasterize_sum_nd_fwd")
                else:
                    _lib.check(lib.gs2d_rasterize_sum_fwd(
                        n, xys.data_ptr(), None if zero_depth else depths.data_ptr(), radii.data_ptr(),
                        conics.data_ptr(), num_tiles_hit.data_ptr(), colors.data_ptr(), opacity.data_ptr(), H, W,
                        cap, state.data_ptr(), lay.total_bytes, out_img.data_ptr(),
                        None if final_idx is None else final_idx.data_ptr(), mail.data_ptr(), ev.cuda_event,
                        stream), "rasterize_sum_fwd")
                ev.synchronize()   # waits for the scan stage only; the rest is already queued
                m = dctx.mail_sync_np[slot]
                num_intersects, overflow, depth_or = int(m[0]), int(m[1]), int(m[2])
                dctx.last_num_intersects = num_intersects
                if overflow:
                    cap = _round_cap(num_intersects + num_intersects // 4)
                    continue
                if num_intersects * 3 > cap * 2 or (H, W) not in dctx.cap:
                    dctx.cap[(H, W)] = max(dctx.cap.get((H, W), 0), cap, _round_cap(2 * num_intersects))
                if depth_or != 0:
                    if nd:
                        raise NotImplementedError("N-channel / untruncated rasterisation needs zero depths "
                                                  "(both projections emit zeros)")
                    general = True   # non-zero depths: keys need the depth bits (64-bit path)
                break

        ctx.num_intersects = num_intersects
        ctx.general = bool(general)

        if num_intersects < 1:
            # rasterize_sum.py:121-129
            out_img = torch.ones(H, W, colors.shape[-1], device=dev) * background
            ctx.save_for_backward(xys, conics, colors, opacity)
        elif general:
            out_img, final_idx, ids_sorted, tile_bins = _forward_general(
                n, num_intersects, xys, depths, radii, conics, num_tiles_hit, colors, opacity, H, W, (tx, ty, 1),
                background)
            ctx.save_for_backward(xys, conics, colors, opacity, ids_sorted, tile_bins, final_idx, background)
        else:
            ctx.cap = cap
            ctx.state_bytes = lay.total_bytes
            if need_grad:
                ctx.save_for_backward(xys, conics, colors, opacity, state, final_idx)
            else:
                ctx.save_for_backward(xys, conics, colors, opacity)

        if return_alpha:
            out_alpha = torch.zeros(H, W, dtype=torch.float32, device=dev)   # 1 - final_Ts, final_Ts == 1
            return out_img, out_alpha
        return out_img

    @staticmethod
    def backward(ctx, v_out_img, v_out_alpha=None):
        saved = ctx.saved_tensors
        xys, conics, colors, opacity = saved[:4]
        n = ctx.num_points
        if ctx.ranked:
            state, final_idx = saved[4], saved[5]
            dev = xys.device
            dctx = _device_ctx(dev)
            if not ctx.captured:
                dctx.poll()      # raises if this forward (or an earlier one) is already known to have overflowed
            v_out_img = v_out_img.contiguous()
            if v_out_img.dtype != torch.float32:
                v_out_img = v_out_img.float()
            grads = torch.empty((n, 12), dtype=torch.float32, device=dev)
            with _C._guard(dev):
                rc = _lib.load().gs2d_rasterize_sum_bwd_ranked(
                    n, ctx.img_height, ctx.img_width, ctx.cap, state.data_ptr(), ctx.state_bytes, final_idx.data_ptr(),
                    v_out_img.data_ptr(), grads.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
                if rc:
                    _lib.check(rc, "rasterize_sum_bwd_ranked")
            return (grads[:, 0:2], None, None, grads[:, 2:5], None, grads[:, 5:8], grads[:, 8:9],
                    None, None, None, None, None, None, None)
        if ctx.num_intersects < 1:
            v_xy = torch.zeros_like(xys)
            v_conic = torch.zeros_like(conics)
            v_colors = torch.zeros_like(colors)
            v_opacity = torch.zeros_like(opacity)
        elif ctx.general:
            ids_sorted, tile_bins, final_idx, background = saved[4:8]
            v_xy, v_conic, v_colors, v_opacity = _C.rasterize_sum_backward(
                ctx.img_height, ctx.img_width, ctx.BLOCK_H, ctx.BLOCK_W, ids_sorted, tile_bins, xys, conics, colors,
                opacity, background, None, final_idx, v_out_img, None)
        else:
            state, final_idx = saved[4], saved[5]
            dev = xys.device
            v_out_img = v_out_img.contiguous()
            if v_out_img.dtype != torch.float32:
                v_out_img = v_out_img.float()
            grads = torch.empty((n, 12), dtype=torch.float32, device=dev)
            with _C._guard(dev):
                if ctx.nd:
                    v_colors = torch.empty((n, ctx.channels), dtype=torch.float32, device=dev)
                    _lib.check(_lib.load().gs2d_rasterize_sum_nd_bwd(
                        n, ctx.channels, ctx.img_height, ctx.img_width, ctx.cap, ctx.max_tile, state.data_ptr(),
                        ctx.state_bytes, colors.data_ptr(), final_idx.data_ptr(), v_out_img.data_ptr(),
                        grads.data_ptr(), v_colors.data_ptr(), torch.cuda.current_stream(dev).cuda_stream),
                        "rasterize_sum_nd_bwd")
                else:
                    _lib.check(_lib.load().gs2d_rasterize_sum_bwd(
                        n, ctx.img_height, ctx.img_width, ctx.cap, state.data_ptr(), ctx.state_bytes,
                        final_idx.data_ptr(), v_out_img.data_ptr(), grads.data_ptr(),
                        torch.cuda.current_stream(dev).cuda_stream), "rasterize_sum_bwd")
                    v_colors = grads[:, 5:8]
            v_xy, v_conic, v_opacity = grads[:, 0:2], grads[:, 2:5], grads[:, 8:9]
        return (v_xy, None, None, v_conic, None, v_colors, v_opacity, None, None, None, None, None, None, None)


def _forward_ranked(ctx, dctx, lib, dev, n, H, W, xys, radii, conics, colors, opacity, background, need_grad):
    """the sync-free forward (module docstring) -> (out_img, final_idx or None, per-call state)"""
    key = (H, W)
    capturing = torch.cuda.is_current_stream_capturing()
    if not capturing:
        dctx.poll()                         # lazily: hints from earlier calls, overflow of an earlier call
    cap = dctx.cap.get(key, 0)
    # a hint learnt from a model of another size says nothing about this one: count again
    first = cap < 1 or (dctx.cap_n.get(key, n) != n and not capturing)
    dctx.cap_n[key] = n
    if cap < 1:
        cap = _round_cap(max(8 * n, 1 << 18) if capturing else max(4 * n, 65536))
    wait = (first or not options.sync_free) and not capturing
    with _C._guard(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        while True:
            nbytes = dctx.opstate_bytes(cap, H, W)
            state = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
            out_img = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
            final_idx = torch.empty((H, W), dtype=torch.int32, device=dev) if need_grad else None
            dctx.seq += 1
            seq = dctx.seq
            if capturing:
                slot = dctx.graph_used % _RING
                dctx.graph_used += 1
                mail_ptr = dctx.graph_mail.data_ptr() + 32 * slot
            else:
                slot = seq % _RING
                if len(dctx.pending) >= _RING - 1:      # the ring is about to wrap: read the oldest first
                    dctx.poll(wait_for=dctx.pending[0][0])
                mail_ptr = dctx.mail_ptr + 32 * slot
            rc = lib.gs2d_rasterize_sum_fwd_ranked(
                n, xys.data_ptr(), radii.data_ptr(), conics.data_ptr(), colors.data_ptr(), opacity.data_ptr(),
                background.data_ptr(), H, W, cap, state.data_ptr(), nbytes, out_img.data_ptr(),
                None if final_idx is None else final_idx.data_ptr(), mail_ptr, seq, stream)
            if rc:
                _lib.check(rc, "rasterize_sum_fwd_ranked")
            if capturing:
                break
            if not wait:
                dctx.pending.append((seq, slot, cap, key))
                break
            # first call of this image size (or options.sync_free = False): learn the count now
            dctx._wait(slot, seq)
            if dctx._read(seq, slot, cap, key):
                cap = dctx.cap[key]             # regrown by _read: redo with room for everything
                continue
            break
    ctx.ranked, ctx.captured = True, capturing
    ctx.cap, ctx.state_bytes = cap, nbytes
    return out_img, final_idx, state


def check_overflow(device=None) -> None:
    """Read every mailbox record that has landed on `device` (default: current) -- including those of
    calls replayed from a CUDA graph, which no Python code sees -- and raise CapacityOverflow if one
    of them outgrew its buffer.  Cheap (pinned host memory); call it now and then around graph replays."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    dctx = _device_ctx(dev)
    dctx.poll()
    for box in dctx.graph_mail_np[:min(dctx.graph_used, _RING)]:
        if int(box[1]) != 0:
            num = int(box[0])
            box[1] = 0
            for key in list(dctx.cap):
                dctx.cap[key] = max(dctx.cap[key], _round_cap(_HEADROOM * num))
            raise CapacityOverflow(f"a captured rasterize_gaussians_sum call had {num} intersections, more than its "
                                   "buffer held (image filled with NaN): capture the graph again")


def _forward_general(n, num_intersects, xys, depths, radii, conics, num_tiles_hit, colors, opacity, H, W,
                     tile_bounds, background):
    """Step-by-step path with the reference's 64-bit (tile | depth) keys, for callers that pass
    non-zero depths (GSVC never does: both projections emit zeros)."""
    cum = _C.cumsum_tiles_hit(num_tiles_hit)
    isect_ids, gaussian_ids = _C.map_gaussian_to_intersects(n, num_intersects, xys, depths, radii, cum, tile_bounds)
    keys_sorted, ids_sorted = _C.sort_pairs(isect_ids, gaussian_ids, 0, 64)
    tiles = tile_bounds[0] * tile_bounds[1]
    tile_bins = _C.get_tile_bin_edges(num_intersects, keys_sorted, rows=max(num_intersects, tiles))
    out_img, _final_Ts, final_idx = _C.rasterize_sum_forward(
        tile_bounds, (16, 16, 1), (W, H, 1), ids_sorted, tile_bins, xys, conics, colors, opacity, background)
    return out_img, final_idx, ids_sorted, tile_bins
