"""One-process-per-GPU partitioning of the rasterizer over a single NVSwitch box.

Two workloads shard naturally; nothing else is distributed:

* render-only: the image's TILE ROWS are split across ranks, Gaussians and camera replicated.
  Each rank runs the full per-Gaussian stage (radii are identical everywhere), bins and blends
  only the instances that fall in its band (`tile_row_begin/end` of the C ABI, include/gsr.h) and
  the bands are exchanged with one all-gather.  No per-Gaussian collective.
* multi-view training: one camera per rank, parameters replicated; each rank's per-Gaussian
  gradients are summed with ONE all-reduce over the flat gradient arena the backward writes
  (the torch binding returns all gradients as views of one buffer), `radii` with a max-reduce.

The collective backend is whatever `torch.distributed` was initialised with: NCCL over
NVLink 5 / NVSwitch on the B200 box, gloo in the CPU tests of the host-side logic.
The reference has no multi-GPU path (train.py is single-device); this is an extension around the
drop-in API and does not change it.
"""
from typing import List, NamedTuple, Optional, Sequence, Tuple

import os
import sys
import torch
import torch.distributed as dist

TILE = 16   # DGR config.h:16-17 (BLOCK_X, BLOCK_Y)


class BandedRasterizationSettings(NamedTuple):
    """GaussianRasterizationSettings (DGR/diff_gaussian_rasterization/__init__.py:168-180) plus the
    tile-row band this rank renders."""
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    tile_rows: Tuple[int, int] = (0, 0)


def tile_rows(image_height: int) -> int:
    return (image_height + TILE - 1) // TILE


def split_tile_rows(n_rows: int, world: int, row_weights: Optional[Sequence[float]] = None) -> List[Tuple[int, int]]:
    """Contiguous bands [begin, end) of tile rows, one per rank, covering [0, n_rows) exactly once.

    Without weights rows are split as evenly as possible (earlier ranks take the remainder).  With
    `row_weights` (e.g. instances per tile row from a previous frame) the cut points are the weighted
    quantiles, so every rank blends about the same number of instances.  A rank may get an empty band
    when world > n_rows.
    """
    if world <= 0:
        raise ValueError("world must be positive")
    if n_rows < 0:
        raise ValueError("n_rows must be non-negative")
    if row_weights is None:
        base, rem = divmod(n_rows, world)
        bands, b = [], 0
        for r in range(world):
            e = b + base + (1 if r < rem else 0)
            bands.append((b, e))
            b = e
        return bands
    w = [max(float(x), 0.0) for x in row_weights]
    if len(w) != n_rows:
        raise ValueError("row_weights must have one entry per tile row")
    total = sum(w)
    if total <= 0.0:
        return split_tile_rows(n_rows, world)
    cuts, acc, row = [0], 0.0, 0
    for r in range(1, world):
        target = total * r / world
        while row < n_rows and acc + 0.5 * w[row] < target:
            acc += w[row]
            row += 1
        cuts.append(row)
    cuts.append(n_rows)
    return [(cuts[i], cuts[i + 1]) for i in range(world)]


def band_pixel_rows(band: Tuple[int, int], image_height: int) -> Tuple[int, int]:
    return min(band[0] * TILE, image_height), min(band[1] * TILE, image_height)


def banded_settings(settings, band: Tuple[int, int]) -> BandedRasterizationSettings:
    return BandedRasterizationSettings(*tuple(settings)[:12], tile_rows=(int(band[0]), int(band[1])))


def gather_bands(local_rows: torch.Tensor, bands: Sequence[Tuple[int, int]], image_height: int,
                 group=None) -> torch.Tensor:
    """All-gather the per-rank (3, rows_r, W) bands into the full (3, H, W) image on every rank.

    Bands differ in height, so each rank pads to the tallest band; the exchange is a single
    all_gather_into_tensor (99.5 MB at 4K in total)."""
    world = dist.get_world_size(group)
    heights = [band_pixel_rows(b, image_height) for b in bands]
    hmax = max(e - b for b, e in heights)
    C, _, W = local_rows.shape
    send = local_rows.new_zeros((C, hmax, W))
    send[:, : local_rows.shape[1]] = local_rows
    recv = local_rows.new_empty((world, C, hmax, W))
    dist.all_gather_into_tensor(recv.view(-1), send.view(-1), group=group)
    out = local_rows.new_empty((C, image_height, W))
    for r, (b, e) in enumerate(heights):
        out[:, b:e] = recv[r, :, : e - b]
    return out


def render_tile_row_sharded(rasterizer_cls, settings, *, means3D, means2D, opacities, shs=None, colors_precomp=None,
                            scales=None, rotations=None, cov3D_precomp=None, row_weights=None, group=None,
                            gather=True):
    """Render-only forward with the image's tile rows split over the ranks of `group`.

    Returns (image, radii, band).  `image` is the full (3,H,W) frame when gather=True, else this
    rank's (3, rows, W) band.  radii are computed on the full image on every rank."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    H = int(settings.image_height)
    bands = split_tile_rows(tile_rows(H), world, row_weights)
    band = bands[rank]
    with torch.no_grad():
        if band[1] > band[0]:
            rast = rasterizer_cls(raster_settings=banded_settings(settings, band))
            color, radii = rast(means3D=means3D, means2D=means2D, opacities=opacities, shs=shs,
                                colors_precomp=colors_precomp, scales=scales, rotations=rotations,
                                cov3D_precomp=cov3D_precomp)
        else:   # more ranks than tile rows: this rank only contributes radii
            rast = rasterizer_cls(raster_settings=banded_settings(settings, (0, 1)))
            color, radii = rast(means3D=means3D, means2D=means2D, opacities=opacities, shs=shs,
                                colors_precomp=colors_precomp, scales=scales, rotations=rotations,
                                cov3D_precomp=cov3D_precomp)
        p0, p1 = band_pixel_rows(band, H)
        local = color[:, p0:p1]
        if not gather or world == 1:
            return (color if world == 1 else local), radii, band
        return gather_bands(local.contiguous(), bands, H, group), radii, band


def flat_span(tensors: Sequence[torch.Tensor]) -> Optional[torch.Tensor]:
    """If every tensor is a dense view into ONE storage (the backward's gradient arena), return the 1-D
    tensor that covers them all, else None."""
    ts = [t for t in tensors if t is not None and t.numel() > 0]
    if not ts:
        return None
    st = ts[0].untyped_storage()
    lo, hi = None, None
    for t in ts:
        if t.untyped_storage().data_ptr() != st.data_ptr() or not t.is_contiguous() or t.dtype != ts[0].dtype:
            return None
        b = t.storage_offset()
        e = b + t.numel()
        lo = b if lo is None else min(lo, b)
        hi = e if hi is None else max(hi, e)
    return torch.empty(0, dtype=ts[0].dtype, device=ts[0].device).set_(st, lo, (hi - lo,), (1,))


def allreduce_gradients(grads: Sequence[torch.Tensor], group=None, average=False) -> int:
    """Sum per-Gaussian gradients over the ranks (one camera each).  One collective when the gradients
    share the backward's flat arena, else one per tensor.  Returns the number of collectives issued."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return 0
    world = dist.get_world_size(group)
    flat = flat_span(grads)
    n = 0
    if flat is not None:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            flat.div_(world)
        return 1
    for g in grads:
        if g is None or g.numel() == 0:
            continue
        dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)
        if average:
            g.div_(world)
        n += 1
    return n


def allreduce_radii(radii: torch.Tensor, group=None) -> torch.Tensor:
    """max_radii2D needs the elementwise max over views (train.py:157)."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(radii, op=dist.ReduceOp.MAX, group=group)
    return radii


def cameras_for_rank(n_views: int, rank: int, world: int) -> List[int]:
    """View c goes to rank c mod world: 8/4/2/1 views per GPU per step at 1/2/4/8 GPUs."""
    return [c for c in range(n_views) if c % world == rank]


def exchange_block_floats(P: int) -> int:
    """Floats of one view's exchange block: campos (4) | dL/dRGB (P,3), padded so that consecutive blocks stay 16-byte aligned
    for any P (the SH kernel reads them with 16-byte loads)."""
    return (4 + 3 * int(P) + 3) & ~3


def exchange_chunks(P: int, n: int) -> List[Tuple[int, int]]:
    """Ranges [begin, end) of Gaussians for the pipelined exchange: n ranges (fewer for small P), every begin a multiple of 128
    (the SH kernel's row block; it also keeps every range's float offset a multiple of 4)."""
    P, n = int(P), max(int(n), 1)
    step = ((P + n - 1) // n + 127) // 128 * 128
    out, b = [], 0
    while b < P:
        out.append((b, min(b + step, P)))
        b += step
    return out or [(0, 0)]


class MultiViewExchange:
    """Gradient exchange for multi-view training: `views_per_rank` cameras per rank and step (SURVEY.md 8(e), C5:
    8 cameras per step = 8/4/2/1 views per GPU at 1/2/4/8 GPUs).

    Per view the SH gradient is an outer product basis(dir_v) (x) dL/dRGB_v, so the 12*M bytes per Gaussian of dL_dsh never
    cross the links.  Every local backward publishes 12 bytes per Gaussian (clamp-masked dL/dRGB) plus its camera position
    into its slot of this rank's block and skips dL_dsh.  The LAST local backward of a step then
      * side stream: all-gathers the blocks (NCCL; or maps the peers' blocks, mode "peer") as soon as its dL/dRGB is
        published -- beside the per-Gaussian backward the main stream is still running -- and ONE kernel
        (`gsr_multiview_sh_grad`) rebuilds every view's SH basis and writes the dL_dsh summed over ALL views of ALL ranks;
      * `end_step()` sums the 11 other live floats per Gaussian (means3D, opacity, scale, rotation: the parameters' .grad
        after autograd has accumulated the local views) with one NCCL all-reduce that overlaps with that kernel.
    means2D.grad is NOT exchanged: the reference's densification statistic is a per-view norm (scene/gaussian_model.py:599-602),
    which each rank accumulates for its own views exactly as the reference does; `reduce_densification_stats(model)` sums
    the accumulators over the ranks (and max-reduces max_radii2D, train.py:157) right before densify_and_prune.
    NVLink traffic per GPU and step at 8 ranks, P = 3 M: 252 MB all-gather + a 132 MB all-reduce, instead of a 744 MB all-reduce.
    """

    def __init__(self, capacity, device, group=None, mode=None, views_per_rank=1):
        import os
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.views_per_rank = int(views_per_rank)
        if self.views_per_rank < 1 or self.world * self.views_per_rank > 16:      # GSR_MAX_VIEWS
            raise ValueError("views_per_rank * world must be in 1..16")
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        # how the 12 B/Gaussian blocks travel: "allgather" (default) = NCCL all-gather into local staging, then the SH
        # kernel on local data; "peer" = plain peer loads inside the SH kernel (symmetric memory).  Measured on 8 B200:
        # 5.18 vs 5.36 ms per step (profiles/r01_multigpu_c5.md) -- the all-gather goes through NVLS.
        self.mode = mode or os.environ.get("GSR_MULTIVIEW_MODE", "allgather")
        self._symm = None
        if self.mode != "allgather" or os.environ.get("GSR_MULTIVIEW_SMALL", "nccl") == "multimem":
            import torch.distributed._symmetric_memory as symm_mem
            self._symm = symm_mem
        # high priority: the exchange's few CTAs must be dispatched WHILE the per-Gaussian backward (tens of thousands of
        # short blocks, launched earlier) is running, not after it.  For the NCCL part set TORCH_NCCL_HIGH_PRIORITY=1
        # before init_process_group (bench.py does).
        self.side = torch.cuda.Stream(device=self.device, priority=-1)
        # one level above the side stream: the reduction's small CTAs must not queue behind the SH kernel's blocks
        self.comm = torch.cuda.Stream(device=self.device, priority=-2)
        self.ev_drgb = torch.cuda.Event()
        self.ev_done = torch.cuda.Event()
        self.ev_small = torch.cuda.Event()
        self.ev_ag = torch.cuda.Event()
        self.ev_chunk = [torch.cuda.Event() for _ in range(4)]     # GSR_MAX_CHUNKS
        with torch.cuda.device(self.device):
            self.ev_drgb.record()                 # torch creates the cudaEvent lazily; the C ABI needs the handle
            self.ev_done.record()
            self.ev_small.record()
            self.ev_ag.record()
            for e in self.ev_chunk:
                e.record()
        # GSR_MV_TIMELINE=1: timed events around every phase of the exchange (diagnostics; see timeline_report)
        self._tl = [] if os.environ.get("GSR_MV_TIMELINE") == "1" else None
        self._tl_cur = None
        # "allgather" mode, GSR_MV_AG_CHUNKS = n > 1: the blocks travel in n ranges of Gaussians and each range's SH rows are
        # written while the next range is still on the links.  Built, correct (2-GPU test) and measured SLOWER on 8 B200
        # (3.75 ms per step with 4 ranges, 3.81 with 8, against 3.36 with one all-gather, profiles/r02_multigpu.md): the step
        # is link-bound from the publish point on, smaller messages use the links less well, and the early SH kernels take
        # SMs and HBM bandwidth from the per-Gaussian backward that gates the all-reduce.  Default: one all-gather.
        self.ag_chunks = int(os.environ.get("GSR_MV_AG_CHUNKS", "1"))
        self.campos_local = torch.zeros(self.views_per_rank, 4, dtype=torch.float32, device=self.device)
        self.campos_all = torch.zeros(self.world * self.views_per_rank, 4, dtype=torch.float32, device=self.device)
        self._view = 0                  # local views whose backward has run in the current step
        self._pending_done = False      # the SH kernel of this step is still running on the side stream
        self._alloc(int(capacity))

    def _mark(self, name, stream=None):
        if self._tl_cur is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream if stream is not None else torch.cuda.current_stream(self.device))
            self._tl_cur.append((name, e))

    def timeline_report(self, skip=5):
        """Mean offsets (ms) of each phase boundary from the start of the backward, over the recorded steps."""
        if not self._tl:
            return {}
        torch.cuda.synchronize(self.device)
        acc = {}
        for marks in self._tl[skip:]:
            if not marks:
                continue
            t0 = marks[0][1]
            for name, e in marks[1:]:
                acc.setdefault(name, []).append(t0.elapsed_time(e))
        return {k: sum(v) / len(v) for k, v in acc.items()}

    def _alloc(self, capacity):
        self.capacity = capacity
        n = exchange_block_floats(capacity) * self.views_per_rank
        # (P,14) small-gradient rows.  "multimem": symmetric allocation bound to an NVLS multicast address, summed by
        # this package's own two-shot kernel through the switch (gsr_multimem_allreduce); "nccl": the parameters' .grad,
        # NCCL all-reduce.  multimem needs NVSwitch multicast support and one view per rank; it falls back by itself
        # (collectively: every rank sees the same answer from the rendezvous).
        self.pack_handle, self.pack_mc, self.pack = None, 0, None
        # Default "nccl": on 8 B200 the multimem kernel needs 0.3-0.4 ms for its 21 MB slice and the step ends later
        # (4.34 ms vs 4.01 ms, profiles/r01_multigpu_c5.md); it stays selectable.
        want = os.environ.get("GSR_MULTIVIEW_SMALL", "nccl")
        if want == "multimem" and self.views_per_rank == 1:
            try:
                self.pack = self._symm.empty(14 * capacity, dtype=torch.float32, device=self.device)
                self.pack.zero_()
                self.pack_handle = self._symm.rendezvous(self.pack, self.group)
                self.pack_mc = int(getattr(self.pack_handle, "multicast_ptr", 0) or 0)
            except Exception as e:      # no symmetric memory / no multicast on this machine
                self.pack_handle, self.pack_mc = None, 0
                if self.rank == 0:
                    print("gsr_multi_gpu: multimem exchange unavailable (%s); using NCCL for the small gradients" % (e,), file=sys.stderr)
        if not self.pack_mc:
            self.pack_handle, self.pack = None, None
        self.small_mode = "multimem" if self.pack_mc else "nccl"
        self.stage_mc = 0
        if self.mode == "multicast":
            # the all-gather done by the switch: the staging buffer is a symmetric allocation bound to an NVLS multicast
            # address; every rank stores its block ONCE through that address (gsr_multimem_broadcast) and the switch
            # replicates it into all copies.  Falls back to the NCCL all-gather, collectively, where multicast is missing.
            try:
                self.block = torch.zeros(n, dtype=torch.float32, device=self.device)
                self.stage = self._symm.empty(self.world * n, dtype=torch.float32, device=self.device)
                self.stage.zero_()
                self.handle = self._symm.rendezvous(self.stage, self.group)
                self.stage_mc = int(getattr(self.handle, "multicast_ptr", 0) or 0)
            except Exception as e:
                self.stage_mc = 0
                if self.rank == 0:
                    print("gsr_multi_gpu: multicast staging unavailable (%s); using the NCCL all-gather" % (e,), file=sys.stderr)
            if not self.stage_mc:
                self.mode = "allgather"
            else:
                self.peer_ptrs = None
        if self.mode == "multicast":
            pass
        elif self.mode == "allgather":
            self.block = torch.zeros(n, dtype=torch.float32, device=self.device)
            self.stage = torch.zeros(self.world * (n + 64 * self.views_per_rank), dtype=torch.float32, device=self.device)
            self.handle, self.peer_ptrs = None, None
        else:
            self.block = self._symm.empty(n, dtype=torch.float32, device=self.device)
            self.block.zero_()
            self.handle = self._symm.rendezvous(self.block, self.group)
            self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
            self.stage = None

    # ---- called by _RasterizeGaussians.backward ------------------------------------------------------------------
    def backward_extra(self, P, campos):
        """Before the backward's kernels: (this view's exchange block with campos written, skip_dsh, event handle, chunk
        events, small pack)."""
        if self._view >= self.views_per_rank:
            raise RuntimeError("MultiViewExchange: more than views_per_rank = %d backward passes in one step; call end_step() "
                               "after the step's last backward" % self.views_per_rank)
        if P > self.capacity:                     # densification grew the model: same P on every rank, so this is collective
            self._alloc(int(P * 1.25) + 1024)
        nb = exchange_block_floats(P)
        blk = self.block[self._view * nb:(self._view + 1) * nb]
        blk[:3].copy_(campos.reshape(-1)[:3], non_blocking=True)
        self.campos_local[self._view, :3].copy_(campos.reshape(-1)[:3], non_blocking=True)
        last = self._view == self.views_per_rank - 1
        if self._tl is not None and last:
            self._tl_cur = []
            self._tl.append(self._tl_cur)
            self._mark("backward_start")
        # chunk events / packed rows only matter for the backward the exchange runs beside: the step's last one
        return (blk, True, int(self.ev_drgb.cuda_event) if last else 0, [int(e.cuda_event) for e in self.ev_chunk] if last else [],
                self.pack if (self.small_mode == "multimem" and last) else None)

    def exchange(self, C, means3D, sh_degree, grad_sh, small_grads):
        """After a local backward's kernels are queued.  Returns the SH gradient autograd gets for this view: None for all but
        the step's last local view (their share arrives with the last one), the sum over ALL views for the last."""
        self._view += 1
        if self._view < self.views_per_rank:
            return None
        P = means3D.shape[0]
        nb = exchange_block_floats(P)
        n = nb * self.views_per_rank
        main = torch.cuda.current_stream(self.device)
        # side stream: starts as soon as dL/dRGB is published (event recorded by gsr_backward right after the blending
        # backward), i.e. it runs beside the per-Gaussian backward that the main stream is still executing
        self._mark("per_gaussian_backward_done", main)
        self.side.wait_event(self.ev_drgb)
        with torch.cuda.stream(self.side):
            self._mark("drgb_published", self.side)
            if self.mode == "allgather" and self.ag_chunks > 1 and P >= 4096:
                # pipelined: the camera positions first (tiny), then per range of Gaussians and local view one all-gather, all
                # queued at once from the comm stream (NCCL runs them in order); the side stream writes each range's SH rows
                # as soon as that range has arrived, while the next range is on the links.
                k = self.views_per_rank
                chunks = exchange_chunks(P, self.ag_chunks)
                self.comm.wait_event(self.ev_drgb)
                works, bufs = [], []
                with torch.cuda.stream(self.comm):
                    dist.all_gather_into_tensor(self.campos_all.view(-1), self.campos_local.view(-1), group=self.group)
                    off = 0
                    for (b, e) in chunks:
                        ln = (3 * (e - b) + 3) & ~3                      # floats per view of this range, 16-byte padded
                        per_view = []
                        for v in range(k):
                            dst = self.stage[off: off + self.world * ln]
                            src = self.block[v * nb + 4 + 3 * b: v * nb + 4 + 3 * b + ln]
                            w = dist.all_gather_into_tensor(dst, src, group=self.group, async_op=True)
                            per_view.append((dst, w))
                            off += self.world * ln
                        bufs.append((b, e, ln, per_view))
                    self.ev_ag.record(self.comm)
                # campos_all rows are in rank-major order (rank r, slot v) -> the same order as the pointers below
                for ci, (b, e, ln, per_view) in enumerate(bufs):
                    for _, w in per_view:
                        w.wait()                                         # the side stream waits for this range only
                    if ci == len(bufs) - 1:
                        self._mark("allgather_done", self.side)
                    ptrs = [per_view[v][0].data_ptr() + 4 * (r * ln) - 4 * (4 + 3 * b) for r in range(self.world) for v in range(k)]
                    C.multiview_sh_grad_range(ptrs, self.campos_all, means3D, grad_sh, int(sh_degree), b, e)
                self._mark("sh_kernel_done", self.side)
            elif self.mode == "allgather":
                out = self.stage[: self.world * n]
                dist.all_gather_into_tensor(out, self.block[:n], group=self.group)
                self.ev_ag.record(self.side)
                self._mark("allgather_done", self.side)
                ptrs = [out[i * nb:].data_ptr() for i in range(self.world * self.views_per_rank)]
                C.multiview_sh_grad(ptrs, means3D, grad_sh, int(sh_degree))
                self._mark("sh_kernel_done", self.side)
            elif self.mode == "multicast":
                C.multimem_broadcast(self.stage_mc + 4 * self.rank * n, self.block, n)
                self.handle.barrier(channel=0)    # every rank's stores have left; all copies hold all blocks
                self.ev_ag.record(self.side)
                self._mark("allgather_done", self.side)
                base = self.stage.data_ptr()
                ptrs = [base + 4 * i * nb for i in range(self.world * self.views_per_rank)]
                C.multiview_sh_grad(ptrs, means3D, grad_sh, int(sh_degree))
                self._mark("sh_kernel_done", self.side)
                self.handle.barrier(channel=1)    # nobody stores the next step's block over one a peer is still reading
            else:
                self.handle.barrier(channel=0)    # every rank's dL/dRGB and campos are in place
                self.ev_ag.record(self.side)
                # start with the next rank's block so that the ranks do not all pull from the same GPU at once
                ptrs = [self.peer_ptrs[(self.rank + 1 + i) % self.world] + 4 * v * nb
                        for i in range(self.world) for v in range(self.views_per_rank)]
                C.multiview_sh_grad(ptrs, means3D, grad_sh, int(sh_degree))
                self.handle.barrier(channel=1)    # nobody overwrites a block a peer is still reading
            self.ev_done.record(self.side)
        self._pending_done = True
        if self.small_mode == "multimem":
            # one view per rank: rows of the symmetric (P,14) block, summed through the switch by this package's own kernel
            # (multimem.ld_reduce + multimem.st on this rank's 1/world slice), then scattered into the five tensors
            g3d, g2d, gop, gsc, grot = small_grads
            g2d_own = g2d.clone()                 # means2D is not an exchanged quantity: keep this view's
            self.comm.wait_event(self.ev_ag)
            self.comm.wait_event(self.ev_chunk[-1])          # the per-Gaussian backward has written every row
            with torch.cuda.stream(self.comm):
                self._mark("small_start", self.comm)
                self.pack_handle.barrier(channel=2)          # ... on every rank
                C.multimem_allreduce(self.pack_mc, 0, 14 * int(P), self.rank, self.world, self.device.index)
                self._mark("small_reduced", self.comm)
                self.pack_handle.barrier(channel=3)          # every rank's slice is reduced and broadcast
                C.unpack_small_grads(0, int(P), self.pack, g3d, g2d, gop, gsc, grot)
            self.ev_small.record(self.comm)
            main.wait_event(self.ev_small)
            g2d.copy_(g2d_own)
        return grad_sh

    # ---- called by the training loop -----------------------------------------------------------------------------
    def end_step(self, small_params):
        """After the step's last `backward()`: sums the .grad of `small_params` (means3D, opacities, scales, rotations --
        whatever is passed, NOT means2D and not the SH tensor) over the ranks with one all-reduce when they share the
        backward's flat arena (else one per tensor), then joins the SH kernel.  Collective."""
        if self._view != self.views_per_rank:
            raise RuntimeError("MultiViewExchange.end_step: %d of %d local views have run" % (self._view, self.views_per_rank))
        main = torch.cuda.current_stream(self.device)
        if self.small_mode != "multimem":
            grads = [p.grad for p in small_params if p.grad is not None]
            # The links carry one thing at a time: first the all-gather (it gates the SH kernel, the longest tail), then
            # this sum -- NCCL runs it behind the all-gather on its own stream; it overlaps with the SH kernel.
            span = flat_span(grads)
            if span is not None and span.numel() <= sum(g.numel() for g in grads) + 16 * len(grads):
                dist.all_reduce(span, op=dist.ReduceOp.SUM, group=self.group)
            else:
                for g in grads:
                    dist.all_reduce(g, op=dist.ReduceOp.SUM, group=self.group)
        self._mark("allreduce_done", main)
        if self._pending_done:
            main.wait_event(self.ev_done)
            self._pending_done = False
        self._mark("exchange_done", main)
        self._view = 0

    def reduce_densification_stats(self, model):
        """Right before densify_and_prune: every rank has accumulated xyz_gradient_accum / denom (per-view norms and counts,
        scene/gaussian_model.py:599-602) and max_radii2D (train.py:157) for ITS views; sums / max-reduces them so that all
        ranks take the same decisions.  Collective; 12 bytes per Gaussian, once per densification interval."""
        dist.all_reduce(model.xyz_gradient_accum, op=dist.ReduceOp.SUM, group=self.group)
        dist.all_reduce(model.denom, op=dist.ReduceOp.SUM, group=self.group)
        dist.all_reduce(model.max_radii2D, op=dist.ReduceOp.MAX, group=self.group)


def enable_multiview_training(capacity, device, group=None, mode=None, views_per_rank=1):
    """Install the fused exchange: every rank runs `views_per_rank` forward/backward passes per step (its cameras) and then
    calls `ex.end_step([means3D, opacities, scales, rotations])`; the SH gradient of the step's last backward and the .grad
    of those parameters are then summed over all views of all ranks of `group`.  Collective.  Returns the exchange object."""
    import diff_gaussian_rasterization as dgr
    ex = MultiViewExchange(capacity, device, group, mode, views_per_rank)
    dgr.set_multiview_exchange(ex)
    return ex


def disable_multiview_training():
    import diff_gaussian_rasterization as dgr
    dgr.set_multiview_exchange(None)
