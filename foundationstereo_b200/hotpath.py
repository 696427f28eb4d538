"""The hot path end to end, as core/foundation_stereo.py:207-242 strings it together:
volumes -> geometry-encoding volume -> `iters` x (lookup + update block).  This is the
call `bench.py` times and the shape of what `patch.install` makes the reference model do.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch

from .geometry import Combined_Geo_Encoding_Volume
from .submodule import build_combined_volume
from .update import SelectiveUpdateBlock

Tensor = torch.Tensor


class StereoHotPath:
    """Owns the update-block weights; `run` processes one batch of pairs.

    Inputs are what the out-of-scope networks of the reference hand to the path:
      fl, fr   (B,C,H,W)    1/4-res features            (core/foundation_stereo.py:201-204)
      pl, pr   (B,12,H,W)   proj_cmb outputs            (:208-209)
      volume   (B,28,D,H,W) filtered cost volume        (:213-215)
      disp     (B,1,H,W)    initial disparity           (:218-220)
      net, inp, att         hidden / context / attention lists at 1/4, 1/8, 1/16 (:222-227)
    """

    def __init__(self, state_dict: Dict[str, Tensor], max_disp: int, corr_levels: int = 2, corr_radius: int = 4,
                 num_groups: int = 8, device: str | torch.device = "cuda", corr_band: bool = False):
        self.device = torch.device(device)
        self.D = max_disp // 4
        self.L = corr_levels
        self.r = corr_radius
        self.G = num_groups
        # opt-in banded init correlation over disparities [0, D] (geometry.Combined_Geo_Encoding_Volume, corr_band=)
        self.corr_band = self.D if corr_band else None
        self.update_block = SelectiveUpdateBlock(state_dict, self.device)
        self.dx = torch.linspace(-corr_radius, corr_radius, 2 * corr_radius + 1).reshape(1, 1, -1, 1)
        self._ws = {}
        # bench.py: CUDA events around the refinement loop of every run() (the loop's share of a timed step)
        self.record_loop = False
        self.loop_events: List[Tuple[torch.cuda.Event, torch.cuda.Event]] = []

    @torch.no_grad()
    def run(self, fl: Tensor, fr: Tensor, pl: Tensor, pr: Tensor, volume: Tensor, disp: Tensor,
            net: Sequence[Tensor], inp: Sequence[Tensor], att: Sequence[Tensor], iters: int
            ) -> Tuple[Tensor, Tensor, Tensor]:
        """-> (final 1/4-res disparity (B,1,H,W) fp32, mask features of the last step,
        the (B,32,D,H,W) gwc+concat volume)."""
        B, _, H, W = fl.shape
        comb = build_combined_volume(fl, fr, pl, pr, self.D, self.G)
        key = (B, volume.shape[1], volume.shape[2], H, W, fr.shape[3])
        ws = self._ws.get(key)
        if ws is None:  # level buffers with stable addresses: the per-iteration graph is captured once
            ws = self._ws[key] = Combined_Geo_Encoding_Volume.make_workspace(B, volume.shape[1], volume.shape[2], H, W,
                                                                           fr.shape[3], self.L, fl.device,
                                                                           corr_band=self.corr_band, radius=self.r)
        geo = Combined_Geo_Encoding_Volume(fl.float(), fr.float(), volume, num_levels=self.L, dx=self.dx, workspace=ws,
                                           corr_band=self.corr_band)
        if self.record_loop:
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
        disp, mask, _ = self.update_block.refine(geo, net, inp, att, disp, iters)
        if self.record_loop:
            ev[1].record()
            self.loop_events.append(ev)
        return disp, mask, comb

    # ---- host-resident inputs: double-buffered upload -----------------------------------------------
    @torch.no_grad()
    def run_pinned(self, batches, iters: int, out_host: Sequence[Tensor] | None = None) -> List[Tensor]:
        """Runs a sequence of batches whose inputs live in PINNED HOST memory (dicts with the keys of
        `run`: fl, fr, pl, pr, volume, disp, net, inp, att) and returns the final disparities in pinned
        host tensors.  Every batch is copied host->device on a copy stream into one of two static
        device slots while the previous batch computes (the 0.16-0.27 GB of inputs per pair take longer over
        PCIe than a third of the compute, so a serial copy costs 25 % of the throughput); the result
        goes device->host asynchronously.  Synchronises once, at the end."""
        batches = list(batches)
        if not batches:
            return []
        main = torch.cuda.current_stream(self.device)
        if not hasattr(self, "_copy"):
            self._copy = torch.cuda.Stream(self.device)
            self._slots = {}
        flat = lambda d: [t for k in ("fl", "fr", "pl", "pr", "volume", "disp") for t in (d[k],)] + \
            list(d["net"]) + list(d["inp"]) + list(d["att"])
        sig = tuple((tuple(t.shape), t.dtype) for t in flat(batches[0]))
        if sig not in self._slots:
            self._slots[sig] = [[torch.empty(t.shape, dtype=t.dtype, device=self.device) for t in flat(batches[0])]
                                for _ in range(2)]
        slots = self._slots[sig]
        ready = [torch.cuda.Event() for _ in range(2)]
        done = [torch.cuda.Event() for _ in range(2)]

        def upload(i: int) -> None:
            s = i & 1
            with torch.cuda.stream(self._copy):
                if i >= 2:
                    self._copy.wait_event(done[s])  # batch i-2 no longer reads the slot
                else:
                    self._copy.wait_stream(main)
                for dst, src in zip(slots[s], flat(batches[i])):
                    assert src.is_pinned(), "run_pinned takes pinned host tensors"
                    dst.copy_(src, non_blocking=True)
                ready[s].record(self._copy)

        outs: List[Tensor] = []
        upload(0)
        for i in range(len(batches)):
            s = i & 1
            main.wait_event(ready[s])
            if i + 1 < len(batches):
                upload(i + 1)
            t = slots[s]
            disp, _, _ = self.run(t[0], t[1], t[2], t[3], t[4], t[5], t[6:9], t[9:12], t[12:15], iters)
            done[s].record(main)
            dst = out_host[i] if out_host is not None else torch.empty(disp.shape, dtype=disp.dtype).pin_memory()
            dst.copy_(disp, non_blocking=True)
            outs.append(dst)
        main.synchronize()
        return outs
