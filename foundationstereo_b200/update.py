"""Update block of the refinement loop: BasicSelectiveMultiUpdateBlock.forward
(core/update.py:141-159) as an engine over the reference's state_dict, running entirely on
libfsb200.so: every convolution / linear layer is the tcgen05 implicit-GEMM kernel
(csrc/conv.cu) with its activation fused, the GRU gate arithmetic (sigmoid, r*h, tanh, the
(1-z)h+zq update and the small/large blend, core/update.py:90-95,119) lives in the gate
epilogues, and the remaining glue (pool2x, interp, the two 7x7 convolutions) are the NHWC
helpers of csrc/nhwc.cu.  No torch operator runs inside a step; there is no cuDNN/eager
fallback.

Arithmetic: 16-bit operands (fp16 by default, like the reference's autocast region,
core/foundation_stereo.py:239), fp32 accumulation in TMEM, fp32 hidden state.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch

from . import _lib
from . import conv as C
from .conv import ConvLaunch, interleave_zr, pack_rows, pack_segment, pad64

Tensor = torch.Tensor

# output channels per CTA of each layer (tuning table; values must divide the padded row count)
DEFAULT_BLOCK_N = {
    "conv0": 128, "conv1": 128, "zr": 128, "q": 128,
    "convc1": 128, "convc2": 128, "convd2": 64, "enc_conv": 128,
    "head0": 128, "pw1": 128, "pw2": 128, "head4": 16, "mask0": 64, "mask2": 32,
    # the coarse levels have few pixel tiles (10 at 30x40, 40 at 60x80): narrower N tiles put
    # more CTAs on the 148 SMs (the K loop of a CTA is serial)
    "gru16.conv0": 32, "gru16.conv1": 32, "gru16.zr": 64, "gru16.q": 32,
    "gru08.zr": 64, "gru08.q": 32,
}


# layers fed through the halo variant of the kernel (FSB_CONV_LAYOUT_HALO: two stacked pixel tiles per
# CTA, one activation box per horizontal tap); the rest use FSB_CONV_LAYOUT_TAP.  The coarse GRU levels
# stay on the tap layout: their few pixel tiles need the narrow-N / many-CTA shape.
DEFAULT_HALO: set = set()


def _down(n: int) -> int:
    return (n + 1) // 2


FORK, JOIN = "fork", "join"
_ENC_D = ("encoder.convd1", "encoder.convd2")


class _Plan:
    """Static buffers + prepared launches for one (B, H, W).

    ``steps`` is the launch list of one refinement step.  Each entry runs on lane 0 (the
    caller's stream) or lane 1 (a side stream): the motion encoder does not depend on
    gru16/gru08 (core/update.py:143-151) and the mask head not on the disparity head
    (:155-158), and the coarse GRU levels have too few pixel tiles to fill 148 SMs, so the
    two chains run concurrently between FORK / JOIN markers (also inside the CUDA graph)."""

    def __init__(self, device: torch.device):
        self.steps: List[Callable[[], None]] = []
        self.names: List[str] = []
        self.lanes: List[int] = []
        self.tags: List[Optional[str]] = []
        self.groups: List[Optional[str]] = []  # "enc" | "coarse" | "fine" | "head": what `run_loop_pipelined` reorders
        self.lane = 0
        self.tag: Optional[str] = None
        self.group: Optional[str] = None
        self.keep: list = []
        self.accum = False
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.iter_graph: Optional[torch.cuda.CUDAGraph] = None
        self.loop_graphs: Dict[int, torch.cuda.CUDAGraph] = {}
        self.iter_key = None
        self.lookup: Optional[Callable[[], None]] = None
        self.fused_c1: Optional[Callable[[], None]] = None
        self.loop_active = False
        self.calls = 0
        self.iter_calls = 0
        self.flops = 0.0
        self.side = torch.cuda.Stream(device)

    def add(self, name: str, fn) -> None:
        self.steps.append(fn)
        self.names.append(name)
        self.lanes.append(self.lane)
        self.tags.append(self.tag)
        self.groups.append(self.group)


class _Convc1Step:
    """The encoder.convc1 step of a plan: the plain 1x1 convolution over the lookup buffer, or -- when `refine` has
    bound ``plan.fused_c1`` -- the lookup fused into it."""

    def __init__(self, plan: _Plan, conv):
        self.plan, self.conv = plan, conv
        self.flops = conv.flops
        self.args = conv.args

    def __call__(self) -> None:
        # only inside the captured, pipelined loop of refine(): the drop-in forward() is handed ITS caller's lookup tensor,
        # and the iteration-by-iteration order (run_steps) runs lookup + convc1 as two launches
        fused = self.plan.fused_c1 if self.plan.loop_active else None
        (fused or self.conv)()


class SelectiveUpdateBlock:
    """Holds packed 16-bit copies of the update-block weights (keys = reference state_dict
    keys) and runs one refinement step on CUDA tensors.

    forward(net, inp, corr, disp, att) keeps the reference contract: ``net`` is mutated in
    place and returned, outputs are fp32 (B,C,H,W) tensors.  The returned tensors are
    channels-last views of the engine's own buffers and stay valid until the next call.
    ``corr`` may be the reference's fp32 (B, L*K*(Cg+1), H, W) tensor or the 16-bit NHWC view
    produced by Combined_Geo_Encoding_Volume(..., nhwc16_out=self.corr_buffer), which is consumed in place.
    """

    def __init__(self, state_dict: Dict[str, Tensor], device: torch.device | str = "cuda",
                 dtype: torch.dtype = torch.float16, use_graph: bool = True,
                 block_n: Optional[Dict[str, int]] = None, halo: Optional[Sequence[str]] = None,
                 fused_mlp: bool = True, copy_outputs: bool = False, dot_head4: bool = True,
                 fused_block: Optional[bool] = None):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("SelectiveUpdateBlock runs on CUDA only (no CPU path)")
        _lib.lib()  # fail loudly if the library is missing
        assert dtype in (torch.float16, torch.bfloat16)
        import os as _os
        self.dtype = dtype
        self.use_graph = use_graph and _os.environ.get("FSB_GRAPH", "1") != "0"
        self.loop_graph = _os.environ.get("FSB_LOOP_GRAPH", "1") != "0"  # refine(): one graph for all iterations
        # refine(): pool2x for the NEXT iteration on the side lane under the disparity head instead of at the head of the
        # iteration.  Measured neutral on the B200 (90.5 / 91.4 with, 91.8 / 91.6 pairs/s without: the extra cross-lane
        # fork / join costs what the 45 us of pooling save), so it is off by default; FSB_TAIL_POOLS=1 switches it on.
        self.tail_pools = _os.environ.get("FSB_TAIL_POOLS", "0") == "1"
        self.pool_fp32 = _os.environ.get("FSB_POOL_FP32", "0") == "1"
        # refine(): the lookup fused into encoder.convc1 (csrc/lookconv.cu: the lookup tensor never reaches HBM, 354 MB of
        # DRAM traffic per 8-pair iteration gone).  B200, 8 pairs: 156 us against 93 + 53 us for the two launches alone,
        # 92.4 vs 91.6 pairs/s in the loop.  Same 16-bit operand values; the GEMM's K order differs (split by level), so
        # results agree to fp32 accumulation order (max |d disparity| 2.6e-5 px after 6 iterations in the test).
        # FSB_FUSED_LOOKUP_CONV=0 restores the two launches; radius != 4, more than two levels, Cg % 4 != 0 or a banded
        # correlation pyramid use them anyway, and so does the iteration-by-iteration order of `run_steps`: the fused launch
        # exists only in the pipelined loop, on the main lane (with it on the side lane of the iteration-by-iteration order a
        # full-size one-pair run stalled twice in nine test-suite runs).
        self.fused_lookup_conv = _os.environ.get("FSB_FUSED_LOOKUP_CONV", "1") != "0"
        # refine(): coarse GRU levels of iteration i+1 under the disparity head of iteration i (run_loop_pipelined).
        # Bit-identical results; B200, 8 pairs: 93.0 -> 93.9 pairs/s (2640 -> 2614 us per iteration) -- the step runs at
        # the 1000 W power cap, and the SM clock under load dropped from 1615 to 1555 MHz with the busier schedule, which
        # is why better overlap buys only 1 %.  FSB_PIPELINE=0 restores the iteration-by-iteration order.
        self.pipelined = _os.environ.get("FSB_PIPELINE", "1") != "0"
        # pipelined loop: convd1 / convd2 on the side lane behind the coarse levels (FSB_ENCD_SIDE=1).  Measured on the B200
        # (8 pairs, A/B twice): 91.6 / 91.7 against 93.5 / 93.2 pairs/s -- the extra cross-lane event and the earlier join
        # cost more than the 85 us they take off the main lane; off.
        self.encd_side = _os.environ.get("FSB_ENCD_SIDE", "0") == "1"
        self.pdl_cross = _os.environ.get("FSB_PDL_CROSS", "0") == "1"  # programmatic edges across lanes too (hung: see run_steps)
        self.two_lanes = _os.environ.get("FSB_TWO_LANES", "1") != "0"  # side stream for encoder / mask head
        self.copy_outputs = copy_outputs  # forward returns fresh tensors instead of views of the plan's buffers
        self.dot_head4 = dot_head4  # False: disp_head.conv.4 as a tensor-core convolution launch + a separate disp += delta
        # depthwise 7x7 + MLP of a disparity-head block as ONE launch (csrc/edgenext.cu; bit-identical to the two launches).
        # Measured on the B200: 178 us per block against 82 + 87 us, 92.9 vs 93.4 pairs/s -- the eight depthwise warps that
        # fit beside the 16 epilogue warps are slower than the stand-alone kernel's 16 (the depthwise part is
        # latency-bound at either occupancy), and with one X slot they only overlap the last 45 % of the previous tile.  Off by default
        # (FSB_FUSED_BLOCK=1 / fused_block=True switch it on).
        self.fused_block = (_os.environ.get("FSB_FUSED_BLOCK", "0") == "1") if fused_block is None else fused_block
        self.fused_mlp = fused_mlp  # False: pwconv1 / pwconv2 of the disparity head as two convolution launches
        self.bn = dict(DEFAULT_BLOCK_N)
        if block_n:
            self.bn.update(block_n)
        self.halo = set(DEFAULT_HALO if halo is None else halo)
        self.replayed_kernels = 0  # libfsb200 kernels launched through CUDA-graph replays (not seen by fsb_launch_count)
        self.fixed_tiles = block_n is not None or halo is not None  # explicit tuning tables switch the heuristics off
        sd = {k: v.detach().to(self.device, torch.float32) for k, v in state_dict.items()}
        # n_gru_layers of the reference module (core/update.py:128-132) = which GRUs its state_dict holds
        self.n_layers = 3 if "gru16.conv0.0.weight" in sd else 2 if "gru08.conv0.0.weight" in sd else 1
        if "gru04.small_gru.convz.weight" not in sd:
            raise _lib.FsbError("state_dict is not a BasicSelectiveMultiUpdateBlock's (no gru04.* keys, core/update.py:132)")
        self.hidden = sd["gru04.small_gru.convz.weight"].shape[0]
        if self.hidden != 128:
            raise _lib.FsbError(f"hidden_dim {self.hidden}: the kernels are built for 128 (core/foundation_stereo.py:137)")
        self.planes = sd["encoder.convc1.weight"].shape[1]
        self.planes_pad = pad64(self.planes)
        self.spec: Dict[str, tuple] = {}                    # weight key -> (launch name, segments, biases, Cout, block_n)
        self.w: Dict[tuple, Tuple[Tensor, Tensor, int]] = {}  # (key, layout, block_n) -> (packed weight, bias fp32, Cout)
        self._sd = sd
        self.f: Dict[str, Tensor] = {}                      # fp32 side parameters
        self._pack(sd)
        self._plans: Dict[Tuple[int, int, int], _Plan] = {}
        self._w1_levels = None   # convc1 weights in the per-level layout of fsb_lookup_convc1

    # ---- weight packing -----------------------------------------------------------------
    def _lay(self, launch: str) -> int:
        return _lib.FSB_CONV_LAYOUT_HALO if launch in self.halo else _lib.FSB_CONV_LAYOUT_TAP

    def _bn(self, key: str, g: Optional[str] = None) -> int:
        return self.bn.get(f"{g}.{key}", self.bn[key]) if g else self.bn[key]

    def _put(self, name: str, launch: str, segs, biases, cout: int, bn: int) -> None:
        """Register the weights of launch `launch` under key `name`: `segs` = one (weight, source
        channels) pair per segment, or a callable block_n -> (segs, biases) when the row order depends
        on the N tile (the interleaved z|r gates).  Packing (K order of the layout, row padding to the
        N tile) happens on first use per (layout, block_n) -- a plan may pick a different tile shape
        for a large batch than for a single pair."""
        self.spec[name] = (launch, segs, biases, cout, bn)

    def _weights(self, name: str, lay: int, bn: int) -> Tuple[Tensor, Tensor, int]:
        key = (name, lay, bn)
        hit = self.w.get(key)
        if hit is None:
            _, segs, biases, cout, _ = self.spec[name]
            if callable(segs):
                segs, biases = segs(bn)
            mats = [pack_segment(w, src, lay) for w, src in segs]
            rows = (max(m.shape[0] for m in mats) + bn - 1) // bn * bn
            w = pack_rows(mats, rows).to(self.dtype).contiguous()
            b = torch.zeros(len(biases), rows, dtype=torch.float32, device=self.device)
            for i, bi in enumerate(biases):
                b[i, :bi.numel()] = bi
            hit = self.w[key] = (w, b.contiguous(), cout)
        return hit

    def _pack(self, sd: Dict[str, Tensor]) -> None:
        hd = self.hidden
        bn = self.bn
        n = self.n_layers
        grus = [("gru04", (2 + (n > 1)) * hd)]                  # input widths: core/update.py:129-132
        if n >= 2:
            grus.append(("gru08", (2 + (n == 3)) * hd))
        if n == 3:
            grus.append(("gru16", 2 * hd))
        for g, cin in grus:
            nx = cin // hd
            self._put(f"{g}.conv0", f"{g}.conv0", [(sd[f"{g}.conv0.0.weight"], [(hd, hd)] * nx)],
                      [sd[f"{g}.conv0.0.bias"]], cin, self._bn("conv0", g))
            self._put(f"{g}.conv1", f"{g}.conv1", [(sd[f"{g}.conv1.0.weight"], [(cin, cin), (hd, hd)])],
                      [sd[f"{g}.conv1.0.bias"]], cin + hd, self._bn("conv1", g))
            # z and r gates of one GRU share a launch: rows interleaved per N tile [z | r]
            def zr(bnz, g=g, cin=cin):
                hx = [(cin + hd, cin + hd)]
                segs, zr_b = [], []
                for gru in ("large_gru", "small_gru"):
                    segs.append((interleave_zr(sd[f"{g}.{gru}.convz.weight"], sd[f"{g}.{gru}.convr.weight"], bnz), hx))
                    zr_b.append(interleave_zr(sd[f"{g}.{gru}.convz.bias"][:, None],
                                              sd[f"{g}.{gru}.convr.bias"][:, None], bnz).flatten())
                return segs, zr_b
            self._put(f"{g}.zr", f"{g}.zr", zr, None, 2 * hd, self._bn("zr", g))
            rhx = [(hd, hd), (cin, cin)]
            self._put(f"{g}.q", f"{g}.q", [(sd[f"{g}.large_gru.convq.weight"], rhx),
                                           (sd[f"{g}.small_gru.convq.weight"], rhx)],
                      [sd[f"{g}.large_gru.convq.bias"], sd[f"{g}.small_gru.convq.bias"]], hd, self._bn("q", g))
        P, Pp = self.planes, self.planes_pad
        self._put("convc1", "encoder.convc1", [(sd["encoder.convc1.weight"], [(P, Pp)])], [sd["encoder.convc1.bias"]],
                  256, bn["convc1"])
        self._put("convc2", "encoder.convc2", [(sd["encoder.convc2.weight"], [(256, 256)])], [sd["encoder.convc2.bias"]],
                  256, bn["convc2"])
        self._put("convd2", "encoder.convd2", [(sd["encoder.convd2.weight"], [(64, 64)])], [sd["encoder.convd2.bias"]],
                  64, bn["convd2"])
        self._put("enc_conv", "encoder.conv", [(sd["encoder.conv.weight"], [(256, 256), (64, 64)])],
                  [sd["encoder.conv.bias"]], hd - 1, bn["enc_conv"])
        self.f["convd1.w"] = sd["encoder.convd1.weight"].reshape(64, 49).contiguous()
        self.f["convd1.b"] = sd["encoder.convd1.bias"].contiguous()
        self._put("head0", "disp_head.conv.0", [(sd["disp_head.conv.0.weight"], [(hd, hd)])],
                  [sd["disp_head.conv.0.bias"]], hd, bn["head0"])
        for i in (2, 3):
            pre = f"disp_head.conv.{i}"
            self.f[f"dw{i}.w"] = sd[pre + ".dwconv.weight"].reshape(hd, 49).contiguous()
            self.f[f"dw{i}.b"] = sd[pre + ".dwconv.bias"].contiguous()
            self.f[f"gamma{i}"] = sd[pre + ".gamma"].contiguous()
            self.f[f"pw1_{i}.w"] = sd[pre + ".pwconv1.weight"].to(self.dtype).contiguous()   # (4hd, hd), nn.Linear layout
            self.f[f"pw1_{i}.b"] = sd[pre + ".pwconv1.bias"].contiguous()
            self.f[f"pw2_{i}.w"] = sd[pre + ".pwconv2.weight"].to(self.dtype).contiguous()   # (hd, 4hd)
            self.f[f"pw2_{i}.b"] = sd[pre + ".pwconv2.bias"].contiguous()
            self._put(f"pw1_{i}", pre + ".pwconv1", [(sd[pre + ".pwconv1.weight"], [(hd, hd)])],
                      [sd[pre + ".pwconv1.bias"]], 4 * hd, bn["pw1"])
            self._put(f"pw2_{i}", pre + ".pwconv2", [(sd[pre + ".pwconv2.weight"], [(4 * hd, 4 * hd)])],
                      [sd[pre + ".pwconv2.bias"]], hd, bn["pw2"])
        # (1, hd, 3, 3) -> [tap][channel] fp32 holding the 16-bit-rounded values the tensor-core path would multiply
        self.f["head4.w"] = sd["disp_head.conv.4.weight"][0].permute(1, 2, 0).reshape(9, hd).to(self.dtype).float().contiguous()
        self.f["head4.b"] = sd["disp_head.conv.4.bias"].contiguous()
        self._put("head4", "disp_head.conv.4", [(sd["disp_head.conv.4.weight"], [(hd, hd)])],
                  [sd["disp_head.conv.4.bias"]], 1, bn["head4"])
        self._put("mask0", "mask.0", [(sd["mask.0.weight"], [(hd, hd)])], [sd["mask.0.bias"]], 64, bn["mask0"])
        self._put("mask2", "mask.2", [(sd["mask.2.weight"], [(64, 64)])], [sd["mask.2.bias"]], 32, bn["mask2"])

    # ---- per-shape plan -------------------------------------------------------------------
    def _conv(self, plan: _Plan, name: str, wname: str, srcs, segs, bn_key: str, epilogue: int, **kw) -> None:
        lay, bn = plan.tile.get(name, (self._lay(name), self.bn.get(name, self.bn[bn_key])))
        if name not in plan.tile and not self.fixed_tiles and lay == _lib.FSB_CONV_LAYOUT_TAP:
            bn = self._pick_block_n(srcs, segs, self.spec[wname][3], epilogue, bn)
        w, b, cout = self._weights(wname, lay, bn)
        launch = ConvLaunch(srcs=srcs, segs=segs, weight=w, bias=b, cout=cout, block_n=bn, epilogue=epilogue, name=name,
                            layout=lay, **kw)
        plan.add(name, launch)
        plan.flops += launch.flops

    @staticmethod
    def _pick_block_n(srcs, segs, cout: int, epilogue: int, default: int, sms: int = 148) -> int:
        """N tile of a persistent tap-layout launch from a cost model of csrc/conv.cu (cycles):
             tile  = K steps x max(MMA 2*bn, operand feed (16 KB + 128*bn B) at ~80 B/cycle from L2) + pipeline fill
             total = ceil(pixel tiles x N tiles / SMs) x tile
        The pixel-tile counts of the three GRU levels (e.g. 38 / 160 / 600 at four 480x640 pairs) make the
        number of rounds, not the tile shape, the first-order term on the coarse levels."""
        B, H, W, _ = srcs[0].shape
        mt = B * min(-(-W // (1 << l)) * -(-H // (128 >> l)) for l in range(2, 8))
        nsteps = sum(k * k * sum(srcs[i].shape[3] // 64 for i in idx) for k, idx in segs)
        nacc = len(segs)
        if epilogue == _lib.FSB_EPI_GATE_ZR:
            # N = 256 (both gates of all 128 hidden channels in one tile, 2 x 256 = all 512 TMEM columns, accumulators NOT
            # double-buffered) is what this model prefers from 12 pairs of 480x640 on -- and on the B200 it HUNG the
            # two-lane, PDL-chained iteration graph there (2 of 2 runs; never with one lane, without PDL, without graphs
            # or with N = 128; profiles/r02f_hang_isolation.txt).  Root cause not established; N = 128 is also the
            # faster choice (98.7 vs 96.4 pairs/s at 12 pairs), so the gate launches stay on double-buffered tiles.
            cand, rows_of = [64, 128], lambda bn: 2 * cout // 2  # rows == 2 * hidden exactly
        elif epilogue == _lib.FSB_EPI_GATE_Q:
            cand, rows_of = [32, 64, 128], lambda bn: cout
        else:
            cand, rows_of = ([16] if cout <= 16 else []) + [32, 64, 96, 128, 192, 256], lambda bn: -(-cout // bn) * bn
        best, best_t = default, None
        for bn in cand:
            rows = rows_of(bn)
            if rows % bn or nacc * bn > 512 or nacc * rows > 1024:
                continue
            if epilogue == _lib.FSB_EPI_SCALE_RES and cout % 32:
                continue
            pair = nsteps >= 24 and mt >= 2 and bn % 16 == 0  # csrc/conv.cu pair mode: half a weight tile per SM
            tile = nsteps * max(2 * bn, (16384 + (64 if pair else 128) * bn) / 70.0) + 3000
            if nacc * bn > 256:          # accumulators not double-buffered: the epilogue is exposed
                tile += 60 * bn
            t = -(-mt * (rows // bn) // sms) * tile
            if best_t is None or t < best_t:
                best, best_t = bn, t
        return best

    def _hsrc(self, plan: _Plan) -> List[Tensor]:
        """What pool2x / interp read: the 16-bit copy of the hidden state (the reference's autocast region pools and
        interpolates fp16 tensors, core/foundation_stereo.py:239-240 -> core/update.py:143-152; half the bytes of the fp32
        master copy) unless FSB_POOL_FP32=1."""
        return plan.h32 if self.pool_fp32 else plan.h16

    def _gru(self, plan: _Plan, g: str, lvl: int, xs: Sequence[Tensor]) -> None:
        """SelectiveConvGRU.forward (core/update.py:112-119) on level `lvl`: 4 launches."""
        hd = self.hidden
        B, H, W = plan.shape[lvl]
        cin = hd * len(xs)
        dev, dt = self.device, self.dtype
        new = lambda c: torch.empty(B, H, W, c, dtype=dt, device=dev)
        x16, hx16 = new(cin), new(cin + hd)
        z_l, z_s, rh_l, rh_s = new(hd), new(hd), new(hd), new(hd)
        h32, h16, att = plan.h32[lvl], plan.h16[lvl], plan.att[lvl]
        n = len(xs)
        self._conv(plan, f"{g}.conv0", f"{g}.conv0", list(xs), [(3, list(range(n)))], "conv0", _lib.FSB_EPI_RELU, out=x16)
        self._conv(plan, f"{g}.conv1", f"{g}.conv1", [x16, h16], [(3, [0, 1])], "conv1", _lib.FSB_EPI_RELU, out=hx16)
        self._conv(plan, f"{g}.zr", f"{g}.zr", [hx16], [(3, [0]), (1, [0])], "zr", _lib.FSB_EPI_GATE_ZR,
                   h32=h32, z_l=z_l, z_s=z_s, rh_l=rh_l, rh_s=rh_s, hidden=hd)
        self._conv(plan, f"{g}.q", f"{g}.q", [rh_l, rh_s, x16], [(3, [0, 2]), (1, [1, 2])], "q", _lib.FSB_EPI_GATE_Q,
                   h32=h32, att=att, z_l=z_l, z_s=z_s, h32_out=h32, h16_out=h16, hidden=hd)

    def _build_plan(self, B: int, H: int, W: int) -> _Plan:
        hd, dev, dt = self.hidden, self.device, self.dtype
        plan = _Plan(dev)
        # tile shape per launch: explicit entries here override the cost model of _pick_block_n (used for every
        # tap-layout launch).  Since the tap kernel became persistent it matches or beats the halo variant on every
        # layer shape of the block (tools/conv_sweep.py: 3x3 512->512 288 vs 295 us, 384->384 182 vs 184 us), so
        # no launch defaults to the halo layout any more; `halo=[...]` still selects it per launch.
        plan.tile = {}
        import os as _os
        for kv in _os.environ.get("FSB_TILE", "").split(","):   # experiments: FSB_TILE="gru04.zr=128,gru16.conv1=128"
            if "=" in kv:
                k, v = kv.split("=")
                plan.tile[k.strip()] = (_lib.FSB_CONV_LAYOUT_TAP, int(v))
        plan.shape = [(B, H, W), (B, _down(H), _down(W)), (B, _down(_down(H)), _down(_down(W)))]
        new = lambda lvl, c, dtype=dt: torch.empty(*plan.shape[lvl], c, dtype=dtype, device=dev)
        plan.h32 = [new(l, hd, torch.float32) for l in range(3)]
        plan.h16 = [new(l, hd) for l in range(3)]
        plan.inp16 = [new(l, hd) for l in range(3)]
        plan.att = [torch.empty(plan.shape[l], dtype=torch.float32, device=dev) for l in range(3)]
        plan.disp = torch.empty(B, 1, H, W, dtype=torch.float32, device=dev)
        plan.delta = torch.empty(B, 1, H, W, dtype=torch.float32, device=dev)
        plan.mask = torch.empty(B, 32, H, W, dtype=torch.float32, device=dev)
        plan.corr16 = torch.zeros(B, H, W, self.planes_pad, dtype=dt, device=dev)
        plan.coords = torch.arange(W, dtype=torch.float32, device=dev).reshape(1, 1, W, 1).repeat(B, H, 1, 1)
        plan.src_key = [None] * 3  # (tensor, version) of the inp/att tensors last converted
        add = plan.add
        n = self.n_layers
        if n == 1:
            # core/update.py:152: with n_gru_layers == 1 the reference's forward never calls gru04 -- net passes through
            # unchanged and the motion encoder's result is unused; only the two heads run.
            pass
        else:
            # encoder (core/update.py:62-70): motion = [conv(cat[cor, d]) (127) | disp (1)]  -- side lane
            add(FORK, FORK)
            plan.lane = 1
            plan.group = "enc"
            c1, c2, d1, d2, motion = new(0, 256), new(0, 256), new(0, 64), new(0, 64), new(0, hd)
            self._conv(plan, "encoder.convc1", "convc1", [plan.corr16], [(1, [0])], "convc1", _lib.FSB_EPI_RELU, out=c1)
            # refine() may bind `plan.fused_c1`: ONE launch that gathers the lookup straight into this GEMM's shared
            # memory operand (csrc/lookconv.cu) -- then `plan.lookup` is a no-op and corr16 is not written at all
            plan.c1 = c1
            plan.steps[-1] = _Convc1Step(plan, plan.steps[-1])
            self._conv(plan, "encoder.convc2", "convc2", [c1], [(3, [0])], "convc2", _lib.FSB_EPI_RELU, out=c2)
            add("encoder.convd1", lambda: C.conv7x7_c1(plan.disp, self.f["convd1.w"], self.f["convd1.b"], d1, 64, motion,
                                                        hd - 1))
            self._conv(plan, "encoder.convd2", "convd2", [d1], [(3, [0])], "convd2", _lib.FSB_EPI_RELU, out=d2)
            self._conv(plan, "encoder.conv", "enc_conv", [c2, d2], [(3, [0, 1])], "enc_conv", _lib.FSB_EPI_RELU, out=motion)
            plan.lane = 0
            plan.group = "coarse"
            plan.keep += [c1, c2, d1, d2, motion]
            if n == 3:
                # gru16: x = [inp[2], pool2x(net[1])]                               (core/update.py:143)
                p1 = new(2, hd)
                plan.tag = "pool_head"
                add("pool2x(net1)", lambda: C.pool2x(self._hsrc(plan)[1], p1))
                plan.tag = None
                self._gru(plan, "gru16", 2, [plan.inp16[2], p1])
                plan.keep.append(p1)
            # gru08: x = [inp[1], pool2x(net[0]) (, interp(net[2]) with three levels)]  (:145-148)
            p0 = new(1, hd)
            plan.tag = "pool_head"
            add("pool2x(net0)", lambda: C.pool2x(self._hsrc(plan)[0], p0))
            plan.tag = None
            xs08 = [plan.inp16[1], p0]
            if n == 3:
                u2 = new(1, hd)
                add("interp(net2)", lambda: C.resize(self._hsrc(plan)[2], u2))
                xs08.append(u2)
            self._gru(plan, "gru08", 1, xs08)
            plan.group = None
            add(JOIN, JOIN)
            plan.group = "fine"
            # gru04: x = [inp[0], motion, interp(net[1])]                           (:150-153)
            u1 = new(0, hd)
            add("interp(net1)", lambda: C.resize(self._hsrc(plan)[1], u1))
            self._gru(plan, "gru04", 0, [plan.inp16[0], motion, u1])
            plan.group = None
            plan.keep += xs08 + [u1]
        if n >= 2:
            # `refine` (which owns the loop) pools the NEXT iteration's inputs here, on the side lane under the disparity
            # head, instead of at the head of the next iteration: pool2x(net[1]) / pool2x(net[0]) only depend on hidden
            # states that are final once gru08 / gru04 have run (core/update.py:143-146), and the two small streaming
            # kernels (45 us at 8 pairs) leave the critical path.  The drop-in `forward` keeps them at the head.
            plan.tag = "pool_tail"
            add(FORK, FORK)
            plan.lane = 1
            if n == 3:
                add("pool2x(net1)'", lambda: C.pool2x(self._hsrc(plan)[1], p1))
            add("pool2x(net0)'", lambda: C.pool2x(self._hsrc(plan)[0], p0))
            plan.lane = 0
            plan.tag = None
        # mask head (core/update.py:134-139,158) -- side lane; tagged so `refine` can leave it out of the iteration
        # graph and run it once after the last iteration (only that one's mask is consumed in test_mode,
        # core/foundation_stereo.py:243-247)
        plan.tag = "mask"
        add(FORK, FORK)
        plan.lane = 1
        m0 = new(0, 64)
        self._conv(plan, "mask.0", "mask0", [plan.h16[0]], [(3, [0])], "mask0", _lib.FSB_EPI_RELU, out=m0)
        self._conv(plan, "mask.2", "mask2", [m0], [(3, [0])], "mask2", _lib.FSB_EPI_RELU, out_f32=plan.mask,
                   out_scale=0.25)
        plan.lane = 0
        plan.tag = None
        # disp head (core/update.py:20-32) with EdgeNextConvEncoder (core/submodule.py:565-591)
        plan.group = "head"
        y = [new(0, hd) for _ in range(3)]
        t, u = new(0, hd), new(0, 4 * hd)
        self._conv(plan, "disp_head.conv.0", "head0", [plan.h16[0]], [(3, [0])], "head0", _lib.FSB_EPI_RELU, out=y[0])
        for k, i in enumerate((2, 3)):
            if self.fused_mlp and self.fused_block:
                # the whole EdgeNeXt block -- depthwise 7x7 -> pwconv1 -> GELU -> pwconv2 -> gamma * . + residual -- in ONE
                # launch (csrc/edgenext.cu): the depthwise tile is produced in shared memory as the first GEMM's operand
                def block(k=k, i=i):
                    C.edgenext_block(y[k], self.f[f"dw{i}.w"], self.f[f"dw{i}.b"], self.f[f"pw1_{i}.w"], self.f[f"pw1_{i}.b"],
                                     self.f[f"pw2_{i}.w"], self.f[f"pw2_{i}.b"], self.f[f"gamma{i}"], y[k + 1])
                block.flops = 2.0 * B * H * W * (2 * hd * 4 * hd)
                add(f"disp_head.conv.{i}.block", block)
                plan.flops += block.flops
                continue
            add(f"disp_head.conv.{i}.dwconv",
                lambda k=k, i=i: C.dwconv7x7(y[k], self.f[f"dw{i}.w"], self.f[f"dw{i}.b"], t))
            if self.fused_mlp:
                # pwconv1 -> GELU -> pwconv2 -> gamma * . + residual in one launch (csrc/mlp.cu): the 512-wide hidden
                # activations never leave the SM
                def mlp(k=k, i=i):
                    C.edgenext_mlp(t, self.f[f"pw1_{i}.w"], self.f[f"pw1_{i}.b"], self.f[f"pw2_{i}.w"],
                                   self.f[f"pw2_{i}.b"], self.f[f"gamma{i}"], y[k], y[k + 1])
                mlp.flops = 2.0 * B * H * W * (2 * hd * 4 * hd)
                add(f"disp_head.conv.{i}.mlp", mlp)
                plan.flops += mlp.flops
                continue
            self._conv(plan, f"disp_head.conv.{i}.pwconv1", f"pw1_{i}", [t], [(1, [0])], "pw1", _lib.FSB_EPI_GELU, out=u)
            self._conv(plan, f"disp_head.conv.{i}.pwconv2", f"pw2_{i}", [u], [(1, [0])], "pw2", _lib.FSB_EPI_SCALE_RES,
                       out=y[k + 1], res=y[k], gamma=self.f[f"gamma{i}"])
        if self.dot_head4:
            # 128 -> 1 channels: a dot product per pixel on CUDA cores (csrc/nhwc.cu), not an N = 16 tensor tile; in the
            # loop (`refine`) the same launch also does disp += delta (core/foundation_stereo.py:242)
            add("disp_head.conv.4", lambda: C.conv3x3_cout1(y[2], self.f["head4.w"], self.f["head4.b"], plan.delta,
                                                             plan.disp if plan.accum else None))
        else:
            self._conv(plan, "disp_head.conv.4", "head4", [y[2]], [(3, [0])], "head4", _lib.FSB_EPI_BIAS,
                       out_f32=plan.delta)
        plan.group = None
        add(JOIN, JOIN)
        plan.keep += [y, t, u, m0]
        return plan

    # ---- ingest ---------------------------------------------------------------------------
    @staticmethod
    def _is_view_of(t: Tensor, buf: Tensor) -> bool:
        """t is the (B,C,H,W) channels-last view of the NHWC buffer `buf`."""
        B, H, W, Cc = buf.shape
        return (t.data_ptr() == buf.data_ptr() and t.dtype == buf.dtype and tuple(t.shape) == (B, Cc, H, W)
                and t.stride() == (H * W * Cc, 1, W * Cc, Cc))

    def _corr_view(self, corr: Tensor) -> Optional[Tensor]:
        """The (B,H,W,Cpad) buffer behind a 16-bit NHWC lookup result, or None."""
        B, P, H, W = corr.shape
        Pp = self.planes_pad
        if corr.dtype == self.dtype and P == self.planes and corr.stride() == (H * W * Pp, 1, W * Pp, Pp):
            return torch.as_strided(corr, (B, H, W, Pp), (H * W * Pp, W * Pp, Pp, 1))
        return None

    def _ingest(self, plan: _Plan, net, inp, att, corr: Optional[Tensor], disp: Tensor) -> None:
        nl = self.n_layers
        if len(net) < nl or len(inp) < nl or len(att) < nl:
            raise _lib.FsbError(f"net / inp / att must hold {nl} pyramid levels (n_gru_layers of this update block), got "
                                f"{len(net)} / {len(inp)} / {len(att)}")
        for l in range(nl):
            n = net[l]
            if tuple(n.shape) != (plan.shape[l][0], self.hidden) + tuple(plan.shape[l][1:]):
                raise _lib.FsbError(f"net[{l}] has shape {tuple(n.shape)}, expected "
                                    f"{(plan.shape[l][0], self.hidden) + tuple(plan.shape[l][1:])}")
            if not self._is_view_of(n, plan.h32[l]):
                C.nchw_to_nhwc(n.float().contiguous(), plan.h16[l], plan.h32[l])
            # the context tensors are converted once and reused while the caller passes the SAME tensor objects,
            # unmodified (the reference loop does).  The key holds the tensors themselves: an address alone can be
            # recycled by the caching allocator for another batch's data.
            # Inference tensors (torch.inference_mode, scripts/train.py:515) keep no version counter: None = never
            # trusted, reconvert every call.
            ver = lambda t: None if t.is_inference() else t._version
            key = (inp[l], ver(inp[l]), att[l], ver(att[l]))
            old = plan.src_key[l]
            if (old is None or key[1] is None or key[3] is None or old[0] is not key[0] or old[2] is not key[2]
                    or old[1] != key[1] or old[3] != key[3]):
                C.nchw_to_nhwc(inp[l].float().contiguous(), plan.inp16[l], None)
                plan.att[l].copy_(att[l].reshape(plan.att[l].shape))
                plan.src_key[l] = key
        if disp.data_ptr() != plan.disp.data_ptr():
            plan.disp.copy_(disp)
        if corr is not None:
            view = self._corr_view(corr)
            if view is None or view.data_ptr() != plan.corr16.data_ptr():
                # the reference's fp32 (B, L*K*(Cg+1), H, W) tensor (or a foreign 16-bit one)
                if corr.shape[1] != self.planes:
                    raise _lib.FsbError(f"corr has {corr.shape[1]} channels, encoder.convc1 takes {self.planes} "
                                        "(corr_levels * (2*corr_radius+1) * (volume_dim+1), core/update.py:55)")
                C.nchw_to_nhwc(corr.float().contiguous(), plan.corr16, None)

    @staticmethod
    def _no_autograd(*tensors) -> None:
        """The engine is inference-only: its outputs carry no graph.  Raise rather than hand detached
        tensors to a caller that is recording one (scripts/train.py:333 differentiates through the block)."""
        if torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors):
            raise _lib.FsbError("the B200 update block is inference-only (no autograd formula): call it under "
                                "torch.no_grad() / torch.inference_mode(), or uninstall() it for training")

    # ---- one step -------------------------------------------------------------------------
    def plan_for(self, B: int, H: int, W: int) -> _Plan:
        key = (B, H, W)
        plan = self._plans.get(key)
        if plan is None:
            with torch.cuda.device(self.device):
                plan = self._plans[key] = self._build_plan(B, H, W)
        return plan

    def corr_buffer(self, B: int, H: int, W: int) -> Tensor:
        """The plan's 16-bit (B,H,W,Cpad) lookup buffer: Combined_Geo_Encoding_Volume's fused
        output writes here and convc1 reads it in place."""
        return self.plan_for(B, H, W).corr16

    def run_steps(self, plan: _Plan, lookup: bool = False, mask: bool = True, only_mask: bool = False,
                  tail_pools: bool = False, only_tag: Optional[str] = None) -> None:
        """The launches of one refinement step, issued on the current stream and the plan's
        side stream (no graph).  With ``lookup`` the bound geometry lookup runs first and the
        disparity update (core/foundation_stereo.py:242) last: one whole loop iteration.
        ``mask=False`` leaves the mask head out; ``only_mask`` / ``only_tag`` run nothing but the steps of that tag (on
        the current stream); ``tail_pools``: the pool2x launches run at the end of the step for the NEXT one (see
        `_build_plan`) instead of at its head."""
        if only_mask:
            only_tag = "mask"
        only = only_tag is not None

        def skip(tag):
            if only:
                return tag != only_tag
            if tag == "mask":
                return not mask
            if tag == "pool_head":
                return tail_pools
            if tag == "pool_tail":
                return not tail_pools
            return False

        main = torch.cuda.current_stream()
        plan.accum = lookup and self.dot_head4 and not only  # disp += delta inside the disp_head.conv.4 launch
        if lookup:
            plan.lookup()
        single = not self.two_lanes
        # A launch that follows a cross-lane wait goes out WITHOUT the programmatic-dependent-launch attribute, so the
        # captured graph keeps a full (completion) edge across lanes and only same-lane edges are programmatic.
        # With every edge programmatic the two-lane iteration graph hung on the B200 from 12 pairs of 480x640 on
        # (profiles/r02f_hang_isolation.txt); FSB_PDL_CROSS=1 restores that behaviour for experiments.
        lib = _lib.lib()
        fence = [False, False]

        def launch(fn, lane):
            if fence[lane] and not self.pdl_cross:
                fence[lane] = False
                old = lib.fsb_set_pdl(0)
                try:
                    fn()
                finally:
                    lib.fsb_set_pdl(old)
            else:
                fn()

        for fn, lane, tag in zip(plan.steps, plan.lanes, plan.tags):
            if skip(tag):
                continue
            if single or only:
                if fn is not FORK and fn is not JOIN:
                    fn()
                continue
            if fn is FORK:
                plan.side.wait_stream(main)
                fence[1] = True
            elif fn is JOIN:
                main.wait_stream(plan.side)
                fence[0] = True
            elif lane == 0:
                launch(fn, 0)
            else:
                with torch.cuda.stream(plan.side):
                    launch(fn, 1)
        if lookup and not plan.accum:
            C.axpy(plan.disp, plan.delta)

    def run_loop_pipelined(self, plan: _Plan, iters: int) -> None:
        """``iters`` whole loop iterations (lookup -> update block -> disp += delta, no mask head) with the coarse GRU
        levels of iteration i+1 issued UNDER the disparity head of iteration i:

            side lane:                 coarse(i+1) = pool2x, gru16, pool2x, interp, gru08
            main lane:  ... fine(i) -> head(i) -> lookup(i+1) -> encoder(i+1) -> [join] -> fine(i+1) = interp, gru04 -> ...

        gru16 / gru08 of the next iteration depend on the hidden states only (core/update.py:143-148), which are final
        once gru04 has run; the disparity head (:155), the lookup (core/foundation_stereo.py:238) and the motion encoder
        (core/update.py:149) neither read what the coarse levels write nor write what they read.  The same kernels on
        the same operands as `run_steps` -> bit-identical results; only the stream order differs: the head's
        latency-bound launches (depthwise 7x7, GELU MLP) share the GPU with the coarse levels' small grids instead of
        each running alone."""
        assert self.n_layers >= 2
        main = torch.cuda.current_stream()
        plan.accum = self.dot_head4
        fused = plan.fused_c1 is not None
        plan.loop_active = fused     # encoder.convc1 = the lookup fused into it (csrc/lookconv.cu); no lookup launch
        try:
            self._emit_pipelined(plan, iters, main, fused)
        finally:
            plan.loop_active = False

    def _emit_pipelined(self, plan: _Plan, iters: int, main, fused: bool) -> None:
        lib = _lib.lib()
        fence = [False, False]

        def launch(fn, lane):
            if fence[lane] and not self.pdl_cross:   # full completion edge across lanes (see run_steps)
                fence[lane] = False
                old = lib.fsb_set_pdl(0)
                try:
                    fn()
                finally:
                    lib.fsb_set_pdl(old)
            else:
                fn()

        def group(name, lane, only=None, skip=()):
            for fn, g, tag, nm in zip(plan.steps, plan.groups, plan.tags, plan.names):
                if g != name or fn is FORK or fn is JOIN or tag in ("mask", "pool_tail"):
                    continue
                if (only is not None and nm not in only) or nm in skip:
                    continue
                if lane == 0:
                    launch(fn, 0)
                else:
                    with torch.cuda.stream(plan.side):
                        launch(fn, 1)

        for i in range(iters):
            plan.side.wait_stream(main)          # fork: the hidden states of iteration i-1 are final
            fence[1] = True
            group("coarse", 1)
            if i > 0:
                group("head", 0)                 # disparity head of iteration i-1, disp += delta in its last launch
                if not plan.accum:
                    C.axpy(plan.disp, plan.delta)
            if self.encd_side:
                # the disparity branch of the motion encoder (convd1 -> convd2, core/update.py:66-67) needs only the new
                # disparity: it follows the coarse levels on the side lane, and the join moves in front of encoder.conv
                if i > 0:
                    ev = torch.cuda.Event()
                    ev.record(main)
                    plan.side.wait_event(ev)
                    fence[1] = True
                group("enc", 1, only=_ENC_D)
                if not fused:
                    plan.lookup()
                group("enc", 0, skip=_ENC_D + ("encoder.conv",))
                main.wait_stream(plan.side)      # join
                fence[0] = True
                group("enc", 0, only=("encoder.conv",))
            else:
                if not fused:
                    plan.lookup()
                group("enc", 0)
                main.wait_stream(plan.side)      # join
                fence[0] = True
            group("fine", 0)
        group("head", 0)
        if not plan.accum:
            C.axpy(plan.disp, plan.delta)

    def _capture(self, plan: _Plan, lookup: bool, mask: bool = True, repeat: int = 1,
                 tail_pools: bool = False, pipelined: bool = False) -> torch.cuda.CUDAGraph:
        # Dead Python cycles that hold CUDA graphs / streams of an earlier engine must not be collected DURING the
        # capture: destroying a graph or a stream from the capturing thread invalidates it
        # (cudaErrorStreamCaptureInvalidated; this torch no longer runs gc.collect() in torch.cuda.graph.__enter__).
        import gc
        g = torch.cuda.CUDAGraph()
        torch.cuda.synchronize()
        gc.collect()
        gc_was_on = gc.isenabled()
        gc.disable()
        n0 = _lib.launch_count()
        try:
            with torch.cuda.graph(g):
                if pipelined:
                    self.run_loop_pipelined(plan, repeat)
                else:
                    for _ in range(repeat):
                        self.run_steps(plan, lookup, mask, tail_pools=tail_pools)
        finally:
            if gc_was_on:
                gc.enable()
        g.fsb_kernels = _lib.launch_count() - n0  # libfsb200 kernels one replay launches
        return g

    def _replay(self, g: torch.cuda.CUDAGraph) -> None:
        g.replay()
        self.replayed_kernels += g.fsb_kernels

    def _execute(self, plan: _Plan) -> None:
        plan.calls += 1
        if not self.use_graph:
            return self.run_steps(plan)
        if plan.graph is None:
            if plan.calls == 1:  # first call eager: sets kernel attributes, warms the allocator
                return self.run_steps(plan)
            plan.graph = self._capture(plan, False)
        self._replay(plan.graph)

    def forward(self, net: List[Tensor], inp: Sequence[Tensor], corr: Tensor, disp: Tensor,
                att: Sequence[Tensor]) -> Tuple[List[Tensor], Tensor, Tensor]:
        """Same contract as core/update.py:141-159: mutates and returns ``net``; returns
        (net, mask, delta_disp) as fp32 tensors.  The results are views of the engine's own buffers,
        valid until the next call on the same (B,H,W) (the reference loop consumes them at once,
        core/foundation_stereo.py:240-247); ``copy_outputs=True`` at construction returns fresh tensors."""
        if not disp.is_cuda:
            raise _lib.FsbError("SelectiveUpdateBlock runs on CUDA tensors only (no CPU path)")
        self._no_autograd(*net, *inp, corr, disp, *att)   # checked in the CALLER's grad mode
        B, _, H, W = disp.shape
        with torch.no_grad(), torch.cuda.device(self.device):
            plan = self.plan_for(B, H, W)
            self._ingest(plan, net, inp, att, corr, disp)
            self._execute(plan)
            for l in range(self.n_layers):
                net[l] = plan.h32[l].permute(0, 3, 1, 2)
            if self.copy_outputs:
                for l in range(self.n_layers):
                    net[l] = net[l].clone()
                return net, plan.mask.clone(), plan.delta.clone()
        return net, plan.mask, plan.delta

    __call__ = forward

    # ---- the whole refinement loop (core/foundation_stereo.py:236-242) -----------------------
    def refine(self, geo, net: Sequence[Tensor], inp: Sequence[Tensor], att: Sequence[Tensor], disp: Tensor,
               iters: int, mask_every_iter: bool = False) -> Tuple[Tensor, Tensor, List[Tensor]]:
        """``iters`` x [geo lookup -> update block -> disp += delta] as ONE CUDA-graph replay for the whole loop (the
        first call on a plan issues the same launches eagerly): the lookup is gathered straight into convc1's
        shared-memory operand (csrc/lookconv.cu; or written as the 16-bit NHWC tensor convc1 reads), the update
        runs on the plan's static buffers and the disparity is updated in place, so there is no
        host work or host sync between iterations.  ``geo`` is a Combined_Geo_Encoding_Volume
        (its level buffers should have stable addresses -- ``workspace=`` -- or the graph is
        re-captured).  Returns (disparity (B,1,H,W) fp32, mask features of the last step, net).

        The mask head (core/update.py:134-139,158) depends on net[0] only and, in test_mode, only the
        last iteration's result is consumed (core/foundation_stereo.py:243-247).  This method owns the
        loop, so by default it runs the mask head ONCE, after the last iteration -- the same values the
        reference's last call returns; ``mask_every_iter=True`` keeps it inside every iteration like the
        drop-in ``forward`` does."""
        self._no_autograd(*net, *inp, disp, *att)
        B, _, H, W = disp.shape
        with torch.no_grad(), torch.cuda.device(self.device):
            plan = self.plan_for(B, H, W)
            self._ingest(plan, list(net), inp, att, None, disp.float())
            key = tuple(t.data_ptr() for t in list(geo.geo_levels) + list(geo.corr_levels)) + (
                geo.radius, mask_every_iter, getattr(geo, "corr_band", None), self.fused_lookup_conv, self.pipelined, self.encd_side)
            if plan.iter_key != key:
                levels = (list(geo.geo_levels), list(geo.corr_levels), geo.radius)
                from . import ops
                band = getattr(geo, "corr_band", None)
                plan.fused_c1 = None
                Cg = levels[0][0].shape[2]
                if (self.fused_lookup_conv and band is None and self.n_layers >= 2 and levels[2] == 4
                        and len(levels[0]) <= 2 and Cg % 4 == 0 and Cg <= 28):
                    # lookup -> convc1 as ONE launch (csrc/lookconv.cu): the lookup tensor never reaches HBM
                    L_ = len(levels[0])
                    if self._w1_levels is None or self._w1_levels[0] != (L_, Cg):
                        self._w1_levels = ((L_, Cg), ops.pack_convc1_weight(self._sd["encoder.convc1.weight"], L_, 9, Cg, self.dtype),
                                           self._sd["encoder.convc1.bias"].float().contiguous())
                    _, w1, b1 = self._w1_levels
                    plan.fused_c1 = lambda: ops.lookup_convc1(levels[0], levels[1], plan.disp, plan.coords, levels[2], w1,
                                                              b1, plan.c1)
                    # only `run_loop_pipelined` (where the launch sits on the MAIN lane) uses it; the iteration-by-iteration
                    # order of `run_steps` (FSB_PIPELINE=0, mask_every_iter) keeps lookup + convc1
                    plan.lookup = lambda: ops.geo_lookup_nhwc16(levels[0], levels[1], plan.disp, plan.coords, levels[2],
                                                                plan.corr16)
                elif band is not None:
                    plan.lookup = lambda: ops.geo_lookup_band(levels[0], levels[1], plan.disp, plan.coords, levels[2], band,
                                                              plan.corr16)
                else:
                    plan.lookup = lambda: ops.geo_lookup_nhwc16(levels[0], levels[1], plan.disp, plan.coords, levels[2],
                                                                plan.corr16)
                plan.iter_key = key
                plan.iter_graph = None
                plan.loop_graphs = {}
                plan.iter_calls = 0
            self._refine_loop(plan, iters, mask_every_iter)
            if not mask_every_iter:
                self.run_steps(plan, only_mask=True)
            return plan.disp.clone(), plan.mask, [plan.h32[l].permute(0, 3, 1, 2) for l in range(self.n_layers)]

    def _refine_loop(self, plan: _Plan, iters: int, mask_every_iter: bool) -> None:
        if True:
            # prologue: the first iteration's pooled inputs; every iteration then pools for its successor (tail_pools)
            tp = self.tail_pools
            if tp:
                self.run_steps(plan, only_tag="pool_head")
            pipe = self.pipelined and self.two_lanes and self.n_layers >= 2 and not mask_every_iter and not tp
            if pipe and not (self.use_graph and self.loop_graph and plan.iter_calls > 0):
                # first call on a plan (or no graphs at all): the SAME pipelined launch order, issued eagerly -- every call
                # then runs the same kernels in the same order (bit-identical results from call to call), and the fused
                # lookup -> convc1 launch only ever runs on the main lane
                self.run_loop_pipelined(plan, iters)
                plan.iter_calls += iters
            elif self.use_graph and self.loop_graph and plan.iter_calls > 0:
                # the WHOLE loop as one graph (iters x 30-odd kernel nodes): programmatic dependent launch keeps
                # overlapping kernel tails across iteration boundaries and the host launches once per pair batch
                g = plan.loop_graphs.get(iters)
                if g is None:
                    g = plan.loop_graphs[iters] = self._capture(plan, True, mask_every_iter, iters, tail_pools=tp,
                                                                pipelined=pipe)
                self._replay(g)
                plan.iter_calls += iters
            else:
                for _ in range(iters):
                    plan.iter_calls += 1
                    if not self.use_graph or plan.iter_calls == 1:
                        self.run_steps(plan, True, mask_every_iter, tail_pools=tp)
                        continue
                    if plan.iter_graph is None:
                        plan.iter_graph = self._capture(plan, True, mask_every_iter, tail_pools=tp)
                    self._replay(plan.iter_graph)
