"""The configuration bench.py times (BASELINE configs[1]: 1/4-res 120x160, C=128, D=104, L=2, r=4, 32 iterations, the
bench's own synthetic tensors and random-init weights) checked against the CPU oracle at FULL size -- the tile shapes
(`update._pick_block_n`), pair / row-reuse modes and the captured iteration graph of that shape are what runs here, not
a shrunken stand-in.  B = 2 DIFFERENT pairs, so a batch-index bug cannot hide.

Tolerance: mean |delta disparity| <= 0.05 px in FULL-RESOLUTION pixels (= 4 x the 1/4-res difference,
core/foundation_stereo.py:189) after 32 free-running iterations; the measured figures go to
gpurun_out/fullsize_parity.json (committed under profiles/)."""
import json
import os

import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = dict(C=128, H=120, W=160, D=104, Cp=12, Cg=28, G=8, L=2, r=4, iters=32)
REPORT = {}


def _save():
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "fullsize_parity.json"), "w") as fh:
        json.dump(REPORT, fh, indent=1, sort_keys=True)


@pytest.mark.parametrize("B,delta_scale", [(2, 1.0), (1, 0.25)])
def test_config1_full_size_32_iterations_vs_oracle(B, delta_scale):
    import foundationstereo_b200 as fsb
    from foundationstereo_b200 import synthetic
    c = CFG
    torch.set_num_threads(os.cpu_count() or 1)
    state = synthetic.random_update_block_state(seed=0, corr_levels=c["L"], corr_radius=c["r"], volume_dim=c["Cg"],
                                                delta_scale=delta_scale)
    x = synthetic.hot_path_inputs(B, c["C"], c["H"], c["W"], c["D"], c["Cp"], c["Cg"], seed=0)
    coords = torch.arange(c["W"], dtype=torch.float32).reshape(1, 1, -1, 1).repeat(B, c["H"], 1, 1)
    with torch.no_grad():
        want_gwc = oracle.gwc_volume(x["fl"], x["fr"], c["D"], c["G"])
        geo = oracle.GeoEncodingVolume(x["fl"], x["fr"], x["volume"], c["L"], c["r"])
        want_feat = geo(x["disp"], coords)
        want, want_mask, want_net = oracle.refine(state, geo, x["net"], x["inp"], x["att"], x["disp"], c["iters"])
    cu = lambda t: t.cuda()
    hp = fsb.StereoHotPath(state, max_disp=4 * c["D"], corr_levels=c["L"], corr_radius=c["r"], num_groups=c["G"])
    args = (cu(x["fl"]), cu(x["fr"]), cu(x["pl"]), cu(x["pr"]), cu(x["volume"]), cu(x["disp"]),
            [cu(t) for t in x["net"]], [cu(t) for t in x["inp"]], [cu(t) for t in x["att"]], c["iters"])
    got, mask, comb = hp.run(*args)
    got2, mask2, _ = hp.run(*args)   # second call = the captured-graph path the bench times
    torch.cuda.synchronize()
    rel = lambda a, b: float((a.double().cpu() - b.double()).abs().max() / b.double().abs().max())
    e_gwc = rel(comb[:, :c["G"]], want_gwc)
    vol = fsb.Combined_Geo_Encoding_Volume(cu(x["fl"]), cu(x["fr"]), cu(x["volume"]), num_levels=c["L"], dx=hp.dx)
    e_feat = rel(vol(cu(x["disp"]), cu(coords)), want_feat)
    mad_q = float((got.cpu() - want).abs().mean())
    mad_q2 = float((got2.cpu() - want).abs().mean())
    per_pair = [float((got[b].cpu() - want[b]).abs().mean()) * 4 for b in range(B)]
    moved = float((want - x["disp"]).abs().mean()) * 4
    REPORT[f"B{B}_delta_scale_{delta_scale:g}"] = {
        "gwc_rel": e_gwc, "lookup_rel": e_feat,
        "disparity_mad_px_fullres_eager_first_call": 4 * mad_q, "disparity_mad_px_fullres_graph_replay": 4 * mad_q2,
        "per_pair_mad_px_fullres": per_pair, "max_abs_px_fullres": 4 * float((got.cpu() - want).abs().max()),
        "oracle_loop_moved_disparity_px_fullres": moved,
        "mask_rel": rel(mask2, want_mask), "net0_rel": None if want_net is None else rel(hp.update_block.plan_for(B, c["H"], c["W"]).h32[0].permute(0, 3, 1, 2), want_net[0]),
    }
    _save()
    print(f"[fullsize] B={B} delta_scale={delta_scale}: " + json.dumps(REPORT[f"B{B}_delta_scale_{delta_scale:g}"]))
    assert e_gwc <= 1e-4 and e_feat <= 1e-4
    assert torch.isfinite(got).all()
    assert 4 * mad_q <= 0.05 and 4 * mad_q2 <= 0.05, (4 * mad_q, 4 * mad_q2)
    assert max(per_pair) <= 0.05, per_pair
