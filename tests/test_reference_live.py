"""CPU tests that run the UNMODIFIED reference (oracle/_ref, or /root/reference in the build container) beside the
oracle restatement on fresh seeded inputs -- the live counterpart of the committed golden vectors -- and check the
harness pieces the GPU end-to-end test relies on (stand-ins, recorder, the restated hot loop)."""
import pytest
import torch

import oracle
from conftest import rel_err
from oracle import refmodel as rm

pytestmark = pytest.mark.skipif(not rm.available(), reason="reference not available (tools/make_ref.py makes oracle/_ref)")


def test_volumes_and_geo_match_reference_live():
    ns = rm.import_reference()
    x = oracle.make_inputs(B=2, C=32, H=9, W=21, D=11, Cg=28, seed=21)
    with torch.no_grad():
        assert torch.equal(ns.build_gwc_volume(x["fl"], x["fr"], 11, 8), oracle.gwc_volume(x["fl"], x["fr"], 11, 8))
        assert torch.equal(ns.build_concat_volume(x["pl"], x["pr"], 11), oracle.concat_volume(x["pl"], x["pr"], 11))
        dx = torch.linspace(-4, 4, 9).reshape(1, 1, 9, 1)
        ref = ns.Combined_Geo_Encoding_Volume(x["fl"], x["fr"], x["volume"], num_levels=2, dx=dx)
        mine = oracle.GeoEncodingVolume(x["fl"], x["fr"], x["volume"], 2, 4)
        assert rel_err(mine(x["disp"], x["coords"]), ref(x["disp"], x["coords"])) <= 1e-5


def test_volume_stem_matches_reference_live():
    """oracle.volume_stem (the checker of the opt-in fused kernel fsb_volume_stem) against the reference's own sequence:
    build_gwc_volume, build_concat_volume, torch.cat and an nn.Conv3d(32, 28, kernel_size=1) as in corr_stem
    (core/foundation_stereo.py:207-213, 164-165)."""
    ns = rm.import_reference()
    x = oracle.make_inputs(B=2, C=32, H=5, W=21, D=11, Cg=28, seed=23)
    torch.manual_seed(5)
    conv0 = torch.nn.Conv3d(8 + 2 * x["pl"].shape[1], 28, kernel_size=1)
    with torch.no_grad():
        want = conv0(torch.cat([ns.build_gwc_volume(x["fl"], x["fr"], 11, 8), ns.build_concat_volume(x["pl"], x["pr"], 11)], 1))
        got = oracle.volume_stem(x["fl"], x["fr"], x["pl"], x["pr"], conv0.weight, conv0.bias, 11, 8)
    assert got.shape == want.shape
    assert rel_err(got, want) <= 1e-5


@pytest.mark.parametrize("n_layers,levels", [(3, 2), (2, 2), (1, 2), (3, 4)])
def test_update_block_matches_reference_live(n_layers, levels):
    """n_gru_layers 1..3 (core/update.py:128-132,142-153; with one layer the reference never calls gru04) and the
    in-tree training default corr_levels = 4 (configs/train/stereo_v1.json)."""
    p = oracle.make_update_block_params(seed=3, corr_levels=levels, n_gru_layers=n_layers)
    ub = rm.make_reference_update_block(p, corr_levels=levels, n_gru_layers=n_layers)
    x = oracle.make_inputs(B=1, C=16, H=16, W=24, D=24, Cg=28, seed=22)
    with torch.no_grad():
        geo = oracle.GeoEncodingVolume(x["fl"], x["fr"], x["volume"], levels, 4)
        corr = geo(x["disp"], x["coords"])
        nets = lambda: [t.clone() for t in x["net"][:n_layers]]
        want = ub(nets(), list(x["inp"]), corr, x["disp"], list(x["att"]))
        got = oracle.update_block(p, nets(), x["inp"], corr, x["disp"], x["att"])
    assert len(got[0]) == len(want[0]) == n_layers
    for a, b in zip(got[0], want[0]):
        assert rel_err(a, b) <= 1e-5
    assert rel_err(got[1], want[1]) <= 1e-5 and rel_err(got[2], want[2]) <= 1e-5


def test_full_model_runs_and_loop_restatement_matches_forward():
    """FoundationStereo (random init, small image) through the harness on the CPU: the stock forward's
    1/4-res trajectory equals `refmodel.reference_loop` fed with the recorded hot-path inputs, i.e. the
    driver the GPU test uses for the noise floor IS the reference's loop."""
    model = rm.build_model("vits", max_disp=64, mixed_precision=False, seed=0)
    left, right = rm.make_pair(1, 64, 96, max_disp=12, seed=2)
    fs = rm.import_model_module()
    with rm.sdpa_instead_of_flash_attn(), torch.no_grad():
        with rm.Recorder(model) as rec:
            up = model(left, right, iters=3, test_mode=True)
        assert rec.calls == 3 and up.shape == (1, 1, 64, 96)
        disp, mask, states = rm.reference_loop(model.update_block, fs.Combined_Geo_Encoding_Volume, rec.geo_args,
                                               rec.first, 3, autocast=False, trace=True)
        assert len(states) == 3
        again = model.upsample_disp(disp, mask.float(), model.stem_2(fs.normalize_image(left)).float())
    assert torch.allclose(again, up, atol=1e-5), float((again - up).abs().max())
