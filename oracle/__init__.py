"""CPU oracle for the FoundationStereo cost-volume + iterative-refinement hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``foundationstereo_b200/`` may import
this package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs do, and there only as the checker
or as the timed CPU baseline -- never as the thing shipped.

Parity status: PINNED against outputs of the reference itself.  The reference
(`/root/reference`, pure PyTorch) was imported in the build container and run
on seeded inputs by ``tests/golden/make_golden.py``; the resulting vectors are
committed under ``tests/golden/*.npz`` and ``tests/test_oracle_golden.py``
checks every oracle function against them.  The reference has no tests or
golden vectors of its own for this path.
"""
from .hotpath import (  # noqa: F401
    l2_normalize,
    gwc_volume,
    concat_volume,
    volume_stem,
    row_correlation,
    halve_last,
    corr_pyramid,
    geo_pyramid,
    lerp_rows,
    geo_lookup,
    GeoEncodingVolume,
    corr_band_geometry,
    band_limit_corr_pyramid,
    update_block,
    selective_gru,
    raft_gru,
    motion_encoder,
    disp_head,
    mask_head,
    pool2x,
    resize_like,
    refine,
    disparity_regression,
    softmax_disparity_regression,
    context_upsample,
    upsample_disp_tail,
)
from .params import make_update_block_params, update_block_param_shapes  # noqa: F401
from .inputs import make_inputs  # noqa: F401
