"""CPU: the Real-ESRGAN oracle (oracle/realesrgan_ref.py) against the committed golden vectors, which
``oracle/make_golden_rrdb.py`` minted by running the reference's own classes (AST-extracted from
models/enhancement/vsr/realesrgan_handler.py); plus the host-side weight plumbing of the GPU twin."""
import json
import os

import numpy as np
import pytest

from oracle import realesrgan_ref as R
import rrdb_emulation as E
from video_enhancer_proto_b200 import realesrgan as RG

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def rrdb_golden():
    return json.load(open(os.path.join(G, "rrdb_manifest.json")))["cases"], np.load(os.path.join(G, "rrdb_cases.npz"))


def test_oracle_reproduces_the_golden_frames(rrdb_golden):
    cases, arr = rrdb_golden
    assert len(cases) >= 5
    for key, c in cases.items():
        model = R.make_network(c["seed"], c["scale"], c["num_block"], c["gain"])
        got = R.enhance_frame(model, arr[key + "__in"], c["scale"], c["tile_size"], c["tile_pad"])
        assert np.array_equal(got, arr[key + "__out_u8"]), key
        assert (int(got.min()), int(got.max())) == (c["out_min"], c["out_max"])


def test_tile_plan_geometry():
    plan = R.tile_plan(70, 45, 32, 10)
    assert len(plan) == 3 * 2
    assert plan[0] == (0, 42, 0, 42, 0, 32, 0, 32)            # first tile: no padding on the image border
    assert plan[-1] == (54, 70, 22, 45, 64, 70, 32, 45)       # last tile: ragged core, padding only towards the inside
    cover = np.zeros((70, 45), int)
    for (_, _, _, _, y0, y1, x0, x1) in plan:
        cover[y0:y1, x0:x1] += 1
    assert (cover == 1).all()


def test_init_state_is_the_reference_procedure_and_state_plumbing():
    for scale, nb, seed in ((4, 2, 0), (2, 1, 5)):
        a = RG.init_state(seed, scale, nb)
        b = R.export_state(R.make_network(seed, scale, nb))
        assert list(a) == list(b) == RG.state_keys(nb, scale)
        for k in a:
            assert np.array_equal(a[k], b[k]) and a[k].shape == RG.expected_shape(k), k
        assert RG.infer_plan(a) == (nb, scale)
        assert len(a) == 2 + 30 * nb + 2 + 2 * (2 if scale == 4 else 1) + 4          # vex_rrdb_load_weights' count
    st = dict(RG.init_state(0, 4, 1))
    bad = dict(st)
    bad["body.0.rdb2.conv3.weight"] = bad["body.0.rdb2.conv3.weight"][:, :64]
    with pytest.raises(ValueError):
        RG.normalize_state(bad, 1, 4)
    bad = dict(st)
    del bad["conv_hr.bias"]
    with pytest.raises(KeyError):
        RG.normalize_state(bad, 1, 4)
    with pytest.raises(KeyError):
        RG.infer_plan({"conv_first.weight": 0})
    assert R.flops_per_input_pixel(4, 23) == 2 * 9 * (192 + 69 * 26624 + 4096 + 36 * 4096 + 16 * 192)


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        return
    from video_enhancer_proto_b200 import _lib
    h = RG.B200RealESRGAN(seed=0, num_block=1)
    with pytest.raises(_lib.VexError):
        h._enhance_frame(np.zeros((8, 8, 3), np.uint8))


def test_fp16_rounding_points_keep_the_23_block_network_inside_the_bar():
    """Why the GPU trunk is fp16 and what bar it can meet: the emulation of its rounding points (tests/rrdb_emulation.py)
    against the fp32 oracle on the reference configuration (23 blocks, scale 4).  Gain 1: within the 2 LSB bar with room;
    gain 6 (full-range output): the same relative error, a few LSB - the bound the full-size GPU test is judged against."""
    img = np.random.default_rng(5).integers(0, 256, (48, 56, 3), dtype=np.uint8)
    for gain, bound in ((1.0, 1), (6.0, 6)):
        model = R.make_network(2, 4, 23, gain)
        d = np.abs(E.process_single(model, img).astype(int) - R.process_single(model, img).astype(int))
        assert d.max() <= bound, (gain, int(d.max()))
        assert d.mean() < 0.1 * gain, (gain, float(d.mean()))
